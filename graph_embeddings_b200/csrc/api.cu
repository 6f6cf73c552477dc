// api.cu — C ABI of libgraphemb.so: context, graph hand-off (A1), matrix import/export (A6).
#include "common.cuh"

#include <stdio.h>
#include <string.h>

#include <thread>

static std::string g_create_err;

namespace {

// per-node fp64 weight totals in CSR order (DirectedWeighted.java:69-75: total += (float)w,
// sequential; UndirectedWeighted.java:59-71: out list first, then in list, one accumulator)
__global__ void node_totals(const int64_t* __restrict__ out_off, const float* __restrict__ out_w,
                            const int64_t* __restrict__ in_off, const float* __restrict__ in_w, int32_t V,
                            int32_t directed, double* __restrict__ out_total, double* __restrict__ in_total,
                            float* __restrict__ out_wmax, float* __restrict__ in_wmax) {
    const int32_t v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= V) return;
    double t = 0;
    float mo = 0.f, mi = 0.f;
    for (int64_t k = out_off[v]; k < out_off[v + 1]; k++) { t = __dadd_rn(t, (double)out_w[k]); mo = fmaxf(mo, out_w[k]); }
    for (int64_t k = in_off[v]; k < in_off[v + 1]; k++) mi = fmaxf(mi, in_w[k]);
    out_wmax[v] = mo;
    in_wmax[v] = mi;
    if (directed) {
        out_total[v] = t;
        double u = 0;
        for (int64_t k = in_off[v]; k < in_off[v + 1]; k++) u = __dadd_rn(u, (double)in_w[k]);
        in_total[v] = u;
    } else {
        for (int64_t k = in_off[v]; k < in_off[v + 1]; k++) t = __dadd_rn(t, (double)in_w[k]);
        out_total[v] = t;  // undirected: one total shared by both lists
        in_total[v] = t;
    }
}

__global__ void max_f32(const float* __restrict__ X, int64_t n, float* __restrict__ out) {
    float m = 0.f;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) m = fmaxf(m, X[k]);
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(reinterpret_cast<int*>(out), __float_as_int(m));
}

// ids of an imported COO must address the store: 0 <= id < V (one bad index from a host caller
// would otherwise become an out-of-bounds device write in the update kernel)
__global__ void check_ids(const int32_t* __restrict__ I, const int32_t* __restrict__ J, int64_t n, int32_t V,
                          int32_t* __restrict__ bad) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    bool b = false;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride)
        b |= (uint32_t)I[k] >= (uint32_t)V || (uint32_t)J[k] >= (uint32_t)V;
    if (b) *bad = 1;
}

// CSR validation on the device (one pass over the arrays already uploaded): offsets start at 0 and
// never decrease, every neighbour id is in range, every list is sorted by id and unique — the form
// getIn/OutNeighborhoods hands over (BookmarkColoring.java:49-50).  code: 1 off[0], 2 monotone,
// 3 id range, 4 order.
__global__ void check_csr_dev(const int64_t* __restrict__ off, const int32_t* __restrict__ nbr, int32_t V, int64_t E,
                              int32_t* __restrict__ code) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x, t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t0 == 0 && (off[0] != 0 || off[V] != E)) atomicMax(code, 1);
    for (int64_t v = t0; v < V; v += stride) {
        const int64_t b = off[v], e = off[v + 1];
        if (e < b || b < 0 || e > E) { atomicMax(code, 2); continue; }
        int64_t prev = -1;
        for (int64_t k = b; k < e; k++) {
            const int64_t u = nbr[k];
            if (u < 0 || u >= V) { atomicMax(code, 3); break; }
            if (u <= prev) { atomicMax(code, 4); break; }
            prev = u;
        }
    }
}

}  // namespace

// Host -> device copy of a pageable buffer through two pinned staging buffers: host threads fill one
// while the other crosses PCIe (the driver's own pageable path stages single-threaded, 3-6 GB/s).
int upload_staged(ge_ctx* ctx, void* dst, const void* src, size_t bytes) {
    if (!bytes) return GE_OK;
    cudaStream_t st = ctx->stream;
    if (bytes < ((size_t)8 << 20)) { GE_CUDA(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st)); return GE_OK; }
    const size_t chunk = (size_t)32 << 20;
    if (!ctx->h_stage[0]) {
        for (int k = 0; k < 2; k++) {
            GE_CUDA(cudaMallocHost(&ctx->h_stage[k], chunk));
            GE_CUDA(cudaEventCreateWithFlags(&ctx->ev_stage[k], cudaEventDisableTiming));
        }
    }
    unsigned nthr = std::thread::hardware_concurrency();
    nthr = nthr < 1 ? 1 : (nthr > 8 ? 8 : nthr);
    const size_t n = (bytes + chunk - 1) / chunk;
    for (size_t c = 0; c < n; c++) {
        const int b = (int)(c & 1);
        const size_t o = c * chunk, len = bytes - o < chunk ? bytes - o : chunk;
        // the copy out of this buffer (two chunks ago, or the tail of the previous upload) must be done
        if (ctx->stage_busy[b]) { GE_CUDA(cudaEventSynchronize(ctx->ev_stage[b])); ctx->stage_busy[b] = false; }
        char* h = (char*)ctx->h_stage[b];
        const char* sp = (const char*)src + o;
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nthr; t++) {
            const size_t lo = len * t / nthr, hi = len * (t + 1) / nthr;
            th.emplace_back([=] { memcpy(h + lo, sp + lo, hi - lo); });
        }
        for (auto& x : th) x.join();
        GE_CUDA(cudaMemcpyAsync((char*)dst + o, h, len, cudaMemcpyHostToDevice, st));
        GE_CUDA(cudaEventRecord(ctx->ev_stage[b], st));
        ctx->stage_busy[b] = true;
    }
    return GE_OK;
}

void ge_free_graph(ge_ctx* ctx) {
    cudaFree(ctx->d_out_off); cudaFree(ctx->d_in_off); cudaFree(ctx->d_out_nbr); cudaFree(ctx->d_in_nbr);
    cudaFree(ctx->d_out_w); cudaFree(ctx->d_in_w); cudaFree(ctx->d_out_total); cudaFree(ctx->d_in_total);
    cudaFree(ctx->d_out_wmax); cudaFree(ctx->d_in_wmax); ctx->d_out_wmax = ctx->d_in_wmax = nullptr;
    ctx->d_out_off = ctx->d_in_off = nullptr; ctx->d_out_nbr = ctx->d_in_nbr = nullptr;
    ctx->d_out_w = ctx->d_in_w = nullptr; ctx->d_out_total = ctx->d_in_total = nullptr;
    ctx->E_out = ctx->E_in = 0;
}

// per-node fp64 weight totals + heaviest edge for the CSR now on the device (both hand-off paths)
int ge_graph_finish(ge_ctx* ctx) {
    const int32_t V = ctx->V;
    cudaFree(ctx->d_out_total); cudaFree(ctx->d_in_total); cudaFree(ctx->d_out_wmax); cudaFree(ctx->d_in_wmax);
    ctx->d_out_total = ctx->d_in_total = nullptr; ctx->d_out_wmax = ctx->d_in_wmax = nullptr;
    GE_CUDA(cudaMalloc(&ctx->d_out_total, (size_t)V * 8));
    GE_CUDA(cudaMalloc(&ctx->d_in_total, (size_t)V * 8));
    GE_CUDA(cudaMalloc(&ctx->d_out_wmax, (size_t)V * 4));
    GE_CUDA(cudaMalloc(&ctx->d_in_wmax, (size_t)V * 4));
    node_totals<<<(V + 127) / 128, 128, 0, ctx->stream>>>(ctx->d_out_off, ctx->d_out_w, ctx->d_in_off, ctx->d_in_w, V,
                                                          ctx->cfg.directed, ctx->d_out_total, ctx->d_in_total,
                                                          ctx->d_out_wmax, ctx->d_in_wmax);
    GE_CUDA(cudaGetLastError());
    GE_CUDA(cudaStreamSynchronize(ctx->stream));
    return GE_OK;
}

extern "C" void ge_config_default(ge_config* c) {
    if (!c) return;
    memset(c, 0, sizeof(*c));
    c->dim = 300;                     // dblp/onstage/saa.config.yml:3
    c->method = GE_METHOD_PGLOVE;     // *.config.yml:2
    c->opt_method = GE_OPT_ADAGRAD;   // opt.method
    c->directed = 1;
    c->normalize = GE_NORM_NONE;
    c->alpha = 0.1;
    c->epsilon = 1e-3;
    c->tolerance = 1e-4;
    c->maxiter = 1000;
    c->seed = 12345;
    c->learning_rate = 0.05f;         // Optimizer.java:26
    c->math = GE_MATH_FAST;
    c->order = GE_ORDER_DEVICE;
    c->init = GE_INIT_JAVA;
}

extern "C" const char* ge_version(void) { return "graphemb-b200 0.1 (sm_100a)"; }

extern "C" const char* ge_last_error(const ge_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

extern "C" int ge_create(const ge_config* cfg, ge_ctx** out) {
    if (!cfg || !out) { g_create_err = "null argument"; return GE_ERR_INVALID; }
    *out = nullptr;
    // Configuration.check (Configuration.java:478-494) for the fields this path reads
    if (cfg->dim <= 0 || cfg->dim > 1024) { g_create_err = "Invalid configuration: No dimension specified (1..1024)"; return GE_ERR_INVALID; }
    if (cfg->method != GE_METHOD_GLOVE && cfg->method != GE_METHOD_PGLOVE) { g_create_err = "Invalid configuration: Invalid method, choose one of: glove, pglove"; return GE_ERR_INVALID; }
    if (!(cfg->alpha > 0) || !(cfg->epsilon > 0)) { g_create_err = "Invalid configuration: Invalid BCA parameters, alpha and epsilon are mandatory"; return GE_ERR_INVALID; }
    if (cfg->opt_method < GE_OPT_ADAGRAD || cfg->opt_method > GE_OPT_AMSGRAD) { g_create_err = "Invalid optimization method"; return GE_ERR_INVALID; }
    if (cfg->opt_method != GE_OPT_ADAGRAD && cfg->dim > 512) { g_create_err = "adam / amsgrad: dim must be <= 512"; return GE_ERR_INVALID; }
    if (cfg->normalize < 0 || cfg->normalize > 2) { g_create_err = "Invalid bca.normalize"; return GE_ERR_INVALID; }
    if (cfg->update < GE_UPDATE_ATOMIC || cfg->update > GE_UPDATE_ATOMIC_ADAM) { g_create_err = "Invalid update mode"; return GE_ERR_INVALID; }
    if (cfg->update == GE_UPDATE_ATOMIC_ADAM && cfg->opt_method == GE_OPT_ADAGRAD) {
        g_create_err = "GE_UPDATE_ATOMIC_ADAM applies to adam / amsgrad only (adagrad: GE_UPDATE_ATOMIC)";
        return GE_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        g_create_err = "no CUDA device: libgraphemb has no CPU fallback";
        return GE_ERR_NODEVICE;
    }
    if (cfg->device < 0 || cfg->device >= ndev) { g_create_err = "device ordinal out of range"; return GE_ERR_INVALID; }
    ge_ctx* ctx = new ge_ctx();
    ctx->cfg = *cfg;
    if (ctx->cfg.learning_rate == 0.f) ctx->cfg.learning_rate = 0.05f;
    cudaError_t e = cudaSetDevice(cfg->device);
    cudaDeviceProp prop;
    if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, cfg->device);
    if (e == cudaSuccess) {
        ctx->sm_count = prop.multiProcessorCount;
        if (prop.major < 10) {
            g_create_err = "device is not sm_100-class: this library is built for sm_100a only";
            delete ctx;
            return GE_ERR_NODEVICE;
        }
    }
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking);
    ctx->stream = ctx->own_stream;
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&ctx->ev1);
    if (e != cudaSuccess) { g_create_err = cudaGetErrorString(e); delete ctx; return GE_ERR_CUDA; }
    *out = ctx;
    return GE_OK;
}

extern "C" int ge_set_stream(ge_ctx* ctx, void* cuda_stream) {
    if (!ctx) return GE_ERR_INVALID;
    cudaSetDevice(ctx->cfg.device);
    cudaStreamSynchronize(ctx->stream);
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return GE_OK;
}

extern "C" void ge_destroy(ge_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->cfg.device);
    ge_dist_free(ctx);
    ge_free_sim(ctx);
    ge_free_store(ctx);
    ge_free_cooc(ctx);
    ge_free_graph(ctx);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    for (int k = 0; k < 2; k++) {
        if (ctx->h_stage[k]) cudaFreeHost(ctx->h_stage[k]);
        if (ctx->ev_stage[k]) cudaEventDestroy(ctx->ev_stage[k]);
    }
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

// device-side checks of the CSR now resident + per-node totals (both hand-off paths end here)
int ge_graph_validate(ge_ctx* ctx) {
    int32_t* d_code = nullptr;
    int32_t code[2] = {0, 0};
    GE_CUDA(cudaMalloc(&d_code, 8));
    cudaError_t e = cudaMemsetAsync(d_code, 0, 8, ctx->stream);
    if (e == cudaSuccess) {
        check_csr_dev<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(ctx->d_out_off, ctx->d_out_nbr, ctx->V, ctx->E_out, d_code);
        check_csr_dev<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(ctx->d_in_off, ctx->d_in_nbr, ctx->V, ctx->E_in, d_code + 1);
        e = cudaMemcpyAsync(code, d_code, 8, cudaMemcpyDeviceToHost, ctx->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_code);
    GE_CUDA(e);
    for (int k = 0; k < 2; k++) {
        if (!code[k]) continue;
        const char* name = k ? "in" : "out";
        const char* what[] = {"", "_off[0] != 0 or _off[V] != E", "_off not monotone", "_nbr id out of range",
                              " neighbour lists must be sorted by id and unique"};
        ctx->err = std::string(name) + what[code[k]];
        ge_free_graph(ctx);
        return GE_ERR_GRAPH;
    }
    return GE_OK;
}

int ge_graph_alloc(ge_ctx* ctx, int32_t V, int64_t E_out, int64_t E_in) {
    ge_free_graph(ctx);
    ge_free_cooc(ctx);
    ge_free_store(ctx);
    ctx->V = V;
    ctx->E_out = E_out;
    ctx->E_in = E_in;
    const size_t ob = (size_t)(V + 1) * sizeof(int64_t);
    GE_CUDA(cudaMalloc(&ctx->d_out_off, ob));
    GE_CUDA(cudaMalloc(&ctx->d_in_off, ob));
    GE_CUDA(cudaMalloc(&ctx->d_out_nbr, (size_t)(E_out + 1) * 4));
    GE_CUDA(cudaMalloc(&ctx->d_out_w, (size_t)(E_out + 1) * 4));
    GE_CUDA(cudaMalloc(&ctx->d_in_nbr, (size_t)(E_in + 1) * 4));
    GE_CUDA(cudaMalloc(&ctx->d_in_w, (size_t)(E_in + 1) * 4));
    return GE_OK;
}

extern "C" int ge_set_graph_csr(ge_ctx* ctx, int32_t V, const int64_t* out_off, const int32_t* out_nbr,
                                const float* out_w, const int64_t* in_off, const int32_t* in_nbr,
                                const float* in_w) {
    if (!ctx) return GE_ERR_INVALID;
    if (V <= 0 || !out_off || !in_off) { ctx->err = "bad graph arguments"; return GE_ERR_INVALID; }
    if (out_off[V] < 0 || in_off[V] < 0 || (out_off[V] > 0 && (!out_nbr || !out_w)) || (in_off[V] > 0 && (!in_nbr || !in_w))) {
        ctx->err = "bad graph arguments"; return GE_ERR_INVALID;
    }
    cudaSetDevice(ctx->cfg.device);
    int r = ge_graph_alloc(ctx, V, out_off[V], in_off[V]);
    if (r) { ge_free_graph(ctx); return r; }
    const size_t ob = (size_t)(V + 1) * sizeof(int64_t);
    const struct { void* d; const void* h; size_t n; } parts[] = {
        {ctx->d_out_off, out_off, ob}, {ctx->d_in_off, in_off, ob},
        {ctx->d_out_nbr, out_nbr, (size_t)ctx->E_out * 4}, {ctx->d_out_w, out_w, (size_t)ctx->E_out * 4},
        {ctx->d_in_nbr, in_nbr, (size_t)ctx->E_in * 4}, {ctx->d_in_w, in_w, (size_t)ctx->E_in * 4}};
    for (const auto& q : parts) {
        r = upload_staged(ctx, q.d, q.h, q.n);
        if (r) { ge_free_graph(ctx); return r; }
    }
    ctx->graph_seconds = 0;
    r = ge_graph_validate(ctx);   // synchronises: the caller's buffers are free again
    if (r) return r;
    return ge_graph_finish(ctx);
}

void ge_free_cooc(ge_ctx* ctx) {
    cudaFree(ctx->d_I); cudaFree(ctx->d_J); cudaFree(ctx->d_X); cudaFree(ctx->d_row_off);
    ctx->d_I = ctx->d_J = nullptr; ctx->d_X = nullptr; ctx->d_row_off = nullptr;
    ctx->nnz = 0; ctx->xmax = 0; ctx->have_cooc = false;
    cudaFree(ctx->d_nz); ctx->d_nz = nullptr;
    cudaFree(ctx->d_perm); ctx->d_perm = nullptr;
    ctx->h_perm.clear();
}

extern "C" int ge_cooc_build_rows(ge_ctx* ctx, int32_t first, int32_t count, int64_t* nnz, double* xmax) {
    if (!ctx) return GE_ERR_INVALID;
    if (!ctx->d_out_off) { ctx->err = "no graph: call ge_set_graph_csr first"; return GE_ERR_STATE; }
    cudaSetDevice(ctx->cfg.device);
    if (count < 0) { first = 0; count = ctx->V; }
    if (first < 0 || first + (int64_t)count > ctx->V) { ctx->err = "row range out of bounds"; return GE_ERR_INVALID; }
    ge_free_cooc(ctx);
    int r = ge_bca_build(ctx, first, count);
    if (r) return r;
    if (nnz) *nnz = ctx->nnz;
    if (xmax) *xmax = ctx->xmax;
    return GE_OK;
}
extern "C" int ge_cooc_build(ge_ctx* ctx, int64_t* nnz, double* xmax) { return ge_cooc_build_rows(ctx, 0, -1, nnz, xmax); }

// rows of a cyclic block layout (common.cuh ge_vid): the `count` virtual ids from `first_vid` on, i.e. whole
// blocks {b, b + nblocks, ...}; the COO comes out in virtual-id order with the real ids in I and J
extern "C" int ge_cooc_build_rows_cyclic(ge_ctx* ctx, int32_t nblocks, int32_t first_vid, int32_t count, int64_t* nnz, double* xmax) {
    if (!ctx) return GE_ERR_INVALID;
    if (!ctx->d_out_off) { ctx->err = "no graph: call ge_set_graph_csr first"; return GE_ERR_STATE; }
    cudaSetDevice(ctx->cfg.device);
    if (nblocks < 1 || nblocks > 32) { ctx->err = "nblocks out of range"; return GE_ERR_INVALID; }
    const int32_t Bs = (ctx->V + nblocks - 1) / nblocks;
    if (first_vid < 0 || count < 0 || first_vid % Bs || count % Bs || first_vid + (int64_t)count > (int64_t)nblocks * Bs) {
        ctx->err = "cyclic row range must be whole blocks"; return GE_ERR_INVALID;
    }
    ge_free_cooc(ctx);
    int r = ge_bca_build(ctx, first_vid, count, nblocks, Bs);
    if (r) return r;
    if (nnz) *nnz = ctx->nnz;
    if (xmax) *xmax = ctx->xmax;
    return GE_OK;
}

extern "C" int ge_cooc_stats(ge_ctx* ctx, int64_t* pops, int64_t* scanned, double* seconds) {
    if (!ctx) return GE_ERR_INVALID;
    if (pops) *pops = ctx->bca_pops;
    if (scanned) *scanned = ctx->bca_scanned;
    if (seconds) *seconds = ctx->bca_seconds;
    return GE_OK;
}

extern "C" int ge_cooc_fetched(ge_ctx* ctx, int64_t* fetched, int64_t* tier2_rows) {
    if (!ctx) return GE_ERR_INVALID;
    if (fetched) *fetched = ctx->bca_fetched;
    if (tier2_rows) *tier2_rows = ctx->bca_overflow_rows;
    return GE_OK;
}

extern "C" int ge_cooc_export(ge_ctx* ctx, int32_t* I, int32_t* J, float* X, int64_t cap) {
    if (!ctx) return GE_ERR_INVALID;
    if (!ctx->have_cooc) { ctx->err = "no co-occurrence matrix"; return GE_ERR_STATE; }
    if (!ctx->d_I) { ctx->err = "the COO arrays were released when the matrix was bucketed for the multi-GPU trainer"; return GE_ERR_STATE; }
    if (cap < ctx->nnz) { ctx->err = "export capacity too small"; return GE_ERR_INVALID; }
    cudaSetDevice(ctx->cfg.device);
    const size_t n = (size_t)ctx->nnz;
    if (I) GE_CUDA(cudaMemcpyAsync(I, ctx->d_I, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (J) GE_CUDA(cudaMemcpyAsync(J, ctx->d_J, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (X) GE_CUDA(cudaMemcpyAsync(X, ctx->d_X, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    GE_CUDA(cudaStreamSynchronize(ctx->stream));
    return GE_OK;
}

extern "C" int ge_cooc_import(ge_ctx* ctx, int32_t V, int64_t nnz, const int32_t* I, const int32_t* J,
                              const float* X, double xmax) {
    if (!ctx) return GE_ERR_INVALID;
    if (V <= 0 || nnz <= 0 || !I || !J || !X) { ctx->err = "bad matrix arguments"; return GE_ERR_INVALID; }
    if (V > GE_ID_MASK) { ctx->err = "vocabulary must stay below 2^27 vertices (packed nonzero ids)"; return GE_ERR_INVALID; }
    cudaSetDevice(ctx->cfg.device);
    if (ctx->V != V) { ge_free_store(ctx); ge_free_graph(ctx); }  // a resident graph of another V must not be walked with this V
    ge_free_cooc(ctx);
    ctx->V = V;
    const size_t n = (size_t)nnz;
    GE_CUDA(cudaMalloc(&ctx->d_I, n * 4));
    GE_CUDA(cudaMalloc(&ctx->d_J, n * 4));
    GE_CUDA(cudaMalloc(&ctx->d_X, n * 4));
    GE_CUDA(cudaMemcpyAsync(ctx->d_I, I, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    GE_CUDA(cudaMemcpyAsync(ctx->d_J, J, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    GE_CUDA(cudaMemcpyAsync(ctx->d_X, X, n * 4, cudaMemcpyHostToDevice, ctx->stream));
    {
        int32_t* d_bad = nullptr;
        int32_t bad = 0;
        GE_CUDA(cudaMalloc(&d_bad, 4));
        cudaError_t e = cudaMemsetAsync(d_bad, 0, 4, ctx->stream);
        if (e == cudaSuccess) {
            check_ids<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(ctx->d_I, ctx->d_J, nnz, V, d_bad);
            e = cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, ctx->stream);
        }
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        cudaFree(d_bad);
        if (e != cudaSuccess || bad) {
            cudaFree(ctx->d_I); cudaFree(ctx->d_J); cudaFree(ctx->d_X);
            ctx->d_I = ctx->d_J = nullptr; ctx->d_X = nullptr;
            GE_CUDA(e);
            ctx->err = "matrix index out of range: every I/J must satisfy 0 <= id < V";
            return GE_ERR_INVALID;
        }
    }
    ctx->nnz = nnz;
    if (xmax > 0) ctx->xmax = xmax;
    else {
        float* d_m = nullptr;
        GE_CUDA(cudaMalloc(&d_m, 4));
        GE_CUDA(cudaMemsetAsync(d_m, 0, 4, ctx->stream));
        max_f32<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(ctx->d_X, nnz, d_m);
        float m = 0;
        GE_CUDA(cudaMemcpyAsync(&m, d_m, 4, cudaMemcpyDeviceToHost, ctx->stream));
        GE_CUDA(cudaStreamSynchronize(ctx->stream));
        cudaFree(d_m);
        ctx->xmax = (double)m;
    }
    GE_CUDA(cudaStreamSynchronize(ctx->stream));
    ctx->have_cooc = true;
    return GE_OK;
}

extern "C" int64_t ge_cooc_count(ge_ctx* ctx) { return ctx ? ctx->nnz : 0; }
extern "C" int32_t ge_vocab_size(ge_ctx* ctx) { return ctx ? ctx->V : 0; }
extern "C" double ge_cooc_max(ge_ctx* ctx) { return ctx ? ctx->xmax : 0; }
