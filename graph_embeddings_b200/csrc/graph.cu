// graph.cu — adjacency flattening on the device.
//
// Replaces what the reference does on the CPU before the first BCA job can run
// (J/bca/BookmarkColoring.java:49-52):
//   graph.getOutNeighborhoods() / getInNeighborhoods()          unique neighbour sets per vertex
//   Out/InEdgeNeighborhoodAlgorithm.compute                      the edge aligned with each neighbour
//   (J/convert/util/EdgeNeighborhoodAlgorithm.java:42-77, getEdge :22-33; weights read through
//    edgeWeights.getValueAsFloat(edge), J/bca/jobs/DirectedWeighted.java:73,81)
// Input: the raw directed edge list (src, dst, f32 weight), parallel edges allowed.  Output: the
// two CSRs ge_set_graph_csr expects (neighbour lists sorted by id, unique).  Parallel edges s->d
// collapse to one neighbour; the weight kept is that of the FIRST such edge in input order (grph
// picks by its internal set order: parity unpinned for multigraphs) — the same edge for
// the out- and the in-list.
//
// Method: pack (major, minor) into one 64-bit key of 2*ceil(log2 V) bits, stable LSD radix sort
// (8-bit digits; per pass: per-tile histogram -> exclusive scan -> stable scatter with warp
// match-any ranking) carrying the edge index, flag group heads, scan, compact, offsets by binary
// search.  All HBM-bound streaming; algorithmic bytes 12 B/edge in + 8 B/edge out per CSR.
#include "common.cuh"

#include <mutex>
#include <utility>

namespace {

constexpr int RS_THREADS = 256;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 keys per block
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_PER_WARP = RS_TILE / RS_WARPS;  // 512 contiguous keys per warp

__global__ void pack_keys(const int32_t* __restrict__ src, const int32_t* __restrict__ dst, int64_t E, int32_t V,
                          int idbits, int by_dst, uint64_t* __restrict__ keys, uint32_t* __restrict__ vals,
                          int32_t* __restrict__ bad) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < E; e += stride) {
        const int32_t s = src[e], d = dst[e];
        if ((uint32_t)s >= (uint32_t)V || (uint32_t)d >= (uint32_t)V) { *bad = 1; keys[e] = 0; vals[e] = (uint32_t)e; continue; }
        const uint64_t major = by_dst ? (uint64_t)d : (uint64_t)s, minor = by_dst ? (uint64_t)s : (uint64_t)d;
        keys[e] = (major << idbits) | minor;
        vals[e] = (uint32_t)e;
    }
}

// per-tile digit histogram: hist[digit * nblocks + block]
__global__ void __launch_bounds__(RS_THREADS) rs_hist(const uint64_t* __restrict__ keys, int64_t n, int shift,
                                                      uint32_t* __restrict__ hist, int nblocks) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * RS_TILE;
#pragma unroll
    for (int it = 0; it < RS_ITEMS; it++) {
        const int64_t k = base + it * RS_THREADS + threadIdx.x;
        if (k < n) atomicAdd(&h[(keys[k] >> shift) & 255], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * nblocks + blockIdx.x] = h[threadIdx.x];
}

// stable scatter: warp w owns keys [w*512, w*512+512) of the tile, walked in batches of 32
__global__ void __launch_bounds__(RS_THREADS) rs_scatter(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                         int64_t n, int shift, const uint32_t* __restrict__ offs, int nblocks,
                                                         uint64_t* __restrict__ okeys, uint32_t* __restrict__ ovals) {
    __shared__ uint32_t cnt[RS_WARPS][256];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (int k = threadIdx.x; k < RS_WARPS * 256; k += RS_THREADS) (&cnt[0][0])[k] = 0;
    __syncthreads();
    const int64_t wbase = (int64_t)blockIdx.x * RS_TILE + (int64_t)w * RS_PER_WARP;
    uint64_t key[RS_ITEMS];
    const unsigned lt = (1u << lane) - 1;
#pragma unroll
    for (int b = 0; b < RS_ITEMS; b++) {
        const int64_t k = wbase + b * 32 + lane;
        const bool live = k < n;
        key[b] = live ? keys[k] : ~0ULL;
        const unsigned act = __ballot_sync(0xffffffffu, live);
        if (live) {
            const uint32_t d = (uint32_t)(key[b] >> shift) & 255;
            const unsigned peers = __match_any_sync(act, d);
            if ((peers & lt) == 0) cnt[w][d] += __popc(peers);  // one leader per digit, no conflict
        }
        __syncwarp();
    }
    __syncthreads();
    {   // digit t: running start for each warp = global offset of (digit, block) + earlier warps
        const int t = threadIdx.x;
        uint32_t run = offs[(size_t)t * nblocks + blockIdx.x];
#pragma unroll
        for (int ww = 0; ww < RS_WARPS; ww++) { const uint32_t c = cnt[ww][t]; cnt[ww][t] = run; run += c; }
    }
    __syncthreads();
#pragma unroll
    for (int b = 0; b < RS_ITEMS; b++) {
        const int64_t k = wbase + b * 32 + lane;
        const bool live = k < n;
        const unsigned act = __ballot_sync(0xffffffffu, live);
        if (live) {
            const uint32_t d = (uint32_t)(key[b] >> shift) & 255;
            const unsigned peers = __match_any_sync(act, d);
            const uint32_t pos = cnt[w][d] + __popc(peers & lt);
            okeys[pos] = key[b];
            ovals[pos] = vals[k];
            __syncwarp(act);
            if ((peers & lt) == 0) cnt[w][d] += __popc(peers);
        }
        __syncwarp();
    }
}

// ---- exclusive scan of uint32 (n < 2^31): tile sums -> single-block scan of sums -> apply
__global__ void __launch_bounds__(RS_THREADS) scan_tile_sums(const uint32_t* __restrict__ in, int64_t n, uint32_t* __restrict__ sums) {
    __shared__ uint32_t ws[RS_WARPS];
    const int64_t base = (int64_t)blockIdx.x * RS_TILE + (int64_t)threadIdx.x * RS_ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (int it = 0; it < RS_ITEMS; it++) if (base + it < n) s += in[base + it];
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) { uint32_t t = 0; for (int k = 0; k < RS_WARPS; k++) t += ws[k]; sums[blockIdx.x] = t; }
}

__global__ void scan_sums(uint32_t* __restrict__ sums, int nb, uint32_t* __restrict__ total) {
    __shared__ uint32_t wsum[32];
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int base = 0; base < nb; base += blockDim.x) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < nb ? sums[i] : 0;
        uint32_t x = v;
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) wsum[wid] = x;
        __syncthreads();
        if (wid == 0) {
            const uint32_t t = lane < nw ? wsum[lane] : 0;
            uint32_t s = t;
            for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, s, o); if (lane >= o) s += y; }
            wsum[lane] = s - t;
        }
        __syncthreads();
        const uint32_t excl = carry + wsum[wid] + (x - v);
        if (i < nb) sums[i] = excl;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0 && total) *total = carry;
}

__global__ void __launch_bounds__(RS_THREADS) scan_apply(const uint32_t* __restrict__ in, int64_t n, const uint32_t* __restrict__ sums,
                                                         uint32_t* __restrict__ out) {
    __shared__ uint32_t ws[RS_WARPS];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int64_t base = (int64_t)blockIdx.x * RS_TILE + (int64_t)threadIdx.x * RS_ITEMS;
    uint32_t v[RS_ITEMS], s = 0;
#pragma unroll
    for (int it = 0; it < RS_ITEMS; it++) { v[it] = base + it < n ? in[base + it] : 0; s += v[it]; }
    uint32_t x = s;
    for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) ws[wid] = x;
    __syncthreads();
    uint32_t wb = 0;
    for (int k = 0; k < wid; k++) wb += ws[k];
    uint32_t run = sums[blockIdx.x] + wb + (x - s);
#pragma unroll
    for (int it = 0; it < RS_ITEMS; it++) { if (base + it < n) out[base + it] = run; run += v[it]; }
}

__global__ void flag_heads(const uint64_t* __restrict__ keys, int64_t n, uint32_t* __restrict__ flag) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride)
        flag[k] = (k == 0 || keys[k] != keys[k - 1]) ? 1u : 0u;
}

// group heads -> compacted CSR body; the head carries the smallest edge index of its group
// (stable sort), i.e. the first parallel edge in input order
__global__ void compact_heads(const uint64_t* __restrict__ keys, const uint32_t* __restrict__ vals, const uint32_t* __restrict__ flag,
                              const uint32_t* __restrict__ pos, int64_t n, int idbits, const float* __restrict__ w,
                              uint64_t* __restrict__ ukeys, int32_t* __restrict__ nbr, float* __restrict__ wout) {
    const uint64_t mask = (1ULL << idbits) - 1;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        if (!flag[k]) continue;
        const uint32_t p = pos[k];
        ukeys[p] = keys[k];
        nbr[p] = (int32_t)(keys[k] & mask);
        wout[p] = w[vals[k]];
    }
}

// off[v] = number of unique keys whose major id < v  (lower bound of v << idbits)
__global__ void offsets_by_search(const uint64_t* __restrict__ ukeys, int64_t m, int32_t V, int idbits, int64_t* __restrict__ off) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v > V) return;
    const uint64_t target = (uint64_t)v << idbits;
    int64_t lo = 0, hi = m;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (ukeys[mid] < target) lo = mid + 1; else hi = mid;
    }
    off[v] = lo;
}

struct Scratch {
    void* p[8] = {};
    ~Scratch() { for (void* q : p) if (q) cudaFree(q); }
};

}  // namespace


// exclusive scan helper (in may alias out); total receives the sum
static int device_scan(ge_ctx* ctx, const uint32_t* in, int64_t n, uint32_t* out, uint32_t* d_sums, uint32_t* d_total) {
    const int nb = (int)((n + RS_TILE - 1) / RS_TILE);
    scan_tile_sums<<<nb, RS_THREADS, 0, ctx->stream>>>(in, n, d_sums);
    scan_sums<<<1, 1024, 0, ctx->stream>>>(d_sums, nb, d_total);
    scan_apply<<<nb, RS_THREADS, 0, ctx->stream>>>(in, n, d_sums, out);
    GE_CUDA(cudaGetLastError());
    return GE_OK;
}

// one direction: returns the CSR in freshly allocated arrays
static int build_direction(ge_ctx* ctx, const int32_t* d_src, const int32_t* d_dst, const float* d_w, int64_t E, int32_t V,
                           int by_dst, int64_t** off_out, int32_t** nbr_out, float** w_out, int64_t* m_out) {
    cudaStream_t st = ctx->stream;
    int idbits = 1;
    while ((1LL << idbits) < V) idbits++;
    const int nb = (int)((E + RS_TILE - 1) / RS_TILE);
    Scratch S;
    uint64_t *k0, *k1;
    uint32_t *v0, *v1, *hist, *sums, *flag;
    int32_t* bad;
    const size_t hist_n = (size_t)256 * nb;
    const int hist_nb = (int)((hist_n + RS_TILE - 1) / RS_TILE);
    const int sums_n = (nb > hist_nb ? nb : hist_nb) + 2;
    GE_CUDA(cudaMalloc(&S.p[0], (size_t)E * 8)); k0 = (uint64_t*)S.p[0];
    GE_CUDA(cudaMalloc(&S.p[1], (size_t)E * 8)); k1 = (uint64_t*)S.p[1];
    GE_CUDA(cudaMalloc(&S.p[2], (size_t)E * 4)); v0 = (uint32_t*)S.p[2];
    GE_CUDA(cudaMalloc(&S.p[3], (size_t)E * 4)); v1 = (uint32_t*)S.p[3];
    GE_CUDA(cudaMalloc(&S.p[4], hist_n * 4)); hist = (uint32_t*)S.p[4];
    GE_CUDA(cudaMalloc(&S.p[5], (size_t)sums_n * 4 + 16)); sums = (uint32_t*)S.p[5];
    GE_CUDA(cudaMalloc(&S.p[6], (size_t)E * 4)); flag = (uint32_t*)S.p[6];
    GE_CUDA(cudaMalloc(&S.p[7], 16)); bad = (int32_t*)S.p[7];
    uint32_t* d_total = (uint32_t*)(bad + 1);
    GE_CUDA(cudaMemsetAsync(bad, 0, 16, st));
    pack_keys<<<ctx->sm_count * 8, 256, 0, st>>>(d_src, d_dst, E, V, idbits, by_dst, k0, v0, bad);
    GE_DBG("pack_keys");
    for (int shift = 0; shift < 2 * idbits; shift += 8) {
        rs_hist<<<nb, RS_THREADS, 0, st>>>(k0, E, shift, hist, nb);
        GE_DBG("rs_hist");
        int r = device_scan(ctx, hist, (int64_t)hist_n, hist, sums, nullptr);
        if (r) return r;
        GE_DBG("scan(hist)");
        rs_scatter<<<nb, RS_THREADS, 0, st>>>(k0, v0, E, shift, hist, nb, k1, v1);
        GE_DBG("rs_scatter");
        std::swap(k0, k1);
        std::swap(v0, v1);
    }
    flag_heads<<<ctx->sm_count * 8, 256, 0, st>>>(k0, E, flag);
    GE_DBG("flag_heads");
    uint32_t* pos = v1;  // the spare value buffer holds the scanned positions
    int r = device_scan(ctx, flag, E, pos, sums, d_total);
    if (r) return r;
    int32_t h_bad[2] = {0, 0};
    GE_CUDA(cudaMemcpyAsync(h_bad, bad, 8, cudaMemcpyDeviceToHost, st));
    GE_CUDA(cudaStreamSynchronize(st));
    if (h_bad[0]) { ctx->err = "edge list: vertex id out of range"; return GE_ERR_GRAPH; }
    const int64_t m = (int64_t)(uint32_t)h_bad[1];
    int64_t* off = nullptr; int32_t* nbr = nullptr; float* w = nullptr;
    GE_CUDA(cudaMalloc(&off, (size_t)(V + 1) * 8));
    if (cudaMalloc(&nbr, (size_t)(m + 1) * 4) != cudaSuccess || cudaMalloc(&w, (size_t)(m + 1) * 4) != cudaSuccess) {
        cudaFree(off); cudaFree(nbr); cudaGetLastError(); ctx->err = "out of device memory for the CSR"; return GE_ERR_NOMEM;
    }
    uint64_t* ukeys = k1;  // spare key buffer
    compact_heads<<<ctx->sm_count * 8, 256, 0, st>>>(k0, v0, flag, pos, E, idbits, d_w, ukeys, nbr, w);
    GE_DBG("compact_heads");
    offsets_by_search<<<(V + 1 + 255) / 256, 256, 0, st>>>(ukeys, m, V, idbits, off);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(st);
    if (e != cudaSuccess) { cudaFree(off); cudaFree(nbr); cudaFree(w); GE_CUDA(e); }
    *off_out = off; *nbr_out = nbr; *w_out = w; *m_out = m;
    return GE_OK;
}

extern "C" int ge_set_graph_edges(ge_ctx* ctx, int32_t V, int64_t E, const int32_t* src, const int32_t* dst, const float* w) {
    if (!ctx) return GE_ERR_INVALID;
    if (V <= 0 || E < 0 || (E > 0 && (!src || !dst || !w))) { ctx->err = "bad edge-list arguments"; return GE_ERR_INVALID; }
    if (E >= 0x7fffffffLL) { ctx->err = "edge list too long (the reference indexes edges with int)"; return GE_ERR_INVALID; }
    cudaSetDevice(ctx->cfg.device);
    ge_free_graph(ctx);
    ge_free_cooc(ctx);
    ge_free_store(ctx);
    cudaStream_t st = ctx->stream;
    Scratch in;
    const size_t n = (size_t)(E > 0 ? E : 1);
    GE_CUDA(cudaMalloc(&in.p[0], n * 4));
    GE_CUDA(cudaMalloc(&in.p[1], n * 4));
    GE_CUDA(cudaMalloc(&in.p[2], n * 4));
    if (E > 0) {
        GE_CUDA(cudaMemcpyAsync(in.p[0], src, n * 4, cudaMemcpyHostToDevice, st));
        GE_CUDA(cudaMemcpyAsync(in.p[1], dst, n * 4, cudaMemcpyHostToDevice, st));
        GE_CUDA(cudaMemcpyAsync(in.p[2], w, n * 4, cudaMemcpyHostToDevice, st));
    }
    GE_CUDA(cudaEventRecord(ctx->ev0, st));
    ctx->V = V;
    if (E == 0) {
        GE_CUDA(cudaMalloc(&ctx->d_out_off, (size_t)(V + 1) * 8));
        GE_CUDA(cudaMalloc(&ctx->d_in_off, (size_t)(V + 1) * 8));
        GE_CUDA(cudaMemsetAsync(ctx->d_out_off, 0, (size_t)(V + 1) * 8, st));
        GE_CUDA(cudaMemsetAsync(ctx->d_in_off, 0, (size_t)(V + 1) * 8, st));
        GE_CUDA(cudaMalloc(&ctx->d_out_nbr, 4)); GE_CUDA(cudaMalloc(&ctx->d_out_w, 4));
        GE_CUDA(cudaMalloc(&ctx->d_in_nbr, 4)); GE_CUDA(cudaMalloc(&ctx->d_in_w, 4));
        ctx->E_out = ctx->E_in = 0;
    } else {
        int r = build_direction(ctx, (int32_t*)in.p[0], (int32_t*)in.p[1], (float*)in.p[2], E, V, 0,
                                &ctx->d_out_off, &ctx->d_out_nbr, &ctx->d_out_w, &ctx->E_out);
        if (r) { ge_free_graph(ctx); return r; }
        r = build_direction(ctx, (int32_t*)in.p[0], (int32_t*)in.p[1], (float*)in.p[2], E, V, 1,
                            &ctx->d_in_off, &ctx->d_in_nbr, &ctx->d_in_w, &ctx->E_in);
        if (r) { ge_free_graph(ctx); return r; }
    }
    GE_CUDA(cudaEventRecord(ctx->ev1, st));
    GE_DBG("flatten");
    int r = ge_graph_finish(ctx);
    if (r) return r;
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->graph_seconds = ms * 1e-3;
    return GE_OK;
}

extern "C" int ge_graph_sizes(ge_ctx* ctx, int64_t* E_out, int64_t* E_in, double* flatten_seconds) {
    if (!ctx) return GE_ERR_INVALID;
    if (!ctx->d_out_off) { ctx->err = "no graph"; return GE_ERR_STATE; }
    if (E_out) *E_out = ctx->E_out;
    if (E_in) *E_in = ctx->E_in;
    if (flatten_seconds) *flatten_seconds = ctx->graph_seconds;
    return GE_OK;
}

extern "C" int ge_graph_export(ge_ctx* ctx, int64_t* out_off, int32_t* out_nbr, float* out_w, int64_t* in_off,
                               int32_t* in_nbr, float* in_w) {
    if (!ctx) return GE_ERR_INVALID;
    if (!ctx->d_out_off) { ctx->err = "no graph"; return GE_ERR_STATE; }
    cudaSetDevice(ctx->cfg.device);
    cudaStream_t st = ctx->stream;
    const size_t ob = (size_t)(ctx->V + 1) * 8;
    if (out_off) GE_CUDA(cudaMemcpyAsync(out_off, ctx->d_out_off, ob, cudaMemcpyDeviceToHost, st));
    if (in_off) GE_CUDA(cudaMemcpyAsync(in_off, ctx->d_in_off, ob, cudaMemcpyDeviceToHost, st));
    if (out_nbr && ctx->E_out) GE_CUDA(cudaMemcpyAsync(out_nbr, ctx->d_out_nbr, (size_t)ctx->E_out * 4, cudaMemcpyDeviceToHost, st));
    if (out_w && ctx->E_out) GE_CUDA(cudaMemcpyAsync(out_w, ctx->d_out_w, (size_t)ctx->E_out * 4, cudaMemcpyDeviceToHost, st));
    if (in_nbr && ctx->E_in) GE_CUDA(cudaMemcpyAsync(in_nbr, ctx->d_in_nbr, (size_t)ctx->E_in * 4, cudaMemcpyDeviceToHost, st));
    if (in_w && ctx->E_in) GE_CUDA(cudaMemcpyAsync(in_w, ctx->d_in_w, (size_t)ctx->E_in * 4, cudaMemcpyDeviceToHost, st));
    GE_CUDA(cudaStreamSynchronize(st));
    return GE_OK;
}

// ---- ge_flatten_edges_static: one direction of the flattening above with NO allocation ------------
// No counterpart in the reference; added for callers that cannot
// call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels as
// build_direction; the sort buffers live in ONE file-scope device array that every GPU loading this
// library holds until exit, inputs and outputs are the caller's device buffers, nothing is synchronised.
// Every call shares the array: two calls that overlap on the GPU (different streams / handles) race.
namespace {

constexpr int64_t ST_MAX_E = GE_STATIC_MAX_EDGES;
constexpr size_t ST_NB = (size_t)((ST_MAX_E + RS_TILE - 1) / RS_TILE);           // radix tiles at the maximum
constexpr size_t ST_HIST_N = 256 * ST_NB;
constexpr size_t ST_HIST_NB = (ST_HIST_N + RS_TILE - 1) / RS_TILE;
constexpr size_t st_up(size_t b) { return (b + 255) & ~(size_t)255; }
constexpr size_t ST_OFF_K0 = 0;
constexpr size_t ST_OFF_K1 = ST_OFF_K0 + st_up((size_t)ST_MAX_E * 8);
constexpr size_t ST_OFF_V0 = ST_OFF_K1 + st_up((size_t)ST_MAX_E * 8);
constexpr size_t ST_OFF_V1 = ST_OFF_V0 + st_up((size_t)ST_MAX_E * 4);
constexpr size_t ST_OFF_FLAG = ST_OFF_V1 + st_up((size_t)ST_MAX_E * 4);
constexpr size_t ST_OFF_HIST = ST_OFF_FLAG + st_up((size_t)ST_MAX_E * 4);
constexpr size_t ST_OFF_SUMS = ST_OFF_HIST + st_up(ST_HIST_N * 4);
constexpr size_t ST_OFF_MISC = ST_OFF_SUMS + st_up(((ST_NB > ST_HIST_NB ? ST_NB : ST_HIST_NB) + 2) * 4 + 16);
constexpr size_t ST_BYTES = ST_OFF_MISC + 256;  // misc: [0] bad-id flag, [1] number of unique keys

static_assert(ST_BYTES == (size_t)GE_STATIC_SCRATCH_BYTES, "GE_STATIC_SCRATCH_BYTES in graphemb.h is stale");

static __device__ __align__(256) unsigned char g_flatten_scratch[ST_BYTES];

constexpr int ST_MAX_DEVICES = 64;
unsigned char* g_flatten_base[ST_MAX_DEVICES] = {};  // DEVICE address of g_flatten_scratch, per device ordinal
std::mutex g_flatten_setup;

// offsets as offsets_by_search, the unique count read from the device; thread 0 publishes {count, bad flag}
__global__ void offsets_by_search_static(const uint64_t* __restrict__ ukeys, const int32_t* __restrict__ misc, int32_t V,
                                         int idbits, int64_t* __restrict__ off, int64_t* __restrict__ info) {
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t m = (int64_t)(uint32_t)misc[1];
    if (v == 0) { info[0] = m; info[1] = misc[0]; }
    if (v > V) return;
    const uint64_t target = (uint64_t)v << idbits;
    int64_t lo = 0, hi = m;
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (ukeys[mid] < target) lo = mid + 1; else hi = mid;
    }
    off[v] = lo;
}

}  // namespace

extern "C" int ge_flatten_edges_static(ge_ctx* ctx, int32_t V, int64_t E, int32_t by_dst, const int32_t* d_src,
                                       const int32_t* d_dst, const float* d_w, int64_t* d_off, int32_t* d_nbr,
                                       float* d_wout, int64_t* d_info) {
    if (!ctx) return GE_ERR_INVALID;
    if (V <= 0 || E < 0 || !d_off || !d_info || (E > 0 && (!d_src || !d_dst || !d_w || !d_nbr || !d_wout))) {
        ctx->err = "bad edge-list arguments";
        return GE_ERR_INVALID;
    }
    if (E > ST_MAX_E) {
        ctx->err = "ge_flatten_edges_static: E exceeds GE_STATIC_MAX_EDGES (" + std::to_string(ST_MAX_E) +
                   " edges fit the static scratch); use ge_set_graph_edges";
        return GE_ERR_INVALID;
    }
    const int dev = ctx->cfg.device;
    if (dev < 0 || dev >= ST_MAX_DEVICES) { ctx->err = "device ordinal beyond the static scratch table"; return GE_ERR_INVALID; }
    GE_CUDA(cudaSetDevice(dev));
    {
        std::lock_guard<std::mutex> lock(g_flatten_setup);
        if (!g_flatten_base[dev]) {
            // set-up, once per device, with that device current: in host code the array's name is only a
            // stand-in for the symbol — the kernels need the address the runtime resolves for THIS device
            void* p = nullptr;
            GE_CUDA(cudaGetSymbolAddress((void**)&p, g_flatten_scratch));
            if (!p) { ctx->err = "cudaGetSymbolAddress returned no address for the static scratch"; return GE_ERR_CUDA; }
            g_flatten_base[dev] = (unsigned char*)p;
        }
    }
    unsigned char* base = g_flatten_base[dev];
    cudaStream_t st = ctx->stream;
    int32_t* misc = (int32_t*)(base + ST_OFF_MISC);
    if (E == 0) {  // also the warm-up call: resolves the address, touches no sort buffer
        GE_CUDA(cudaMemsetAsync(d_off, 0, (size_t)(V + 1) * 8, st));
        GE_CUDA(cudaMemsetAsync(d_info, 0, 16, st));
        return GE_OK;
    }
    uint64_t *k0 = (uint64_t*)(base + ST_OFF_K0), *k1 = (uint64_t*)(base + ST_OFF_K1);
    uint32_t *v0 = (uint32_t*)(base + ST_OFF_V0), *v1 = (uint32_t*)(base + ST_OFF_V1);
    uint32_t *flag = (uint32_t*)(base + ST_OFF_FLAG), *hist = (uint32_t*)(base + ST_OFF_HIST), *sums = (uint32_t*)(base + ST_OFF_SUMS);
    int idbits = 1;
    while ((1LL << idbits) < V) idbits++;
    const int nb = (int)((E + RS_TILE - 1) / RS_TILE);
    const size_t hist_n = (size_t)256 * nb;
    // the array keeps the last call's contents: the bad-id flag and the unique count must start at zero
    // (keys, values, flags, histograms and tile sums are fully rewritten for the first E / nb entries)
    GE_CUDA(cudaMemsetAsync(misc, 0, 256, st));
    pack_keys<<<ctx->sm_count * 8, 256, 0, st>>>(d_src, d_dst, E, V, idbits, by_dst ? 1 : 0, k0, v0, misc);
    for (int shift = 0; shift < 2 * idbits; shift += 8) {
        rs_hist<<<nb, RS_THREADS, 0, st>>>(k0, E, shift, hist, nb);
        int r = device_scan(ctx, hist, (int64_t)hist_n, hist, sums, nullptr);
        if (r) return r;
        rs_scatter<<<nb, RS_THREADS, 0, st>>>(k0, v0, E, shift, hist, nb, k1, v1);
        std::swap(k0, k1);
        std::swap(v0, v1);
    }
    flag_heads<<<ctx->sm_count * 8, 256, 0, st>>>(k0, E, flag);
    uint32_t* pos = v1;
    int r = device_scan(ctx, flag, E, pos, sums, (uint32_t*)(misc + 1));
    if (r) return r;
    uint64_t* ukeys = k1;
    compact_heads<<<ctx->sm_count * 8, 256, 0, st>>>(k0, v0, flag, pos, E, idbits, d_w, ukeys, d_nbr, d_wout);
    offsets_by_search_static<<<(V + 1 + 255) / 256, 256, 0, st>>>(ukeys, misc, V, idbits, d_off, d_info);
    GE_CUDA(cudaGetLastError());
    return GE_OK;
}
