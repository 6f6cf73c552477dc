// dist.cu — multi-GPU trainer: nb x nb block grid with ring rotation of the C blocks
// (K7 block bucketing).  One process per GPU; the only exchange step is one
// ncclSend/ncclRecv of a C block (rows + AdaGrad state + bias, one contiguous slab) per
// sub-epoch, plus one all-reduce of the fp64 loss per epoch.
//
// Ids are cut into nb contiguous blocks of Bs = ceil(V/nb) ids.  Rank r (of p, p | nb, m = nb/p)
// owns row-blocks r*m .. r*m+m-1: their W rows never move.  Think of nb virtual ranks; virtual
// rank v owns row-block v and in sub-epoch s updates block (v, (v+s) mod nb).  No two virtual
// ranks share a W or a C block within a sub-epoch, so cross-GPU updates are race-free; inside a
// block the update kernel is the single-GPU one (Hogwild with atomic row reductions).
// Rank r therefore needs the cyclic window of m C blocks starting at (r*m+s) mod nb; moving from
// s to s+1 it hands the window's first block to rank r-1 and receives a new last block from
// rank r+1.  After nb sub-epochs every C block is home again.
//
// Replaces the reference's thread slicing of the shuffled nonzeros (J/opt/Optimizer.java:59-63,
// J/opt/grad/Adagrad.java:47-49), which has no multi-device counterpart.
//
// Rotation data plane (world > 1), default GE_DIST_ROT=ce: every column block is cut in two
// halves; as soon as the update of the window's first block has finished a half, that half is
// pushed into the neighbour's spare ring slot by the COPY ENGINE (cudaMemcpyAsync on a peer
// mapping obtained through CUDA IPC, no SMs involved) while the SMs update the other half / the
// other window blocks.  Hand-shake = two flag words per half in each rank's memory, written by the
// neighbours' copy streams with 4-byte DMA copies and awaited with stream memory operations
// (cuStreamWaitValue32): `ready` (rank+1 -> me: half p of transfer q has landed) and `free`
// (rank-1 -> me: the half-slot I am about to overwrite has been sent on).  GE_DIST_ROT=nccl keeps
// the round-1 plane (whole block, ncclSend/ncclRecv in the compute stream), GE_DIST_ROT=pipe the
// half-block NCCL variant.
#include "common.cuh"
#include "shm_group.h"

#include <dlfcn.h>

#include <cstdio>
#include <cstring>

namespace {

// ---- NCCL through dlopen (torch already has libnccl.so.2 loaded; no link-time dependency) ----
typedef struct ncclComm* ncclComm_t;
struct NcclId { char internal[GE_NCCL_ID_BYTES]; };
enum { NCCL_UINT8 = 1, NCCL_INT64 = 4, NCCL_FLOAT64 = 8, NCCL_SUM = 0, NCCL_MAX = 2 };
struct NcclApi {
    void* h = nullptr;
    int (*GetUniqueId)(NcclId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, NcclId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool load(std::string& err) {
        if (h) return true;
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) { h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
        if (!h) { err = std::string("cannot load libnccl: ") + dlerror(); return false; }
#define SYM(f, s) f = reinterpret_cast<decltype(f)>(dlsym(h, s)); if (!f) { err = std::string("libnccl lacks ") + s; return false; }
        SYM(GetUniqueId, "ncclGetUniqueId") SYM(CommInitRank, "ncclCommInitRank") SYM(CommDestroy, "ncclCommDestroy")
        SYM(Send, "ncclSend") SYM(Recv, "ncclRecv") SYM(AllReduce, "ncclAllReduce")
        SYM(AllGather, "ncclAllGather") SYM(Broadcast, "ncclBroadcast")
        SYM(GroupStart, "ncclGroupStart") SYM(GroupEnd, "ncclGroupEnd") SYM(GetErrorString, "ncclGetErrorString")
#undef SYM
        return true;
    }
};
NcclApi g_nccl;

}  // namespace

enum { ROT_NCCL = 0, ROT_PIPE = 1, ROT_CE = 2 };
constexpr uint32_t SEQTAB = 1u << 20;   // sub-epochs a prepared trainer can run (131 072 epochs of 8)
typedef int (*WaitValue32Fn)(cudaStream_t, unsigned long long, uint32_t, unsigned int);

struct DistState {
    int rank = 0, world = 1;
    ncclComm_t comm = nullptr;
    int nb = 0, m = 0;            // blocks per side, blocks per rank
    int32_t Bs = 0;               // ids per block
    int32_t row_base = 0, row_cnt = 0;  // first (virtual) id of the rank's rows; REAL rows it owns
    ShmGroup* shm = nullptr;          // single-node host plumbing (shm_group.h); null = NCCL carries it
    unsigned char id[GE_NCCL_ID_BYTES] = {};  // the job's unique id, kept for a late ncclCommInitRank
    int32_t layout = 0;               // 0 = contiguous blocks, 1 = cyclic (common.cuh ge_vid)
    int32_t lay_nb() const { return layout ? nb : 0; }
    int64_t nnz_global = 0;
    std::vector<int64_t> bucket_off;  // m*nb + 1 offsets into d_nzb, bucket = local_rowblock*nb + colblock
    NzEntry* d_nzb = nullptr;         // bucketed nonzeros, block-local indices
    float* d_Cring = nullptr;         // (m+1) slots x Bs rows x S floats (m slots when world == 1)
    int P = 1;                        // parts per column block (2 when world > 1: the rotation is pipelined
                                      // half a block at a time behind the compute), Hs = rows in part 0
    int32_t Hs = 0;
    cudaStream_t cs = nullptr;        // high-priority stream the ncclSend/ncclRecv pairs run on
    cudaEvent_t ev_done[2] = {nullptr, nullptr}, ev_recv[2] = {nullptr, nullptr};
    unsigned long long* d_queue = nullptr;  // one batch counter per update launch of an epoch
    int head = 0;                     // slot of the window's first block
    int sub = 0;                      // sub-epochs done in the current rotation (block ids derive from it)
    double comm_seconds = 0;
    // --- copy-engine rotation (GE_DIST_ROT=ce) ---
    int rot = ROT_NCCL;               // data plane in use
    float* peer_ring = nullptr;       // rank-1's ring allocation, mapped through CUDA IPC
    float* next_ring = nullptr;       // rank+1's ring allocation (only its flag words are written); == peer_ring when world == 2
    uint32_t* flags = nullptr;        // local flag words, in the tail of the ring allocation: ready[0..1], free[2..3]
    uint32_t* prev_flags = nullptr;   // rank-1's flags (I write its ready[p])
    uint32_t* next_flags = nullptr;   // rank+1's flags (I write its free[p])
    uint32_t* seqtab = nullptr;       // seqtab[k] = k: source of the 4-byte flag copies
    uint32_t q = 0;                   // transfers issued per half so far (= sub-epochs done)
    std::vector<cudaEvent_t> xfer_ev; // 2 events per transfer of an epoch (copy alone, no waits)
    // a large part is cut into `split` chunks pushed from as many streams, so that several copy engines
    // carry it (one engine alone moved the d=300 blocks of configs[4] at 488 GB/s of NVLink's 900)
    int split = 1;
    cudaStream_t cs_extra[7] = {};
    cudaEvent_t ev_go = nullptr, ev_chunk[7] = {};
    std::vector<int32_t> bucket_hot;  // bucket has entries with damping exponents -> HOT kernel variant
    std::vector<int64_t> bucket_cap;  // warps the bucket's launch may use (train.cu ge_warp_cap); 0 = no cap
    cudaEvent_t c0 = nullptr, c1 = nullptr;
};

void ge_dist_free(ge_ctx* ctx) {
    DistState* d = ctx->dist;
    if (!d) return;
    cudaFree(d->d_nzb);
    cudaFree(d->d_Cring);
    cudaFree(d->d_queue);
    if (d->next_ring && d->next_ring != d->peer_ring) cudaIpcCloseMemHandle(d->next_ring);
    if (d->peer_ring) cudaIpcCloseMemHandle(d->peer_ring);
    cudaFree(d->seqtab);
    for (cudaEvent_t e : d->xfer_ev) cudaEventDestroy(e);
    for (int p = 0; p < 2; p++) { if (d->ev_done[p]) cudaEventDestroy(d->ev_done[p]); if (d->ev_recv[p]) cudaEventDestroy(d->ev_recv[p]); }
    if (d->cs) cudaStreamDestroy(d->cs);
    for (int c = 0; c < 7; c++) { if (d->cs_extra[c]) cudaStreamDestroy(d->cs_extra[c]); if (d->ev_chunk[c]) cudaEventDestroy(d->ev_chunk[c]); }
    if (d->ev_go) cudaEventDestroy(d->ev_go);
    if (d->c0) cudaEventDestroy(d->c0);
    if (d->c1) cudaEventDestroy(d->c1);
    if (d->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(d->comm);
    if (d->shm) { d->shm->close(); delete d->shm; }
    delete d;
    ctx->dist = nullptr;
}

namespace {

#define GE_NCCL(call)                                                                    \
    do {                                                                                 \
        int r__ = (call);                                                                \
        if (r__ != 0) {                                                                  \
            ctx->err = std::string(#call) + ": " + g_nccl.GetErrorString(r__);           \
            return GE_ERR_DIST;                                                          \
        }                                                                                \
    } while (0)

// ---- host-level collectives of the ranks: shared memory on one node, else NCCL through a device word
int ensure_comm(ge_ctx* ctx) {   // collective: every rank calls it at the same point
    DistState* d = ctx->dist;
    if (d->comm || d->world == 1) return GE_OK;
    if (!g_nccl.load(ctx->err)) return GE_ERR_DIST;
    NcclId id;
    memcpy(&id, d->id, GE_NCCL_ID_BYTES);
    const auto t0 = std::chrono::steady_clock::now();
    GE_NCCL(g_nccl.CommInitRank(&d->comm, d->world, id, d->rank));
    if (getenv("GE_DEBUG"))
        fprintf(stderr, "[dist init] ncclCommInitRank %.1f ms\n", std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    return GE_OK;
}

int plumb_allgather(ge_ctx* ctx, const void* mine, void* all, size_t n) {
    DistState* d = ctx->dist;
    if (d->world == 1) { memcpy(all, mine, n); return GE_OK; }
    if (d->shm) return d->shm->allgather(mine, all, n, ctx->err) ? GE_OK : GE_ERR_DIST;
    cudaStream_t st = ctx->stream;
    unsigned char* dv = nullptr;
    GE_CUDA(cudaMalloc(&dv, n * (size_t)(d->world + 1)));
    cudaError_t e = cudaMemcpyAsync(dv + n * (size_t)d->world, mine, n, cudaMemcpyHostToDevice, st);
    int nr = 0;
    if (e == cudaSuccess) nr = g_nccl.AllGather(dv + n * (size_t)d->world, dv, n, NCCL_UINT8, d->comm, st);
    if (e == cudaSuccess && !nr) e = cudaMemcpyAsync(all, dv, n * (size_t)d->world, cudaMemcpyDeviceToHost, st);
    if (e == cudaSuccess && !nr) e = cudaStreamSynchronize(st);
    cudaFree(dv);
    GE_CUDA(e);
    GE_NCCL(nr);
    return GE_OK;
}

// v[0..n) <- sum (op 0) or max (op 1) over the ranks, n <= 8
int plumb_allreduce(ge_ctx* ctx, double* v, int n, int op) {
    DistState* d = ctx->dist;
    if (d->world == 1) return GE_OK;
    double all[8 * 64];
    if (n > 8 || d->world > 64) { ctx->err = "plumb_allreduce: too large"; return GE_ERR_INVALID; }
    int r = plumb_allgather(ctx, v, all, sizeof(double) * (size_t)n);
    if (r) return r;
    for (int k = 0; k < n; k++) {
        double a = all[k];
        for (int q = 1; q < d->world; q++) { const double b = all[(size_t)q * n + k]; a = op ? (b > a ? b : a) : a + b; }
        v[k] = a;   // every rank folds in rank order: identical bits everywhere
    }
    return GE_OK;
}

int plumb_barrier(ge_ctx* ctx) {
    double z = 0;
    return plumb_allreduce(ctx, &z, 1, 0);
}

__global__ void bucket_count(const int32_t* __restrict__ I, const int32_t* __restrict__ J, int64_t nnz,
                             int32_t row_base, int32_t Bs, int32_t nb, int32_t P, int32_t Hs,
                             unsigned long long* __restrict__ cnt) {
    __shared__ unsigned int local[1024];
    for (int k = threadIdx.x; k < 1024; k += blockDim.x) local[k] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
        const int cb = J[k] / Bs;
        const int part = (P > 1 && J[k] - cb * Bs >= Hs) ? 1 : 0;
        const int b = (((I[k] - row_base) / Bs) * nb + cb) * P + part;
        atomicAdd(&local[b], 1u);
    }
    __syncthreads();
    for (int k = threadIdx.x; k < 1024; k += blockDim.x)
        if (local[k]) atomicAdd(&cnt[k], (unsigned long long)local[k]);
}

// scatter into buckets with block-local indices; warp-aggregated cursor claims
__global__ void bucket_scatter(const int32_t* __restrict__ I, const int32_t* __restrict__ J,
                               const float* __restrict__ X, int64_t nnz, int32_t row_base, int32_t Bs, int32_t nb,
                               int32_t P, int32_t Hs, int32_t method, double xmax, unsigned long long* __restrict__ cursor,
                               const int32_t* __restrict__ cnt_i, const int32_t* __restrict__ cnt_j,
                               const int32_t* __restrict__ hot_cnt, int32_t m, int32_t keep_i, int32_t keep_j,
                               int32_t* __restrict__ any_hot, NzEntry* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
        const int32_t i = I[k] - row_base, j = J[k];
        const int cb = j / Bs;
        const int part = (P > 1 && j - cb * Bs >= Hs) ? 1 : 0;
        const int b = ((i / Bs) * nb + cb) * P + part;
        const float x = X[k];
        NzEntry e;
        e.i = i;             // W index: local to the rank's row range
        e.j = j - cb * Bs;   // C index: local to the column block
        // damping exponents of hubs (train.cu ge_hot_count, common.cuh ge_damp_bits): expected occurrences inside THIS bucket — a row's
        // nonzeros spread over nb column blocks, a column's over the rank's m row blocks
        const int32_t ki = ge_damp_bits(cnt_i[i] / (nb * P), hot_cnt[b]), kj = ge_damp_bits(cnt_j[j] / m, hot_cnt[b]);
        if (ki) { e.i |= ki << GE_ID_BITS; any_hot[b] = 1; }
        if (kj) { e.j |= kj << GE_ID_BITS; any_hot[b] = 1; }
        if (cnt_i[i] >= keep_i) e.i |= GE_ID_KEEP;  // L2-resident set (train.cu ge_keep_counts)
        if (cnt_j[j] >= keep_j) e.j |= GE_ID_KEEP;
        if (method == GE_METHOD_GLOVE) {
            e.a = (float)log((double)x);
            e.b = ((double)x > xmax) ? 1.0f : (float)pow((double)x / xmax, 0.75);
        } else {
            e.a = (float)log((double)(x / (1.0f - x)));
            e.b = x;
        }
        const unsigned long long at = atomicAdd(&cursor[b], 1ULL);
        out[at] = e;
    }
}

// per bucket: nonzeros that touch an id damped in that bucket's (uncapped) launch, estimated like
// bucket_scatter does — a row's count spread over nb * P buckets, a column's over the m row blocks.
// side 0: cnt over the rank's m * Bs rows; side 1: cnt over all nb * Bs columns.
__global__ void bucket_hot_mass(const int32_t* __restrict__ cnt, int64_t n, int side, int32_t Bs, int32_t nb, int32_t P,
                                int32_t Hs, int32_t m, const int32_t* __restrict__ hot, unsigned long long* __restrict__ mass) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const int32_t c = cnt[k];
        if (c == 0) continue;
        const int blk = (int)(k / Bs);
        if (side == 0) {
            const int32_t est = c / (nb * P);
            for (int q = 0; q < nb * P; q++) { const int b = blk * nb * P + q; if (est >= hot[b]) atomicAdd(mass + b, (unsigned long long)est); }
        } else {
            const int32_t est = c / m;
            const int part = (P > 1 && k - (int64_t)blk * Bs >= Hs) ? 1 : 0;
            for (int q = 0; q < m; q++) { const int b = (q * nb + blk) * P + part; if (est >= hot[b]) atomicAdd(mass + b, (unsigned long long)est); }
        }
    }
}

// largest occurrence count per row block (P = 1) or per half column block (input of ge_warp_cap)
__global__ void segment_max(const int32_t* __restrict__ cnt, int64_t n, int32_t Bs, int32_t P, int32_t Hs,
                            int32_t* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const int32_t c = cnt[k];
        const int blk = (int)(k / Bs);
        const int seg = blk * P + ((P > 1 && k - (int64_t)blk * Bs >= Hs) ? 1 : 0);
        if (c > out[seg]) atomicMax(out + seg, c);   // the plain read only filters
    }
}

// COO ids -> virtual ids (in place; the arrays are dropped after bucketing), and the check that every row
// belongs to this rank: a foreign row would index outside the bucket table
__global__ void to_virtual_ids(int32_t* __restrict__ I, int32_t* __restrict__ J, int64_t nnz, int32_t lay_nb, int32_t Bs,
                               int32_t first, int32_t count, int32_t* __restrict__ bad) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
        const int32_t i = ge_vid(I[k], lay_nb, Bs), j = ge_vid(J[k], lay_nb, Bs);
        if (i < first || i >= first + count) *bad = 1;
        I[k] = i; J[k] = j;
    }
}

// rows of block b that are real vertices (the rest of its Bs slab rows is padding)
__host__ __device__ inline int64_t block_rows(int64_t b, int32_t nb, int32_t Bs, int32_t V, int32_t cyclic) {
    int64_t n = cyclic ? (b < V ? ((int64_t)V - b + nb - 1) / nb : 0) : (int64_t)V - b * Bs;
    if (n < 0) n = 0;
    return n > Bs ? Bs : n;
}

// real ids of the slab rows, block by block (padding skipped): the order ge_dist_extract emits
__global__ void local_ids(int32_t* __restrict__ out, int32_t m, int32_t rb0, int32_t nb, int32_t Bs, int32_t V, int32_t cyclic) {
    const int64_t total = (int64_t)m * Bs;
    for (int64_t l = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; l < total; l += (int64_t)gridDim.x * blockDim.x) {
        const int32_t k = (int32_t)(l / Bs), t = (int32_t)(l % Bs), rb = rb0 + k;
        const int64_t id = cyclic ? (int64_t)rb + (int64_t)nb * t : (int64_t)rb * Bs + t;
        if (id >= V) continue;
        int64_t before = 0;
        for (int q = 0; q < k; q++) before += block_rows(rb0 + q, nb, Bs, V, cyclic);
        out[before + t] = (int32_t)id;
    }
}

struct IpcPair { cudaIpcMemHandle_t ring; int32_t ok; int32_t pad[15]; };

void dist_unmap_peers(DistState* d) {
    if (d->next_ring && d->next_ring != d->peer_ring) cudaIpcCloseMemHandle(d->next_ring);
    if (d->peer_ring) cudaIpcCloseMemHandle(d->peer_ring);
    d->peer_ring = d->next_ring = nullptr;
    d->prev_flags = d->next_flags = nullptr;
}

WaitValue32Fn g_wait32 = nullptr;

// Exchange CUDA IPC handles of every rank's ring and flag words (one ncclAllGather) and map the
// neighbours'.  Falls back to the NCCL plane when IPC or stream memory operations are unavailable.
int dist_map_peers(ge_ctx* ctx) {
    DistState* d = ctx->dist;
    cudaStream_t st = ctx->stream;
    dist_unmap_peers(d);
    if (!d->seqtab) {
        GE_CUDA(cudaMalloc(&d->seqtab, (size_t)SEQTAB * 4));
        std::vector<uint32_t> h(SEQTAB);
        for (uint32_t k = 0; k < SEQTAB; k++) h[k] = k;
        GE_CUDA(cudaMemcpyAsync(d->seqtab, h.data(), (size_t)SEQTAB * 4, cudaMemcpyHostToDevice, st));
        GE_CUDA(cudaStreamSynchronize(st));
    }
    d->q = 0;
    bool ok = true;
    if (!g_wait32) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuStreamWaitValue32", &fn, cudaEnableDefault, &qr) != cudaSuccess || !fn) { cudaGetLastError(); ok = false; }
        else g_wait32 = reinterpret_cast<WaitValue32Fn>(fn);
    }
    IpcPair mine{};
    if (ok && cudaIpcGetMemHandle(&mine.ring, d->d_Cring) != cudaSuccess) { cudaGetLastError(); ok = false; }
    mine.ok = ok ? 1 : 0;
    // every rank takes part in the gather even when its own set-up failed
    std::vector<IpcPair> all(d->world);
    {
        int r = plumb_allgather(ctx, &mine, all.data(), sizeof(IpcPair));
        if (r) return r;
    }
    for (int k = 0; k < d->world; k++) ok = ok && all[k].ok == 1;
    const int prev = (d->rank + d->world - 1) % d->world, next = (d->rank + 1) % d->world;
    if (ok) {
        // the flag words sit behind the last slot of every ring allocation (same geometry on all ranks)
        const size_t ring_floats = (size_t)(d->m + 1) * d->Bs * ctx->S;
        void *pr = nullptr, *nr2 = nullptr;
        if (cudaIpcOpenMemHandle(&pr, all[prev].ring, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = false;
        if (ok && next != prev && cudaIpcOpenMemHandle(&nr2, all[next].ring, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) ok = false;
        if (!ok) {
            cudaGetLastError();
            if (pr) cudaIpcCloseMemHandle(pr);
        } else {
            d->peer_ring = (float*)pr;
            d->next_ring = next != prev ? (float*)nr2 : (float*)pr;
            d->prev_flags = reinterpret_cast<uint32_t*>(d->peer_ring + ring_floats);
            d->next_flags = reinterpret_cast<uint32_t*>(d->next_ring + ring_floats);
        }
    }
    // agree on the plane: one rank that could not map its neighbour sends everybody to NCCL
    {
        double hv = ok ? 0.0 : 1.0;
        int r = plumb_allreduce(ctx, &hv, 1, 0);
        if (r) return r;
        if (hv != 0.0) {
            if (getenv("GE_DEBUG")) fprintf(stderr, "[graphemb] rank %d: CUDA IPC / stream memory operations unavailable, rotation falls back to NCCL send/recv\n", d->rank);
            d->rot = ROT_PIPE;
            r = ensure_comm(ctx);   // every rank got the same sum: a collective call
            if (r) return r;
        }
    }
    if (d->rot == ROT_CE && d->xfer_ev.empty()) {
        d->xfer_ev.resize((size_t)4 * d->nb);
        for (auto& ev : d->xfer_ev) GE_CUDA(cudaEventCreate(&ev));
    }
    return GE_OK;
}

}  // namespace

extern "C" int ge_dist_unique_id(void* id128) {
    std::string err;
    if (!id128 || !g_nccl.load(err)) return GE_ERR_DIST;
    NcclId id;
    if (g_nccl.GetUniqueId(&id) != 0) return GE_ERR_DIST;
    memcpy(id128, &id, GE_NCCL_ID_BYTES);
    return GE_OK;
}

extern "C" int ge_dist_init(ge_ctx* ctx, int32_t rank, int32_t world, const void* id128) {
    if (!ctx) return GE_ERR_INVALID;
    if (world < 1 || rank < 0 || rank >= world) { ctx->err = "bad rank/world"; return GE_ERR_INVALID; }
    cudaSetDevice(ctx->cfg.device);
    ge_dist_free(ctx);
    DistState* d = new DistState();
    d->rank = rank; d->world = world;
    ctx->dist = d;
    if (world > 1) {
        if (!id128) { ctx->err = "ge_dist_init: unique id required for world > 1"; return GE_ERR_INVALID; }
        memcpy(d->id, id128, GE_NCCL_ID_BYTES);
        // Plumbing.  With the copy-engine rotation NCCL would carry a few hundred bytes per epoch, and bringing it
        // up costs 2.1 s (ncclCommInitRank, with or without NVLS) + 1.0 s (first collective) at 8 ranks: the ranks
        // of one node meet in a shared-memory segment instead (shm_group.h) and NCCL is brought up only when a
        // plane that moves data with it is selected (GE_DIST_ROT=nccl|pipe, IPC unavailable) or on request
        // (GE_DIST_PLUMBING=nccl).
        const char* rote = getenv("GE_DIST_ROT");
        const char* plumb = getenv("GE_DIST_PLUMBING");
        const bool want_nccl_data = (rote && (!strcmp(rote, "nccl") || !strcmp(rote, "pipe"))) || (!rote && getenv("GE_DIST_PIPE") && atoi(getenv("GE_DIST_PIPE")) == 1);
        if (!(plumb && !strcmp(plumb, "nccl"))) {
            d->shm = new ShmGroup();
            std::string why;
            if (!d->shm->open(id128, rank, world, why)) {   // fails on every rank or on none (the attach barrier)
                if (getenv("GE_DEBUG")) fprintf(stderr, "[graphemb] rank %d: %s; plumbing falls back to NCCL\n", rank, why.c_str());
                d->shm->close(); delete d->shm; d->shm = nullptr;
            }
        }
        if (!d->shm || want_nccl_data) { int r = ensure_comm(ctx); if (r) return r; }
        int lo = 0, hi = 0;
        cudaDeviceGetStreamPriorityRange(&lo, &hi);
        GE_CUDA(cudaStreamCreateWithPriority(&d->cs, cudaStreamNonBlocking, hi));
        {
            const char* e = getenv("GE_DIST_CE_SPLIT");
            d->split = e ? atoi(e) : 4;
            if (d->split < 1) d->split = 1;
            if (d->split > 8) d->split = 8;
            GE_CUDA(cudaEventCreateWithFlags(&d->ev_go, cudaEventDisableTiming));
            for (int x = 0; x + 1 < d->split; x++) {
                GE_CUDA(cudaStreamCreateWithPriority(&d->cs_extra[x], cudaStreamNonBlocking, hi));
                GE_CUDA(cudaEventCreateWithFlags(&d->ev_chunk[x], cudaEventDisableTiming));
            }
        }
        for (int p = 0; p < 2; p++) {
            GE_CUDA(cudaEventCreateWithFlags(&d->ev_done[p], cudaEventDisableTiming));
            GE_CUDA(cudaEventCreateWithFlags(&d->ev_recv[p], cudaEventDisableTiming));
        }
    }
    GE_CUDA(cudaEventCreate(&d->c0));
    GE_CUDA(cudaEventCreate(&d->c1));
    return GE_OK;
}

// rank's id range for an nb-block grid over V ids
extern "C" int ge_dist_row_range(ge_ctx* ctx, int32_t V, int32_t nblocks, int32_t* first, int32_t* count) {
    if (!ctx || !ctx->dist) { if (ctx) ctx->err = "ge_dist_init first"; return GE_ERR_STATE; }
    DistState* d = ctx->dist;
    if (nblocks < 1 || nblocks > 32 || nblocks % d->world) { ctx->err = "nblocks must be a multiple of the world size (<= 32)"; return GE_ERR_INVALID; }
    const int32_t Bs = (V + nblocks - 1) / nblocks, m = nblocks / d->world;
    int64_t lo = (int64_t)d->rank * m * Bs, hi = lo + (int64_t)m * Bs;
    if (!d->layout) {   // cyclic: a range of VIRTUAL ids, whole blocks (padding included)
        if (lo > V) lo = V;
        if (hi > V) hi = V;
    }
    if (first) *first = (int32_t)lo;
    if (count) *count = (int32_t)(hi - lo);
    return GE_OK;
}

extern "C" int ge_dist_prepare(ge_ctx* ctx, int32_t nblocks) {
    if (!ctx || !ctx->dist) { if (ctx) ctx->err = "ge_dist_init first"; return GE_ERR_STATE; }
    if (!ctx->have_cooc) { ctx->err = "no co-occurrence rows: build this rank's rows first"; return GE_ERR_STATE; }
    DistState* d = ctx->dist;
    const ge_config& c = ctx->cfg;
    if (c.math != GE_MATH_FAST) { ctx->err = "the block-rotation trainer runs the FAST math path"; return GE_ERR_INVALID; }
    cudaSetDevice(c.device);
    cudaStream_t st = ctx->stream;
    int32_t first = 0, count = 0;
    int r = ge_dist_row_range(ctx, ctx->V, nblocks, &first, &count);
    if (r) return r;
    if (ctx->d_row_off && (ctx->row_first != first || ctx->row_count != count || (ctx->row_lay_nb > 1) != (d->layout != 0 && nblocks > 1) ||
                           (ctx->row_lay_nb > 1 && ctx->row_lay_nb != nblocks))) {
        ctx->err = "co-occurrence rows do not match this rank's block range (use ge_dist_row_range)";
        return GE_ERR_STATE;
    }
    d->nb = nblocks; d->m = nblocks / d->world;
    d->Bs = (ctx->V + nblocks - 1) / nblocks;
    // data plane of the rotation (see the file header).  Half blocks (P = 2) for every world size,
    // so that a 1-GPU run of the schedule walks exactly the buckets (and per-bucket order keys) of
    // an N-GPU run: that is what the N-rank-vs-1-rank loss check compares.
    {
        const char* e = getenv("GE_DIST_ROT");
        const char* legacy = getenv("GE_DIST_PIPE");
        d->rot = ROT_CE;
        if (e && !strcmp(e, "nccl")) d->rot = ROT_NCCL;
        else if ((e && !strcmp(e, "pipe")) || (!e && legacy && atoi(legacy) == 1)) d->rot = ROT_PIPE;
    }
    d->P = d->rot == ROT_NCCL ? 1 : 2;
    d->Hs = d->P > 1 ? (d->Bs + 1) / 2 : d->Bs;
    if (d->m * d->nb * d->P > 1024) { ctx->err = "too many buckets (nblocks^2 / world * 2 must be <= 1024)"; return GE_ERR_INVALID; }
    d->row_base = first; d->row_cnt = count;
    if (d->layout) {   // real rows among the m * Bs slab rows
        int64_t n = 0;
        for (int k = 0; k < d->m; k++) n += block_rows((int64_t)d->rank * d->m + k, nblocks, d->Bs, ctx->V, 1);
        d->row_cnt = (int32_t)n;
    }
    {   // ids -> virtual ids; every row must be one of this rank's
        int32_t* d_bad = nullptr;
        GE_CUDA(cudaMalloc(&d_bad, 4));
        cudaMemsetAsync(d_bad, 0, 4, st);
        if (ctx->nnz > 0)
            to_virtual_ids<<<ctx->sm_count * 8, 256, 0, st>>>(ctx->d_I, ctx->d_J, ctx->nnz, d->lay_nb(), d->Bs, first, count, d_bad);
        int32_t bad = 0;
        cudaError_t e = cudaMemcpyAsync(&bad, d_bad, 4, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        cudaFree(d_bad);
        GE_CUDA(e);
        if (bad) {
            ge_free_cooc(ctx);   // the arrays now hold virtual ids: unusable for anything else
            ctx->err = "a nonzero's row is not one of this rank's rows (ge_dist_row_range / ge_dist_local_ids)";
            return GE_ERR_INVALID;
        }
    }

    GeLap lap(st, "dist prepare");
    // global max (f(x) weight uses it, GloveCost.java:19) and global nnz (cost normaliser)
    if (d->rot != ROT_CE) { r = ensure_comm(ctx); if (r) return r; }   // same environment on every rank: collective
    double gmax = ctx->xmax, gnnz = (double)ctx->nnz;   // nnz < 2^53: exact in a double
    r = plumb_allreduce(ctx, &gmax, 1, 1);
    if (!r) r = plumb_allreduce(ctx, &gnnz, 1, 0);
    if (r) return r;
    ctx->xmax = gmax;
    d->nnz_global = (long long)gnnz;

    // ---- K7: bucket the local nonzeros by (row-block, col-block)
    const int nbuckets = d->m * d->nb * d->P;
    unsigned long long* d_cnt = nullptr;
    GE_CUDA(cudaMalloc(&d_cnt, 1024 * 8));
    GE_CUDA(cudaMemsetAsync(d_cnt, 0, 1024 * 8, st));
    if (ctx->nnz > 0)
        lap.lap("ids + all-reduces");
        bucket_count<<<ctx->sm_count * 4, 256, 0, st>>>(ctx->d_I, ctx->d_J, ctx->nnz, first, d->Bs, d->nb, d->P, d->Hs, d_cnt);
    std::vector<unsigned long long> hc(1024);
    GE_CUDA(cudaMemcpyAsync(hc.data(), d_cnt, 1024 * 8, cudaMemcpyDeviceToHost, st));
    GE_CUDA(cudaStreamSynchronize(st));
    d->bucket_off.assign(nbuckets + 1, 0);
    std::vector<unsigned long long> cur(1024, 0);
    for (int b = 0; b < nbuckets; b++) { d->bucket_off[b + 1] = d->bucket_off[b] + (int64_t)hc[b]; cur[b] = (unsigned long long)d->bucket_off[b]; }
    cudaFree(d->d_nzb); d->d_nzb = nullptr;
    GE_CUDA(cudaMalloc(&d->d_nzb, (size_t)(ctx->nnz > 0 ? ctx->nnz : 1) * sizeof(NzEntry)));
    GE_CUDA(cudaMemcpyAsync(d_cnt, cur.data(), 1024 * 8, cudaMemcpyHostToDevice, st));
    {
        int32_t *cnt_i = nullptr, *cnt_j = nullptr, *d_hot = nullptr;
        std::vector<int32_t> hot(2048, 0x7fffffff);   // [0,1024): thresholds, [1024,2048): any-hot flags (out)
        const bool flags = c.update == GE_UPDATE_ATOMIC && c.opt_method == GE_OPT_ADAGRAD;
        for (int b = 0; b < 1024; b++) hot[1024 + b] = 0;
        r = ge_count_ids(ctx, ctx->d_I, ctx->nnz, first, (int64_t)d->m * d->Bs, &cnt_i);
        if (!r) r = ge_count_ids(ctx, ctx->d_J, ctx->nnz, 0, (int64_t)d->nb * d->Bs, &cnt_j);
        // per bucket: the launch's warp cap (an extreme hub inside it) and, from the warps it then has,
        // the occurrence count from which an id is damped
        d->bucket_cap.assign(nbuckets, 0);
        ctx->hub_kmax = 0; ctx->hub_mass_max = 0; ctx->capped_launches = 0; ctx->warp_cap = 0;
        if (!r && flags) {
            int32_t* d_seg = nullptr;
            int32_t* d_hot0 = nullptr;
            unsigned long long* d_mass = nullptr;
            const int nseg_i = d->m, nseg_j = d->nb * d->P;
            std::vector<int32_t> seg(nseg_i + nseg_j, 0);
            std::vector<int32_t> hot0(nbuckets);
            std::vector<unsigned long long> mass(nbuckets, 0);
            for (int b = 0; b < nbuckets; b++) hot0[b] = ge_hot_count(ctx, (int64_t)hc[b], d->Bs, 0);   // uncapped launches
            cudaError_t e = cudaMalloc(&d_seg, seg.size() * 4);
            if (e == cudaSuccess) e = cudaMalloc(&d_hot0, (size_t)nbuckets * 4);
            if (e == cudaSuccess) e = cudaMalloc(&d_mass, (size_t)nbuckets * 8);
            if (e == cudaSuccess) e = cudaMemsetAsync(d_seg, 0, seg.size() * 4, st);
            if (e == cudaSuccess) e = cudaMemsetAsync(d_mass, 0, (size_t)nbuckets * 8, st);
            if (e == cudaSuccess) e = cudaMemcpyAsync(d_hot0, hot0.data(), (size_t)nbuckets * 4, cudaMemcpyHostToDevice, st);
            if (e == cudaSuccess) {
                segment_max<<<ctx->sm_count * 4, 256, 0, st>>>(cnt_i, (int64_t)d->m * d->Bs, d->Bs, 1, 0, d_seg);
                segment_max<<<ctx->sm_count * 4, 256, 0, st>>>(cnt_j, (int64_t)d->nb * d->Bs, d->Bs, d->P, d->Hs, d_seg + nseg_i);
                bucket_hot_mass<<<ctx->sm_count * 4, 256, 0, st>>>(cnt_i, (int64_t)d->m * d->Bs, 0, d->Bs, d->nb, d->P, d->Hs, d->m, d_hot0, d_mass);
                bucket_hot_mass<<<ctx->sm_count * 4, 256, 0, st>>>(cnt_j, (int64_t)d->nb * d->Bs, 1, d->Bs, d->nb, d->P, d->Hs, d->m, d_hot0, d_mass);
                e = cudaMemcpyAsync(seg.data(), d_seg, seg.size() * 4, cudaMemcpyDeviceToHost, st);
                if (e == cudaSuccess) e = cudaMemcpyAsync(mass.data(), d_mass, (size_t)nbuckets * 8, cudaMemcpyDeviceToHost, st);
            }
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
            cudaFree(d_seg); cudaFree(d_hot0); cudaFree(d_mass);
            if (e != cudaSuccess) { cudaFree(cnt_i); cudaFree(cnt_j); cudaFree(d_cnt); GE_CUDA(e); }
            for (int b = 0; b < nbuckets; b++) {
                const int lrb = b / (d->nb * d->P), cbp = b % (d->nb * d->P);
                const int64_t mi = seg[lrb] / (d->nb * d->P), mj = seg[nseg_i + cbp] / d->m;   // bucket_scatter's estimates
                const double share = hc[b] ? (double)mass[b] / (double)hc[b] : 0.0;
                d->bucket_cap[b] = ge_warp_cap(ctx, (int64_t)hc[b], mi > mj ? mi : mj, share, true);
                hot[b] = ge_hot_count(ctx, (int64_t)hc[b], d->Bs, d->bucket_cap[b]);
                if (d->bucket_cap[b] > 0 && (ctx->warp_cap == 0 || d->bucket_cap[b] < ctx->warp_cap)) ctx->warp_cap = d->bucket_cap[b];  // smallest, for ge_train_hub_info
            }
        }
        int32_t keep_i = 0x7fffffff, keep_j = 0x7fffffff;
        // one launch = one (row-block, half column-block) bucket: it sees 1/(nb*P) of a row's and 1/m of a
        // column's local nonzeros; 1/m of the rank's flagged W rows and 1/(nb*P) of the flagged C rows are live
        if (!r) ge_keep_counts(ctx, cnt_i, (int64_t)d->m * d->Bs, (double)d->nb * d->P, 1.0 / d->m,
                               cnt_j, (int64_t)d->nb * d->Bs, (double)d->m, 1.0 / ((double)d->nb * d->P), &keep_i, &keep_j);
        cudaError_t e = r ? cudaSuccess : cudaMalloc(&d_hot, 2048 * 4);
        if (!r && e == cudaSuccess) e = cudaMemcpyAsync(d_hot, hot.data(), 2048 * 4, cudaMemcpyHostToDevice, st);
        if (!r && e == cudaSuccess && ctx->nnz > 0)
            bucket_scatter<<<ctx->sm_count * 4, 256, 0, st>>>(ctx->d_I, ctx->d_J, ctx->d_X, ctx->nnz, first, d->Bs, d->nb,
                                                             d->P, d->Hs, c.method, ctx->xmax, d_cnt, cnt_i, cnt_j, d_hot, d->m, keep_i, keep_j, d_hot + 1024, d->d_nzb);
        if (!r && e == cudaSuccess) e = cudaGetLastError();
        if (!r && e == cudaSuccess) e = cudaMemcpyAsync(hot.data() + 1024, d_hot + 1024, 1024 * 4, cudaMemcpyDeviceToHost, st);
        if (!r && e == cudaSuccess) e = cudaStreamSynchronize(st);
        cudaFree(cnt_i); cudaFree(cnt_j); cudaFree(d_hot);
        d->bucket_hot.assign(hot.begin() + 1024, hot.begin() + 1024 + nbuckets);
        if (r) { cudaFree(d_cnt); return r; }
        if (e != cudaSuccess) { cudaFree(d_cnt); GE_CUDA(e); }
    }
    cudaFree(d_cnt);
    lap.lap("histograms + bucketing");
    // the COO arrays are no longer needed on the device once bucketed
    cudaFree(ctx->d_I); cudaFree(ctx->d_J); cudaFree(ctx->d_X);
    ctx->d_I = ctx->d_J = nullptr; ctx->d_X = nullptr;

    // ---- store: W rows of this rank, ring of C block slots
    if (d->world > 1 && d->peer_ring) {
        // re-prepare: nobody may free a ring a neighbour still has mapped -> unmap, then a barrier
        dist_unmap_peers(d);
        GE_CUDA(cudaStreamSynchronize(st));
        r = plumb_barrier(ctx);
        if (r) return r;
    }
    ge_free_store(ctx);
    ge_store_geometry(ctx);
    const size_t slab = (size_t)d->Bs * ctx->S;  // floats per block
    const int slots = d->world > 1 ? d->m + 1 : d->m;
    GE_CUDA(cudaMalloc(&ctx->d_W, (size_t)d->m * slab * sizeof(float)));
    cudaFree(d->d_Cring); d->d_Cring = nullptr;
    GE_CUDA(cudaMalloc(&d->d_Cring, (size_t)slots * slab * sizeof(float) + 256));  // + the rotation's flag words
    d->flags = reinterpret_cast<uint32_t*>(d->d_Cring + (size_t)slots * slab);
    GE_CUDA(cudaMemsetAsync(ctx->d_W, 0, (size_t)d->m * slab * sizeof(float), st));
    GE_CUDA(cudaMemsetAsync(d->d_Cring, 0, (size_t)slots * slab * sizeof(float) + 256, st));
    lap.lap("frees + store allocation");
    r = ge_store_init_rows(ctx, ctx->d_W, d->d_Cring, d->layout ? d->m * d->Bs : count, first, d->lay_nb(), d->Bs);
    lap.lap("store init");
    if (r) return r;
    if (!ctx->d_cost) GE_CUDA(cudaMalloc(&ctx->d_cost, sizeof(double)));
    cudaFree(d->d_queue); d->d_queue = nullptr;
    GE_CUDA(cudaMalloc(&d->d_queue, (size_t)nbuckets * sizeof(unsigned long long)));
    if (d->world > 1 && d->rot == ROT_CE) {
        r = dist_map_peers(ctx);
        if (r) return r;
        lap.lap("IPC mapping of the rings");
    }
    d->head = 0; d->sub = 0;
    ctx->epoch = 0;
    ctx->history.clear();
    ctx->have_store = true;
    GE_CUDA(cudaStreamSynchronize(st));
    return GE_OK;
}

extern "C" int ge_dist_train_epoch(ge_ctx* ctx, double* cost) {
    if (!ctx || !ctx->dist || !ctx->dist->d_nzb) { if (ctx) ctx->err = "ge_dist_prepare first"; return GE_ERR_STATE; }
    DistState* d = ctx->dist;
    const ge_config& c = ctx->cfg;
    cudaSetDevice(c.device);
    cudaStream_t st = ctx->stream;
    const size_t slab = (size_t)d->Bs * ctx->S;
    const int slots = d->world > 1 ? d->m + 1 : d->m;
    uint64_t ekey = (uint64_t)c.seed * 0x9E3779B97F4A7C15ULL + (uint64_t)(ctx->epoch + 1) * 0xD6E8FEB86659FD93ULL;
    GE_CUDA(cudaMemsetAsync(ctx->d_cost, 0, sizeof(double), st));
    GE_CUDA(cudaEventRecord(ctx->ev0, st));
    const int P = d->P;
    const int rot = d->world > 1 ? d->rot : -1;
    const bool piped = rot == ROT_PIPE;
    if (rot == ROT_CE && (uint64_t)d->q + (uint64_t)d->nb + 1 >= SEQTAB) { ctx->err = "rotation sequence table exhausted: ge_dist_prepare again"; return GE_ERR_STATE; }
    if (piped) GE_CUDA(cudaMemsetAsync(d->d_queue, 0, (size_t)d->m * d->nb * P * sizeof(unsigned long long), st));
    // byte ranges of the P parts of a block slab (part 0 = rows [0, Hs), part 1 = rows [Hs, Bs))
    const size_t part_off[2] = {0, (size_t)d->Hs * ctx->S};
    const size_t part_len[2] = {(size_t)d->Hs * ctx->S, (size_t)(d->Bs - d->Hs) * ctx->S};
    const int to = (d->rank + d->world - 1) % d->world, from = (d->rank + 1) % d->world;
    int nxfer = 0;
    for (int s = 0; s < d->nb; s++) {
        // window of C blocks held now: block (rank*m + s + k) mod nb sits in slot (head + k) mod slots.
        // The window's FIRST block is updated first, part by part; as soon as a part is done it leaves
        // for rank-1 while the next part / the other window blocks are being updated, and the matching
        // part of the block arriving from rank+1 lands in the spare slot.  That block is the window's
        // LAST one in the next sub-epoch, whose updates wait on the part's arrival only.
        for (int k = 0; k < d->m; k++) {
            const int cb = (d->rank * d->m + s + k) % d->nb;
            const int lrb = k;  // virtual rank v = rank*m + k processes column block (v + s) mod nb
            float* Cslot = d->d_Cring + (size_t)((d->head + k) % slots) * slab;
            for (int p = 0; p < P; p++) {
                const int bucket = (lrb * d->nb + cb) * P + p;
                const int64_t off = d->bucket_off[bucket], cnt = d->bucket_off[bucket + 1] - off;
                if (k == d->m - 1 && d->q > 0) {  // this block arrived during the previous sub-epoch (transfer q-1)
                    if (piped) GE_CUDA(cudaStreamWaitEvent(st, d->ev_recv[p], 0));
                    else if (rot == ROT_CE && g_wait32(st, (unsigned long long)(uintptr_t)(d->flags + p), d->q, 0 /* GEQ */) != 0) {
                        ctx->err = "cuStreamWaitValue32 (ready flag) failed"; return GE_ERR_CUDA;
                    }
                }
                ctx->launch_max_warps = d->bucket_cap[bucket];
                ctx->launch_pf = d->bucket_cap[bucket] > 0 ? 8 : -1;
                int r = ge_launch_update(ctx, d->d_nzb + off, cnt, ctx->d_W, Cslot, nullptr,
                                         ekey ^ ((uint64_t)bucket * 0xA24BAED4963EE407ULL), ctx->d_cost, st, d->Bs,
                                         d->bucket_hot[bucket] != 0, piped ? d->d_queue + bucket : nullptr);
                ctx->launch_max_warps = 0;
                ctx->launch_pf = -1;
                if (r) return r;
                if (k != 0 || (!piped && rot != ROT_CE)) continue;
                const size_t src_off = (size_t)d->head * slab + part_off[p];
                const size_t dst_off = (size_t)((d->head + d->m) % slots) * slab + part_off[p];
                GE_CUDA(cudaEventRecord(d->ev_done[p], st));
                GE_CUDA(cudaStreamWaitEvent(d->cs, d->ev_done[p], 0));
                if (piped) {
                    GE_NCCL(g_nccl.GroupStart());
                    GE_NCCL(g_nccl.Send(d->d_Cring + src_off, part_len[p] * sizeof(float), NCCL_UINT8, to, d->comm, d->cs));
                    GE_NCCL(g_nccl.Recv(d->d_Cring + dst_off, part_len[p] * sizeof(float), NCCL_UINT8, from, d->comm, d->cs));
                    GE_NCCL(g_nccl.GroupEnd());
                    GE_CUDA(cudaEventRecord(d->ev_recv[p], d->cs));
                } else {
                    // transfer q, part p: my slot `head` -> rank-1's slot (head + m) % slots.  That slot is
                    // the one rank-1 sent FROM in transfer q-1: wait until it says so (free >= q).
                    if (g_wait32(d->cs, (unsigned long long)(uintptr_t)(d->flags + 2 + p), d->q, 0) != 0) {
                        ctx->err = "cuStreamWaitValue32 (free flag) failed"; return GE_ERR_CUDA;
                    }
                    cudaEvent_t e0 = d->xfer_ev[(size_t)2 * nxfer], e1 = d->xfer_ev[(size_t)2 * nxfer + 1];
                    nxfer++;
                    GE_CUDA(cudaEventRecord(e0, d->cs));
                    // chunks of whole rows; below 8 MB a part stays one copy
                    const size_t bytes = part_len[p] * sizeof(float);
                    const int nch = bytes >= ((size_t)8 << 20) ? d->split : 1;
                    const size_t rows_p = part_len[p] / ctx->S, rows_c = (rows_p + nch - 1) / nch;
                    if (nch > 1) GE_CUDA(cudaEventRecord(d->ev_go, d->cs));
                    for (int x = 0; x < nch; x++) {
                        const size_t r0 = (size_t)x * rows_c, r1 = r0 + rows_c < rows_p ? r0 + rows_c : rows_p;
                        if (r0 >= r1) continue;
                        cudaStream_t xs = x == 0 ? d->cs : d->cs_extra[x - 1];
                        if (x > 0) GE_CUDA(cudaStreamWaitEvent(xs, d->ev_go, 0));
                        GE_CUDA(cudaMemcpyAsync(d->peer_ring + dst_off + r0 * ctx->S, d->d_Cring + src_off + r0 * ctx->S,
                                                (r1 - r0) * ctx->S * sizeof(float), cudaMemcpyDeviceToDevice, xs));
                        if (x > 0) {
                            GE_CUDA(cudaEventRecord(d->ev_chunk[x - 1], xs));
                            GE_CUDA(cudaStreamWaitEvent(d->cs, d->ev_chunk[x - 1], 0));
                        }
                    }
                    GE_CUDA(cudaEventRecord(e1, d->cs));
                    // flags travel as 4-byte DMA copies behind the data on the same stream (no SM involved)
                    GE_CUDA(cudaMemcpyAsync(d->prev_flags + p, d->seqtab + d->q + 1, 4, cudaMemcpyDeviceToDevice, d->cs));      // ready at rank-1
                    GE_CUDA(cudaMemcpyAsync(d->next_flags + 2 + p, d->seqtab + d->q + 1, 4, cudaMemcpyDeviceToDevice, d->cs));  // free at rank+1
                }
            }
        }
        if (rot == ROT_NCCL) {  // whole-block rotation on the compute stream
            float* sendp = d->d_Cring + (size_t)d->head * slab;
            float* recvp = d->d_Cring + (size_t)((d->head + d->m) % slots) * slab;
            GE_NCCL(g_nccl.GroupStart());
            GE_NCCL(g_nccl.Send(sendp, slab * sizeof(float), NCCL_UINT8, to, d->comm, st));
            GE_NCCL(g_nccl.Recv(recvp, slab * sizeof(float), NCCL_UINT8, from, d->comm, st));
            GE_NCCL(g_nccl.GroupEnd());
        }
        d->head = (d->head + 1) % slots;  // one GPU: the rotation is this pointer shift alone
        if (d->world > 1) d->q++;
    }
    // every block home before the epoch ends: the last transfer's parts have landed here
    if (piped) for (int p = 0; p < P; p++) GE_CUDA(cudaStreamWaitEvent(st, d->ev_recv[p], 0));
    if (rot == ROT_CE) for (int p = 0; p < P; p++)
        if (g_wait32(st, (unsigned long long)(uintptr_t)(d->flags + p), d->q, 0) != 0) { ctx->err = "cuStreamWaitValue32 (epoch end) failed"; return GE_ERR_CUDA; }
    if (d->world == 1) d->head = (d->head + slots - (d->nb % slots)) % slots;  // nb shifts of an nb-ring: identity
    GE_CUDA(cudaEventRecord(ctx->ev1, st));
    double total = 0;
    if (d->world > 1 && !d->shm) GE_NCCL(g_nccl.AllReduce(ctx->d_cost, ctx->d_cost, 1, NCCL_FLOAT64, NCCL_SUM, d->comm, st));
    GE_CUDA(cudaMemcpyAsync(&total, ctx->d_cost, sizeof(double), cudaMemcpyDeviceToHost, st));
    GE_CUDA(cudaStreamSynchronize(st));
    if (d->world > 1 && d->shm) { int r = plumb_allreduce(ctx, &total, 1, 0); if (r) return r; }
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->update_seconds += ms * 1e-3;
    if (rot == ROT_CE) {  // every transfer timed on the copy stream behind its waits: the DMA alone
        GE_CUDA(cudaStreamSynchronize(d->cs));
        for (int x = 0; x < nxfer; x++) {
            float t = 0;
            if (cudaEventElapsedTime(&t, d->xfer_ev[(size_t)2 * x], d->xfer_ev[(size_t)2 * x + 1]) == cudaSuccess) d->comm_seconds += t * 1e-3;
        }
    }
    const double local = total / (double)d->nnz_global;
    ctx->history.push_back(local);
    ctx->epoch++;
    if (cost) *cost = local;
    return GE_OK;
}

// (W + C)/2 for this rank's id range; valid between epochs, when every C block is home.  Same pinned
// double-buffered path as ge_extract (train.cu).
template <typename T>
static int dist_extract_impl(ge_ctx* ctx, T* out, int64_t cap) {
    if (!ctx || !ctx->dist || !ctx->have_store) { if (ctx) ctx->err = "ge_dist_prepare first"; return GE_ERR_STATE; }
    DistState* d = ctx->dist;
    const int32_t dim = ctx->cfg.dim;
    if (!out || cap < (int64_t)d->row_cnt * dim) { ctx->err = "extract capacity too small"; return GE_ERR_INVALID; }
    cudaSetDevice(ctx->cfg.device);
    const size_t slab = (size_t)d->Bs * ctx->S;
    const int slots = d->world > 1 ? d->m + 1 : d->m;
    std::vector<const float*> W, C;
    std::vector<int64_t> rows;
    for (int k = 0; k < d->m; k++) {
        const int64_t n = block_rows((int64_t)d->rank * d->m + k, d->nb, d->Bs, ctx->V, d->layout);
        if (n <= 0) continue;
        W.push_back(ctx->d_W + (size_t)k * slab);
        C.push_back(d->d_Cring + (size_t)((d->head + k) % slots) * slab);
        rows.push_back(n);
    }
    ge_advise_result_buffer(out, (size_t)d->row_cnt * dim * sizeof(T));
    return ge_extract_blocks<T>(ctx, (int)rows.size(), W.data(), C.data(), rows.data(), out);
}
// layout of the block grid: 0 = contiguous id blocks (default), 1 = cyclic (block b = ids b, b + nb, ...);
// call before ge_dist_row_range / the build / ge_dist_prepare
extern "C" int ge_dist_set_layout(ge_ctx* ctx, int32_t layout) {
    if (!ctx || !ctx->dist) { if (ctx) ctx->err = "ge_dist_init first"; return GE_ERR_STATE; }
    if (layout != 0 && layout != 1) { ctx->err = "layout: 0 contiguous, 1 cyclic"; return GE_ERR_INVALID; }
    if (ctx->dist->d_nzb && ctx->dist->layout != layout) { ctx->err = "layout is fixed once prepared"; return GE_ERR_STATE; }
    ctx->dist->layout = layout;
    return GE_OK;
}

// real vertex ids of this rank's rows, in the order ge_dist_extract emits them; returns their number
extern "C" int64_t ge_dist_local_ids(ge_ctx* ctx, int32_t* ids, int64_t cap) {
    if (!ctx || !ctx->dist || !ctx->dist->d_nzb) { if (ctx) ctx->err = "ge_dist_prepare first"; return GE_ERR_STATE; }
    DistState* d = ctx->dist;
    if (!ids || cap < d->row_cnt) { ctx->err = "ge_dist_local_ids: capacity too small"; return GE_ERR_INVALID; }
    if (d->row_cnt == 0) return 0;
    cudaSetDevice(ctx->cfg.device);
    int32_t* dv = nullptr;
    GE_CUDA(cudaMalloc(&dv, (size_t)d->row_cnt * 4));
    local_ids<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(dv, d->m, d->rank * d->m, d->nb, d->Bs, ctx->V, d->layout);
    cudaError_t e = cudaMemcpyAsync(ids, dv, (size_t)d->row_cnt * 4, cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(dv);
    GE_CUDA(e);
    return d->row_cnt;
}

extern "C" int ge_dist_extract_f32(ge_ctx* ctx, float* out, int64_t cap) { return dist_extract_impl<float>(ctx, out, cap); }
extern "C" int ge_dist_extract(ge_ctx* ctx, double* out, int64_t cap) { return dist_extract_impl<double>(ctx, out, cap); }

// A1 hand-off for N ranks: ONE host -> device copy on `root`, then ncclBroadcast of the six CSR arrays over
// NVLink, instead of N uploads of the same graph across PCIe.  Non-root ranks pass V = 0 and null arrays.
extern "C" int ge_dist_set_graph_csr(ge_ctx* ctx, int32_t root, int32_t V, const int64_t* out_off, const int32_t* out_nbr,
                                     const float* out_w, const int64_t* in_off, const int32_t* in_nbr, const float* in_w) {
    if (!ctx || !ctx->dist) { if (ctx) ctx->err = "ge_dist_init first"; return GE_ERR_STATE; }
    DistState* d = ctx->dist;
    if (d->world == 1) return ge_set_graph_csr(ctx, V, out_off, out_nbr, out_w, in_off, in_nbr, in_w);
    if (root < 0 || root >= d->world) { ctx->err = "bad root"; return GE_ERR_INVALID; }
    const bool me = d->rank == root;
    cudaSetDevice(ctx->cfg.device);
    cudaStream_t st = ctx->stream;
    GeLap lap(st, "dist hand-off");
    long long hdr[4] = {0, 0, 0, 0};  // V, E_out, E_in, ok
    if (me) {
        const bool ok = V > 0 && out_off && in_off && out_off[V] >= 0 && in_off[V] >= 0 &&
                        (out_off[V] == 0 || (out_nbr && out_w)) && (in_off[V] == 0 || (in_nbr && in_w));
        hdr[0] = V; hdr[1] = ok ? out_off[V] : 0; hdr[2] = ok ? in_off[V] : 0; hdr[3] = ok;
    }
    if (d->shm) {
        long long allh[4 * 64];
        if (d->world > 64) { ctx->err = "world too large"; return GE_ERR_INVALID; }
        int rr = plumb_allgather(ctx, hdr, allh, sizeof(hdr));
        if (rr) return rr;
        memcpy(hdr, allh + 4 * (size_t)root, sizeof(hdr));
    } else {
        long long* d_hdr = nullptr;
        GE_CUDA(cudaMalloc(&d_hdr, sizeof(hdr)));
        cudaError_t e = cudaMemcpyAsync(d_hdr, hdr, sizeof(hdr), cudaMemcpyHostToDevice, st);
        int nr = e == cudaSuccess ? g_nccl.Broadcast(d_hdr, d_hdr, sizeof(hdr), NCCL_UINT8, root, d->comm, st) : 0;
        if (e == cudaSuccess && !nr) e = cudaMemcpyAsync(hdr, d_hdr, sizeof(hdr), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess && !nr) e = cudaStreamSynchronize(st);
        cudaFree(d_hdr);
        GE_CUDA(e);
        GE_NCCL(nr);
    }
    if (!hdr[3]) { ctx->err = "bad graph arguments on the root rank"; return GE_ERR_INVALID; }
    lap.lap("header broadcast");
    int r = ge_graph_alloc(ctx, (int32_t)hdr[0], hdr[1], hdr[2]);
    if (r) { ge_free_graph(ctx); return r; }
    lap.lap("device allocation");
    const size_t ob = (size_t)(ctx->V + 1) * sizeof(int64_t);
    const struct { void* dv; const void* h; size_t n; } parts[] = {
        {ctx->d_out_off, out_off, ob}, {ctx->d_in_off, in_off, ob},
        {ctx->d_out_nbr, out_nbr, (size_t)ctx->E_out * 4}, {ctx->d_out_w, out_w, (size_t)ctx->E_out * 4},
        {ctx->d_in_nbr, in_nbr, (size_t)ctx->E_in * 4}, {ctx->d_in_w, in_w, (size_t)ctx->E_in * 4}};
    if (me)
        for (const auto& q : parts)
            if (q.n) { r = upload_staged(ctx, q.dv, q.h, q.n); if (r) { ge_free_graph(ctx); return r; } }
    // peers pull the arrays out of the root's device memory (CUDA IPC, NVLink) when the node plumbing is up;
    // else ncclBroadcast
    bool via_ipc = d->shm != nullptr;
    if (via_ipc) {
        struct Handles { cudaIpcMemHandle_t h[6]; int32_t ok; int32_t pad[31]; };
        Handles mine{};
        std::vector<Handles> all(d->world);
        if (me) {
            mine.ok = 1;
            GE_CUDA(cudaStreamSynchronize(st));   // uploads done before anybody reads
            for (int k = 0; k < 6; k++)
                if (parts[k].n && cudaIpcGetMemHandle(&mine.h[k], parts[k].dv) != cudaSuccess) { cudaGetLastError(); mine.ok = 0; }
        }
        r = plumb_allgather(ctx, &mine, all.data(), sizeof(Handles));
        if (r) return r;
        void* src[6] = {};
        double bad = all[root].ok ? 0.0 : 1.0;
        if (!me && bad == 0.0)
            for (int k = 0; k < 6; k++)
                if (parts[k].n && cudaIpcOpenMemHandle(&src[k], all[root].h[k], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); src[k] = nullptr; bad = 1.0; }
        r = plumb_allreduce(ctx, &bad, 1, 0);
        if (r) return r;
        cudaError_t e = cudaSuccess;
        if (bad == 0.0 && !me) {
            for (int k = 0; k < 6 && e == cudaSuccess; k++)
                if (parts[k].n) e = cudaMemcpyAsync(parts[k].dv, src[k], parts[k].n, cudaMemcpyDeviceToDevice, st);
            if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        }
        for (int k = 0; k < 6; k++) if (src[k]) cudaIpcCloseMemHandle(src[k]);
        GE_CUDA(e);
        if (bad != 0.0) via_ipc = false;
        else { r = plumb_barrier(ctx); if (r) return r; }   // the root's arrays stay untouched until every peer has its copy
    }
    if (!via_ipc) {
        r = ensure_comm(ctx);
        if (r) return r;
        for (const auto& q : parts)
            if (q.n) GE_NCCL(g_nccl.Broadcast(q.dv, q.dv, q.n, NCCL_UINT8, root, d->comm, st));
    }
    lap.lap("H2D on root + broadcasts");
    ctx->graph_seconds = 0;
    r = ge_graph_validate(ctx);
    if (r) return r;
    lap.lap("CSR validation");
    r = ge_graph_finish(ctx);
    lap.lap("node totals / wmax");
    return r;
}

// data plane of the rotation actually in use: 0 = NCCL send/recv in the compute stream, 1 = half-block
// NCCL on a second stream, 2 = half blocks pushed by the copy engine into the neighbour's IPC-mapped
// ring; -1 = single GPU (pointer shift)
extern "C" int ge_dist_rotation_mode(ge_ctx* ctx) {
    if (!ctx || !ctx->dist) return GE_ERR_STATE;
    return ctx->dist->world > 1 ? ctx->dist->rot : -1;
}

// per-rank bookkeeping: nnz of each (local row-block, col-block) bucket; comm_seconds = device time of
// this rank's outgoing copies (copy-engine plane only: each DMA timed behind its hand-shake waits)
extern "C" int ge_dist_stats(ge_ctx* ctx, int64_t* bucket_nnz, int32_t cap, double* comm_seconds, int64_t* nnz_global) {
    if (!ctx || !ctx->dist) return GE_ERR_STATE;
    DistState* d = ctx->dist;
    const int n = d->m * d->nb;
    if (bucket_nnz) for (int b = 0; b < n && b < cap; b++) bucket_nnz[b] = d->bucket_off[(size_t)(b + 1) * d->P] - d->bucket_off[(size_t)b * d->P];
    if (comm_seconds) *comm_seconds = d->comm_seconds;
    if (nnz_global) *nnz_global = d->nnz_global;
    return n;
}
