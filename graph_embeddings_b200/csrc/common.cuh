// common.cuh — shared declarations of libgraphemb.so (sm_100a only; no CPU fallback).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "../../include/graphemb.h"

#define GE_CUDA(call)                                                                        \
    do {                                                                                     \
        cudaError_t e__ = (call);                                                            \
        if (e__ != cudaSuccess) {                                                            \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e__);                  \
            return e__ == cudaErrorMemoryAllocation ? GE_ERR_NOMEM : GE_ERR_CUDA;            \
        }                                                                                    \
    } while (0)

// GE_DEBUG=1 in the environment: synchronise after the named step and report which one faulted
#define GE_DBG(name)                                                                         \
    do {                                                                                     \
        static const bool dbg__ = getenv("GE_DEBUG") != nullptr;                             \
        if (dbg__) {                                                                         \
            cudaError_t e__ = cudaStreamSynchronize(ctx->stream);                            \
            if (e__ == cudaSuccess) e__ = cudaGetLastError();                                \
            if (e__ != cudaSuccess) {                                                        \
                ctx->err = std::string(name) + ": " + cudaGetErrorString(e__);               \
                return GE_ERR_CUDA;                                                          \
            }                                                                                \
        }                                                                                    \
    } while (0)

// ---- java.util.Random as an affine map over Z/2^48 (JDK Javadoc; the reference draws through
// J/util/rnd/ExtendedRandom.java:27-35).  Usable on host and device. ---------------------------
struct JavaLcg {
    static constexpr uint64_t MULT = 0x5DEECE66DULL, ADD = 0xBULL, MASK = (1ULL << 48) - 1;
    uint64_t s;
    __host__ __device__ static JavaLcg seeded(int64_t seed) { return JavaLcg{((uint64_t)seed ^ MULT) & MASK}; }
    __host__ __device__ inline int32_t next(int bits) {
        s = (s * MULT + ADD) & MASK;
        return (int32_t)(int64_t)(s >> (48 - bits));
    }
    // O(log n) skip of n draws: compose the affine map with itself by squaring.
    __host__ __device__ inline void jump(uint64_t n) {
        uint64_t a = MULT, c = ADD, A = 1, C = 0;
        while (n) {
            if (n & 1) { A = (A * a) & MASK; C = (C * a + c) & MASK; }
            c = ((a + 1) * c) & MASK;
            a = (a * a) & MASK;
            n >>= 1;
        }
        s = (A * s + C) & MASK;
    }
    __host__ __device__ inline float nextFloat() { return (float)next(24) / (float)(1 << 24); }
    // Random.nextInt(bound), JDK 8 (power-of-two fast path + rejection loop)
    __host__ inline int32_t nextInt(int32_t bound) {
        int32_t r = next(31);
        const int32_t m = bound - 1;
        if ((bound & m) == 0) return (int32_t)(((int64_t)bound * (int64_t)r) >> 31);
        for (int32_t u = r;; u = next(31)) {
            r = u % bound;
            if ((int32_t)((uint32_t)u - (uint32_t)r + (uint32_t)m) >= 0) return r;
        }
    }
};

// One co-occurrence nonzero as the update kernel reads it (16 B, one LDG.128).  Bits 27..30 of i / j
// carry the concurrency class k of a hub row / column (ge_damp_bits): the FAST/ATOMIC kernel corrects
// that side's delta for the ~K = GE_HOT_WRITERS * 2^((k-0.5)/2) warps that read the same stale row
// (train.cu).  Bit 31 marks the rows kept resident in L2.  Ids therefore stay below 2^27 (GE_ID_MASK)
// — no restriction next to the reference, whose int-indexed focus[] needs V * dim < 2^31.
//   FAST math : a = (float)log(x) [glove] or (float)log(x/(1-x)) [pglove], b = f(x) weight
//   JAVA math : a = x, b unused (log/pow are evaluated in fp64 inside the kernel)
struct __align__(16) NzEntry {
    int32_t i, j;
    float a, b;
};
#define GE_ID_BITS 27
#define GE_ID_MASK 0x07ffffff
#define GE_ID_KEEP ((int32_t)0x80000000)  // bit 31: row belongs to the L2-resident set (evict_last cache policy)
// Concurrency class of an id that occurs `cnt` times where `hot_count` occurrences mean GE_HOT_WRITERS
// expected concurrent writers: the smallest k >= 1 with cnt * 2^(-k/2) < hot_count (capped at 15), i.e.
// cnt / hot_count lies in [2^((k-1)/2), 2^(k/2)).  0 = not a hub.
__host__ __device__ inline int32_t ge_damp_bits(int32_t cnt, int32_t hot_count) {
    if (cnt < hot_count || hot_count <= 0) return 0;
    const int32_t k = (int32_t)floorf(2.0f * log2f((float)cnt / (float)hot_count)) + 1;
    return k > 15 ? 15 : k;
}

struct DistState;  // dist.cu
struct SimState;   // sim.cu

struct ge_ctx {
    ge_config cfg;
    std::string err;
    int sm_count = 148;
    cudaStream_t stream = nullptr;      // stream all work is issued on (own_stream or the caller's)
    cudaStream_t own_stream = nullptr;
    void* h_stage[2] = {nullptr, nullptr};        // pinned staging for pageable hand-off buffers (api.cu upload_staged)
    cudaEvent_t ev_stage[2] = {nullptr, nullptr};
    bool stage_busy[2] = {false, false};          // a copy out of the staging buffer may still be in flight

    // --- graph (A1) ---
    int32_t V = 0;
    int64_t E_out = 0, E_in = 0;
    double graph_seconds = 0;  // device time of the last edge-list flattening
    int64_t *d_out_off = nullptr, *d_in_off = nullptr;
    int32_t *d_out_nbr = nullptr, *d_in_nbr = nullptr;
    float *d_out_w = nullptr, *d_in_w = nullptr;
    double *d_out_total = nullptr, *d_in_total = nullptr;  // per-node fp64 weight sums
    float *d_out_wmax = nullptr, *d_in_wmax = nullptr;      // per-node heaviest edge (pruning bound)

    // --- co-occurrence matrix (A6) ---
    int64_t nnz = 0;
    double xmax = 0;
    bool have_cooc = false;
    int32_t *d_I = nullptr, *d_J = nullptr;
    float* d_X = nullptr;
    int64_t* d_row_off = nullptr;  // row_count+1 offsets when built here (rows contiguous), else null
    int32_t row_first = 0, row_count = 0, row_lay_nb = 0;  // virtual-id range when row_lay_nb > 1 (ge_vid)
    int64_t bca_overflow_rows = 0;  // rows that needed the global-memory table tier
    int64_t bca_pops = 0, bca_scanned = 0, bca_fetched = 0;
    double bca_seconds = 0;

    // --- embedding store (B1): two arrays of V rows x S floats,
    //     row = [vec(D4) | gradsq(D4) | bias | gradsq_bias | pad], S = roundup(2*D4+2, 8) ---
    int32_t D4 = 0, S = 0, NS = 1, BO = 0;  // NS state vectors per row, BO = offset of the bias block
    float *d_W = nullptr, *d_C = nullptr;
    bool have_store = false;
    NzEntry* d_nz = nullptr;      // packed nonzeros
    int32_t* d_perm = nullptr;    // GE_ORDER_JAVA only
    std::vector<int32_t> h_perm;  // GE_ORDER_JAVA host replay
    JavaLcg rng{0};               // stream position after init (GE_ORDER_JAVA)
    double* d_cost = nullptr;     // epoch cost accumulator
    bool any_hot = false;         // packed nonzeros carry hub damping exponents
    int32_t keep_count = 0x7fffffff;  // ids occurring at least this often are in the L2-resident set
    double hub_mass_max = 0;          // largest share of a launch's nonzeros that touch a damped id
    double hub_kmax = 0;              // hottest id: warps of a full wave expected on it at once
    int32_t capped_launches = 0;
    int64_t warp_cap = 0;             // extreme hub present: launches use at most this many warps (ge_warp_cap); 0 = none
    int64_t launch_max_warps = 0, launch_reserve_blocks = 0;  // per-launch overrides (train.cu shape_launch)
    int32_t launch_pf = -1;
    int32_t epoch = 0;
    std::vector<double> history;
    double update_seconds = 0;
    int64_t launches = 0, updates = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;

    DistState* dist = nullptr;
    SimState* sim = nullptr;  // result of the last similarity compare
};

// train.cu
int ge_train_prepare(ge_ctx* ctx);  // pack nonzeros, allocate cost/perm
int ge_launch_update(ge_ctx* ctx, const NzEntry* nz, int64_t count, float* W, float* C,
                     const int32_t* perm, uint64_t order_key, double* d_cost, cudaStream_t st,
                     int64_t span /* ids addressable on the smaller side: bounds concurrency */,
                     bool any_hot /* some entries carry hub damping exponents: use the HOT kernel variant */,
                     unsigned long long* queue /* zeroed device counter: dynamic batch claiming; nullptr = static */);
void ge_free_store(ge_ctx* ctx);
// (W + C) / 2 of `nblocks` slab pairs, rows[k] rows each, concatenated into host memory through pinned double buffers
template <typename T>
int ge_extract_blocks(ge_ctx* ctx, int nblocks, const float* const* W, const float* const* C, const int64_t* rows, T* out);
int ge_count_ids(ge_ctx* ctx, const int32_t* d_ids, int64_t nnz, int32_t base, int64_t n, int32_t** d_cnt);
int32_t ge_hot_count(ge_ctx* ctx, int64_t launch_nnz, int64_t span, int64_t warp_cap);
// GE_DEBUG stopwatch: lap("name") synchronises the stream and prints the time since the previous lap
struct GeLap {
    cudaStream_t st; const char* what; bool on; std::chrono::steady_clock::time_point t;
    GeLap(cudaStream_t s, const char* w) : st(s), what(w), on(getenv("GE_DEBUG") != nullptr), t(std::chrono::steady_clock::now()) {}
    void lap(const char* name) {
        if (!on) return;
        cudaStreamSynchronize(st);
        const auto n = std::chrono::steady_clock::now();
        fprintf(stderr, "[%s] %-28s %8.1f ms\n", what, name, std::chrono::duration<double, std::milli>(n - t).count());
        t = n;
    }
};
void ge_advise_result_buffer(void* p, size_t bytes);  // MADV_HUGEPAGE on a large host result buffer
int64_t ge_warp_cap(ge_ctx* ctx, int64_t launch_nnz, int64_t max_count, double hub_mass, bool block_schedule);  // warps a launch may use when one id holds max_count of its nonzeros (0 = no cap)
  // occurrences from which an id is a hub (damped deltas)
void ge_keep_counts(ge_ctx* ctx, const int32_t* cnt_i, int64_t ni, double div_i, double share_i, const int32_t* cnt_j, int64_t nj,
                    double div_j, double share_j, int32_t* keep_i, int32_t* keep_j);  // L2-resident set thresholds (raw occurrence counts)
void ge_store_geometry(ge_ctx* ctx);  // D4, NS, BO, S from cfg
int ge_store_init_rows(ge_ctx* ctx, float* W, float* C, int32_t rows, int32_t row0, int32_t lay_nb = 0, int32_t lay_Bs = 0);
// sim.cu
void ge_free_sim(ge_ctx* ctx);
// api.cu
void ge_free_graph(ge_ctx* ctx);
int ge_graph_finish(ge_ctx* ctx);
int ge_graph_alloc(ge_ctx* ctx, int32_t V, int64_t E_out, int64_t E_in);   // frees what depends on the old graph, allocates the CSR arrays
int ge_graph_validate(ge_ctx* ctx);   // sorted / unique / in-range neighbour lists, monotone offsets (device pass)
int upload_staged(ge_ctx* ctx, void* dst, const void* src, size_t bytes);   // pageable host -> device through pinned double buffers
// bca.cu
// Id layout of the block schedule.  Contiguous (nb <= 1): block b = ids [b*Bs, (b+1)*Bs).  Cyclic: block
// b = ids {b, b + nb, b + 2nb, ...} — an id-ordered graph (node types in consecutive id ranges, like the
// DBLP-shaped configs) then puts the same mix of types into every block.  Both are expressed through a
// "virtual id" under which blocks are contiguous: vid(v) = (v mod nb) * Bs + v div nb, Bs = ceil(V / nb);
// vids whose inverse is >= V are padding (at most one per block, at its end).
__host__ __device__ inline int32_t ge_vid(int32_t v, int32_t nb, int32_t Bs) { return nb > 1 ? (v % nb) * Bs + v / nb : v; }
__host__ __device__ inline int32_t ge_vid_inv(int32_t vid, int32_t nb, int32_t Bs) { return nb > 1 ? vid / Bs + nb * (vid % Bs) : vid; }
// rows = the `count` virtual ids from `first` on (lay_nb <= 1: plain ids)
int ge_bca_build(ge_ctx* ctx, int32_t first, int32_t count, int32_t lay_nb = 0, int32_t lay_Bs = 0);
void ge_free_cooc(ge_ctx* ctx);
// dist.cu
void ge_dist_free(ge_ctx* ctx);
