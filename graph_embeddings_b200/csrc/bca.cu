// bca.cu — co-occurrence builder (K1-K3 of SURVEY §2.3): the reference's Bookmark Coloring push.
//
// Replaces new BookmarkColoring(graph, config) (J/bca/BookmarkColoring.java:32-120):
//   DirectedWeighted.doWork   (J/bca/jobs/DirectedWeighted.java:31-101, forward + reverse pass)
//   UndirectedWeighted.doWork (J/bca/jobs/UndirectedWeighted.java:31-114)
//   BCV add / merge / toUnity / toCounts / max (J/bca/util/BCV.java:35-107)
//   consumer loop: normalise, setMax, append (I,J,V) (BookmarkColoring.java:76-113)
//
// One warp owns one bookmark.  The TreeMap frontier and the BCV HashMap of the reference become
// ONE per-warp record table {node id | active bit, wet paint (fp64), f32 sum fwd, f32 sum rev}
// with an open-addressing index and a compact list of the records currently in the frontier.
// pollFirstEntry() = warp-wide min over the keys of that list (REDUX.MIN): O(frontier / 32) per
// pop, however many records the bookmark has retired.  Because a bookmark's events are all produced by its own warp, in pop order, the
// float32 "add in pop order" of BCV.add needs no global sort: results are bit-exact by
// construction (fp64 paint with explicit round-to-nearest ops, no FMA contraction).
// Table tier 1 lives in shared memory (cap 512 records); bookmarks that outgrow it are re-run
// by the same code on a global-memory table sized by the bound 2/(alpha*eps).
#include "common.cuh"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstring>

namespace {

constexpr uint32_t INACTIVE = 0x80000000u;
constexpr int T1_CAP = 104;          // tier-1 records per warp (shared memory): 25 B each + 512 B index -> 16 blocks of 2 warps per SM (register-bound)
constexpr int T1_WARPS = 2;          // warps per block in tier 1
constexpr int MAX_ARENAS = 8;

struct DirGraph {
    const int64_t* off;
    const int32_t* nbr;
    const float* w;
    const double* total;
    const float* wmax;
};

struct BcaParams {
    int32_t V;
    DirGraph out, in;
    double alpha, eps, one_minus_alpha;
    int32_t directed, normalize;
    // work: bookmarks work[0..nwork) or, when work == nullptr, first .. first+nwork
    const int32_t* work;
    int32_t nwork, first, row_base;  // row_base: index 0 of row_start/row_cnt
    int32_t lay_nb, lay_Bs;          // cyclic block layout: first / row_base are virtual ids (common.cuh ge_vid)
    int32_t* next;                   // dynamic work counter
    // table geometry
    int32_t cap, hcap_log2;
    // tier-2 scratch (global), per warp: laid out by the kernel from this base
    uint8_t* scratch;
    size_t scratch_per_warp;
    int32_t ntables;                 // tier 2: tables in `scratch`; warps beyond it do nothing
    // outputs
    int32_t *aI, *aJ;
    float* aX;
    int64_t arena_cap, arena_base;   // arena covers unified indices [arena_base, arena_base+arena_cap)
    unsigned long long* cursor;      // entries used in this arena
    int64_t* row_start;              // unified index of the row's first entry, -1 = not emitted yet
    int32_t* row_cnt;
    int32_t* overflow; int32_t* n_overflow;   // table too small -> next tier
    int32_t* deferred; int32_t* n_deferred;   // arena full -> next arena
    int32_t* gmax_bits;                       // global max (non-negative float as int)
    unsigned long long *pops, *scanned, *fetched;
};

__device__ __forceinline__ uint16_t cas_slot(uint16_t* a, uint16_t v) {
    return ::atomicCAS(reinterpret_cast<unsigned short*>(a), (unsigned short)0, (unsigned short)v);
}
__device__ __forceinline__ uint32_t cas_slot(uint32_t* a, uint32_t v) { return ::atomicCAS(a, 0u, v); }

template <typename HT>
struct Table {
    uint32_t* key;   // node id | INACTIVE
    double* wet;     // wet paint while in the frontier, 0 otherwise
    float* accf;     // BCV value, forward (or undirected) pass
    float* accr;     // BCV value, reverse pass
    uint8_t* flg;    // bit0: popped in fwd, bit1: popped in rev
    HT* hash;        // record index + 1, 0 = empty
    HT* ord;         // record indices in first-pop order (= the order BCV entries are created)
    HT* act;         // record indices of the frontier (TreeMap contents), unordered
    int32_t cap;
    uint32_t hshift, hmask;

    __device__ __forceinline__ uint32_t slot0(uint32_t id) const { return (id * 0x9E3779B1u) >> hshift; }
    // returns record index or -1
    __device__ __forceinline__ int find(uint32_t id) const {
        uint32_t s = slot0(id);
        for (;;) {
            const uint32_t h = hash[s];
            if (h == 0) return -1;
            if ((key[h - 1] & ~INACTIVE) == id) return (int)h - 1;
            s = (s + 1) & hmask;
        }
    }
    // slot that holds record `rec` (it is in the index)
    __device__ __forceinline__ uint32_t slot_of(int rec) const {
        uint32_t s = slot0(key[rec] & ~INACTIVE);
        while (hash[s] != (HT)(rec + 1)) s = (s + 1) & hmask;
        return s;
    }
    __device__ __forceinline__ void insert(uint32_t id, int rec) {
        uint32_t s = slot0(id);
        for (;;) {
            if (cas_slot(&hash[s], (HT)(rec + 1)) == 0) return;
            s = (s + 1) & hmask;
        }
    }
};

// Runs one pass (DirectedWeighted.doWork with reverse = (pass==1), or UndirectedWeighted.doWork)
// on table T holding n records; returns the new n, or -1 on table overflow.
template <typename HT>
__device__ int bca_pass(const BcaParams& P, Table<HT>& T, int n, int pass, int lane, int& npop,
                        unsigned long long& pops, unsigned long long& scanned, unsigned long long& fetched) {
    const unsigned FULL = 0xffffffffu;
    const uint8_t bit = (uint8_t)(1 << pass);
    const unsigned lt = (1u << lane) - 1;
    // the frontier holds the bookmark alone (record 0) when a pass starts (:39)
    if (lane == 0) T.act[0] = 0;
    int nact = 1;
    __syncwarp();
    while (nact > 0) {
        // pollFirstEntry(): smallest node id in the frontier (DirectedWeighted.java:48)
        uint32_t best = 0xffffffffu;
        int bpos = -1;
        for (int c = lane; c < nact; c += 32) {
            const uint32_t k = T.key[T.act[c]];
            if (k < best) { best = k; bpos = c; }
        }
        const uint32_t m = __reduce_min_sync(FULL, best);
        const unsigned who = __ballot_sync(FULL, best == m);
        const int pos = __shfl_sync(FULL, bpos, __ffs(who) - 1);
        const int idx = (int)T.act[pos];
        const int32_t focus = (int32_t)m;
        const double paint = T.wet[idx];
        const bool first_pop = T.flg[idx] == 0;
        const HT last = T.act[nact - 1];
        __syncwarp();
        if (lane == 0) {
            if (first_pop) T.ord[npop] = (HT)idx;
            T.act[pos] = last;  // remove from the frontier
            T.wet[idx] = 0.0;
            T.key[idx] = m | INACTIVE;
            // bcv.add(focusNode, alpha * wetPaint) (:53) -> BCV.add(int,float) :35-37
            const float v = (float)__dmul_rn(P.alpha, paint);
            float* acc = pass ? T.accr : T.accf;
            const uint8_t f = T.flg[idx];
            acc[idx] = (f & bit) ? __fadd_rn(acc[idx], v) : __fadd_rn(0.0f, v);
            T.flg[idx] = f | bit;
        }
        nact--;
        pops++;
        npop += first_pop ? 1 : 0;
        __syncwarp();
        if (paint < P.eps) continue;  // :56
        const double base = __dmul_rn(P.one_minus_alpha, paint);
        const int nlists = P.directed ? 1 : 2;
        for (int li = 0; li < nlists; li++) {
            const DirGraph& g = P.directed ? (pass ? P.in : P.out) : (li ? P.in : P.out);
            const int64_t b = g.off[focus], e = g.off[focus + 1];
            if (e == b) continue;  // :66 (undirected: empty list contributes nothing)
            const double total = g.total[focus];
            scanned += (unsigned long long)(e - b);
            if (P.directed && total == 0.0) continue;  // :77
            // No neighbour can reach epsilon when even the heaviest edge cannot: division and
            // multiplication by a positive constant are monotone under round-to-nearest, so
            // partial <= pmax for every edge.  (NaN compares false -> falls through.)
            const double pmax = __dmul_rn(base, __ddiv_rn((double)g.wmax[focus], total));
            if (pmax < P.eps) continue;
            fetched += (unsigned long long)(e - b);  // entries this kernel really reads (scanned counts the reference's)
            for (int64_t k0 = b; k0 < e; k0 += 32) {
                const int64_t k = k0 + lane;
                bool take = false;
                uint32_t nb = 0;
                double partial = 0.0;
                if (k < e) {
                    nb = (uint32_t)g.nbr[k];
                    const float w = g.w[k];
                    // (1 - alpha) * wetPaint * (weight / totalWeight)  (:82), left to right
                    partial = __dmul_rn(base, __ddiv_rn((double)w, total));
                    take = !(partial < P.eps);  // :85
                }
                int rec = -1;
                bool wake = false;  // known node that is not in the frontier: put() after an earlier pop (:92-96)
                if (take) {
                    rec = T.find(nb);
                    if (rec >= 0) {  // :89-91 addPaint, or :92-96
                        wake = (T.key[rec] & INACTIVE) != 0;
                        T.wet[rec] = __dadd_rn(T.wet[rec], partial);
                        T.key[rec] = nb;
                    }
                }
                const unsigned newmask = __ballot_sync(FULL, take && rec < 0);
                const int cnt = __popc(newmask);
                if (cnt) {
                    if (n + cnt > T.cap) return -1;
                    if (take && rec < 0) {
                        rec = n + __popc(newmask & lt);
                        T.key[rec] = nb;
                        T.wet[rec] = partial;
                        T.accf[rec] = 0.f;
                        T.accr[rec] = 0.f;
                        T.flg[rec] = 0;
                        T.insert(nb, rec);
                        wake = true;
                    }
                    n += cnt;
                }
                // neighbour lists are sorted and unique: no two lanes of a batch see the same record
                const unsigned wakemask = __ballot_sync(FULL, wake);
                if (wake) T.act[nact + __popc(wakemask & lt)] = (HT)rec;
                nact += __popc(wakemask);
                __syncwarp();
            }
        }
    }
    return n;
}

// MINB = resident blocks per SM the register allocation must allow (tier 2 trades a few spilled
// registers for more warps in flight: 16 -> 64 registers, 24 -> 40, 32 -> 32)
template <typename HT, bool SMEM, int MINB = 1>
__global__ void __launch_bounds__(T1_WARPS * 32, MINB) bca_push(BcaParams P) {
    extern __shared__ __align__(16) uint8_t smem[];
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int hcap = 1 << P.hcap_log2;
    Table<HT> T;
    {
        uint8_t* p;
        if (SMEM) {
            const size_t per = (size_t)P.cap * (8 + 4 + 4 + 4 + 1 + 2 * sizeof(HT)) + (size_t)hcap * sizeof(HT);
            p = smem + (size_t)wib * ((per + 15) & ~(size_t)15);
        } else {
            const size_t gw = (size_t)blockIdx.x * (blockDim.x >> 5) + wib;
            if (gw >= (size_t)P.ntables) return;  // grid rounds up to whole blocks: no table behind this warp
            p = P.scratch + gw * P.scratch_per_warp;
        }
        T.wet = reinterpret_cast<double*>(p); p += (size_t)P.cap * 8;
        T.key = reinterpret_cast<uint32_t*>(p); p += (size_t)P.cap * 4;
        T.accf = reinterpret_cast<float*>(p); p += (size_t)P.cap * 4;
        T.accr = reinterpret_cast<float*>(p); p += (size_t)P.cap * 4;
        T.hash = reinterpret_cast<HT*>(p); p += (size_t)hcap * sizeof(HT);
        T.ord = reinterpret_cast<HT*>(p); p += (size_t)P.cap * sizeof(HT);
        T.act = reinterpret_cast<HT*>(p); p += (size_t)P.cap * sizeof(HT);
        T.flg = p;
        T.cap = P.cap; T.hshift = 32 - P.hcap_log2; T.hmask = hcap - 1;
    }
    const int hwords = (int)((size_t)hcap * sizeof(HT) / 4);
    uint32_t* hwipe = reinterpret_cast<uint32_t*>(T.hash);
    for (int s = lane; s < hwords; s += 32) hwipe[s] = 0;
    __syncwarp();

    unsigned long long pops = 0, scanned = 0, fetched = 0;
    float wmaxv = 0.f;
    for (;;) {
        int wi = 0;
        if (lane == 0) wi = atomicAdd(P.next, 1);
        wi = __shfl_sync(FULL, wi, 0);
        if (wi >= P.nwork) break;
        const int32_t bookmark = P.work ? P.work[wi] : ge_vid_inv(P.first + wi, P.lay_nb, P.lay_Bs);
        if (bookmark >= P.V) continue;  // padding slot of a cyclic block
        const int32_t slot = ge_vid(bookmark, P.lay_nb, P.lay_Bs) - P.row_base;

        // nodeTree.put(bookmark, new PaintedNode(bookmark, 1)) (:39)
        if (lane == 0) {
            T.key[0] = (uint32_t)bookmark; T.wet[0] = 1.0; T.accf[0] = 0.f; T.accr[0] = 0.f; T.flg[0] = 0;
            T.insert((uint32_t)bookmark, 0);
        }
        __syncwarp();
        unsigned long long bp = 0, bs = 0, bf = 0;
        int npop = 0;
        int n = bca_pass(P, T, 1, 0, lane, npop, bp, bs, bf);
        if (n > 0 && P.directed) {  // BCAJob.call :32-36 — reverse pass, merged below
            if (lane == 0) { T.key[0] = (uint32_t)bookmark; T.wet[0] = 1.0; }
            __syncwarp();
            n = bca_pass(P, T, n, 1, lane, npop, bp, bs, bf);
        }
        if (n < 0) {  // table overflow: hand the bookmark to the next tier
            if (lane == 0) P.overflow[atomicAdd(P.n_overflow, 1)] = bookmark;
            for (int s = lane; s < hwords; s += 32) hwipe[s] = 0;
            __syncwarp();
            continue;
        }
        pops += bp; scanned += bs; fetched += bf;

        // ---- BCV.merge (:105-107): fwd + rev per key; then normalise (BookmarkColoring.java:81-91)
        // value of record r:
        auto value = [&](int q) -> float {  // q-th BCV entry in creation order
            const int r = (int)T.ord[q];
            const uint8_t f = T.flg[r];
            if (!P.directed) return T.accf[r];
            return (f == 3) ? __fadd_rn(T.accf[r], T.accr[r]) : ((f & 1) ? T.accf[r] : T.accr[r]);
        };
        int first_out = 0;  // records [first_out, n) are emitted
        float scale_den = 1.f, usum = 1.f;
        if (P.normalize == GE_NORM_COUNTS) {  // BCV.toCounts :52-59
            float mx = -INFINITY, mn = INFINITY;
            for (int r = lane; r < n; r += 32) { const float v = value(r); mx = fmaxf(mx, v); mn = fminf(mn, v); }
            for (int o = 16; o > 0; o >>= 1) {
                mx = fmaxf(mx, __shfl_xor_sync(FULL, mx, o));
                mn = fminf(mn, __shfl_xor_sync(FULL, mn, o));
            }
            scale_den = __fdiv_rn(__fsub_rn(mx, mn), __fsub_rn(1000.0f, 1.0f));
            first_out = 1;  // remove(rootNode)
        } else if (P.normalize == GE_NORM_UNITY) {  // BCV.toUnity :64-70, sum in discovery order
            float s = 0.f;
            if (lane == 0) {
                for (int r = 1; r < n; r++) s = (r == 1) ? value(r) : __fadd_rn(s, value(r));
            }
            usum = __shfl_sync(FULL, s, 0);
            first_out = 1;
        }
        const int cnt = n - first_out;
        long long start = -1;
        if (lane == 0) {
            const unsigned long long at = atomicAdd(P.cursor, (unsigned long long)cnt);
            if ((long long)(at + cnt) <= P.arena_cap) start = (long long)at;
            else P.deferred[atomicAdd(P.n_deferred, 1)] = bookmark;
        }
        start = __shfl_sync(FULL, start, 0);
        float rmax = -INFINITY;
        for (int r = first_out + lane; r < n; r += 32) {
            float v = value(r);
            if (P.normalize == GE_NORM_COUNTS) v = __fadd_rn(__fdiv_rn(v, scale_den), 1.0f);
            else if (P.normalize == GE_NORM_UNITY) v = __fsub_rn(__fdiv_rn(v, usum), 1e-6f);
            rmax = fmaxf(rmax, v);
            if (start >= 0) {
                const long long o = start + (r - first_out);
                P.aI[o] = bookmark;
                P.aJ[o] = (int32_t)(T.key[T.ord[r]] & ~INACTIVE);
                P.aX[o] = v;
            }
        }
        for (int o = 16; o > 0; o >>= 1) rmax = fmaxf(rmax, __shfl_xor_sync(FULL, rmax, o));
        if (cnt == 0) rmax = 1.0f;  // BCV.max() of an empty vector (:82-84)
        if (start >= 0) {
            wmaxv = fmaxf(wmaxv, rmax);
            if (lane == 0) {
                P.row_start[slot] = P.arena_base + start;
                P.row_cnt[slot] = cnt;
            }
        } else {
            pops -= bp; scanned -= bs; fetched -= bf;  // will be recounted when the row is re-run
        }
        // wipe the index.  Tier 2's index is 256 KB per warp (65 536 slots for the 20 002-record bound): zeroing
        // all of it after every row wrote 0.5 TB at C4 — clear the n used slots instead (found read-only first,
        // because clearing while other lanes still probe would cut their chains; `act` is free by now)
        if (SMEM) {
            for (int s = lane; s < hwords; s += 32) hwipe[s] = 0;
        } else {
            for (int r = lane; r < n; r += 32) T.act[r] = (HT)T.slot_of(r);
            __syncwarp();
            for (int r = lane; r < n; r += 32) T.hash[T.act[r]] = 0;
        }
        __syncwarp();
    }
    if (lane == 0) {
        if (pops) atomicAdd(P.pops, pops);
        if (scanned) atomicAdd(P.scanned, scanned);
        if (fetched) atomicAdd(P.fetched, fetched);
        if (wmaxv > 0.f) atomicMax(P.gmax_bits, __float_as_int(wmaxv));
    }
}

// gather rows from the arenas into CSR order (row r of the range -> [off[r], off[r+1]))
struct ArenaSet {
    const int32_t* I[MAX_ARENAS];
    const int32_t* J[MAX_ARENAS];
    const float* X[MAX_ARENAS];
    int64_t cap;  // all arenas have the same capacity; unified index = arena * cap + offset
};

__global__ void compact_rows(ArenaSet A, const int64_t* __restrict__ row_start, const int64_t* __restrict__ row_off,
                             int32_t nrows, int32_t* __restrict__ I, int32_t* __restrict__ J, float* __restrict__ X) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < nrows; r += nwarps) {
        const int64_t src = row_start[r], dst = row_off[r];
        const int32_t cnt = (int32_t)(row_off[r + 1] - dst);
        const int a = (int)(src / A.cap);
        const int64_t o = src - (int64_t)a * A.cap;
        for (int32_t k = lane; k < cnt; k += 32) {
            I[dst + k] = A.I[a][o + k];
            J[dst + k] = A.J[a][o + k];
            X[dst + k] = A.X[a][o + k];
        }
    }
}

// exclusive scan of row counts -> row offsets (one block of 1024 threads walks the array)
__global__ void scan_counts(const int32_t* __restrict__ cnt, int32_t n, int64_t* __restrict__ off) {
    __shared__ long long warp_excl[32];
    __shared__ long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = blockDim.x >> 5;
    for (int64_t base = 0; base < n; base += blockDim.x) {
        const int64_t i = base + threadIdx.x;
        const long long v = i < n ? cnt[i] : 0;
        long long x = v;  // inclusive scan within the warp
        for (int o = 1; o < 32; o <<= 1) {
            const long long y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_excl[wid] = x;
        __syncthreads();
        if (wid == 0) {
            const long long t = lane < nw ? warp_excl[lane] : 0;
            long long s = t;
            for (int o = 1; o < 32; o <<= 1) {
                const long long y = __shfl_up_sync(0xffffffffu, s, o);
                if (lane >= o) s += y;
            }
            warp_excl[lane] = s - t;
        }
        __syncthreads();
        const long long excl = carry + warp_excl[wid] + (x - v);
        if (i < n) off[i] = excl;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) off[n] = carry;
}

}  // namespace

// ==========================================================================================
// host driver
// ==========================================================================================
namespace {
struct DevBuf {  // frees on scope exit
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    template <typename T> T* as() { return reinterpret_cast<T*>(p); }
};
}  // namespace

int ge_bca_build(ge_ctx* ctx, int32_t first, int32_t count, int32_t lay_nb, int32_t lay_Bs) {
    const ge_config& c = ctx->cfg;
    cudaStream_t st = ctx->stream;
    BcaParams P{};
    P.V = ctx->V;
    P.out = DirGraph{ctx->d_out_off, ctx->d_out_nbr, ctx->d_out_w, ctx->d_out_total, ctx->d_out_wmax};
    P.in = DirGraph{ctx->d_in_off, ctx->d_in_nbr, ctx->d_in_w, ctx->d_in_total, ctx->d_in_wmax};
    P.alpha = c.alpha; P.eps = c.epsilon; P.one_minus_alpha = 1 - c.alpha;  // Java: (1 - alpha), fp64
    P.directed = c.directed; P.normalize = c.normalize;
    P.first = first; P.row_base = first;
    P.lay_nb = lay_nb; P.lay_Bs = lay_Bs;

    // bound on records per bookmark: every pop but the root retires >= alpha*eps paint per pass
    double bound = 2.0 / (c.alpha * c.epsilon) + 2.0;
    if (bound > (double)ctx->V) bound = (double)ctx->V;
    const int64_t cap2 = (int64_t)bound < 64 ? 64 : (int64_t)bound;

    DevBuf row_start, row_cnt, overflow, deferred, counters;
    GE_CUDA(cudaMalloc(&row_start.p, (size_t)count * 8));
    GE_CUDA(cudaMalloc(&row_cnt.p, (size_t)count * 4));
    GE_CUDA(cudaMalloc(&overflow.p, (size_t)count * 4));
    GE_CUDA(cudaMalloc(&deferred.p, (size_t)count * 4));
    // counters: [0] next, [1] n_overflow, [2] n_deferred, [3] gmax bits, then u64: cursor, pops, scanned, fetched
    GE_CUDA(cudaMalloc(&counters.p, 64));
    GE_CUDA(cudaMemsetAsync(counters.p, 0, 64, st));
    GE_CUDA(cudaMemsetAsync(row_start.p, 0xff, (size_t)count * 8, st));
    GE_CUDA(cudaMemsetAsync(row_cnt.p, 0, (size_t)count * 4, st));
    int32_t* d32 = counters.as<int32_t>();
    unsigned long long* d64 = reinterpret_cast<unsigned long long*>(d32 + 4);
    P.next = d32; P.n_overflow = d32 + 1; P.n_deferred = d32 + 2; P.gmax_bits = d32 + 3;
    P.cursor = d64; P.pops = d64 + 1; P.scanned = d64 + 2; P.fetched = d64 + 3;
    P.row_start = row_start.as<int64_t>(); P.row_cnt = row_cnt.as<int32_t>();
    P.overflow = overflow.as<int32_t>(); P.deferred = deferred.as<int32_t>();

    // arena capacity: a share of free memory, never more than count * cap2
    size_t free_b = 0, total_b = 0;
    GE_CUDA(cudaMemGetInfo(&free_b, &total_b));
    int64_t arena_cap = (int64_t)((double)free_b * 0.30 / 12.0);
    const double ub = (double)count * (double)cap2;
    if ((double)arena_cap > ub) arena_cap = (int64_t)ub;
    if (arena_cap < 1024) arena_cap = 1024;

    // tier-1 launch geometry
    // records per warp in tier 1: fewer records = more resident warps for this latency-bound kernel, but more
    // rows that overflow into tier 2 and run twice (GE_BCA_T1_CAP, sweep in profiles/r02_bca_t1cap.md)
    int t1_cap = getenv("GE_BCA_T1_CAP") ? atoi(getenv("GE_BCA_T1_CAP")) : T1_CAP;
    t1_cap = t1_cap < 32 ? 32 : (t1_cap > 2040 ? 2040 : t1_cap);
    t1_cap &= ~7;   // keeps every array of the table 8-byte aligned
    int hcap1_log2 = 6;
    while ((1 << hcap1_log2) < 2 * t1_cap) hcap1_log2++;   // index at most half full
    const size_t per1 = (((size_t)t1_cap * 25 + ((size_t)1 << hcap1_log2) * 2) + 15) & ~(size_t)15;
    const size_t smem1 = per1 * T1_WARPS;
    GE_CUDA(cudaFuncSetAttribute(bca_push<uint16_t, true, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem1));
    int bps1 = 1;
    GE_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps1, bca_push<uint16_t, true, 16>, T1_WARPS * 32, smem1));
    if (bps1 < 1) bps1 = 1;

    ArenaSet AS{};
    AS.cap = arena_cap;
    std::vector<void*> arenas;
    auto free_arenas = [&]() { for (void* a : arenas) cudaFree(a); arenas.clear(); };

    GE_CUDA(cudaEventRecord(ctx->ev0, st));
    DevBuf work;  // deferred rows of the previous round
    int32_t nwork = count;
    int64_t overflow_rows = 0;
    int round = 0;
    int rc = GE_OK;
    for (; nwork > 0; round++) {
        if (round >= MAX_ARENAS) { ctx->err = "co-occurrence matrix does not fit in device memory"; rc = GE_ERR_NOMEM; break; }
        void *aI = nullptr, *aJ = nullptr, *aX = nullptr;
        if (cudaMalloc(&aI, (size_t)arena_cap * 4) != cudaSuccess || cudaMalloc(&aJ, (size_t)arena_cap * 4) != cudaSuccess ||
            cudaMalloc(&aX, (size_t)arena_cap * 4) != cudaSuccess) {
            cudaFree(aI); cudaFree(aJ); cudaFree(aX); cudaGetLastError();
            ctx->err = "out of device memory for the co-occurrence arena"; rc = GE_ERR_NOMEM; break;
        }
        arenas.push_back(aI); arenas.push_back(aJ); arenas.push_back(aX);
        AS.I[round] = (int32_t*)aI; AS.J[round] = (int32_t*)aJ; AS.X[round] = (float*)aX;
        P.aI = (int32_t*)aI; P.aJ = (int32_t*)aJ; P.aX = (float*)aX;
        P.arena_cap = arena_cap; P.arena_base = (int64_t)round * arena_cap;
        if (cudaMemsetAsync(d32, 0, 12, st) != cudaSuccess ||   // next, n_overflow, n_deferred
            cudaMemsetAsync(P.cursor, 0, 8, st) != cudaSuccess) {
            ctx->err = std::string("bca round reset: ") + cudaGetErrorString(cudaGetLastError()); rc = GE_ERR_CUDA; break;
        }
        P.work = round == 0 ? nullptr : work.as<int32_t>();
        P.nwork = nwork;
        // ---- tier 1: shared-memory tables
        P.cap = t1_cap; P.hcap_log2 = hcap1_log2; P.scratch = nullptr; P.scratch_per_warp = 0;
        const auto tick0 = std::chrono::steady_clock::now();
        {
            int64_t blocks = ((int64_t)nwork + T1_WARPS - 1) / T1_WARPS;
            const int64_t full = (int64_t)ctx->sm_count * bps1;
            if (blocks > full) blocks = full;
            bca_push<uint16_t, true, 16><<<(unsigned)blocks, T1_WARPS * 32, smem1, st>>>(P);
        }
        int32_t h[3];
        if (cudaMemcpyAsync(h, d32, 12, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
            ctx->err = std::string("bca_push tier 1: ") + cudaGetErrorString(cudaGetLastError()); rc = GE_ERR_CUDA; break;
        }
        int32_t n_over = h[1];
        const auto tick1 = std::chrono::steady_clock::now();
        int32_t tier_rows[2] = {0, 0};
        double tier_ms[2] = {0, 0};
        // ---- global-memory tiers for the rows that overflowed: a mid tier with small tables (they stay
        // L2-resident: 64 warps/SM x 36 KB), then tables sized by the bound 2/(alpha*eps) for the few that remain
        int64_t mid_cap = getenv("GE_BCA_MID_CAP") ? atoll(getenv("GE_BCA_MID_CAP")) : 0;
        if (mid_cap >= cap2 || mid_cap <= t1_cap) mid_cap = 0;   // 0 = no mid tier
        overflow_rows += n_over;
        for (int tier = mid_cap ? 0 : 1; tier < 2 && n_over > 0 && rc == GE_OK; tier++) {
            const auto tk = std::chrono::steady_clock::now();
            const int64_t capt = tier == 0 ? mid_cap : cap2;
            int hl = 1;
            while (((int64_t)1 << hl) < 2 * capt) hl++;
            const size_t per2 = (((size_t)capt * 29 + ((size_t)1 << hl) * 4) + 15) & ~(size_t)15;
            // resident warps per SM (GE_BCA_T2_WARPS): 64 registers allow 32, 40 allow 48, 32 allow 64 — the kernel
            // is bound by the latency of its dependent loads, and a few spilled registers cost less than idle
            // issue slots (sweep in profiles/r02_bca_tiers.md)
            int t2_per_sm = getenv("GE_BCA_T2_WARPS") ? atoi(getenv("GE_BCA_T2_WARPS")) : 64;
            t2_per_sm = t2_per_sm < 2 ? 2 : (t2_per_sm > 64 ? 64 : t2_per_sm);
            int64_t warps = (int64_t)ctx->sm_count * t2_per_sm;
            if (warps > n_over) warps = n_over;
            size_t fb = 0, tb = 0;
            cudaMemGetInfo(&fb, &tb);
            while (warps > 1 && (double)warps * (double)per2 > (double)fb * 0.5) warps /= 2;
            DevBuf scratch, list;
            if (cudaMalloc(&scratch.p, (size_t)warps * per2) != cudaSuccess || cudaMalloc(&list.p, (size_t)n_over * 4) != cudaSuccess) {
                cudaGetLastError(); ctx->err = "out of device memory for global-memory BCA tables"; rc = GE_ERR_NOMEM; break;
            }
            if (cudaMemcpyAsync(list.p, P.overflow, (size_t)n_over * 4, cudaMemcpyDeviceToDevice, st) != cudaSuccess ||
                cudaMemsetAsync(d32, 0, 8, st) != cudaSuccess) {  // next, n_overflow (n_deferred keeps counting)
                ctx->err = std::string("bca global tier set-up: ") + cudaGetErrorString(cudaGetLastError()); rc = GE_ERR_CUDA; break;
            }
            BcaParams Q = P;
            Q.work = list.as<int32_t>(); Q.nwork = n_over;
            Q.cap = (int32_t)capt; Q.hcap_log2 = hl;
            Q.scratch = scratch.as<uint8_t>(); Q.scratch_per_warp = per2; Q.ntables = (int32_t)warps;
            const int64_t blocks = (warps + T1_WARPS - 1) / T1_WARPS;
            if (t2_per_sm <= 32) bca_push<uint32_t, false, 16><<<(unsigned)blocks, T1_WARPS * 32, 0, st>>>(Q);
            else if (t2_per_sm <= 48) bca_push<uint32_t, false, 24><<<(unsigned)blocks, T1_WARPS * 32, 0, st>>>(Q);
            else bca_push<uint32_t, false, 32><<<(unsigned)blocks, T1_WARPS * 32, 0, st>>>(Q);
            if (cudaMemcpyAsync(h, d32, 12, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
                ctx->err = std::string("bca_push global tier: ") + cudaGetErrorString(cudaGetLastError()); rc = GE_ERR_CUDA; break;
            }
            tier_rows[tier] = n_over;
            tier_ms[tier] = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - tk).count();
            n_over = h[1];
            if (tier == 1 && n_over != 0) { ctx->err = "internal: BCA record bound exceeded"; rc = GE_ERR_CUDA; break; }
        }
        if (rc) break;
        const int32_t n_def = h[2];
        if (getenv("GE_DEBUG"))
            fprintf(stderr, "[bca] round %d: %d rows, shared tier (%d records/warp, %d blocks/SM) %.1f ms, mid tier (%lld records) %d rows %.1f ms, "
                            "bound tier %d rows %.1f ms, deferred %d\n",
                    round, nwork, t1_cap, bps1, std::chrono::duration<double, std::milli>(tick1 - tick0).count(), (long long)mid_cap,
                    tier_rows[0], tier_ms[0], tier_rows[1], tier_ms[1], n_def);
        if (n_def > 0) {
            if (!work.p && cudaMalloc(&work.p, (size_t)count * 4) != cudaSuccess) { cudaGetLastError(); rc = GE_ERR_NOMEM; ctx->err = "out of device memory"; break; }
            if (cudaMemcpyAsync(work.p, P.deferred, (size_t)n_def * 4, cudaMemcpyDeviceToDevice, st) != cudaSuccess) {
                ctx->err = std::string("bca deferred list: ") + cudaGetErrorString(cudaGetLastError()); rc = GE_ERR_CUDA; break;
            }
        }
        nwork = n_def;
    }
    if (rc) { free_arenas(); return rc; }

    // ---- row offsets, exact-size final arrays, CSR-ordered compaction
    if (cudaMalloc(&ctx->d_row_off, (size_t)(count + 1) * 8) != cudaSuccess) { cudaGetLastError(); free_arenas(); ctx->err = "out of device memory"; return GE_ERR_NOMEM; }
    scan_counts<<<1, 1024, 0, st>>>(P.row_cnt, count, ctx->d_row_off);
    int64_t nnz = 0;
    unsigned long long h64[4];
    int32_t gbits = 0;
    cudaMemcpyAsync(&nnz, ctx->d_row_off + count, 8, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(h64, d64, 32, cudaMemcpyDeviceToHost, st);
    cudaMemcpyAsync(&gbits, P.gmax_bits, 4, cudaMemcpyDeviceToHost, st);
    if (cudaStreamSynchronize(st) != cudaSuccess) { free_arenas(); ctx->err = std::string("scan: ") + cudaGetErrorString(cudaGetLastError()); return GE_ERR_CUDA; }
    const size_t nn = (size_t)(nnz > 0 ? nnz : 1);
    if (cudaMalloc(&ctx->d_I, nn * 4) != cudaSuccess || cudaMalloc(&ctx->d_J, nn * 4) != cudaSuccess || cudaMalloc(&ctx->d_X, nn * 4) != cudaSuccess) {
        cudaGetLastError(); free_arenas(); ctx->err = "out of device memory for the co-occurrence matrix"; return GE_ERR_NOMEM;
    }
    compact_rows<<<ctx->sm_count * 8, 256, 0, st>>>(AS, P.row_start, ctx->d_row_off, count, ctx->d_I, ctx->d_J, ctx->d_X);
    cudaEventRecord(ctx->ev1, st);
    if (cudaStreamSynchronize(st) != cudaSuccess) { free_arenas(); ctx->err = std::string("compact: ") + cudaGetErrorString(cudaGetLastError()); return GE_ERR_CUDA; }
    free_arenas();
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->bca_seconds = ms * 1e-3;
    ctx->bca_pops = (int64_t)h64[1];
    ctx->bca_scanned = (int64_t)h64[2];
    ctx->bca_fetched = (int64_t)h64[3];
    ctx->bca_overflow_rows = overflow_rows;
    ctx->nnz = nnz;
    float gm;
    memcpy(&gm, &gbits, 4);
    ctx->xmax = (double)gm;  // BookmarkColoring.max starts at 0 (:26) and only grows (:162-164)
    ctx->row_first = first; ctx->row_count = count; ctx->row_lay_nb = lay_nb > 1 ? lay_nb : 0;
    ctx->have_cooc = true;
    return GE_OK;
}
