// sim.cu — similarity edges between literals.
//
// Replaces the compare loop of J/convert/Rdf2GrphConverter.java:127-186: one CompareJob per source
// literal (J/compare/CompareJob.java:32-53) walks every target j >= startIndex, skips the literal's
// own vertex, and keeps (otherVert, (float) similarity) when metric.similarity(s1, s2) >= threshold.
// Here the group's literals arrive as the numbers the metrics read (profiles as sorted id lists,
// parsed ints, LocalDate fields; include/graphemb.h ge_sim_group) and the O(|source| x |target|)
// compare runs on the device:
//
//   pass 1  sim_pass<false>  one warp per (source row, segment of SEG targets): lane = one target,
//                            hits counted with ballot; rows whose job would have died are flagged
//   scan                     segment counts -> 64-bit offsets (order = (row, target index))
//   pass 2  sim_pass<true>   same walk, hits written at offset + rank inside the warp
//
// Integer work (intersection sizes, dot products, date differences) is exact; the one division /
// sqrt per pair uses IEEE fp64 round-to-nearest like Java; Math.pow is CUDA's pow (<= 2 ulp).
#include "common.cuh"

#include <math.h>

namespace {

constexpr int SEG = 1024;       // targets per work item
constexpr int SCAN_TILE = 1024; // counts per scan block (256 threads x 4)

struct SimDev {
    int32_t method, upper, time_dir, n_cut;
    double threshold, smooth, distance;
    const int32_t *vert, *sid;
    const uint8_t* flags;
    const int64_t* prof_off;
    const int32_t *prof_key, *prof_cnt;
    const double* norm;
    const int32_t *cut_slot, *num_val;
    const uint8_t* num_status;
    const int64_t *epoch_day, *packed_month;
    const int32_t *source, *target;
    int32_t n_source, n_target, nseg;
    // JAROWINKLER / LEVENSHTEIN: labels as UTF-16 code units + per-thread scratch words, interleaved
    // by thread (word k of thread t at scratch[k * nthreads + t]) so that lanes in step coalesce
    const int64_t* char_off;
    const uint16_t* chars;
    uint32_t* scratch;
    int64_t nthreads;
};

// Euclidean norm of a profile: Math.sqrt(sum of Math.pow(v, 2)) (TokenCosine.java:30-32,
// PreComputedNgramCosine.java:36-38); the squares are integers, so the sum is order-independent
__global__ void profile_norms(const int64_t* __restrict__ off, const int32_t* __restrict__ cnt, int32_t n,
                              double* __restrict__ norm) {
    const int32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    double s = 0.0;
    for (int64_t p = off[k]; p < off[k + 1]; p++) s = __dadd_rn(s, __dmul_rn((double)cnt[p], (double)cnt[p]));
    norm[k] = __dsqrt_rn(s);
}

// ---- info.debatty java-string-similarity (pom.xml:50-51, version "RELEASE"; restated from the
// published 2.0.0 source, not under /root/reference) ------------------------------------------------

// NormalizedLevenshtein.similarity = 1.0 - Levenshtein.distance(s1, s2) / max(len1, len2) (double / int);
// the distance is the textbook unit-cost one, so any exact evaluation gives the reference's value.
// One DP row per thread in scratch (in-place update with the diagonal carried in a register).
__device__ __forceinline__ double levenshtein_similarity(const SimDev& g, int32_t a, int32_t b, int64_t tid) {
    const uint16_t *s1 = g.chars + g.char_off[a], *s2 = g.chars + g.char_off[b];
    const int32_t n1 = (int32_t)(g.char_off[a + 1] - g.char_off[a]), n2 = (int32_t)(g.char_off[b + 1] - g.char_off[b]);
    const int32_t mx = n1 > n2 ? n1 : n2, gap = n1 > n2 ? n1 - n2 : n2 - n1;
    // distance >= |n1 - n2| and every step below is monotone under correct rounding: skip the DP
    // when even that bound misses the threshold
    if (!(__dsub_rn(1.0, __ddiv_rn((double)gap, (double)mx)) >= g.threshold)) return __longlong_as_double(0x7ff8000000000000LL);
    int32_t dist;
    if (n1 == 0) dist = n2;
    else if (n2 == 0) dist = n1;
    else {
        uint32_t* row = g.scratch + tid;
        const int64_t st = g.nthreads;
        for (int32_t j = 0; j <= n2; j++) row[(int64_t)j * st] = (uint32_t)j;
        for (int32_t i = 0; i < n1; i++) {
            const uint16_t c1 = __ldg(s1 + i);
            uint32_t diag = (uint32_t)i, left = (uint32_t)i + 1;  // v0[0], v1[0]
            row[0] = left;
            for (int32_t j = 0; j < n2; j++) {
                const uint32_t up = row[(int64_t)(j + 1) * st];
                const uint32_t sub = diag + (c1 == __ldg(s2 + j) ? 0u : 1u);
                uint32_t v = left + 1 < up + 1 ? left + 1 : up + 1;
                v = v < sub ? v : sub;
                row[(int64_t)(j + 1) * st] = v;
                diag = up;
                left = v;
            }
        }
        dist = (int32_t)row[(int64_t)n2 * st];
    }
    return __dsub_rn(1.0, __ddiv_rn((double)dist, (double)mx));
}

// JaroWinkler.similarity: matches() pairs every character of the shorter string with the first free
// equal character of the longer one inside the window max(len/2 - 1, 0); transpositions = half the
// matched characters that differ in order; prefix = common prefix (not capped).  The Jaro term is
// evaluated in FLOAT (m is a float in the source), the Winkler boost in double.
__device__ __forceinline__ double jaro_winkler_similarity(const SimDev& g, int32_t a, int32_t b, int64_t tid) {
    const uint16_t *s1 = g.chars + g.char_off[a], *s2 = g.chars + g.char_off[b];
    const int32_t n1 = (int32_t)(g.char_off[a + 1] - g.char_off[a]), n2 = (int32_t)(g.char_off[b + 1] - g.char_off[b]);
    const uint16_t *mx = n1 > n2 ? s1 : s2, *mn = n1 > n2 ? s2 : s1;   // s1.length() > s2.length() ? s1 : s2
    const int32_t nmx = n1 > n2 ? n1 : n2, nmn = n1 > n2 ? n2 : n1;
    if (nmn == 0) return 0.0;  // no matches: return 0f
    const int32_t range = nmx / 2 - 1 > 0 ? nmx / 2 - 1 : 0;
    const int64_t st = g.nthreads;
    uint32_t* bmax = g.scratch + tid;                                      // match_flags over max, one bit each
    uint32_t* bmin = g.scratch + tid + (int64_t)((nmx + 31) >> 5) * st;    // match_indexes[i] != -1 over min
    for (int32_t k = 0; k < (nmx + 31) >> 5; k++) bmax[(int64_t)k * st] = 0u;
    for (int32_t k = 0; k < (nmn + 31) >> 5; k++) bmin[(int64_t)k * st] = 0u;
    int32_t matches = 0;
    for (int32_t mi = 0; mi < nmn; mi++) {
        const uint16_t c1 = __ldg(mn + mi);
        const int32_t lo = mi - range > 0 ? mi - range : 0, hi = mi + range + 1 < nmx ? mi + range + 1 : nmx;
        for (int32_t xi = lo; xi < hi; xi++) {
            const uint32_t w = bmax[(int64_t)(xi >> 5) * st];
            if (!((w >> (xi & 31)) & 1u) && c1 == __ldg(mx + xi)) {
                bmax[(int64_t)(xi >> 5) * st] = w | (1u << (xi & 31));
                bmin[(int64_t)(mi >> 5) * st] |= 1u << (mi & 31);
                matches++;
                break;
            }
        }
    }
    if (matches == 0) return 0.0;
    int32_t differ = 0, xi = 0;  // walk the matched characters of both strings in order
    for (int32_t mi = 0; mi < nmn; mi++) {
        if (!((bmin[(int64_t)(mi >> 5) * st] >> (mi & 31)) & 1u)) continue;
        while (!((bmax[(int64_t)(xi >> 5) * st] >> (xi & 31)) & 1u)) xi++;
        if (__ldg(mn + mi) != __ldg(mx + xi)) differ++;
        xi++;
    }
    int32_t prefix = 0;
    for (int32_t k = 0; k < nmn; k++) { if (__ldg(s1 + k) == __ldg(s2 + k)) prefix++; else break; }
    const float m = (float)matches;
    const float jf = __fdiv_rn(__fadd_rn(__fadd_rn(__fdiv_rn(m, (float)n1), __fdiv_rn(m, (float)n2)),
                                         __fdiv_rn(__fsub_rn(m, (float)(differ / 2)), m)), 3.0f);
    const double j = (double)jf;
    if (!(j > 0.7)) return j;
    const double coef = fmin(0.1, __ddiv_rn(1.0, (double)nmx));
    return __dadd_rn(j, __dmul_rn(__dmul_rn(coef, (double)prefix), __dsub_rn(1.0, j)));
}

// metric.similarity(label[a], label[b]); oob = the reference would have thrown out of the job
__device__ __forceinline__ double pair_similarity(const SimDev& g, int32_t a, int32_t b, bool& oob, int64_t tid) {
    const uint8_t fa = g.flags[a], fb = g.flags[b];
    const bool same = g.sid[a] == g.sid[b];
    if (g.method >= GE_SIM_JAROWINKLER) {
        if (same) return 1.0;  // s1.equals(s2)
        return g.method == GE_SIM_LEVENSHTEIN ? levenshtein_similarity(g, a, b, tid) : jaro_winkler_similarity(g, a, b, tid);
    }
    if (g.method <= GE_SIM_TOKEN_JACCARD) {
        if (same) return 1.0;  // s1.equals(s2)
        const bool jaccard = g.method == GE_SIM_NGRAM_JACCARD || g.method == GE_SIM_TOKEN_JACCARD;
        if (g.method == GE_SIM_NGRAM_COSINE && !((fa & GE_SIM_FLAG_LONG) && (fb & GE_SIM_FLAG_LONG))) return 0.0;
        int64_t pa = g.prof_off[a], pb = g.prof_off[b];
        const int64_t ea = g.prof_off[a + 1], eb = g.prof_off[b + 1];
        const int32_t na = (int32_t)(ea - pa), nb = (int32_t)(eb - pb);
        if (jaccard) {
            // inter <= min and union >= max, and a correctly rounded quotient is monotone: when
            // even min/max misses the threshold the exact value does too (0/0 = NaN misses as well)
            const int32_t lo = na < nb ? na : nb, hi = na < nb ? nb : na;
            if (!(__ddiv_rn((double)lo, (double)hi) >= g.threshold)) return __longlong_as_double(0x7ff8000000000000LL);
        }
        int32_t inter = 0;
        uint32_t dot = 0;  // Java int arithmetic wraps
        while (pa < ea && pb < eb) {
            const int32_t ka = __ldg(g.prof_key + pa), kb = __ldg(g.prof_key + pb);
            if (ka == kb) {
                inter++;
                dot += (uint32_t)__ldg(g.prof_cnt + pa) * (uint32_t)__ldg(g.prof_cnt + pb);
                pa++; pb++;
            } else if (ka < kb) pa++;
            else pb++;
        }
        if (jaccard) return __ddiv_rn((double)inter, (double)(na + nb - inter));
        return __ddiv_rn((double)(int32_t)dot, __dmul_rn(g.norm[a], g.norm[b]));
    }
    if ((fa | fb) & GE_SIM_FLAG_EMPTY) return 0.0;  // Numeric.java:26, Date.java:38
    if (same) return 1.0;
    double base;
    if (g.method == GE_SIM_NUMERIC) {
        const int32_t slot = g.cut_slot[a];  // both labels are cut at s1's '^' (Numeric.java:29-33)
        const uint8_t sb = g.num_status[(int64_t)b * g.n_cut + slot];
        if (sb == GE_SIM_NUM_OOB) { oob = true; return 0.0; }
        if (g.num_status[(int64_t)a * g.n_cut + slot] != GE_SIM_NUM_OK || sb != GE_SIM_NUM_OK) return 0.0;
        const uint32_t d = (uint32_t)g.num_val[(int64_t)a * g.n_cut + slot] - (uint32_t)g.num_val[(int64_t)b * g.n_cut + slot];
        int32_t ad = (int32_t)d;
        if (ad < 0) ad = (int32_t)(0u - d);  // Math.abs(int): MIN_VALUE stays negative
        base = (double)ad;
    } else {
        if (!((fa & GE_SIM_FLAG_PARSED) && (fb & GE_SIM_FLAG_PARSED))) return 0.0;  // DateTimeParseException
        const int64_t d1 = g.epoch_day[a], d2 = g.epoch_day[b];
        if (g.time_dir == GE_SIM_TIME_BACKWARDS && d1 > d2) return 0.0;  // d1.isAfter(d2)
        if (g.time_dir == GE_SIM_TIME_FORWARDS && d1 < d2) return 0.0;   // d1.isBefore(d2)
        int64_t between;
        if (g.method == GE_SIM_DATE_DAYS) between = d2 - d1;
        else {
            between = (g.packed_month[b] - g.packed_month[a]) / 32;  // LocalDate.monthsUntil
            if (g.method == GE_SIM_DATE_YEARS) between /= 12;
        }
        base = fabs((double)between);
    }
    return pow(__dadd_rn(fabs(__dsub_rn(base, g.distance)), 1.0), __dsub_rn(g.smooth, 1.0));
}

template <bool FILL>
__global__ void __launch_bounds__(256) sim_pass(SimDev g, uint32_t* __restrict__ seg_count,
                                                const unsigned long long* __restrict__ seg_off,
                                                uint8_t* __restrict__ row_dead, int32_t* __restrict__ out_v,
                                                int32_t* __restrict__ out_o, float* __restrict__ out_s) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int64_t nwork = (int64_t)g.n_source * g.nseg;
    for (int64_t w = warp; w < nwork; w += nwarps) {
        const int32_t r = (int32_t)(w / g.nseg), s = (int32_t)(w - (int64_t)r * g.nseg);
        const int32_t start = g.upper ? r + 1 : 0;  // CompareJob.java:28
        const int32_t lo = s * SEG, hi = min(lo + SEG, g.n_target);
        if (FILL && row_dead[r]) continue;
        if (hi <= start) { if (!FILL && lane == 0) seg_count[w] = 0; continue; }
        const int32_t a = g.source[r];
        const int32_t va = g.vert[a];
        unsigned long long at = FILL ? seg_off[w] : 0ULL;
        uint32_t total = 0;
        bool dead = false;
        for (int32_t j0 = lo; j0 < hi; j0 += 32) {
            const int32_t j = j0 + lane;
            bool hit = false;
            double sim = 0.0;
            int32_t vb = 0;
            if (j < hi && j >= start) {
                const int32_t b = g.target[j];
                vb = g.vert[b];
                if (vb != va) {  // CompareJob.java:40
                    bool oob = false;
                    sim = pair_similarity(g, a, b, oob, (int64_t)blockIdx.x * blockDim.x + threadIdx.x);
                    dead |= oob;
                    hit = !oob && sim >= g.threshold;  // NaN fails, as in Java
                }
            }
            const uint32_t m = __ballot_sync(0xffffffffu, hit);
            if (FILL && hit) {
                const unsigned long long p = at + __popc(m & ((1u << lane) - 1u));
                out_v[p] = va; out_o[p] = vb; out_s[p] = (float)sim;
            }
            at += __popc(m);
            total += __popc(m);
        }
        if (!FILL) {
            if (lane == 0) seg_count[w] = total;
            if (__any_sync(0xffffffffu, dead) && lane == 0) row_dead[r] = 1;
        }
    }
}

// a source row whose job threw keeps nothing (its CompareResult never reaches the consumer loop)
__global__ void drop_dead_rows(uint32_t* __restrict__ seg_count, const uint8_t* __restrict__ row_dead, int32_t nseg, int64_t nwork) {
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w < nwork && row_dead[w / nseg]) seg_count[w] = 0;
}

// ---- exclusive scan u32 -> u64: tile sums, one block over the tile sums, tile rescans ----
__device__ __forceinline__ unsigned long long block_exclusive(unsigned long long v, unsigned long long* total) {
    __shared__ unsigned long long wsum[8];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    unsigned long long x = v;
    for (int o = 1; o < 32; o <<= 1) {
        const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    __syncthreads();  // wsum reuse across calls
    if (lane == 31) wsum[wid] = x;
    __syncthreads();
    unsigned long long before = 0, all = 0;
    for (int k = 0; k < 8; k++) { if (k < wid) before += wsum[k]; all += wsum[k]; }
    *total = all;
    return before + x - v;
}

__global__ void __launch_bounds__(256) scan_tile_totals(const uint32_t* __restrict__ in, int64_t n, unsigned long long* __restrict__ tile_sum) {
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + threadIdx.x * 4;
    unsigned long long v = 0;
    for (int k = 0; k < 4; k++) if (base + k < n) v += in[base + k];
    unsigned long long all;
    block_exclusive(v, &all);
    if (threadIdx.x == 0) tile_sum[blockIdx.x] = all;
}

__global__ void __launch_bounds__(256) scan_tile_sums(unsigned long long* __restrict__ tile_sum, int64_t ntile, unsigned long long* __restrict__ grand) {
    unsigned long long carry = 0;  // one block walks the tile sums 256 at a time
    for (int64_t base = 0; base < ntile; base += 256) {
        const int64_t k = base + threadIdx.x;
        const unsigned long long v = k < ntile ? tile_sum[k] : 0ULL;
        unsigned long long all;
        const unsigned long long ex = block_exclusive(v, &all);
        if (k < ntile) tile_sum[k] = carry + ex;
        carry += all;
    }
    if (threadIdx.x == 0) *grand = carry;
}

__global__ void __launch_bounds__(256) scan_tile_apply(const uint32_t* __restrict__ in, int64_t n, const unsigned long long* __restrict__ tile_sum,
                                                       unsigned long long* __restrict__ out) {
    const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + threadIdx.x * 4;
    uint32_t x[4];
    unsigned long long v = 0;
    for (int k = 0; k < 4; k++) { x[k] = base + k < n ? in[base + k] : 0u; v += x[k]; }
    unsigned long long all;
    unsigned long long ex = tile_sum[blockIdx.x] + block_exclusive(v, &all);
    for (int k = 0; k < 4; k++) { if (base + k < n) out[base + k] = ex; ex += x[k]; }
}

}  // namespace

struct SimState {
    int64_t count = 0;
    int32_t *d_v = nullptr, *d_o = nullptr;
    float* d_s = nullptr;
};

void ge_free_sim(ge_ctx* ctx) {
    if (!ctx->sim) return;
    cudaFree(ctx->sim->d_v); cudaFree(ctx->sim->d_o); cudaFree(ctx->sim->d_s);
    delete ctx->sim;
    ctx->sim = nullptr;
}

namespace {
struct Scratch {  // device buffers of one call, released on every exit path
    std::vector<void*> p;
    ~Scratch() { for (void* q : p) cudaFree(q); }
    template <typename T>
    cudaError_t up(const T* host, size_t n, const T** dev, cudaStream_t st) {
        *dev = nullptr;
        if (!host || n == 0) return cudaSuccess;
        void* d = nullptr;
        cudaError_t e = cudaMalloc(&d, n * sizeof(T));
        if (e != cudaSuccess) return e;
        p.push_back(d);
        *dev = (const T*)d;
        return cudaMemcpyAsync(d, host, n * sizeof(T), cudaMemcpyHostToDevice, st);
    }
    template <typename T>
    cudaError_t make(size_t n, T** dev) {
        void* d = nullptr;
        cudaError_t e = cudaMalloc(&d, (n ? n : 1) * sizeof(T));
        if (e != cudaSuccess) return e;
        p.push_back(d);
        *dev = (T*)d;
        return cudaSuccess;
    }
};
}  // namespace

extern "C" int ge_similarity_pairs(ge_ctx* ctx, const ge_sim_group* grp, int64_t* count, double* seconds) {
    if (!ctx || !grp) return GE_ERR_INVALID;
    const int32_t n = grp->n_items, method = grp->method;
    if (method < GE_SIM_NGRAM_COSINE || method > GE_SIM_LEVENSHTEIN) { ctx->err = "unsupported similarity method"; return GE_ERR_INVALID; }
    if (n < 0 || grp->n_source < 0 || grp->n_target < 0 || (n > 0 && (!grp->vert || !grp->string_id || !grp->flags)) ||
        (grp->n_source > 0 && !grp->source) || (grp->n_target > 0 && !grp->target)) { ctx->err = "bad similarity group arguments"; return GE_ERR_INVALID; }
    const bool profile = method <= GE_SIM_TOKEN_JACCARD, numeric = method == GE_SIM_NUMERIC;
    const bool date = method >= GE_SIM_DATE_DAYS && method <= GE_SIM_DATE_YEARS, text = method >= GE_SIM_JAROWINKLER;
    if (n > 0 && text && (!grp->char_off || (grp->char_off[n] > 0 && !grp->chars))) { ctx->err = "JAROWINKLER / LEVENSHTEIN without labels"; return GE_ERR_INVALID; }
    if (n > 0 && profile && (!grp->prof_off || (grp->prof_off[n] > 0 && (!grp->prof_key || !grp->prof_cnt)))) { ctx->err = "profile metric without profiles"; return GE_ERR_INVALID; }
    if (n > 0 && numeric && (grp->n_cut <= 0 || !grp->cut_slot || !grp->num_val || !grp->num_status)) { ctx->err = "NUMERIC without parsed values"; return GE_ERR_INVALID; }
    if (n > 0 && date && (!grp->epoch_day || !grp->packed_month)) { ctx->err = "DATE_* without parsed dates"; return GE_ERR_INVALID; }
    if (grp->time_dir < 0 || grp->time_dir > 2) { ctx->err = "bad time direction"; return GE_ERR_INVALID; }
    for (int32_t k = 0; k < grp->n_source; k++) if (grp->source[k] < 0 || grp->source[k] >= n) { ctx->err = "source item out of range"; return GE_ERR_INVALID; }
    for (int32_t k = 0; k < grp->n_target; k++) if (grp->target[k] < 0 || grp->target[k] >= n) { ctx->err = "target item out of range"; return GE_ERR_INVALID; }
    if (numeric) for (int32_t k = 0; k < n; k++) if (grp->cut_slot[k] < 0 || grp->cut_slot[k] >= grp->n_cut) { ctx->err = "cut slot out of range"; return GE_ERR_INVALID; }
    cudaSetDevice(ctx->cfg.device);
    ge_free_sim(ctx);
    ctx->sim = new SimState();
    if (count) *count = 0;
    if (seconds) *seconds = 0.0;
    if (grp->n_source == 0 || grp->n_target == 0) return GE_OK;

    cudaStream_t st = ctx->stream;
    Scratch sc;
    SimDev g{};
    g.method = method; g.upper = grp->upper_triangle ? 1 : 0; g.time_dir = grp->time_dir; g.n_cut = numeric ? grp->n_cut : 1;
    g.threshold = grp->threshold; g.smooth = grp->smooth; g.distance = grp->distance;
    g.n_source = grp->n_source; g.n_target = grp->n_target;
    g.nseg = (grp->n_target + SEG - 1) / SEG;
    GE_CUDA(sc.up(grp->vert, (size_t)n, &g.vert, st));
    GE_CUDA(sc.up(grp->string_id, (size_t)n, &g.sid, st));
    GE_CUDA(sc.up(grp->flags, (size_t)n, &g.flags, st));
    GE_CUDA(sc.up(grp->source, (size_t)grp->n_source, &g.source, st));
    GE_CUDA(sc.up(grp->target, (size_t)grp->n_target, &g.target, st));
    if (profile) {
        const size_t len = (size_t)grp->prof_off[n];
        GE_CUDA(sc.up(grp->prof_off, (size_t)n + 1, &g.prof_off, st));
        GE_CUDA(sc.up(grp->prof_key, len, &g.prof_key, st));
        GE_CUDA(sc.up(grp->prof_cnt, len, &g.prof_cnt, st));
        double* d_norm = nullptr;
        GE_CUDA(sc.make((size_t)n, &d_norm));
        profile_norms<<<(n + 255) / 256, 256, 0, st>>>(g.prof_off, g.prof_cnt, n, d_norm);
        g.norm = d_norm;
    } else if (numeric) {
        GE_CUDA(sc.up(grp->cut_slot, (size_t)n, &g.cut_slot, st));
        GE_CUDA(sc.up(grp->num_val, (size_t)n * grp->n_cut, &g.num_val, st));
        GE_CUDA(sc.up(grp->num_status, (size_t)n * grp->n_cut, &g.num_status, st));
    } else if (date) {
        GE_CUDA(sc.up(grp->epoch_day, (size_t)n, &g.epoch_day, st));
        GE_CUDA(sc.up(grp->packed_month, (size_t)n, &g.packed_month, st));
    }
    int64_t scratch_words = 0;  // per thread
    if (text) {
        GE_CUDA(sc.up(grp->char_off, (size_t)n + 1, &g.char_off, st));
        GE_CUDA(sc.up(grp->chars, (size_t)grp->char_off[n], &g.chars, st));
        int64_t longest = 0;
        for (int32_t k = 0; k < n; k++) {
            const int64_t len = grp->char_off[k + 1] - grp->char_off[k];
            if (len < 0) { ctx->err = "label offsets must not decrease"; return GE_ERR_INVALID; }
            longest = len > longest ? len : longest;
        }
        // LEVENSHTEIN: one DP row; JAROWINKLER: two bit sets
        scratch_words = method == GE_SIM_LEVENSHTEIN ? longest + 1 : 2 * ((longest + 31) / 32);
    }
    const int64_t nwork = (int64_t)g.n_source * g.nseg;
    const int64_t ntile = (nwork + SCAN_TILE - 1) / SCAN_TILE;
    uint32_t* d_cnt = nullptr;
    unsigned long long *d_off = nullptr, *d_tile = nullptr, *d_total = nullptr;
    uint8_t* d_dead = nullptr;
    GE_CUDA(sc.make((size_t)nwork, &d_cnt));
    GE_CUDA(sc.make((size_t)nwork, &d_off));
    GE_CUDA(sc.make((size_t)ntile, &d_tile));
    GE_CUDA(sc.make((size_t)1, &d_total));
    GE_CUDA(sc.make((size_t)g.n_source, &d_dead));
    GE_CUDA(cudaMemsetAsync(d_dead, 0, (size_t)g.n_source, st));

    if (!ctx->ev0) { GE_CUDA(cudaEventCreate(&ctx->ev0)); GE_CUDA(cudaEventCreate(&ctx->ev1)); }
    GE_CUDA(cudaEventRecord(ctx->ev0, st));
    const int64_t want = (nwork * 32 + 255) / 256;
    int grid = (int)(want < (int64_t)ctx->sm_count * 8 ? (want < 1 ? 1 : want) : (int64_t)ctx->sm_count * 8);
    if (scratch_words > 0) {  // bound the scratch to 2 GB by running fewer blocks when labels are very long
        const int64_t fit = ((int64_t)2 << 30) / (256 * 4 * scratch_words);
        if (grid > fit) grid = fit < 1 ? 1 : (int)fit;
        g.nthreads = (int64_t)grid * 256;
        GE_CUDA(sc.make((size_t)(g.nthreads * scratch_words), &g.scratch));
    }
    sim_pass<false><<<grid, 256, 0, st>>>(g, d_cnt, nullptr, d_dead, nullptr, nullptr, nullptr);
    drop_dead_rows<<<(unsigned)((nwork + 255) / 256), 256, 0, st>>>(d_cnt, d_dead, g.nseg, nwork);
    scan_tile_totals<<<(unsigned)ntile, 256, 0, st>>>(d_cnt, nwork, d_tile);
    scan_tile_sums<<<1, 256, 0, st>>>(d_tile, ntile, d_total);
    scan_tile_apply<<<(unsigned)ntile, 256, 0, st>>>(d_cnt, nwork, d_tile, d_off);
    GE_CUDA(cudaGetLastError());
    unsigned long long total = 0;
    GE_CUDA(cudaMemcpyAsync(&total, d_total, sizeof total, cudaMemcpyDeviceToHost, st));
    GE_CUDA(cudaStreamSynchronize(st));
    SimState* s = ctx->sim;
    s->count = (int64_t)total;
    if (total > 0) {
        GE_CUDA(cudaMalloc(&s->d_v, (size_t)total * sizeof(int32_t)));
        GE_CUDA(cudaMalloc(&s->d_o, (size_t)total * sizeof(int32_t)));
        GE_CUDA(cudaMalloc(&s->d_s, (size_t)total * sizeof(float)));
        sim_pass<true><<<grid, 256, 0, st>>>(g, nullptr, d_off, d_dead, s->d_v, s->d_o, s->d_s);
        GE_CUDA(cudaGetLastError());
    }
    GE_CUDA(cudaEventRecord(ctx->ev1, st));
    GE_CUDA(cudaStreamSynchronize(st));
    float ms = 0.f;
    GE_CUDA(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    if (count) *count = s->count;
    if (seconds) *seconds = (double)ms * 1e-3;
    GE_DBG("ge_similarity_pairs");
    return GE_OK;
}

extern "C" int ge_similarity_export(ge_ctx* ctx, int64_t cap, int32_t* vert, int32_t* other, float* sim) {
    if (!ctx) return GE_ERR_INVALID;
    if (!ctx->sim) { ctx->err = "no similarity result: call ge_similarity_pairs first"; return GE_ERR_STATE; }
    const SimState* s = ctx->sim;
    if (cap < s->count) { ctx->err = "similarity export: capacity below the pair count"; return GE_ERR_INVALID; }
    if (s->count == 0) return GE_OK;
    cudaSetDevice(ctx->cfg.device);
    if (vert) GE_CUDA(cudaMemcpyAsync(vert, s->d_v, (size_t)s->count * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    if (other) GE_CUDA(cudaMemcpyAsync(other, s->d_o, (size_t)s->count * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    if (sim) GE_CUDA(cudaMemcpyAsync(sim, s->d_s, (size_t)s->count * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    GE_CUDA(cudaStreamSynchronize(ctx->stream));
    return GE_OK;
}
