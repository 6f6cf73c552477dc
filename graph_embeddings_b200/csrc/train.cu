// train.cu — embedding store + GloVe / pGloVe AdaGrad update (K4, K5, K6, K8 of SURVEY §2.3).
//
// Replaces J/opt/Optimizer.java (store :45-57, extractResult :130-140) and the per-nonzero
// loop of J/opt/grad/Adagrad.java:49-95 with GloveCost.java:7-20 / PGloveCost.java:7-20.
//
// K6 is HBM-bound gather / FMA / scatter: one warp owns one nonzero at a time, the two
// store rows (W_i|Gw_i|b_i|Gb_i and C_j|Gc_j|b_j|Gb_j) are each ONE contiguous block of S
// floats read with 128-bit loads that bypass L1 (ld.global.cg: Hogwild readers must see other
// SMs' writes) and written back with 128-bit stores.  Algorithmic bytes per nonzero:
// 32*dim + 44.
#include "common.cuh"

#include <math.h>
#include <string.h>

#include <sys/mman.h>
#include <thread>

namespace {

__device__ __forceinline__ float4 ldcg4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }
// order-pinned variants (asm volatile statements keep their program order): used to fix WHEN the
// row loads issue relative to the dot-product reduction (GE_LOAD_ORDER, see the kernel)
__device__ __forceinline__ float4 ldcg4v(const float* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float2 ldcg2v(const float* p) {
    float2 v;
    asm volatile("ld.global.cg.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// L2 residency control.  The store is far larger than the 126 MB L2 and rows are visited in a uniformly
// shuffled order, so under plain LRU a row survives between two of its visits only if it is touched
// every ~50 us (C4: degree > ~19 000): the read hit rate is 13 % (profiles/r02_update_c4_ncu.md).
// Rows whose occurrence count puts them among the ~60 000 most visited (bit 31 of the packed id, set
// at pack time) are loaded and reduced with an evict_last cache policy; everything else uses the
// plain instructions (measured on B200: routing EVERY access through a createpolicy descriptor, even
// evict_normal, costs 17 % at C4 — profiles/r02_l2_keep.md).
__device__ __forceinline__ uint64_t l2_policy_keep() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ uint64_t l2_policy_stream() {
    uint64_t p;
    asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
    return p;
}
__device__ __forceinline__ float4 ldcg4p(const float* p, uint64_t pol) {
    float4 v;
    asm volatile("ld.global.cg.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p), "l"(pol) : "memory");
    return v;
}
__device__ __forceinline__ float2 ldcg2p(const float* p, uint64_t pol) {
    float2 v;
    asm volatile("ld.global.cg.L2::cache_hint.v2.f32 {%0, %1}, [%2], %3;" : "=f"(v.x), "=f"(v.y) : "l"(p), "l"(pol) : "memory");
    return v;
}
__device__ __forceinline__ void red4p(float* p, float a, float b, float c, float d, uint64_t pol) {
    asm volatile("red.relaxed.gpu.global.add.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d), "l"(pol) : "memory");
}
__device__ __forceinline__ void red2p(float* p, float a, float b, uint64_t pol) {
    asm volatile("red.relaxed.gpu.global.add.L2::cache_hint.v2.f32 [%0], {%1, %2}, %3;" ::"l"(p), "f"(a), "f"(b), "l"(pol) : "memory");
}

// vector float reduction at L2 (sm_90+): no value returns to the SM, updates are never lost
__device__ __forceinline__ void red4(float* p, float a, float b, float c, float d) {
    asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red2(float* p, float a, float b) {
    asm volatile("red.relaxed.gpu.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}

// Bulk L2 prefetch of one whole store row (vec | gradsq | bias block: S*4 contiguous bytes, 16-B
// aligned) issued by ONE lane through the bulk-copy (TMA) engine: the row of the nonzero `pf`
// steps ahead is in L2 when its LDG.128s issue, which raises memory-level parallelism without
// spending registers on a second row set.
__device__ __forceinline__ void prefetch_row_l2(const float* row, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(row), "r"(bytes) : "memory");
}

__device__ __forceinline__ void prefetch_line_l2(const float* p) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(p) : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// GloveCost.innerCost :9-11 — `inner += focus[u+d] * context[v+d]` is a SEQUENTIAL f32 sum over d.
// JAVA math reproduces that order exactly: lanes publish their four products, every lane folds
// them in dimension order (redundantly), so the parity trajectory does not depend on a tree sum.
template <int NT>
__device__ __forceinline__ float dot_sequential(const float4 (&w)[NT], const float4 (&c)[NT], int nchunk) {
    float acc = 0.f;
#pragma unroll
    for (int q = 0; q < NT; q++) {
        const float px = __fmul_rn(w[q].x, c[q].x), py = __fmul_rn(w[q].y, c[q].y);
        const float pz = __fmul_rn(w[q].z, c[q].z), pw = __fmul_rn(w[q].w, c[q].w);
        const int live = min(32, nchunk - 32 * q);
        for (int s = 0; s < live; s++) {
            acc = __fadd_rn(acc, __shfl_sync(0xffffffffu, px, s));
            acc = __fadd_rn(acc, __shfl_sync(0xffffffffu, py, s));
            acc = __fadd_rn(acc, __shfl_sync(0xffffffffu, pz, s));
            acc = __fadd_rn(acc, __shfl_sync(0xffffffffu, pw, s));
        }
    }
    return acc;
}

// Keyed bijection on [0, n): 4-round unbalanced Feistel over ceil(log2 n) bits + cycle walking.
// Stands in for the reference's Fisher-Yates epoch shuffle (Permutation.java:21-23) in
// GE_ORDER_DEVICE mode: a fresh pseudo-random order per epoch with no permutation array.
struct Bijection {
    uint64_t n;
    uint32_t abits, bbits;
    uint64_t key;
    __host__ __device__ static Bijection make(uint64_t n, uint64_t key) {
        uint32_t bits = 1;
        while ((1ULL << bits) < n) bits++;
        if (bits < 2) bits = 2;
        Bijection b;
        b.n = n; b.abits = bits / 2; b.bbits = bits - bits / 2; b.key = key;
        return b;
    }
    __host__ __device__ inline uint32_t f(uint32_t x, uint32_t round) const {
        uint64_t h = (uint64_t)x * 0x9E3779B97F4A7C15ULL + key + round * 0xD1B54A32D192ED03ULL;
        h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ULL; h ^= h >> 32;
        return (uint32_t)h;
    }
    __host__ __device__ inline uint64_t operator()(uint64_t k) const {
        const uint32_t ma = (1u << abits) - 1, mb = (uint32_t)((1ULL << bbits) - 1);
        do {
            uint32_t L = (uint32_t)(k >> bbits) & ma, R = (uint32_t)k & mb;
            L ^= f(R, 0) & ma;
            R ^= f(L, 1) & mb;
            L ^= f(R, 2) & ma;
            R ^= f(L, 3) & mb;
            k = ((uint64_t)L << bbits) | R;
        } while (k >= n);
        return k;
    }
};

// ------------------------------------------------------------------------------------------
// K6 — the update kernel.  NT = float4 chunks per lane per vector (dim <= 128*NT).
// JAVA: reference promotion pattern (fp64 islands); otherwise f32 + rsqrt.
// ------------------------------------------------------------------------------------------
template <int NT, bool JAVA, bool ATOMIC, bool TP, bool HOT, bool DYN>
__global__ void __launch_bounds__(256, (NT == 1 && !JAVA) ? 4 : 1)
glove_adagrad_update(const NzEntry* __restrict__ nz, int64_t count, float* __restrict__ W,
                     float* __restrict__ C, const int32_t* __restrict__ perm, Bijection bij,
                     int32_t D4, int32_t S, int32_t method, double xmax, float lr, int32_t pf,
                     unsigned long long* __restrict__ queue, double* __restrict__ cost_out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int nchunk = D4 >> 2;  // float4 chunks per vector
    const double lr_d = (double)lr;
    double cost_d = 0.0;
    float cost_f = 0.0f;  // JAVA: the reference's f32 running cost (Adagrad.java:46,60)
    const uint64_t pol_keep = l2_policy_keep();

    auto fetch = [&](int64_t pos) -> NzEntry {  // lane l fetches the nonzero at shuffled position pos
        NzEntry e = {0, 0, 0.f, 0.f};
        if (pos < count) {
            const int64_t idx = perm ? (int64_t)perm[pos] : (int64_t)bij((uint64_t)pos);
            const int4 raw = __ldg(reinterpret_cast<const int4*>(nz + idx));
            e.i = raw.x; e.j = raw.y; e.a = __int_as_float(raw.z); e.b = __int_as_float(raw.w);
        }
        return e;
    };
    const uint32_t row_bytes = (uint32_t)S * 4u;
    constexpr bool two_phase = TP;  // launcher: D4 >= 128 (rows at least one warp of float4 wide)
    const bool pf_lines = pf >= 100;  // GE_PF_DIST = 100 + distance selects prefetch.global.L2 per line
    if (pf_lines) pf -= 100;
    // Batches of 32 nonzeros.  !DYN: static grid-stride ownership (single-GPU epochs: one resident
    // wave, every warp starts at once).  DYN: warps claim chunks of QCH batches from a global counter,
    // so blocks that become resident late (the pipelined multi-GPU trainer runs NCCL copy kernels
    // beside this one) simply claim less instead of stretching the tail.
    constexpr int64_t QCH = 8;
    auto claim = [&]() -> int64_t {  // first position of a freshly claimed chunk
        unsigned long long c = 0;
        if (lane == 0) c = atomicAdd(queue, (unsigned long long)QCH);
        return (int64_t)__shfl_sync(0xffffffffu, c, 0) * 32;
    };
    int64_t chunk_end = 0, next_chunk = 0;
    int64_t base0 = warp * 32;
    if (DYN) { base0 = claim(); chunk_end = base0 + QCH * 32; next_chunk = claim(); }
    NzEntry mine = fetch(base0 + lane);
    for (int64_t base = base0, next_base; base < count; base = next_base) {
        if (DYN) {
            next_base = base + 32;
            if (next_base == chunk_end) { next_base = next_chunk; chunk_end = next_base + QCH * 32; next_chunk = claim(); }
        } else {
            next_base = base + nwarps * 32;
        }
        const NzEntry nxt = fetch(next_base + lane);  // next batch: issued now, needed 32 nonzeros later
        const int nvalid = (int)min((int64_t)32, count - base);
        for (int t = 0; t < nvalid; t++) {
            if (pf > 0) {  // warp-uniform
                const int tt = t + pf;
                int32_t pi, pj;
                bool ok;
                if (tt < 32) {
                    pi = __shfl_sync(0xffffffffu, mine.i, tt); pj = __shfl_sync(0xffffffffu, mine.j, tt);
                    ok = base + tt < count;
                } else {
                    pi = __shfl_sync(0xffffffffu, nxt.i, tt - 32); pj = __shfl_sync(0xffffffffu, nxt.j, tt - 32);
                    ok = next_base + (tt - 32) < count;
                }
                pi &= GE_ID_MASK; pj &= GE_ID_MASK;
                if (ok) {
                    if (pf_lines) {  // per-line variant: lanes 0..15 cover the W row, 16..31 the C row (128 B each)
                        const int l = lane & 15;
                        if ((uint32_t)l * 128u < row_bytes)
                            prefetch_line_l2((lane < 16 ? W + (size_t)pi * S : C + (size_t)pj * S) + l * 32);
                    } else if (lane < 2) {
                        prefetch_row_l2(lane == 0 ? W + (size_t)pi * S : C + (size_t)pj * S, row_bytes);
                    }
                }
            }
            int32_t bi = __shfl_sync(0xffffffffu, mine.i, t);
            int32_t bj = __shfl_sync(0xffffffffu, mine.j, t);
            // bits 27..30 = concurrency class k of a hub id (pack time): its deltas are scaled by 2^(-k/2)
            // so that the K stale steps the L2 adds up act like at most GE_HOT_WRITERS of them
            float si = 1.f, sj = 1.f;
            if (HOT) {  // HOT kernels exist for ATOMIC only
                si = exp2f(-0.5f * (float)((bi >> GE_ID_BITS) & 15));
                sj = exp2f(-0.5f * (float)((bj >> GE_ID_BITS) & 15));
            }
            // bit 31 = row is in the L2-resident set (pack time): warp-uniform choice of the access flavour
            const bool keepw = bi < 0, keepc = bj < 0;
            bi &= GE_ID_MASK; bj &= GE_ID_MASK;
            const float ea = __shfl_sync(0xffffffffu, mine.a, t);
            const float eb = __shfl_sync(0xffffffffu, mine.b, t);
            float* wrow = W + (size_t)bi * S;
            float* crow = C + (size_t)bj * S;

            float4 w[NT], c[NT], gw[NT], gc[NT];
            float2 bw, bc;
            float dot = 0.f;
            if (JAVA) {
#pragma unroll
                for (int q = 0; q < NT; q++) {
                    const int ch = lane + 32 * q;
                    if (ch < nchunk) {
                        w[q] = ldcg4(wrow + 4 * ch);
                        c[q] = ldcg4(crow + 4 * ch);
                        gw[q] = ldcg4(wrow + D4 + 4 * ch);
                        gc[q] = ldcg4(crow + D4 + 4 * ch);
                    } else {
                        w[q] = c[q] = make_float4(0.f, 0.f, 0.f, 0.f);
                        gw[q] = gc[q] = make_float4(1.f, 1.f, 1.f, 1.f);
                    }
                }
                bw = __ldcg(reinterpret_cast<const float2*>(wrow + 2 * D4));  // (bias, gradsq)
                bc = __ldcg(reinterpret_cast<const float2*>(crow + 2 * D4));
                dot = dot_sequential<NT>(w, c, nchunk);
            } else {
                // Issue order of the row loads is pinned (asm volatile).  two_phase: the vectors
                // first, the AdaGrad rows + bias pairs only after the dot-product reduction —
                // measured on B200 (C3 graph): d=128 134.4 vs 141.8 ms, d=300 325 vs 370 ms per
                // epoch; for rows narrower than a warp (d=100: 119.8 vs 107.7 ms) everything up
                // front is faster, so the launcher picks by D4 (profiles/r01_load_order.md).
                auto ldv = [&](float4 (&a)[NT], const float* base, float fill, bool keep) {
#pragma unroll
                    for (int q = 0; q < NT; q++) {
                        const int ch = lane + 32 * q;
                        a[q] = ch < nchunk ? (keep ? ldcg4p(base + 4 * ch, pol_keep) : ldcg4v(base + 4 * ch)) : make_float4(fill, fill, fill, fill);
                    }
                };
                auto ldb = [&](const float* p, bool keep) -> float2 { return keep ? ldcg2p(p, pol_keep) : ldcg2v(p); };
                ldv(w, wrow, 0.f, keepw);
                ldv(c, crow, 0.f, keepc);
                if (!two_phase) {
                    ldv(gw, wrow + D4, 1.f, keepw); ldv(gc, crow + D4, 1.f, keepc);
                    bw = ldb(wrow + 2 * D4, keepw); bc = ldb(crow + 2 * D4, keepc);
                }
#pragma unroll
                for (int q = 0; q < NT; q++)
                    dot += w[q].x * c[q].x + w[q].y * c[q].y + w[q].z * c[q].z + w[q].w * c[q].w;
                dot = warp_sum(dot);
                if (two_phase) {
                    ldv(gw, wrow + D4, 1.f, keepw); ldv(gc, crow + D4, 1.f, keepc);
                    bw = ldb(wrow + 2 * D4, keepw); bc = ldb(crow + 2 * D4, keepc);
                }
            }

            float inner, wc;
            if (JAVA) {
                // GloveCost.java:13 / PGloveCost.java:13, weightedCost :19
                const float x = ea;
                double target;
                if (method == GE_METHOD_GLOVE) target = log((double)x);
                else target = log((double)__fdiv_rn(x, __fsub_rn(1.0f, x)));
                inner = (float)__dadd_rn((double)dot, __dsub_rn((double)__fadd_rn(bw.x, bc.x), target));
                if (method == GE_METHOD_GLOVE)
                    wc = ((double)x > xmax) ? inner : __fmul_rn((float)pow(__ddiv_rn((double)x, xmax), 0.75), inner);
                else
                    wc = __fmul_rn(x, inner);
                // Adagrad.java:60 — cost = (float)(cost + 0.5*wc*inner), product in fp64
                cost_f = (float)__dadd_rn((double)cost_f, __dmul_rn(__dmul_rn(0.5, (double)wc), (double)inner));
            } else {
                inner = dot + (bw.x + bc.x) - ea;
                wc = eb * inner;
                cost_d += 0.5 * (double)wc * (double)inner;
            }

#pragma unroll
            for (int q = 0; q < NT; q++) {
                const int ch = lane + 32 * q;
                if (ch < nchunk) {
                    float* wv = reinterpret_cast<float*>(&w[q]);
                    float* cv = reinterpret_cast<float*>(&c[q]);
                    float* gwv = reinterpret_cast<float*>(&gw[q]);
                    float* gcv = reinterpret_cast<float*>(&gc[q]);
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        if (JAVA) {  // Adagrad.java:72-80
                            const float g1 = __fmul_rn(wc, cv[e]);
                            const float g2 = __fmul_rn(wc, wv[e]);
                            wv[e] = (float)__dsub_rn((double)wv[e], __dmul_rn(__ddiv_rn((double)g1, __dsqrt_rn((double)gwv[e])), lr_d));
                            cv[e] = (float)__dsub_rn((double)cv[e], __dmul_rn(__ddiv_rn((double)g2, __dsqrt_rn((double)gcv[e])), lr_d));
                            gwv[e] = __fadd_rn(gwv[e], __fmul_rn(g1, g1));
                            gcv[e] = __fadd_rn(gcv[e], __fmul_rn(g2, g2));
                        } else if (ATOMIC) {  // deltas, applied by the L2 with red.add
                            const float g1 = wc * cv[e];
                            const float g2 = wc * wv[e];
                            float dw = -g1 * rsqrtf(gwv[e]) * lr, dc = -g2 * rsqrtf(gcv[e]) * lr;
                            // gradsq keeps the full g^2, as every reference update adds it.  (A state-dependent
                            // factor 2 / (1 + sqrt(1 + K g^2 / G)) — the exact ratio for K equal stale AdaGrad
                            // steps — was measured instead of the constant: better late epochs, but it ignores the
                            // curvature of the bias step and diverged inside column blocks; profiles/r02_hubs.md.)
                            if (HOT) { dw *= si; dc *= sj; }
                            wv[e] = dw;
                            cv[e] = dc;
                            gwv[e] = g1 * g1;
                            gcv[e] = g2 * g2;
                        } else {
                            const float g1 = wc * cv[e];
                            const float g2 = wc * wv[e];
                            wv[e] -= g1 * rsqrtf(gwv[e]) * lr;
                            cv[e] -= g2 * rsqrtf(gcv[e]) * lr;
                            gwv[e] += g1 * g1;
                            gcv[e] += g2 * g2;
                        }
                    }
                    if (ATOMIC) {
                        if (keepw) {
                            red4p(wrow + 4 * ch, w[q].x, w[q].y, w[q].z, w[q].w, pol_keep);
                            red4p(wrow + D4 + 4 * ch, gw[q].x, gw[q].y, gw[q].z, gw[q].w, pol_keep);
                        } else {
                            red4(wrow + 4 * ch, w[q].x, w[q].y, w[q].z, w[q].w);
                            red4(wrow + D4 + 4 * ch, gw[q].x, gw[q].y, gw[q].z, gw[q].w);
                        }
                        if (keepc) {
                            red4p(crow + 4 * ch, c[q].x, c[q].y, c[q].z, c[q].w, pol_keep);
                            red4p(crow + D4 + 4 * ch, gc[q].x, gc[q].y, gc[q].z, gc[q].w, pol_keep);
                        } else {
                            red4(crow + 4 * ch, c[q].x, c[q].y, c[q].z, c[q].w);
                            red4(crow + D4 + 4 * ch, gc[q].x, gc[q].y, gc[q].z, gc[q].w);
                        }
                    } else {
                        st4(wrow + 4 * ch, w[q]);
                        st4(crow + 4 * ch, c[q]);
                        st4(wrow + D4 + 4 * ch, gw[q]);
                        st4(crow + D4 + 4 * ch, gc[q]);
                    }
                }
            }
            if (lane == 0) {  // Adagrad.java:88-93 — bias step has no learning rate
                if (JAVA) {
                    bw.x = (float)__dsub_rn((double)bw.x, __ddiv_rn((double)wc, __dsqrt_rn((double)bw.y)));
                    bc.x = (float)__dsub_rn((double)bc.x, __ddiv_rn((double)wc, __dsqrt_rn((double)bc.y)));
                    const float w2 = __fmul_rn(wc, wc);
                    bw.y = __fadd_rn(bw.y, w2);
                    bc.y = __fadd_rn(bc.y, w2);
                } else if (ATOMIC) {
                    const float ab1 = wc * rsqrtf(bw.y) * si, ab2 = wc * rsqrtf(bc.y) * sj;
                    if (keepw) red2p(wrow + 2 * D4, -ab1, wc * wc, pol_keep); else red2(wrow + 2 * D4, -ab1, wc * wc);
                    if (keepc) red2p(crow + 2 * D4, -ab2, wc * wc, pol_keep); else red2(crow + 2 * D4, -ab2, wc * wc);
                } else {
                    bw.x -= wc * rsqrtf(bw.y);
                    bc.x -= wc * rsqrtf(bc.y);
                    bw.y += wc * wc;
                    bc.y += wc * wc;
                }
                if (!ATOMIC || JAVA) {
                    *reinterpret_cast<float2*>(wrow + 2 * D4) = bw;
                    *reinterpret_cast<float2*>(crow + 2 * D4) = bc;
                }
            }
            if (JAVA) __syncwarp();  // serial order: lane 0's bias store precedes the next loads
        }
        mine = nxt;
    }
    if (lane == 0) {
        const double cst = JAVA ? (double)cost_f : cost_d;
        if (cst != 0.0) atomicAdd(cost_out, cst);
    }
}

// ------------------------------------------------------------------------------------------
// Adam / AMSGrad update: J/opt/grad/Adam.java:75-146, AMSGrad.java:91-162.
// Same gather skeleton as K6; row = [vec | M1 | M2 | bias, M1b, M2b].  Algorithmic bytes per
// nonzero: 48*dim + 60 (6 rows each way + 6 bias scalars each way + the triple).
// AMS: second moment = max(old, new) (AMSGrad.java:129-130), no bias correction, step
// lr / (sqrt(v)+eps) * m (:133); Adam: step corr * m / (sqrt(v)+eps) (:118) with the per-epoch
// `corr` of Adam.java:84 computed on the host in fp64.
// FAST: f32 arithmetic, plain 128-bit stores — the reference's own racy read-modify-write
// (last writer wins).  Unlike AdaGrad's, an Adam step is ~lr*m/sqrt(v): bounded and nearly
// independent of the gradient's size, so ADDING the deltas of many concurrent warps on one hub
// row (red.add, GE_UPDATE_ATOMIC_ADAM, kept for experiments) overshoots by the number of
// concurrent writers; losing all but one of them does not.  JAVA: the reference's promotion
// pattern (f32 moments, fp64 step).
// ------------------------------------------------------------------------------------------
template <int NT, bool JAVA, bool AMS>
__global__ void __launch_bounds__(256)
adam_update(const NzEntry* __restrict__ nz, int64_t count, float* __restrict__ W, float* __restrict__ C,
            const int32_t* __restrict__ perm, Bijection bij, int32_t D4, int32_t S, int32_t method,
            double xmax, float lr, double corr_d, int32_t atomic_p, double* __restrict__ cost_out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int nchunk = D4 >> 2;
    const int32_t BO = 3 * D4;
    const float b1 = 0.9f, b2 = 0.999f, epsf = 1e-7f;               // Adam.java:45-53
    const float omb1 = __fsub_rn(1.0f, b1), omb2 = __fsub_rn(1.0f, b2);
    const double lr_d = (double)lr, eps_d = (double)epsf;
    const float stepf = AMS ? lr : (float)corr_d;
    double cost_d = 0.0;
    float cost_f = 0.0f;

    for (int64_t base = warp * 32; base < count; base += nwarps * 32) {
        NzEntry mine = {0, 0, 0.f, 0.f};
        const int64_t pos = base + lane;
        if (pos < count) {
            const int64_t idx = perm ? (int64_t)perm[pos] : (int64_t)bij((uint64_t)pos);
            const int4 raw = __ldg(reinterpret_cast<const int4*>(nz + idx));
            mine.i = raw.x; mine.j = raw.y; mine.a = __int_as_float(raw.z); mine.b = __int_as_float(raw.w);
        }
        const int nvalid = (int)min((int64_t)32, count - base);
        for (int t = 0; t < nvalid; t++) {
            const int32_t bi = __shfl_sync(0xffffffffu, mine.i, t) & GE_ID_MASK;  // damping bits unused here
            const int32_t bj = __shfl_sync(0xffffffffu, mine.j, t) & GE_ID_MASK;
            const float ea = __shfl_sync(0xffffffffu, mine.a, t);
            const float eb = __shfl_sync(0xffffffffu, mine.b, t);
            float* wrow = W + (size_t)bi * S;
            float* crow = C + (size_t)bj * S;

            float4 w[NT], c[NT], mw[NT], mc[NT], vw[NT], vc[NT];
#pragma unroll
            for (int q = 0; q < NT; q++) {
                const int ch = lane + 32 * q;
                if (ch < nchunk) {
                    w[q] = ldcg4(wrow + 4 * ch);
                    c[q] = ldcg4(crow + 4 * ch);
                    mw[q] = ldcg4(wrow + D4 + 4 * ch);
                    mc[q] = ldcg4(crow + D4 + 4 * ch);
                    vw[q] = ldcg4(wrow + 2 * D4 + 4 * ch);
                    vc[q] = ldcg4(crow + 2 * D4 + 4 * ch);
                } else {
                    w[q] = c[q] = mw[q] = mc[q] = vw[q] = vc[q] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            // bias block: (bias, M1b, M2b, pad) — 16-B aligned because BO = 3*D4 and S % 4 == 0
            float4 bw = ldcg4(wrow + BO), bc = ldcg4(crow + BO);

            float dot = 0.f;
            if (JAVA) {
                dot = dot_sequential<NT>(w, c, nchunk);
            } else {
#pragma unroll
                for (int q = 0; q < NT; q++)
                    dot += w[q].x * c[q].x + w[q].y * c[q].y + w[q].z * c[q].z + w[q].w * c[q].w;
                dot = warp_sum(dot);
            }

            float inner, wc;
            if (JAVA) {
                const float x = ea;
                double target;
                if (method == GE_METHOD_GLOVE) target = log((double)x);
                else target = log((double)__fdiv_rn(x, __fsub_rn(1.0f, x)));
                inner = (float)__dadd_rn((double)dot, __dsub_rn((double)__fadd_rn(bw.x, bc.x), target));
                if (method == GE_METHOD_GLOVE)
                    wc = ((double)x > xmax) ? inner : __fmul_rn((float)pow(__ddiv_rn((double)x, xmax), 0.75), inner);
                else
                    wc = __fmul_rn(x, inner);
                cost_f = (float)__dadd_rn((double)cost_f, __dmul_rn(__dmul_rn(0.5, (double)wc), (double)inner));
            } else {
                inner = dot + (bw.x + bc.x) - ea;
                wc = eb * inner;
                cost_d += 0.5 * (double)wc * (double)inner;
            }

            // one scalar step; returns the new parameter (JAVA) or its delta (FAST)
            auto upd = [&](float p, float g, float& m, float& v) -> float {
                if (JAVA) {
                    const float m1 = __fadd_rn(__fmul_rn(b1, m), __fmul_rn(omb1, g));
                    float v1 = __fadd_rn(__fmul_rn(b2, v), __fmul_rn(omb2, __fmul_rn(g, g)));
                    if (AMS) v1 = (v <= v1) ? v1 : v;  // FastMath.max(old, new)
                    m = m1; v = v1;
                    const double den = __dadd_rn(__dsqrt_rn((double)v1), eps_d);
                    if (AMS) return (float)__dsub_rn((double)p, __dmul_rn(__ddiv_rn(lr_d, den), (double)m1));
                    return (float)__dsub_rn((double)p, __ddiv_rn(__dmul_rn(corr_d, (double)m1), den));
                } else {
                    const float m1 = b1 * m + omb1 * g;
                    float v1 = b2 * v + omb2 * (g * g);
                    if (AMS) v1 = fmaxf(v, v1);
                    m = m1; v = v1;
                    const float delta = -stepf * m1 / (sqrtf(v1) + epsf);
                    return atomic_p ? delta : p + delta;
                }
            };

#pragma unroll
            for (int q = 0; q < NT; q++) {
                const int ch = lane + 32 * q;
                if (ch < nchunk) {
                    float* wv = reinterpret_cast<float*>(&w[q]);
                    float* cv = reinterpret_cast<float*>(&c[q]);
                    float* mwv = reinterpret_cast<float*>(&mw[q]);
                    float* mcv = reinterpret_cast<float*>(&mc[q]);
                    float* vwv = reinterpret_cast<float*>(&vw[q]);
                    float* vcv = reinterpret_cast<float*>(&vc[q]);
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const float gu = JAVA ? __fmul_rn(wc, cv[e]) : wc * cv[e];   // pre-update values
                        const float gv = JAVA ? __fmul_rn(wc, wv[e]) : wc * wv[e];
                        wv[e] = upd(wv[e], gu, mwv[e], vwv[e]);
                        cv[e] = upd(cv[e], gv, mcv[e], vcv[e]);
                    }
                    if (JAVA || !atomic_p) {
                        st4(wrow + 4 * ch, w[q]);
                        st4(crow + 4 * ch, c[q]);
                    } else {
                        red4(wrow + 4 * ch, w[q].x, w[q].y, w[q].z, w[q].w);
                        red4(crow + 4 * ch, c[q].x, c[q].y, c[q].z, c[q].w);
                    }
                    st4(wrow + D4 + 4 * ch, mw[q]);
                    st4(crow + D4 + 4 * ch, mc[q]);
                    st4(wrow + 2 * D4 + 4 * ch, vw[q]);
                    st4(crow + 2 * D4 + 4 * ch, vc[q]);
                }
            }
            if (lane == 0) {  // Adam.java:128-143 / AMSGrad.java:144-159 — biases: gradient = weightedCost
                const float nbw = upd(bw.x, wc, bw.y, bw.z);
                const float nbc = upd(bc.x, wc, bc.y, bc.z);
                if (JAVA || !atomic_p) {
                    bw.x = nbw; bc.x = nbc;
                    st4(wrow + BO, bw);
                    st4(crow + BO, bc);
                } else {
                    atomicAdd(wrow + BO, nbw);  // result unused -> RED.ADD.F32: delta on the bias
                    atomicAdd(crow + BO, nbc);
                    wrow[BO + 1] = bw.y; wrow[BO + 2] = bw.z;
                    crow[BO + 1] = bc.y; crow[BO + 2] = bc.z;
                }
            }
            if (JAVA) __syncwarp();
        }
    }
    if (lane == 0) {
        const double cst = JAVA ? (double)cost_f : cost_d;
        if (cst != 0.0) atomicAdd(cost_out, cst);
    }
}

// ------------------------------------------------------------------------------------------
// pack: (I, J, X) -> NzEntry.  FAST math folds log / pow (per-nonzero constants) into the
// entry once, in fp64, so the hot loop has no transcendental (GloveCost.java:13,19).
// ------------------------------------------------------------------------------------------
__global__ void pack_nonzeros(const int32_t* __restrict__ I, const int32_t* __restrict__ J,
                              const float* __restrict__ X, int64_t nnz, int32_t method, int32_t java,
                              double xmax, const int32_t* __restrict__ cnt_i, const int32_t* __restrict__ cnt_j,
                              int32_t hot_count, int32_t keep_i, int32_t keep_j, int32_t* __restrict__ any_hot,
                              NzEntry* __restrict__ out) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < nnz; k += stride) {
        const float x = X[k];
        NzEntry e;
        e.i = I[k]; e.j = J[k];
        if (cnt_i) {  // hub row: damping exponent; L2-resident set: bit 31
            const int32_t c = cnt_i[e.i], k = ge_damp_bits(c, hot_count);
            if (k) { e.i |= k << GE_ID_BITS; *any_hot = 1; }
            if (c >= keep_i) e.i |= GE_ID_KEEP;
        }
        if (cnt_j) {  // hub column
            const int32_t c = cnt_j[e.j], k = ge_damp_bits(c, hot_count);
            if (k) { e.j |= k << GE_ID_BITS; *any_hot = 1; }
            if (c >= keep_j) e.j |= GE_ID_KEEP;
        }
        if (java) { e.a = x; e.b = 0.f; }
        else if (method == GE_METHOD_GLOVE) {
            e.a = (float)log((double)x);
            e.b = ((double)x > xmax) ? 1.0f : (float)pow((double)x / xmax, 0.75);
        } else {
            e.a = (float)log((double)(x / (1.0f - x)));
            e.b = x;
        }
        out[k] = e;
    }
}

// sum of the occurrence counts of the ids at or above `threshold` (nonzeros that touch a damped id)
__global__ void hot_mass(const int32_t* __restrict__ cnt, int64_t n, int32_t threshold, unsigned long long* __restrict__ out) {
    unsigned long long m = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) { const int32_t c = cnt[k]; if (c >= threshold) m += (unsigned long long)c; }
    for (int o = 16; o > 0; o >>= 1) m += __shfl_xor_sync(0xffffffffu, m, o);
    if ((threadIdx.x & 31) == 0 && m) atomicAdd(out, m);
}

// largest occurrence count of any id (input of the concurrency cap, ge_warp_cap)
__global__ void max_count(const int32_t* __restrict__ cnt, int64_t n, int32_t* __restrict__ out) {
    int32_t m = 0;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) m = max(m, cnt[k]);
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m > 0) atomicMax(out, m);
}

// ------------------------------------------------------------------------------------------
// K4 — store init.  One thread per row replays the row's (2 + 2*dim) draws; GE_INIT_JAVA jumps
// the LCG to the row's stream position, so the result equals Optimizer.java:50-57 bit for bit.
// value = (float)((double)nextFloat - 0.5) / (float)dim ; gradsq = 1 (Adagrad.java:27-33)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t counter24(uint64_t seed, uint64_t idx) {
    uint64_t h = idx * 0x9E3779B97F4A7C15ULL + seed;
    h ^= h >> 30; h *= 0xBF58476D1CE4E5B9ULL; h ^= h >> 27; h *= 0x94D049BB133111EBULL; h ^= h >> 31;
    return (uint32_t)(h >> 40);
}

__global__ void store_init(float* __restrict__ W, float* __restrict__ C, int32_t V, int32_t dim,
                           int32_t D4, int32_t S, int64_t seed, int32_t java, int32_t row0,
                           int32_t NS, float g0, int32_t lay_nb, int32_t lay_Bs, int32_t Vglobal) {
    const int32_t r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= V) return;
    const int32_t id = ge_vid_inv(row0 + r, lay_nb, lay_Bs);  // slab row r holds this vertex
    if (id >= Vglobal) return;                                 // padding row of a cyclic block: stays zero, never addressed
    const uint64_t per_row = 2 + 2 * (uint64_t)dim;
    const uint64_t first = (uint64_t)id * per_row;  // global draw index of this row
    JavaLcg g = JavaLcg::seeded(seed);
    if (java) g.jump(first);
    uint64_t n = first;
    const float fd = (float)dim;
    auto draw = [&]() -> float {
        const float u = java ? g.nextFloat() : (float)counter24((uint64_t)seed, n) / (float)(1 << 24);
        n++;
        return __fdiv_rn((float)((double)u - 0.5), fd);
    };
    float* w = W + (size_t)r * S;
    float* c = C + (size_t)r * S;
    const int32_t BO = (1 + NS) * D4;  // bias block: [bias | state scalars(NS)]
    w[BO] = draw();
    c[BO] = draw();
    for (int s = 1; s <= NS; s++) { w[BO + s] = g0; c[BO + s] = g0; }
    for (int d = 0; d < dim; d++) {
        w[d] = draw();
        c[d] = draw();
    }
    for (int d = dim; d < D4; d++) w[d] = c[d] = 0.f;
    // state vectors: AdaGrad gradsq = 1 (Adagrad.java:27-33); Adam/AMSGrad moments = 0 (`new float[]`)
    for (int d = D4; d < BO; d++) { w[d] = g0; c[d] = g0; }
    for (int d = BO + 1 + NS; d < S; d++) { w[d] = 0.f; c[d] = 0.f; }
}

// K8 — extractResult (Optimizer.java:130-140): (focus + context) / 2 in f32, widened
template <typename T>
__global__ void extract_avg(const float* __restrict__ W, const float* __restrict__ C, int64_t row0, int32_t rows,
                            int32_t dim, int32_t S, T* __restrict__ out) {
    const int64_t total = (int64_t)rows * dim, stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += stride) {
        const int64_t r = k / dim;
        const int32_t d = (int32_t)(k - r * dim);
        out[k] = (T)__fdiv_rn(__fadd_rn(W[(row0 + r) * S + d], C[(row0 + r) * S + d]), 2.0f);
    }
}

// de-interleave one component of the store into the reference's array layout (validation)
__global__ void store_component(const float* __restrict__ A, int32_t V, int32_t dim, int32_t S,
                                int32_t off, int32_t width, float* __restrict__ out) {
    const int64_t total = (int64_t)V * width, stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < total; k += stride) {
        const int64_t r = k / width;
        out[k] = A[r * S + off + (k - r * width)];
    }
}

__global__ void materialise_order(Bijection bij, int64_t n, int32_t* __restrict__ out, int32_t base) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride)
        out[k] = base + (int32_t)bij((uint64_t)k);
}

uint64_t epoch_key(int64_t seed, int32_t epoch) {
    uint64_t h = (uint64_t)seed * 0x9E3779B97F4A7C15ULL + (uint64_t)(epoch + 1) * 0xD6E8FEB86659FD93ULL;
    h ^= h >> 32; h *= 0xD6E8FEB86659FD93ULL; h ^= h >> 32;
    return h;
}

}  // namespace

// ==========================================================================================
// host side
// ==========================================================================================
// rows prefetched into L2 this many nonzeros ahead (GE_PF_DIST overrides; 0 = off; max 32)
static int ge_prefetch_distance() {
    const char* e = getenv("GE_PF_DIST");  // read per launch: a development sweep can change it between epochs
    const int v = e ? atoi(e) : 0;
    return v < 0 ? 0 : ((v % 100) > 32 ? 0 : v);
}

// grid for one resident wave of kernel `k`, bounded for tiny problems (see below)
template <typename K>
static dim3 wave_grid(ge_ctx* ctx, K k, int64_t count, int64_t span, int ids_per_warp = 16) {
    int bps = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k, 256, 0);
    if (bps < 1) bps = 1;
    int64_t want = (count + 32 * 8 - 1) / (32 * 8);  // blocks if every warp took one batch
    int64_t full = (int64_t)ctx->sm_count * bps;     // one resident wave
    // Staleness bound for tiny problems: never more than one concurrent warp per 16 ids of
    // the smaller side (256 for adam / amsgrad, whose bounded sign-like steps tolerate far less
    // staleness), so that a test-sized matrix sees the collision rate of a large one (at bench
    // scale span/16 — or span/256 — is far above a full wave and this never binds).
    int64_t cap = span / ((int64_t)ids_per_warp * 8);
    if (cap < 1) cap = 1;
    if (full > cap) full = cap;
    return dim3((unsigned)(want < full ? (want < 1 ? 1 : want) : full));
}

// Per-launch overrides set by the caller around ge_launch_update (hub segment / its companion launch):
// at most `launch_max_warps` warps (0 = no cap), `launch_reserve_blocks` blocks of the resident wave left
// to a kernel running beside this one.
static void shape_launch(const ge_ctx* ctx, dim3& grid, dim3& block) {
    if (ctx->launch_reserve_blocks > 0) {
        const int64_t g = (int64_t)grid.x - ctx->launch_reserve_blocks;
        grid.x = (unsigned)(g < 1 ? 1 : g);
    }
    if (ctx->launch_max_warps > 0) {
        const int64_t w = ctx->launch_max_warps;
        if (w < 8) { block.x = (unsigned)(32 * w); grid.x = 1; }
        else if ((int64_t)grid.x > w / 8) grid.x = (unsigned)(w / 8);
    }
}

template <bool JAVA, bool AMS>
static int launch_adam_nt(ge_ctx* ctx, int nt, cudaStream_t st, const NzEntry* nz, int64_t count, float* W,
                          float* C, const int32_t* perm, Bijection bij, double* d_cost, int64_t span) {
    const ge_config& c = ctx->cfg;
    // Adam.java:84 — learningRate * sqrt(1 - beta2^(it+1)) / (1 - beta1^(it+1)), fp64, float betas widened
    const double it1 = (double)(ctx->epoch + 1);
    const double corr = (double)c.learning_rate * sqrt(1 - pow((double)0.999f, it1)) / (1 - pow((double)0.9f, it1));
#define GE_LAUNCH(NT_)                                                                             \
    {                                                                                              \
        auto k = adam_update<NT_, JAVA, AMS>;                                                      \
        dim3 block(256), grid;                                                                     \
        if (c.serial) { block = dim3(32); grid = dim3(1); }                                        \
        else grid = wave_grid(ctx, k, count, span, 256);                                           \
        k<<<grid, block, 0, st>>>(nz, count, W, C, perm, bij, ctx->D4, ctx->S, c.method, ctx->xmax, \
                                  c.learning_rate, corr, c.update == GE_UPDATE_ATOMIC_ADAM, d_cost); \
    }
    switch (nt) {
        case 1: GE_LAUNCH(1); break;
        case 2: GE_LAUNCH(2); break;
        case 3: GE_LAUNCH(3); break;
        case 4: GE_LAUNCH(4); break;
        default: ctx->err = "adam / amsgrad: dim too large (max 512)"; return GE_ERR_INVALID;
    }
#undef GE_LAUNCH
    GE_CUDA(cudaGetLastError());
    return GE_OK;
}

int ge_launch_update(ge_ctx* ctx, const NzEntry* nz, int64_t count, float* W, float* C,
                     const int32_t* perm, uint64_t order_key, double* d_cost, cudaStream_t st, int64_t span,
                     bool any_hot, unsigned long long* queue) {
    if (count <= 0) return GE_OK;
    int nt = (ctx->D4 / 4 + 31) / 32;
    if (nt == 5) nt = 6;
    if (nt == 7) nt = 8;
    const bool java = ctx->cfg.math == GE_MATH_JAVA;
    const Bijection bij = Bijection::make((uint64_t)count, order_key);
    ctx->launches++;
    ctx->updates += count;
    if (ctx->cfg.opt_method != GE_OPT_ADAGRAD) {
        const bool ams = ctx->cfg.opt_method == GE_OPT_AMSGRAD;
        if (java) return ams ? launch_adam_nt<true, true>(ctx, nt, st, nz, count, W, C, perm, bij, d_cost, span)
                             : launch_adam_nt<true, false>(ctx, nt, st, nz, count, W, C, perm, bij, d_cost, span);
        return ams ? launch_adam_nt<false, true>(ctx, nt, st, nz, count, W, C, perm, bij, d_cost, span)
                   : launch_adam_nt<false, false>(ctx, nt, st, nz, count, W, C, perm, bij, d_cost, span);
    }
    const int pf = ctx->launch_pf >= 0 ? ctx->launch_pf : ge_prefetch_distance();
    const ge_config& c = ctx->cfg;
    // one instantiation per (NT, math, update mode, load order); grid = one resident wave
#define GE_GO2(NT_, JAVA_, ATOMIC_, TP_, HOT_, DYN_)                                                      \
    {                                                                                                     \
        auto k = glove_adagrad_update<NT_, JAVA_, ATOMIC_, TP_, HOT_, DYN_>;                              \
        dim3 block(256), grid;                                                                            \
        if (c.serial) { block = dim3(32); grid = dim3(1); }                                               \
        else { grid = wave_grid(ctx, k, count, span); shape_launch(ctx, grid, block); }                   \
        k<<<grid, block, 0, st>>>(nz, count, W, C, perm, bij, ctx->D4, ctx->S, c.method, ctx->xmax,       \
                                  c.learning_rate, pf, queue, d_cost);                                    \
    }
#define GE_GO(NT_, JAVA_, ATOMIC_, TP_, HOT_)                                                             \
    {                                                                                                     \
        if (queue && (ATOMIC_)) GE_GO2(NT_, JAVA_, ATOMIC_, TP_, HOT_, (ATOMIC_))                         \
        else GE_GO2(NT_, JAVA_, ATOMIC_, TP_, HOT_, false)                                                \
    }
#define GE_MODE(NT_, TP_)                                                                                 \
    if (java) GE_GO(NT_, true, false, false, false)                                                       \
    else if (c.update == GE_UPDATE_STORE) GE_GO(NT_, false, false, TP_, false)                            \
    else if (any_hot) GE_GO(NT_, false, true, TP_, true)                                                  \
    else GE_GO(NT_, false, true, TP_, false)
    switch (nt) {
        case 1: if (ctx->D4 >= 128) { GE_MODE(1, true) } else { GE_MODE(1, false) } break;
        case 2: GE_MODE(2, true) break;
        case 3: GE_MODE(3, true) break;
        case 4: GE_MODE(4, true) break;
        case 6: GE_MODE(6, true) break;
        case 8: GE_MODE(8, true) break;
        default: ctx->err = "dim too large (max 1024)"; return GE_ERR_INVALID;
    }
#undef GE_MODE
#undef GE_GO
#undef GE_GO2
    GE_CUDA(cudaGetLastError());
    return GE_OK;
}

void ge_free_store(ge_ctx* ctx) {
    cudaFree(ctx->d_W); cudaFree(ctx->d_C); cudaFree(ctx->d_nz); cudaFree(ctx->d_perm); cudaFree(ctx->d_cost);
    ctx->d_W = ctx->d_C = nullptr; ctx->d_nz = nullptr; ctx->d_perm = nullptr; ctx->d_cost = nullptr;
    ctx->have_store = false;
    ctx->h_perm.clear(); ctx->h_perm.shrink_to_fit();
}

// Row geometry: [vec(D4) | NS state vectors(D4 each) | bias | NS state scalars | pad to 8 floats].
// AdaGrad NS = 1 (gradsq); Adam / AMSGrad NS = 2 (first moment, second moment).
void ge_store_geometry(ge_ctx* ctx) {
    ctx->D4 = (ctx->cfg.dim + 3) & ~3;
    ctx->NS = ctx->cfg.opt_method == GE_OPT_ADAGRAD ? 1 : 2;
    ctx->BO = (1 + ctx->NS) * ctx->D4;
    ctx->S = (ctx->BO + 1 + ctx->NS + 7) & ~7;
}

// store_init on caller-provided slabs: `rows` rows whose (virtual, common.cuh ge_vid) ids start at row0 (dist.cu)
int ge_store_init_rows(ge_ctx* ctx, float* W, float* C, int32_t rows, int32_t row0, int32_t lay_nb, int32_t lay_Bs) {
    const ge_config& c = ctx->cfg;
    if (rows <= 0) return GE_OK;
    store_init<<<(rows + 127) / 128, 128, 0, ctx->stream>>>(W, C, rows, c.dim, ctx->D4, ctx->S, c.seed,
                                                          c.init == GE_INIT_JAVA, row0, ctx->NS,
                                                          c.opt_method == GE_OPT_ADAGRAD ? 1.0f : 0.0f, lay_nb, lay_Bs, ctx->V);
    GE_CUDA(cudaGetLastError());
    return GE_OK;
}

extern "C" int ge_store_init(ge_ctx* ctx) {
    if (!ctx) return GE_ERR_INVALID;
    if (ctx->V <= 0) { ctx->err = "no vocabulary: set a graph or import a matrix first"; return GE_ERR_STATE; }
    if (ctx->V > GE_ID_MASK) { ctx->err = "vocabulary must stay below 2^27 vertices (packed nonzero ids)"; return GE_ERR_INVALID; }
    const ge_config& c = ctx->cfg;
    cudaFree(ctx->d_W); cudaFree(ctx->d_C); ctx->d_W = ctx->d_C = nullptr;
    ge_store_geometry(ctx);
    const size_t bytes = (size_t)ctx->V * ctx->S * sizeof(float);
    GE_CUDA(cudaMalloc(&ctx->d_W, bytes));
    GE_CUDA(cudaMalloc(&ctx->d_C, bytes));
    int ri = ge_store_init_rows(ctx, ctx->d_W, ctx->d_C, ctx->V, 0);
    if (ri) return ri;
    // stream position after the init draws: the epoch shuffles continue on the same stream
    // (Optimizer.random and Permutation.random are one thread-local instance)
    ctx->rng = JavaLcg::seeded(c.seed);
    ctx->rng.jump((uint64_t)ctx->V * (2 + 2 * (uint64_t)c.dim));
    ctx->h_perm.clear();
    ctx->epoch = 0;
    ctx->history.clear();
    ctx->have_store = true;
    return GE_OK;
}

namespace {
__global__ void count_ids(const int32_t* __restrict__ ids, int64_t n, int32_t base, int32_t* __restrict__ cnt) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) atomicAdd(&cnt[ids[k] - base], 1);
}
}  // namespace

namespace {
// rows per occurrence-count class (8 classes per octave): input of the L2-resident-set threshold
__global__ void count_classes(const int32_t* __restrict__ cnt, int64_t n, unsigned long long* __restrict__ hist) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) {
        const int32_t c = cnt[k];
        if (c > 0) atomicAdd(&hist[min(255, (int)(8.0f * log2f((float)c)))], 1ULL);
    }
}
}  // namespace

// Occurrence counts from which a row of W (keep_i) / of C (keep_j) joins the L2-resident set: rows
// that occur at least that often are loaded / reduced with an evict_last policy (bit 31 of the packed
// id).  One launch sees about cnt / div of an id's occurrences and `share` of the flagged rows of
// that side (single GPU: div = share = 1; block trainer: a row's nonzeros spread over the column
// blocks, and only one row-block / column-block pair is live per launch).  The threshold is on the
// per-launch count, the same for both sides, as low as the budget GE_L2_KEEP_MB (default 32; 0 = off)
// allows.  0x7fffffff = nothing qualifies.
void ge_keep_counts(ge_ctx* ctx, const int32_t* cnt_i, int64_t ni, double div_i, double share_i,
                    const int32_t* cnt_j, int64_t nj, double div_j, double share_j, int32_t* keep_i, int32_t* keep_j) {
    const double mb = getenv("GE_L2_KEEP_MB") ? atof(getenv("GE_L2_KEEP_MB")) : 32.0;  // read per pack: sweeps change it
    *keep_i = *keep_j = 0x7fffffff;
    if (mb <= 0 || !cnt_i || !cnt_j) return;
    const double budget_rows = mb * 1048576.0 / ((double)ctx->S * 4.0);
    unsigned long long* d_h = nullptr;
    unsigned long long h[512];
    if (cudaMalloc(&d_h, sizeof(h)) != cudaSuccess) { cudaGetLastError(); return; }
    cudaMemsetAsync(d_h, 0, sizeof(h), ctx->stream);
    count_classes<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(cnt_i, ni, d_h);
    count_classes<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(cnt_j, nj, d_h + 256);
    cudaError_t e = cudaMemcpyAsync(h, d_h, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d_h);
    if (e != cudaSuccess) { cudaGetLastError(); return; }
    // rows of a side whose raw count is at least `raw` (classes are 1/8 octave wide: round the class up)
    auto at_least = [&](const unsigned long long* hh, double raw) -> double {
        int cls = raw <= 1.0 ? 0 : (int)ceil(8.0 * log2(raw));
        double rows = 0;
        for (int c = cls < 0 ? 0 : cls; c < 256; c++) rows += (double)hh[c];
        return rows;
    };
    double best = -1;
    for (int t = 8 * 31; t >= 0; t--) {  // per-launch count threshold 2^(t/8), descending
        const double tau = exp2((double)t / 8.0);
        const double rows = share_i * at_least(h, tau * div_i) + share_j * at_least(h + 256, tau * div_j);
        if (rows > budget_rows) break;
        best = tau;
    }
    if (best < 0) return;
    const double ki = ceil(exp2(ceil(8.0 * log2(fmax(1.0, best * div_i))) / 8.0)), kj = ceil(exp2(ceil(8.0 * log2(fmax(1.0, best * div_j))) / 8.0));
    *keep_i = ki > 2e9 ? 0x7fffffff : (int32_t)ki;
    *keep_j = kj > 2e9 ? 0x7fffffff : (int32_t)kj;
}

// occurrences of every id in [base, base+n) among ids[0..nnz): input of the hot-id flags
int ge_count_ids(ge_ctx* ctx, const int32_t* d_ids, int64_t nnz, int32_t base, int64_t n, int32_t** d_cnt) {
    *d_cnt = nullptr;
    GE_CUDA(cudaMalloc(d_cnt, (size_t)(n > 0 ? n : 1) * 4));
    GE_CUDA(cudaMemsetAsync(*d_cnt, 0, (size_t)(n > 0 ? n : 1) * 4, ctx->stream));
    if (nnz > 0) count_ids<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(d_ids, nnz, base, *d_cnt);
    GE_CUDA(cudaGetLastError());
    return GE_OK;
}

// An id is "hot" when, with a full wave of warps drawing nonzeros uniformly from a launch of
// `launch_nnz` nonzeros, more than GE_HOT_WRITERS (default 16) warps are expected to sit on it at
// once: red.add would then ADD that many stale-gradient steps (measured on Zipf(1) columns: clean at 16,
// first-epoch loss 3x too high at 24 inside column blocks, NaN from 32; R-MAT matrices tolerate ~40-70; NaN at
// ~100-250 on Zipf columns), whereas the reference's racy plain stores keep about one of them
// (Adagrad.java:76-80).  Hot ids get a damping exponent (common.cuh ge_damp_bits) that scales their
// deltas so the sum stays below GE_HOT_WRITERS full steps.  Returns the occurrence count from which
// an id is hot.
int32_t ge_hot_count(ge_ctx* ctx, int64_t launch_nnz, int64_t span, int64_t warp_cap) {
    const double writers = getenv("GE_HOT_WRITERS") ? atof(getenv("GE_HOT_WRITERS")) : 16.0;
    double warps = (double)ctx->sm_count * 32.0;        // 4 blocks x 8 warps per SM at most
    if (warps > (double)span / 16.0) warps = (double)span / 16.0;   // wave_grid's cap for tiny problems
    if (warp_cap > 0 && warps > (double)warp_cap) warps = (double)warp_cap;
    if (warps < 1) warps = 1;
    double c = writers * (double)launch_nnz / warps;
    if (c < 1) c = 1;
    return c > 2e9 ? 0x7fffffff : (int32_t)c;
}

// Extreme hubs.  Scaling a hub's deltas (above) is accurate while a few dozen warps share the row —
// every BCA-built matrix of the BASELINE configs — but not when one id holds percents of all
// nonzeros: on a Zipf(1) matrix (5.5 % on one column, ~260 of the 4 736 resident warps on it) the damped path is
// 10-30 % off the single-thread loss in the first epoch, and moving such nonzeros into a launch of
// their own changes the visiting order and with it the online loss (profiles/r02_hubs.md).  The
// reference never has more writers on a row than threads.  So the launch itself is capped — the uniform
// shuffled order is kept — to as many warps as leave a bounded number of writers on the hottest id.  Measured vs
// the single-thread oracle on that matrix, first epoch / later epochs: 16 on the hub 2.0 % / <= 0.6 %, 8 on
// the hub 0.23 % / <= 0.13 %.  Inside the block schedule's buckets, where the hub is 12-50 % of a launch
// instead of 5 %, the same count is less accurate (16: 38 %, 8: 2.0 % / <= 0.4 %), so the buckets use half
// of it (4: 0.45 % / <= 0.14 % against the oracle walking the block order).
// When: the error of damping grows with the share of the launch's nonzeros that touch a damped id, not with
// the hottest id alone — C4's hottest column has ~85 warps on it inside its buckets but the damped ids hold
// at most 3.2 % of a bucket (damped there, 2.8e-5 from the 1-rank walk); on the Zipf matrix they hold
// 25 % of the launch and up to 66 % of a bucket.  The cap engages when that share (`hub_mass`, from the id
// histograms) reaches GE_HUB_MASS (default 0.05).  The epoch then runs at the hub's own pace (count x
// per-nonzero latency / writers), which is what bounded staleness on that row costs.  GE_HUB_SERIAL=0
// switches the cap off.  Returns 0 = no cap.
// How many writers: 8 in the single launch, 4 in a bucket of the block schedule (GE_HUB_SERIAL, GE_HUB_SERIAL_BLOCKS).
// Scaling the number with the hottest id's share of the launch instead (GE_HUB_LOAD / share, between 2 and
// GE_HOT_WRITERS; off unless GE_HUB_LOAD is set) was measured: fine for the single launch (9 writers, 0.38 %) but
// 5.9 % in the first epoch of the block schedule, whose buckets hold several mid-sized hubs each.  The price of
// the fixed 4: a mildly skewed matrix cut very finely (C3 in 8x8x2 buckets: hottest id 2.9 % of a bucket, damped
// ids 12.7 %) runs its buckets on 137 warps.
int64_t ge_warp_cap(ge_ctx* ctx, int64_t launch_nnz, int64_t max_count, double hub_mass, bool block_schedule) {
    const double theta = getenv("GE_HUB_MASS") ? atof(getenv("GE_HUB_MASS")) : 0.05;
    if (max_count <= 0 || launch_nnz <= 0) return 0;
    const double full = (double)ctx->sm_count * 32.0;
    const double share = (double)max_count / (double)launch_nnz;
    const double k = full * share;   // warps of a full wave on the hottest id
    if (k > ctx->hub_kmax) ctx->hub_kmax = k;
    if (hub_mass > ctx->hub_mass_max) ctx->hub_mass_max = hub_mass;
    double serial = getenv("GE_HUB_SERIAL") ? atof(getenv("GE_HUB_SERIAL")) : 8.0;
    if (block_schedule) serial = getenv("GE_HUB_SERIAL_BLOCKS") ? atof(getenv("GE_HUB_SERIAL_BLOCKS")) : 0.5 * serial;
    if (getenv("GE_HUB_LOAD")) {
        const double writers = getenv("GE_HOT_WRITERS") ? atof(getenv("GE_HOT_WRITERS")) : 16.0;
        serial = atof(getenv("GE_HUB_LOAD")) / share;
        serial = serial < 2 ? 2 : (serial > writers ? writers : serial);
    }
    if (serial <= 0 || hub_mass < theta || k <= serial) return 0;
    const double w = serial / share;
    ctx->capped_launches++;
    return w < 1 ? 1 : (int64_t)w;
}

// hottest id over all launches shaped since the last prepare: expected concurrent warps on it in an
// uncapped wave, and how many launches (1 for the single-GPU trainer, buckets for the block schedule) are capped
extern "C" int ge_train_hub_info(ge_ctx* ctx, double* max_concurrency, double* damped_share, int32_t* capped_launches, int64_t* warp_cap) {
    if (!ctx) return GE_ERR_INVALID;
    if (max_concurrency) *max_concurrency = ctx->hub_kmax;
    if (damped_share) *damped_share = ctx->hub_mass_max;
    if (capped_launches) *capped_launches = ctx->capped_launches;
    if (warp_cap) *warp_cap = ctx->warp_cap;
    return GE_OK;
}

int ge_train_prepare(ge_ctx* ctx) {
    if (!ctx->have_cooc || ctx->nnz <= 0) { ctx->err = "no co-occurrence matrix: build or import one first"; return GE_ERR_STATE; }
    if (!ctx->d_nz && !ctx->d_I) { ctx->err = "the COO arrays were released by ge_dist_prepare: use ge_dist_train_epoch"; return GE_ERR_STATE; }
    if (!ctx->have_store) { int r = ge_store_init(ctx); if (r) return r; }
    const ge_config& c = ctx->cfg;
    if (!ctx->d_nz) {
        GE_CUDA(cudaMalloc(&ctx->d_nz, (size_t)ctx->nnz * sizeof(NzEntry)));
        int32_t *cnt_i = nullptr, *cnt_j = nullptr;
        if (c.math == GE_MATH_FAST && c.update == GE_UPDATE_ATOMIC && c.opt_method == GE_OPT_ADAGRAD) {
            int r = ge_count_ids(ctx, ctx->d_I, ctx->nnz, 0, ctx->V, &cnt_i);
            if (!r) r = ge_count_ids(ctx, ctx->d_J, ctx->nnz, 0, ctx->V, &cnt_j);
            if (r) { cudaFree(cnt_i); cudaFree(cnt_j); return r; }
        }
        int32_t keep_i, keep_j;
        ge_keep_counts(ctx, cnt_i, ctx->V, 1.0, 1.0, cnt_j, ctx->V, 1.0, 1.0, &keep_i, &keep_j);
        ctx->keep_count = keep_i;
        int32_t* d_flags = nullptr;   // [0] any damped id, [1] max occurrence count, [2..3] u64 damped mass
        GE_CUDA(cudaMalloc(&d_flags, 16));
        cudaMemsetAsync(d_flags, 0, 16, ctx->stream);
        int32_t fl[4] = {0, 0, 0, 0};
        cudaError_t pe = cudaSuccess;
        ctx->warp_cap = 0;
        ctx->hub_kmax = 0; ctx->hub_mass_max = 0; ctx->capped_launches = 0;
        if (cnt_i) {
            const int32_t hot0 = ge_hot_count(ctx, ctx->nnz, ctx->V, 0);   // uncapped wave
            unsigned long long* d_mass = reinterpret_cast<unsigned long long*>(d_flags + 2);
            max_count<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(cnt_i, ctx->V, d_flags + 1);
            max_count<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(cnt_j, ctx->V, d_flags + 1);
            hot_mass<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(cnt_i, ctx->V, hot0, d_mass);
            hot_mass<<<ctx->sm_count * 4, 256, 0, ctx->stream>>>(cnt_j, ctx->V, hot0, d_mass);
            pe = cudaMemcpyAsync(fl, d_flags, 16, cudaMemcpyDeviceToHost, ctx->stream);
            if (pe == cudaSuccess) pe = cudaStreamSynchronize(ctx->stream);
            unsigned long long mass = 0;
            memcpy(&mass, fl + 2, 8);
            if (pe == cudaSuccess) ctx->warp_cap = ge_warp_cap(ctx, ctx->nnz, fl[1], (double)mass / (double)ctx->nnz, false);
            fl[0] = 0;
            if (pe == cudaSuccess) pe = cudaMemsetAsync(d_flags, 0, 4, ctx->stream);
        }
        if (pe == cudaSuccess) {
            pack_nonzeros<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(ctx->d_I, ctx->d_J, ctx->d_X, ctx->nnz, c.method,
                                                                     c.math == GE_MATH_JAVA, ctx->xmax, cnt_i, cnt_j,
                                                                     ge_hot_count(ctx, ctx->nnz, ctx->V, ctx->warp_cap), keep_i, keep_j,
                                                                     d_flags, ctx->d_nz);
            pe = cudaGetLastError();
        }
        if (pe == cudaSuccess) pe = cudaMemcpyAsync(fl, d_flags, 4, cudaMemcpyDeviceToHost, ctx->stream);
        if (pe == cudaSuccess) pe = cudaStreamSynchronize(ctx->stream);
        cudaFree(cnt_i); cudaFree(cnt_j); cudaFree(d_flags);
        GE_CUDA(pe);
        ctx->any_hot = fl[0] != 0;
    }
    if (!ctx->d_cost) GE_CUDA(cudaMalloc(&ctx->d_cost, sizeof(double)));
    if (c.order == GE_ORDER_JAVA) {
        if (ctx->nnz > 0x7fffffffLL) { ctx->err = "GE_ORDER_JAVA needs nnz < 2^31 (the reference indexes with int)"; return GE_ERR_INVALID; }
        if (ctx->h_perm.empty()) {
            ctx->h_perm.resize(ctx->nnz);
            for (int64_t k = 0; k < ctx->nnz; k++) ctx->h_perm[k] = (int32_t)k;  // Permutation.java:11-15
        }
        if (!ctx->d_perm) GE_CUDA(cudaMalloc(&ctx->d_perm, (size_t)ctx->nnz * sizeof(int32_t)));
    }
    return GE_OK;
}

extern "C" int ge_train_epoch(ge_ctx* ctx, double* cost) {
    if (!ctx) return GE_ERR_INVALID;
    int r = ge_train_prepare(ctx);
    if (r) return r;
    const ge_config& c = ctx->cfg;
    const int32_t* perm = nullptr;
    if (c.order == GE_ORDER_JAVA) {
        // coMatrix.shuffle() (Optimizer.java:79): cumulative Fisher-Yates, ExtendedRandom.java:398-407
        const int32_t n = (int32_t)ctx->nnz;
        int32_t* a = ctx->h_perm.data();
        for (int32_t i = 0; i < n; i++) {
            const int32_t rr = i + ctx->rng.nextInt(n - i);
            const int32_t t = a[i]; a[i] = a[rr]; a[rr] = t;
        }
        GE_CUDA(cudaMemcpyAsync(ctx->d_perm, a, (size_t)n * sizeof(int32_t), cudaMemcpyHostToDevice, ctx->stream));
        perm = ctx->d_perm;
    }
    GE_CUDA(cudaMemsetAsync(ctx->d_cost, 0, sizeof(double), ctx->stream));
    GE_CUDA(cudaEventRecord(ctx->ev0, ctx->stream));
    ctx->launch_max_warps = ctx->warp_cap;           // 0 = none; see ge_warp_cap
    ctx->launch_pf = ctx->warp_cap > 0 ? 8 : -1;     // few warps are latency-bound: pull the rows 8 nonzeros ahead into L2
    r = ge_launch_update(ctx, ctx->d_nz, ctx->nnz, ctx->d_W, ctx->d_C, perm, epoch_key(c.seed, ctx->epoch),
                         ctx->d_cost, ctx->stream, ctx->V, ctx->any_hot, nullptr);
    ctx->launch_max_warps = 0;
    ctx->launch_pf = -1;
    if (r) return r;
    GE_CUDA(cudaEventRecord(ctx->ev1, ctx->stream));
    double total = 0;
    GE_CUDA(cudaMemcpyAsync(&total, ctx->d_cost, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    GE_CUDA(cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    ctx->update_seconds += ms * 1e-3;
    const double local = total / (double)ctx->nnz;  // Optimizer.java:96
    ctx->history.push_back(local);
    ctx->epoch++;
    if (cost) *cost = local;
    return GE_OK;
}

extern "C" int ge_train(ge_ctx* ctx, ge_epoch_cb cb, void* user, double* final_cost, int32_t* epochs) {
    if (!ctx) return GE_ERR_INVALID;
    // Optimizer.optimize :67-120
    double finalCost = 0, prevCost = 0;
    int32_t it = 0;
    for (; it < ctx->cfg.maxiter; it++) {
        double local = 0;
        const double before = ctx->update_seconds;
        int r = ge_train_epoch(ctx, &local);
        if (r) return r;
        const double diff = fabs(prevCost - local);  // :98
        prevCost = local;                             // :102
        const bool stop = cb && cb(user, it, local, ctx->update_seconds - before);
        if (diff <= ctx->cfg.tolerance) { finalCost = local; it++; break; }  // :104-107
        if (stop) { it++; break; }
    }
    if (final_cost) *final_cost = finalCost;
    if (epochs) *epochs = it;
    return GE_OK;
}

extern "C" int ge_cost_history(ge_ctx* ctx, double* out, int32_t cap) {
    if (!ctx) return GE_ERR_INVALID;
    const int32_t n = (int32_t)ctx->history.size();
    for (int32_t k = 0; k < n && k < cap; k++) out[k] = ctx->history[k];
    return n;
}

// The result goes to pageable host memory (a Java double[] / numpy array), which the driver would
// copy at 3-4 GB/s through its own staging.  Instead: row chunks through two pinned staging buffers;
// while chunk c crosses PCIe, host threads move chunk c-1 from staging into `out` (they also take the
// first-touch page faults of a freshly allocated destination in parallel).  The device always produces
// f32 — (float)(f + c) / 2 IS the reference's value (Optimizer.java:130-140) — and the widening to the
// double[] the seam returns happens in that host pass: half the PCIe bytes, identical bits.
// The result buffer is fresh anonymous memory of many GB (double[V * dim]): written through 4 KB first-touch
// faults the host side of the extract ran at ~5 GB/s (2.9 s of a 4.4 s configs[4] job).  Ask for
// transparent huge pages on the 2 MB-aligned interior; a no-op where THP is off.
void ge_advise_result_buffer(void* p, size_t bytes) {
    const uintptr_t huge = (uintptr_t)2 << 20;
    const uintptr_t a = ((uintptr_t)p + huge - 1) & ~(huge - 1), b = ((uintptr_t)p + bytes) & ~(huge - 1);
    if (bytes >= ((size_t)64 << 20) && b > a) madvise((void*)a, b - a, MADV_HUGEPAGE);
}

struct ExtractStage {
    float* d[2] = {nullptr, nullptr};
    float* h[2] = {nullptr, nullptr};
    cudaEvent_t ev[2] = {nullptr, nullptr};
    int64_t chunk_rows = 0;
    cudaError_t init(int32_t dim, int64_t max_rows) {
        chunk_rows = (int64_t)((size_t)(48u << 20) / ((size_t)dim * sizeof(float)));
        if (chunk_rows < 1) chunk_rows = 1;
        if (chunk_rows > max_rows) chunk_rows = max_rows > 0 ? max_rows : 1;
        cudaError_t e = cudaSuccess;
        for (int k = 0; k < 2 && e == cudaSuccess; k++) {
            e = cudaMalloc(&d[k], (size_t)chunk_rows * dim * sizeof(float));
            if (e == cudaSuccess) e = cudaMallocHost(&h[k], (size_t)chunk_rows * dim * sizeof(float));
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&ev[k], cudaEventDisableTiming);
        }
        return e;
    }
    ~ExtractStage() {
        for (int k = 0; k < 2; k++) {
            cudaFree(d[k]);
            cudaFreeHost(h[k]);
            if (ev[k]) cudaEventDestroy(ev[k]);
        }
    }
};

// rows [0, rows) of the slabs W / C -> out (rows x dim, row-major)
template <typename T>
static cudaError_t extract_rows(ge_ctx* ctx, ExtractStage& S, const float* W, const float* C, int64_t rows, T* out) {
    const int32_t dim = ctx->cfg.dim;
    // host threads of the drain: a handle that owns the machine (single-GPU trainer) uses up to 16, a rank of the
    // block schedule shares the host with its peers and stays at 8; GE_HOST_THREADS overrides both
    unsigned nthr = std::thread::hardware_concurrency();
    const unsigned cap = ctx->dist ? 8u : 16u;
    nthr = nthr < 1 ? 1 : (nthr > cap ? cap : nthr);
    if (const char* ev = getenv("GE_HOST_THREADS")) { const int v = atoi(ev); if (v >= 1 && v <= 64) nthr = (unsigned)v; }
    auto drain = [&](int64_t c) {  // chunk c: staging -> out
        const int64_t row0 = c * S.chunk_rows;
        const size_t n = (size_t)((rows - row0) < S.chunk_rows ? (rows - row0) : S.chunk_rows) * dim;
        const float* src = S.h[c & 1];
        T* dst = out + (size_t)row0 * dim;
        auto move = [=](size_t lo, size_t hi) {
            if (sizeof(T) == sizeof(float)) memcpy((void*)(dst + lo), src + lo, (hi - lo) * sizeof(float));
            else for (size_t k = lo; k < hi; k++) dst[k] = (T)src[k];
        };
        if (n < ((size_t)2 << 20) || nthr == 1) { move(0, n); return; }
        std::vector<std::thread> th;
        for (unsigned t = 0; t < nthr; t++) th.emplace_back(move, n * t / nthr, n * (t + 1) / nthr);
        for (auto& x : th) x.join();
    };
    const int64_t nchunk = (rows + S.chunk_rows - 1) / S.chunk_rows;
    cudaError_t e = cudaSuccess;
    for (int64_t c = 0; c < nchunk && e == cudaSuccess; c++) {
        const int64_t row0 = c * S.chunk_rows;
        const int32_t nr = (int32_t)((rows - row0) < S.chunk_rows ? (rows - row0) : S.chunk_rows);
        // slot c&1 was drained two iterations ago (drain runs on this thread), so it is free
        extract_avg<float><<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(W, C, row0, nr, dim, ctx->S, S.d[c & 1]);
        e = cudaMemcpyAsync(S.h[c & 1], S.d[c & 1], (size_t)nr * dim * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaEventRecord(S.ev[c & 1], ctx->stream);
        if (c > 0 && e == cudaSuccess) {
            e = cudaEventSynchronize(S.ev[(c - 1) & 1]);
            if (e == cudaSuccess) drain(c - 1);
        }
    }
    if (e == cudaSuccess && nchunk > 0) {
        e = cudaEventSynchronize(S.ev[(nchunk - 1) & 1]);
        if (e == cudaSuccess) drain(nchunk - 1);
    }
    return e;
}

template <typename T>
static int extract_impl(ge_ctx* ctx, T* out) {
    if (!ctx || !out) return GE_ERR_INVALID;
    if (!ctx->have_store) { ctx->err = "no embedding store"; return GE_ERR_STATE; }
    if (ctx->dist) { ctx->err = "block-rotation trainer: use ge_dist_extract / ge_dist_extract_f32"; return GE_ERR_STATE; }
    cudaSetDevice(ctx->cfg.device);
    ExtractStage S;
    ge_advise_result_buffer(out, (size_t)ctx->V * ctx->cfg.dim * sizeof(T));
    cudaError_t e = S.init(ctx->cfg.dim, ctx->V);
    if (e == cudaSuccess) e = extract_rows<T>(ctx, S, ctx->d_W, ctx->d_C, ctx->V, out);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    GE_CUDA(e);
    return GE_OK;
}
extern "C" int ge_extract(ge_ctx* ctx, double* out) { return extract_impl<double>(ctx, out); }
extern "C" int ge_extract_f32(ge_ctx* ctx, float* out) { return extract_impl<float>(ctx, out); }

// block trainer (dist.cu): `nblocks` (W slab, C slab) pairs of `rows[k]` rows each, concatenated into out
template <typename T>
int ge_extract_blocks(ge_ctx* ctx, int nblocks, const float* const* W, const float* const* C, const int64_t* rows, T* out) {
    int64_t mx = 1;
    for (int k = 0; k < nblocks; k++) mx = rows[k] > mx ? rows[k] : mx;
    ExtractStage S;
    cudaError_t e = S.init(ctx->cfg.dim, mx);
    for (int k = 0; k < nblocks && e == cudaSuccess; k++) {
        e = extract_rows<T>(ctx, S, W[k], C[k], rows[k], out);
        out += (size_t)rows[k] * ctx->cfg.dim;
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    GE_CUDA(e);
    return GE_OK;
}
template int ge_extract_blocks<float>(ge_ctx*, int, const float* const*, const float* const*, const int64_t*, float*);
template int ge_extract_blocks<double>(ge_ctx*, int, const float* const*, const float* const*, const int64_t*, double*);

extern "C" int ge_store_export(ge_ctx* ctx, int32_t which, float* out) {
    if (!ctx || !out || which < 0 || which > 11) return GE_ERR_INVALID;
    if (!ctx->have_store) { ctx->err = "no embedding store"; return GE_ERR_STATE; }
    const int32_t dim = ctx->cfg.dim, D4 = ctx->D4, BO = ctx->BO, NS = ctx->NS;
    const float* A = (which & 1) ? ctx->d_C : ctx->d_W;
    int32_t off, width;
    switch (which >> 1) {
        case 0: off = 0; width = dim; break;             // focus / context
        case 1: off = BO; width = 1; break;              // fBias / cBias
        case 2: off = NS * D4; width = dim; break;       // gradSq (AdaGrad) or second moment M2 (Adam/AMSGrad)
        case 3: off = BO + NS; width = 1; break;         // same, bias scalars
        case 4: off = D4; width = dim; break;            // first moment M1 (Adam/AMSGrad only)
        default: off = BO + 1; width = 1; break;         // M1 of the biases
    }
    if (which >= 8 && NS < 2) { ctx->err = "first moments exist only for adam / amsgrad"; return GE_ERR_INVALID; }
    const size_t n = (size_t)ctx->V * width;
    float* d = nullptr;
    GE_CUDA(cudaMalloc(&d, n * sizeof(float)));
    store_component<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(A, ctx->V, dim, ctx->S, off, width, d);
    cudaError_t e = cudaMemcpyAsync(out, d, n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    GE_CUDA(e);
    return GE_OK;
}

extern "C" int ge_epoch_order(ge_ctx* ctx, int32_t epoch, int32_t* perm, int64_t cap) {
    if (!ctx || !perm) return GE_ERR_INVALID;
    if (!ctx->have_cooc || cap < ctx->nnz || ctx->nnz > 0x7fffffffLL) { ctx->err = "bad capacity / no matrix"; return GE_ERR_INVALID; }
    if (ctx->cfg.order == GE_ORDER_JAVA) {
        if (ctx->h_perm.empty()) { ctx->err = "GE_ORDER_JAVA: order exists only after an epoch has run"; return GE_ERR_STATE; }
        for (int64_t k = 0; k < ctx->nnz; k++) perm[k] = ctx->h_perm[k];
        return GE_OK;
    }
    int32_t* d = nullptr;
    GE_CUDA(cudaMalloc(&d, (size_t)ctx->nnz * sizeof(int32_t)));
    materialise_order<<<ctx->sm_count * 8, 256, 0, ctx->stream>>>(Bijection::make((uint64_t)ctx->nnz, epoch_key(ctx->cfg.seed, epoch)), ctx->nnz, d, 0);
    cudaError_t e = cudaMemcpyAsync(perm, d, (size_t)ctx->nnz * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
    cudaFree(d);
    GE_CUDA(e);
    return GE_OK;
}

// forget the packed nonzeros and the embedding store: the next epoch re-packs the matrix (re-reading the
// tuning environment: GE_L2_KEEP_MB, GE_HOT_WRITERS) and starts from a fresh initialisation
extern "C" int ge_train_reset(ge_ctx* ctx) {
    if (!ctx) return GE_ERR_INVALID;
    cudaSetDevice(ctx->cfg.device);
    ge_free_store(ctx);
    ctx->epoch = 0;
    ctx->history.clear();
    return GE_OK;
}

extern "C" int ge_train_stats(ge_ctx* ctx, double* update_seconds, int64_t* launches, int64_t* updates) {
    if (!ctx) return GE_ERR_INVALID;
    if (update_seconds) *update_seconds = ctx->update_seconds;
    if (launches) *launches = ctx->launches;
    if (updates) *updates = ctx->updates;
    return GE_OK;
}
