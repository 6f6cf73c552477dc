// Random-row bandwidth microbenchmark for the access pattern of glove_adagrad_update (train.cu):
// every "nonzero" reads two rows of S floats (S*4 = 1 056 B at d=128) at unrelated addresses of a
// store much larger than L2 and adds S floats back into each.  No arithmetic worth the name: this is
// the ceiling of the pattern, to be set beside the update kernel's achieved bytes/s
// (profiles/r02_random_rows.md).  Variants:
//   ldg  : warp per nonzero, ld.global.cg.v4 + red.global.add.v4.f32 (what the kernel issues), U nonzeros
//          in flight per warp
//   tma  : warp per nonzero stream, rows staged global -> shared with cp.async.bulk + mbarrier in a ring of
//          G stages per warp, written back with cp.reduce.async.bulk (.add.f32): no row registers, one
//          instruction per row each way
//   read : ldg variant without the write-back (read-only ceiling)
// Build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o scripts/rowbench scripts/rowbench.cu
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_)); exit(1); } } while (0)

__device__ __forceinline__ uint64_t mix(uint64_t x) {
    x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
    return x;
}
__device__ __forceinline__ float4 ldcg4(const float* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void red4(float* p, float4 v) {
    asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// ---- ldg / read ---------------------------------------------------------------------------------
template <int U, bool WRITE>
__global__ void __launch_bounds__(256) k_ldg(float* __restrict__ W, float* __restrict__ C, int64_t V, int S, int64_t n,
                                             uint64_t key, float* __restrict__ sink) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nw = ((int64_t)gridDim.x * blockDim.x) >> 5;
    const int per = S / 4;  // float4 per row
    float acc = 0.f;
    for (int64_t k0 = warp * U; k0 < n; k0 += nw * U) {
        float4 a[U][3], b[U][3];   // S <= 384 floats: up to 3 float4 per lane
        float* pw[U]; float* pc[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const uint64_t h = mix((uint64_t)(k0 + u) ^ key);
            pw[u] = W + (int64_t)(h % (uint64_t)V) * S;
            pc[u] = C + (int64_t)(mix(h) % (uint64_t)V) * S;
#pragma unroll
            for (int t = 0; t < 3; t++) {
                const int q = lane + 32 * t;
                if (q < per) { a[u][t] = ldcg4(pw[u] + 4 * q); b[u][t] = ldcg4(pc[u] + 4 * q); }
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
#pragma unroll
            for (int t = 0; t < 3; t++) {
                const int q = lane + 32 * t;
                if (q < per) {
                    if (WRITE) {
                        float4 x = a[u][t], y = b[u][t];
                        x.x *= 1e-9f; x.y *= 1e-9f; x.z *= 1e-9f; x.w *= 1e-9f;
                        y.x *= 1e-9f; y.y *= 1e-9f; y.z *= 1e-9f; y.w *= 1e-9f;
                        red4(pw[u] + 4 * q, y); red4(pc[u] + 4 * q, x);
                    } else acc += a[u][t].x + b[u][t].w;
                }
            }
        }
    }
    if (!WRITE && acc == 123.456f) *sink = acc;
}

// ---- tma ----------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n.reg .pred p;\nWAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}\n" :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_red_s2g(void* dst, const void* src, uint32_t bytes) {
    asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;"
                 :: "l"(dst), "r"(smem_u32(src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" :: "n"(N) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ring of G stages per warp; a stage = the two rows of one nonzero (2*S floats).  Lane 0 is the
// producer (bulk loads G-1 nonzeros ahead) and the issuer of the bulk reductions; all lanes "compute".
template <int G, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k_tma(float* __restrict__ W, float* __restrict__ C, int64_t V, int S, int64_t n,
                                                    uint64_t key) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const int64_t warp = (int64_t)blockIdx.x * WARPS + wib, nw = (int64_t)gridDim.x * WARPS;
    const uint32_t row_bytes = (uint32_t)S * 4;
    float* ring = reinterpret_cast<float*>(smem) + (size_t)wib * G * 2 * S;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)WARPS * G * 2 * S * 4) + wib * G;
    if (lane == 0) for (int g = 0; g < G; g++) mbar_init(bars + g, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    const int64_t mine = warp < n ? (n - warp + nw - 1) / nw : 0;   // nonzeros warp, warp+nw, ...
    auto rows = [&](int64_t t, float*& pw, float*& pc) {
        const uint64_t h = mix((uint64_t)(warp + t * nw) ^ key);
        pw = W + (int64_t)(h % (uint64_t)V) * S;
        pc = C + (int64_t)(mix(h) % (uint64_t)V) * S;
    };
    auto issue = [&](int64_t t) {  // lane 0 only
        const int g = (int)(t % G);
        float *pw, *pc;
        rows(t, pw, pc);
        mbar_expect(bars + g, 2 * row_bytes);
        bulk_g2s(ring + (size_t)g * 2 * S, pw, row_bytes, bars + g);
        bulk_g2s(ring + (size_t)g * 2 * S + S, pc, row_bytes, bars + g);
    };
    if (lane == 0) for (int64_t t = 0; t < G - 1 && t < mine; t++) issue(t);
    for (int64_t t = 0; t < mine; t++) {
        const int g = (int)(t % G);
        if (lane == 0 && t + G - 1 < mine) {
            // stage (t+G-1)%G == (t-1)%G was handed to a bulk reduction in iteration t-1: its reads of
            // shared memory must be over before the next load lands there
            bulk_wait_read<0>();
            issue(t + G - 1);
        }
        mbar_wait(bars + g, (uint32_t)((t / G) & 1));
        float* st = ring + (size_t)g * 2 * S;
        // "compute": scale in place (generic proxy), swap roles like the kernel (delta of W from C's row)
        for (int q = lane; q < 2 * S / 4; q += 32) {
            float4 v = reinterpret_cast<float4*>(st)[q];
            v.x *= 1e-9f; v.y *= 1e-9f; v.z *= 1e-9f; v.w *= 1e-9f;
            reinterpret_cast<float4*>(st)[q] = v;
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
            float *pw, *pc;
            rows(t, pw, pc);
            bulk_red_s2g(pw, st, row_bytes);
            bulk_red_s2g(pc, st + S, row_bytes);
            bulk_commit();
        }
    }
    if (lane == 0) bulk_wait_read<0>();
}

template <typename F>
static float time_ms(F launch, int reps) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(); launch();
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int r = 0; r < reps; r++) launch();
    CK(cudaEventRecord(e1));
    CK(cudaDeviceSynchronize());
    float ms = 0;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    return ms / reps;
}

int main(int argc, char** argv) {
    const int64_t V = argc > 1 ? atoll(argv[1]) : 10000000;
    const int dim = argc > 2 ? atoi(argv[2]) : 128;
    const int64_t n = argc > 3 ? atoll(argv[3]) : 400000000;
    const int S = (2 * dim + 9) / 8 * 8;   // train.cu row layout: values, gradsq, bias, bias gradsq, padded to 32 B
    int dev = 0, sms = 0;
    CK(cudaGetDevice(&dev));
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    float *W, *C, *sink;
    CK(cudaMalloc(&W, (size_t)V * S * 4)); CK(cudaMalloc(&C, (size_t)V * S * 4)); CK(cudaMalloc(&sink, 4));
    CK(cudaMemset(W, 0, (size_t)V * S * 4)); CK(cudaMemset(C, 0, (size_t)V * S * 4));
    printf("{\"V\": %lld, \"dim\": %d, \"row_bytes\": %d, \"nonzeros\": %lld, \"store_gb\": %.1f, \"sms\": %d, \"results\": [\n",
           (long long)V, dim, S * 4, (long long)n, 2.0 * V * S * 4 / 1e9, sms);
    const double rd = 2.0 * S * 4, rw = 4.0 * S * 4;
    uint64_t key = 0x9E3779B97F4A7C15ULL;
    bool first = true;
    auto report = [&](const char* name, int bps, float ms, double bytes) {
        printf("%s  {\"variant\": \"%s\", \"blocks_per_sm\": %d, \"ms\": %.2f, \"gbs\": %.1f, \"gnz_per_s\": %.3f}", first ? "" : ",\n", name, bps, ms,
               bytes * n / ms / 1e6, n / ms / 1e6);
        first = false;
        fflush(stdout);
    };
#define LDG(U_, WRITE_, NAME_)                                                                                   \
    {                                                                                                            \
        int bps = 0;                                                                                             \
        CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_ldg<U_, WRITE_>, 256, 0));                      \
        float ms = time_ms([&] { k_ldg<U_, WRITE_><<<sms * bps, 256>>>(W, C, V, S, n, key++, sink); }, 3);       \
        CK(cudaGetLastError());                                                                                  \
        report(NAME_, bps, ms, WRITE_ ? rw : rd);                                                                \
    }
    if (S <= 384) {
        LDG(1, true, "ldg_red_u1") LDG(2, true, "ldg_red_u2")
        LDG(1, false, "ldg_read_u1") LDG(2, false, "ldg_read_u2")
    }
#define TMA(G_, WARPS_, NAME_)                                                                                   \
    {                                                                                                            \
        const size_t sm = (size_t)WARPS_ * G_ * 2 * S * 4 + (size_t)WARPS_ * G_ * 8;                             \
        if (sm <= 227 * 1024) {                                                                                  \
            CK(cudaFuncSetAttribute(k_tma<G_, WARPS_>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));   \
            int bps = 0;                                                                                         \
            CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&bps, k_tma<G_, WARPS_>, WARPS_ * 32, sm));         \
            if (bps > 0) {                                                                                       \
                float ms = time_ms([&] { k_tma<G_, WARPS_><<<sms * bps, WARPS_ * 32, sm>>>(W, C, V, S, n, key++); }, 3); \
                CK(cudaGetLastError());                                                                          \
                report(NAME_, bps, ms, rw);                                                                      \
            }                                                                                                    \
        }                                                                                                        \
    }
    TMA(2, 8, "tma_g2_w8") TMA(3, 8, "tma_g3_w8") TMA(4, 8, "tma_g4_w8") TMA(4, 4, "tma_g4_w4") TMA(6, 4, "tma_g6_w4") TMA(8, 4, "tma_g8_w4")
    TMA(4, 16, "tma_g4_w16") TMA(2, 16, "tma_g2_w16")
    printf("\n]}\n");
    return 0;
}
