// microbench.cu -- measured per-pipe issue rates on the box, used as roofline denominators where
// MEASURED_PEAKS.json has none (integer POPC for the Hamming kernel, fp64 for RANSAC, alu/fma/TMEM-read for the
// tensor-core epilogue).  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o microbench microbench.cu
// Output: one JSON object on stdout (profiles/r1_microbench.json is a copy of one run).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

constexpr int ITERS = 4096;
constexpr int UNROLL = 8;      // independent chains per thread

__global__ void k_popc(uint32_t* out, uint32_t seed)
{
    uint32_t a[UNROLL], acc[UNROLL];
    for (int i = 0; i < UNROLL; ++i) { a[i] = seed + threadIdx.x * 977u + i * 131u; acc[i] = i; }
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < UNROLL; ++i) { acc[i] = __popc(a[i] ^ acc[i]) + acc[i]; }   // POPC + LOP + IADD per step
    }
    uint32_t s = 0;
    for (int i = 0; i < UNROLL; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void k_popc_only(uint32_t* out, uint32_t seed)
{
    uint32_t acc[UNROLL];
    for (int i = 0; i < UNROLL; ++i) acc[i] = seed + threadIdx.x * 977u + i * 131u;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < UNROLL; ++i) acc[i] = __popc(acc[i]) | 0x10000u;            // dependent POPC chain x UNROLL (LOP is cheap)
    }
    uint32_t s = 0;
    for (int i = 0; i < UNROLL; ++i) s += acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>   // 0 DFMA, 1 DADD, 2 DMUL, 3 DDIV (IEEE, __ddiv_rn)
__global__ void k_f64(double* out, double seed)
{
    double a[UNROLL];
    const double b = 1.0000001 + seed, c = 1e-9;
    for (int i = 0; i < UNROLL; ++i) a[i] = seed + threadIdx.x * 1e-3 + i;
    const int n = OP == 3 ? ITERS / 16 : ITERS;
    for (int it = 0; it < n; ++it) {
#pragma unroll
        for (int i = 0; i < UNROLL; ++i) {
            if (OP == 0) a[i] = fma(a[i], b, c);
            if (OP == 1) a[i] = __dadd_rn(a[i], c);
            if (OP == 2) a[i] = __dmul_rn(a[i], b);
            if (OP == 3) a[i] = __ddiv_rn(b, a[i]);
        }
    }
    double s = 0;
    for (int i = 0; i < UNROLL; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int OP>   // 0 FMNMX, 1 FMNMX3, 2 FADD, 3 FFMA, 4 SHF funnel
__global__ void k_f32(float* out, float seed)
{
    float a[UNROLL];
    const float b = 1.0000001f + seed, c = 1e-9f;
    for (int i = 0; i < UNROLL; ++i) a[i] = seed + threadIdx.x * 1e-3f + i;
    for (int it = 0; it < ITERS; ++it) {
#pragma unroll
        for (int i = 0; i < UNROLL; ++i) {
            if (OP == 0) a[i] = fminf(a[i], b + (float)it);
            if (OP == 1) { asm volatile("min.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(b), "f"(c)); }
            if (OP == 2) a[i] = __fadd_rn(a[i], c);
            if (OP == 3) a[i] = __fmaf_rn(a[i], b, c);
            if (OP == 4) a[i] = __uint_as_float(__funnelshift_l(__float_as_uint(b), __float_as_uint(a[i]), 1));
        }
    }
    float s = 0;
    for (int i = 0; i < UNROLL; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// TMEM read rate: every warp of the CTA re-reads 32 lanes x 32 columns (tcgen05.ld.32x32b.x32) of its own quadrant.
__global__ void k_ldtm(uint32_t* out, int iters)
{
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t base = slot + ((uint32_t)((warp & 3) * 32) << 16);
    uint32_t acc = 0;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            uint32_t r[32];
            asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                         "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                         "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                         : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                           "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                           "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                           "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                         : "r"(base + ((c * 32 + it * 128) & 511)) : "memory");
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            acc ^= r[0] ^ r[31];
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(slot), "r"(512) : "memory");
}

// tcgen05.mma issue rate: one thread issues `n` back-to-back MMAs (kind::f16, bf16 operands, M = 128, K = 16) on zeroed
// shared-memory operands (K-major, 128-byte swizzle), `chains` accumulators used round-robin, then one commit.
template <int TF32>
__global__ void k_umma(long long* cycles, int n, int N, int chains, int commit_every = 0, const unsigned char* gsrc = nullptr, int tma_kb_per_mma_x16 = 0)
{
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);
    __shared__ uint32_t slot;
    __shared__ uint64_t bar, bar2;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1000000;" ::"r"((uint32_t)__cvta_generic_to_shared(&bar2)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    __shared__ uint64_t tbar;
    __shared__ volatile int s_done;
    if (threadIdx.x == 32) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tbar)) : "memory"); s_done = 0; }
    __syncthreads();
    if (warp == 1 && gsrc && threadIdx.x == 32) {
        // concurrent TMA traffic into another shared-memory region: 16 KB bulk copies back to back until the MMA warp is done
        // (tma_kb_per_mma_x16 = pacing: after each copy wait so that the average is x/16 KB per MMA issued at 64 clk)
        unsigned char* dst = smem + 16384 + 32768;
        uint32_t ph = 0;
        while (!s_done) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&tbar)), "r"(16384) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(gsrc + (size_t)blockIdx.x * 16384), "r"(16384), "r"((uint32_t)__cvta_generic_to_shared(&tbar)) : "memory");
            asm volatile("{\n\t.reg .pred p;\n\tWT: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra DT;\n\tbra WT;\n\tDT:\n\t}"
                         ::"r"((uint32_t)__cvta_generic_to_shared(&tbar)), "r"(ph) : "memory");
            ph ^= 1;
            if (tma_kb_per_mma_x16 > 0) { const long long t1 = clock64() + (long long)(16 * 16 * 64) / tma_kb_per_mma_x16 ; while (clock64() < t1 && !s_done) {} }
        }
    }
    if (warp == 0) {                                   // whole warp runs the loop (uniform), one elected lane issues
        const uint32_t a = (uint32_t)__cvta_generic_to_shared(smem), b = a + 16384;
        const uint64_t hi = ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
        const uint64_t ad = hi | (uint64_t)((a >> 4) & 0x3FFF), bd = hi | (uint64_t)((b >> 4) & 0x3FFF);
        const uint32_t fmt = TF32 ? 2u : 1u;           // a / b format: 1 = bf16 (kind::f16), 2 = tf32 (kind::tf32, K = 8 per MMA)
        const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        const uint32_t d1 = tmem + (chains > 1 ? (uint32_t)N : 0u);
        const long long t0 = clock64();
        for (int i = 0; i < n; i += 4) {
            uint32_t pred;
            asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
            if (pred) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (TF32)
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                     "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                                     ::"r"((k & 1) ? d1 : tmem), "l"(ad + 2 * k), "l"(bd + 2 * k), "r"(idesc), "r"(1u) : "memory");
                    else
                        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                     "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                                     ::"r"((k & 1) ? d1 : tmem), "l"(ad + 2 * k), "l"(bd + 2 * k), "r"(idesc), "r"(1u) : "memory");
                }
                if (commit_every && ((i >> 2) % commit_every) == 0)
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(&bar2)) : "memory");
            }
            __syncwarp();
        }
        if (threadIdx.x == 0) {
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(&bar)) : "memory");
            asm volatile("{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}"
                         ::"r"((uint32_t)__cvta_generic_to_shared(&bar)) : "memory");
            cycles[blockIdx.x] = clock64() - t0;
            s_done = 1;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
}

template <typename F>
static float time_ms(F launch)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(e0); launch(); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int clk_khz = 0; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    void* buf; CK(cudaMalloc(&buf, (size_t)sms * 8 * 1024 * 8));
    const int blocks = sms * 8, threads = 256;          // 2048 threads per SM resident
    const double nthr = (double)blocks * threads;
    printf("{\"gpu\": \"%s\", \"sm_count\": %d, \"clock_khz_max\": %d", prop.name, sms, clk_khz);
    float ms;
    ms = time_ms([&] { k_popc<<<blocks, threads>>>((uint32_t*)buf, 1u); });
    printf(", \"popc_xor_add_gops\": %.1f", nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_popc_only<<<blocks, threads>>>((uint32_t*)buf, 1u); });
    printf(", \"popc_gops\": %.1f", nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_f64<0><<<blocks, threads>>>((double*)buf, 0.5); });
    printf(", \"dfma_ginst\": %.1f", nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_f64<1><<<blocks, threads>>>((double*)buf, 0.5); });
    printf(", \"dadd_ginst\": %.1f", nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_f64<2><<<blocks, threads>>>((double*)buf, 0.5); });
    printf(", \"dmul_ginst\": %.1f", nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_f64<3><<<blocks, threads>>>((double*)buf, 0.5); });
    printf(", \"ddiv_rn_gops\": %.1f", nthr * (ITERS / 16) * UNROLL / ms * 1e-6);
    const char* names[5] = {"fmnmx_ginst", "fmnmx3_ginst", "fadd_ginst", "ffma_ginst", "shf_funnel_ginst"};
    ms = time_ms([&] { k_f32<0><<<blocks, threads>>>((float*)buf, 0.5f); }); printf(", \"%s\": %.1f", names[0], nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_f32<1><<<blocks, threads>>>((float*)buf, 0.5f); }); printf(", \"%s\": %.1f", names[1], nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_f32<2><<<blocks, threads>>>((float*)buf, 0.5f); }); printf(", \"%s\": %.1f", names[2], nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_f32<3><<<blocks, threads>>>((float*)buf, 0.5f); }); printf(", \"%s\": %.1f", names[3], nthr * ITERS * UNROLL / ms * 1e-6);
    ms = time_ms([&] { k_f32<4><<<blocks, threads>>>((float*)buf, 0.5f); }); printf(", \"%s\": %.1f", names[4], nthr * ITERS * UNROLL / ms * 1e-6);
    for (int warps : {4, 8, 16}) {
        const int iters = 2000;
        ms = time_ms([&] { k_ldtm<<<sms, warps * 32>>>((uint32_t*)buf, iters); });
        // bytes per SM per clock at the max clock (32 lanes x 32 columns x 4 B per warp-load)
        const double bytes = (double)warps * iters * 4 * 4096;
        printf(", \"ldtm_x32_%dwarps_B_per_clk_per_sm\": %.1f", warps, bytes / (ms * 1e-3 * clk_khz * 1e3));
    }
    cudaFuncSetAttribute(k_umma<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    cudaFuncSetAttribute(k_umma<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int N : {64, 128, 256}) {
        for (int chains : {1, 2}) {
            if (N * chains > 512) continue;
            const int n = 2000;
            long long* cyc = (long long*)buf;
            k_umma<0><<<sms, 64, 64 * 1024>>>(cyc, n, N, chains);
            if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "k_umma failed\n"); break; }
            long long h[148]; cudaMemcpy(h, cyc, sizeof(long long) * (sms < 148 ? sms : 148), cudaMemcpyDeviceToHost);
            printf(", \"umma_m128_n%d_k16_bf16_chains%d_cycles_per_mma\": %.1f", N, chains, (double)h[0] / n);
        }
    }
    for (int N : {128, 256}) {                          // kind::tf32: M128 x N x K8 per MMA -> tf32 tensor peak = 2*128*N*8 / cycles
        const int n = 2000;
        long long* cyc = (long long*)buf;
        k_umma<1><<<sms, 64, 64 * 1024>>>(cyc, n, N, 1, 0);
        if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "k_umma tf32 failed\n"); break; }
        long long h[1]; cudaMemcpy(h, cyc, sizeof(long long), cudaMemcpyDeviceToHost);
        const double cpm = (double)h[0] / n;
        printf(", \"umma_m128_n%d_k8_tf32_cycles_per_mma\": %.1f", N, cpm);
        if (N == 256) printf(", \"tf32_tflops_issue_bound\": %.1f", 2.0 * 128 * N * 8 / cpm * sms * clk_khz * 1e3 / 1e12);
    }
    {   // MMA rate while TMA bulk copies stream into the same SM's shared memory (the ranking kernel's situation: 36 KB of B per 18 MMAs = 2 KB per MMA)
        unsigned char* gsrc; cudaMalloc(&gsrc, (size_t)sms * 16384); cudaMemset(gsrc, 0, (size_t)sms * 16384);
        cudaFuncSetAttribute(k_umma<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
        for (int N : {128, 256}) {
            for (int pace : {0, 32, 64}) {              // 0 = copies back to back (as fast as TMA goes); 32 / 64 = 2 / 4 KB per 64 clocks
                const int n = 4000;
                long long* cyc = (long long*)buf;
                k_umma<0><<<sms, 64, 96 * 1024>>>(cyc, n, N, 1, 0, gsrc, pace);
                if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "k_umma+tma failed\n"); break; }
                long long h[1]; cudaMemcpy(h, cyc, sizeof(long long), cudaMemcpyDeviceToHost);
                printf(", \"umma_m128_n%d_with_tma_pace%d_cycles_per_mma\": %.1f", N, pace, (double)h[0] / n);
            }
        }
    }
    for (int chains : {1, 2}) {
        for (int ce : {1, 2, 4, 8, 16}) {
            const int n = 4096;
            long long* cyc = (long long*)buf;
            k_umma<0><<<sms, 64, 64 * 1024>>>(cyc, n, 128, chains, ce);
            if (cudaDeviceSynchronize() != cudaSuccess) { fprintf(stderr, "k_umma failed\n"); break; }
            long long h[1]; cudaMemcpy(h, cyc, sizeof(long long), cudaMemcpyDeviceToHost);
            printf(", \"umma_m128_n128_chains%d_commit_every_%d_mma_cycles_per_mma\": %.1f", chains, 4 * ce, (double)h[0] / n);
        }
    }
    printf("}\n");
    return 0;
}
