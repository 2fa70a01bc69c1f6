// ransac5pt.cu -- RANSAC 5-point essential-matrix kernels for sm_100a.
//
// Replaces the hypothesise-and-verify loop of CvModelEstimator3::runRANSAC (poselib/source/five-point-nister/
// modelest.cpp:343-474) with three data-parallel stages over PRE-DRAWN minimal samples:
//   fivept_solve_kernel    one thread = one minimal sample -> up to 10 essential matrices (fivept_device.cuh;
//                          reference: CvEMEstimator::run5Point, five-point.cpp:366-471)
//   compact_*              exclusive scan of the per-sample model counts + gather into a dense model list
//                          (deterministic order: sample-major, solver root order)
//   sampson_score_kernel   one thread = one model, the CTA streams all correspondences through shared memory
//                          (1-D TMA bulk copies, 2-stage ring, warp-wide broadcast reads) and evaluates the
//                          reference's Sampson error (five-point.cpp:476-503) in ITS operation order with
//                          explicit round-to-nearest fp64 ops (no FMA contraction), casts to float, compares
//                          err <= th*th (modelest.cpp:69-83), counts inliers and accumulates sum(err).
// The sequential selection / tie-break / adaptive-stop rule (modelest.cpp:394-415) is then replayed on the
// host over the (count, errsum) arrays (pose_host.cu), exactly as the reference applies it model by model.
//
// Roofline: fp64 pipe.  One Sampson evaluation = 32 fp64 mul/add + 1 divide + 1 cvt.
#include "mpl_common.cuh"
#include "fivept_device.cuh"

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// corr[i] = (x1, y1, x2, y2): 32-byte records so that one correspondence is two 128-bit shared loads
__global__ void pack_corr_kernel(const double* __restrict__ p1, const double* __restrict__ p2, int n, double4* __restrict__ corr)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) corr[i] = make_double4(p1[2 * i], p1[2 * i + 1], p2[2 * i], p2[2 * i + 1]);
}

template <bool EARLY>
__global__ void __launch_bounds__(64)
fivept_solve_kernel(const double4* __restrict__ corr, int n, const int32_t* __restrict__ samples, int h,
                    double* __restrict__ models /* [h][10][9] */, int32_t* __restrict__ n_models)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= h) return;
    double x1[5], y1[5], x2[5], y2[5];
    bool ok = true;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int id = samples[5 * s + j];
        if (id < 0 || id >= n) { ok = false; x1[j] = y1[j] = x2[j] = y2[j] = 0; continue; }
        const double4 c = corr[id];
        x1[j] = c.x; y1[j] = c.y; x2[j] = c.z; y2[j] = c.w;
    }
    int cnt = 0;
    if (ok) {
        double ee[4][9];
        fivept::null_space_5x9(x1, y1, x2, y2, ee);
        cnt = fivept::solve_from_basis<EARLY>(ee, models + (size_t)s * 90);
    }
    n_models[s] = cnt;
}

// single-block exclusive scan of n_models -> offsets[h+1]: each thread owns a contiguous chunk (local sum, one block
// scan of the 1024 partial sums, then the chunk's running prefix) -- two barriers in total instead of four per 1024
// elements
__global__ void __launch_bounds__(1024) scan_counts_kernel(const int32_t* __restrict__ counts, int h, int32_t* __restrict__ offsets)
{
    __shared__ int warp_sums[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = (h + 1023) / 1024;
    const int start = min(h, tid * per), end = min(h, start + per);
    int sum = 0;
    for (int i = start; i < end; ++i) sum += counts[i];
    int x = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_sums[wid] = x;
    __syncthreads();
    if (wid == 0) {
        int w = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
        warp_sums[lane] = w;
    }
    __syncthreads();
    int running = (wid ? warp_sums[wid - 1] : 0) + x - sum;
    for (int i = start; i < end; ++i) { offsets[i] = running; running += counts[i]; }
    if (tid == 1023) offsets[h] = running;
}

__global__ void gather_models_kernel(const double* __restrict__ models, const int32_t* __restrict__ n_models,
                                     const int32_t* __restrict__ offsets, int h, double* __restrict__ dense)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;      // one thread per (sample, slot, element)
    const int s = t / 90, r = t - s * 90;
    if (s >= h) return;
    const int j = r / 9;
    if (j < n_models[s]) dense[(size_t)(offsets[s] + j) * 9 + (r - j * 9)] = models[(size_t)s * 90 + r];
}

#ifndef MPL_SCORE_THREADS
#define MPL_SCORE_THREADS 128
#endif
#ifndef MPL_SCORE_CHUNK
#define MPL_SCORE_CHUNK 256
#endif
constexpr int kScoreThreads = MPL_SCORE_THREADS;
constexpr int kChunk = MPL_SCORE_CHUNK;      // correspondences per TMA stage (32 B each)
#ifndef MPL_SCORE_STAGES
#define MPL_SCORE_STAGES 2
#endif
constexpr int kScoreStages = MPL_SCORE_STAGES;

// One Sampson evaluation in the reference's operation order (five-point.cpp:490-503).
__device__ __forceinline__ float sampson_err(const double (&E)[9], const double4 c)
{
    const double x1 = c.x, y1 = c.y, x2 = c.z, y2 = c.w;
    const double Ex0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], x1), __dmul_rn(E[1], y1)), E[2]);
    const double Ex1 = __dadd_rn(__dadd_rn(__dmul_rn(E[3], x1), __dmul_rn(E[4], y1)), E[5]);
    const double Ex2 = __dadd_rn(__dadd_rn(__dmul_rn(E[6], x1), __dmul_rn(E[7], y1)), E[8]);
    const double Et0 = __dadd_rn(__dadd_rn(__dmul_rn(E[0], x2), __dmul_rn(E[3], y2)), E[6]);
    const double Et1 = __dadd_rn(__dadd_rn(__dmul_rn(E[1], x2), __dmul_rn(E[4], y2)), E[7]);
    const double r = __dadd_rn(__dadd_rn(__dmul_rn(x2, Ex0), __dmul_rn(y2, Ex1)), Ex2);
    const double den = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(Ex0, Ex0), __dmul_rn(Ex1, Ex1)), __dmul_rn(Et0, Et0)), __dmul_rn(Et1, Et1));
    return __double2float_rn(__ddiv_rn(__dmul_rn(r, r), den));
}

// Few models (one image pair: ~5 000; a refit: <= 10) leave most SMs idle and every thread with a long latency-bound
// loop, so the correspondences are then split over the S CTAs of a thread-block cluster (S = 2, 4, 8): CTA rank r scores
// the models of its block against correspondences [n r / S, n (r+1) / S), and rank 0 adds the partial (count, sum) pairs
// in rank order through distributed shared memory -- deterministic, no scratch, no atomics.  S = 1 (the C3 regime, one
// CTA per 128 models) is the plain sequential sum.
__global__ void __launch_bounds__(kScoreThreads)
sampson_score_kernel(const double* __restrict__ models, int m, const double4* __restrict__ corr, int n, double th2,
                     int32_t* __restrict__ counts, double* __restrict__ errsums, int S)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double4* tile = reinterpret_cast<double4*>(smem_raw);                 // [kScoreStages][kChunk]
    __shared__ __align__(8) uint64_t full_bar[kScoreStages];
    __shared__ double part_sum[kScoreThreads];
    __shared__ int part_good[kScoreThreads];
    const int tid = threadIdx.x;
    const int rank = S > 1 ? (int)(blockIdx.x % S) : 0;
    const int mi = (int)(blockIdx.x / S) * kScoreThreads + tid;
    const int i_lo = (int)((long long)n * rank / S), i_hi = (int)((long long)n * (rank + 1) / S);
    const int n_loc = i_hi - i_lo;
    const int n_chunks = (n_loc + kChunk - 1) / kChunk;
    if (tid == 0) {
        for (int s = 0; s < kScoreStages; ++s) mbar_init(&full_bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int ci) {
        const int s = ci % kScoreStages;
        const int cnt = min(kChunk, n_loc - ci * kChunk);
        mbar_expect_tx(&full_bar[s], (uint32_t)cnt * 32u);
        tma_bulk_g2s(tile + s * kChunk, corr + i_lo + (size_t)ci * kChunk, (uint32_t)cnt * 32u, &full_bar[s]);
    };
    if (tid == 0) for (int i = 0; i < kScoreStages && i < n_chunks; ++i) issue(i);

    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = mi < m ? models[(size_t)mi * 9 + k] : 0.0;
    int good = 0;
    double esum = 0.0;
    for (int ci = 0; ci < n_chunks; ++ci) {
        const int s = ci % kScoreStages;
        mbar_wait(&full_bar[s], (uint32_t)((ci / kScoreStages) & 1));
        const double4* t = tile + s * kChunk;
        const int cnt = min(kChunk, n_loc - ci * kChunk);
#pragma unroll 4
        for (int i = 0; i < cnt; ++i) {
            const float e = sampson_err(E, t[i]);                        // two LDS.128, warp-wide broadcast
            good += ((double)e <= th2) ? 1 : 0;
            esum = __dadd_rn(esum, (double)e);
        }
        __syncthreads();
        if (tid == 0 && ci + kScoreStages < n_chunks) issue(ci + kScoreStages);
    }
    if (S > 1) {
        part_good[tid] = good; part_sum[tid] = esum;
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (rank == 0) {
            for (int r = 1; r < S; ++r) {
                uint32_t ag, as;
                asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ag) : "r"(smem_u32(&part_good[tid])), "r"(r));
                asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(as) : "r"(smem_u32(&part_sum[tid])), "r"(r));
                int g2; double s2;
                asm volatile("ld.shared::cluster.s32 %0, [%1];" : "=r"(g2) : "r"(ag) : "memory");
                asm volatile("ld.shared::cluster.f64 %0, [%1];" : "=d"(s2) : "r"(as) : "memory");
                good += g2; esum = __dadd_rn(esum, s2);
            }
        }
        // the partials of ranks > 0 must stay alive until rank 0 has read them
        asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        if (rank != 0) return;
    }
    if (mi < m) { counts[mi] = good; if (errsums) errsums[mi] = esum; }
}

// ---- LMedS (runLMeDS, modelest.cpp:483-564) ------------------------------------------------------------
// The reference sorts the n float errors of a model through their bit patterns as ints (icvSortDistances over
// (int*)err.data, modelest.cpp:477-480, 539) and takes the middle one / the float sum of the two middle ones times 0.5.
// Here one CTA per model finds the order statistics lo = (n-1)/2 and hi = n/2 of the same keys by a 4-pass radix select
// (8 bits per pass, per-warp histograms in shared memory); the errors are recomputed in every pass from the
// L2-resident correspondences (44 fp64 issue slots each) instead of being stored.  A NaN error (0/0) is the x86
// default NaN 0xFFC00000 in the reference, a negative int that sorts first; the key mapping reproduces that.
constexpr int kMedThreads = 256;
__device__ __forceinline__ uint32_t lmeds_key(float e)
{
    const uint32_t b = (e != e) ? 0xFFC00000u : __float_as_uint(e);
    return b ^ 0x80000000u;                                               // signed int order -> unsigned order
}
__device__ __forceinline__ float lmeds_unkey(uint32_t k) { return __uint_as_float(k ^ 0x80000000u); }

__global__ void __launch_bounds__(kMedThreads)
lmeds_median_kernel(const double* __restrict__ models, int m, const double4* __restrict__ corr, int n, double* __restrict__ medians)
{
    __shared__ unsigned hist[kMedThreads / 32][256];
    __shared__ unsigned s_prefix, s_rank, s_eq;
    __shared__ unsigned s_min[kMedThreads / 32];
    const int mi = blockIdx.x, tid = threadIdx.x, wid = tid >> 5, lane = tid & 31;
    if (mi >= m) return;
    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = models[(size_t)mi * 9 + k];
    const unsigned lo = (unsigned)(n - 1) / 2, hi = (unsigned)n / 2;
    if (tid == 0) { s_prefix = 0; s_rank = lo; }
    for (int pass = 0; pass < 4; ++pass) {
        const int shift = 24 - 8 * pass;
        for (int i = tid; i < (kMedThreads / 32) * 256; i += kMedThreads) (&hist[0][0])[i] = 0;
        __syncthreads();
        const unsigned prefix = s_prefix;
        for (int i = tid; i < n; i += kMedThreads) {
            const uint32_t k = lmeds_key(sampson_err(E, corr[i]));
            if (pass == 0 || (k >> (shift + 8)) == prefix) atomicAdd(&hist[wid][(k >> shift) & 255u], 1u);
        }
        __syncthreads();
        unsigned c = 0;                                                   // thread = bin
#pragma unroll
        for (int w = 0; w < kMedThreads / 32; ++w) c += hist[w][tid];
        // exclusive scan over the 256 bins: warp scan + warp totals
        unsigned incl = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const unsigned v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        __syncthreads();                                                  // all reads of hist[][] done before it is reused
        if (lane == 31) hist[0][wid] = incl;
        __syncthreads();
        unsigned base = 0;
        for (int w = 0; w < wid; ++w) base += hist[0][w];
        const unsigned excl = base + incl - c, r = s_rank;
        __syncthreads();
        if (c > 0 && r >= excl && r < excl + c) { s_prefix = (prefix << 8) | (unsigned)tid; s_rank = r - excl; s_eq = c; }
        __syncthreads();
    }
    const uint32_t k_lo = s_prefix;
    uint32_t k_hi = k_lo;
    // elements <= k_lo: (lo - s_rank) strictly smaller + s_eq equal ones; rank hi lies beyond them -> the smallest larger key
    if (hi != lo && (lo - s_rank) + s_eq <= hi) {
        unsigned mn = 0xFFFFFFFFu;
        for (int i = tid; i < n; i += kMedThreads) {
            const uint32_t k = lmeds_key(sampson_err(E, corr[i]));
            if (k > k_lo) mn = min(mn, k);
        }
        mn = __reduce_min_sync(0xffffffffu, mn);
        if (lane == 0) s_min[wid] = mn;
        __syncthreads();
        if (tid == 0) { for (int w = 1; w < kMedThreads / 32; ++w) mn = min(mn, s_min[w]); s_min[0] = mn; }
        __syncthreads();
        k_hi = s_min[0];
    }
    if (tid == 0) {
        const float f_lo = lmeds_unkey(k_lo), f_hi = lmeds_unkey(k_hi);
        medians[mi] = (n & 1) ? (double)f_lo : __dmul_rn((double)__fadd_rn(f_lo, f_hi), 0.5);     // modelest.cpp:541-542
    }
}

// findInliers (modelest.cpp:69-83) of one model: mask + count
__global__ void __launch_bounds__(256)
inlier_mask_kernel(const double* __restrict__ model, const double4* __restrict__ corr, int n, double th2, uint8_t* __restrict__ mask,
                   int* __restrict__ count)
{
    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = model[k];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = i < n && (double)sampson_err(E, corr[i]) <= th2;
    if (i < n) mask[i] = in ? 1 : 0;
    const unsigned b = __ballot_sync(0xffffffffu, in);
    if ((threadIdx.x & 31) == 0 && b) atomicAdd(count, __popc(b));
}

}  // namespace

// ---- launch wrappers (declared in mpl_ransac.h) ------------------------------------------------------
#include "mpl_ransac.h"

int mpl_launch_pack_corr(const double* p1, const double* p2, int n, void* corr, cudaStream_t st)
{
    if (n <= 0) return MPL_OK;
    pack_corr_kernel<<<(n + 255) / 256, 256, 0, st>>>(p1, p2, n, (double4*)corr);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_fivept_solve(const void* corr, int n, const int32_t* samples, int h, double* models, int32_t* n_models, cudaStream_t st)
{
    if (h <= 0) return MPL_OK;
    MplProfScope prof(MPL_PROF_SOLVE, st);
    if (mpl_solver_early_exit()) fivept_solve_kernel<true><<<(h + 63) / 64, 64, 0, st>>>((const double4*)corr, n, samples, h, models, n_models);
    else                         fivept_solve_kernel<false><<<(h + 63) / 64, 64, 0, st>>>((const double4*)corr, n, samples, h, models, n_models);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_compact_models(const double* models, const int32_t* n_models, int h, int32_t* offsets, double* dense, cudaStream_t st)
{
    if (h <= 0) return MPL_OK;
    scan_counts_kernel<<<1, 1024, 0, st>>>(n_models, h, offsets);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    const long long total = (long long)h * 90;
    gather_models_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(models, n_models, offsets, h, dense);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_sampson_score(const double* models, int m, const void* corr, int n, double th, int32_t* counts, double* errsums, cudaStream_t st)
{
    if (m <= 0) return MPL_OK;
    const size_t smem = (size_t)kScoreStages * kChunk * 32;
    // per launch: the attribute is per device (mpl_init may re-target; per-thread contexts call concurrently)
    MPL_CUDA_CHECK(cudaFuncSetAttribute(sampson_score_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int blocks = (m + kScoreThreads - 1) / kScoreThreads;
    int S = 1;                                                    // cluster size: fill ~2 CTAs per SM, >= 64 correspondences per rank
    while (S < 8 && blocks * S < 2 * 148 && n / (2 * S) >= 64) S *= 2;
    MplProfScope prof(MPL_PROF_SCORE, st);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(blocks * S)); cfg.blockDim = dim3(kScoreThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)S; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    MPL_CUDA_CHECK(cudaLaunchKernelEx(&cfg, sampson_score_kernel, models, m, (const double4*)corr, n, th * th, counts, errsums, S));
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_lmeds_medians(const double* models, int m, const void* corr, int n, double* medians, cudaStream_t st)
{
    if (m <= 0) return MPL_OK;
    MplProfScope prof(MPL_PROF_SCORE, st);
    lmeds_median_kernel<<<m, kMedThreads, 0, st>>>(models, m, (const double4*)corr, n, medians);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_inlier_mask(const double* model, const void* corr, int n, double th, uint8_t* mask, int* count, cudaStream_t st)
{
    MPL_CUDA_CHECK(cudaMemsetAsync(count, 0, sizeof(int), st));
    inlier_mask_kernel<<<(n + 255) / 256, 256, 0, st>>>(model, (const double4*)corr, n, th * th, mask, count);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}
