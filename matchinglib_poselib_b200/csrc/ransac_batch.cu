// ransac_batch.cu -- RANSAC 5-point for a BATCH of image pairs with no host round trip: chunked solve / score, the
// reference's sequential selection rule replayed on the device, and the least-squares refit on the device.
//
// Reference (paths relative to matchinglib_poselib/source/poselib/source/five-point-nister):
//   runRANSAC loop, take / tie-break rule          modelest.cpp:377-416
//   cvRANSACUpdateNumIters1                        modelest.cpp:86-109
//   least-squares refit on all inliers             modelest.cpp:420-464  (runKernel on n x 9 rows: five-point.cpp:375-388)
//   final mask / result                            modelest.cpp:466-473
//
// Layout: one MplPairState per pair; correspondences [pair][corr_stride] 32-byte records; minimal samples [pair][h][5];
// model slots [pair][h][10][9] fp64 (slot j of sample s valid iff j < n_models[pair][s]) with (count, errsum) per slot.
//
// Kernels (all take the whole batch; a pair that is finished -- state.niters <= first sample of the chunk -- costs nothing):
//   solve_chunk_kernel          thread = (pair, sample): fivept::solve from the minimal sample (fivept_device.cuh)
//   score_slots_kernel          CTA = (pair, group of 16 samples, rank in a cluster that splits the correspondences): compacts
//                               the valid slots of its samples in shared memory, one thread per model, correspondences
//                               streamed through a TMA ring exactly like sampson_score_kernel (same sampson_err)
//   replay_chunk_kernel         one thread per pair walks its slots in (sample, root) order applying modelest.cpp:394-415
//   refit_basis_kernel          CTA per pair: inlier mask of the selected model; Givens QR of the inliers' n x 9 constraint
//                               rows (every thread folds its rows into a 9 x 9 triangle in registers, binary tree over the CTA),
//                               then a one-sided Jacobi SVD of the 9 x 9 triangle by one warp -> the four right singular
//                               vectors of the smallest singular values.  Q = (orthogonal) R, so R has Q's singular values and
//                               right singular vectors: nothing is squared (round 1 formed Q^T Q and eigen-solved it).
//   refit_solve_kernel          thread per pair: fivept::solve_from_basis
//   refit_select_kernel         CTA per pair: scores the <= 10 refit models, applies modelest.cpp:446-462, final mask
#include "mpl_common.cuh"
#include "fivept_device.cuh"
#include "mpl_ransac.h"

#include <float.h>

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

// One Sampson evaluation in the reference's operation order (five-point.cpp:490-503); identical to ransac5pt.cu's.
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

// cvRANSACUpdateNumIters1 (modelest.cpp:86-109)
__device__ int update_num_iters_dev(double p, double ep, int model_points, int max_iters)
{
    p = fmax(p, 0.); p = fmin(p, 1.);
    ep = fmax(ep, 0.); ep = fmin(ep, 1.);
    double num = fmax(1. - p, DBL_MIN);
    double denom = 1. - pow(1. - ep, (double)model_points);
    if (denom < DBL_MIN) return 0;
    num = log(num); denom = log(denom);
    return denom >= 0 || -num >= (double)max_iters * (-denom) ? max_iters : (int)round(num / denom);
}

// ---- minimal samples ------------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t splitmix64(uint64_t& s)
{
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// thread = (pair, sample): getSubset (modelest.cpp:567-610) = five distinct indices drawn from the pair's own counter-based
// stream (seed, sample id); checkSubset never rejects in the reference (pose_host.cu has the argument).
__global__ void draw_samples_batch_kernel(const MplPairState* __restrict__ st, int n_pairs, int h, int32_t* __restrict__ samples)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int p = (int)(t / h), s = (int)(t - (long long)p * h);
    if (p >= n_pairs) return;
    const int n = st[p].n;
    int idx[5] = {0, 1, 2, 3, 4};
    if (n >= 5) {
        uint64_t rs = st[p].seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(s + 1));
        for (int i = 0; i < 5;) {
            const int c = (int)(splitmix64(rs) % (uint64_t)n);
            int j = 0;
            for (; j < i; ++j) if (c == idx[j]) break;
            if (j < i) continue;
            idx[i++] = c;
        }
    }
    int32_t* o = samples + ((size_t)p * h + s) * 5;
    for (int j = 0; j < 5; ++j) o[j] = idx[j];
}

// ---- solve ----------------------------------------------------------------------------------------------------
template <bool EARLY>
__global__ void __launch_bounds__(64)
solve_chunk_kernel(const double4* __restrict__ corr, size_t corr_stride, const MplPairState* __restrict__ st, int n_pairs,
                   const int32_t* __restrict__ samples, int h, int s0, int c, double* __restrict__ models, int32_t* __restrict__ n_models)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int p = (int)(t / c), s = s0 + (int)(t - (long long)p * c);
    if (p >= n_pairs || s >= h) return;
    const int n = st[p].n;
    if (n < 5 || s >= st[p].niters) return;              // the loop of modelest.cpp:377 can never reach this sample
    const int32_t* sm = samples + ((size_t)p * h + s) * 5;
    const double4* cp = corr + (size_t)p * corr_stride;
    double x1[5], y1[5], x2[5], y2[5];
    bool ok = true;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const int id = sm[j];
        if (id < 0 || id >= n) { ok = false; x1[j] = y1[j] = x2[j] = y2[j] = 0; continue; }
        const double4 v = cp[id];
        x1[j] = v.x; y1[j] = v.y; x2[j] = v.z; y2[j] = v.w;
    }
    int cnt = 0;
    if (ok) {
        double ee[4][9];
        fivept::null_space_5x9(x1, y1, x2, y2, ee);
        cnt = fivept::solve_from_basis<EARLY>(ee, models + ((size_t)p * h + s) * 90);
    }
    n_models[(size_t)p * h + s] = cnt;
}

// ---- score ----------------------------------------------------------------------------------------------------
constexpr int kScoreThreads = 128;
#ifndef MPL_SLOTS_CHUNK
#define MPL_SLOTS_CHUNK 512
#endif
constexpr int kChunk = MPL_SLOTS_CHUNK;      // correspondences per TMA stage (32 B each)
constexpr int kScoreStages = 2;
constexpr int kGroupSamples = 16;            // samples per CTA: <= 160 slots, ~77 valid on average (4.8 real roots per sample)

__global__ void __launch_bounds__(kScoreThreads)
score_slots_kernel(const double4* __restrict__ corr, size_t corr_stride, const MplPairState* __restrict__ st,
                   const double* __restrict__ models, const int32_t* __restrict__ n_models, int h, int s0, int c, double th2,
                   int32_t* __restrict__ counts, double* __restrict__ errsums, int S)
{
    __shared__ __align__(128) double4 tile[kScoreStages][kChunk];
    __shared__ __align__(8) uint64_t full_bar[kScoreStages];
    __shared__ double part_sum[kScoreThreads];
    __shared__ int part_good[kScoreThreads];
    __shared__ short slot_list[kGroupSamples * 10];
    __shared__ int s_nvalid;
    const int tid = threadIdx.x;
    const int p = blockIdx.y;
    const int rank = S > 1 ? (int)(blockIdx.x % S) : 0;
    const int sg = s0 + (int)(blockIdx.x / S) * kGroupSamples;             // first sample of this CTA's group
    const int n = st[p].n;
    const int niters = st[p].niters;
    // uniform over the cluster: every CTA of a cluster has the same (p, sg)
    if (n < 5 || sg >= niters || sg >= h || sg >= s0 + c) return;
    const int s_end = min(min(sg + kGroupSamples, s0 + c), min(h, niters));
    if (tid == 0) {
        int nv = 0;
        for (int s = sg; s < s_end; ++s) {
            const int nm = n_models[(size_t)p * h + s];
            for (int j = 0; j < nm; ++j) slot_list[nv++] = (short)((s - sg) * 10 + j);
        }
        s_nvalid = nv;
        for (int s = 0; s < kScoreStages; ++s) mbar_init(&full_bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const int nvalid = s_nvalid;
    const double4* cp = corr + (size_t)p * corr_stride;
    const int i_lo = (int)((long long)n * rank / S), i_hi = (int)((long long)n * (rank + 1) / S);
    const int n_loc = i_hi - i_lo;
    const int n_chunks = (n_loc + kChunk - 1) / kChunk;
    const int passes = (nvalid + kScoreThreads - 1) / kScoreThreads;       // 1, rarely 2 (uniform over the cluster)
    int issued = 0, consumed = 0;
    auto issue = [&](int ci) {
        const int s = issued % kScoreStages;
        const int cnt = min(kChunk, n_loc - ci * kChunk);
        mbar_expect_tx(&full_bar[s], (uint32_t)cnt * 32u);
        tma_bulk_g2s(&tile[s][0], cp + i_lo + (size_t)ci * kChunk, (uint32_t)cnt * 32u, &full_bar[s]);
        ++issued;
    };
    for (int pass = 0; pass < passes; ++pass) {
        const int e = pass * kScoreThreads + tid;
        const bool live = e < nvalid;
        const int slot = live ? slot_list[e] : 0;
        const size_t gslot = ((size_t)p * h + sg) * 10 + slot;
        double E[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) E[k] = live ? models[gslot * 9 + k] : 0.0;
        if (tid == 0) for (int i = 0; i < kScoreStages && i < n_chunks; ++i) issue(i);
        int good = 0;
        double esum = 0.0;
        for (int ci = 0; ci < n_chunks; ++ci) {
            const int s = consumed % kScoreStages;
            mbar_wait(&full_bar[s], (uint32_t)((consumed / kScoreStages) & 1));
            ++consumed;
            const double4* t = &tile[s][0];
            const int cnt = min(kChunk, n_loc - ci * kChunk);
#pragma unroll 4
            for (int i = 0; i < cnt; ++i) {
                const float er = sampson_err(E, t[i]);
                good += ((double)er <= th2) ? 1 : 0;
                esum = __dadd_rn(esum, (double)er);
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
            asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
        }
        if (rank == 0 && live) { counts[gslot] = good; errsums[gslot] = esum; }
    }
}

// ---- selection replay (modelest.cpp:377-416) over one chunk of samples ------------------------------------
__global__ void replay_chunk_kernel(MplPairState* __restrict__ st, int n_pairs, const double* __restrict__ models,
                                    const int32_t* __restrict__ n_models, const int32_t* __restrict__ counts,
                                    const double* __restrict__ errsums, int h, int s0, int c, double conf)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    MplPairState S = st[p];
    if (S.n < 5) return;
    int it = s0;
    const int s_hi = min(s0 + c, h);
    if (it >= S.niters) return;                                   // already past the adaptive stop
    for (; it < s_hi && it < S.niters; ++it) {
        const int nm = n_models[(size_t)p * h + it];
        for (int j = 0; j < nm; ++j) {
            const size_t g = ((size_t)p * h + it) * 10 + j;
            const int good = counts[g];
            if (good > max(S.best_count, 4)) {
                S.best_slot = it * 10 + j; S.best_count = good;
                S.niters = update_num_iters_dev(conf, (double)(S.n - good) / S.n, 5, S.niters);
                S.errminsum = errsums[g];
            } else if (good == max(S.best_count, 5) && S.errminsum < DBL_MAX && S.errminsum > errsums[g]) {
                S.best_slot = it * 10 + j; S.errminsum = errsums[g];
            }
        }
    }
    S.iters_run = it;
    if (S.best_slot >= 0) {
        const double* m = models + ((size_t)p * h * 10 + S.best_slot) * 9;
#pragma unroll
        for (int k = 0; k < 9; ++k) S.E[k] = m[k];
    }
    st[p] = S;
}

// ---- refit --------------------------------------------------------------------------------------------------
constexpr int kRefitThreads = 128;
__host__ __device__ constexpr int tri(int k, int j) { return k * 9 - (k * (k - 1)) / 2 + (j - k); }   // packed upper triangle, j >= k

// fold one row (zeros before column K0) into the triangle R by Givens rotations
template <int K0>
__device__ __forceinline__ void qr_add_row(double (&R)[45], double (&r)[9])
{
#pragma unroll
    for (int k = K0; k < 9; ++k) {
        const double b = r[k];
        if (b != 0.0) {
            const double a = R[tri(k, k)];
            const double rho = sqrt(fma(a, a, b * b));
            const double cs = a / rho, sn = b / rho;
            R[tri(k, k)] = rho;
#pragma unroll
            for (int j = k + 1; j < 9; ++j) {
                const double t1 = R[tri(k, j)], t2 = r[j];
                R[tri(k, j)] = fma(cs, t1, sn * t2);
                r[j] = fma(cs, t2, -(sn * t1));
            }
        }
    }
}
template <int K>
__device__ __forceinline__ void merge_rows(double (&R)[45], const double* __restrict__ other)
{
    if constexpr (K < 9) {
        double r[9];
#pragma unroll
        for (int j = 0; j < 9; ++j) r[j] = j >= K ? other[tri(K, j)] : 0.0;
        qr_add_row<K>(R, r);
        merge_rows<K + 1>(R, other);
    }
}

// sum over the whole warp, identical in every lane (the Jacobi control flow below must stay warp-uniform)
__device__ __forceinline__ double warp_sum16(double v)
{
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// mask[i] = Sampson(E_best, corr i) <= th^2 for all i; basis[pair][4][9] = the refit's null-space basis (refine only)
__global__ void __launch_bounds__(kRefitThreads)
refit_basis_kernel(const double4* __restrict__ corr, size_t corr_stride, const MplPairState* __restrict__ st, double th2,
                   int refine, uint8_t* __restrict__ mask, size_t mask_stride, double* __restrict__ basis)
{
    __shared__ double sR[kRefitThreads][45];
    const int p = blockIdx.x, tid = threadIdx.x;
    const MplPairState& S = st[p];
    const int n = S.n;
    if (n < 5 || S.best_slot < 0) return;
    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = S.E[k];
    const double4* cp = corr + (size_t)p * corr_stride;
    uint8_t* mp = mask + (size_t)p * mask_stride;
    double R[45];
#pragma unroll
    for (int k = 0; k < 45; ++k) R[k] = 0.0;
    for (int i = tid; i < n; i += kRefitThreads) {
        const double4 c = cp[i];
        const bool in = (double)sampson_err(E, c) <= th2;
        mp[i] = in ? 1 : 0;
        if (in && refine) {
            double r[9] = {c.x * c.z, c.y * c.z, c.z, c.x * c.w, c.y * c.w, c.w, c.x, c.y, 1.0};     // five-point.cpp:375-383
            qr_add_row<0>(R, r);
        }
    }
    if (!refine) return;
#pragma unroll
    for (int k = 0; k < 45; ++k) sR[tid][k] = R[k];
    __syncthreads();
    for (int stride = kRefitThreads / 2; stride >= 1; stride >>= 1) {
        if (tid < stride) {
            merge_rows<0>(R, sR[tid + stride]);
#pragma unroll
            for (int k = 0; k < 45; ++k) sR[tid][k] = R[k];
        }
        __syncthreads();
    }
    if (tid >= 32) return;
    // one-sided (Hestenes) Jacobi SVD of the 9 x 9 triangle by one warp: lane i holds component i of every column u_j of
    // R and of every right singular vector v_j
    const int lane = tid;
    double U[9], V[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) {
        U[j] = (lane <= j) ? sR[0][tri(lane, j)] : 0.0;                  // R[i][j] = 0 below the diagonal; lanes >= 9 hold zeros
        V[j] = (lane == j) ? 1.0 : 0.0;
    }
    for (int sweep = 0; sweep < 60; ++sweep) {
        bool rotated = false;
#pragma unroll
        for (int pp = 0; pp < 8; ++pp)
#pragma unroll
            for (int q = pp + 1; q < 9; ++q) {
                const double a = warp_sum16(U[pp] * U[pp]), b = warp_sum16(U[q] * U[q]), c = warp_sum16(U[pp] * U[q]);
                if (fabs(c) <= 1e-300 || fabs(c) <= 2.220446049250313e-16 * sqrt(a * b)) continue;
                rotated = true;
                const double zeta = (b - a) / (2.0 * c);
                const double tt = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
                const double up = U[pp], uq = U[q], vp = V[pp], vq = V[q];
                U[pp] = cs * up - sn * uq; U[q] = sn * up + cs * uq;
                V[pp] = cs * vp - sn * vq; V[q] = sn * vp + cs * vq;
            }
        if (!rotated) break;
    }
    double w[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) w[j] = warp_sum16(U[j] * U[j]);
    // descending singular values (cv::SVD's order); Vt rows 5..8 span the null space used by run5Point (five-point.cpp:388)
#pragma unroll
    for (int j = 0; j < 9; ++j) {
        int rank = 0;
#pragma unroll
        for (int k = 0; k < 9; ++k) rank += (w[k] > w[j] || (w[k] == w[j] && k < j)) ? 1 : 0;
        if (rank >= 5 && lane < 9) basis[(size_t)p * 36 + (rank - 5) * 9 + lane] = V[j];
    }
}

__global__ void refit_solve_kernel(const MplPairState* __restrict__ st, int n_pairs, const double* __restrict__ basis,
                                   double* __restrict__ rmodels /* [pair][90] */, int32_t* __restrict__ rnm)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n_pairs) return;
    if (st[p].n < 5 || st[p].best_slot < 0) { rnm[p] = -1; return; }
    double ee[4][9];
    for (int b = 0; b < 4; ++b) for (int k = 0; k < 9; ++k) ee[b][k] = basis[(size_t)p * 36 + 9 * b + k];
    rnm[p] = fivept::solve_from_basis(ee, rmodels + (size_t)p * 90);
}

// scores the refit models of one pair, applies modelest.cpp:446-462, rewrites the mask if a refit model is taken and
// publishes the pair's final (n_inl, E).  refine == 0: only the final bookkeeping.
constexpr int kSelThreads = 256;
__global__ void __launch_bounds__(kSelThreads)
refit_select_kernel(const double4* __restrict__ corr, size_t corr_stride, MplPairState* __restrict__ st, double th2, int refine,
                    const double* __restrict__ rmodels, const int32_t* __restrict__ rnm, uint8_t* __restrict__ mask, size_t mask_stride)
{
    __shared__ int s_good[kSelThreads / 32][10];
    __shared__ double s_sum[kSelThreads / 32][10];
    __shared__ int s_pick;
    const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    MplPairState& S = st[p];
    const int n = S.n;
    if (n < 5 || S.best_slot < 0) { if (tid == 0) { S.n_inl = 0; S.refit_nm = 0; } return; }
    if (!refine) { if (tid == 0) { S.n_inl = S.best_count; S.refit_nm = 0; } return; }
    const int nm = min(rnm[p], 10);
    if (nm <= 0) {                       // modelest.cpp:443-445: runKernel on the inliers gave nothing -> runRANSAC returns false
        if (tid == 0) { S.n_inl = 0; S.refit_nm = 0; }
        return;
    }
    const double4* cp = corr + (size_t)p * corr_stride;
    const double* rm = rmodels + (size_t)p * 90;
    int good[10]; double esum[10];
#pragma unroll
    for (int m = 0; m < 10; ++m) { good[m] = 0; esum[m] = 0.0; }
    for (int i = tid; i < n; i += kSelThreads) {
        const double4 c = cp[i];
#pragma unroll
        for (int m = 0; m < 10; ++m) {
            if (m < nm) {
                double E[9];
#pragma unroll
                for (int k = 0; k < 9; ++k) E[k] = rm[9 * m + k];
                const float er = sampson_err(E, c);
                good[m] += ((double)er <= th2) ? 1 : 0;
                esum[m] += (double)er;
            }
        }
    }
#pragma unroll
    for (int m = 0; m < 10; ++m) {
        int g = good[m]; double e = esum[m];
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) { g += __shfl_xor_sync(0xffffffffu, g, o); e += __shfl_xor_sync(0xffffffffu, e, o); }
        if (lane == 0) { s_good[wid][m] = g; s_sum[wid][m] = e; }
    }
    __syncthreads();
    if (tid == 0) {
        int best_count = S.best_count; double errminsum = S.errminsum; int pick = -1;
        for (int m = 0; m < nm; ++m) {
            int g = 0; double e = 0;
            for (int w = 0; w < kSelThreads / 32; ++w) { g += s_good[w][m]; e += s_sum[w][m]; }
            if (g > max(best_count, 4)) { pick = m; best_count = g; errminsum = e; }
            else if (g == best_count && errminsum < DBL_MAX && errminsum > e) { pick = m; errminsum = e; }
        }
        s_pick = pick;
        S.n_inl = best_count; S.best_count = best_count; S.errminsum = errminsum; S.refit_nm = nm;
        if (pick >= 0) for (int k = 0; k < 9; ++k) S.E[k] = rm[9 * pick + k];
    }
    __syncthreads();
    const int pick = s_pick;
    if (pick < 0) return;
    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = rm[9 * pick + k];
    uint8_t* mp = mask + (size_t)p * mask_stride;
    for (int i = tid; i < n; i += kSelThreads) mp[i] = ((double)sampson_err(E, cp[i]) <= th2) ? 1 : 0;
}

}  // namespace

// ---- launch wrappers (mpl_ransac.h) ---------------------------------------------------------------------------
int mpl_launch_draw_samples_batch(const MplPairState* st, int n_pairs, int h, int32_t* samples, cudaStream_t s)
{
    if (n_pairs <= 0 || h <= 0) return MPL_OK;
    const long long total = (long long)n_pairs * h;
    draw_samples_batch_kernel<<<(unsigned)((total + 127) / 128), 128, 0, s>>>(st, n_pairs, h, samples);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_solve_chunk(const void* corr, size_t corr_stride, const MplPairState* st, int n_pairs, const int32_t* samples, int h,
                           int s0, int c, double* models, int32_t* n_models, cudaStream_t s)
{
    if (n_pairs <= 0 || c <= 0) return MPL_OK;
    MplProfScope prof(MPL_PROF_SOLVE, s);
    const long long total = (long long)n_pairs * c;
    if (mpl_solver_early_exit()) solve_chunk_kernel<true><<<(unsigned)((total + 63) / 64), 64, 0, s>>>((const double4*)corr, corr_stride, st, n_pairs, samples, h, s0, c, models, n_models);
    else                         solve_chunk_kernel<false><<<(unsigned)((total + 63) / 64), 64, 0, s>>>((const double4*)corr, corr_stride, st, n_pairs, samples, h, s0, c, models, n_models);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_score_slots(const void* corr, size_t corr_stride, const MplPairState* st, int n_pairs, const double* models,
                           const int32_t* n_models, int h, int s0, int c, double th, int n_hint, int32_t* counts, double* errsums, cudaStream_t s)
{
    if (n_pairs <= 0 || c <= 0) return MPL_OK;
    const int groups = (c + kGroupSamples - 1) / kGroupSamples;
    int S = 1;                                                    // cluster size: split the correspondences when the grid is small
    while (S < 8 && (long long)groups * n_pairs * S < 2 * 148 && n_hint / (2 * S) >= 64) S *= 2;
    MplProfScope prof(MPL_PROF_SCORE, s);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(groups * S), (unsigned)n_pairs); cfg.blockDim = dim3(kScoreThreads); cfg.dynamicSmemBytes = 0; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)S; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    MPL_CUDA_CHECK(cudaLaunchKernelEx(&cfg, score_slots_kernel, (const double4*)corr, corr_stride, st, models, n_models, h, s0, c, th * th, counts, errsums, S));
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_replay_chunk(MplPairState* st, int n_pairs, const double* models, const int32_t* n_models, const int32_t* counts,
                            const double* errsums, int h, int s0, int c, double conf, cudaStream_t s)
{
    if (n_pairs <= 0 || c <= 0) return MPL_OK;
    replay_chunk_kernel<<<(n_pairs + 31) / 32, 32, 0, s>>>(st, n_pairs, models, n_models, counts, errsums, h, s0, c, conf);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_refit(const void* corr, size_t corr_stride, MplPairState* st, int n_pairs, double th, int refine, uint8_t* mask,
                     size_t mask_stride, double* basis, double* rmodels, int32_t* rnm, cudaStream_t s)
{
    if (n_pairs <= 0) return MPL_OK;
    refit_basis_kernel<<<n_pairs, kRefitThreads, 0, s>>>((const double4*)corr, corr_stride, st, th * th, refine, mask, mask_stride, basis);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    if (refine) {
        MplProfScope prof(MPL_PROF_SOLVE, s);
        refit_solve_kernel<<<(n_pairs + 31) / 32, 32, 0, s>>>(st, n_pairs, basis, rmodels, rnm);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    }
    refit_select_kernel<<<n_pairs, kSelThreads, 0, s>>>((const double4*)corr, corr_stride, st, th * th, refine, rmodels, rnm, mask, mask_stride);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    return MPL_OK;
}
