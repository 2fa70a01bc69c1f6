// usac.cu -- poselib::estimateEssentialOrPoseUSAC, the DEFAULT RobMethod of poselib-test.
//
// Reference (paths relative to matchinglib_poselib/source/poselib):
//   estimateEssentialOrPoseUSAC                          source/pose_estim.cpp:1737-2262 (outer flow, SPRT / PROSAC initial values)
//   estimateEssentialMatUsac                             source/usac/usac_estimations.cpp:283-735 (the configuration)
//   USAC<ProblemType>::solve and its helpers             include/usac/estimators/USAC.h:345-1230 (Raguram et al., PAMI 2013):
//       uniform / PROSAC sampling :620-644, :716-852; PROSAC stopping :860-910; SPRT design :918-942, history :1186-1200;
//       LOSAC :948-1060; standard / SPRT stopping :1086-1160; storeSolution :1208-1228
//   EssentialMatEstimator                                include/usac/estimators/EssentialMatEstimator.h:
//       initProblem :189-352, generateMinimalSampleModels :384-523, generateRefinedModel (REFINE_WEIGHTS) :540-597,
//       validateSample :1043-1077, validateModel :1085-1104, evaluateModel :1110-1174, findWeights :2366-2389, storeModel :2436
//   FTools (normalizePoints, computeDataMatrix, formCovMat, singulF, computeEpipole, getOriSign)
//                                                        source/usac/utils/FundmatrixFunctions.cpp:7-85, 314-380
//
// USAC is a sequential, adaptive schedule: every hypothesis can change the SPRT thresholds, the stopping count and (through
// local optimisation) the best model.  What is data parallel is (1) turning minimal samples into validated models, (2) the
// Sampson test of every model against every correspondence, and (3) the local optimisation's chain of covariance
// accumulations / evaluations over all correspondences.  So:
//   * the host draws a BATCH of minimal samples with the reference's sampler (glibc rand() restated: the stream of a seed is
//     the reference's), pre-validates them (validateSample), and one launch pair turns the batch into models (5-point solver,
//     oriented-epipolar validateModel) and their inlier BIT MATRIX (one bit per model and correspondence);
//   * the host walks USAC::solve over those bits exactly as USAC.h states it -- the SPRT of evaluateModel is a pure function
//     of a model's bits in evaluation-pool order, of the pool cursor and of the current (delta, epsilon, A);
//   * whenever a model becomes the best one and the local optimisation will draw from the random stream (or PROSAC's stopping
//     length moves), the rest of the batch is DROPPED and the sampler rewound to the state it had after that hypothesis'
//     sample (so the stream is consumed in the reference's order: sample, LO draws, next sample); the local optimisation
//     runs as ONE launch (a CTA per inner-RANSAC repetition doing the whole fit / evaluate / re-weight chain, SPRT walk
//     included) and the next batch is sized by the new stopping count.
//
// Configuration served: estimator NISTER or STEWENIUS (both are answered by the device 5-point solver of fivept_device.cuh:
// the same ten-root polynomial system; opengv's two numerical routes to it are not restated), inner refinement
// REFINE_WEIGHTS (8-point with Torr weights, cfgUSAC 6th digit 0), no degeneracy test (4th digit 0).  Anything else is
// refused (MPL_ERR_BAD_ARG): there is no CPU fallback.
//
// Parity: statistical.  Sampler, evaluation pool, SPRT, LOSAC and the stopping rules are the reference's, restated; the
// models differ in rounding and in their order within a sample (own solver), and the repetitions of one local optimisation
// run side by side from the same pool cursor (the reference's cursor only moves between them when an SPRT rejects).
// One reference DEFECT is not restated: in some scenes its local optimisation never improves a model (singulF's 3 x 3 SVD
// port fails on some essential-like matrices; include/mplb200.h, lo_mode, has the measurement).  lo_mode 0 (default) runs the
// optimisation as stated, on the device; lo_mode 1 draws the inner samples and treats their models as rejected, which is
// what the reference does in those scenes.  Scene by scene the reference's runs equal one mode or the other.
// tests/test_usac_gpu.py compares with the reference's own estimator compiled into oracle/_ref/libusac_ref.so.
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "mpl_ransac.h"
#include "small_svd.cuh"
#include "fivept_device.cuh"
#include "glibc_rand.h"
#include "../../include/mplb200.h"

#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

namespace {

// ------------------------------------------------------------------------------------------------------------------
// device side
// ------------------------------------------------------------------------------------------------------------------

// FTools::normalizePoints' two similarity transforms T = [s 0 -s*mx; 0 s -s*my; 0 0 1] (FundmatrixFunctions.cpp:42-53)
struct NormT { double s1, tx1, ty1, s2, tx2, ty2; };          // tx = -s * mean_x

__device__ __forceinline__ void normalise_pt(const NormT& T, const double4 c, double& x1, double& y1, double& x2, double& y2)
{
    x1 = T.s1 * c.x + T.tx1; y1 = T.s1 * c.y + T.ty1; x2 = T.s2 * c.z + T.tx2; y2 = T.s2 * c.w + T.ty2;     // MathTools::vmul rows
}

// models_ = T2^-T * models_denorm_ * T1^-1 (EssentialMatEstimator.h:511) for similarity T: T^-1 = [1/s 0 mx; 0 1/s my; 0 0 1]
__device__ void normalise_model(const NormT& T, const double* E, double* F)
{
    const double i1 = 1.0 / T.s1, m1x = -T.tx1 * i1, m1y = -T.ty1 * i1, i2 = 1.0 / T.s2, m2x = -T.tx2 * i2, m2y = -T.ty2 * i2;
    double M[9];                                                // E * T1^-1
    for (int r = 0; r < 3; ++r) { M[3 * r] = E[3 * r] * i1; M[3 * r + 1] = E[3 * r + 1] * i1; M[3 * r + 2] = E[3 * r] * m1x + E[3 * r + 1] * m1y + E[3 * r + 2]; }
    for (int c = 0; c < 3; ++c) { F[c] = i2 * M[c]; F[3 + c] = i2 * M[3 + c]; F[6 + c] = m2x * M[c] + m2y * M[3 + c] + M[6 + c]; }
}

// models_denorm_ = T2^T * models_ * T1 (EssentialMatEstimator.h:586-589)
__device__ void denormalise_model(const NormT& T, const double* F, double* E)
{
    double M[9];                                                // F * T1
    for (int r = 0; r < 3; ++r) { M[3 * r] = F[3 * r] * T.s1; M[3 * r + 1] = F[3 * r + 1] * T.s1; M[3 * r + 2] = F[3 * r] * T.tx1 + F[3 * r + 1] * T.ty1 + F[3 * r + 2]; }
    for (int c = 0; c < 3; ++c) { E[c] = T.s2 * M[c]; E[3 + c] = T.s2 * M[3 + c]; E[6 + c] = T.tx2 * M[c] + T.ty2 * M[3 + c] + M[6 + c]; }
}

// evaluateModel's Sampson error (EssentialMatEstimator.h:1135-1143) in its operation order, without contraction:
// r = x1 . (E^T x2), error = r^2 / (|E^T x2|_xy^2 + |E x1|_xy^2)
__device__ __forceinline__ double usac_err(const double (&m)[9], const double4 c)
{
    const double x1 = c.x, y1 = c.y, x2 = c.z, y2 = c.w;
    const double rxc = __dadd_rn(__dadd_rn(__dmul_rn(m[0], x2), __dmul_rn(m[3], y2)), m[6]);
    const double ryc = __dadd_rn(__dadd_rn(__dmul_rn(m[1], x2), __dmul_rn(m[4], y2)), m[7]);
    const double rwc = __dadd_rn(__dadd_rn(__dmul_rn(m[2], x2), __dmul_rn(m[5], y2)), m[8]);
    const double r = __dadd_rn(__dadd_rn(__dmul_rn(x1, rxc), __dmul_rn(y1, ryc)), rwc);
    const double rx = __dadd_rn(__dadd_rn(__dmul_rn(m[0], x1), __dmul_rn(m[1], y1)), m[2]);
    const double ry = __dadd_rn(__dadd_rn(__dmul_rn(m[3], x1), __dmul_rn(m[4], y1)), m[5]);
    const double den = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(rxc, rxc), __dmul_rn(ryc, ryc)), __dmul_rn(rx, rx)), __dmul_rn(ry, ry));
    return __ddiv_rn(__dmul_rn(r, r), den);
}

// validateModel (EssentialMatEstimator.h:1085-1104): the oriented epipolar constraint on the sample, with the NORMALISED model
// and points (models_, input_points_): computeEpipole / getOriSign of FundmatrixFunctions.cpp:362-380.
__device__ bool usac_valid_model(const NormT& T, const double* E, const double4* pts, int n)
{
    double F[9];
    normalise_model(T, E, F);
    const double xeps = 1.9984e-15;
    double e[3] = {F[1] * F[8] - F[2] * F[7], F[2] * F[6] - F[0] * F[8], F[0] * F[7] - F[1] * F[6]};        // row0 x row2
    if (!(e[0] > xeps || e[0] < -xeps || e[1] > xeps || e[1] < -xeps || e[2] > xeps || e[2] < -xeps)) {
        e[0] = F[4] * F[8] - F[5] * F[7]; e[1] = F[5] * F[6] - F[3] * F[8]; e[2] = F[3] * F[7] - F[4] * F[6];   // row1 x row2
    }
    double sig1 = 0;
    for (int i = 0; i < n; ++i) {
        double x1, y1, x2, y2;
        normalise_pt(T, pts[i], x1, y1, x2, y2);
        const double s1 = F[0] * x2 + F[3] * y2 + F[6], s2 = e[1] - e[2] * y1;                            // pt[5] = pt[2] = 1
        const double sig = s1 * s2;
        if (i == 0) sig1 = sig; else if (sig1 * sig < 0) return false;
    }
    return true;
}

// One thread per minimal sample: 5-point solve (generateMinimalSampleModels) + validateModel of every solution.
// samples[s][0] < 0 marks a sample the host rejected (validateSample): no model.  counts[s] = n_solutions | valid_bits << 8.
__global__ void __launch_bounds__(64)
usac_solve_kernel(const double4* __restrict__ corr, int n, const int32_t* __restrict__ samples /* [S][5] */, int S, NormT T,
                  double* __restrict__ models /* [S][10][9] */, int32_t* __restrict__ counts)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    if (samples[s * 5] < 0) { counts[s] = 0; return; }
    double4 pts[5];
    double x1[5], y1[5], x2[5], y2[5];
    for (int j = 0; j < 5; ++j) {
        pts[j] = corr[min(max(samples[s * 5 + j], 0), n - 1)];
        x1[j] = pts[j].x; y1[j] = pts[j].y; x2[j] = pts[j].z; y2[j] = pts[j].w;
    }
    double ee[4][9], tmp[90];
    fivept::null_space_5x9(x1, y1, x2, y2, ee);
    // Durand-Kerner with the early exit (every root moved < 2^-40 relative): 0 differing root sets in 10^6 samples
    // (profiles/r2_solver_early_exit.json), 1.3x less latency -- and latency is what a batch of this schedule costs
    const int cnt = fivept::solve_from_basis<true>(ee, tmp);
    double* out = models + (size_t)s * 90;
    int valid = 0;
    for (int m = 0; m < cnt; ++m) {
        for (int k = 0; k < 9; ++k) out[9 * m + k] = tmp[9 * m + k];
        if (usac_valid_model(T, tmp + 9 * m, pts, 5)) valid |= 1 << m;
    }
    counts[s] = cnt | (valid << 8);
}

// Inlier bits of every valid model: bit i of row (s, m) = usac_err(model, corr[i]) < th2 (strict, EssentialMatEstimator.h:1145).
// A warp per model, 32 correspondences per step.
__global__ void __launch_bounds__(256)
usac_bits_kernel(const double* __restrict__ models, const int32_t* __restrict__ counts, int S, const double4* __restrict__ corr, int n,
                 double th2, uint32_t* __restrict__ bits /* [S][10][W] */, int W)
{
    const int slot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (slot >= S * 10) return;
    const int s = slot / 10, j = slot - s * 10;
    if (!((counts[s] >> (8 + j)) & 1)) return;
    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = models[(size_t)slot * 9 + k];
    for (int w = 0; w < W; ++w) {
        const int i = w * 32 + lane;
        const bool in = i < n && usac_err(E, corr[i]) < th2;
        const unsigned b = __ballot_sync(0xffffffffu, in);
        if (lane == 0) bits[(size_t)slot * W + w] = b;
    }
}

// ---- local optimisation (USAC::locallyOptimizeSolution, USAC.h:948-1060): one CTA per inner-RANSAC repetition -------------
constexpr int kLoThreads = 512;
constexpr int kLoWarps = kLoThreads / 32;
constexpr int kLoGroups = kLoThreads / 45;          // 11 groups of 45 threads: thread (g, k) sums product k of the covariance over a tile's rows g, g + 11, ...
constexpr int kLoMaxSample = 16;

struct LoParams {
    int n, W, sample_size, n_steps, pool_start;
    double th2, mult;                      // usac_inlier_threshold_ (squared), lo_threshold_multiplier_
    double delta, epsilon, A;              // the SPRT in force (evaluateModel uses, never updates, them)
    NormT T;
};
struct LoOut { double E[9]; int status, count, shift, n_eval; long long clk[4]; };     // status 0: the repetition reached its final evaluation and passed

struct LoShared {
    double E[9], F[9];                     // models_denorm_[0], models_[0]
    double part[kLoGroups][45];
    double C[45];                          // upper triangle of the 9 x 9 covariance, row-major
    int cnt[kLoWarps];
    int flag, tested;
    double pa[33], pb[33];                 // (delta / epsilon)^i, ((1 - delta) / (1 - epsilon))^o
    long long clk[4];                      // phase clocks (thread 0): fit accumulation, eigen + projection, errors, SPRT walk
};

__host__ __device__ constexpr int tri9(int a, int b) { return a * 9 - (a * (a - 1)) / 2 + (b - a); }     // a <= b

__device__ __forceinline__ double half_warp_sum(double v)     // called by lanes 0..15 only; identical in all of them (control flow stays uniform)
{
#pragma unroll
    for (int o = 8; o >= 1; o >>= 1) v += __shfl_xor_sync(0x0000ffffu, v, o);
    return v;
}

// Eigenvector of the smallest eigenvalue of the symmetric 9 x 9 covariance in S.C -- what generateRefinedModel takes from
// MathTools::svdu1v (EssentialMatEstimator.h:566-583) -- by a one-sided Jacobi SVD run by lanes 0..15 of a warp: lane i holds row i of every
// column (a symmetric positive semi-definite matrix' right singular vectors are its eigenvectors).  Result in S.F (unnormalised F).
__device__ void warp_smallest_eigvec(LoShared& S)
{
    const int lane = threadIdx.x & 31;
    double U[9], V[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) {
        U[j] = lane < 9 ? S.C[lane <= j ? tri9(lane, j) : tri9(j, lane)] : 0.0;
        V[j] = lane == j ? 1.0 : 0.0;
    }
    // Round-robin ordering: in round r the four disjoint column pairs ((r + k) mod 9, (r - k) mod 9), k = 1..4, are rotated
    // together -- their twelve reductions and four chains of scalar fp64 (a divide, a square root, a reciprocal square root
    // each) overlap instead of queueing behind one another; 9 rounds visit all 36 pairs once.
    for (int sweep = 0; sweep < 40; ++sweep) {
        bool rotated = false;
#pragma unroll
        for (int r = 0; r < 9; ++r) {
            double a[4], b[4], c[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int p = (r + k + 1) % 9, q = (r + 8 - k) % 9;
                a[k] = U[p] * U[p]; b[k] = U[q] * U[q]; c[k] = U[p] * U[q];
            }
#pragma unroll
            for (int o = 8; o >= 1; o >>= 1)
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    a[k] += __shfl_xor_sync(0x0000ffffu, a[k], o); b[k] += __shfl_xor_sync(0x0000ffffu, b[k], o); c[k] += __shfl_xor_sync(0x0000ffffu, c[k], o);
                }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int p = (r + k + 1) % 9, q = (r + 8 - k) % 9;
                const bool skip = c[k] * c[k] <= 4.930380657631324e-32 * (a[k] * b[k]) || fabs(c[k]) <= 1e-300;      // |c| <= eps * sqrt(a b)
                rotated |= !skip;
                const double tau = b[k] - a[k];
                const double t = skip ? 0.0 : (tau >= 0 ? 2.0 : -2.0) * c[k] / (fabs(tau) + sqrt(tau * tau + 4.0 * c[k] * c[k]));
                const double cs = rsqrt(1.0 + t * t), sn = cs * t;
                const double up = U[p], uq = U[q], vp = V[p], vq = V[q];
                U[p] = cs * up - sn * uq; U[q] = sn * up + cs * uq;
                V[p] = cs * vp - sn * vq; V[q] = sn * vp + cs * vq;
            }
        }
        if (!rotated) break;
    }
    double wbest = 0; int best = 0;
#pragma unroll
    for (int j = 0; j < 9; ++j) { const double w = half_warp_sum(U[j] * U[j]); if (j == 0 || w < wbest) { wbest = w; best = j; } }     // :570-577
    double x = 0;
#pragma unroll
    for (int j = 0; j < 9; ++j) if (j == best) x = V[j];
    if (lane < 9) S.F[lane] = x;
}

// covariance of the (weighted) data-matrix rows of the selected correspondences -> models_[0], models_denorm_[0]
// (generateRefinedModel, REFINE_WEIGHTS).  Selection: sample list (sel_n > 0), else the first `limit` correspondences in
// index order with err < thr (USAC.h:1019-1021 passes temp_inliers, the count at the inlier threshold, as numPoints for a
// list found at a looser threshold).  Returns the number of points used; < 5 leaves the model untouched (:535-537).
// A tile of kLoThreads rows goes through shared memory (`rows`); thread (g, k) accumulates product k over the tile.
__device__ int lo_fit(LoShared& S, const LoParams& P, const double4* __restrict__ corr, const double* __restrict__ err, double* rows,
                      const int32_t* sel, int sel_n, double thr, int limit, bool weighted)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = tid / 45, k = tid - g * 45;
    const bool summer = g < kLoGroups;
    int ka = 0, kb = 0;
    { int r = k; for (ka = 0; ka < 9; ++ka) { if (r < 9 - ka) { kb = ka + r; break; } r -= 9 - ka; } }
    double F[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) F[j] = S.F[j];
    double acc = 0.0;
    auto put_row = [&](int i, bool take) {
        double r[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        if (take) {
            double x1, y1, x2, y2;
            normalise_pt(P.T, corr[i], x1, y1, x2, y2);
            double w = 1.0;
            if (weighted) {                                                                // findWeights (EssentialMatEstimator.h:2376-2389)
                const double rxc = F[0] * x2 + F[3] * y2 + F[6], ryc = F[1] * x2 + F[4] * y2 + F[7];
                const double rx = F[0] * x1 + F[1] * y1 + F[2], ry = F[3] * x1 + F[4] * y1 + F[5];
                w = 1.0 / sqrt(rxc * rxc + ryc * ryc + rx * rx + ry * ry);
            }
            r[0] = x2 * x1 * w; r[1] = x2 * y1 * w; r[2] = x2 * w; r[3] = y2 * x1 * w; r[4] = y2 * y1 * w; r[5] = y2 * w; r[6] = x1 * w; r[7] = y1 * w; r[8] = w;   // computeDataMatrix
        }
#pragma unroll
        for (int j = 0; j < 9; ++j) rows[tid * 9 + j] = r[j];
    };
    int used = 0;
    const long long c0 = clock64();
    if (sel_n > 0) {
        put_row(tid < sel_n ? sel[tid] : 0, tid < sel_n);
        __syncthreads();
        if (summer) for (int pp = g; pp < sel_n; pp += kLoGroups) acc += rows[pp * 9 + ka] * rows[pp * 9 + kb];
        __syncthreads();
        used = sel_n;
    } else {
        int base = 0;
        for (int t0 = 0; t0 < P.n && base < limit; t0 += kLoThreads) {
            const int i = t0 + tid;
            const bool p = i < P.n && err[i] < thr;
            const unsigned m = __ballot_sync(0xffffffffu, p);
            if (lane == 0) S.cnt[warp] = __popc(m);
            __syncthreads();
            int before = 0, total = 0;
#pragma unroll
            for (int w = 0; w < kLoWarps; ++w) { const int c = S.cnt[w]; total += c; if (w < warp) before += c; }
            const int rank = base + before + __popc(m & ((1u << lane) - 1u));
            put_row(i, p && rank < limit);
            base += total;
            __syncthreads();
            const int tile_n = min(kLoThreads, P.n - t0);
            if (summer) for (int pp = g; pp < tile_n; pp += kLoGroups) acc += rows[pp * 9 + ka] * rows[pp * 9 + kb];
            __syncthreads();
        }
        used = min(base, limit);
    }
    if (used < 5) return used;
    const long long c1 = clock64();
    if (summer) S.part[g][k] = acc;
    __syncthreads();
    if (tid < 45) { double v = 0; for (int gg = 0; gg < kLoGroups; ++gg) v += S.part[gg][tid]; S.C[tid] = v; }
    __syncthreads();
    if (tid < 16) warp_smallest_eigvec(S);
    __syncthreads();
    if (tid == 0) {
        // FTools::singulF (FundmatrixFunctions.cpp:334-360): drop the smallest singular value, F - (F v3) v3^T
        double F0[3][3], v[3];
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) F0[a][b] = S.F[3 * a + b];
        jacobi_null_vector<3>(F0, v);
        double Fv[3];
        for (int a = 0; a < 3; ++a) Fv[a] = F0[a][0] * v[0] + F0[a][1] * v[1] + F0[a][2] * v[2];
        for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) S.F[3 * a + b] = F0[a][b] - Fv[a] * v[b];
        denormalise_model(P.T, S.F, S.E);
        S.clk[0] += c1 - c0; S.clk[1] += clock64() - c1;
    }
    __syncthreads();
    return used;
}

// evaluateModel(0, ...) (EssentialMatEstimator.h:1110-1174): errors of models_denorm_[0] for every correspondence, then the
// SPRT over them in evaluation-pool order from the cursor.  On acceptance *count = inliers at the threshold.
__device__ bool lo_eval(LoShared& S, const LoParams& P, const double4* __restrict__ corr, const int32_t* __restrict__ pool,
                        double* __restrict__ err, uint32_t* sseq, int& cursor, int* count, int& n_eval)
{
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = S.E[k];
    const long long c0 = clock64();
    int c = 0;
    for (int i = tid; i < P.n; i += kLoThreads) { const double e = usac_err(E, corr[i]); err[i] = e; c += e < P.th2 ? 1 : 0; }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if (lane == 0) S.cnt[warp] = c;
    __syncthreads();                                                                       // err[] visible to the whole CTA
    for (int k0 = 0; k0 < P.W * 32; k0 += kLoThreads) {
        const int k = k0 + tid;
        bool in = false;
        if (k < P.n) { int pos = cursor + k; if (pos >= P.n) pos -= P.n; in = err[pool[pos]] < P.th2; }
        const unsigned m = __ballot_sync(0xffffffffu, in);
        if (lane == 0 && (k >> 5) < P.W) sseq[k >> 5] = m;
    }
    __syncthreads();
    if (tid == 0) {
        const long long c1 = clock64();
        const double fa = P.delta / P.epsilon, fb = (1 - P.delta) / (1 - P.epsilon), floor_v = DBL_EPSILON * 10;
        // The bit-serial walk of evaluateModel for the first words (where lam starts at 1 and most rejections happen), then a
        // word-level pass: with i inliers and o outliers in a word, no value inside the word exceeds lam_in * fb^o and the value
        // after it is at most max(lam_in * fa^i * fb^o, floor * fb^o) (a clamp inside restarts from the floor).  If that bound
        // never passes A the test accepts -- the common case for a refitted model -- at W steps instead of n; if it does,
        // the bit-serial walk resumes from where it stopped.
        double lam = 1.0;
        int tested = P.n, good = 1;
        auto walk = [&](int w0, int w1) {
            for (int w = w0; w < w1 && good; ++w) {
                unsigned word = sseq[w];
                const int nb = min(32, P.n - 32 * w);
                for (int b = 0; b < nb; ++b, word >>= 1) {
                    lam *= (word & 1u) ? fa : fb;
                    if (lam <= DBL_EPSILON) lam = floor_v;
                    if (lam > P.A) { good = 0; tested = 32 * w + b + 1; break; }
                }
            }
        };
        const int head = min(P.W, 4);
        walk(0, head);
        if (good && head < P.W) {
            double ub = lam;
            bool sure = true;
            for (int w = head; w < P.W; ++w) {
                const int nb = min(32, P.n - 32 * w);
                const int o = nb - __popc(sseq[w]), i = nb - o;
                const double up = S.pb[o];
                if (ub * up * 1.000001 > P.A) { sure = false; break; }
                ub = fmax(ub * S.pa[i] * up, floor_v * up) * 1.000001;
            }
            if (!sure) walk(head, P.W);
        }
        S.flag = good; S.tested = tested;
        S.clk[2] += c1 - c0; S.clk[3] += clock64() - c1;
    }
    __syncthreads();
    ++n_eval;
    const bool good = S.flag != 0;
    if (!good) { cursor += S.tested; if (cursor >= P.n) cursor -= P.n; }
    int total = 0;
#pragma unroll
    for (int w = 0; w < kLoWarps; ++w) total += S.cnt[w];
    *count = total;
    __syncthreads();
    return good;
}

__global__ void __launch_bounds__(kLoThreads)
usac_lo_kernel(const double4* __restrict__ corr, const int32_t* __restrict__ pool, const int32_t* __restrict__ samples /* [reps][kLoMaxSample] */,
               LoParams P, double* __restrict__ err_all /* [reps][n] */, LoOut* __restrict__ out, uint32_t* __restrict__ bits_out /* [reps][W] */)
{
    extern __shared__ double lo_smem[];                 // kLoThreads x 9 row tile, then W words of SPRT bits
    double* rows = lo_smem;
    uint32_t* sseq = (uint32_t*)(lo_smem + kLoThreads * 9);
    __shared__ LoShared S;
    const int rep = blockIdx.x, tid = threadIdx.x;
    double* err = err_all + (size_t)rep * P.n;
    const int32_t* sel = samples + rep * kLoMaxSample;
    int cursor = P.pool_start, n_eval = 0, status = 0, cnt = 0;
    if (tid < 9) { S.E[tid] = 0; S.F[tid] = 0; }
    if (tid < 4) S.clk[tid] = 0;
    if (tid < 33) { S.pa[tid] = pow(P.delta / P.epsilon, (double)tid); S.pb[tid] = pow((1 - P.delta) / (1 - P.epsilon), (double)tid); }
    __syncthreads();
    const double thr_lo = P.mult * P.th2, step = (thr_lo - P.th2) / P.n_steps;           // USAC.h:989-990
    do {
        lo_fit(S, P, corr, err, rows, sel, P.sample_size, 0.0, 0, false);                       // :999 (sample_size >= 5 is the caller's check)
        if (!lo_eval(S, P, corr, pool, err, sseq, cursor, &cnt, n_eval)) { status = 1; break; }     // :1004-1007
        if (lo_fit(S, P, corr, err, rows, nullptr, 0, thr_lo, INT_MAX, false) < 5) { status = 2; break; } // :1008-1018
        for (int j = 0; j < P.n_steps; ++j) {                                              // :1022-1036
            if (!lo_eval(S, P, corr, pool, err, sseq, cursor, &cnt, n_eval)) continue;
            lo_fit(S, P, corr, err, rows, nullptr, 0, thr_lo - (j + 1) * step, cnt, true);
        }
        if (!lo_eval(S, P, corr, pool, err, sseq, cursor, &cnt, n_eval)) { status = 3; break; }     // :1039-1043
    } while (false);
    if (status == 0)
        for (int w0 = 0; w0 < P.W * 32; w0 += kLoThreads) {
            const int i = w0 + tid;
            const unsigned m = __ballot_sync(0xffffffffu, i < P.n && err[i] < P.th2);
            if ((tid & 31) == 0 && (i >> 5) < P.W) bits_out[(size_t)rep * P.W + (i >> 5)] = m;
        }
    if (tid == 0) {
        LoOut o;
        for (int k = 0; k < 9; ++k) o.E[k] = S.E[k];
        int shift = cursor - P.pool_start; if (shift < 0) shift += P.n;
        o.status = status; o.count = cnt; o.shift = shift; o.n_eval = n_eval;
        for (int k = 0; k < 4; ++k) o.clk[k] = S.clk[k];
        out[rep] = o;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// host side: the reference's random stream, samplers, SPRT, stopping rules and USAC::solve over device batches
// ------------------------------------------------------------------------------------------------------------------

// generateUniformRandomSample (USAC.h:620-636)
void uniform_sample(GlibcRand& g, unsigned data_size, unsigned sample_size, unsigned* sample)
{
    unsigned count = 0;
    do {
        const unsigned index = (unsigned)g.next() % data_size;
        if (std::find(sample, sample + count, index) == sample + count) sample[count++] = index;
    } while (count < sample_size);
}

struct UsacHost {
    // configuration (ConfigParams.h; values filled by the caller as usac_estimations.cpp:351-441 does)
    int n = 0, min_sample = 5;
    unsigned max_hyp = 50000;
    double conf = 0.99, th2 = 0;
    bool prosac = false, prevalidate_sample = true;
    unsigned prosac_max_samples = 1000, prosac_min_stop_len = 20; double prosac_beta = 0.09, prosac_non_rand_conf = 0.99;
    const unsigned* sorted = nullptr;
    double tM = 2314.0, mS = 8.5, delta = 0.05, epsilon = 0.15;
    unsigned lo_sample = 14, lo_reps = 5, lo_steps = 4; double lo_mult = 2.0; int lo_mode = 0;
    // PROSAC state (USAC.h:145-150)
    unsigned subset_size = 0, largest_size = 0, stop_len = 0;
    std::vector<unsigned> growth, non_random_inliers, maximality;
    // SPRT state (:153-163)
    double A = 0;
    struct Hist { double epsilon, delta, A; unsigned k; };
    std::vector<Hist> history; unsigned last_update = 0;              // newest last (the reference's list is newest first)
    std::vector<int32_t> pool; unsigned cursor = 0;                   // evaluation_pool_, eval_pool_index_
    // results (UsacResults)
    unsigned hyp_count = 0, model_count = 0, rejected_samples = 0, rejected_models = 0, points_verified = 0, num_lo = 0, best = 0;
    std::vector<uint8_t> flags; double E[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    GlibcRand rng;
    std::vector<double> pn;                                           // input_points_ (normalised), 4 per correspondence

    void design_sprt()                                                // designSPRTTest (:918-942)
    {
        const double C = (1 - delta) * std::log((1 - delta) / (1 - epsilon)) + delta * std::log(delta / epsilon);
        const double K = (tM * C) / mS + 1;
        double An_1 = K, An = 0;
        for (int i = 0; i < 10; ++i) { An = K + std::log(An_1); if (An - An_1 < 1.5e-8) break; An_1 = An; }
        A = An;
    }
    void add_history() { history.push_back(Hist{epsilon, delta, A, hyp_count - last_update}); last_update = hyp_count; }   // :1186-1200

    void init_prosac()                                                // initPROSAC (:716-811)
    {
        const unsigned N = (unsigned)n, m = (unsigned)min_sample;
        growth.assign(N, 0);
        double T_n = prosac_max_samples;
        unsigned T_n_p = 1;
        for (unsigned i = 0; i < m; ++i) T_n *= (double)(m - i) / (N - i);
        for (unsigned i = 0; i < N; ++i) {
            if (i + 1 <= m) { growth[i] = T_n_p; continue; }
            const double temp = (double)(i + 1) * T_n / (i + 1 - m);
            growth[i] = T_n_p + (unsigned)std::ceil(temp - T_n);
            T_n = temp; T_n_p = growth[i];
        }
        non_random_inliers.assign(N, 0);
        std::vector<double> v(N + 1, 0.0);
        for (unsigned nn = m + 1; nn <= N; ++nn) {
            if (nn - 1 > 1000) { non_random_inliers[nn - 1] = non_random_inliers[nn - 2]; continue; }
            std::fill(v.begin(), v.begin() + nn, 0.0);
            v[m] = prosac_beta * std::pow(1.0 - prosac_beta, (double)nn - m - 1) * (nn - m);
            double pn_i = v[m];
            for (unsigned i = m + 2; i <= nn; ++i) {
                if (i == nn) { v[nn - 1] = std::pow(prosac_beta, (double)nn - m); break; }
                v[i - 1] = pn_i * (prosac_beta / (1 - prosac_beta)) * ((double)(nn - i) / (i - m + 1));
                pn_i = v[i - 1];
            }
            double acc = 0.0; unsigned i_min = 0;
            for (unsigned i = nn; i >= m + 1; --i) { acc += v[i - 1]; if (acc < 1 - prosac_non_rand_conf) i_min = i; else break; }
            non_random_inliers[nn - 1] = i_min;
        }
        maximality.assign(N, max_hyp);
        largest_size = m; subset_size = m; stop_len = N;
    }

    void prosac_sample(unsigned hyp, unsigned* sample)               // generatePROSACMinSample (:818-852)
    {
        const unsigned m = (unsigned)min_sample;
        if (hyp > prosac_max_samples) { uniform_sample(rng, (unsigned)n, m, sample); return; }
        if (subset_size > stop_len) uniform_sample(rng, stop_len, m, sample);              // overwritten below, as in the reference
        if (hyp > growth[subset_size - 1]) {
            ++subset_size;
            if (subset_size > (unsigned)n) subset_size = (unsigned)n;
            if (largest_size < subset_size) largest_size = subset_size;
        }
        uniform_sample(rng, subset_size - 1, m - 1, sample);
        sample[m - 1] = subset_size - 1;
        for (unsigned i = 0; i < m; ++i) sample[i] = sorted[sample[i]];
    }

    bool validate_sample(const unsigned* s) const                    // validateSample (EssentialMatEstimator.h:1043-1077): image 1 points
    {
        int i, j, k;
        for (i = 0; i < min_sample; i++) {
            for (j = 0; j < i; j++) {
                const double pix = pn[4 * s[i]], piy = pn[4 * s[i] + 1], pjx = pn[4 * s[j]], pjy = pn[4 * s[j] + 1];
                const double dx1 = pjx - pix, dy1 = pjy - piy;
                for (k = 0; k < j; k++) {
                    const double dx2 = pn[4 * s[k]] - pix, dy2 = pn[4 * s[k] + 1] - piy;
                    if (std::fabs(dx2 * dy1 - dy2 * dx1) <= FLT_EPSILON * (std::fabs(dx1) + std::fabs(dy1) + std::fabs(dx2) + std::fabs(dy2))) break;
                }
                if (k < j) break;
            }
            if (j < i) break;
        }
        return i >= min_sample - 1;
    }

    unsigned clamp_count(double v) const { return (v >= 0 && v < (double)max_hyp) ? (unsigned)v : max_hyp; }   // beyond max_hyp the loop bound decides anyway

    unsigned standard_stopping(unsigned num_inliers, unsigned tot, unsigned sample_size) const     // updateStandardStopping (:1086-1112)
    {
        double ni = 1.0, np = 1.0;
        for (unsigned i = 0; i < sample_size; ++i) { ni *= (double)num_inliers - i; np *= (double)tot - i; }
        const double p = ni / np;
        if (p < DBL_EPSILON) return max_hyp;
        if (1 - p < DBL_EPSILON) return 1;
        return clamp_count(std::ceil(std::log(1 - conf) / std::log(1 - p)));
    }

    unsigned prosac_stopping(unsigned hyp)                            // updatePROSACStopping (:860-910)
    {
        unsigned max_samples = maximality[stop_len - 1];
        unsigned inlier_count = 0;
        for (unsigned i = 0; i < prosac_min_stop_len; ++i) inlier_count += flags[sorted[i]];
        for (unsigned i = prosac_min_stop_len; i < (unsigned)n; ++i) {
            inlier_count += flags[sorted[i]];
            if (non_random_inliers[i] < inlier_count) {
                non_random_inliers[i] = inlier_count;
                if (i == (unsigned)n - 1 || (flags[sorted[i]] && !flags[sorted[i + 1]])) {
                    unsigned new_samples = standard_stopping(inlier_count, i + 1, (unsigned)min_sample);
                    if (i + 1 < largest_size) new_samples += hyp - growth[i];
                    if (new_samples < maximality[i]) {
                        maximality[i] = new_samples;
                        if (new_samples < max_samples || (new_samples == max_samples && i + 1 >= stop_len)) { stop_len = i + 1; max_samples = new_samples; }
                    }
                }
            }
        }
        return max_samples;
    }

    static double exp_sprt(double new_eps, double eps, double dlt)   // computeExpSPRT (:1166-1180)
    {
        const double al = std::log(dlt / eps), be = std::log((1.0 - dlt) / (1.0 - eps));
        const double x0 = std::log(1.0 / (1.0 - new_eps)) / be, v0 = new_eps * std::exp(x0 * al);
        const double x1 = std::log((1.0 - 2.0 * v0) / (1.0 - new_eps)) / be, v1 = new_eps * std::exp(x1 * al) + (1.0 - new_eps) * std::exp(x1 * be);
        return x0 - (x0 - x1) / (1.0 + v0 - v1) * v0;
    }

    unsigned sprt_stopping(unsigned num_inliers, unsigned tot) const  // updateSPRTStopping (:1119-1160)
    {
        double ni = 1.0, np = 1.0, k = 0.0, log_eta = 0.0;
        const double new_eps = (double)num_inliers / tot;
        for (int i = 0; i < min_sample; ++i) { ni *= (double)num_inliers - i; np *= (double)tot - i; }
        const double p = ni / np;
        if (p < DBL_EPSILON) return max_hyp;
        if (1.0 - p < DBL_EPSILON) return 1;
        for (size_t t = history.size(); t-- > 0;) {
            const Hist& h = history[t];
            k += h.k;
            const double hh = exp_sprt(new_eps, h.epsilon, h.delta);
            const double prob_reject_good = 1.0 / std::exp(hh * std::log(h.A));
            log_eta += (double)h.k * std::log(1.0 - p * (1.0 - prob_reject_good));
        }
        const double ns = k + (std::log(1.0 - conf) - log_eta) / std::log(1.0 - p * (1.0 - (1.0 / A)));
        return clamp_count(std::ceil(ns));
    }

    // evaluateModel's SPRT (EssentialMatEstimator.h:1120-1173) over a model's inlier bits, from the pool cursor
    bool evaluate(const uint32_t* bits, unsigned* inliers, unsigned* tested)
    {
        const double fa = delta / epsilon, fb = (1 - delta) / (1 - epsilon);
        double lam = 1.0;
        unsigned cnt = 0;
        for (unsigned i = 0; i < (unsigned)n; ++i) {
            if (cursor > (unsigned)n - 1) cursor = 0;
            const unsigned pt = (unsigned)pool[cursor++];
            const bool in = (bits[pt >> 5] >> (pt & 31)) & 1u;
            if (in) { ++cnt; lam *= fa; } else lam *= fb;
            if (lam <= DBL_EPSILON) lam = DBL_EPSILON * 10;
            if (lam > A) { *inliers = cnt; *tested = i + 1; return false; }
        }
        *inliers = cnt; *tested = (unsigned)n;
        return true;
    }
};

struct UsacDevice {
    cudaStream_t st; MplBuf* B;
    const double4* d_corr; const int32_t* d_pool; int n, W; NormT T;
    int err = 0; long long launches = 0; int batches = 0, lo_launches = 0;
    void ck(cudaError_t e) { if (e != cudaSuccess && !err) { mpl_set_error("usac: %s", cudaGetErrorString(e)); err = MPL_ERR_CUDA; } }

    // one batch of minimal samples -> models, (n_solutions | valid bits << 8), inlier bits; results land in pinned host memory
    void generate(const int32_t* samples, int S, double th2, const double** models, const int32_t** counts, const uint32_t** bits)
    {
        const size_t o_s = 0, o_m = ((size_t)S * 20 + 255) & ~(size_t)255, o_c = o_m + (size_t)S * 720, o_b = (o_c + (size_t)S * 4 + 255) & ~(size_t)255,
                     total = o_b + (size_t)S * 10 * W * 4;
        if (err || (err = B[MPL_BUF_R3].ensure(total))) return;
        char* D = (char*)B[MPL_BUF_R3].ptr;
        char* Hs = nullptr;
        if ((err = mpl_stage_ensure(total, (void**)&Hs))) return;
        memcpy(Hs + o_s, samples, (size_t)S * 20);
        ck(cudaMemcpyAsync(D + o_s, Hs + o_s, (size_t)S * 20, cudaMemcpyHostToDevice, st));
        usac_solve_kernel<<<(S + 63) / 64, 64, 0, st>>>(d_corr, n, (const int32_t*)(D + o_s), S, T, (double*)(D + o_m), (int32_t*)(D + o_c));
        ck(cudaGetLastError()); mpl_count_launch();
        usac_bits_kernel<<<(S * 10 + 7) / 8, 256, 0, st>>>((const double*)(D + o_m), (const int32_t*)(D + o_c), S, d_corr, n, th2, (uint32_t*)(D + o_b), W);
        ck(cudaGetLastError()); mpl_count_launch();
        ck(cudaMemcpyAsync(Hs + o_m, D + o_m, total - o_m, cudaMemcpyDeviceToHost, st));
        ck(cudaStreamSynchronize(st));
        launches += 2; ++batches;
        *models = (const double*)(Hs + o_m); *counts = (const int32_t*)(Hs + o_c); *bits = (const uint32_t*)(Hs + o_b);
    }

    // one local optimisation: `reps` repetitions side by side; results in the SECOND pinned area (the batch's stay valid)
    void local_opt(const int32_t* samples, int reps, const LoParams& P, const LoOut** out, const uint32_t** bits)
    {
        const size_t o_s = 0, o_e = 1024, o_o = (o_e + (size_t)reps * n * 8 + 255) & ~(size_t)255, o_b = (o_o + (size_t)reps * sizeof(LoOut) + 255) & ~(size_t)255,
                     total = o_b + (size_t)reps * W * 4;
        if (err || (err = B[MPL_BUF_R4].ensure(total))) return;
        char* D = (char*)B[MPL_BUF_R4].ptr;
        char* Hs = nullptr;
        const size_t h_total = 1024 + (total - o_o);
        if ((err = mpl_stage2_ensure(h_total, (void**)&Hs))) return;
        const size_t lo_smem_bytes = (size_t)kLoThreads * 9 * 8 + (size_t)W * 4;
        ck(cudaFuncSetAttribute(usac_lo_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lo_smem_bytes));
        memcpy(Hs, samples, (size_t)reps * kLoMaxSample * 4);
        ck(cudaMemcpyAsync(D + o_s, Hs, (size_t)reps * kLoMaxSample * 4, cudaMemcpyHostToDevice, st));
        usac_lo_kernel<<<reps, kLoThreads, lo_smem_bytes, st>>>(d_corr, d_pool, (const int32_t*)(D + o_s), P, (double*)(D + o_e), (LoOut*)(D + o_o), (uint32_t*)(D + o_b));
        ck(cudaGetLastError()); mpl_count_launch();
        ck(cudaMemcpyAsync(Hs + 1024, D + o_o, total - o_o, cudaMemcpyDeviceToHost, st));
        ck(cudaStreamSynchronize(st));
        ++launches; ++lo_launches;
        *out = (const LoOut*)(Hs + 1024); *bits = (const uint32_t*)(Hs + 1024 + (o_b - o_o));
    }
};

bool usac_trace() { static const bool on = getenv("MPL_USAC_TRACE") != nullptr; return on; }

// USAC::locallyOptimizeSolution (USAC.h:948-1060)
unsigned locally_optimise(UsacHost& H, UsacDevice& G, unsigned best_inliers)
{
    if (best_inliers < 2 * H.lo_sample) return 0;                                          // :951-954
    const unsigned sz = std::min(H.lo_sample, best_inliers / 2);
    std::vector<unsigned> orig;                                                            // findInliers(err_ptr_[1], threshold): the best model's inliers
    orig.reserve(best_inliers);
    for (int i = 0; i < H.n; ++i) if (H.flags[i]) orig.push_back((unsigned)i);
    ++H.num_lo;
    const int reps = (int)H.lo_reps;
    std::vector<int32_t> samples((size_t)reps * kLoMaxSample, 0);
    unsigned tmp[kLoMaxSample];
    for (int r = 0; r < reps; ++r) {                                                       // :995-1001
        uniform_sample(H.rng, best_inliers, sz, tmp);
        for (unsigned j = 0; j < sz; ++j) samples[(size_t)r * kLoMaxSample + j] = (int32_t)orig[tmp[j]];
    }
    if (H.lo_mode == 1) return best_inliers;          // what the reference does where its singulF fails (include/mplb200.h, lo_mode): no repetition improves
    LoParams P;
    P.n = H.n; P.W = G.W; P.sample_size = (int)sz; P.n_steps = (int)H.lo_steps; P.pool_start = (int)(H.cursor > (unsigned)H.n - 1 ? 0 : H.cursor);
    P.th2 = H.th2; P.mult = H.lo_mult; P.delta = H.delta; P.epsilon = H.epsilon; P.A = H.A; P.T = G.T;
    const LoOut* out = nullptr; const uint32_t* bits = nullptr;
    G.local_opt(samples.data(), reps, P, &out, &bits);
    if (G.err) return 0;
    unsigned lo_inliers = best_inliers, shift = 0;
    for (int r = 0; r < reps; ++r) {
        shift += (unsigned)out[r].shift;
        if (usac_trace()) fprintf(stderr, "  LO rep %d: status %d count %d evals %d shift %d; clocks fit %lld eig %lld err %lld sprt %lld\n", r, out[r].status, out[r].count, out[r].n_eval, out[r].shift, out[r].clk[0], out[r].clk[1], out[r].clk[2], out[r].clk[3]);
        if (out[r].status == 0 && (unsigned)out[r].count > lo_inliers) {                   // :1045-1050 storeSolution(0, lo_inliers)
            lo_inliers = (unsigned)out[r].count;
            H.best = lo_inliers;
            for (int i = 0; i < H.n; ++i) H.flags[i] = (bits[(size_t)r * G.W + (i >> 5)] >> (i & 31)) & 1u;
            memcpy(H.E, out[r].E, 72);
        }
    }
    H.cursor = (unsigned)((P.pool_start + shift) % (unsigned)H.n);
    return lo_inliers;
}

// USAC::solve (USAC.h:345-612) over device batches.  Returns false like the reference (too few points).
bool usac_solve(UsacHost& H, UsacDevice& G)
{
    const unsigned N = (unsigned)H.n;
    unsigned adaptive = H.max_hyp;
    bool update_sprt_stopping = true;
    if (N < (unsigned)H.min_sample || (H.prosac && N < H.prosac_min_stop_len)) return false;          // :351-356
    const unsigned max2 = H.max_hyp / 2, max3 = 2 * H.max_hyp / 3;
    struct Snap { GlibcRand rng; unsigned subset, largest; };
    std::vector<int32_t> samples; std::vector<Snap> snaps;
    // A batch costs one solver latency (~1.2 ms) whatever its size, so batches are as large as the stopping count allows, bounded
    // by the size of the inlier bit matrix that comes back (8 MiB) and, before a first model exists, by what is plausibly needed.
    const unsigned cap_bytes = std::max(16u, (unsigned)((8u << 20) / ((size_t)40 * ((N + 31) / 32))));
    unsigned batch_cap = std::min(H.prosac ? 64u : 512u, cap_bytes);
    while (H.hyp_count < adaptive && H.hyp_count < H.max_hyp) {
        const unsigned next = H.hyp_count + 1;
        if (next == max2 && H.best == 0) H.th2 *= 1.33;                                   // :367-373 (a batch never runs across these counts)
        else if (next == max3 && H.best == 0) H.th2 *= 1.13;
        unsigned B = std::min(std::min(adaptive, H.max_hyp) - H.hyp_count, batch_cap);
        if (next < max2) B = std::min(B, max2 - next); else if (next < max3) B = std::min(B, max3 - next);
        B = std::max(B, 1u);
        samples.assign((size_t)B * 5, 0); snaps.resize(B);
        for (unsigned b = 0; b < B; ++b) {                                                 // 1. generate sample (:378-414), 2. validate it (:418-428)
            unsigned s[5];
            if (H.prosac) H.prosac_sample(next + b, s); else uniform_sample(H.rng, N, 5, s);
            snaps[b] = Snap{H.rng, H.subset_size, H.largest_size};
            const bool ok = !H.prevalidate_sample || H.validate_sample(s);
            for (int j = 0; j < 5; ++j) samples[(size_t)b * 5 + j] = ok ? (int32_t)s[j] : -1;
        }
        const double* models = nullptr; const int32_t* counts = nullptr; const uint32_t* bits = nullptr;
        G.generate(samples.data(), (int)B, H.th2, &models, &counts, &bits);               // 3. generate models (:432-434) + their inlier bits
        if (G.err) return true;
        bool dropped = false;
        for (unsigned b = 0; b < B && !dropped; ++b) {
            if (!(H.hyp_count < adaptive && H.hyp_count < H.max_hyp)) break;
            ++H.hyp_count;
            if (samples[(size_t)b * 5] < 0) { ++H.rejected_samples; continue; }
            const unsigned nsol = (unsigned)(counts[b] & 0xff), valid = (unsigned)counts[b] >> 8;
            H.model_count += nsol;
            bool update_best = false;
            for (unsigned i = 0; i < nsol; ++i) {                                          // 4-5. validate + evaluate (:438-504)
                if (!((valid >> i) & 1u)) { ++H.rejected_models; continue; }
                unsigned inl = 0, tested = 0;
                const uint32_t* row = bits + ((size_t)b * 10 + i) * G.W;
                const bool good = H.evaluate(row, &inl, &tested);
                if (!good) {
                    H.points_verified += tested;
                    const double delta_new = (double)inl / tested;
                    if (delta_new > 0 && std::fabs(H.delta - delta_new) / H.delta > 0.1) { H.add_history(); H.delta = delta_new; H.design_sprt(); }
                } else {
                    H.points_verified += N;
                    if (inl > H.best) {
                        update_best = true;
                        H.best = inl;
                        H.add_history();
                        H.epsilon = (double)H.best / N;
                        H.design_sprt();
                        update_sprt_stopping = true;
                        for (unsigned p = 0; p < N; ++p) H.flags[p] = (row[p >> 5] >> (p & 31)) & 1u;        // storeSolution (:1208-1228)
                        memcpy(H.E, models + ((size_t)b * 10 + i) * 9, 72);
                    }
                }
            }
            if (update_best) {
                // When the local optimisation draws (or PROSAC's stopping length moves, which changes what the sampler consumes)
                // the sampler continues from the state it had after THIS hypothesis' sample -- the reference's stream is
                // sample, LO draws, next sample -- and the rest of the batch is dropped and drawn again.
                auto rewind = [&]() { if (!dropped) { H.rng = snaps[b].rng; H.subset_size = snaps[b].subset; H.largest_size = snaps[b].largest; dropped = true; } };
                if (H.best >= 2 * H.lo_sample) rewind();
                if (usac_trace()) fprintf(stderr, "hyp %u: new best %u (eps %.4f delta %.4f A %.3f)\n", H.hyp_count, H.best, H.epsilon, H.delta, H.A);
                const unsigned lo = locally_optimise(H, G, H.best);                       // 7. (:526-566); no degeneracy test in this build
                if (G.err) return true;
                if (lo > H.best) H.best = lo;
                const unsigned stop_len_before = H.stop_len;
                if (H.prosac && H.hyp_count <= H.prosac_max_samples) adaptive = H.prosac_stopping(H.hyp_count);       // 8. (:570-582)
                else adaptive = H.standard_stopping(H.best, N, (unsigned)H.min_sample);
                if (H.stop_len != stop_len_before) rewind();
                if (usac_trace()) fprintf(stderr, "hyp %u: after LO best %u, stopping count %u (stop_len %u)\n", H.hyp_count, H.best, adaptive, H.stop_len);
            }
            if (!H.prosac && H.hyp_count >= adaptive && update_sprt_stopping) {           // :584-591
                adaptive = H.sprt_stopping(H.best, N);
                update_sprt_stopping = false;
                if (usac_trace()) fprintf(stderr, "hyp %u: SPRT stopping count %u\n", H.hyp_count, adaptive);
            }
        }
        batch_cap = std::min(H.best ? 1024u : 512u, cap_bytes);
    }
    return true;
}

}  // namespace

extern "C" int mpl_usac_rand_stream(unsigned seed, int count, int* out)
{
    if (!out || count < 0) return MPL_ERR_BAD_ARG;
    GlibcRand g; g.seed(seed);
    for (int i = 0; i < count; ++i) out[i] = g.next();
    return MPL_OK;
}

extern "C" void mpl_usac_params_default(mpl_usac_params* p, double th)
{
    if (!p) return;
    memset(p, 0, sizeof(*p));
    p->th = th; p->conf = 0.99; p->max_hypotheses = 50000; p->prevalidate_sample = 1;                     // usac_estimations.cpp:351-366
    p->prosac_beta = 0.09; p->prosac_non_rand_conf = 0.99; p->prosac_max_samples = 1000; p->prosac_min_stop_len = 20;     // :389-396, pose_estim.cpp:1754
    p->sprt_delta = 0.05; p->sprt_epsilon = 0.15; p->sprt_tM = 2314.0; p->sprt_mS = 8.5;                  // :399-430
    p->lo_inner_sample_size = 14; p->lo_inner_reps = 5; p->lo_num_steps = 4; p->lo_threshold_multiplier = 2.0;            // :380-386
    p->estimator = 0; p->refine = 0; p->degeneracy = 0; p->lo_mode = 0; p->seed = 1;
}

extern "C" int mpl_usac5pt(const double* p1, const double* p2, int n, const unsigned* sorted_idx, const mpl_usac_params* prm,
                           double* E, uint8_t* mask, mpl_usac_result* res)
{
    if (!p1 || !p2 || !prm || !E || n < 0) { mpl_set_error("mpl_usac5pt: bad argument"); return MPL_ERR_BAD_ARG; }
    if (prm->refine != 0 || prm->degeneracy != 0 || (prm->estimator != 0 && prm->estimator != 2)) {
        mpl_set_error("mpl_usac5pt: only estimator NISTER / STEWENIUS with REFINE_WEIGHTS and no degeneracy test is built"); return MPL_ERR_BAD_ARG;
    }
    if (prm->lo_inner_sample_size > kLoMaxSample || prm->lo_inner_sample_size < 5 || prm->lo_inner_reps < 1 || prm->lo_inner_reps > 64 ||
        prm->lo_num_steps < 1 || prm->max_hypotheses < 1 || prm->lo_mode < 0 || prm->lo_mode > 1 || !(prm->th > 0)) { mpl_set_error("mpl_usac5pt: parameter out of range"); return MPL_ERR_BAD_ARG; }
    if (res) memset(res, 0, sizeof(*res));
    for (int k = 0; k < 9; ++k) E[k] = 0.0;
    if (mask && n > 0) memset(mask, 0, (size_t)n);
    if (sorted_idx) for (int i = 0; i < n; ++i) if (sorted_idx[i] >= (unsigned)n) { mpl_set_error("mpl_usac5pt: sorted_idx out of range"); return MPL_ERR_BAD_ARG; }

    UsacHost H;
    H.n = n; H.max_hyp = (unsigned)prm->max_hypotheses; H.conf = prm->conf; H.th2 = prm->th * prm->th;    // USAC.h:218 squares the threshold
    H.prosac = sorted_idx != nullptr; H.sorted = sorted_idx; H.prevalidate_sample = prm->prevalidate_sample != 0;
    H.prosac_max_samples = (unsigned)prm->prosac_max_samples; H.prosac_min_stop_len = (unsigned)prm->prosac_min_stop_len;
    H.prosac_beta = prm->prosac_beta; H.prosac_non_rand_conf = prm->prosac_non_rand_conf;
    H.tM = prm->sprt_tM; H.mS = prm->sprt_mS; H.delta = prm->sprt_delta; H.epsilon = prm->sprt_epsilon;
    H.lo_sample = (unsigned)prm->lo_inner_sample_size; H.lo_reps = (unsigned)prm->lo_inner_reps; H.lo_steps = (unsigned)prm->lo_num_steps; H.lo_mult = prm->lo_threshold_multiplier; H.lo_mode = prm->lo_mode;
    if (n > (1 << 20)) { mpl_set_error("mpl_usac5pt: at most 2^20 correspondences"); return MPL_ERR_LIMIT; }
    if (n < 5 || (H.prosac && (unsigned)n < H.prosac_min_stop_len)) return MPL_OK;         // solve() == false: res->ok stays 0
    H.rng.seed(prm->seed);
    // initDataUSAC (USAC.h:274-336): PROSAC tables, the first SPRT design, the evaluation pool (std::random_shuffle over rand())
    if (H.prosac) H.init_prosac();
    H.design_sprt();
    H.pool.resize(n);
    for (int i = 0; i < n; ++i) H.pool[i] = i;
    for (int i = 1; i < n; ++i) { const int j = H.rng.next() % (i + 1); if (i != j) std::swap(H.pool[i], H.pool[j]); }
    H.flags.assign(n, 0);
    // initProblem (EssentialMatEstimator.h:189-352): FTools::normalizePoints
    NormT T;
    {
        double m1x = 0, m1y = 0, m2x = 0, m2y = 0;
        for (int i = 0; i < n; ++i) { m1x += p1[2 * i]; m1y += p1[2 * i + 1]; m2x += p2[2 * i]; m2y += p2[2 * i + 1]; }
        m1x /= (double)n; m2x /= (double)n; m1y /= (double)n; m2y /= (double)n;
        double d1 = 0, d2 = 0;
        for (int i = 0; i < n; ++i) {
            d1 += std::sqrt((p1[2 * i] - m1x) * (p1[2 * i] - m1x) + (p1[2 * i + 1] - m1y) * (p1[2 * i + 1] - m1y));
            d2 += std::sqrt((p2[2 * i] - m2x) * (p2[2 * i] - m2x) + (p2[2 * i + 1] - m2y) * (p2[2 * i + 1] - m2y));
        }
        d1 /= (double)n; d2 /= (double)n;
        const double s1 = std::sqrt(2.0) / d1, s2 = std::sqrt(2.0) / d2;
        T = NormT{s1, -s1 * m1x, -s1 * m1y, s2, -s2 * m2x, -s2 * m2y};
        H.pn.resize((size_t)n * 4);
        for (int i = 0; i < n; ++i) {
            H.pn[4 * i] = s1 * p1[2 * i] + T.tx1; H.pn[4 * i + 1] = s1 * p1[2 * i + 1] + T.ty1;
            H.pn[4 * i + 2] = s2 * p2[2 * i] + T.tx2; H.pn[4 * i + 3] = s2 * p2[2 * i + 1] + T.ty2;
        }
    }

    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx(); cudaStream_t st = ctx->stream; MplBuf* B = ctx->bufs;
    if ((rc = B[MPL_BUF_R0].ensure((size_t)n * 32))) return rc;
    if ((rc = B[MPL_BUF_R1].ensure((size_t)n * 32))) return rc;
    if ((rc = B[MPL_BUF_R2].ensure((size_t)n * 4))) return rc;
    double* d_p1 = (double*)B[MPL_BUF_R0].ptr; double* d_p2 = d_p1 + 2 * (size_t)n;
    void* d_corr = B[MPL_BUF_R1].ptr;
    MPL_CUDA_CHECK(cudaMemcpyAsync(d_p1, p1, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(d_p2, p2, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(B[MPL_BUF_R2].ptr, H.pool.data(), (size_t)n * 4, cudaMemcpyHostToDevice, st));
    if ((rc = mpl_launch_pack_corr(d_p1, d_p2, n, d_corr, st))) return rc;

    UsacDevice G;
    G.st = st; G.B = B; G.d_corr = (const double4*)d_corr; G.d_pool = (const int32_t*)B[MPL_BUF_R2].ptr; G.n = n; G.W = (n + 31) / 32; G.T = T;
    G.launches = 1;
    const bool ok = usac_solve(H, G);
    if (G.err) return G.err;
    memcpy(E, H.E, 72);
    if (mask) memcpy(mask, H.flags.data(), (size_t)n);
    if (res) {
        res->ok = ok ? 1 : 0;
        res->hyp_count = H.hyp_count; res->model_count = H.model_count; res->rejected_sample_count = H.rejected_samples;
        res->rejected_model_count = H.rejected_models; res->total_points_verified = H.points_verified;
        res->num_local_optimizations = H.num_lo; res->best_inlier_count = H.best;
        if (!H.history.empty()) { res->sprt_delta = H.history.back().delta; res->sprt_epsilon = H.history.back().epsilon; }     // USAC.h:596-600
        res->n_batches = G.batches; res->n_launches = (int)G.launches;
    }
    return MPL_OK;
}
