// arrsac.cu -- poselib::estimateEssentialMat(..., method = "ARRSAC", ...), the API's DEFAULT method.
//
// Reference (paths relative to matchinglib_poselib/source/poselib):
//   estimateEssentialMat -> findEssentialMat(ARRSAC)     source/pose_estim.cpp:866-869, five-point-nister/five-point.cpp:119-123
//   CvModelEstimator3::runARRSAC                          source/five-point-nister/modelest.cpp:197-341
//   theia::Arrsac::Estimate / GenerateInitialHypothesisSet  include/arrsac/arrsac.h:206-350, 352-545 (Raguram et al., ECCV 2008)
//   ProsacSampler / RandomSampler (static cv::RNG)        include/arrsac/prosac_sampler.h:79-161, random_sampler.h:58-75
//   SPRT                                                  include/arrsac/sequential_probability_ratio.h:84-126, source/arrsac/...ratio.cc:51-75
//   EssentialMatEstimatorTheia (EstimateModel / ...Nonminimal / Error)   modelest.cpp:111-195
//   CvEMEstimator::ValidModel (oriented epipolar constraint)             five-point.cpp:534-601
//   robustEssentialRefine (pseudo-Huber IRLS)             source/pose_estim.cpp:337-837; SampsonL1 pose_helper.cpp:115-143;
//                                                         costPseudoHuber BA_driver.cpp:2639-2648; getClosestE pose_helper.cpp:152-176
//
// ARRSAC is a sequential, adaptive schedule over small work items (SPRT parameters, the hypothesis budget and the sampling
// universe change after every hypothesis).  What is data parallel is (1) generating hypotheses -- 5-point solves of minimal
// samples, 8-point solves of the inner-RANSAC samples, the oriented-epipolar model test -- and (2) verifying them: the
// Sampson inlier test of every hypothesis against every correspondence.  So the device produces, per BATCH of samples, the
// models and their inlier BIT MATRIX (one bit per hypothesis and correspondence), and the host walks the reference's schedule
// over those bits exactly as arrsac.h states it: the SPRT is a pure function of a hypothesis' first bits, the preemptive
// breadth-first scoring is a running popcount.  Batches are speculative: the whole PROSAC budget (sample k for every k the
// schedule may reach) is drawn and solved in one launch, the 20 inner-RANSAC samples of a support in another (dropped, with
// the random sampler rewound, when a better support appears mid-way).  Samplers are cv::RNG's multiply-with-carry
// generator with its default state, one per sampler like the reference's function-local statics; the PROSAC stream is
// consumed for every k, also those the schedule spends on inner RANSAC (the reference draws only the k it uses).
//
// Parity: statistical.  The schedule, thresholds and samplers are the reference's; given the same hypotheses it takes the same
// decisions (std::sort included).  The hypotheses differ in rounding (own 5-point solver, see fivept_device.cuh; the
// non-minimal solve is a Hartley-normalised 8-point DLT like cv::findFundamentalMat(FM_8POINT) but through a QR / one-sided
// Jacobi factorisation instead of OpenCV's eigen-decomposition of A^T A), so a different model can win near-ties.
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "mpl_ransac.h"
#include "small_svd.cuh"
#include "fivept_device.cuh"
#include "../../include/mplb200.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

namespace {

// ------------------------------------------------------------------------------------------------------------------
// device side
// ------------------------------------------------------------------------------------------------------------------
constexpr int MAX_SAMPLE = 14;            // nonmin_sample_size (modelest.cpp:267: Arrsac(5, th^2, 500, 100, 14, 8))

__device__ __forceinline__ float sampson_err(const double (&E)[9], const double4 c)     // five-point.cpp:490-503, as ransac5pt.cu
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

__device__ __forceinline__ bool is_zero(double d) { return d < 1e-3 && d > -1e-3; }      // five-point.hpp:76-81

// CvEMEstimator::ValidModel (five-point.cpp:534-601) on the sample's own points.
__device__ bool valid_model(const double* Ein, const double4* pts, int n)
{
    double E[9];
    for (int k = 0; k < 9; ++k) E[k] = Ein[k];
    double At[3][3], x[3], V[3][3], w[3];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) At[i][j] = E[3 * j + i];      // svd of E^T (:556)
    jacobi_null_vector<3>(At, x, V, w);
    double s[3] = {w[0], w[1], w[2]};
    for (int a = 0; a < 2; ++a) for (int b = a + 1; b < 3; ++b) if (s[b] > s[a]) { const double t = s[a]; s[a] = s[b]; s[b] = t; }
    if (s[0] / s[1] > 1.2) return false;                                                   // :558-559
    if (!is_zero(0.01 * s[2] / s[1])) return false;                                        // :560-561
    const double e2[3] = {x[0], x[1], x[2]};                                               // V.col(2): null vector of E^T (sign free: both signs of E are tried)
    bool emult = false;
    int fail = 0;
    for (int pass = 0; pass < 2; ++pass) {
        bool restart = false;
        fail = emult ? 0 : fail;
        for (int i = 0; i < n && !restart; ++i) {
            const double x1 = pts[i].x, y1 = pts[i].y, x2 = pts[i].z, y2 = pts[i].w;
            const double l1[3] = {e2[1] * 1.0 - e2[2] * y2, e2[2] * x2 - e2[0] * 1.0, e2[0] * y2 - e2[1] * x2};      // e2 x x2
            const double l2[3] = {E[0] * x1 + E[1] * y1 + E[2], E[3] * x1 + E[4] * y1 + E[5], E[6] * x1 + E[7] * y1 + E[8]};
            for (int j = 0; j < 3; ++j) {
                if (is_zero(0.1 * l1[j]) || is_zero(0.1 * l2[j])) continue;
                if (l1[j] * l2[j] < 0) {
                    if (!emult) { emult = true; fail = 0; for (int k = 0; k < 9; ++k) E[k] = -E[k]; restart = true; break; }     // :583-589 goto tryocagain
                    ++fail;
                    break;
                }
            }
        }
        if (!restart) break;
    }
    return !((float)fail / (float)n >= 0.4f);                                              // :597-598
}

// One thread per sample.  Minimal samples (size 5): fivept solver.  Larger samples (inner RANSAC, size >= 8): 8-point DLT
// with Hartley normalisation and rank-2 enforcement (cv::findFundamentalMat(FM_8POINT), modelest.cpp:169-178).  Sizes 6, 7
// (possible when few inliers are known: arrsac.h:297-299) go through the n-point form of the 5-point solver like size 5
// does in the reference (random_sampler size < nonmin_sample_size_min -> EstimateModel).  Every model then passes ValidModel.
__global__ void __launch_bounds__(64)
arrsac_solve_kernel(const double4* __restrict__ corr, int n, const int32_t* __restrict__ samples /* [S][MAX_SAMPLE] */,
                    const int32_t* __restrict__ sizes, int S, double* __restrict__ models /* [S][10][9] */, int32_t* __restrict__ n_models)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const int sz = sizes[s];
    double4 pts[MAX_SAMPLE];
    for (int j = 0; j < sz; ++j) { const int id = samples[s * MAX_SAMPLE + j]; pts[j] = corr[min(max(id, 0), n - 1)]; }
    double* out = models + (size_t)s * 90;
    double tmp[90];
    int cnt = 0;
    if (sz < 8) {
        double ee[4][9];
        if (sz == 5) {
            double x1[5], y1[5], x2[5], y2[5];
            for (int j = 0; j < 5; ++j) { x1[j] = pts[j].x; y1[j] = pts[j].y; x2[j] = pts[j].z; y2[j] = pts[j].w; }
            fivept::null_space_5x9(x1, y1, x2, y2, ee);
        } else {                                              // 6 or 7 points: right singular vectors of the 4 smallest singular values
            double A[9][9], x[9], V[9][9], w[9];
            for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) A[i][j] = 0.0;
            for (int i = 0; i < sz; ++i) {
                const double r[9] = {pts[i].x * pts[i].z, pts[i].y * pts[i].z, pts[i].z, pts[i].x * pts[i].w, pts[i].y * pts[i].w, pts[i].w, pts[i].x, pts[i].y, 1.0};
                for (int j = 0; j < 9; ++j) A[i][j] = r[j];
            }
            jacobi_null_vector<9>(A, x, V, w);
            int ord[9]; for (int i = 0; i < 9; ++i) ord[i] = i;
            for (int a = 0; a < 8; ++a) for (int b = a + 1; b < 9; ++b) if (w[ord[b]] > w[ord[a]]) { const int t = ord[a]; ord[a] = ord[b]; ord[b] = t; }
            for (int b = 0; b < 4; ++b) for (int k = 0; k < 9; ++k) ee[b][k] = V[ord[5 + b]][k];
        }
        // early exit of the Durand-Kerner sweeps (every root moved < 2^-40 relative): same root sets in 10^6 measured samples, a fifth of
        // the 300-sweep latency; ARRSAC's parity with the reference is statistical anyway (own sampler streams), USAC does the same
        cnt = fivept::solve_from_basis<true>(ee, tmp);
    } else {
        // run8Point: isotropic normalisation of both point sets, DLT rows (x2 x1, x2 y1, x2, y2 x1, y2 y1, y2, x1, y1, 1)
        double m1x = 0, m1y = 0, m2x = 0, m2y = 0;
        for (int i = 0; i < sz; ++i) { m1x += pts[i].x; m1y += pts[i].y; m2x += pts[i].z; m2y += pts[i].w; }
        const double t = 1.0 / sz;
        m1x *= t; m1y *= t; m2x *= t; m2y *= t;
        double sc1 = 0, sc2 = 0;
        for (int i = 0; i < sz; ++i) {
            sc1 += sqrt((pts[i].x - m1x) * (pts[i].x - m1x) + (pts[i].y - m1y) * (pts[i].y - m1y));
            sc2 += sqrt((pts[i].z - m2x) * (pts[i].z - m2x) + (pts[i].w - m2y) * (pts[i].w - m2y));
        }
        sc1 *= t; sc2 *= t;
        if (sc1 >= FLT_EPSILON && sc2 >= FLT_EPSILON) {
            sc1 = sqrt(2.0) / sc1; sc2 = sqrt(2.0) / sc2;
            // Givens QR of the sz x 9 system into a 9 x 9 triangle, then its smallest right singular vector
            double R[9][9];
            for (int i = 0; i < 9; ++i) for (int j = 0; j < 9; ++j) R[i][j] = 0.0;
            for (int i = 0; i < sz; ++i) {
                const double x1 = (pts[i].x - m1x) * sc1, y1 = (pts[i].y - m1y) * sc1, x2 = (pts[i].z - m2x) * sc2, y2 = (pts[i].w - m2y) * sc2;
                double r[9] = {x2 * x1, x2 * y1, x2, y2 * x1, y2 * y1, y2, x1, y1, 1.0};
                for (int k = 0; k < 9; ++k) {
                    if (r[k] == 0.0) continue;
                    const double a = R[k][k], rho = sqrt(a * a + r[k] * r[k]), cs = a / rho, sn = r[k] / rho;
                    R[k][k] = rho;
                    for (int j = k + 1; j < 9; ++j) { const double t1 = R[k][j], t2 = r[j]; R[k][j] = cs * t1 + sn * t2; r[j] = cs * t2 - sn * t1; }
                }
            }
            double f[9], V[9][9], w[9];
            jacobi_null_vector<9>(R, f, V, w);
            // rank 2: remove the component along the smallest right singular vector (F0 - (F0 v) v^T)
            double F0[3][3], v[3];
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F0[i][j] = f[3 * i + j];
            jacobi_null_vector<3>(F0, v);
            double Fv[3];
            for (int i = 0; i < 3; ++i) Fv[i] = F0[i][0] * v[0] + F0[i][1] * v[1] + F0[i][2] * v[2];
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F0[i][j] -= Fv[i] * v[j];
            // F = T2^T F0 T1 with T = [s 0 -s m; 0 s -s m; 0 0 1]
            const double T1[3][3] = {{sc1, 0, -sc1 * m1x}, {0, sc1, -sc1 * m1y}, {0, 0, 1}}, T2[3][3] = {{sc2, 0, -sc2 * m2x}, {0, sc2, -sc2 * m2y}, {0, 0, 1}};
            double M[3][3], F[3][3];
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { double a = 0; for (int k = 0; k < 3; ++k) a += T2[k][i] * F0[k][j]; M[i][j] = a; }
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { double a = 0; for (int k = 0; k < 3; ++k) a += M[i][k] * T1[k][j]; F[i][j] = a; }
            if (fabs(F[2][2]) > FLT_EPSILON) { const double inv = 1.0 / F[2][2]; for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) F[i][j] *= inv; }   // make F(3,3) = 1
            for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) tmp[3 * i + j] = F[i][j];
            cnt = 1;
        }
    }
    int kept = 0;
    for (int m = 0; m < cnt; ++m) {
        if (!valid_model(tmp + 9 * m, pts, sz)) continue;                                  // modelest.cpp:140-141, 174-175
        for (int k = 0; k < 9; ++k) out[9 * kept + k] = tmp[9 * m + k];
        ++kept;
    }
    n_models[s] = kept;
}

// Inlier bits: bit i of row m = Error(correspondence i, model m) < th^2 (strict: arrsac.h:541 / ...ratio.h:104; Error returns
// the float Sampson error widened to double, modelest.cpp:193).  Warp = one model x 32 correspondences per step (ballot).
__global__ void __launch_bounds__(256)
arrsac_bits_kernel(const double* __restrict__ models /* [S][10][9] */, const int32_t* __restrict__ n_models, int S, const double4* __restrict__ corr,
                   int n, double th2, uint32_t* __restrict__ bits /* [S][10][W] */, int W)
{
    const int slot = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (slot >= S * 10) return;
    const int s = slot / 10, j = slot - s * 10;
    if (j >= n_models[s]) return;
    double E[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) E[k] = models[(size_t)slot * 9 + k];
    for (int w = 0; w < W; ++w) {
        const int i = w * 32 + lane;
        const bool in = i < n && (double)sampson_err(E, corr[i]) < th2;
        const unsigned b = __ballot_sync(0xffffffffu, in);
        if (lane == 0) bits[(size_t)slot * W + w] = b;
    }
}

// robustEssentialRefine (pose_estim.cpp:337-837; model 0, no normalisation, makeClosestE, iters = 0 -> at most 50 sweeps):
// iteratively re-weighted least squares of the epipolar constraint with pseudo-Huber weights on the Sampson L1 distance.
// One CTA: every thread folds its weighted rows into a 9 x 9 triangle (Givens), binary tree over the CTA, the triangle's
// smallest right singular vector (= the eigenvector of A^T A the reference extracts, :613-640, without squaring the
// condition number), closest essential matrix (:152-176), convergence tests (:808-817) by thread 0.
constexpr int kRT = 128;
__host__ __device__ constexpr int tri(int k, int j) { return k * 9 - (k * (k - 1)) / 2 + (j - k); }

template <int K0>
__device__ __forceinline__ void qr_add_row(double (&R)[45], double (&r)[9])
{
#pragma unroll
    for (int k = K0; k < 9; ++k) {
        const double b = r[k];
        if (b != 0.0) {
            const double a = R[tri(k, k)], rho = sqrt(fma(a, a, b * b)), cs = a / rho, sn = b / rho;
            R[tri(k, k)] = rho;
#pragma unroll
            for (int j = k + 1; j < 9; ++j) { const double t1 = R[tri(k, j)], t2 = r[j]; R[tri(k, j)] = fma(cs, t1, sn * t2); r[j] = fma(cs, t2, -(sn * t1)); }
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

struct RefineOut { double E[9]; double err_init, err_final; int iters; int status; };   // status 0 ok, 1 "refinement failed" (:607-611)

__global__ void __launch_bounds__(kRT)
robust_refine_kernel(const double4* __restrict__ corr, const uint8_t* __restrict__ mask, int n, const double* __restrict__ E_init, double th,
                     RefineOut* __restrict__ out)
{
    __shared__ double sR[kRT][45];
    __shared__ double sW[kRT];
    __shared__ double sF[9];
    __shared__ int s_stop;
    const int tid = threadIdx.x;
    if (tid < 9) sF[tid] = E_init[tid];
    if (tid == 0) s_stop = 0;
    __syncthreads();
    int npts = 0;
    for (int i = tid; i < n; i += kRT) npts += mask[i] ? 1 : 0;
    sW[tid] = (double)npts;
    __syncthreads();
    for (int st = kRT / 2; st >= 1; st >>= 1) { if (tid < st) sW[tid] += sW[tid + st]; __syncthreads(); }
    const int np = (int)sW[0];
    __syncthreads();
    const double minDiff = th / 10, minErr = th * th / 100 * np;                          // :448-449
    double err_old = 1e12, err = -9999.0, err_init = 0;
    int j = 0, status = 0;
    for (; j < 50; ++j) {                                                                  // maxIters (:450)
        double F[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) F[k] = sF[k];
        double R[45];
#pragma unroll
        for (int k = 0; k < 45; ++k) R[k] = 0.0;
        double wn = 0.0;
        for (int i = tid; i < n; i += kRT) {
            if (!mask[i]) continue;
            const double4 c = corr[i];
            const double x0 = c.x, y0 = c.y, x1 = c.z, y1 = c.w;
            // SampsonL1 (pose_helper.cpp:115-143)
            const double xe0 = x1 * F[0] + y1 * F[3] + F[6], xe1 = x1 * F[1] + y1 * F[4] + F[7], xe2 = x1 * F[2] + y1 * F[5] + F[8];
            const double num = xe0 * x0 + xe1 * y0 + xe2;
            const double e0 = F[0] * x0 + F[1] * y0 + F[2], e1 = F[3] * x0 + F[4] * y0 + F[5];
            const double denom1 = 1.0 / (sqrt(e0 * e0 + e1 * e1 + xe0 * xe0 + xe1 * xe1) + 1e-8);
            // costPseudoHuber (BA_driver.cpp:2639-2648)
            const double d_abs = fabs(num * denom1) + 1e-12, q = d_abs / th;
            const double weight = sqrt(2 * th * th * (sqrt(1 + q * q) - 1)) / d_abs;
            const double f = denom1 * weight;
            wn += f * f;
            double r[9] = {x1 * x0 * f, x1 * y0 * f, x1 * f, y1 * x0 * f, y1 * y0 * f, y1 * f, x0 * f, y0 * f, f};    // :565-573, 599
            qr_add_row<0>(R, r);
        }
        sW[tid] = wn;
#pragma unroll
        for (int k = 0; k < 45; ++k) sR[tid][k] = R[k];
        __syncthreads();
        for (int st = kRT / 2; st >= 1; st >>= 1) {
            if (tid < st) {
                merge_rows<0>(R, sR[tid + st]);
#pragma unroll
                for (int k = 0; k < 45; ++k) sR[tid][k] = R[k];
                sW[tid] += sW[tid + st];
            }
            __syncthreads();
        }
        if (tid == 0) {
            double A[9][9], x[9], V[9][9], w[9];
            for (int a = 0; a < 9; ++a) for (int b = 0; b < 9; ++b) A[a][b] = b >= a ? sR[0][tri(a, b)] : 0.0;
            jacobi_null_vector<9>(A, x, V, w);
            const double wnorm2 = sW[0];                                                   // weightnorm^2 (:535-540): the rows above are unscaled
            int small = 0; double smin = w[0];
            for (int a = 0; a < 9; ++a) { if (w[a] * w[a] / wnorm2 < DBL_EPSILON) ++small; if (w[a] < smin) smin = w[a]; }
            if (small >= 2) { status = 1; s_stop = 2; }                                    // :601-611 "Refinement failed": E_init is returned
            else {
                // F2 = lastCol reshaped row-major (:632-633), getClosestE (pose_helper.cpp:152-176)
                double F2[3][3], v[3], Vs[3][3], ws[3];
                for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) F2[a][b] = x[3 * a + b];
                jacobi_null_vector<3>(F2, v, Vs, ws);
                double s[3] = {ws[0], ws[1], ws[2]};
                for (int a = 0; a < 2; ++a) for (int b = a + 1; b < 3; ++b) if (s[b] > s[a]) { const double t = s[a]; s[a] = s[b]; s[b] = t; }
                const bool bad = !(s[2] < 1e-3 && s[2] > -1e-3) || s[0] / s[1] > 1.5 || s[0] / s[1] < 0.66;
                if (bad) s_stop = 1;                                                       // :636-643: keep the last valid E
                else {
                    double Fv[3];
                    for (int a = 0; a < 3; ++a) Fv[a] = F2[a][0] * v[0] + F2[a][1] * v[1] + F2[a][2] * v[2];
                    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) sF[3 * a + b] = F2[a][b] - Fv[a] * v[b];       // U diag(s0, s1, 0) V^T
                    err = smin * smin / wnorm2;                                            // (A1 * lastCol).squaredNorm() (:808)
                    if (j == 0) err_init = err;
                    if (j > 1 && (fabs(err_old - err) < minDiff || err < minErr)) s_stop = 1;       // :813-817
                    err_old = err;
                }
            }
        }
        __syncthreads();
        if (s_stop) break;
        __syncthreads();
    }
    if (tid == 0) {
        RefineOut o;
        for (int k = 0; k < 9; ++k) o.E[k] = status ? E_init[k] : sF[k];
        o.err_init = err_init; o.err_final = err; o.iters = j; o.status = status;
        *out = o;
    }
}

// ------------------------------------------------------------------------------------------------------------------
// host side: samplers, SPRT and the ARRSAC schedule over device-generated hypotheses
// ------------------------------------------------------------------------------------------------------------------
struct CvRng {                                   // cv::RNG: multiply-with-carry, default state 0xffffffff; uniform(a, b) = a + next() % (b - a)
    uint64_t state = 0xffffffffull;
    unsigned next() { state = (uint64_t)(unsigned)state * 4164903690u + (unsigned)(state >> 32); return (unsigned)state; }
    int uniform(int a, int b) { return a == b ? a : (int)(next() % (unsigned)(b - a) + a); }
};
thread_local CvRng g_prosac_rng, g_random_rng;  // the reference's are function-local statics: they keep running across calls

// theia::ProsacSampler::Sample for the k-th sample (prosac_sampler.h:90-150), indices into the first `size` data points
void prosac_sample(CvRng& rng, int k, int size, int m, int* out)
{
    double t_n = 200000.0;                       // ransac_convergence_iterations
    int n = m;
    for (int i = 0; i < m; ++i) t_n *= (double)(n - i) / (size - i);
    double t_n_prime = 1.0;
    for (int t = 1; t <= k; ++t) {
        if (t > t_n_prime && n < size) {
            const double t_n_plus1 = (t_n * ((double)n + 1.0)) / ((double)n + 1.0 - (double)m);
            t_n_prime += std::ceil(t_n_plus1 - t_n);
            t_n = t_n_plus1;
            n++;
        }
    }
    int cnt = 0;
    auto used = [&](int v) { for (int i = 0; i < cnt; ++i) if (out[i] == v) return true; return false; };
    if (t_n_prime < k) {
        for (int i = 0; i < m; ++i) { int r; while (used(r = rng.uniform(0, n))) {} out[cnt++] = r; }
    } else {
        for (int i = 0; i < m - 1; ++i) { int r; while (used(r = rng.uniform(0, n - 1))) {} out[cnt++] = r; }
        out[cnt++] = n - 1;
    }
}
// theia::RandomSampler::Sample (random_sampler.h:58-75): sz distinct positions in [0, universe)
void random_sample(CvRng& rng, int universe, int sz, int* out)
{
    int cnt = 0;
    auto used = [&](int v) { for (int i = 0; i < cnt; ++i) if (out[i] == v) return true; return false; };
    for (int i = 0; i < sz; ++i) { int r; while (used(r = rng.uniform(0, universe))) {} out[cnt++] = r; }
}

double sprt_threshold(double sigma, double epsilon, double tm, int num_models)             // sequential_probability_ratio.cc:51-75
{
    const double c = (1.0 - sigma) * std::log((1.0 - sigma) / (1.0 - epsilon)) + sigma * std::log(sigma / epsilon);
    const double a_0 = tm * c / (double)num_models + 1.0;
    double A = a_0;
    for (int i = 0; i < 1000; ++i) {
        const double nA = a_0 + std::log(A);
        const double step = std::fabs(nA - A);
        A = nA;
        if (step < 1e-4) break;
    }
    return A;
}

struct Hyp { int id; double score; };             // ScoredData<Model>: id indexes the hypothesis pool
bool hyp_greater(Hyp a, Hyp b) { return a.score > b.score; }                               // CompareScoredData (arrsac.h:82-85)

struct ArrsacEngine {
    MplCtx* ctx; cudaStream_t st;
    const double4* d_corr; int n, W;
    double th2;
    // hypothesis pool (host copies): model + inlier bits over all n correspondences
    std::vector<double> models; std::vector<uint32_t> bits; int n_hyp = 0;
    long long launches = 0; int batches = 0;
    // device scratch for one batch
    MplBuf* B;
    int err = 0;

    bool bit(int h, int i) const { return (bits[(size_t)h * W + (i >> 5)] >> (i & 31)) & 1u; }

    // Solve + validate + verify a batch of samples on the device; appends the surviving models to the pool and returns, per
    // sample, the pool ids of its hypotheses (in solver order).
    void generate(const std::vector<int>& samp /* S x MAX_SAMPLE */, const std::vector<int>& sizes, std::vector<std::vector<int>>& ids)
    {
        const int S = (int)sizes.size();
        ids.assign(S, {});
        if (S == 0 || err) return;
        const size_t o_s = 0, o_z = o_s + (size_t)S * MAX_SAMPLE * 4, o_m = ((o_z + (size_t)S * 4 + 255) & ~(size_t)255), o_nm = o_m + (size_t)S * 720,
                     o_b = ((o_nm + (size_t)S * 4 + 255) & ~(size_t)255), total = o_b + (size_t)S * 10 * W * 4;
        if ((err = B[MPL_BUF_R3].ensure(total))) return;
        char* D = (char*)B[MPL_BUF_R3].ptr;
        auto ck = [&](cudaError_t e) { if (e != cudaSuccess && !err) { mpl_set_error("arrsac: %s", cudaGetErrorString(e)); err = MPL_ERR_CUDA; } };
        ck(cudaMemcpyAsync(D + o_s, samp.data(), (size_t)S * MAX_SAMPLE * 4, cudaMemcpyHostToDevice, st));
        ck(cudaMemcpyAsync(D + o_z, sizes.data(), (size_t)S * 4, cudaMemcpyHostToDevice, st));
        arrsac_solve_kernel<<<(S + 63) / 64, 64, 0, st>>>(d_corr, n, (const int32_t*)(D + o_s), (const int32_t*)(D + o_z), S, (double*)(D + o_m), (int32_t*)(D + o_nm));
        ck(cudaGetLastError()); mpl_count_launch();
        arrsac_bits_kernel<<<(S * 10 + 7) / 8, 256, 0, st>>>((const double*)(D + o_m), (const int32_t*)(D + o_nm), S, d_corr, n, th2, (uint32_t*)(D + o_b), W);
        ck(cudaGetLastError()); mpl_count_launch();
        launches += 2; ++batches;
        std::vector<double> hm((size_t)S * 90); std::vector<int32_t> hn(S); std::vector<uint32_t> hb((size_t)S * 10 * W);
        ck(cudaMemcpyAsync(hm.data(), D + o_m, hm.size() * 8, cudaMemcpyDeviceToHost, st));
        ck(cudaMemcpyAsync(hn.data(), D + o_nm, hn.size() * 4, cudaMemcpyDeviceToHost, st));
        ck(cudaMemcpyAsync(hb.data(), D + o_b, hb.size() * 4, cudaMemcpyDeviceToHost, st));
        ck(cudaStreamSynchronize(st));
        if (err) return;
        for (int s = 0; s < S; ++s)
            for (int j = 0; j < hn[s] && j < 10; ++j) {
                models.insert(models.end(), hm.begin() + (size_t)(s * 10 + j) * 9, hm.begin() + (size_t)(s * 10 + j + 1) * 9);
                bits.insert(bits.end(), hb.begin() + (size_t)(s * 10 + j) * W, hb.begin() + (size_t)(s * 10 + j + 1) * W);
                ids[s].push_back(n_hyp++);
            }
    }

    // SequentialProbabilityRatioTest (sequential_probability_ratio.h:84-126) over the first `len` correspondences
    bool sprt(int h, int len, double sigma, double epsilon, double A, double* ratio, int* n_inl) const
    {
        int inl = 0; double lr = 1.0;
        for (int i = 0; i < len; ++i) {
            if (bit(h, i)) { lr *= sigma / epsilon; ++inl; } else lr *= (1.0 - sigma) / (1.0 - epsilon);
            if (lr > A) { *ratio = (double)inl / (double)(i + 1); *n_inl = inl; return false; }
        }
        *ratio = (double)inl / (double)len; *n_inl = inl;
        return true;
    }
};

// parameters and running state of theia::Arrsac (arrsac.h:100-118, 206-227)
struct ArrsacState {
    int min_sample = 5, max_cand = 500, block = 100, nonmin = 14, nonmin_min = 8, max_inner_its = 20;
    double sigma = 0.05, epsilon = 0.1, conf = 0.95, tm_ratio = 250.0;
    int verified_accum = 0, n_rejected = 0; double rejected_accum = 0.0;
};

// Arrsac::GenerateInitialHypothesisSet (arrsac.h:206-350) on the first `len` correspondences.
int initial_hypothesis_set(ArrsacEngine& G, ArrsacState& P, int len, std::vector<Hyp>& accepted)
{
    int k = 1, k2 = 0, m_prime = P.max_cand, inner_its = 0, max_inl = 0, rs_size = P.nonmin;
    bool inner = false;
    std::vector<int> universe;                                     // U_in: positions (within the first len) supporting the best hypothesis
    // speculative batches: PROSAC samples k .. k+nb-1 with the sampler state before each; inner samples likewise
    struct Spec { std::vector<std::vector<int>> ids; std::vector<CvRng> before; int k0 = 0, nb = 0; bool valid = false; } sp, si;
    auto fill_prosac = [&](int k0) {
        const int nb = std::min(512, m_prime - k0 + 1);          // the whole remaining PROSAC budget in one launch (the solve's latency does not depend on the count)
        std::vector<int> samp((size_t)nb * MAX_SAMPLE, 0), sizes(nb, 5);
        sp.before.assign(nb + 1, g_prosac_rng);
        for (int j = 0; j < nb; ++j) { sp.before[j] = g_prosac_rng; prosac_sample(g_prosac_rng, k0 + j, len, 5, &samp[(size_t)j * MAX_SAMPLE]); }
        sp.before[nb] = g_prosac_rng;
        G.generate(samp, sizes, sp.ids);
        sp.k0 = k0; sp.nb = nb; sp.valid = true;
    };
    auto fill_inner = [&](int its_left) {
        const int nb = its_left;
        std::vector<int> samp((size_t)nb * MAX_SAMPLE, 0), sizes(nb, rs_size);
        si.before.assign(nb + 1, g_random_rng);
        for (int j = 0; j < nb; ++j) {
            si.before[j] = g_random_rng;
            int pos[MAX_SAMPLE];
            random_sample(g_random_rng, (int)universe.size(), rs_size, pos);
            for (int q = 0; q < rs_size; ++q) samp[(size_t)j * MAX_SAMPLE + q] = universe[pos[q]];
        }
        si.before[nb] = g_random_rng;
        G.generate(samp, sizes, si.ids);
        si.k0 = 0; si.nb = nb; si.valid = true;
    };
    int si_pos = 0;
    while (k <= m_prime && !G.err) {
        std::vector<int> hyps;
        if (!inner) {
            if (!sp.valid || k < sp.k0 || k >= sp.k0 + sp.nb) fill_prosac(k);
            hyps = sp.ids[k - sp.k0];                              // the k-th PROSAC sample (SetSampleNumber(k), prosac_sampler.h:86)
            if (hyps.empty()) { k2++; k++; continue; }
        } else {
            if (!si.valid) { fill_inner(P.max_inner_its - inner_its); si_pos = 0; }
            hyps = si.ids[si_pos];
            g_random_rng = si.before[si_pos + 1];
            ++si_pos;
            inner_its++;
            if (inner_its == P.max_inner_its) { inner_its = 0; inner = false; si.valid = false; }
            if (hyps.empty()) { k2++; k++; continue; }
        }
        P.verified_accum += (int)hyps.size();
        const double A = sprt_threshold(P.sigma, P.epsilon, P.tm_ratio, P.verified_accum / (k - k2));
        for (size_t j = 0; j < hyps.size(); ++j) {
            double ratio; int inl;
            const bool ok = G.sprt(hyps[j], len, P.sigma, P.epsilon, A, &ratio, &inl);
            if (!ok) {
                P.rejected_accum += ratio; P.n_rejected++;
                const double st = P.rejected_accum / (double)P.n_rejected;
                if (st > 0) P.sigma = st;
            } else if (inl > max_inl) {
                max_inl = inl;
                accepted.push_back(Hyp{hyps[j], (double)inl});
                if (inl > P.min_sample) {
                    // force inner RANSAC on the support of this hypothesis (arrsac.h:291-320); pending inner samples of the
                    // previous support are dropped and the random sampler rewound to where the schedule stands
                    if (si.valid) { g_random_rng = si.before[si_pos]; si.valid = false; }
                    inner = true; inner_its = 0;
                    rs_size = std::max(std::min(P.nonmin, (int)std::floor((float)max_inl / 2.0f)), P.min_sample);
                    universe.clear();
                    for (int i = 0; i < len; ++i) if (G.bit(hyps[j], i)) universe.push_back(i);
                    P.epsilon = ratio == 1.0 ? 0.9999 : ratio;
                    m_prime = (int)std::ceil(std::log(1.0 - P.conf) / std::log(1.0 - std::pow(P.epsilon, (double)P.min_sample)));
                    m_prime = std::min(P.max_cand, m_prime);
                }
            } else accepted.push_back(Hyp{hyps[j], (double)inl});
        }
        k++;
    }
    if (accepted.empty()) return 0;
    return k - k2 - 1;
}

// Arrsac::Estimate (arrsac.h:352-545).  Returns the pool id of the best hypothesis or -1.
int arrsac_estimate(ArrsacEngine& G, ArrsacState& P, int n)
{
    const int sub_block = (int)std::floor((float)P.block / 5.0), kill = (int)std::floor(7.0 * (float)sub_block / 12.0);
    std::vector<Hyp> H;
    const int len0 = std::min(n, P.block);
    int k = initial_hypothesis_set(G, P, len0, H);
    if (k == 0 || G.err) return -1;
    if (n <= P.block) {
        double best = 0.0; int bi = 0;
        for (size_t i = 0; i < H.size(); ++i) if (H[i].score > best) { best = H[i].score; bi = (int)i; }
        return H[bi].id;
    }
    int nh = (int)H.size();
    for (int i = P.block; i < n; ++i) {
        if ((i + 1) % P.block == 0) {
            std::sort(H.begin(), H.end(), hyp_greater);
            double max_inliers = H[0].score;
            P.epsilon = max_inliers / (double)(i + 1);
            if (P.epsilon == 1.0) P.epsilon = 0.9999;
            int tmax = (int)std::ceil(std::log(1 - P.conf) / std::log(1 - std::pow(P.epsilon, i + 1)));
            tmax = std::min(P.max_cand, tmax);
            if (tmax > k) {
                int k2 = 0;
                // M' - k new hypotheses from random minimal samples of ALL data, verified by SPRT on the first i+1 points
                const int want = tmax - k;
                std::vector<int> samp((size_t)want * MAX_SAMPLE, 0), sizes(want, 5);
                std::vector<CvRng> before(want + 1);
                for (int j = 0; j < want; ++j) { before[j] = g_random_rng; random_sample(g_random_rng, n, 5, &samp[(size_t)j * MAX_SAMPLE]); }
                before[want] = g_random_rng;
                std::vector<std::vector<int>> ids;
                G.generate(samp, sizes, ids);
                if (G.err) return -1;
                int j = 0;
                for (; j < tmax - k; ++j) {
                    g_random_rng = before[j + 1];
                    const std::vector<int>& est = ids[j];
                    if (est.empty()) { k2++; continue; }
                    P.verified_accum += (int)est.size();
                    const double A = sprt_threshold(P.sigma, P.epsilon, P.tm_ratio, P.verified_accum / (k + j + 1 - k2));
                    for (size_t m = 0; m < est.size(); ++m) {
                        double ratio; int inl;
                        const bool ok = G.sprt(est[m], i + 1, P.sigma, P.epsilon, A, &ratio, &inl);
                        if (!ok) {
                            P.rejected_accum += ratio; P.n_rejected++;
                            const double stt = P.rejected_accum / (double)P.n_rejected;
                            if (stt > 0) P.sigma = stt;
                        } else if (inl > (int)max_inliers) {
                            H.push_back(Hyp{est[m], (double)inl});
                            max_inliers = (double)inl;
                            P.epsilon = ratio == 1.0 ? 0.9999 : ratio;
                            tmax = (int)std::ceil(std::log(1 - P.conf) / std::log(1 - std::pow(P.epsilon, i + 1)));
                            tmax = std::min(P.max_cand, tmax);
                            if (tmax <= (k + j + 1)) break;                                // arrsac.h:484-485: leaves the model loop only
                        } else H.push_back(Hyp{est[m], (double)inl});
                    }
                    k++;
                }
                nh = (int)H.size();
            } else {
                const int n1 = std::max(1, (int)std::floor((float)k * std::pow(2.0, -1.0 * std::floor((float)(i + 1) / (float)P.block))));
                if (n1 < (int)H.size()) { H.resize(n1); nh = n1; }
            }
        } else if ((i + 1) % sub_block == 0) {
            std::sort(H.begin(), H.end(), hyp_greater);
            int j = nh - 1;
            for (; j > 0; j--) if (H[j - 1].score - H[j].score > kill) break;
            nh = nh - j;
            H.resize(nh);
        }
        if (nh == 1) break;
        for (size_t j = 0; j < H.size(); ++j) if (G.bit(H[j].id, i)) H[j].score += 1.0;     // Error(data[i], h) < error_thresh (arrsac.h:538-542)
    }
    return H[0].id;
}

}  // namespace

extern "C" int mpl_arrsac_set_seed(unsigned long long prosac_state, unsigned long long random_state)
{
    g_prosac_rng.state = prosac_state ? prosac_state : 0xffffffffull;
    g_random_rng.state = random_state ? random_state : 0xffffffffull;
    return MPL_OK;
}

// CvModelEstimator3::runARRSAC (modelest.cpp:197-341) with refineEssential = robustEssentialRefine.  See include/mplb200.h.
extern "C" int mpl_arrsac5pt(const double* p1, const double* p2, int n, double th, int refine, double* E, uint8_t* mask, int* n_inl,
                             int* n_hypotheses, int* n_batches)
{
    if (!p1 || !p2 || n < 5 || !E || !n_inl) { mpl_set_error("mpl_arrsac5pt: bad argument"); return MPL_ERR_BAD_ARG; }
    *n_inl = 0;
    if (n_hypotheses) *n_hypotheses = 0;
    if (n_batches) *n_batches = 0;
    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx(); cudaStream_t st = ctx->stream; MplBuf* B = ctx->bufs;
    if ((rc = B[MPL_BUF_R0].ensure((size_t)n * 32))) return rc;
    if ((rc = B[MPL_BUF_R1].ensure((size_t)n * 32))) return rc;
    if ((rc = B[MPL_BUF_R5].ensure((size_t)n + 8192))) return rc;
    double* d_p1 = (double*)B[MPL_BUF_R0].ptr; double* d_p2 = d_p1 + 2 * (size_t)n;
    void* d_corr = B[MPL_BUF_R1].ptr;
    MPL_CUDA_CHECK(cudaMemcpyAsync(d_p1, p1, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(d_p2, p2, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    if ((rc = mpl_launch_pack_corr(d_p1, d_p2, n, d_corr, st))) return rc;

    ArrsacEngine G;
    G.ctx = ctx; G.st = st; G.d_corr = (const double4*)d_corr; G.n = n; G.W = (n + 31) / 32; G.th2 = th * th; G.B = B;
    ArrsacState P;
    int best = -1;
    if (n == 5) {
        // modelest.cpp:222-265: one kernel run on the five points, best of its models by (count, error sum).  Served through
        // the same generate() (ValidModel is not applied there in the reference: the count rule below decides anyway).
        std::vector<int> samp(MAX_SAMPLE, 0), sizes(1, 5); for (int i = 0; i < 5; ++i) samp[i] = i;
        std::vector<std::vector<int>> ids;
        G.generate(samp, sizes, ids);
        if (G.err) return G.err;
        int bc = 0;
        for (int id : ids[0]) { int c = 0; for (int i = 0; i < 5; ++i) c += G.bit(id, i); if (c > std::max(bc, 4)) { bc = c; best = id; } }
    } else {
        best = arrsac_estimate(G, P, n);
        if (G.err) return G.err;
    }
    if (n_hypotheses) *n_hypotheses = G.n_hyp;
    if (n_batches) *n_batches = G.batches;
    if (best < 0) return MPL_OK;                                   // Estimate returned false
    // findInliers of the best model (err <= th^2, modelest.cpp:69-83) through the refit kernels with refine = 0
    char* hs = nullptr;
    if ((rc = mpl_stage_ensure(2 * sizeof(MplPairState) + sizeof(RefineOut) + (size_t)n + 256, (void**)&hs))) return rc;
    MplPairState* h_state = (MplPairState*)hs; MplPairState* h_out = h_state + 1;
    RefineOut* h_ref = (RefineOut*)(h_out + 1); uint8_t* h_mask = (uint8_t*)(h_ref + 1);
    char* misc = (char*)B[MPL_BUF_R5].ptr;
    MplPairState* d_state = (MplPairState*)misc; double* d_E = (double*)(misc + 256); RefineOut* d_ref = (RefineOut*)(misc + 512);
    uint8_t* d_mask = (uint8_t*)(misc + 1024);
    memset(h_state, 0, sizeof(*h_state));
    h_state->n = n; h_state->best_slot = 0; h_state->best_count = 0; h_state->errminsum = DBL_MAX;
    memcpy(h_state->E, &G.models[(size_t)best * 9], 72);
    MPL_CUDA_CHECK(cudaMemcpyAsync(d_state, h_state, sizeof(MplPairState), cudaMemcpyHostToDevice, st));
    if ((rc = mpl_launch_refit(d_corr, (size_t)n, d_state, 1, th, 0, d_mask, (size_t)n, nullptr, nullptr, nullptr, st))) return rc;
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_mask, d_mask, (size_t)n, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    int good = 0;
    for (int i = 0; i < n; ++i) good += h_mask[i];
    if (n > 5 && (((good < 50) && (n > 200)) || (good < 15))) return MPL_OK;             // modelest.cpp:274-277: result = false
    double Ebest[9];
    memcpy(Ebest, &G.models[(size_t)best * 9], 72);
    if (refine && n > 5 && good >= 50) {                                                   // :279-313 (fewer than 50 inliers: the model is returned as is)
        MPL_CUDA_CHECK(cudaMemcpyAsync(d_E, Ebest, 72, cudaMemcpyHostToDevice, st));
        robust_refine_kernel<<<1, kRT, 0, st>>>((const double4*)d_corr, d_mask, n, d_E, th / 50.0, d_ref);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        MPL_CUDA_CHECK(cudaMemcpyAsync(h_ref, d_ref, sizeof(RefineOut), cudaMemcpyDeviceToHost, st));
        MPL_CUDA_CHECK(cudaStreamSynchronize(st));
        // :315-321: goodCount_tmp is computed from the UNrefined model, so the ratio test always passes and E_refined is taken;
        // the mask stays the one of the unrefined model (mask.copyTo(mask_tmp) copies the other way)
        memcpy(Ebest, h_ref->E, 72);
    }
    memcpy(E, Ebest, 72);
    if (mask) memcpy(mask, h_mask, (size_t)n);
    *n_inl = good;
    return MPL_OK;
}
