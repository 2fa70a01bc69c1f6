// recover_pose.cu -- poselib::getPoseTriangPts: decompose E, triangulate with the four (R, +-t) combinations, cheirality
// + distance masks, pick the combination with most valid points.
//
// Replaces (reference paths relative to matchinglib_poselib/source/poselib/source):
//   getPoseTriangPts           pose_estim.cpp:913-946
//   recoverPose                five-point-nister/five-point.cpp:150-338
//   decomposeEssentialMat      five-point-nister/five-point.cpp:340-352
//   getTfromTransEssential     pose_helper.cpp:422-433
//   cv::triangulatePoints      OpenCV 4.2.0 calib3d/src/triangulate.cpp (per point: 4x4 DLT matrix, SVD, last row of Vt) --
//                              third party, not under /root/reference; restated from its published algorithm.
//
// The 3x3 decomposition of E runs on the host (one SVD); the per-point work -- up to four 4x4 SVDs, the two depth
// tests, the distance test, the AND with the caller's mask and the four counts -- is one thread per correspondence
// on the device (SURVEY.md section 8(f) N4 brought forward: the step right after the RANSAC path).
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "small_svd.cuh"
#include "../../include/mplb200.h"

#include <math.h>
#include <string.h>

namespace {

struct PoseCombos { double P[4][12]; int n_combos; };   // row-major 3x4 second-camera matrices; first camera is [I | 0]

// thread = one correspondence: for every combination, DLT triangulation (triangulate.cpp: rows x*P[2]-P[0], y*P[2]-P[1]
// for both cameras), masks Z*W > 0 in camera 1, (P*X).z * W > 0 in camera 2, Z/W < dist (five-point.cpp:201-210), AND
// with the caller's mask (:284-301), counts by warp ballot + one atomic per warp.
__global__ void __launch_bounds__(128)
triangulate_cheirality_kernel(const double* __restrict__ p1, const double* __restrict__ p2, int n, PoseCombos C, double dist,
                              const uint8_t* __restrict__ mask_in, double* __restrict__ Q /* [4][n][3] */,
                              uint8_t* __restrict__ masks /* [4][n] */, int* __restrict__ counts /* [4] */)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < n;
    double x1 = 0, y1 = 0, x2 = 0, y2 = 0;
    if (live) { x1 = p1[2 * i]; y1 = p1[2 * i + 1]; x2 = p2[2 * i]; y2 = p2[2 * i + 1]; }
    const bool keep = live && (mask_in == nullptr || mask_in[i] != 0);
    for (int c = 0; c < C.n_combos; ++c) {
        const double* P = C.P[c];
        double A[4][4], X[4];
        // camera 1 = [I | 0]: x*P0[2] - P0[0] = (-1, 0, x, 0), y*P0[2] - P0[1] = (0, -1, y, 0)
        A[0][0] = -1.0; A[0][1] = 0.0; A[0][2] = x1; A[0][3] = 0.0;
        A[1][0] = 0.0; A[1][1] = -1.0; A[1][2] = y1; A[1][3] = 0.0;
        for (int k = 0; k < 4; ++k) { A[2][k] = x2 * P[8 + k] - P[k]; A[3][k] = y2 * P[8 + k] - P[4 + k]; }
        jacobi_null_vector<4>(A, X);
        const double w = X[3];
        const double z2 = P[8] * X[0] + P[9] * X[1] + P[10] * X[2] + P[11] * X[3];
        const double qx = X[0] / w, qy = X[1] / w, qz = X[2] / w;
        const bool ok = keep && (X[2] * w > 0.0) && (z2 * w > 0.0) && (qz < dist);
        if (live) {
            double* q = Q + ((size_t)c * n + i) * 3;
            q[0] = qx; q[1] = qy; q[2] = qz;
            masks[(size_t)c * n + i] = ok ? (mask_in ? mask_in[i] : (uint8_t)1) : (uint8_t)0;   // bitwise_and(mask, 255-mask) (:284-301)
        }
        const unsigned b = __ballot_sync(0xffffffffu, ok);
        if ((threadIdx.x & 31) == 0 && b) atomicAdd(&counts[c], __popc(b));
    }
}

void mat3_mul(const double A[9], const double B[9], double C[9])
{
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
double det3(const double M[9])
{
    return M[0] * (M[4] * M[8] - M[5] * M[7]) - M[1] * (M[3] * M[8] - M[5] * M[6]) + M[2] * (M[3] * M[7] - M[4] * M[6]);
}

// decomposeEssentialMat (five-point.cpp:340-352): E = U D Vt, singular values descending (cv::SVD order); flip U / Vt to
// det +1; R1 = U W Vt, R2 = U W^T Vt, t = U[:, 2].
void decompose_essential(const double E[9], double R1[9], double R2[9], double t[3])
{
    double A[3][3], x[3], V[3][3], w[3];
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) A[i][j] = E[3 * i + j];
    jacobi_null_vector<3>(A, x, V, w);
    int ord[3] = {0, 1, 2};                       // sort by singular value, descending
    for (int a = 0; a < 2; ++a) for (int b = a + 1; b < 3; ++b) if (w[ord[b]] > w[ord[a]]) { int s = ord[a]; ord[a] = ord[b]; ord[b] = s; }
    double U[9], Vt[9];
    for (int k = 0; k < 3; ++k) {
        const int j = ord[k];
        for (int i = 0; i < 3; ++i) Vt[3 * k + i] = V[j][i];
        if (k == 2) break;
        double u[3];                                                         // u_k = E v_k / sigma_k
        for (int i = 0; i < 3; ++i) u[i] = E[3 * i] * V[j][0] + E[3 * i + 1] * V[j][1] + E[3 * i + 2] * V[j][2];
        const double nu = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]);
        for (int i = 0; i < 3; ++i) U[3 * i + k] = nu > 0 ? u[i] / nu : (i == k ? 1.0 : 0.0);
    }
    {   // E has rank 2: the third left singular vector is u1 x u2 (its sign is the SVD's free choice; the set
        // {R1, R2} x {t, -t} does not depend on it)
        const double a[3] = {U[0], U[3], U[6]}, b[3] = {U[1], U[4], U[7]};
        const double c[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
        const double nc = sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]);
        for (int i = 0; i < 3; ++i) U[3 * i + 2] = c[i] / nc;
    }
    if (det3(U) < 0) for (int i = 0; i < 9; ++i) U[i] = -U[i];
    if (det3(Vt) < 0) for (int i = 0; i < 9; ++i) Vt[i] = -Vt[i];
    const double W[9] = {0, 1, 0, -1, 0, 0, 0, 0, 1}, Wt[9] = {0, -1, 0, 1, 0, 0, 0, 0, 1};
    double T[9];
    mat3_mul(U, W, T); mat3_mul(T, Vt, R1);
    mat3_mul(U, Wt, T); mat3_mul(T, Vt, R2);
    t[0] = U[2]; t[1] = U[5]; t[2] = U[8];
}

void set_P(double P[12], const double R[9], const double t[3], double sign)
{
    for (int i = 0; i < 3; ++i) { P[4 * i] = R[3 * i]; P[4 * i + 1] = R[3 * i + 1]; P[4 * i + 2] = R[3 * i + 2]; P[4 * i + 3] = sign * t[i]; }
}

}  // namespace

// See include/mplb200.h.  Returns the number of valid 3-D points (>= 0) like the reference, or a negative MPL_ERR_*.
extern "C" int mpl_get_pose_triang_pts_api(const double* E, const double* p1, const double* p2, int n, uint8_t* mask, int mask_given,
                                           double dist, int translatE, double* R, double* t, double* Q)
{
    if (!R || !t || !Q) return -1;                                           // pose_estim.cpp:925-926
    if (!E || n < 0 || (n > 0 && (!p1 || !p2))) { mpl_set_error("getPoseTriangPts: bad argument"); return MPL_ERR_BAD_ARG; }
    double R1[9], R2[9], tt[3];
    PoseCombos C;
    if (translatE) {                                                         // getTfromTransEssential, pose_helper.cpp:422-433
        auto near_zero = [](double d) { return d < 1e-3 && d > -1e-3; };
        if (!(near_zero(E[1] / E[3] + 1.0) && near_zero(E[2] / E[6] + 1.0) && near_zero(E[5] / E[7] + 1.0))) {
            mpl_set_error("CV_Assert: not a translational essential matrix"); return MPL_ERR_ASSERT;
        }
        tt[0] = E[5]; tt[1] = E[6]; tt[2] = E[1];
        const double nt = sqrt(tt[0] * tt[0] + tt[1] * tt[1] + tt[2] * tt[2]);
        if (fabs(nt - 1.0) > 1e-3) for (int i = 0; i < 3; ++i) tt[i] /= nt;
        const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        memcpy(R1, I, sizeof(I)); memcpy(R2, I, sizeof(I));
        set_P(C.P[0], R1, tt, 1.0); set_P(C.P[1], R1, tt, -1.0);             // combinations 1 and 3 only (five-point.cpp:181-194)
        C.n_combos = 2;
    } else {
        decompose_essential(E, R1, R2, tt);
        set_P(C.P[0], R1, tt, 1.0); set_P(C.P[1], R2, tt, 1.0); set_P(C.P[2], R1, tt, -1.0); set_P(C.P[3], R2, tt, -1.0);
        C.n_combos = 4;
    }
    int good[4] = {0, 0, 0, 0};
    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx(); cudaStream_t st = ctx->stream; MplBuf* B = ctx->bufs;
    char* base = nullptr; double* d_Q = nullptr; uint8_t* d_masks = nullptr;
    if (n > 0) {
        const size_t np = ((size_t)n + 255) & ~(size_t)255;
        const size_t bytes = 256 + np * 32 + np + 4 * np * 24 + 4 * np;
        if ((rc = B[MPL_BUF_R5].ensure(bytes))) return rc;
        base = (char*)B[MPL_BUF_R5].ptr;
        int* d_counts = (int*)base;
        double* d_p1 = (double*)(base + 256); double* d_p2 = d_p1 + 2 * np;
        uint8_t* d_min = (uint8_t*)(d_p2 + 2 * np);
        d_Q = (double*)(d_min + np); d_masks = (uint8_t*)(d_Q + 12 * np);
        MPL_CUDA_CHECK(cudaMemsetAsync(d_counts, 0, 16, st));
        MPL_CUDA_CHECK(cudaMemcpyAsync(d_p1, p1, (size_t)n * 16, cudaMemcpyHostToDevice, st));
        MPL_CUDA_CHECK(cudaMemcpyAsync(d_p2, p2, (size_t)n * 16, cudaMemcpyHostToDevice, st));
        const bool use_mask = mask && mask_given;
        if (use_mask) MPL_CUDA_CHECK(cudaMemcpyAsync(d_min, mask, (size_t)n, cudaMemcpyHostToDevice, st));
        triangulate_cheirality_kernel<<<(n + 127) / 128, 128, 0, st>>>(d_p1, d_p2, n, C, dist, use_mask ? d_min : nullptr, d_Q, d_masks, d_counts);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        int cnt[4];
        MPL_CUDA_CHECK(cudaMemcpyAsync(cnt, d_counts, 16, cudaMemcpyDeviceToHost, st));
        MPL_CUDA_CHECK(cudaStreamSynchronize(st));
        if (translatE) { good[0] = cnt[0]; good[2] = cnt[1]; }
        else for (int c = 0; c < 4; ++c) good[c] = cnt[c];
    }
    // selection, five-point.cpp:299-336 (good1..good4 = combinations (R1,t), (R2,t), (R1,-t), (R2,-t))
    const int g1 = good[0], g2 = good[1], g3 = good[2], g4 = good[3];
    int pick, ret;
    if (g1 >= g2 && g1 >= g3 && g1 >= g4) { pick = 0; ret = g1; }
    else if (g2 && g2 >= g1 && g2 >= g3 && g2 >= g4) { pick = 1; ret = g2; }
    else if (g3 >= g1 && g3 >= g2 && g3 >= g4) { pick = 2; ret = g3; }
    else { pick = g4 ? 3 : -1; ret = g4; }
    if (pick < 0) {
        const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
        memcpy(R, I, sizeof(I)); t[0] = t[1] = t[2] = 0.0;
        if (mask) memset(mask, 0, (size_t)n);
        memset(Q, 0, (size_t)n * 24);
        return ret;
    }
    memcpy(R, (pick & 1) ? R2 : R1, 72);
    const double sg = (pick >= 2) ? -1.0 : 1.0;
    for (int i = 0; i < 3; ++i) t[i] = sg * tt[i];
    if (n > 0) {
        const int slot = translatE ? (pick == 0 ? 0 : 1) : pick;
        MPL_CUDA_CHECK(cudaMemcpyAsync(Q, d_Q + (size_t)slot * n * 3, (size_t)n * 24, cudaMemcpyDeviceToHost, st));
        if (mask) MPL_CUDA_CHECK(cudaMemcpyAsync(mask, d_masks + (size_t)slot * n, (size_t)n, cudaMemcpyDeviceToHost, st));
        MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    }
    return ret;
}
