// vfc.cu -- filterWithVFC: the vector-field-consensus post-filter of matchinglib::getMatches(..., VFCrefine = true)
// (second half).
//
// Reference (paths relative to matchinglib_poselib/source/matchinglib):
//   getMatches, VFCrefine arm                 source/matchers.cpp:721-733
//   filterWithVFC                             source/vfcMatches.cpp:49-86
//   class VFC (Ma et al., TIP 2014), SparseVFC -- the only method the reference selects (source/vfc.cpp:41):
//       setData :60-72, normalize :560-620, selectSubset :130-150, constructIntraKernel / InterKernel :528-558, initialize :507-526,
//       optimizeSparseVFC :95-128, getP :490-505, calculateTraceCKC :468-488, calculateC_SparseVFC :152-189, calculateV :394-445,
//       calculateSigmaSquare :383-392, calculateGamma :369-381
//   cv::solve(DECOMP_LU) on the 16 x 16 system: OpenCV's LUImpl (partial pivoting, eps = 10 FLT_EPSILON)
//
// Split: the O(n) set-up that fixes the problem -- match -> point pairs, normalize() with its sequential float sums, the
// rand()-driven choice of the (at most 16) control points and their 16 x 16 kernel -- runs on the host exactly as written
// (glibc's rand() restated, glibc_rand.h).  Everything that is O(n) PER EM ITERATION runs in ONE kernel launch of one CTA:
// the 16 x n kernel matrix U, and up to 50 iterations of E-step (posterior P), the 168 weighted sums U diag(P) U^T and
// U diag(P) Y, the 16 x 16 LU solve, V = U^T C, sigma^2, gamma, with the loop's exit tests (iterations, energy change rate,
// sigma^2) evaluated on the device.
//
// Parity: float tolerance, not bitwise -- the reference accumulates its float sums sequentially over the n matches, the device in
// a fixed tree, and expf differs in the last ulp; multiplications and additions are kept separate (no FMA contraction) and in
// the reference's order within a term.  A match whose posterior ends within ~1e-4 of the 0.75 threshold can flip;
// tests/test_vfc_gpu.py measures it against the reference's own vfc.cpp (oracle/_ref/libvfc_ref.so).
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "glibc_rand.h"
#include "../../include/mplb200.h"

#include <cfloat>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

namespace {

constexpr int kCtrl = 16;                      // _numCtrlPts (vfc.cpp:51)
constexpr int kVfcThreads = 1024;
constexpr int kSums = kCtrl * (kCtrl + 1) / 2 + 2 * kCtrl;     // 136 entries of U diag(P) U^T + 32 of U diag(P) Y
constexpr int kGroups = kVfcThreads / kSums;   // 6 groups of 168 threads share the n matches of a sum

struct VfcJob {
    const float2* X; const float2* Y;          // normalised left points, displacement vectors
    int n, m;                                  // matches, control points actually chosen (<= 16)
    float ctrl[kCtrl][2];
    float K[kCtrl][kCtrl];                     // intra-kernel of the control points (host)
    float* Ut;                                 // [n][16] inter-kernel, point-major
    float* P; float2* V;
    uint8_t* keep; int* out;                   // out[0] = kept, out[1] = iterations
};

__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }

// sum over the CTA, identical in every thread (fixed tree: warp butterflies, then the 32 warp sums in order)
__device__ float block_sum(float v, float* red)
{
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v = fadd(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kVfcThreads / 32; ++w) s = fadd(s, red[w]);
    return s;
}

// |a - b|^2 as the reference writes it: pow(float, int) promotes to double, the sum is narrowed to float (vfc.cpp:386, 494, 551)
__device__ __forceinline__ float sqdist(float ax, float ay, float bx, float by)
{
    const double dx = (double)__fsub_rn(ax, bx), dy = (double)__fsub_rn(ay, by);
    return (float)__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
}

__global__ void __launch_bounds__(kVfcThreads)
vfc_sparse_kernel(const VfcJob* __restrict__ job)
{
    __shared__ float red[32];
    __shared__ float part[kGroups][kSums];
    __shared__ float A[kCtrl][kCtrl], Bm[kCtrl][2], C[kCtrl][2], Ks[kCtrl][kCtrl], ctrl[kCtrl][2];
    __shared__ float s_sigma2, s_gamma, s_E, s_tecr;
    __shared__ int s_iter, s_go;
    const VfcJob& J = *job;
    const int tid = threadIdx.x, n = J.n, m = J.m;
    const float beta = 0.1f, lambda = 3.0f, theta = 0.75f, a_par = 10.0f, ecr = 1e-5f, minP = 1e-5f;      // vfc.cpp:43-50
    if (tid < kCtrl * kCtrl) Ks[tid / kCtrl][tid % kCtrl] = J.K[tid / kCtrl][tid % kCtrl];
    if (tid < kCtrl * 2) { ctrl[tid / 2][tid & 1] = J.ctrl[tid / 2][tid & 1]; C[tid / 2][tid & 1] = 0.f; }
    __syncthreads();
    // constructInterKernel(_ctrlPts, _X) (:545-558) and initialize() (:507-526): V = 0, P = 1, sigma^2 from them
    float sg = 0.f, sp = 0.f;
    for (int i = tid; i < n; i += kVfcThreads) {
        const float2 x = J.X[i], y = J.Y[i];
        for (int j = 0; j < kCtrl; ++j)
            J.Ut[(size_t)i * kCtrl + j] = j < m ? expf(fmul(-beta, sqdist(ctrl[j][0], ctrl[j][1], x.x, x.y))) : 0.f;
        J.V[i] = make_float2(0.f, 0.f); J.P[i] = 1.f;
        sg = fadd(sg, sqdist(y.x, y.y, 0.f, 0.f)); sp = fadd(sp, 1.f);
    }
    sg = block_sum(sg, red); sp = block_sum(sp, red);
    if (tid == 0) { s_sigma2 = sg / fmul(sp, 2.f); s_gamma = 0.9f; s_E = 1.f; s_tecr = 1.f; s_iter = 0; }
    __syncthreads();
    // which of the 168 sums this thread owns
    const int g = tid / kSums, o = tid - g * kSums;
    int oi = 0, oj = 0;                                       // o < 136: (oi, oj), oi <= oj; else row oi = (o - 136) / 2, column oj = (o - 136) & 1 of U diag(P) Y
    if (o < kSums - 2 * kCtrl) { int r = o; for (oi = 0; oi < kCtrl; ++oi) { if (r < kCtrl - oi) { oj = oi + r; break; } r -= kCtrl - oi; } }
    else { oi = (o - (kSums - 2 * kCtrl)) >> 1; oj = (o - (kSums - 2 * kCtrl)) & 1; }
    const bool is_y = o >= kSums - 2 * kCtrl;
    for (;;) {
        if (tid == 0) s_go = (s_iter < 50 && s_tecr > ecr && (double)s_sigma2 > 1e-8) ? 1 : 0;      // :106
        __syncthreads();
        if (!s_go) break;
        const float sigma2 = s_sigma2, gamma = s_gamma, E_old = s_E;
        // ---- E-step: getP (:490-505)
        const float temp2 = fmul(fmul(6.283185f, sigma2), (1.f - gamma)) / fmul(gamma, a_par);      // pow(x, 1.0f) == x
        float sumP = 0.f, E = 0.f;
        for (int i = tid; i < n; i += kVfcThreads) {
            const float2 y = J.Y[i], v = J.V[i];
            const float t = sqdist(y.x, y.y, v.x, v.y);
            const float temp1 = expf(-t / fmul(2.f, sigma2));
            const float p = temp1 / fadd(temp1, temp2);
            J.P[i] = fmaxf(minP, p);
            sumP = fadd(sumP, p); E = fadd(E, fmul(p, t));
        }
        sumP = block_sum(sumP, red); E = block_sum(E, red);
        E = E / fmul(2.f, sigma2);
        E = fadd(E, fmul(fmul(sumP, logf(sigma2)), 2.f) / 2.f);
        // ---- calculateTraceCKC (:468-488), energy change rate (:111-112)
        if (tid == 0) {
            float tr = 0.f;
            for (int i = 0; i < m; ++i) {
                float t1 = 0.f, t2 = 0.f;
                for (int j = 0; j < m; ++j) { t1 = fadd(t1, fmul(Ks[i][j], C[j][0])); t2 = fadd(t2, fmul(Ks[i][j], C[j][1])); }
                tr = fadd(tr, fadd(fmul(C[i][0], t1), fmul(C[i][1], t2)));
            }
            const float En = fadd(E, fmul(lambda / 2.f, tr));
            s_E = En; s_tecr = fabsf((En - E_old) / En);
        }
        // ---- M-step: calculateC_SparseVFC (:152-189).  Thread (g, o) sums its term over matches g, g + 6, ...
        float acc = 0.f;
        if (g < kGroups && oi < m && (is_y || oj < m))
            for (int k = g; k < n; k += kGroups) {
                const float pu = fmul(J.P[k], J.Ut[(size_t)k * kCtrl + oi]);
                const float w = is_y ? (oj ? J.Y[k].y : J.Y[k].x) : J.Ut[(size_t)k * kCtrl + oj];
                acc = fadd(acc, fmul(pu, w));
            }
        if (g < kGroups) part[g][o] = acc;
        __syncthreads();
        if (tid < kSums) {
            float s = 0.f;
            for (int gg = 0; gg < kGroups; ++gg) s = fadd(s, part[gg][tid]);
            if (!is_y) { s = fadd(s, fmul(fmul(lambda, sigma2), Ks[oi][oj])); A[oi][oj] = s; A[oj][oi] = s; }
            else Bm[oi][oj] = s;
        }
        __syncthreads();
        if (tid == 0) {                                        // cv::solve(K, Y, C, DECOMP_LU): LUImpl, float
            bool ok = true;
            for (int i = 0; i < m && ok; i++) {
                int k = i;
                for (int j = i + 1; j < m; j++) if (fabsf(A[j][i]) > fabsf(A[k][i])) k = j;
                if (fabsf(A[k][i]) < FLT_EPSILON * 10) { ok = false; break; }
                if (k != i) {
                    for (int j = i; j < m; j++) { const float t = A[i][j]; A[i][j] = A[k][j]; A[k][j] = t; }
                    for (int j = 0; j < 2; j++) { const float t = Bm[i][j]; Bm[i][j] = Bm[k][j]; Bm[k][j] = t; }
                }
                const float d = -1.f / A[i][i];
                for (int j = i + 1; j < m; j++) {
                    const float alpha = fmul(A[j][i], d);
                    for (int kk = i + 1; kk < m; kk++) A[j][kk] = fadd(A[j][kk], fmul(alpha, A[i][kk]));
                    for (int kk = 0; kk < 2; kk++) Bm[j][kk] = fadd(Bm[j][kk], fmul(alpha, Bm[i][kk]));
                }
            }
            if (ok)
                for (int i = m - 1; i >= 0; i--)
                    for (int j = 0; j < 2; j++) {
                        float s = Bm[i][j];
                        for (int k = i + 1; k < m; k++) s = __fsub_rn(s, fmul(A[i][k], Bm[k][j]));
                        Bm[i][j] = s / A[i][i];
                    }
            for (int i = 0; i < kCtrl; ++i) for (int j = 0; j < 2; ++j) C[i][j] = (ok && i < m) ? Bm[i][j] : 0.f;      // a singular system leaves cv::solve's output zero
        }
        __syncthreads();
        // ---- calculateV (:432-444), calculateSigmaSquare (:383-392), calculateGamma (:369-381)
        float sg2 = 0.f, sp2 = 0.f, cnt = 0.f;
        for (int i = tid; i < n; i += kVfcThreads) {
            float t1 = 0.f, t2 = 0.f;
            for (int j = 0; j < m; ++j) { const float u = J.Ut[(size_t)i * kCtrl + j]; t1 = fadd(t1, fmul(u, C[j][0])); t2 = fadd(t2, fmul(u, C[j][1])); }
            J.V[i] = make_float2(t1, t2);
            const float2 y = J.Y[i];
            const float p = J.P[i];
            sg2 = fadd(sg2, fmul(p, sqdist(y.x, y.y, t1, t2))); sp2 = fadd(sp2, p);
            cnt += p > theta ? 1.f : 0.f;
        }
        sg2 = block_sum(sg2, red); sp2 = block_sum(sp2, red); cnt = block_sum(cnt, red);
        if (tid == 0) {
            s_sigma2 = sg2 / fmul(sp2, 2.f);
            s_gamma = fmaxf(fminf(cnt / (float)n, 0.95f), 0.05f);
            s_iter = s_iter + 1;
        }
        __syncthreads();
    }
    // obtainCorrectMatch: the P > theta set of the last calculateGamma; all matches when no iteration ran (:66-68)
    const bool none = s_iter == 0;
    float kept = 0.f;
    for (int i = tid; i < n; i += kVfcThreads) { const bool k = none || J.P[i] > theta; J.keep[i] = k ? 1 : 0; kept += k ? 1.f : 0.f; }
    kept = block_sum(kept, red);
    if (tid == 0) { J.out[0] = (int)kept; J.out[1] = s_iter; }
}

thread_local GlibcRand t_vfc_rng;
thread_local bool t_vfc_seeded = false;

}  // namespace

extern "C" int mpl_vfc_set_seed(unsigned seed) { t_vfc_rng.seed(seed); t_vfc_seeded = true; return MPL_OK; }

// See include/mplb200.h.
extern "C" int mpl_vfc_filter(const void* kp1, int n1, size_t kp_stride1, const void* kp2, int n2, size_t kp_stride2, const mpl_dmatch* matches,
                              int n_matches, uint8_t* keep, int* n_keep, int* vfc_code, int* n_iterations)
{
    if (n_matches < 0 || n1 < 0 || n2 < 0 || !n_keep || !vfc_code || kp_stride1 < 8 || kp_stride2 < 8 ||
        (n_matches > 0 && (!matches || !keep || !kp1 || !kp2))) { mpl_set_error("mpl_vfc_filter: bad argument"); return MPL_ERR_BAD_ARG; }
    *n_keep = 0; *vfc_code = 0;
    if (n_iterations) *n_iterations = 0;
    const int n = n_matches;
    for (int i = 0; i < n; ++i)
        if (matches[i].queryIdx < 0 || matches[i].queryIdx >= n1 || matches[i].trainIdx < 0 || matches[i].trainIdx >= n2) {
            mpl_set_error("mpl_vfc_filter: match %d refers to keypoints (%d, %d) outside [0, %d) x [0, %d)", i, matches[i].queryIdx, matches[i].trainIdx, n1, n2);
            return MPL_ERR_BAD_ARG;
        }
    if (n > 0) memset(keep, 0, (size_t)n);
    if (n < 5) { *vfc_code = -1; return MPL_OK; }                                         // setData (:61-62) -> "Too less matches"
    // ---- host, exactly as written: point pairs (vfcMatches.cpp:56-61), normalize (:560-620)
    std::vector<float> lx(n), ly(n), rx(n), ry(n);
    for (int i = 0; i < n; ++i) {
        const float* a = (const float*)((const char*)kp1 + (size_t)matches[i].queryIdx * kp_stride1);
        const float* b = (const float*)((const char*)kp2 + (size_t)matches[i].trainIdx * kp_stride2);
        lx[i] = a[0]; ly[i] = a[1]; rx[i] = b[0]; ry[i] = b[1];
    }
    float x1 = 0, y1 = 0, x2 = 0, y2 = 0;
    for (int i = 0; i < n; i++) { x1 += lx[i]; y1 += ly[i]; x2 += rx[i]; y2 += ry[i]; }
    x1 /= n; y1 /= n; x2 /= n; y2 /= n;
    double s1 = 0, s2 = 0;
    for (int i = 0; i < n; i++) {
        lx[i] -= x1; ly[i] -= y1; rx[i] -= x2; ry[i] -= y2;
    }
    for (int i = 0; i < n; i++) { s1 += (double)lx[i] * (double)lx[i]; s1 += (double)ly[i] * (double)ly[i]; s2 += (double)rx[i] * (double)rx[i]; s2 += (double)ry[i] * (double)ry[i]; }
    s1 /= n; s2 /= n; s1 = std::sqrt(s1); s2 = std::sqrt(s2);
    auto all_kept = [&]() { memset(keep, 1, (size_t)n); *n_keep = n; *vfc_code = 0; return MPL_OK; };   // optimize() returned early: _matchIdx is everything
    if ((float)s1 < 0.1 || (float)s2 < 0.1) return all_kept();                           // MIN_STANDARD_DEVIATION (:605-606)
    std::vector<float> X(2 * (size_t)n), Y(2 * (size_t)n);
    for (int i = 0; i < n; i++) {
        lx[i] /= (float)s1; ly[i] /= (float)s1; rx[i] /= (float)s2; ry[i] /= (float)s2;
        X[2 * i] = lx[i]; X[2 * i + 1] = ly[i]; Y[2 * i] = rx[i] - lx[i]; Y[2 * i + 1] = ry[i] - ly[i];
    }
    // ---- selectSubset (:130-150) with rand(), constructIntraKernel (:528-543)
    if (!t_vfc_seeded) { t_vfc_rng.seed(1); t_vfc_seeded = true; }
    VfcJob J;
    memset(&J, 0, sizeof(J));
    const int want = n < kCtrl ? n : kCtrl;
    int cnt = 0, iter = 0;
    while (cnt < want && iter < want * 3) {
        const int idx = t_vfc_rng.next() % n;
        float dist = (float)0x3f3f3f3f;
        for (int i = 0; i < cnt; i++) { const float tmp = std::fabs(J.ctrl[i][0] - X[2 * idx]) + std::fabs(J.ctrl[i][1] - X[2 * idx + 1]); dist = std::min(tmp, dist); }
        if (dist > 1e-3) { J.ctrl[cnt][0] = X[2 * idx]; J.ctrl[cnt][1] = X[2 * idx + 1]; cnt++; }
        iter++;
    }
    J.n = n; J.m = cnt;
    for (int i = 0; i < cnt; i++) {
        J.K[i][i] = 1;
        for (int j = i + 1; j < cnt; j++) {
            const double dx = (double)(J.ctrl[i][0] - J.ctrl[j][0]), dy = (double)(J.ctrl[i][1] - J.ctrl[j][1]);
            float t = (float)(dx * dx + dy * dy);
            t *= -0.1f;
            t = expf(t);
            J.K[i][j] = t; J.K[j][i] = t;
        }
    }
    // ---- device: U and the EM iterations
    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx(); cudaStream_t st = ctx->stream; MplBuf* B = ctx->bufs;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t nn = (size_t)n;
    const size_t o_job = 0, o_x = al(sizeof(VfcJob)), o_y = o_x + al(nn * 8), o_u = o_y + al(nn * 8), o_p = o_u + al(nn * kCtrl * 4), o_v = o_p + al(nn * 4),
                 o_keep = o_v + al(nn * 8), o_out = o_keep + al(nn), total = o_out + 256;
    if ((rc = B[MPL_BUF_RAW].ensure(total))) return rc;
    char* D = (char*)B[MPL_BUF_RAW].ptr;
    char* hs = nullptr;
    if ((rc = mpl_stage_ensure(o_u + al(nn) + 256, (void**)&hs))) return rc;
    J.X = (const float2*)(D + o_x); J.Y = (const float2*)(D + o_y); J.Ut = (float*)(D + o_u); J.P = (float*)(D + o_p); J.V = (float2*)(D + o_v);
    J.keep = (uint8_t*)(D + o_keep); J.out = (int*)(D + o_out);
    memcpy(hs + o_job, &J, sizeof(J)); memcpy(hs + o_x, X.data(), nn * 8); memcpy(hs + o_y, Y.data(), nn * 8);
    MPL_CUDA_CHECK(cudaMemcpyAsync(D, hs, o_u, cudaMemcpyHostToDevice, st));
    vfc_sparse_kernel<<<1, kVfcThreads, 0, st>>>((const VfcJob*)(D + o_job));
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    char* h_keep = hs + o_u; int* h_out = (int*)(h_keep + al(nn));
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_keep, D + o_keep, nn, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_out, D + o_out, 8, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    memcpy(keep, h_keep, nn);
    *n_keep = h_out[0];
    if (n_iterations) *n_iterations = h_out[1];
    if ((double)*n_keep / (double)n < 0.1) *vfc_code = -2;                               // vfcMatches.cpp:82-83 "Maybe VFC failed"
    return MPL_OK;
}
