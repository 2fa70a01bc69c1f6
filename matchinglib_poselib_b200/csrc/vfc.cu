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
// the 16 x n kernel matrix U, and up to 50 iterations of E-step (posterior P; per match, spread over the CTA), the 168
// weighted sums U diag(P) U^T and U diag(P) Y plus the four scalar sums (172 threads, each one sum, sequential in match order),
// the 16 x 16 LU solve, V = U^T C, sigma^2, gamma, with the loop's exit tests evaluated on the device.
//
// Parity: the kept set is the reference's for the same rand() seed.  The EM iteration is numerically chaotic (sigma^2 settles near
// 1e-5, the energy keeps jumping, all 50 iterations run -- also in the reference), so "close" arithmetic gives kept sets that
// differ in up to 2 % of the matches; the device therefore does the reference's ARITHMETIC: float sums sequential in match order
// (one thread per sum), multiplications and additions separate and in the written order (no FMA contraction), pow(x, 2) in
// double, expf / logf as the once-rounded double results glibc returns.  What can still differ: glibc's expf is not correctly
// rounded for about one argument in 10^8.  tests/test_vfc_gpu.py compares with the reference's own vfc.cpp
// (oracle/_ref/libvfc_ref.so, built with -ffp-contract=off like an x86-64 baseline build).
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

struct VfcJob {
    const float2* X; const float2* Y;          // normalised left points, displacement vectors
    int n, m;                                  // matches, control points actually chosen (<= 16)
    float ctrl[kCtrl][2];
    float K[kCtrl][kCtrl];                     // intra-kernel of the control points (host)
    float* U;                                  // [16][n] inter-kernel, row j = control point j (the reference's _U)
    float* P; float* Praw; float* T; float2* V;
    uint8_t* keep; int* out;                   // out[0] = kept, out[1] = iterations
    unsigned long long exp_tab[32];            // glibc's __exp2f_data.tab: bits of 2^(i/32) minus (i << 47), filled on the host
};

__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
// expf as glibc computes it (sysdeps/ieee754/flt-32/e_expf.c, the algorithm of ARM's optimized routines): x * 32 / ln 2 = k + r,
// 2^(k/32) from a 32-entry table, a cubic in r, all in double, ONE rounding to float at the end.  That last value is not always
// the correctly rounded one (6 of 10^4 arguments differ from (float)exp((double)x)), and the EM iteration amplifies single ulps,
// so the algorithm is restated; against the host's expf it agreed on 2 * 10^7 of 2 * 10^7 random arguments in [-60, 0].
// Arguments here are never positive; below -103.97 glibc returns 0.
__device__ __forceinline__ float expf_glibc(float x, const unsigned long long* __restrict__ tab)
{
    if (!(x >= -0x1.9fe368p6f)) return x != x ? x : 0.0f;
    const double N = 32.0, InvLn2N = 0x1.71547652b82fep+0 * N, SHIFT = 0x1.8p+52;
    const double C0 = 0x1.c6af84b912394p-5 / N / N / N, C1 = 0x1.ebfce50fac4f3p-3 / N / N, C2 = 0x1.62e42ff0c52d6p-1 / N;
    const double z = __dmul_rn(InvLn2N, (double)x);
    double kd = __dadd_rn(z, SHIFT);
    const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd = __dsub_rn(kd, SHIFT);
    const double r = __dsub_rn(z, kd);
    const double sc = __longlong_as_double((long long)(tab[ki & 31] + (ki << 47)));
    const double zz = __fma_rn(C0, r, C1), r2 = __dmul_rn(r, r);
    double y = __fma_rn(C2, r, 1.0);
    y = __fma_rn(zz, r2, y);
    return (float)__dmul_rn(y, sc);
}
__device__ __forceinline__ float logf_like_glibc(float x) { return (float)log((double)x); }     // only feeds the energy change rate

// |a - b|^2 as the reference writes it: pow(float, int) promotes to double, the sum is narrowed to float (vfc.cpp:386, 494, 551)
__device__ __forceinline__ float sqdist(float ax, float ay, float bx, float by)
{
    const double dx = (double)__fsub_rn(ax, bx), dy = (double)__fsub_rn(ay, by);
    return (float)__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
}

// The reference's float sums run sequentially over the n matches; float addition is not associative and the EM iteration
// amplifies every rounding difference (sigma^2 settles near 1e-5 and the energy keeps jumping: all 50 iterations run), so the
// sums here are sequential too -- one THREAD per sum, 172 sums side by side, fed through shared-memory tiles that the whole CTA
// loads -- and only the per-match work (posteriors, V, the kernel matrix) is spread over the CTA.
// acc += term(0) + term(1) + ... strictly in order, with the (independent) terms of eight elements formed before their adds so
// that only the add chain is serial
template <class F>
__device__ __forceinline__ float seq_sum(float acc, int len, F term)
{
    int k = 0;
    for (; k + 8 <= len; k += 8) {
        const float x0 = term(k), x1 = term(k + 1), x2 = term(k + 2), x3 = term(k + 3), x4 = term(k + 4), x5 = term(k + 5), x6 = term(k + 6), x7 = term(k + 7);
        acc = fadd(acc, x0); acc = fadd(acc, x1); acc = fadd(acc, x2); acc = fadd(acc, x3); acc = fadd(acc, x4); acc = fadd(acc, x5); acc = fadd(acc, x6); acc = fadd(acc, x7);
    }
    for (; k < len; ++k) acc = fadd(acc, term(k));
    return acc;
}

// the same, with the operands of four consecutive matches fetched by one 16-byte shared-memory load per array (the sums are
// bound by shared-memory wavefronts: six warps read the same tiles)
template <class F4, class F1>
__device__ __forceinline__ float seq_sum4(float acc, int len, F4 term4, F1 term1)
{
    int k = 0;
    if (len >= 8) {
        float4 a = term4(0), b = term4(4);                     // the terms of the NEXT eight matches are formed while the adds of these eight queue up
        for (; k + 16 <= len; k += 8) {
            const float4 na = term4(k + 8), nb = term4(k + 12);
            acc = fadd(acc, a.x); acc = fadd(acc, a.y); acc = fadd(acc, a.z); acc = fadd(acc, a.w);
            acc = fadd(acc, b.x); acc = fadd(acc, b.y); acc = fadd(acc, b.z); acc = fadd(acc, b.w);
            a = na; b = nb;
        }
        acc = fadd(acc, a.x); acc = fadd(acc, a.y); acc = fadd(acc, a.z); acc = fadd(acc, a.w);
        acc = fadd(acc, b.x); acc = fadd(acc, b.y); acc = fadd(acc, b.z); acc = fadd(acc, b.w);
        k += 8;
    }
    for (; k < len; ++k) acc = fadd(acc, term1(k));
    return acc;
}
__device__ __forceinline__ float4 ld4(const float* p, int k) { return *reinterpret_cast<const float4*>(p + k); }
__device__ __forceinline__ float4 mul3(float4 p, float4 u, float4 w) { return make_float4(fmul(fmul(p.x, u.x), w.x), fmul(fmul(p.y, u.y), w.y), fmul(fmul(p.z, u.z), w.z), fmul(fmul(p.w, u.w), w.w)); }
__device__ __forceinline__ float4 mul2(float4 p, float4 t) { return make_float4(fmul(p.x, t.x), fmul(p.y, t.y), fmul(p.z, t.z), fmul(p.w, t.w)); }

constexpr int kSig = 960;                                 // matches per tile of the two sigma^2 sums (960 loader threads)
constexpr int kTile = 256, kTileLd = kTile + 4;          // row stride 260: 16-byte aligned rows, a warp's 16 rows x 16 bytes = two wavefronts

__global__ void __launch_bounds__(kVfcThreads)
vfc_sparse_kernel(const VfcJob* __restrict__ job)
{
    __shared__ __align__(16) float sU[2][kCtrl][kTileLd], sP[2][kTile], sA[2][kTile], sB[2][kTile], sC[2][kTile], sD[2][kTile];     // two tile buffers: warps 8.. load tile t + 1 while warps 0..7 sum tile t
    __shared__ float A[kCtrl][kCtrl], Bm[kCtrl][2], C[kCtrl][2], Ks[kCtrl][kCtrl], ctrl[kCtrl][2];
    __shared__ float s_sigma2, s_gamma, s_E, s_tecr, s_sumP, s_Esum, s_trace, s_sg, s_sp;
    __shared__ int s_iter, s_go, s_cnt, s_ok;
    __shared__ long long s_clk[5];                  // thread 0's clocks per phase: posteriors, M-step sums, energy + solve, V, sigma^2 sums
    __shared__ unsigned long long tab[32];
    const VfcJob& J = *job;
    const int tid = threadIdx.x, n = J.n, m = J.m;
    const float beta = 0.1f, lambda = 3.0f, theta = 0.75f, a_par = 10.0f, ecr = 1e-5f, minP = 1e-5f;      // vfc.cpp:43-50
    if (tid < kCtrl * kCtrl) Ks[tid / kCtrl][tid % kCtrl] = J.K[tid / kCtrl][tid % kCtrl];
    if (tid < kCtrl * 2) { ctrl[tid / 2][tid & 1] = J.ctrl[tid / 2][tid & 1]; C[tid / 2][tid & 1] = 0.f; }
    if (tid < 32) tab[tid] = J.exp_tab[tid];
    __syncthreads();
    // constructInterKernel(_ctrlPts, _X) (:545-558) and initialize() (:507-526): V = 0, P = 1, sigma^2 from them
    for (int i = tid; i < n; i += kVfcThreads) {
        const float2 x = J.X[i], y = J.Y[i];
        for (int j = 0; j < m; ++j) J.U[(size_t)j * n + i] = expf_glibc(fmul(-beta, sqdist(ctrl[j][0], ctrl[j][1], x.x, x.y)), tab);
        J.V[i] = make_float2(0.f, 0.f); J.P[i] = 1.f;
        J.T[i] = sqdist(y.x, y.y, 0.f, 0.f);
    }
    if (tid == 0) { s_gamma = 0.9f; s_E = 1.f; s_tecr = 1.f; s_iter = 0; for (int k = 0; k < 5; ++k) s_clk[k] = 0; }
    __syncthreads();
    // which of the 168 M-step sums this thread owns: tid < 136: (oi, oj), oi <= oj; tid < 168: row oi, column oj of U diag(P) Y
    int oi = 0, oj = 0;
    if (tid < kSums - 2 * kCtrl) { int r = tid; for (oi = 0; oi < kCtrl; ++oi) { if (r < kCtrl - oi) { oj = oi + r; break; } r -= kCtrl - oi; } }
    else if (tid < kSums) { oi = (tid - (kSums - 2 * kCtrl)) >> 1; oj = (tid - (kSums - 2 * kCtrl)) & 1; }
    const bool is_y = tid >= kSums - 2 * kCtrl;
    const bool live = tid < kSums && oi < m && (is_y || oj < m);
    bool first = true;                                         // the first pass only computes sigma^2 of initialize()
    for (;;) {
        if (!first) {
            if (tid == 0) { s_go = (s_iter < 50 && s_tecr > ecr && (double)s_sigma2 > 1e-8) ? 1 : 0; s_cnt = 0; }      // :106
            __syncthreads();
            if (!s_go) break;
        }
        const float sigma2 = first ? 1.f : s_sigma2, gamma = s_gamma, E_old = s_E;
        long long c0 = clock64(), c1;
        auto lap = [&](int k) { c1 = clock64(); if (tid == 0) s_clk[k] += c1 - c0; c0 = c1; };
        if (!first) {
            // ---- E-step, per match: getP (:490-505)
            const float temp2 = fmul(fmul(6.283185f, sigma2), (1.f - gamma)) / fmul(gamma, a_par);      // pow(x, 1.0f) == x
            for (int i = tid; i < n; i += kVfcThreads) {
                const float2 y = J.Y[i], v = J.V[i];
                const float t = sqdist(y.x, y.y, v.x, v.y);
                const float temp1 = expf_glibc(-t / fmul(2.f, sigma2), tab);
                const float p = temp1 / fadd(temp1, temp2);
                J.Praw[i] = p; J.P[i] = fmaxf(minP, p); J.T[i] = t;
            }
            __syncthreads();
            lap(0);
            // ---- the sequential sums of this half: 168 of calculateC_SparseVFC (:152-178), sum p and sum p t of getP, calculateTraceCKC
            float acc = 0.f;
            auto load_tile = [&](int buf, int t0, int first_thread, int n_threads) {     // every global load is issued before the first shared store
                const int len = min(kTile, n - t0), lt = tid - first_thread;
                float v[6]; int idx[6];
#pragma unroll
                for (int q = 0; q < 6; ++q) {
                    const int e = lt + q * n_threads, r = e / kTile, c = e - r * kTile;
                    const bool ok = e < m * kTile && c < len;
                    idx[q] = ok ? r * kTileLd + c : -1;
                    v[q] = ok ? J.U[(size_t)r * n + t0 + c] : 0.f;
                }
                float p0 = 0.f, p1 = 0.f; float2 y = make_float2(0.f, 0.f);
                const int c = lt & (kTile - 1);
                const bool okA = lt < kTile && c < len, okB = lt >= kTile && lt < 2 * kTile && c < len;
                if (okA) { p0 = J.P[t0 + c]; y = J.Y[t0 + c]; }
                if (okB) { p0 = J.Praw[t0 + c]; p1 = J.T[t0 + c]; }
#pragma unroll
                for (int q = 0; q < 6; ++q) if (idx[q] >= 0) (&sU[buf][0][0])[idx[q]] = v[q];
                if (okA) { sP[buf][c] = p0; sA[buf][c] = y.x; sB[buf][c] = y.y; }
                if (okB) { sC[buf][c] = p0; sD[buf][c] = p1; }
            };
            load_tile(0, 0, 0, kVfcThreads);
            __syncthreads();
            int buf = 0;
            for (int t0 = 0; t0 < n; t0 += kTile, buf ^= 1) {
                const int len = min(kTile, n - t0);
                if (tid >= 256) { if (t0 + kTile < n) load_tile(buf ^ 1, t0 + kTile, 256, kVfcThreads - 256); }
                else if (live) {
                    const float* ui = sU[buf][oi];
                    const float* w = is_y ? (oj ? sB[buf] : sA[buf]) : sU[buf][oj];
                    const float* pp = sP[buf];
                    acc = seq_sum4(acc, len, [&](int k) { return mul3(ld4(pp, k), ld4(ui, k), ld4(w, k)); }, [&](int k) { return fmul(fmul(pp[k], ui[k]), w[k]); });
                } else if (tid == 192) {
                    const float* c = sC[buf];
                    acc = seq_sum4(acc, len, [&](int k) { return ld4(c, k); }, [&](int k) { return c[k]; });
                } else if (tid == 224) {
                    const float* c = sC[buf]; const float* d = sD[buf];
                    acc = seq_sum4(acc, len, [&](int k) { return mul2(ld4(c, k), ld4(d, k)); }, [&](int k) { return fmul(c[k], d[k]); });
                }
                __syncthreads();
            }
            lap(1);
            if (tid < kSums) {
                if (!is_y) { if (live) { acc = fadd(acc, fmul(fmul(lambda, sigma2), Ks[oi][oj])); A[oi][oj] = acc; A[oj][oi] = acc; } }
                else if (live) Bm[oi][oj] = acc;
            } else if (tid == 192) s_sumP = acc;
            else if (tid == 224) s_Esum = acc;
            else if (tid >= 256 && tid < 288) {                // calculateTraceCKC: lane i forms row i's term, lane 0 adds the terms in row order
                const int i = tid - 256;
                float term = 0.f;
                if (i < m) {
                    float t1 = 0.f, t2 = 0.f;
                    for (int j = 0; j < m; ++j) { t1 = fadd(t1, fmul(Ks[i][j], C[j][0])); t2 = fadd(t2, fmul(Ks[i][j], C[j][1])); }
                    term = fadd(fmul(C[i][0], t1), fmul(C[i][1], t2));
                }
                float tr = 0.f;
                for (int r = 0; r < m; ++r) { const float tv = __shfl_sync(0xffffffffu, term, r); tr = fadd(tr, tv); }
                if (i == 0) s_trace = tr;
            }
            __syncthreads();
            if (tid == 64) {                                   // energy (:501-502, :110-112)
                float E = s_Esum / fmul(2.f, sigma2);
                E = fadd(E, fmul(fmul(s_sumP, logf_like_glibc(sigma2)), 2.f) / 2.f);
                const float En = fadd(E, fmul(lambda / 2.f, s_trace));
                s_E = En; s_tecr = fabsf((En - E_old) / En);
            }
            // ---- cv::solve(K, Y, C, DECOMP_LU): OpenCV's LUImpl in float; lane j of warp 0 owns row j (every element sees the same
            //      operations in the same order as in the serial loop)
            if (tid < 32) {
                const int lane = tid;
                if (lane == 0) s_ok = 1;
                __syncwarp();
                for (int i = 0; i < m; i++) {
                    float best = (lane >= i && lane < m) ? fabsf(A[lane][i]) : -1.f;
                    int k = lane;
#pragma unroll
                    for (int off = 16; off >= 1; off >>= 1) {
                        const float ob = __shfl_xor_sync(0xffffffffu, best, off); const int ok2 = __shfl_xor_sync(0xffffffffu, k, off);
                        if (ob > best || (ob == best && ok2 < k)) { best = ob; k = ok2; }      // the FIRST largest |a| (strict > in the reference)
                    }
                    if (best < FLT_EPSILON * 10) { if (lane == 0) s_ok = 0; break; }
                    if (k != i) {
                        if (lane >= i && lane < m) { const float t = A[i][lane]; A[i][lane] = A[k][lane]; A[k][lane] = t; }
                        if (lane < 2) { const float t = Bm[i][lane]; Bm[i][lane] = Bm[k][lane]; Bm[k][lane] = t; }
                    }
                    __syncwarp();
                    const float d = -1.f / A[i][i];
                    if (lane > i && lane < m) {
                        const float alpha = fmul(A[lane][i], d);
                        for (int kk = i + 1; kk < m; kk++) A[lane][kk] = fadd(A[lane][kk], fmul(alpha, A[i][kk]));
                        for (int kk = 0; kk < 2; kk++) Bm[lane][kk] = fadd(Bm[lane][kk], fmul(alpha, Bm[i][kk]));
                    }
                    __syncwarp();
                }
                __syncwarp();
                if (s_ok && lane < 2)
                    for (int i = m - 1; i >= 0; i--) {
                        float sv = Bm[i][lane];
                        for (int k = i + 1; k < m; k++) sv = __fsub_rn(sv, fmul(A[i][k], Bm[k][lane]));
                        Bm[i][lane] = sv / A[i][i];
                    }
                __syncwarp();
                if (lane < kCtrl) { C[lane][0] = (s_ok && lane < m) ? Bm[lane][0] : 0.f; C[lane][1] = (s_ok && lane < m) ? Bm[lane][1] : 0.f; }   // singular: cv::solve zeroes its output
            }
            __syncthreads();
            lap(2);
            // ---- per match: calculateV (:432-444), the terms of calculateSigmaSquare (:383-392), calculateGamma's count (:369-381)
            int cnt = 0;
            for (int i = tid; i < n; i += kVfcThreads) {
                float t1 = 0.f, t2 = 0.f;
                float u[kCtrl];
#pragma unroll
                for (int j = 0; j < kCtrl; ++j) u[j] = j < m ? J.U[(size_t)j * n + i] : 0.f;
#pragma unroll
                for (int j = 0; j < kCtrl; ++j) if (j < m) { t1 = fadd(t1, fmul(u[j], C[j][0])); t2 = fadd(t2, fmul(u[j], C[j][1])); }
                J.V[i] = make_float2(t1, t2);
                const float2 y = J.Y[i];
                J.T[i] = sqdist(y.x, y.y, t1, t2);
                cnt += J.P[i] > theta ? 1 : 0;
            }
            if (cnt) atomicAdd(&s_cnt, cnt);
            __syncthreads();
            lap(3);
        }
        // ---- calculateSigmaSquare's two sequential sums (:383-392), 1024 matches per tile
        float a1 = 0.f;
        {
            float* base = &sU[0][0][0];                                    // 2 x (1024 + 1024) floats of the (idle) U tile buffers
            auto load2 = [&](int b, int t0, int first_thread) {
                const int i = t0 + tid - first_thread;
                if (tid >= first_thread && tid - first_thread < kVfcThreads - 64 && i < n && tid - first_thread < kSig) { base[b * 2 * kSig + tid - first_thread] = J.P[i]; base[b * 2 * kSig + kSig + tid - first_thread] = J.T[i]; }
            };
            load2(0, 0, 0);
            __syncthreads();
            int b = 0;
            for (int t0 = 0; t0 < n; t0 += kSig, b ^= 1) {
                const int len = min(kSig, n - t0);
                const float* sPP = base + b * 2 * kSig; const float* sTT = sPP + kSig;
                if (tid >= 64) { if (t0 + kSig < n) load2(b ^ 1, t0 + kSig, 64); }
                else if (tid == 0) a1 = seq_sum4(a1, len, [&](int k) { return mul2(ld4(sPP, k), ld4(sTT, k)); }, [&](int k) { return fmul(sPP[k], sTT[k]); });
                else if (tid == 32) a1 = seq_sum4(a1, len, [&](int k) { return ld4(sPP, k); }, [&](int k) { return sPP[k]; });
                __syncthreads();
            }
        }
        if (tid == 0) s_sg = a1; else if (tid == 32) s_sp = a1;
        __syncthreads();
        lap(4);
        if (tid == 0) {
            s_sigma2 = s_sg / fmul(s_sp, 2.f);
            if (!first) { s_gamma = fmaxf(fminf((float)s_cnt / (float)n, 0.95f), 0.05f); s_iter = s_iter + 1; }
        }
        first = false;
        __syncthreads();
    }
    // obtainCorrectMatch: the P > theta set of the last calculateGamma; all matches when no iteration ran (:66-68)
    const bool none = s_iter == 0;
    __syncthreads();
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    int kept = 0;
    for (int i = tid; i < n; i += kVfcThreads) { const bool k = none || J.P[i] > theta; J.keep[i] = k ? 1 : 0; kept += k ? 1 : 0; }
    if (kept) atomicAdd(&s_cnt, kept);
    __syncthreads();
    if (tid == 0) { J.out[0] = s_cnt; J.out[1] = s_iter; for (int k = 0; k < 5; ++k) J.out[2 + k] = (int)(s_clk[k] >> 10); }
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
    for (int i = 0; i < 32; ++i) { const double v = exp2(i / 32.0); unsigned long long u; memcpy(&u, &v, 8); J.exp_tab[i] = u - ((unsigned long long)i << 47); }
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
    const size_t o_job = 0, o_x = al(sizeof(VfcJob)), o_y = o_x + al(nn * 8), o_u = o_y + al(nn * 8), o_p = o_u + al(nn * kCtrl * 4), o_pr = o_p + al(nn * 4),
                 o_t = o_pr + al(nn * 4), o_v = o_t + al(nn * 4), o_keep = o_v + al(nn * 8), o_out = o_keep + al(nn), total = o_out + 256;
    if ((rc = B[MPL_BUF_RAW].ensure(total))) return rc;
    char* D = (char*)B[MPL_BUF_RAW].ptr;
    char* hs = nullptr;
    if ((rc = mpl_stage_ensure(o_u + al(nn) + 256, (void**)&hs))) return rc;
    J.X = (const float2*)(D + o_x); J.Y = (const float2*)(D + o_y); J.U = (float*)(D + o_u); J.P = (float*)(D + o_p); J.Praw = (float*)(D + o_pr); J.T = (float*)(D + o_t); J.V = (float2*)(D + o_v);
    J.keep = (uint8_t*)(D + o_keep); J.out = (int*)(D + o_out);
    memcpy(hs + o_job, &J, sizeof(J)); memcpy(hs + o_x, X.data(), nn * 8); memcpy(hs + o_y, Y.data(), nn * 8);
    MPL_CUDA_CHECK(cudaMemcpyAsync(D, hs, o_u, cudaMemcpyHostToDevice, st));
    vfc_sparse_kernel<<<1, kVfcThreads, 0, st>>>((const VfcJob*)(D + o_job));
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    char* h_keep = hs + o_u; int* h_out = (int*)(h_keep + al(nn));
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_keep, D + o_keep, nn, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_out, D + o_out, 32, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    memcpy(keep, h_keep, nn);
    *n_keep = h_out[0];
    if (n_iterations) *n_iterations = h_out[1];
    if (getenv("MPL_VFC_TRACE")) fprintf(stderr, "vfc: n %d, %d iterations, kilo-clocks: posteriors %d, M-step sums %d, energy + solve %d, V %d, sigma sums %d\n", n, h_out[1], h_out[2], h_out[3], h_out[4], h_out[5], h_out[6]);
    if ((double)*n_keep / (double)n < 0.1) *vfc_code = -2;                               // vfcMatches.cpp:82-83 "Maybe VFC failed"
    return MPL_OK;
}
