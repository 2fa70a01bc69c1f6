// gms.cu -- GMS (grid-based motion statistics) match filter on the device: the optional post-filter right after the ratio
// test in matchinglib::getCorrespondences.
//
// Replaces (paths relative to matchinglib_poselib/source/matchinglib):
//   filterMatchesGMS                      source/gms.cpp:54-60, called at source/correspondences.cpp:386
//   MatchGMS (getInlierMask, run, assignMatchPairs, verifyCellPairs, setScale, getN9, getGridIndexLeft/Right)
//                                         thirdparty/gms-1.0/src/MatchGMS.cpp:55-325
// Integer / grid work, bit-exact against the reference: the same float normalisation (x * (1.0f / width)), the same
// floor(x * grid [+ 0.5f]) cell indices, first arg-max per left cell, score < 6 * sqrt(mean population) in double.
//
// One CTA (1024 threads) per image pair runs the whole filter: for every (scale, rotation) and each of the 4 shifted grid
// types -- zero the 400 x cells_r motion-statistics matrix (global scratch, L2 resident: 640 KB at scale 1), histogram the
// matches with atomics, one warp per left cell for the row sum / first arg-max / 3 x 3 neighbourhood test, mark the matches.
// Algorithmic bytes are tiny (8 B per keypoint, 16 B per match); the time is the 4 x (zero + scan) passes over the matrix.
// Deviation shared with the oracle: a right-grid index >= cells (keypoint on / beyond the image border; out-of-bounds write
// in the reference) makes the match invalid.
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "../../include/mplb200.h"

#include <math.h>
#include <mutex>
#include <string.h>

namespace {

constexpr int GW = 20, GH = 20, CELLS_L = GW * GH;                // MatchGMS.cpp:64
constexpr int MAX_CELLS_R = 40 * 40;                              // scale ratio 2.0
__constant__ int c_rot[8][9] = {
    {1,2,3,4,5,6,7,8,9}, {4,1,2,7,5,3,8,9,6}, {7,4,1,8,5,2,9,6,3}, {8,7,4,9,5,1,6,3,2},
    {9,8,7,6,5,4,3,2,1}, {6,9,8,3,5,7,2,1,4}, {3,6,9,2,5,8,1,4,7}, {2,3,6,1,5,9,4,7,8}};

__device__ __forceinline__ int n9_at(int cell, int gw, int gh, int k)          // getN9()[k]; k = xi + 4 + yi * 3
{
    const int yi = k / 3 - 1, xi = k % 3 - 1;
    const int xx = cell % gw + xi, yy = cell / gw + yi;
    if (xx < 0 || xx >= gw || yy < 0 || yy >= gh) return -1;
    return xx + yy * gw;
}

__device__ __forceinline__ int grid_left(float x, float y, int type)
{
    int ix, iy;
    if (type == 1)      { ix = (int)floorf(__fmul_rn(x, (float)GW));                       iy = (int)floorf(__fmul_rn(y, (float)GH)); }
    else if (type == 2) { ix = (int)floorf(__fadd_rn(__fmul_rn(x, (float)GW), 0.5f));      iy = (int)floorf(__fmul_rn(y, (float)GH)); }
    else if (type == 3) { ix = (int)floorf(__fmul_rn(x, (float)GW));                       iy = (int)floorf(__fadd_rn(__fmul_rn(y, (float)GH), 0.5f)); }
    else                { ix = (int)floorf(__fadd_rn(__fmul_rn(x, (float)GW), 0.5f));      iy = (int)floorf(__fadd_rn(__fmul_rn(y, (float)GH), 0.5f)); }
    if (ix >= GW || iy >= GH) return -1;
    return ix + iy * GW;
}

struct GmsJob {
    const char* kp1; const char* kp2; size_t kp_stride1, kp_stride2;      // cv::KeyPoint::pt lives at the start of every record
    const mpl_dmatch* matches; int nm;
    float w1i, h1i, w2i, h2i;
    int use_scale, use_rotation;
    int* stat;                      // CELLS_L x MAX_CELLS_R
    int* pl; int* pr;               // per match: left / right cell of the current run
    uint8_t* cur; uint8_t* best;    // per match: inlier flags of the current run / of the best run
    int* n_inliers;
};

__global__ void __launch_bounds__(1024) gms_filter_kernel(const GmsJob* __restrict__ jobs)
{
    __shared__ int cell_pairs[CELLS_L], npts[CELLS_L];
    __shared__ int s_cnt;
    const GmsJob J = jobs[blockIdx.x];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int nm = J.nm;
    const double scale_ratio[5] = {1.0, 0.5, 0.70710678118654757, 1.4142135623730951, 2.0};
    const int n_scales = J.use_scale ? 5 : 1, n_rot = J.use_rotation ? 8 : 1;
    int best = 0;
    for (int i = tid; i < nm; i += 1024) J.best[i] = 0;
    for (int sl = 0; sl < n_scales; ++sl) {
        const int gwr = (int)(GW * scale_ratio[sl]), ghr = (int)(GH * scale_ratio[sl]), cells_r = gwr * ghr;     // setScale
        for (int rt = 0; rt < n_rot; ++rt) {
            for (int i = tid; i < nm; i += 1024) { J.cur[i] = 0; J.pl[i] = 0; J.pr[i] = 0; }
            for (int type = 1; type <= 4; ++type) {
                for (int i = tid; i < CELLS_L * cells_r; i += 1024) J.stat[i] = 0;
                for (int c = tid; c < CELLS_L; c += 1024) { cell_pairs[c] = -1; npts[c] = 0; }
                __syncthreads();
                for (int i = tid; i < nm; i += 1024) {                               // assignMatchPairs (MatchGMS.cpp:173-201)
                    const mpl_dmatch m = J.matches[i];
                    const float2 a = *reinterpret_cast<const float2*>(J.kp1 + (size_t)m.queryIdx * J.kp_stride1);
                    const float2 b = *reinterpret_cast<const float2*>(J.kp2 + (size_t)m.trainIdx * J.kp_stride2);
                    const int l = grid_left(__fmul_rn(a.x, J.w1i), __fmul_rn(a.y, J.h1i), type);
                    int r;
                    if (type == 1) {
                        r = (int)floorf(__fmul_rn(__fmul_rn(b.x, J.w2i), (float)gwr)) + (int)floorf(__fmul_rn(__fmul_rn(b.y, J.h2i), (float)ghr)) * gwr;
                        if (r >= cells_r) r = -1;
                    } else r = J.pr[i];
                    if (l < 0 || r < 0) { J.pl[i] = -1; J.pr[i] = -1; continue; }
                    J.pl[i] = l; J.pr[i] = r;
                    atomicAdd(&J.stat[l * cells_r + r], 1);
                    atomicAdd(&npts[l], 1);
                }
                __syncthreads();
                for (int c = warp; c < CELLS_L; c += 32) {                           // verifyCellPairs (:203-248), one warp per left cell
                    const int* row = J.stat + c * cells_r;
                    int sum = 0, mx = 0, arg = 0x7FFFFFFF;
                    for (int j = lane; j < cells_r; j += 32) {
                        const int v = row[j];
                        sum += v;
                        if (v > mx) { mx = v; arg = j; }                             // first maximum of this lane's columns
                    }
#pragma unroll
                    for (int o = 16; o >= 1; o >>= 1) {
                        sum += __shfl_xor_sync(0xffffffffu, sum, o);
                        const int omx = __shfl_xor_sync(0xffffffffu, mx, o), oarg = __shfl_xor_sync(0xffffffffu, arg, o);
                        if (omx > mx || (omx == mx && oarg < arg)) { mx = omx; arg = oarg; }
                    }
                    if (sum == 0) { if (lane == 0) cell_pairs[c] = -1; continue; }
                    int sc = 0, pop = 0, np = 0;
                    if (lane < 9) {
                        const int ll = n9_at(c, GW, GH, lane), rr = n9_at(arg, gwr, ghr, c_rot[rt][lane] - 1);
                        if (ll != -1 && rr != -1) { sc = J.stat[ll * cells_r + rr]; pop = npts[ll]; np = 1; }
                    }
#pragma unroll
                    for (int o = 8; o >= 1; o >>= 1) {
                        sc += __shfl_xor_sync(0xffffffffu, sc, o); pop += __shfl_xor_sync(0xffffffffu, pop, o); np += __shfl_xor_sync(0xffffffffu, np, o);
                    }
                    if (lane == 0) {
                        const double th = 6.0 * sqrt((double)pop / (double)np);     // THRESH_FACTOR 6 (MatchGMS.h:7)
                        cell_pairs[c] = ((double)sc < th) ? -2 : arg;
                    }
                }
                __syncthreads();
                for (int i = tid; i < nm; i += 1024) {                               // mark inliers (:158-166)
                    const int l = J.pl[i], r = J.pr[i];
                    if (l == -1 && r == -1) continue;
                    if (cell_pairs[l] == r) J.cur[i] = 1;
                }
                __syncthreads();
            }
            if (tid == 0) s_cnt = 0;
            __syncthreads();
            int cnt = 0;
            for (int i = tid; i < nm; i += 1024) cnt += J.cur[i];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
            if (lane == 0 && cnt) atomicAdd(&s_cnt, cnt);
            __syncthreads();
            const int total = s_cnt;
            if (total > best) {                                                      // getInlierMask (:89-96): strict '>'
                best = total;
                for (int i = tid; i < nm; i += 1024) J.best[i] = J.cur[i];
            }
            __syncthreads();
        }
    }
    if (tid == 0) *J.n_inliers = best;
}

}  // namespace

// See include/mplb200.h.
extern "C" int mpl_gms_filter(const void* kp1, int n1, size_t kp_stride1, int width1, int height1, const void* kp2, int n2, size_t kp_stride2,
                              int width2, int height2, const mpl_dmatch* matches, int n_matches, int use_scale, int use_rotation,
                              uint8_t* inlier_mask, int* n_inliers)
{
    if (n_matches < 0 || n1 < 0 || n2 < 0 || !n_inliers || width1 <= 0 || height1 <= 0 || width2 <= 0 || height2 <= 0 || kp_stride1 < 8 || kp_stride2 < 8 ||
        (n_matches > 0 && (!matches || !inlier_mask || !kp1 || !kp2))) { mpl_set_error("mpl_gms_filter: bad argument"); return MPL_ERR_BAD_ARG; }
    *n_inliers = 0;
    if (n_matches == 0) return MPL_OK;
    for (int i = 0; i < n_matches; ++i)
        if (matches[i].queryIdx < 0 || matches[i].queryIdx >= n1 || matches[i].trainIdx < 0 || matches[i].trainIdx >= n2) {
            mpl_set_error("mpl_gms_filter: match %d refers to keypoints (%d, %d) outside [0, %d) x [0, %d)", i, matches[i].queryIdx, matches[i].trainIdx, n1, n2);
            return MPL_ERR_BAD_ARG;
        }
    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx(); cudaStream_t st = ctx->stream; MplBuf* B = ctx->bufs;
    const size_t nm = (size_t)n_matches;
    auto al = [](size_t x) { return (x + 255) & ~(size_t)255; };
    const size_t o_job = 0, o_kp1 = al(sizeof(GmsJob)), o_kp2 = o_kp1 + al((size_t)n1 * 8), o_m = o_kp2 + al((size_t)n2 * 8),
                 o_stat = o_m + al(nm * sizeof(mpl_dmatch)), o_pl = o_stat + al((size_t)CELLS_L * MAX_CELLS_R * 4), o_pr = o_pl + al(nm * 4),
                 o_cur = o_pr + al(nm * 4), o_best = o_cur + al(nm), o_n = o_best + al(nm), total = o_n + 256;
    if ((rc = B[MPL_BUF_RAW].ensure(total))) return rc;
    char* D = (char*)B[MPL_BUF_RAW].ptr;
    char* hs = nullptr;
    if ((rc = mpl_stage_ensure(al(sizeof(GmsJob)) + al(nm) + 64, (void**)&hs))) return rc;
    // keypoints: only the pt field (8 bytes) of every record travels
    MPL_CUDA_CHECK(cudaMemcpy2DAsync(D + o_kp1, 8, kp1, kp_stride1, 8, (size_t)n1, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpy2DAsync(D + o_kp2, 8, kp2, kp_stride2, 8, (size_t)n2, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(D + o_m, matches, nm * sizeof(mpl_dmatch), cudaMemcpyHostToDevice, st));
    GmsJob* J = (GmsJob*)hs;
    J->kp1 = D + o_kp1; J->kp2 = D + o_kp2; J->kp_stride1 = 8; J->kp_stride2 = 8; J->matches = (const mpl_dmatch*)(D + o_m); J->nm = n_matches;
    J->w1i = 1.0f / (float)width1; J->h1i = 1.0f / (float)height1; J->w2i = 1.0f / (float)width2; J->h2i = 1.0f / (float)height2;     // MatchGMS.cpp:106-107
    J->use_scale = use_scale ? 1 : 0; J->use_rotation = use_rotation ? 1 : 0;
    J->stat = (int*)(D + o_stat); J->pl = (int*)(D + o_pl); J->pr = (int*)(D + o_pr); J->cur = (uint8_t*)(D + o_cur); J->best = (uint8_t*)(D + o_best);
    J->n_inliers = (int*)(D + o_n);
    MPL_CUDA_CHECK(cudaMemcpyAsync(D + o_job, J, sizeof(GmsJob), cudaMemcpyHostToDevice, st));
    gms_filter_kernel<<<1, 1024, 0, st>>>((const GmsJob*)(D + o_job));
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    char* h_mask = hs + al(sizeof(GmsJob));
    int* h_n = (int*)(h_mask + al(nm));
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_mask, D + o_best, nm, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_n, D + o_n, 4, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    memcpy(inlier_mask, h_mask, nm);
    *n_inliers = *h_n;
    return MPL_OK;
}
