// pipeline.cu -- both halves of the hot path for a BATCH of image pairs, device resident from the descriptor upload to the
// final (matches, E, inlier mask): SURVEY.md section 8(e) "batch of image pairs" and section 8(f) N3.
//
// The reference driver does, per image pair and on the host (tests/poselib-test/main.cpp:1461-1599, 1717):
//   getCorrespondences -> getMatches (k-NN + ratio test)           matchers.cpp:525-714 / nmslib_matchers.h:159-424
//   points from kp1[queryIdx].pt / kp2[trainIdx].pt                main.cpp:1522-1527
//   ImgToCamCoordTrans (float)((x - cx) / fx)                      pose_helper.cpp:1100-1109
//   Remove_LensDist -> LensDist_Oulu (10 iterations, proof test)   pose_helper.cpp:1169-1279
//   p1, p2 <- n x 2 CV_64F                                         main.cpp:1588-1596
//   estimateEssentialMat(E, p1, p2, "RANSAC", th, refine, mask)    pose_estim.cpp:857-890 -> modelest.cpp:343-474
// Here a call takes any number of pairs; they are processed in sub-batches on three pipeline lanes (own stream, device arena
// and pinned host area each), so the uploads of one sub-batch overlap the kernels of the other and ONE host thread keeps
// the GPU busy.  Per sub-batch: H2D of descriptors / keypoints -> batched row packing -> Hamming scan with blockIdx.z = pair
// (float descriptors: the tensor-core L2 path per pair) -> ratio test + cv::DMatch compaction + keypoint gather + camera
// normalisation + lens undistortion (one CTA per pair) -> minimal samples -> chunked solve / score / device selection
// replay -> refit -> one D2H of the results.  No host synchronisation in between.
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "mpl_ransac.h"
#include "../../include/mplb200.h"

#include <float.h>
#include <string.h>
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <vector>

namespace {

struct CamParams {          // by value to the gather kernel
    double fx[2], fy[2], cx[2], cy[2];
    double k[2][8];         // OpenCV order k1 k2 p1 p2 k3 k4 k5 k6
    int pts_mode;           // 0: double camera-normalised coordinates; 1: float pixel coordinates (cv::KeyPoint::pt)
    int undistort;          // Remove_LensDist's own switch: !(nearZero(sum(dist1)) && nearZero(sum(dist2)))
};

// LensDist_Oulu (pose_helper.cpp:1232-1279), iters = 10 (declaration default, pose_helper.cpp:67): operation order and the
// float <-> double conversions are the reference's (the library is compiled without FMA contraction).
__device__ bool lens_dist_oulu(const float dx, const float dy, float& cx, float& cy, const double* d)
{
    const double k1 = d[0], k2 = d[1], p1 = d[2], p2 = d[3], k3 = d[4], k4 = d[5], k5 = d[6], k6 = d[7];
    double r2, _2xy, rad_corr, delta0, delta1;
    for (int i = 0; i < 10; ++i) {
        r2 = (double)cx * (double)cx + (double)cy * (double)cy;
        _2xy = 2.0 * (double)cx * (double)cy;
        rad_corr = (1.0 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1.0 + ((k6 * r2 + k5) * r2 + k4) * r2);
        delta0 = p1 * _2xy + p2 * (r2 + 2.0 * (double)cx * (double)cx);
        delta1 = p1 * (r2 + 2.0 * (double)cy * (double)cy) + p2 * _2xy;
        cx = (float)(((double)dx - delta0) / rad_corr);
        cy = (float)(((double)dy - delta1) / rad_corr);
    }
    r2 = (double)cx * (double)cx + (double)cy * (double)cy;
    _2xy = 2.0 * (double)cx * (double)cy;
    rad_corr = (1.0 + ((k3 * r2 + k2) * r2 + k1) * r2) / (1.0 + ((k6 * r2 + k5) * r2 + k4) * r2);
    delta0 = p1 * _2xy + p2 * (r2 + 2.0 * (double)cx * (double)cx);
    delta1 = p1 * (r2 + 2.0 * (double)cy * (double)cy) + p2 * _2xy;
    const float px = (float)((double)cx * rad_corr + delta0 - (double)dx);
    const float py = (float)((double)cy * rad_corr + delta1 - (double)dy);
    return !(sqrtf(px * px + py * py) > 0.25f);
}

// correspondence of (query qi, train ti): camera coordinates as the driver forms them; false = dropped by Remove_LensDist
__device__ __forceinline__ bool make_corr(const CamParams& C, const void* pts1, const void* pts2, int qi, int ti, double4& out)
{
    if (C.pts_mode == 0) {
        const double2 a = ((const double2*)pts1)[qi], b = ((const double2*)pts2)[ti];
        out = make_double4(a.x, a.y, b.x, b.y);
        return true;
    }
    const float2 a = ((const float2*)pts1)[qi], b = ((const float2*)pts2)[ti];
    // ImgToCamCoordTrans (pose_helper.cpp:1100-1109)
    float x1 = (float)(((double)a.x - C.cx[0]) / C.fx[0]), y1 = (float)(((double)a.y - C.cy[0]) / C.fy[0]);
    float x2 = (float)(((double)b.x - C.cx[1]) / C.fx[1]), y2 = (float)(((double)b.y - C.cy[1]) / C.fy[1]);
    bool ok = true;
    if (C.undistort) {                                             // Remove_LensDist (pose_helper.cpp:1189-1199)
        const float d1x = x1, d1y = y1, d2x = x2, d2y = y2;
        ok = lens_dist_oulu(d1x, d1y, x1, y1, C.k[0]);
        if (ok) ok = lens_dist_oulu(d2x, d2y, x2, y2, C.k[1]);
    }
    out = make_double4((double)x1, (double)y1, (double)x2, (double)y2);   // main.cpp:1588-1596
    return ok;
}

// One CTA per pair, order-preserving double compaction: (a) queries that pass the ratio test -> cv::DMatch list
// (matchers.cpp:601-625 / nmslib_matchers.h:399-421), (b) of those, the ones whose points survive undistortion -> RANSAC's
// 32-byte correspondence records + their index in the match list.  Also initialises the pair's RANSAC state.
__global__ void __launch_bounds__(1024)
pair_gather_kernel(const ulonglong2* __restrict__ keys, const int2* __restrict__ dims, const unsigned long long* __restrict__ seeds,
                   int nq_stride, int nt_stride, int ratio_test, int is_float, CamParams C, const char* __restrict__ pts1,
                   const char* __restrict__ pts2, mpl_dmatch* __restrict__ matches, double4* __restrict__ corr, int32_t* __restrict__ corr_src,
                   MplPairState* __restrict__ st, int h)
{
    __shared__ int wsum_m[32], wsum_c[32];
    const int p = blockIdx.x;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int nq = dims[p].x, nt = dims[p].y;
    keys += (size_t)p * nq_stride; matches += (size_t)p * nq_stride; corr += (size_t)p * nq_stride; corr_src += (size_t)p * nq_stride;
    const size_t pe = C.pts_mode == 0 ? 16 : 8;
    const void* P1 = pts1 + (size_t)p * nq_stride * pe;
    const void* P2 = pts2 + (size_t)p * nt_stride * pe;
    const int per = ((nq + 1023) / 1024) * 32;                   // queries per warp, lanes read consecutive keys
    const int start = min(nq, wid * per), end = min(nq, start + per);
    auto dist_f = [&](mpl_key64 k) {
        if (k == MPL_KEY_INVALID) return is_float ? 3.402823466e+38f : (float)0x7FFFFFFF;
        const uint32_t d = (uint32_t)(k >> 32);
        return is_float ? __uint_as_float(d) : (float)(int32_t)d;
    };
    auto passes = [&](const ulonglong2 k) {
        if (k.x == MPL_KEY_INVALID) return false;
        return !ratio_test || dist_f(k.x) < 0.75f * dist_f(k.y);
    };
    int sum_m = 0, sum_c = 0;
    for (int i = start; i < end; i += 32) {
        ulonglong2 k = make_ulonglong2(MPL_KEY_INVALID, MPL_KEY_INVALID);
        const bool pm = (i + lane < end) && passes(k = keys[i + lane]);
        double4 c;
        const bool pc = pm && make_corr(C, P1, P2, i + lane, (int)(uint32_t)(k.x & 0xFFFFFFFFull), c);
        sum_m += __popc(__ballot_sync(0xffffffffu, pm));
        sum_c += __popc(__ballot_sync(0xffffffffu, pc));
    }
    if (lane == 0) { wsum_m[wid] = sum_m; wsum_c[wid] = sum_c; }
    __syncthreads();
    if (wid == 0) {
        int a = wsum_m[lane], b = wsum_c[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int ya = __shfl_up_sync(0xffffffffu, a, o), yb = __shfl_up_sync(0xffffffffu, b, o);
            if (lane >= o) { a += ya; b += yb; }
        }
        wsum_m[lane] = a; wsum_c[lane] = b;
    }
    __syncthreads();
    const int total_m = wsum_m[31], total_c = wsum_c[31];
    // Remove_LensDist returns false when fewer than 16 correspondences survive (pose_helper.cpp:1203-1204): the driver skips the pair
    const bool undist_failed = C.pts_mode == 1 && C.undistort && total_c < 16;
    int pos_m = wid ? wsum_m[wid - 1] : 0, pos_c = wid ? wsum_c[wid - 1] : 0;
    for (int i = start; i < end; i += 32) {
        ulonglong2 k = make_ulonglong2(MPL_KEY_INVALID, MPL_KEY_INVALID);
        const bool pm = (i + lane < end) && passes(k = keys[i + lane]);
        double4 c = make_double4(0, 0, 0, 0);
        const int ti = (int)(uint32_t)(k.x & 0xFFFFFFFFull);
        const bool pc = pm && make_corr(C, P1, P2, i + lane, ti, c);
        const unsigned bm = __ballot_sync(0xffffffffu, pm), bc = __ballot_sync(0xffffffffu, pc);
        const unsigned below = (1u << lane) - 1u;
        const int om = pos_m + __popc(bm & below);
        if (pm) {
            mpl_dmatch m; m.queryIdx = i + lane; m.trainIdx = ti; m.imgIdx = -1; m.distance = dist_f(k.x);
            matches[om] = m;
        }
        if (pc && !undist_failed) {
            const int oc = pos_c + __popc(bc & below);
            corr[oc] = c; corr_src[oc] = om;
        }
        pos_m += __popc(bm); pos_c += __popc(bc);
    }
    if (tid == 0) {
        MplPairState S;
        S.n1 = nq; S.n2 = nt; S.n_matches = total_m; S.n = undist_failed ? 0 : total_c; S.niters = h; S.iters_run = 0; S.best_count = 0;
        S.best_slot = -1; S.refit_nm = 0; S.n_inl = 0; S.status = undist_failed ? 1 : 0; S.pad = 0; S.errminsum = DBL_MAX;
        for (int k = 0; k < 9; ++k) S.E[k] = 0.0;
        S.seed = seeds[p];
        st[p] = S;
    }
}

// descriptor rows of every (pair, side) from the raw upload area into the scan layout (w16 x 16-byte rows, zero padded,
// `used` payload bytes: BRUTEFORCENMS drops the last packed word)
__global__ void pack_pairs_kernel(const uint8_t* __restrict__ raw, const unsigned long long* __restrict__ offs, const int2* __restrict__ dims,
                                  int nbytes, int used, int row_bytes, int nq_stride, int nt_stride, uint8_t* __restrict__ dq, uint8_t* __restrict__ dt)
{
    const int p = blockIdx.y >> 1, side = blockIdx.y & 1;
    const int n = side ? dims[p].y : dims[p].x;
    const uint8_t* src = raw + offs[2 * p + side];
    uint8_t* dst = side ? dt + (size_t)p * nt_stride * row_bytes : dq + (size_t)p * nq_stride * row_bytes;
    const size_t total = (size_t)n * row_bytes;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / row_bytes;
        const int c = (int)(i - r * row_bytes);
        dst[i] = c < used ? src[r * nbytes + c] : (uint8_t)0;
    }
}

// Upload of a sub-batch without the copy engine: when every buffer of the sub-batch lies in mpl_host_alloc memory (pinned and mapped), ONE
// kernel reads the host buffers over PCIe (16-byte loads, thousands in flight) instead of four cudaMemcpyAsync per pair, whose set-up
// cost kept the link at half its rate for copies of 80 - 300 KB.  One grid row per buffer.
// host-side time of a call by phase (MPL_PAIR_TRACE=1 prints it): where the single host thread spends the call
thread_local double g_trace_wait = 0, g_trace_enqueue = 0, g_trace_total = 0;

struct CopyItem { const char* src; unsigned long long dst_off; unsigned long long bytes; };
__global__ void __launch_bounds__(256)
gather_copy_kernel(const CopyItem* __restrict__ items, char* __restrict__ base)
{
    const CopyItem it = items[blockIdx.y];
    const size_t n16 = it.bytes >> 4;
    const uint4* s4 = reinterpret_cast<const uint4*>(it.src);
    uint4* d4 = reinterpret_cast<uint4*>(base + it.dst_off);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) {
        uint4 v;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(s4 + i));
        d4[i] = v;
    }
    if (blockIdx.x == 0) for (size_t i = (n16 << 4) + threadIdx.x; i < it.bytes; i += blockDim.x) base[it.dst_off + i] = it.src[i];
}

// inlier mask over the MATCH list (what the caller indexes with its cv::DMatch vector): matches dropped by the undistortion
// never reach RANSAC and read 0
__global__ void __launch_bounds__(256)
mask_to_matches_kernel(const MplPairState* __restrict__ st, const uint8_t* __restrict__ mask_corr, const int32_t* __restrict__ corr_src,
                       int nq_stride, uint8_t* __restrict__ mask_match)
{
    const int p = blockIdx.x;
    const int nm = st[p].n_matches, n = st[p].n;
    const bool have = st[p].n_inl > 0;
    uint8_t* mm = mask_match + (size_t)p * nq_stride;
    for (int i = threadIdx.x; i < nm; i += blockDim.x) mm[i] = 0;
    __syncthreads();
    if (!have) return;
    const uint8_t* mc = mask_corr + (size_t)p * nq_stride;
    const int32_t* src = corr_src + (size_t)p * nq_stride;
    for (int i = threadIdx.x; i < n; i += blockDim.x) mm[src[i]] = mc[i];
}

inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

struct Arena {              // offsets inside a lane's device arena for a sub-batch of `bs` pairs
    size_t dims, offs, seeds, raw, dq, dt, keys, pts1, pts2, matches, corr, src, state, samples, models, nm, counts, errs,
           maskc, maskm, basis, rmodels, rnm, l2scr, part, cptab, total;
};

struct BatchPlan {
    int cv_type, cols, used, w16, row_bytes;       // binary: row_bytes = w16 * 16; float: row_bytes = ld * 4
    int is_float, sqrt_out, ignore_tail;
    int nq_stride, nt_stride, h, bs;
    int splits;                                     // train splits of the Hamming scan (> 1 only for batches too small to fill the GPU)
    int use_tc;                                     // Hamming scan on the tensor cores (mpl_hamming_tc_top2_batch): one ranking launch per group of pairs
    size_t pe;                                      // bytes per keypoint
    size_t l2_scratch;
    CamParams cam;
};

Arena make_arena(const BatchPlan& P, size_t raw_bytes)
{
    Arena A; size_t o = 0;
    const size_t bs = (size_t)P.bs, nq = (size_t)P.nq_stride, nt = (size_t)P.nt_stride, h = (size_t)(P.h > 0 ? P.h : 1);
    auto take = [&](size_t bytes) { const size_t at = o; o = al256(o + bytes); return at; };
    A.dims = take(bs * 8); A.offs = take(bs * 16); A.seeds = take(bs * 8);
    A.raw = take(raw_bytes);
    A.dq = take(bs * nq * P.row_bytes); A.dt = take(bs * nt * P.row_bytes);
    A.keys = take(bs * nq * 16);
    A.pts1 = take(bs * nq * P.pe); A.pts2 = take(bs * nt * P.pe);
    A.matches = take(bs * nq * sizeof(mpl_dmatch)); A.corr = take(bs * nq * 32); A.src = take(bs * nq * 4);
    A.state = take(bs * sizeof(MplPairState));
    A.samples = take(bs * h * 20); A.models = take(bs * h * 720); A.nm = take(bs * h * 4);
    A.counts = take(bs * h * 40); A.errs = take(bs * h * 80);
    A.maskc = take(bs * nq); A.maskm = take(bs * nq);
    A.basis = take(bs * 36 * 8); A.rmodels = take(bs * 90 * 8); A.rnm = take(bs * 4);
    A.l2scr = take(P.l2_scratch);
    A.part = take(P.splits > 1 ? bs * (size_t)P.splits * nq * 16 : 16);
    A.cptab = take(bs * 4 * sizeof(CopyItem));
    A.total = o;
    return A;
}

struct HostArea { size_t dims, offs, seeds, cptab, state, matches, maskm, samples, total; };
HostArea make_host_area(const BatchPlan& P, bool want_samples)
{
    HostArea H; size_t o = 0;
    const size_t bs = (size_t)P.bs, nq = (size_t)P.nq_stride, h = (size_t)(P.h > 0 ? P.h : 1);
    auto take = [&](size_t bytes) { const size_t at = o; o = al256(o + bytes); return at; };
    H.dims = take(bs * 8); H.offs = take(bs * 16); H.seeds = take(bs * 8); H.cptab = take(bs * 4 * sizeof(CopyItem));
    H.state = take(bs * sizeof(MplPairState)); H.matches = take(bs * nq * sizeof(mpl_dmatch)); H.maskm = take(bs * nq);
    H.samples = take(want_samples ? bs * h * 20 : 16);
    H.total = o;
    return H;
}

// Enqueue one sub-batch (pairs [p0, p0 + nb)) on a lane: uploads, every kernel, downloads.  Nothing synchronises.
int enqueue_sub_batch(MplLane& L, const BatchPlan& P, const mpl_pair_io* pairs, int p0, int nb, const mpl_pair_params* prm, bool want_samples)
{
    int rc;
    cudaStream_t st = L.stream;
    size_t raw_bytes = 0;
    const size_t elem = P.is_float ? 4 : 1;
    for (int i = 0; i < nb; ++i) raw_bytes += al256((size_t)pairs[p0 + i].n1 * P.cols * elem) + al256((size_t)pairs[p0 + i].n2 * P.cols * elem);
    const Arena A = make_arena(P, P.is_float ? 16 : raw_bytes);
    const HostArea H = make_host_area(P, want_samples);
    if ((rc = L.arena.ensure(A.total))) return rc;
    if ((rc = L.ensure_host(H.total))) return rc;
    char* D = (char*)L.arena.ptr; char* HS = (char*)L.host;
    int2* h_dims = (int2*)(HS + H.dims); unsigned long long* h_offs = (unsigned long long*)(HS + H.offs);
    unsigned long long* h_seeds = (unsigned long long*)(HS + H.seeds);
    // ---- uploads: cudaMemcpyAsync per buffer.  Experiment, off by default (MPL_PAIR_ZEROCOPY=1): one gather kernel that reads the host
    // buffers directly when every buffer is device-readable pinned memory (mpl_host_alloc), contiguous and 16-byte aligned.  Measured at
    // C4: matching-only 31.8 k -> 32.6 k pairs/s (the link, not the per-copy set-up, is the bound), whole pipeline 29.8 k -> 22.0 k (the
    // gather's CTAs wait on PCIe while holding SMs the RANSAC kernels want; the copy engine works for free).
    bool zc = false;
    if (const char* e = getenv("MPL_PAIR_ZEROCOPY")) zc = atoi(e) != 0 && !P.is_float;
    for (int i = 0; i < nb && zc; ++i) {
        const mpl_pair_io& pr = pairs[p0 + i];
        const size_t rowb = (size_t)P.cols * elem;
        const void* bp[4] = {pr.desc1, pr.desc2, pr.pts1, pr.pts2};
        const size_t bb[4] = {(size_t)pr.n1 * rowb, (size_t)pr.n2 * rowb, (size_t)pr.n1 * P.pe, (size_t)pr.n2 * P.pe};
        if ((pr.n1 && pr.step1 != rowb) || (pr.n2 && pr.step2 != rowb)) zc = false;
        for (int b = 0; b < 4 && zc; ++b)
            if (bb[b] && ((((uintptr_t)bp[b]) & 15) != 0 || !mpl_host_is_mapped(bp[b], bb[b]))) zc = false;
    }
    CopyItem* h_cp = (CopyItem*)(HS + H.cptab);
    size_t off = 0;
    for (int i = 0; i < nb; ++i) {
        const mpl_pair_io& pr = pairs[p0 + i];
        h_dims[i] = make_int2(pr.n1, pr.n2);
        h_seeds[i] = pr.seed;
        const void* src[2] = {pr.desc1, pr.desc2}; const int n[2] = {pr.n1, pr.n2}; const size_t stp[2] = {pr.step1, pr.step2};
        for (int s = 0; s < 2; ++s) {
            const size_t rowb = (size_t)P.cols * elem;
            char* dst; size_t dpitch;
            if (P.is_float) {                                      // straight into the scan layout (row pitch ld floats)
                dst = D + (s ? A.dt + (size_t)i * P.nt_stride * P.row_bytes : A.dq + (size_t)i * P.nq_stride * P.row_bytes);
                dpitch = (size_t)P.row_bytes;
                h_offs[2 * i + s] = 0;
            } else {
                h_offs[2 * i + s] = off; dst = D + A.raw + off; dpitch = rowb; off += al256((size_t)n[s] * rowb);
            }
            if (zc) { h_cp[4 * i + s] = CopyItem{(const char*)src[s], (unsigned long long)(dst - D), (unsigned long long)((size_t)n[s] * rowb)}; continue; }
            if (n[s] == 0) continue;
            if (stp[s] == rowb && dpitch == rowb) MPL_CUDA_CHECK(cudaMemcpyAsync(dst, src[s], (size_t)n[s] * rowb, cudaMemcpyHostToDevice, st));
            else MPL_CUDA_CHECK(cudaMemcpy2DAsync(dst, dpitch, src[s], stp[s], rowb, n[s], cudaMemcpyHostToDevice, st));
        }
        if (zc) {
            h_cp[4 * i + 2] = CopyItem{(const char*)pr.pts1, (unsigned long long)(A.pts1 + (size_t)i * P.nq_stride * P.pe), (unsigned long long)((size_t)pr.n1 * P.pe)};
            h_cp[4 * i + 3] = CopyItem{(const char*)pr.pts2, (unsigned long long)(A.pts2 + (size_t)i * P.nt_stride * P.pe), (unsigned long long)((size_t)pr.n2 * P.pe)};
            continue;
        }
        if (pr.n1) MPL_CUDA_CHECK(cudaMemcpyAsync(D + A.pts1 + (size_t)i * P.nq_stride * P.pe, pr.pts1, (size_t)pr.n1 * P.pe, cudaMemcpyHostToDevice, st));
        if (pr.n2) MPL_CUDA_CHECK(cudaMemcpyAsync(D + A.pts2 + (size_t)i * P.nt_stride * P.pe, pr.pts2, (size_t)pr.n2 * P.pe, cudaMemcpyHostToDevice, st));
    }
    if (zc) {
        MPL_CUDA_CHECK(cudaMemcpyAsync(D + A.cptab, HS + H.cptab, (size_t)nb * 4 * sizeof(CopyItem), cudaMemcpyHostToDevice, st));
        gather_copy_kernel<<<dim3(16, 4 * nb), 256, 0, st>>>((const CopyItem*)(D + A.cptab), D);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    }
    MPL_CUDA_CHECK(cudaMemcpyAsync(D + A.dims, HS + H.dims, (size_t)nb * 8, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(D + A.offs, HS + H.offs, (size_t)nb * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(D + A.seeds, HS + H.seeds, (size_t)nb * 8, cudaMemcpyHostToDevice, st));
    const int2* d_dims = (const int2*)(D + A.dims);
    mpl_key64* d_keys = (mpl_key64*)(D + A.keys);
    // ---- k-NN
    if (!P.is_float) {
        int blocks = (int)(((size_t)std::max(P.nq_stride, P.nt_stride) * P.row_bytes + 255) / 256);
        blocks = std::max(1, std::min(blocks, 64));
        pack_pairs_kernel<<<dim3(blocks, 2 * nb), 256, 0, st>>>((const uint8_t*)(D + A.raw), (const unsigned long long*)(D + A.offs), d_dims,
                                                                 P.cols, P.used, P.row_bytes, P.nq_stride, P.nt_stride, (uint8_t*)(D + A.dq), (uint8_t*)(D + A.dt));
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        if (P.use_tc) {
            if ((rc = mpl_hamming_tc_top2_batch((const uint8_t*)(D + A.dq), P.nq_stride, (const uint8_t*)(D + A.dt), P.nt_stride, P.w16, nb, d_dims,
                                                D + A.l2scr, P.l2_scratch, d_keys, mpl_ctx()->sm_count, st, mpl_hamming_path() == 3 ? 1 : 2))) return rc;
        } else if (P.splits > 1) {
            // a batch of a few pairs cannot fill the GPU with one CTA row per pair: split the train rows as the single-pair
            // matcher does and merge the per-split keys (one merge launch per pair; only taken for nb <= 8)
            mpl_key64* d_part = (mpl_key64*)(D + A.part);
            if ((rc = mpl_launch_hamming_top2((const uint8_t*)(D + A.dq), P.nq_stride, (const uint8_t*)(D + A.dt), P.nt_stride, P.w16, P.splits, nb, 0,
                                              d_part, st, d_dims))) return rc;
            for (int i = 0; i < nb; ++i)
                if ((rc = mpl_launch_merge_top2(d_part + (size_t)i * P.splits * P.nq_stride * 2, P.splits, P.nq_stride, 2, nullptr, nullptr, -1,
                                                d_keys + (size_t)i * P.nq_stride * 2, st))) return rc;
        } else if ((rc = mpl_launch_hamming_top2((const uint8_t*)(D + A.dq), P.nq_stride, (const uint8_t*)(D + A.dt), P.nt_stride, P.w16, 1, nb, 0,
                                                 d_keys, st, d_dims))) return rc;
    } else {
        const size_t ld = (size_t)P.row_bytes / 4;
        MPL_CUDA_CHECK(cudaMemsetAsync(d_keys, 0xFF, (size_t)nb * P.nq_stride * 16, st));
        for (int i = 0; i < nb; ++i) {
            const mpl_pair_io& pr = pairs[p0 + i];
            mpl_key64* ko = d_keys + (size_t)i * P.nq_stride * 2;
            if (pr.n1 == 0) continue;
            if ((rc = mpl_knn2_l2f32_dev(D + A.dq + (size_t)i * P.nq_stride * P.row_bytes, pr.n1, ld, D + A.dt + (size_t)i * P.nt_stride * P.row_bytes,
                                         pr.n2, ld, P.cols, P.sqrt_out, 0, D + A.l2scr, P.l2_scratch, ko, (void*)st))) return rc;
        }
    }
    // ---- ratio test, DMatch list, correspondences, RANSAC state
    MplPairState* d_state = (MplPairState*)(D + A.state);
    pair_gather_kernel<<<nb, 1024, 0, st>>>((const ulonglong2*)d_keys, d_dims, (const unsigned long long*)(D + A.seeds), P.nq_stride, P.nt_stride,
                                            prm->ratio_test ? 1 : 0, P.is_float, P.cam, D + A.pts1, D + A.pts2, (mpl_dmatch*)(D + A.matches),
                                            (double4*)(D + A.corr), (int32_t*)(D + A.src), d_state, P.h);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    // ---- RANSAC
    const int h = P.h;
    if (h > 0) {
        int32_t* d_samples = (int32_t*)(D + A.samples); double* d_models = (double*)(D + A.models); int32_t* d_nm = (int32_t*)(D + A.nm);
        int32_t* d_counts = (int32_t*)(D + A.counts); double* d_errs = (double*)(D + A.errs);
        const void* d_corr = D + A.corr;
        if ((rc = mpl_launch_draw_samples_batch(d_state, nb, h, d_samples, st))) return rc;
        // solve in two chunks: the first fills about one wave of the solver (its latency does not depend on the thread count up
        // to there), the second takes whatever the adaptive stop leaves; score / replay in growing sub-chunks so that pairs that
        // stop early (modelest.cpp:403-404) never pay for the samples behind their stop
        int c0 = ((32768 / nb + 63) / 64) * 64;
        c0 = std::max(64, std::min(c0, h));
        for (int s0 = 0; s0 < h;) {
            const int c = s0 == 0 ? std::min(c0, h) : h - s0;
            if ((rc = mpl_launch_solve_chunk(d_corr, (size_t)P.nq_stride, d_state, nb, d_samples, h, s0, c, d_models, d_nm, st))) return rc;
            int sc = 64;
            for (int u = s0; u < s0 + c;) {
                const int cc = std::min(sc, s0 + c - u);
                if ((rc = mpl_launch_score_slots(d_corr, (size_t)P.nq_stride, d_state, nb, d_models, d_nm, h, u, cc, prm->th, P.nq_stride / 2,
                                                 d_counts, d_errs, st))) return rc;
                if ((rc = mpl_launch_replay_chunk(d_state, nb, d_models, d_nm, d_counts, d_errs, h, u, cc, prm->conf, st))) return rc;
                u += cc;
                if (u >= 128) sc = std::min(sc * 2, 512);
            }
            s0 += c;
        }
        if ((rc = mpl_launch_refit(d_corr, (size_t)P.nq_stride, d_state, nb, prm->th, prm->refine ? 1 : 0, (uint8_t*)(D + A.maskc), (size_t)P.nq_stride,
                                   (double*)(D + A.basis), (double*)(D + A.rmodels), (int32_t*)(D + A.rnm), st))) return rc;
    }
    mask_to_matches_kernel<<<nb, 256, 0, st>>>(d_state, (const uint8_t*)(D + A.maskc), (const int32_t*)(D + A.src), P.nq_stride, (uint8_t*)(D + A.maskm));
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    // ---- downloads
    MPL_CUDA_CHECK(cudaMemcpyAsync(HS + H.state, D + A.state, (size_t)nb * sizeof(MplPairState), cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(HS + H.matches, D + A.matches, (size_t)nb * P.nq_stride * sizeof(mpl_dmatch), cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(HS + H.maskm, D + A.maskm, (size_t)nb * P.nq_stride, cudaMemcpyDeviceToHost, st));
    if (want_samples && h > 0) MPL_CUDA_CHECK(cudaMemcpyAsync(HS + H.samples, D + A.samples, (size_t)nb * h * 20, cudaMemcpyDeviceToHost, st));
    return MPL_OK;
}

// Wait for a lane's sub-batch and hand its results to the caller's buffers.
int finish_sub_batch(MplLane& L, const BatchPlan& P, const mpl_pair_io* pairs, int p0, int nb, bool want_samples, mpl_pair_result* results)
{
    const auto t_a = std::chrono::steady_clock::now();
    MPL_CUDA_CHECK(cudaStreamSynchronize(L.stream));
    g_trace_wait += std::chrono::duration<double>(std::chrono::steady_clock::now() - t_a).count();
    const HostArea H = make_host_area(P, want_samples);
    const char* HS = (const char*)L.host;
    const MplPairState* S = (const MplPairState*)(HS + H.state);
    for (int i = 0; i < nb; ++i) {
        const mpl_pair_io& pr = pairs[p0 + i];
        mpl_pair_result& R = results[p0 + i];
        R.n_matches = S[i].n_matches; R.n_corr = S[i].n; R.n_samples = (S[i].n >= 5) ? P.h : 0; R.n_inl = S[i].n_inl;
        R.iters_run = S[i].iters_run; R.status = S[i].status;
        memcpy(R.E, S[i].E, sizeof(R.E));
        if (R.n_inl <= 0) memset(R.E, 0, sizeof(R.E));
        if (pr.matches && R.n_matches > 0) memcpy(pr.matches, HS + H.matches + (size_t)i * P.nq_stride * sizeof(mpl_dmatch), (size_t)R.n_matches * sizeof(mpl_dmatch));
        if (pr.mask && R.n_matches > 0) memcpy(pr.mask, HS + H.maskm + (size_t)i * P.nq_stride, (size_t)R.n_matches);
        if (pr.samples && R.n_samples > 0) memcpy(pr.samples, HS + H.samples + (size_t)i * P.h * 20, (size_t)P.h * 20);
    }
    return MPL_OK;
}

}  // namespace

// See include/mplb200.h.
extern "C" int mpl_match_pose_batch(const mpl_pair_io* pairs, int n_pairs, const mpl_pair_params* prm, mpl_pair_result* results)
{
    if (n_pairs < 0 || (n_pairs > 0 && (!pairs || !results)) || !prm || !prm->matcher_name) { mpl_set_error("mpl_match_pose_batch: null argument"); return MPL_ERR_BAD_ARG; }
    const bool linear = strcmp(prm->matcher_name, "LINEAR") == 0, nms = strcmp(prm->matcher_name, "BRUTEFORCENMS") == 0;
    if (!linear && !nms) { mpl_set_error("mpl_match_pose_batch: matcher %s is not served (no CPU fallback)", prm->matcher_name); return MPL_ERR_BAD_ARG; }
    if ((prm->cv_type != 0 && prm->cv_type != 5) || prm->cols <= 0 || prm->h < 0 || (prm->pts_mode != 0 && prm->pts_mode != 1)) {
        mpl_set_error("mpl_match_pose_batch: bad parameters (cv_type %d, cols %d, h %d, pts_mode %d)", prm->cv_type, prm->cols, prm->h, prm->pts_mode);
        return MPL_ERR_BAD_ARG;
    }
    BatchPlan P;
    memset(&P, 0, sizeof(P));
    P.cv_type = prm->cv_type; P.cols = prm->cols; P.is_float = prm->cv_type == 5; P.h = prm->h;
    P.sqrt_out = (P.is_float && nms) ? 1 : 0;
    P.ignore_tail = (!P.is_float && nms) ? ((P.cols & 1) ? 1 : 2) : 0;
    P.pe = prm->pts_mode == 0 ? 16 : 8;
    if (P.is_float) {
        P.row_bytes = ((P.cols + 3) / 4) * 16; P.used = P.cols; P.w16 = 0;
    } else {
        P.used = P.cols - P.ignore_tail;
        if (P.used <= 0 || P.used > 256) { mpl_set_error("mpl_match_pose_batch: unsupported descriptor length %d", P.cols); return MPL_ERR_LIMIT; }
        int w = (P.used + 15) / 16; P.w16 = 1; while (P.w16 < w) P.w16 *= 2;
        P.row_bytes = P.w16 * 16;
    }
    bool want_samples = false;
    int nq_max = 1, nt_max = 1;
    const size_t min_step = (size_t)P.cols * (P.is_float ? 4 : 1);
    for (int i = 0; i < n_pairs; ++i) {
        const mpl_pair_io& pr = pairs[i];
        if (pr.n1 < 0 || pr.n2 < 0 || (pr.n1 > 0 && (!pr.desc1 || !pr.pts1 || pr.step1 < min_step)) || (pr.n2 > 0 && (!pr.desc2 || !pr.pts2 || pr.step2 < min_step))) {
            mpl_set_error("mpl_match_pose_batch: pair %d: bad descriptor / keypoint arguments", i); return MPL_ERR_BAD_ARG;
        }
        if (pr.n2 > (1 << 20)) { mpl_set_error("mpl_match_pose_batch: pair %d: more than 2^20 train rows", i); return MPL_ERR_LIMIT; }
        nq_max = std::max(nq_max, pr.n1); nt_max = std::max(nt_max, pr.n2);
        want_samples |= pr.samples != nullptr;
        memset(&results[i], 0, sizeof(results[i]));
    }
    if (n_pairs == 0) return MPL_OK;
    P.nq_stride = (nq_max + 7) & ~7; P.nt_stride = (nt_max + 7) & ~7;
    // camera model (pts_mode 1)
    P.cam.pts_mode = prm->pts_mode;
    const double* K[2] = {prm->K0, prm->K1}; const double* dd[2] = {prm->dist0, prm->dist1};
    double dsum[2] = {0, 0};
    for (int c = 0; c < 2; ++c) {
        P.cam.fx[c] = K[c][0]; P.cam.cx[c] = K[c][2]; P.cam.fy[c] = K[c][4]; P.cam.cy[c] = K[c][5];
        for (int k = 0; k < 8; ++k) { P.cam.k[c][k] = dd[c][k]; dsum[c] += dd[c][k]; }
    }
    auto near_zero = [](double d) { return d < 1e-3 && d > -1e-3; };          // pose_helper.h:82-87
    P.cam.undistort = (near_zero(dsum[0]) && near_zero(dsum[1])) ? 0 : 1;     // pose_helper.cpp:1176
    if (prm->pts_mode == 1 && (P.cam.fx[0] == 0 || P.cam.fy[0] == 0 || P.cam.fx[1] == 0 || P.cam.fy[1] == 0)) {
        mpl_set_error("mpl_match_pose_batch: pts_mode 1 needs the camera matrices K0, K1"); return MPL_ERR_BAD_ARG;
    }

    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx();
    P.l2_scratch = P.is_float ? mpl_knn_l2_scratch_bytes(P.nq_stride, P.nt_stride, P.cols) + 256 : 16;
    // sub-batch size: the model slots (720 B per sample) are the big array -- keep a lane's arena under ~1 GiB
    const size_t per_pair = (size_t)std::max(P.h, 1) * (720 + 20 + 4 + 120) + (size_t)P.nq_stride * (P.row_bytes + 16 + P.pe + 16 + 32 + 4 + 2 + P.cols)
                            + (size_t)P.nt_stride * (P.row_bytes + P.pe + P.cols);
    // sub-batch cap: MPL_PAIR_BS (experiments; default 256) and an arena budget of 1 GiB per lane (2 GiB when the cap is raised)
    size_t bs_cap = 256;
    if (const char* e = getenv("MPL_PAIR_BS")) { const long v = atol(e); if (v >= 1 && v <= 4096) bs_cap = (size_t)v; }
    int bs = (int)std::min<size_t>(bs_cap, std::max<size_t>(1, ((size_t)(bs_cap > 256 ? 2 : 1) << 30) / per_pair));
    if (P.is_float) bs = std::min(bs, 32);                                       // one tensor-core pipeline per pair on the lane's stream
    bs = std::min(bs, n_pairs);
    // lanes in flight (MPL_PAIR_LANES, 1 .. 4): a sub-batch is a chain of dependent launches (uploads, scan, RANSAC chunks with a solve
    // latency each, refit, downloads) of ~18 ms whatever its size, so a call of a few hundred pairs wants several sub-batches side by side
    int n_lanes = 3;                                                             // measured at C4: 2 lanes 25.7 k, 3 lanes 29.4 k, 4 lanes 29.6 k pairs/s (4 hurts calls of ~500 pairs)
    if (const char* e = getenv("MPL_PAIR_LANES")) { const long v = atol(e); if (v >= 1 && v <= 4) n_lanes = (int)v; }
    // the lanes need at least one sub-batch each; split small batches so that uploads and kernels overlap
    if (n_pairs >= 8 * n_lanes && n_pairs <= bs * (n_lanes - 1) + bs / 2) bs = std::min(bs, (n_pairs + n_lanes - 1) / n_lanes);
    P.bs = bs;
    P.splits = 1;
    // Hamming scan: tensor cores (bits as bf16, knn_l2_tc.cu) when the rows are 16 / 32 / 64 bytes and the pairs are big enough to
    // fill the 256 x 128 tiles; the popc batch kernel otherwise (mpl_set_hamming_path: 1 forces popc, 2 the tensor cores)
    P.use_tc = 0;
    if (!P.is_float && mpl_hamming_tc_batch_supported(P.nq_stride, P.nt_stride, P.w16)) {
        const int path = mpl_hamming_path();
        if (path >= 2 || (path == 0 && P.nq_stride >= 512 && P.nt_stride >= 512)) P.use_tc = 1;
    }
    if (P.use_tc) P.l2_scratch = mpl_hamming_tc_batch_scratch_bytes(P.nq_stride, P.nt_stride, P.w16, bs, ctx->sm_count) + 256;
    if (!P.is_float && !P.use_tc && bs <= 8) {
        const int sp = mpl_hamming_pick_splits(P.nq_stride, P.nt_stride, P.w16, std::max(1, ctx->sm_count / bs));
        if (sp > 1 && (P.nt_stride + sp - 1) / sp >= 32) P.splits = sp;
    }
    const int n_sub = (n_pairs + bs - 1) / bs;
    const auto t_call = std::chrono::steady_clock::now();
    g_trace_wait = g_trace_enqueue = 0;
    for (int i = 0; i < n_lanes; ++i) if (!ctx->lanes[i].stream) MPL_CUDA_CHECK(cudaStreamCreateWithFlags(&ctx->lanes[i].stream, cudaStreamNonBlocking));
    int inflight_p0[4] = {0, 0, 0, 0}, inflight_nb[4] = {0, 0, 0, 0};
    rc = MPL_OK;
    for (int k = 0; k < n_sub && rc == MPL_OK; ++k) {
        const int li = k % n_lanes;
        MplLane& L = ctx->lanes[li];
        if (L.busy) { rc = finish_sub_batch(L, P, pairs, inflight_p0[li], inflight_nb[li], want_samples, results); L.busy = 0; if (rc) break; }
        const int p0 = k * bs, nb = std::min(bs, n_pairs - p0);
        const auto t_e = std::chrono::steady_clock::now();
        rc = enqueue_sub_batch(L, P, pairs, p0, nb, prm, want_samples);
        g_trace_enqueue += std::chrono::duration<double>(std::chrono::steady_clock::now() - t_e).count();
        if (rc == MPL_OK) { L.busy = 1; inflight_p0[li] = p0; inflight_nb[li] = nb; }
    }
    for (int j = 0; j < n_lanes; ++j) {                             // drain in submission order
        const int lane = (n_sub + j) % n_lanes;
        MplLane& L = ctx->lanes[lane];
        if (!L.busy) continue;
        if (rc == MPL_OK) rc = finish_sub_batch(L, P, pairs, inflight_p0[lane], inflight_nb[lane], want_samples, results);
        else cudaStreamSynchronize(L.stream);
        L.busy = 0;
    }
    if (getenv("MPL_PAIR_TRACE")) {
        g_trace_total = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_call).count();
        fprintf(stderr, "mpl_match_pose_batch: %d pairs, %d sub-batches of %d on %d lanes: %.1f ms; host thread: enqueue %.1f ms, waiting for the GPU %.1f ms, copy-out %.1f ms\n",
                n_pairs, n_sub, bs, n_lanes, g_trace_total * 1e3, g_trace_enqueue * 1e3, g_trace_wait * 1e3, (g_trace_total - g_trace_enqueue - g_trace_wait) * 1e3);
    }
    return rc;
}

// One pair = a batch of one (same kernels, same results); see include/mplb200.h.
extern "C" int mpl_match_pose_pair(const uint8_t* desc1, int n1, size_t step1, const uint8_t* desc2, int n2, size_t step2, int nbytes,
                                   const double* pts1, const double* pts2, const char* matcher_name, int ratio_test,
                                   double th, int h, double conf, int refine, unsigned long long seed,
                                   mpl_dmatch* matches, int* n_matches, int32_t* samples_out, int* n_samples,
                                   double* E, uint8_t* mask, int* n_inl, int* iters_run)
{
    if (!desc1 || !desc2 || !pts1 || !pts2 || !matcher_name || !matches || !n_matches || !E || !n_inl || n1 < 0 || n2 < 0 || nbytes <= 0 ||
        h < 0 || step1 < (size_t)nbytes || step2 < (size_t)nbytes) { mpl_set_error("mpl_match_pose_pair: bad argument"); return MPL_ERR_BAD_ARG; }
    *n_matches = 0; *n_inl = 0;
    if (n_samples) *n_samples = 0;
    if (iters_run) *iters_run = 0;
    if (n1 == 0 || n2 == 0) return MPL_OK;
    mpl_pair_io io; memset(&io, 0, sizeof(io));
    io.desc1 = desc1; io.desc2 = desc2; io.step1 = step1; io.step2 = step2; io.pts1 = pts1; io.pts2 = pts2; io.n1 = n1; io.n2 = n2;
    io.seed = seed; io.matches = matches; io.mask = mask; io.samples = samples_out;
    mpl_pair_params prm; memset(&prm, 0, sizeof(prm));
    prm.cv_type = 0; prm.cols = nbytes; prm.matcher_name = matcher_name; prm.ratio_test = ratio_test; prm.pts_mode = 0;
    prm.th = th; prm.h = h; prm.conf = conf; prm.refine = refine;
    mpl_pair_result res;
    const int rc = mpl_match_pose_batch(&io, 1, &prm, &res);
    if (rc) return rc;
    *n_matches = res.n_matches; *n_inl = res.n_inl;
    if (n_samples) *n_samples = res.n_samples;
    if (iters_run) *iters_run = res.iters_run;
    if (res.n_inl > 0) memcpy(E, res.E, 72);
    return MPL_OK;
}
