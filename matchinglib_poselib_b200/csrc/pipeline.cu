// pipeline.cu -- device-resident hand-off between the two halves of the hot path:
//   binary-descriptor k-NN  ->  Lowe ratio test + cv::DMatch compaction  ->  keypoint gather into RANSAC's
//   correspondence records  ->  minimal-sample drawing  ->  5-point solve + Sampson scoring  ->  selection.
// In the reference these are separate host steps with std::vector traffic in between (tests/poselib-test/main.cpp:
// 1522-1599: getCorrespondences -> ratio-filtered matches -> ImgToCamCoordTrans -> estimateEssentialMat).  Here the matches
// never leave the device between the scan and the scoring; the host sees the match list, the inlier mask and E at the end
// (one 8-byte read of the match count in the middle, needed to size the launches).
//
// Reference semantics kept:
//   ratio test + DMatch emission, ascending queryIdx   matchers.cpp:601-625 (LINEAR), nmslib_matchers.h:399-421
//   minimal samples: 5 distinct indices, every prefix passing checkSubset in both images, <= 300 attempts
//                                                      modelest.cpp:567-650 (the generator is NOT std::rand: the reference
//                                                      seeds with time(), modelest.cpp:58, so no sequence is promised)
//   everything from the solve on                       pose_host.cu (mpl_ransac_run_locked)
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "mpl_ransac.h"
#include "../../include/mplb200.h"

#include <float.h>
#include <string.h>

int mpl_hamming_keys_locked(const uint8_t* q, int nq, size_t q_step, const uint8_t* t, int nt, size_t t_step, int nbytes,
                            int ignore_tail_bytes, mpl_key64** d_keys);
int mpl_ransac_device_inputs_locked(MplCtx* ctx, int n, int h_use, void** d_corr, int32_t** d_samples);
int mpl_ransac_run_locked(MplCtx* ctx, int n, int h_use, double th, int max_iters, double conf, int refine, double* E, uint8_t* mask,
                          int* n_inl, int* iters_run, double* models_out, int32_t* n_models_out, int32_t* counts_out, double* errsums_out);

namespace {

struct PipeCounters { int n_matches; int first_failed_sample; };

// One CTA, order-preserving compaction: every warp owns a contiguous chunk of queries (ballot counts, block scan, write).
// Hamming keys are (dist << 32 | trainIdx); a query passes iff (float)d0 < 0.75f * (float)d1 (ratio != 0), or always.
// Writes the cv::DMatch records AND the 32-byte correspondence records (x1, y1, x2, y2) of the survivors.
__global__ void __launch_bounds__(1024)
ratio_compact_gather_kernel(const ulonglong2* __restrict__ keys, int nq, int ratio_test, const double2* __restrict__ pts1,
                            const double2* __restrict__ pts2, mpl_dmatch* __restrict__ matches, double4* __restrict__ corr,
                            PipeCounters* __restrict__ ctr)
{
    __shared__ int warp_sums[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int per = ((nq + 1023) / 1024) * 32;                   // queries per warp, lanes read consecutive keys
    const int start = min(nq, wid * per), end = min(nq, start + per);
    auto passes = [&](const ulonglong2 k) {
        if (k.x == MPL_KEY_INVALID) return false;
        if (!ratio_test) return true;
        if (k.y == MPL_KEY_INVALID) return false;
        return (float)(uint32_t)(k.x >> 32) < 0.75f * (float)(uint32_t)(k.y >> 32);
    };
    int sum = 0;
    for (int i = start; i < end; i += 32) {
        const bool p = (i + lane < end) && passes(keys[i + lane]);
        sum += __popc(__ballot_sync(0xffffffffu, p));
    }
    if (lane == 0) warp_sums[wid] = sum;
    __syncthreads();
    if (wid == 0) {
        int w = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
        warp_sums[lane] = w;
    }
    __syncthreads();
    int pos = wid ? warp_sums[wid - 1] : 0;
    for (int i = start; i < end; i += 32) {
        ulonglong2 k = make_ulonglong2(MPL_KEY_INVALID, MPL_KEY_INVALID);
        const bool p = (i + lane < end) && passes(k = keys[i + lane]);
        const unsigned b = __ballot_sync(0xffffffffu, p);
        if (p) {
            const int o = pos + __popc(b & ((1u << lane) - 1u));
            const int ti = (int)(uint32_t)(k.x & 0xFFFFFFFFull);
            mpl_dmatch m; m.queryIdx = i + lane; m.trainIdx = ti; m.imgIdx = -1; m.distance = (float)(uint32_t)(k.x >> 32);
            matches[o] = m;
            const double2 a = pts1[i + lane], c = pts2[ti];
            corr[o] = make_double4(a.x, a.y, c.x, c.y);
        }
        pos += __popc(b);
    }
    if (tid == 1023) { ctr->n_matches = pos; ctr->first_failed_sample = 0x7FFFFFFF; }
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t& s)
{
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// checkSubset for the LAST point of the prefix (modelest.cpp:613-650), on one image's coordinates
__device__ __forceinline__ bool subset_ok_dev(const double* x, const double* y, int count)
{
    const int i = count - 1;
    for (int j = 0; j < i; ++j) {
        const double dx1 = x[j] - x[i], dy1 = y[j] - y[i];
        for (int k = 0; k < j; ++k) {
            const double dx2 = x[k] - x[i], dy2 = y[k] - y[i];
            if (fabs(dx2 * dy1 - dy2 * dx1) <= (double)FLT_EPSILON * (fabs(dx1) + fabs(dy1) + fabs(dx2) + fabs(dy2))) return false;
        }
    }
    return true;
}

// thread = one minimal sample, its own counter-based random stream (seed, sample id): getSubset's rules
// (modelest.cpp:567-610).  A sample that cannot be completed within 300 attempts records its index; the host then
// truncates the run there, like the reference's loop ends when getSubset fails.
__global__ void draw_samples_kernel(const double4* __restrict__ corr, int n, int h, unsigned long long seed, int32_t* __restrict__ samples,
                                    PipeCounters* __restrict__ ctr)
{
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= h) return;
    uint64_t st = seed ^ (0xD1B54A32D192ED03ull * (uint64_t)(s + 1));
    int idx[5]; double x1[5], y1[5], x2[5], y2[5];
    int i = 0, iters = 0;
    const int max_attempts = 300;
    for (; iters < max_attempts; ++iters) {
        for (i = 0; i < 5 && iters < max_attempts;) {
            const int c = (int)(splitmix64(st) % (uint64_t)n);
            int j = 0;
            for (; j < i; ++j) if (c == idx[j]) break;
            if (j < i) continue;
            const double4 p = corr[c];
            idx[i] = c; x1[i] = p.x; y1[i] = p.y; x2[i] = p.z; y2[i] = p.w;
            if (!subset_ok_dev(x1, y1, i + 1) || !subset_ok_dev(x2, y2, i + 1)) { ++iters; continue; }
            ++i;
        }
        break;
    }
    if (!(i == 5 && iters < max_attempts)) { atomicMin(&ctr->first_failed_sample, s); for (int j = 0; j < 5; ++j) idx[j] = j; }
    for (int j = 0; j < 5; ++j) samples[5 * s + j] = idx[j];
}

}  // namespace

// See include/mplb200.h.
extern "C" int mpl_match_pose_pair(const uint8_t* desc1, int n1, size_t step1, const uint8_t* desc2, int n2, size_t step2, int nbytes,
                                   const double* pts1, const double* pts2, const char* matcher_name, int ratio_test,
                                   double th, int h, double conf, int refine, unsigned long long seed,
                                   mpl_dmatch* matches, int* n_matches, int32_t* samples_out, int* n_samples,
                                   double* E, uint8_t* mask, int* n_inl, int* iters_run)
{
    if (!desc1 || !desc2 || !pts1 || !pts2 || !matcher_name || !matches || !n_matches || !E || !n_inl || n1 < 0 || n2 < 0 || nbytes <= 0 ||
        h < 0 || step1 < (size_t)nbytes || step2 < (size_t)nbytes) { mpl_set_error("mpl_match_pose_pair: bad argument"); return MPL_ERR_BAD_ARG; }
    const bool linear = strcmp(matcher_name, "LINEAR") == 0, nms = strcmp(matcher_name, "BRUTEFORCENMS") == 0;
    if (!linear && !nms) { mpl_set_error("mpl_match_pose_pair: matcher %s is not served (no CPU fallback)", matcher_name); return MPL_ERR_BAD_ARG; }
    // BRUTEFORCENMS drops the last packed word: 2 bytes for even, 1 byte for odd descriptor lengths
    const int ignore_tail = nms ? ((nbytes & 1) ? 1 : 2) : 0;
    if (ignore_tail >= nbytes || nbytes - ignore_tail > 256) { mpl_set_error("mpl_match_pose_pair: unsupported descriptor length"); return MPL_ERR_LIMIT; }
    *n_matches = 0; *n_inl = 0;
    if (n_samples) *n_samples = 0;
    if (iters_run) *iters_run = 0;
    if (n1 == 0 || n2 == 0) return MPL_OK;
    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx(); cudaStream_t st = ctx->stream; MplBuf* B = ctx->bufs;
    // RANSAC scratch first, at its largest possible size (n1 correspondences), so that nothing is reallocated later
    void* d_corr = nullptr; int32_t* d_samples = nullptr;
    const int h_alloc = h > 0 ? h : 1;
    if ((rc = mpl_ransac_device_inputs_locked(ctx, n1 > 5 ? n1 : 5, h_alloc, &d_corr, &d_samples))) return rc;
    if ((rc = B[MPL_BUF_IDX].ensure(((size_t)n1 + n2) * 16))) return rc;                     // keypoints of both images
    if ((rc = B[MPL_BUF_DIST].ensure(256 + (size_t)n1 * sizeof(mpl_dmatch)))) return rc;     // counters | matches
    double2* d_pts1 = (double2*)B[MPL_BUF_IDX].ptr; double2* d_pts2 = d_pts1 + n1;
    PipeCounters* d_ctr = (PipeCounters*)B[MPL_BUF_DIST].ptr;
    mpl_dmatch* d_matches = (mpl_dmatch*)((char*)B[MPL_BUF_DIST].ptr + 256);
    MPL_CUDA_CHECK(cudaMemcpyAsync(d_pts1, pts1, (size_t)n1 * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(d_pts2, pts2, (size_t)n2 * 16, cudaMemcpyHostToDevice, st));
    mpl_key64* d_keys = nullptr;
    if ((rc = mpl_hamming_keys_locked(desc1, n1, step1, desc2, n2, step2, nbytes, ignore_tail, &d_keys))) return rc;
    ratio_compact_gather_kernel<<<1, 1024, 0, st>>>((const ulonglong2*)d_keys, n1, ratio_test ? 1 : 0, d_pts1, d_pts2, d_matches,
                                                    (double4*)d_corr, d_ctr);
    MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    // pinned staging of this call's results: | counters 256 B | matches n1 x 16 | samples h x 20 | mask n1 |  (true async copies;
    // handed to the caller's buffers by memcpy after the last wait)
    char* h_st = nullptr;
    const size_t o_m = 256, o_s = o_m + (size_t)n1 * sizeof(mpl_dmatch), o_k = o_s + (((size_t)h_alloc * 20 + 15) & ~(size_t)15);
    if ((rc = mpl_stage2_ensure(o_k + (size_t)n1 + 16, (void**)&h_st))) return rc;
    PipeCounters* h_c = (PipeCounters*)h_st;
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_c, d_ctr, sizeof(PipeCounters), cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));                                               // the one mid-pipeline read: 8 bytes
    PipeCounters c = *h_c;
    const int n = c.n_matches;
    *n_matches = n;
    if (n > 0) MPL_CUDA_CHECK(cudaMemcpyAsync(h_st + o_m, d_matches, (size_t)n * sizeof(mpl_dmatch), cudaMemcpyDeviceToHost, st));
    int h_use = 0;
    if (n >= 5 && h > 0) {
        draw_samples_kernel<<<(h + 127) / 128, 128, 0, st>>>((const double4*)d_corr, n, h, seed, d_samples, d_ctr);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        MPL_CUDA_CHECK(cudaMemcpyAsync(h_c, d_ctr, sizeof(PipeCounters), cudaMemcpyDeviceToHost, st));
        MPL_CUDA_CHECK(cudaStreamSynchronize(st));
        c = *h_c;
        h_use = c.first_failed_sample < h ? c.first_failed_sample : h;
        if (samples_out && h_use > 0) MPL_CUDA_CHECK(cudaMemcpyAsync(h_st + o_s, d_samples, (size_t)h_use * 20, cudaMemcpyDeviceToHost, st));
    }
    if (n_samples) *n_samples = h_use;
    if (h_use > 0) {
        // max_iters = h: the pre-drawn samples ARE the iteration budget (reference: 1000, five-point.cpp:117)
        rc = mpl_ransac_run_locked(ctx, n, h_use, th, h_use, conf, refine, E, mask ? (uint8_t*)(h_st + o_k) : nullptr, n_inl, iters_run,
                                   nullptr, nullptr, nullptr, nullptr);
        if (rc) return rc;
    }
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    if (n > 0) memcpy(matches, h_st + o_m, (size_t)n * sizeof(mpl_dmatch));
    if (samples_out && h_use > 0) memcpy(samples_out, h_st + o_s, (size_t)h_use * 20);
    if (mask && h_use > 0 && *n_inl > 0) memcpy(mask, h_st + o_k, (size_t)n);
    return MPL_OK;
}
