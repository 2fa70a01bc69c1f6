// pose_host.cu -- host side of the pose half of the hot path (C ABI in include/mplb200.h):
//   mpl_ransac5pt              runRANSAC over pre-drawn (or internally drawn) minimal samples
//   mpl_estimate_essential_api poselib::estimateEssentialMat(..., "RANSAC", ...) over raw arrays
//   mpl_ransac_*_dev           device-pointer stages for the benchmark / device-resident pipelines
// Host logic here is what the reference also does sequentially and cheaply: the model selection replay
// (modelest.cpp:394-415), cvRANSACUpdateNumIters1 (:86-109), getSubset (:567-610).  Every O(models x correspondences) and
// O(samples) computation, and the least-squares refit, runs in ransac5pt.cu / ransac_batch.cu on the device.  (The pair
// batch path replays the selection on the device too: ransac_batch.cu replay_chunk_kernel; this host replay is the
// independent statement of the same rule that the tests cross-check it against.)
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "mpl_ransac.h"
#include "../../include/mplb200.h"

#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

// Experiment switch (default OFF): the minimal solver leaves cv::solvePoly's Durand-Kerner loop once a sweep moved every root
// by < 2^-40 relative instead of running all 300 sweeps.  Not the shipping default: see DESIGN.md section 4.4.
static std::atomic<int> g_solver_early{0};
bool mpl_solver_early_exit() { return g_solver_early.load(std::memory_order_relaxed) != 0; }
extern "C" int mpl_set_solver_early_exit(int on) { g_solver_early.store(on ? 1 : 0, std::memory_order_relaxed); return MPL_OK; }

// cvRANSACUpdateNumIters1 (modelest.cpp:86-109)
static int update_num_iters(double p, double ep, int model_points, int max_iters)
{
    p = std::fmax(p, 0.); p = std::fmin(p, 1.);
    ep = std::fmax(ep, 0.); ep = std::fmin(ep, 1.);
    double num = std::fmax(1. - p, DBL_MIN);
    double denom = 1. - std::pow(1. - ep, model_points);
    if (denom < DBL_MIN) return 0;
    num = std::log(num); denom = std::log(denom);
    return denom >= 0 || -num >= (double)max_iters * (-denom) ? max_iters : (int)std::round(num / denom);
}

struct Replay { int best_model = -1; int best_count = 0; int iters_run = 0; double errminsum = DBL_MAX; };

// modelest.cpp:377-416 applied to precomputed per-model (count, errsum); models of sample s are
// offsets[s] .. offsets[s+1]-1 in solver order.
static Replay replay_selection(const int32_t* counts, const double* errsums, const int32_t* offsets, int h, int n,
                               double conf, int max_iters)
{
    Replay r;
    int niters = max_iters;
    int it = 0;
    for (; it < niters && it < h; ++it) {
        for (int m = offsets[it]; m < offsets[it + 1]; ++m) {
            const int good = counts[m];
            if (good > std::max(r.best_count, 4)) {
                r.best_model = m; r.best_count = good;
                niters = update_num_iters(conf, (double)(n - good) / n, 5, niters);
                r.errminsum = errsums[m];
            } else if (good == std::max(r.best_count, 5) && r.errminsum < DBL_MAX && r.errminsum > errsums[m]) {
                r.best_model = m; r.errminsum = errsums[m];
            }
        }
    }
    r.iters_run = it;
    return r;
}

// ---- minimal-sample drawing: getSubset + checkSubset (modelest.cpp:567-650) ---------------------------
static thread_local uint64_t g_seed = 0x9E3779B97F4A7C15ull;   // per calling thread, like the contexts
extern "C" int mpl_ransac_set_seed(unsigned long long seed) { g_seed = seed ? seed : 0x9E3779B97F4A7C15ull; return MPL_OK; }
static inline uint32_t next_rand(uint64_t& s) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return (uint32_t)(s >> 16) & 0x7FFFFFFF; }

// checkSubset (modelest.cpp:613-650) is called with checkPartialSubsets = true (set in the constructor, :57), i.e. with
// i0 = i1 = count - 1, and ends in `return i >= i1`: whether or not its collinearity loop breaks, i == i1 at that point, so
// the reference ACCEPTS EVERY SUBSET and getSubset reduces to "5 distinct indices" (its 300-attempt failure path is dead
// code for count >= 5).  Reproduced as is: no collinearity rejection here either (round 1 rejected, which truncated runs on
// real match lists with many-to-one matches where the reference proceeds).
// Draws h samples (always succeeds for n >= 5; *n_drawn = h).
extern "C" int mpl_ransac_draw_samples(const double* p1, const double* p2, int n, int h, int32_t* samples, int* n_drawn)
{
    if (!p1 || !p2 || !samples || !n_drawn || n < 5 || h < 0) { mpl_set_error("draw_samples: bad argument"); return MPL_ERR_BAD_ARG; }
    uint64_t s = g_seed;
    for (int d = 0; d < h; ++d) {
        int idx[5];
        for (int i = 0; i < 5;) {
            const int c = (int)(next_rand(s) % (uint32_t)n);      // modelest.cpp:585 idx_i = rand() % count
            int j = 0;
            for (; j < i; ++j) if (c == idx[j]) break;
            if (j < i) continue;                                   // :586-590 duplicate: draw again
            idx[i++] = c;
        }
        for (int j = 0; j < 5; ++j) samples[5 * d + j] = idx[j];
    }
    g_seed = s;
    *n_drawn = h;
    return MPL_OK;
}

// ---- device-pointer stages -----------------------------------------------------------------------------
extern "C" int mpl_ransac_pack_corr_dev(const void* d_p1, const void* d_p2, int n, void* d_corr, void* stream)
{
    return mpl_launch_pack_corr((const double*)d_p1, (const double*)d_p2, n, d_corr, (cudaStream_t)stream);
}
extern "C" int mpl_ransac_solve_dev(const void* d_corr, int n, const void* d_samples, int h, void* d_models, void* d_n_models,
                                    void* d_offsets, void* d_dense, void* stream)
{
    int rc = mpl_launch_fivept_solve(d_corr, n, (const int32_t*)d_samples, h, (double*)d_models, (int32_t*)d_n_models, (cudaStream_t)stream);
    if (rc) return rc;
    return mpl_launch_compact_models((const double*)d_models, (const int32_t*)d_n_models, h, (int32_t*)d_offsets, (double*)d_dense, (cudaStream_t)stream);
}
extern "C" int mpl_ransac_score_dev(const void* d_dense, int m, const void* d_corr, int n, double th, void* d_counts, void* d_errsums, void* stream)
{
    return mpl_launch_sampson_score((const double*)d_dense, m, d_corr, n, th, (int32_t*)d_counts, (double*)d_errsums, (cudaStream_t)stream);
}

// ---- runRANSAC ---------------------------------------------------------------------------------------
// Scratch layout of one RANSAC run inside the context's buffers (n correspondences, h samples).
struct RansacBufs {
    double *d_p1, *d_p2; void* d_corr; int32_t *d_samples, *d_nm, *d_off; double *d_models, *d_dense, *d_errsums;
    int32_t* d_counts; double* d_misc; uint8_t* d_mask;
};
static int ransac_bufs_locked(MplCtx* ctx, int n, int h_use, RansacBufs& R)
{
    int rc;
    MplBuf* B = ctx->bufs;
    if ((rc = B[MPL_BUF_R0].ensure((size_t)n * 32 + (size_t)n * 32))) return rc;     // p1|p2 staging
    if ((rc = B[MPL_BUF_R1].ensure((size_t)n * 32))) return rc;                      // corr
    if ((rc = B[MPL_BUF_R2].ensure((size_t)h_use * 20 + (size_t)(h_use + 1) * 8))) return rc;   // samples | n_models | offsets
    if ((rc = B[MPL_BUF_R3].ensure((size_t)h_use * 90 * 8))) return rc;              // models [h][10][9]
    if ((rc = B[MPL_BUF_R4].ensure((size_t)h_use * 90 * 8))) return rc;              // dense models
    if ((rc = B[MPL_BUF_R5].ensure((size_t)h_use * 10 * 12 + (size_t)n * 8 + 8192))) return rc;   // counts | errsums | misc (state, basis, refit models) | mask
    R.d_p1 = (double*)B[MPL_BUF_R0].ptr; R.d_p2 = R.d_p1 + 2 * (size_t)n;
    R.d_corr = B[MPL_BUF_R1].ptr;
    R.d_samples = (int32_t*)B[MPL_BUF_R2].ptr;
    R.d_nm = R.d_samples + 5 * (size_t)h_use;
    R.d_off = R.d_nm + h_use;
    R.d_models = (double*)B[MPL_BUF_R3].ptr;
    R.d_dense = (double*)B[MPL_BUF_R4].ptr;
    R.d_errsums = (double*)B[MPL_BUF_R5].ptr;
    R.d_counts = (int32_t*)(R.d_errsums + (size_t)h_use * 10);
    R.d_misc = (double*)(((uintptr_t)(R.d_counts + (size_t)h_use * 10) + 63) & ~(uintptr_t)63);  // 64-byte aligned: state (256 B) | basis[36] | models[90] | nm
    R.d_mask = (uint8_t*)(R.d_misc + 32 + 36 + 90 + 8);
    return MPL_OK;
}

// runRANSAC from the solve on: correspondences (R.d_corr) and samples (R.d_samples) are already on the device.
int mpl_ransac_run_locked(MplCtx* ctx, int n, int h_use, double th, int max_iters, double conf, int refine, double* E, uint8_t* mask,
                          int* n_inl, int* iters_run, double* models_out, int32_t* n_models_out, int32_t* counts_out, double* errsums_out);

int mpl_ransac_device_inputs_locked(MplCtx* ctx, int n, int h_use, void** d_corr, int32_t** d_samples)
{
    RansacBufs R;
    const int rc = ransac_bufs_locked(ctx, n, h_use, R);
    if (rc) return rc;
    *d_corr = R.d_corr; *d_samples = R.d_samples;
    return MPL_OK;
}

extern "C" int mpl_ransac5pt(const double* p1, const double* p2, int n, const int32_t* samples, int h, double th,
                             int max_iters, double conf, int refine, double* E, uint8_t* mask, int* n_inl, int* iters_run,
                             double* models_out, int32_t* n_models_out, int32_t* counts_out, double* errsums_out)
{
    if (!p1 || !p2 || n < 5 || h < 0 || !E || !n_inl || max_iters < 0) { mpl_set_error("mpl_ransac5pt: bad argument"); return MPL_ERR_BAD_ARG; }
    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx();
    cudaStream_t st = ctx->stream;
    *n_inl = 0;
    if (iters_run) *iters_run = 0;
    std::vector<int32_t> drawn;
    if (!samples) {                                           // draw like getSubset does
        h = max_iters;
        drawn.resize((size_t)5 * (h > 0 ? h : 1));
        int nd = 0;
        if ((rc = mpl_ransac_draw_samples(p1, p2, n, h, drawn.data(), &nd))) return rc;
        h = nd; samples = drawn.data();
    }
    const bool want_all = models_out || n_models_out || counts_out || errsums_out;
    const int h_use = want_all ? h : std::min(h, max_iters);   // later samples can never be reached by the loop
    if (h_use == 0) return MPL_OK;
    RansacBufs R;
    if ((rc = ransac_bufs_locked(ctx, n, h_use, R))) return rc;
    MPL_CUDA_CHECK(cudaMemcpyAsync(R.d_p1, p1, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(R.d_p2, p2, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(R.d_samples, samples, (size_t)h_use * 20, cudaMemcpyHostToDevice, st));
    if ((rc = mpl_launch_pack_corr(R.d_p1, R.d_p2, n, R.d_corr, st))) return rc;
    return mpl_ransac_run_locked(ctx, n, h_use, th, max_iters, conf, refine, E, mask, n_inl, iters_run, models_out, n_models_out, counts_out, errsums_out);
}

int mpl_ransac_run_locked(MplCtx* ctx, int n, int h_use, double th, int max_iters, double conf, int refine, double* E, uint8_t* mask,
                          int* n_inl, int* iters_run, double* models_out, int32_t* n_models_out, int32_t* counts_out, double* errsums_out)
{
    int rc;
    cudaStream_t st = ctx->stream;
    RansacBufs R;
    if ((rc = ransac_bufs_locked(ctx, n, h_use, R))) return rc;                      // no growth here: same sizes as the caller's
    void* d_corr = R.d_corr; int32_t* d_samples = R.d_samples; int32_t* d_nm = R.d_nm; int32_t* d_off = R.d_off;
    double* d_models = R.d_models; double* d_dense = R.d_dense; double* d_errsums = R.d_errsums; int32_t* d_counts = R.d_counts;
    double* d_misc = R.d_misc; uint8_t* d_mask = R.d_mask;
    *n_inl = 0;
    if (iters_run) *iters_run = 0;
    // every device -> host read of this function lands in the thread's pinned staging area:
    // | off (h+1) i32 | counts (<= 10 h) i32 | errsums (<= 10 h) f64 | gram partials | small (bestE, refit counts / sums) |
    const size_t m_max = (size_t)10 * h_use;
    const size_t o_cnt = (((size_t)h_use + 1) * 4 + 15) & ~(size_t)15, o_err = o_cnt + ((m_max * 4 + 15) & ~(size_t)15),
                 o_small = (o_err + m_max * 8 + 15) & ~(size_t)15;
    char* h_stage = nullptr;
    if ((rc = mpl_stage_ensure(o_small + 2 * sizeof(MplPairState) + 64, (void**)&h_stage))) return rc;
    int32_t* off = (int32_t*)h_stage; int32_t* counts = (int32_t*)(h_stage + o_cnt); double* errsums = (double*)(h_stage + o_err);
    if ((rc = mpl_launch_fivept_solve(d_corr, n, d_samples, h_use, d_models, d_nm, st))) return rc;
    if ((rc = mpl_launch_compact_models(d_models, d_nm, h_use, d_off, d_dense, st))) return rc;
    MPL_CUDA_CHECK(cudaMemcpyAsync(off, d_off, ((size_t)h_use + 1) * 4, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    const int m = off[h_use];
    if (m < 0 || (size_t)m > m_max) { mpl_set_error("ransac: model count %d out of range", m); return MPL_ERR_CUDA; }
    if (m > 0) {
        if ((rc = mpl_launch_sampson_score(d_dense, m, d_corr, n, th, d_counts, d_errsums, st))) return rc;
        MPL_CUDA_CHECK(cudaMemcpyAsync(counts, d_counts, (size_t)m * 4, cudaMemcpyDeviceToHost, st));
        MPL_CUDA_CHECK(cudaMemcpyAsync(errsums, d_errsums, (size_t)m * 8, cudaMemcpyDeviceToHost, st));
    }
    if (models_out) MPL_CUDA_CHECK(cudaMemcpyAsync(models_out, d_dense, (size_t)m * 72, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    if (n_models_out) for (int s = 0; s < h_use; ++s) n_models_out[s] = off[s + 1] - off[s];
    if (counts_out) memcpy(counts_out, counts, (size_t)m * 4);
    if (errsums_out) memcpy(errsums_out, errsums, (size_t)m * 8);

    Replay r = replay_selection(counts, errsums, off, h_use, n, conf, max_iters);
    if (iters_run) *iters_run = r.iters_run;
    if (r.best_model < 0) return MPL_OK;                      // runRANSAC returns false (maxGoodCount == 0)
    // inlier mask of the selected model and the least-squares refit (modelest.cpp:420-464) on the device: the same kernels
    // the pair-batch path uses (ransac_batch.cu), here for a batch of one.  The state record carries the selection.
    MplPairState* h_state = (MplPairState*)(h_stage + o_small);
    MplPairState* d_state = (MplPairState*)d_misc;
    double* d_basis = d_misc + 32; double* d_rmodels = d_basis + 36; int32_t* d_rnm = (int32_t*)(d_rmodels + 90);
    memset(h_state, 0, sizeof(*h_state));
    h_state->n = n; h_state->niters = max_iters; h_state->iters_run = r.iters_run; h_state->best_count = r.best_count; h_state->best_slot = 0;
    h_state->errminsum = r.errminsum;
    MPL_CUDA_CHECK(cudaMemcpyAsync(d_state, h_state, sizeof(MplPairState), cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(&d_state->E[0], d_dense + (size_t)r.best_model * 9, 72, cudaMemcpyDeviceToDevice, st));
    if ((rc = mpl_launch_refit(d_corr, (size_t)n, d_state, 1, th, refine ? 1 : 0, d_mask, (size_t)n, d_basis, d_rmodels, d_rnm, st))) return rc;
    MplPairState* h_out = h_state + 1;
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_out, d_state, sizeof(MplPairState), cudaMemcpyDeviceToHost, st));
    if (mask) MPL_CUDA_CHECK(cudaMemcpyAsync(mask, d_mask, (size_t)n, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    // modelest.cpp:443-445: a refit that yields no model makes runRANSAC return false (n_inl = 0; E and mask are then not
    // results -- findEssentialMat returns before writing them, five-point.cpp:117-118)
    *n_inl = h_out->n_inl;
    if (h_out->n_inl > 0) memcpy(E, h_out->E, 72);
    return MPL_OK;
}

// ---- runLMeDS (modelest.cpp:483-564) ------------------------------------------------------------------
// niters = round(log(1 - conf) / log(1 - (1 - 0.45)^5)) clamped to [3, max_iters] (:518-519; 134 for the 0.999 / 2000 that
// findEssentialMat passes, five-point.cpp:126): a FIXED number of minimal samples, each model's median of the n float errors,
// strict '<' keeps the first minimal model, then sigma = 2.5 * 1.4826 * (1 + 5 / (n - 5)) * sqrt(minMedian) (>= 0.001) and
// findInliers with it; result = count >= 5.  No refit.  One solve launch, one median launch (CTA per model), one mask launch.
static int lmeds_niters(double conf, int max_iters)
{
    int niters = (int)std::round(std::log(1. - conf) / std::log(1. - std::pow(1. - 0.45, 5.0)));
    niters = std::max(niters, 3);
    return std::min(niters, max_iters);
}

extern "C" int mpl_lmeds5pt(const double* p1, const double* p2, int n, const int32_t* samples, int h, double conf, int max_iters,
                            double* E, uint8_t* mask, int* n_inl, double* min_median, double* sigma_out, double* medians_out,
                            int32_t* n_models_out)
{
    if (!p1 || !p2 || n < 5 || h < 0 || !E || !n_inl || max_iters < 0) { mpl_set_error("mpl_lmeds5pt: bad argument"); return MPL_ERR_BAD_ARG; }
    *n_inl = 0;
    if (min_median) *min_median = DBL_MAX;
    if (n == 5) { mpl_set_error("mpl_lmeds5pt: n == 5 is the solver-only case of findEssentialMat (mpl_estimate_essential_api)"); return MPL_ERR_BAD_ARG; }
    std::lock_guard<std::mutex> g(mpl_ctx_mutex());
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    MplCtx* ctx = mpl_ctx();
    cudaStream_t st = ctx->stream;
    std::vector<int32_t> drawn;
    if (!samples) {
        h = lmeds_niters(conf, max_iters);
        drawn.resize((size_t)5 * (h > 0 ? h : 1));
        int nd = 0;
        if ((rc = mpl_ransac_draw_samples(p1, p2, n, h, drawn.data(), &nd))) return rc;
        h = nd; samples = drawn.data();
    }
    if (h == 0) return MPL_OK;
    RansacBufs R;
    if ((rc = ransac_bufs_locked(ctx, n, h, R))) return rc;
    MPL_CUDA_CHECK(cudaMemcpyAsync(R.d_p1, p1, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(R.d_p2, p2, (size_t)n * 16, cudaMemcpyHostToDevice, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(R.d_samples, samples, (size_t)h * 20, cudaMemcpyHostToDevice, st));
    if ((rc = mpl_launch_pack_corr(R.d_p1, R.d_p2, n, R.d_corr, st))) return rc;
    const size_t m_max = (size_t)10 * h;
    const size_t o_med = (((size_t)h + 1) * 4 + 15) & ~(size_t)15;
    char* h_stage = nullptr;
    if ((rc = mpl_stage_ensure(o_med + m_max * 8 + 64, (void**)&h_stage))) return rc;
    int32_t* off = (int32_t*)h_stage; double* med = (double*)(h_stage + o_med);
    if ((rc = mpl_launch_fivept_solve(R.d_corr, n, R.d_samples, h, R.d_models, R.d_nm, st))) return rc;
    if ((rc = mpl_launch_compact_models(R.d_models, R.d_nm, h, R.d_off, R.d_dense, st))) return rc;
    MPL_CUDA_CHECK(cudaMemcpyAsync(off, R.d_off, ((size_t)h + 1) * 4, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    const int m = off[h];
    if (m < 0 || (size_t)m > m_max) { mpl_set_error("lmeds: model count %d out of range", m); return MPL_ERR_CUDA; }
    if (n_models_out) for (int s = 0; s < h; ++s) n_models_out[s] = off[s + 1] - off[s];
    if (m == 0) return MPL_OK;
    double* d_med = R.d_errsums;                                     // m doubles: the slot the RANSAC path keeps its error sums in
    if ((rc = mpl_launch_lmeds_medians(R.d_dense, m, R.d_corr, n, d_med, st))) return rc;
    MPL_CUDA_CHECK(cudaMemcpyAsync(med, d_med, (size_t)m * 8, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    if (medians_out) memcpy(medians_out, med, (size_t)m * 8);
    double minMedian = DBL_MAX; int best = -1;
    for (int i = 0; i < m; ++i) if (med[i] < minMedian) { minMedian = med[i]; best = i; }      // modelest.cpp:544-548
    if (min_median) *min_median = minMedian;
    if (best < 0) return MPL_OK;                                     // every median NaN / inf: runLMeDS returns false
    double sigma = 2.5 * 1.4826 * (1 + 5. / (n - 5)) * std::sqrt(minMedian);                   // :553-554
    if (!(sigma > 0.001)) sigma = 0.001;
    if (sigma_out) *sigma_out = sigma;
    int* d_count = (int*)R.d_misc;
    if ((rc = mpl_launch_inlier_mask(R.d_dense + (size_t)best * 9, R.d_corr, n, sigma, R.d_mask, d_count, st))) return rc;
    int* h_count = (int*)(h_stage + o_med + m_max * 8);
    MPL_CUDA_CHECK(cudaMemcpyAsync(h_count, d_count, 4, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync(E, R.d_dense + (size_t)best * 9, 72, cudaMemcpyDeviceToHost, st));
    if (mask) MPL_CUDA_CHECK(cudaMemcpyAsync(mask, R.d_mask, (size_t)n, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    *n_inl = *h_count;                                               // the caller applies result = count >= 5 (:557)
    return MPL_OK;
}

// poselib::estimateEssentialMat (pose_estim.cpp:857-890) -> findEssentialMat (five-point.cpp:69-148), method "RANSAC".
// Returns 1 (true) / 0 (false) like the reference's bool, or a negative MPL_ERR_*.  E_out holds 9 doubles, or
// 9 * (*n_E) doubles (up to 90) when n == 5 (solver only, mask all ones, five-point.cpp:108-114).
extern "C" int mpl_estimate_essential_api(const double* p1, const double* p2, int n, const char* method, double threshold,
                                          int refine, double* E_out, int* n_E, uint8_t* mask)
{
    if (!method) { mpl_set_error("estimateEssentialMat: null method"); return MPL_ERR_BAD_ARG; }
    const std::string m(method);
    if (m != "RANSAC" && m != "ARRSAC" && m != "LMEDS") {
        // pose_estim.cpp:878-887: USAC / unknown methods print and exit(1)
        mpl_set_error("estimateEssentialMat: method %s is not served here (RANSAC, ARRSAC, LMEDS; USAC has its own entry point mpl_usac5pt; no CPU fallback)", method);
        return MPL_ERR_BAD_ARG;
    }
    if (n < 5 || !p1 || !p2) { mpl_set_error("CV_Assert(npoints >= 5 && points2.checkVector(2) == npoints)"); return MPL_ERR_ASSERT; }
    if (!E_out) return 0;                                      // five-point.cpp:137-144: E not requested -> false
    if (n_E) *n_E = 1;
    if (n == 5) {
        std::lock_guard<std::mutex> g(mpl_ctx_mutex());
        int rc = mpl_ensure_ctx_locked();
        if (rc) return rc;
        MplCtx* ctx = mpl_ctx(); cudaStream_t st = ctx->stream; MplBuf* B = ctx->bufs;
        if ((rc = B[MPL_BUF_R0].ensure(4096))) return rc;
        char* base = (char*)B[MPL_BUF_R0].ptr;
        double* d_p1 = (double*)base; double* d_p2 = d_p1 + 10; void* d_corr = base + 256; int32_t* d_s = (int32_t*)(base + 512);
        double* d_models = (double*)(base + 1024); int32_t* d_nm = (int32_t*)(base + 2048);
        const int32_t s[5] = {0, 1, 2, 3, 4};
        MPL_CUDA_CHECK(cudaMemcpyAsync(d_p1, p1, 80, cudaMemcpyHostToDevice, st));
        MPL_CUDA_CHECK(cudaMemcpyAsync(d_p2, p2, 80, cudaMemcpyHostToDevice, st));
        MPL_CUDA_CHECK(cudaMemcpyAsync(d_s, s, 20, cudaMemcpyHostToDevice, st));
        if ((rc = mpl_launch_pack_corr(d_p1, d_p2, 5, d_corr, st))) return rc;
        if ((rc = mpl_launch_fivept_solve(d_corr, 5, d_s, 1, d_models, d_nm, st))) return rc;
        int cnt = 0;
        MPL_CUDA_CHECK(cudaMemcpyAsync(&cnt, d_nm, 4, cudaMemcpyDeviceToHost, st));
        MPL_CUDA_CHECK(cudaStreamSynchronize(st));
        MPL_CUDA_CHECK(cudaMemcpyAsync(E_out, d_models, (size_t)cnt * 72, cudaMemcpyDeviceToHost, st));
        MPL_CUDA_CHECK(cudaStreamSynchronize(st));
        if (n_E) *n_E = cnt;
        if (mask) memset(mask, 1, 5);
        return 1;
    }
    int n_inl = 0, iters = 0;
    if (m == "ARRSAC") {                                       // pose_estim.cpp:866-869
        int rc = mpl_arrsac5pt(p1, p2, n, threshold, refine, E_out, mask, &n_inl, nullptr, nullptr);
        if (rc) return rc;
        return n_inl > 0 ? 1 : 0;
    }
    if (m == "LMEDS") {                                        // pose_estim.cpp:874-877 -> five-point.cpp:124-128: runLMeDS(prob 0.999, 2000), no refit
        int rc = mpl_lmeds5pt(p1, p2, n, nullptr, 0, 0.999, 2000, E_out, mask, &n_inl, nullptr, nullptr, nullptr, nullptr);
        if (rc) return rc;
        return n_inl >= 5 ? 1 : 0;
    }
    int rc = mpl_ransac5pt(p1, p2, n, nullptr, 0, threshold, 1000, 0.999, refine, E_out, mask, &n_inl, &iters,
                           nullptr, nullptr, nullptr, nullptr);
    if (rc) return rc;
    return n_inl > 0 ? 1 : 0;
}
