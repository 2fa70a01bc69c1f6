// mpl_ransac.h -- launch wrappers of ransac5pt.cu (device pointers, enqueue only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

bool mpl_solver_early_exit();   // pose_host.cu: mpl_set_solver_early_exit (experiment switch, default off)
int mpl_launch_pack_corr(const double* p1, const double* p2, int n, void* corr /* n x 32 B */, cudaStream_t st);
int mpl_launch_fivept_solve(const void* corr, int n, const int32_t* samples, int h, double* models /* h x 90 */,
                            int32_t* n_models /* h */, cudaStream_t st);
int mpl_launch_compact_models(const double* models, const int32_t* n_models, int h, int32_t* offsets /* h+1 */,
                              double* dense /* up to 10h x 9 */, cudaStream_t st);
int mpl_launch_sampson_score(const double* models, int m, const void* corr, int n, double th, int32_t* counts,
                             double* errsums /* may be null */, cudaStream_t st);

// LMedS: median of the n float Sampson errors of every model (modelest.cpp:537-542); inlier mask + count of one model
int mpl_launch_lmeds_medians(const double* models, int m, const void* corr, int n, double* medians, cudaStream_t st);
int mpl_launch_inlier_mask(const double* model, const void* corr, int n, double th, uint8_t* mask, int* count, cudaStream_t st);

// ---- ransac_batch.cu: device-driven RANSAC for a batch of pairs (no host round trip) -----------------------------------
// One record per pair.  Written by the gather kernel (pipeline.cu) or by the host (mpl_ransac5pt's refit), advanced by the
// replay / refit kernels, read back at the end.
struct MplPairState {
    int n1, n2;              // descriptor rows of the pair
    int n_matches;           // matches after the ratio test
    int n;                   // correspondences (matches that survived lens undistortion)
    int niters;              // adaptive iteration bound (starts at the sample budget h)
    int iters_run;
    int best_count;          // maxGoodCount
    int best_slot;           // sample * 10 + root of the model held in E; -1 = none
    int refit_nm;            // models of the least-squares refit (0: none / not run)
    int n_inl;               // final inlier count (0: runRANSAC returned false)
    int status;              // 0 ok; 1: Remove_LensDist failed (< 16 valid points)
    int pad;
    double errminsum;
    double E[9];
    unsigned long long seed;
};
int mpl_launch_draw_samples_batch(const MplPairState* st, int n_pairs, int h, int32_t* samples, cudaStream_t s);
int mpl_launch_solve_chunk(const void* corr, size_t corr_stride, const MplPairState* st, int n_pairs, const int32_t* samples, int h,
                           int s0, int c, double* models /* [pair][h][10][9] */, int32_t* n_models /* [pair][h] */, cudaStream_t s);
int mpl_launch_score_slots(const void* corr, size_t corr_stride, const MplPairState* st, int n_pairs, const double* models,
                           const int32_t* n_models, int h, int s0, int c, double th, int n_hint, int32_t* counts, double* errsums, cudaStream_t s);
int mpl_launch_replay_chunk(MplPairState* st, int n_pairs, const double* models, const int32_t* n_models, const int32_t* counts,
                            const double* errsums, int h, int s0, int c, double conf, cudaStream_t s);
// inlier mask of every pair's selected model, then (refine) the least-squares refit: basis [pair][36], rmodels [pair][90], rnm [pair]
int mpl_launch_refit(const void* corr, size_t corr_stride, MplPairState* st, int n_pairs, double th, int refine, uint8_t* mask,
                     size_t mask_stride, double* basis, double* rmodels, int32_t* rnm, cudaStream_t s);
