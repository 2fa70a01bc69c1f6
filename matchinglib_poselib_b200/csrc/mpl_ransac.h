// mpl_ransac.h -- launch wrappers of ransac5pt.cu (device pointers, enqueue only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

int mpl_launch_pack_corr(const double* p1, const double* p2, int n, void* corr /* n x 32 B */, cudaStream_t st);
int mpl_launch_fivept_solve(const void* corr, int n, const int32_t* samples, int h, double* models /* h x 90 */,
                            int32_t* n_models /* h */, cudaStream_t st);
int mpl_launch_fivept_from_basis(const double* basis /* 4 x 9 */, double* models /* 90 */, int32_t* n_models, cudaStream_t st);
int mpl_launch_compact_models(const double* models, const int32_t* n_models, int h, int32_t* offsets /* h+1 */,
                              double* dense /* up to 10h x 9 */, cudaStream_t st);
int mpl_launch_sampson_score(const double* models, int m, const void* corr, int n, double th, int32_t* counts,
                             double* errsums /* may be null */, cudaStream_t st);
int mpl_launch_sampson_mask(const double* model, const void* corr, int n, double th, uint8_t* mask, float* err, cudaStream_t st);
int mpl_gram_blocks(int n);
int mpl_launch_gram(const void* corr, const uint8_t* mask, int n, double* partial /* blocks x 45 */, cudaStream_t st);
