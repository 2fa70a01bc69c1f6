// mpl_common.cuh -- shared declarations for the sm_100a kernels behind the C-ABI (include/mplb200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <stdio.h>

// ---- status codes of the C-ABI (negative = failure; mirrors include/mplb200.h) -----------------
#define MPL_OK              0
#define MPL_ERR_BAD_ARG    -1
#define MPL_ERR_CUDA      -100
#define MPL_ERR_NO_DEVICE -101
#define MPL_ERR_LIMIT     -102
#define MPL_ERR_ASSERT    -103   /* a CV_Assert of the reference would have thrown */

void mpl_set_error(const char* fmt, ...);
void mpl_count_launch();   // bumps the counter behind mpl_kernel_launch_count()

// ---- per-kernel-class device timers (CUDA events on the launching stream; off by default) -------
// bench.py enables them to time the dominant kernel of each path live (roofline.achieved).
enum { MPL_PROF_HAMMING = 0, MPL_PROF_L2_EXACT, MPL_PROF_L2_TC, MPL_PROF_L2_RERANK, MPL_PROF_MERGE,
       MPL_PROF_SOLVE, MPL_PROF_SCORE, MPL_PROF_L2_PREP, MPL_PROF_COUNT };
void mpl_prof_begin(int id, cudaStream_t st);
void mpl_prof_end(int id, cudaStream_t st);
struct MplProfScope {
    int id; cudaStream_t st;
    MplProfScope(int id_, cudaStream_t st_) : id(id_), st(st_) { mpl_prof_begin(id, st); }
    ~MplProfScope() { mpl_prof_end(id, st); }
};

#define MPL_CUDA_CHECK(expr)                                                                   \
    do {                                                                                       \
        cudaError_t _e = (expr);                                                               \
        if (_e != cudaSuccess) {                                                               \
            mpl_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return MPL_ERR_CUDA;                                                               \
        }                                                                                      \
    } while (0)

// A top-2 candidate is ordered by the 64-bit key (dist << 32) | trainIdx: unsigned comparison of
// keys == lexicographic (dist, trainIdx), the order of cvflann's KNNUniqueResultSet (ties -> lowest
// train index).  Float distances are non-negative, so their IEEE bit patterns order like uints.
typedef unsigned long long mpl_key64;
#define MPL_KEY_INVALID 0xFFFFFFFFFFFFFFFFull

// ---- knn_hamming.cu ---------------------------------------------------------------------------
// rows: w16 x 16-byte units per descriptor (zero padded), row stride = w16*16 bytes.
int mpl_launch_pack_rows_u8(const uint8_t* src, int n, size_t src_step, int nbytes_used, uint8_t* dst, int w16,
                            cudaStream_t st);
// partial[(split*nq + q)*2 + {0,1}] = best two keys of query q inside train split `split`.
// pair_dims (device, optional): int2 (nq, nt) per pair of a batch stored at the common strides nq / nt rows (n_splits = 1)
int mpl_launch_hamming_top2(const uint8_t* q, int nq, const uint8_t* t, int nt, int w16, int n_splits, int n_pairs,
                            long long train_offset, mpl_key64* partial, cudaStream_t st, const void* pair_dims = nullptr);
int mpl_hamming_pick_splits(int nq, int nt, int w16, int sm_count);
// merge n_parts partial lists per query (layout [part][nq][2]) into idx/dist (k = 1|2 columns).
int mpl_launch_merge_top2(const mpl_key64* partial, int n_parts, int nq, int k, int32_t* idx, void* dist,
                          int float_sqrt, mpl_key64* merged_keys, cudaStream_t st);
// all-gather over peer memory fused with the merge (knn_hamming.cu: allgather_merge_kernel)
int mpl_launch_allgather_merge(const mpl_key64* local_keys, int nq, int rank, int world, void* const* peer_gather, void* const* peer_sync,
                               int k, int32_t* idx, void* dist, int float_sqrt, mpl_key64* merged, cudaStream_t st);
// ratio test + cv::DMatch compaction of merged keys [nq][2]: out = 16-byte header {count} + up to nq 16-byte matches.
int mpl_launch_ratio_dmatch(const mpl_key64* keys, int nq, int ratio_test, int is_float, void* out, cudaStream_t st);

// ---- knn_l2.cu --------------------------------------------------------------------------------
int mpl_launch_l2_exact_top2(const float* q, int nq, size_t q_ld, const float* t, int nt, size_t t_ld, int dim,
                             int order_mode, int n_splits, long long train_offset, mpl_key64* partial,
                             cudaStream_t st);
int mpl_l2_pick_splits(int nq, int nt, int sm_count);
size_t mpl_l2_list_scratch_entries(int nq, int sm_count);
int mpl_launch_l2_exact_list(const float* q, const int* qlist, const unsigned* count, size_t q_ld, const float* t, int nt, size_t t_ld,
                             int dim, int order_mode, long long train_offset, mpl_key64* partial, mpl_key64* keys_out, int sm_count,
                             unsigned* ticket /* zeroed by the caller */, cudaStream_t st);
int mpl_launch_l2_rerank(const float* q, int nq, size_t q_ld, const float* t, size_t t_ld, int dim,
                         const int32_t* cand, int n_cand, int order_mode, long long train_offset,
                         mpl_key64* out, cudaStream_t st);
