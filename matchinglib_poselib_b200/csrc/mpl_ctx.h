// mpl_ctx.h -- process-wide device context of libmplb200 (one device, one stream, growable scratch).
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <mutex>
#include "mpl_common.cuh"

struct MplBuf {
    void*  ptr = nullptr;
    size_t cap = 0;
    int  ensure(size_t bytes);   // grow (never shrinks); contents are NOT preserved
    void release();
};

enum { MPL_BUF_Q = 0, MPL_BUF_T, MPL_BUF_RAW, MPL_BUF_PART, MPL_BUF_KEYS, MPL_BUF_IDX, MPL_BUF_DIST,
       MPL_BUF_R0, MPL_BUF_R1, MPL_BUF_R2, MPL_BUF_R3, MPL_BUF_R4, MPL_BUF_R5, MPL_BUF_COUNT };

// One of the two pipeline lanes of the pair-batch path (pipeline.cu): its own stream, one device arena and one pinned
// host area, so that the uploads of sub-batch k+1 overlap the kernels of sub-batch k.
struct MplLane {
    cudaStream_t stream = nullptr;
    MplBuf arena;
    void*  host = nullptr;       // cudaHostAlloc: pair table (H2D source) + results (D2H destination)
    size_t host_cap = 0;
    int    busy = 0;             // a sub-batch is in flight
    int  ensure_host(size_t bytes);
    void release();
};

struct MplCtx {
    bool ready = false;
    int device = -1, sm_count = 148, cc_major = 0, cc_minor = 0;
    size_t total_mem = 0;
    cudaStream_t stream = nullptr;
    MplBuf bufs[MPL_BUF_COUNT];
    MplLane lanes[4];               // MPL_MAX_LANES; pipeline.cu uses the first n_lanes of them
};

MplCtx*     mpl_ctx();
std::mutex& mpl_ctx_mutex();
int         mpl_ensure_ctx_locked();   // caller holds mpl_ctx_mutex()
// The calling thread's pinned host staging area, grown to at least `bytes` (contents are NOT preserved when it grows).
// Device -> host reads that land here are true asynchronous copies; a copy into pageable memory (a std::vector, the stack)
// is staged by the driver and blocks inside the call.
int         mpl_stage_ensure(size_t bytes, void** ptr);
int         mpl_stage2_ensure(size_t bytes, void** ptr);   // a second area with the same rules

// ---- knn_l2_tc.cu: tcgen05 ranking kernel + exact re-rank --------------------------------------
bool   mpl_l2tc_supported(int nq, int nt, int dim);
size_t mpl_l2tc_scratch_bytes(int nq, int nt, int dim, int sm_count);
int    mpl_l2tc_top2(const float* d_q, int nq, size_t q_ld, const float* d_t, int nt, size_t t_ld, int dim, int sqrt_out,
                     long long train_offset, void* d_scratch, size_t scratch_bytes, mpl_key64* d_keys_out,
                     int sm_count, cudaStream_t st);
int    mpl_l2tc_last_fallbacks(const void* d_scratch, unsigned* count, unsigned* tf32_mode, cudaStream_t st);
bool   mpl_host_is_mapped(const void* p, size_t bytes);   // capi.cu: inside an mpl_host_alloc region (device-readable pinned memory)
int    mpl_hamming_path();             // capi.cu: mpl_set_hamming_path (0 auto, 1 popc, 2 tensor cores)
// the same for a batch of pairs with a common geometry in one ranking launch per group of pairs (dims[pair] = (n1, n2))
bool   mpl_hamming_tc_batch_supported(int nq_stride, int nt_stride, int w16);
size_t mpl_hamming_tc_batch_scratch_bytes(int nq_stride, int nt_stride, int w16, int nb, int sm_count);
int    mpl_hamming_tc_top2_batch(const uint8_t* d_q, int nq_stride, const uint8_t* d_t, int nt_stride, int w16, int nb, const int2* d_dims,
                                 void* d_scratch, size_t scratch_bytes, mpl_key64* d_keys_out, int sm_count, cudaStream_t st, int fmt);
// Hamming distances as a bf16 contraction on the same ranking kernel (128 / 256 / 512-bit rows: w16 = 1, 2, 4)
bool   mpl_hamming_tc_supported(int nq, int nt, int w16);
size_t mpl_hamming_tc_scratch_bytes(int nq, int nt, int w16, int sm_count);
int    mpl_hamming_tc_top2(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int w16, long long train_offset, void* d_scratch,
                           size_t scratch_bytes, mpl_key64* d_keys_out, int sm_count, cudaStream_t st, int fmt /* 2 = e4m3, 1 = bf16 */);
