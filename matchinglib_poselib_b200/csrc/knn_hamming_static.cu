// knn_hamming_static.cu -- mpl_knn2_hamming_static_dev: mpl_knn2_hamming_dev (capi.cu) for callers that cannot allocate device
// memory and have none to lend: CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// This entry point has no counterpart in the reference.
//
// The train-split path of the Hamming top-2 kernel (knn_hamming.cu: hamming_top2_kernel writes [split][query][2] partial keys,
// merge_top2_kernel folds them) needs scratch for those partial keys.  Here the scratch is a file-scope device array of
// MPL_KNN_STATIC_SCRATCH_BYTES (4 MiB) that every GPU loading the library holds until the process exits.  The call makes no
// cudaMalloc / cudaFree and calls no allocating library routine: launches, one cudaGetSymbolAddress per device, and (for an
// empty train set) a cudaMemsetAsync on the caller's output.
//
// The array's name is a host-side stand-in in host code, not a device address; the address is fetched once per device, with
// that device current, by cudaGetSymbolAddress and cached.  All calls on a device share the array: it keeps the previous call's
// partial keys, which is harmless because the top-2 kernel overwrites every [split][query][2] entry the merge kernel reads --
// nothing needs clearing -- but two calls that overlap on the GPU (different streams, or concurrent host threads) RACE on it.
// Serialise them, or use mpl_knn2_hamming_dev with one scratch buffer per stream.
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "../../include/mplb200.h"

#include <algorithm>
#include <mutex>

namespace {

static __device__ __align__(256) unsigned char g_knn_static_scratch[MPL_KNN_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
void* g_scratch_addr[kMaxDevices] = {};
std::mutex g_scratch_mutex;

// Device address of g_knn_static_scratch on the CURRENT device (set-up once per device).
int static_scratch_ptr(void** out)
{
    int dev = -1;
    MPL_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) { mpl_set_error("hamming_static: device ordinal %d out of range", dev); return MPL_ERR_LIMIT; }
    std::lock_guard<std::mutex> g(g_scratch_mutex);
    if (!g_scratch_addr[dev]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_knn_static_scratch);
        if (e != cudaSuccess || !p) { mpl_set_error("hamming_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e)); return MPL_ERR_CUDA; }
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return MPL_OK;
}

}  // namespace

extern "C" size_t mpl_knn_hamming_static_max_bytes(void) { return (size_t)MPL_KNN_STATIC_SCRATCH_BYTES; }

extern "C" int mpl_knn2_hamming_static_dev(const void* d_q, int nq, const void* d_t, int nt, int w16, long long train_offset,
                                           void* d_keys_out, void* stream)
{
    if (nq <= 0) return MPL_OK;
    if (!d_q || !d_keys_out || (nt > 0 && !d_t)) { mpl_set_error("hamming_static: null pointer"); return MPL_ERR_BAD_ARG; }
    cudaStream_t st = (cudaStream_t)stream;
    if (nt <= 0) { MPL_CUDA_CHECK(cudaMemsetAsync(d_keys_out, 0xFF, (size_t)nq * 2 * sizeof(mpl_key64), st)); return MPL_OK; }
    const MplCtx* ctx = mpl_ctx();
    const int sm = ctx->ready ? ctx->sm_count : 148;
    int splits = mpl_hamming_pick_splits(nq, nt, w16, sm);                          // the split count mpl_knn2_hamming_dev would use
    const size_t per_split = (size_t)nq * 2 * sizeof(mpl_key64);
    if ((size_t)splits * per_split > (size_t)MPL_KNN_STATIC_SCRATCH_BYTES) {
        // Fewer splits give the same keys (the merge is an exact minimum over (distance, train index)), only less parallelism:
        // take the largest reachable count that fits, never below ceil(nt / 2^20) (a split's rows are indexed with 20 bits).
        const int min_splits = (int)(((long long)nt + (1 << 20) - 1) >> 20);
        int fit = (int)std::min<size_t>((size_t)MPL_KNN_STATIC_SCRATCH_BYTES / per_split, (size_t)splits);
        for (; fit >= 2 && fit >= min_splits; --fit) { const int rps = (nt + fit - 1) / fit; if ((nt + rps - 1) / rps == fit) break; }
        if (fit >= 2 && fit >= min_splits) splits = fit;
        else if (min_splits <= 1) splits = 1;
        else {
            mpl_set_error("hamming_static: %d train rows need at least %d splits, and %d splits x %d queries need %zu bytes of partial keys; the "
                          "static scratch holds %zu (MPL_KNN_STATIC_SCRATCH_BYTES); use mpl_knn2_hamming_dev with caller scratch",
                          nt, min_splits, min_splits, nq, (size_t)min_splits * per_split, (size_t)MPL_KNN_STATIC_SCRATCH_BYTES);
            return MPL_ERR_LIMIT;
        }
    }
    if (splits == 1)
        return mpl_launch_hamming_top2((const uint8_t*)d_q, nq, (const uint8_t*)d_t, nt, w16, 1, 1, train_offset, (mpl_key64*)d_keys_out, st);
    void* scratch = nullptr;
    int rc = static_scratch_ptr(&scratch);
    if (rc) return rc;
    rc = mpl_launch_hamming_top2((const uint8_t*)d_q, nq, (const uint8_t*)d_t, nt, w16, splits, 1, train_offset, (mpl_key64*)scratch, st);
    if (rc) return rc;
    return mpl_launch_merge_top2((const mpl_key64*)scratch, splits, nq, 2, nullptr, nullptr, -1, (mpl_key64*)d_keys_out, st);
}
