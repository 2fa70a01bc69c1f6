// capi.cu -- the C ABI (include/mplb200.h): context, scratch memory, host-pointer entry points and the
// getMatches mirror.  Host logic only; kernels live in knn_hamming.cu / knn_l2_exact.cu / knn_l2_tc.cu /
// ransac5pt.cu.  There is no CPU compute path here: every entry point needs a CUDA device.
#include "mpl_common.cuh"
#include "mpl_ctx.h"
#include "../../include/mplb200.h"

#include <atomic>
#include <cstdarg>
#include <map>
#include <sys/mman.h>
#include <cstring>
#include <cstdlib>
#include <climits>
#include <cfloat>
#include <mutex>
#include <string>
#include <vector>

// ------------------------------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void mpl_set_error(const char* fmt, ...)
{
    va_list ap; va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
void mpl_count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

// ---- device timers ----------------------------------------------------------------------------
struct ProfSlot { std::vector<std::pair<cudaEvent_t, cudaEvent_t>> used, pool; };
static ProfSlot g_prof[MPL_PROF_COUNT];
static std::atomic<int> g_prof_on{0};
static std::mutex g_prof_mu;

void mpl_prof_begin(int id, cudaStream_t st)
{
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> g(g_prof_mu);
    ProfSlot& s = g_prof[id];
    std::pair<cudaEvent_t, cudaEvent_t> ev;
    if (!s.pool.empty()) { ev = s.pool.back(); s.pool.pop_back(); }
    else { cudaEventCreate(&ev.first); cudaEventCreate(&ev.second); }
    cudaEventRecord(ev.first, st);
    s.used.push_back(ev);
}
void mpl_prof_end(int id, cudaStream_t st)
{
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    std::lock_guard<std::mutex> g(g_prof_mu);
    ProfSlot& s = g_prof[id];
    if (!s.used.empty()) cudaEventRecord(s.used.back().second, st);
}
extern "C" int mpl_prof_enable(int on) { g_prof_on.store(on ? 1 : 0); return MPL_OK; }
// Sum of the device durations recorded for kernel class `id` since the last read (waits for them).
extern "C" int mpl_prof_read(int id, double* total_ms, long long* launches)
{
    if (id < 0 || id >= MPL_PROF_COUNT) { mpl_set_error("prof_read: bad id"); return MPL_ERR_BAD_ARG; }
    std::lock_guard<std::mutex> g(g_prof_mu);
    ProfSlot& s = g_prof[id];
    double tot = 0; long long n = 0;
    for (auto& ev : s.used) {
        float ms = 0;
        if (cudaEventSynchronize(ev.second) == cudaSuccess && cudaEventElapsedTime(&ms, ev.first, ev.second) == cudaSuccess) { tot += ms; ++n; }
        s.pool.push_back(ev);
    }
    s.used.clear();
    (void)cudaGetLastError();
    if (total_ms) *total_ms = tot;
    if (launches) *launches = n;
    return MPL_OK;
}

// One context per HOST THREAD: its own stream, scratch buffers and pinned staging, created on first use on the device the
// last mpl_init() named.  Calls from different threads therefore run concurrently on the GPU (independent image pairs of a
// batch: the per-pair RANSAC is latency-bound and fills a fraction of the SMs); calls of one thread are synchronous.
static thread_local MplCtx g_ctx;
static thread_local std::mutex g_mu;            // the entry points keep their lock_guard; never contended across threads
static std::atomic<int> g_device{0};

MplCtx* mpl_ctx() { return &g_ctx; }
std::mutex& mpl_ctx_mutex() { return g_mu; }

int MplBuf::ensure(size_t bytes)
{
    if (bytes <= cap) return MPL_OK;
    if (ptr) { MPL_CUDA_CHECK(cudaFree(ptr)); ptr = nullptr; cap = 0; }
    size_t want = bytes + bytes / 4 + 256;
    MPL_CUDA_CHECK(cudaMalloc(&ptr, want));
    cap = want;
    return MPL_OK;
}
void MplBuf::release() { if (ptr) cudaFree(ptr); ptr = nullptr; cap = 0; }

int MplLane::ensure_host(size_t bytes)
{
    if (bytes <= host_cap) return MPL_OK;
    if (host) { cudaFreeHost(host); host = nullptr; host_cap = 0; }
    const size_t want = bytes + bytes / 4 + 4096;
    MPL_CUDA_CHECK(cudaHostAlloc(&host, want, cudaHostAllocDefault));
    host_cap = want;
    return MPL_OK;
}
void MplLane::release()
{
    if (stream) { cudaStreamSynchronize(stream); cudaStreamDestroy(stream); stream = nullptr; }
    arena.release();
    if (host) cudaFreeHost(host);
    host = nullptr; host_cap = 0; busy = 0;
}

// Pinned host staging for results: device -> host copies land here at full PCIe rate (a copy into pageable memory is
// staged by the driver in small synchronous pieces), then one memcpy hands them to the caller's buffer.
static thread_local void*  g_stage = nullptr;
static thread_local size_t g_stage_cap = 0;
static int stage_ensure(size_t bytes)
{
    if (bytes <= g_stage_cap) return MPL_OK;
    if (g_stage) { cudaFreeHost(g_stage); g_stage = nullptr; g_stage_cap = 0; }
    const size_t want = bytes + bytes / 4 + 4096;
    MPL_CUDA_CHECK(cudaHostAlloc(&g_stage, want, cudaHostAllocDefault));
    g_stage_cap = want;
    return MPL_OK;
}
// second, independent area (the pair pipeline's own results, alive across mpl_ransac_run_locked's use of the first)
static thread_local void*  g_stage2 = nullptr;
static thread_local size_t g_stage2_cap = 0;
int mpl_stage2_ensure(size_t bytes, void** ptr)
{
    if (bytes > g_stage2_cap) {
        if (g_stage2) { cudaFreeHost(g_stage2); g_stage2 = nullptr; g_stage2_cap = 0; }
        const size_t want = bytes + bytes / 4 + 4096;
        MPL_CUDA_CHECK(cudaHostAlloc(&g_stage2, want, cudaHostAllocDefault));
        g_stage2_cap = want;
    }
    *ptr = g_stage2;
    return MPL_OK;
}
int mpl_stage_ensure(size_t bytes, void** ptr) { int rc = stage_ensure(bytes); if (rc) return rc; *ptr = g_stage; return MPL_OK; }
static void stage_release()
{
    if (g_stage) cudaFreeHost(g_stage);
    g_stage = nullptr; g_stage_cap = 0;
    if (g_stage2) cudaFreeHost(g_stage2);
    g_stage2 = nullptr; g_stage2_cap = 0;
}

// frees the calling thread's streams, scratch and staging (caller has made the context's device current)
static void ctx_release()
{
    if (!g_ctx.ready) return;
    cudaStreamSynchronize(g_ctx.stream);
    for (auto& b : g_ctx.bufs) b.release();
    for (auto& l : g_ctx.lanes) l.release();
    stage_release();
    cudaStreamDestroy(g_ctx.stream);
    g_ctx.ready = false;
    (void)cudaGetLastError();
}
// ... when the thread ends
struct CtxGuard {
    ~CtxGuard() {
        if (!g_ctx.ready) return;
        if (cudaSetDevice(g_ctx.device) != cudaSuccess) return;
        ctx_release();
    }
};

static int init_locked(int device)
{
    static thread_local CtxGuard guard;
    (void)guard;
    if (g_ctx.ready && g_ctx.device == device) return MPL_OK;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        mpl_set_error("no CUDA device available (%s); libmplb200 has no CPU fallback",
                      e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
        return MPL_ERR_NO_DEVICE;
    }
    if (device < 0 || device >= n) { mpl_set_error("device %d out of range (0..%d)", device, n - 1); return MPL_ERR_BAD_ARG; }
    if (g_ctx.ready) { cudaSetDevice(g_ctx.device); ctx_release(); }
    MPL_CUDA_CHECK(cudaSetDevice(device));
    cudaDeviceProp p;
    MPL_CUDA_CHECK(cudaGetDeviceProperties(&p, device));
    if (p.major != 10) {
        mpl_set_error("device %d is sm_%d%d; libmplb200 is built for sm_100a (B200) only", device, p.major, p.minor);
        return MPL_ERR_NO_DEVICE;
    }
    g_ctx.device = device; g_ctx.sm_count = p.multiProcessorCount; g_ctx.cc_major = p.major; g_ctx.cc_minor = p.minor;
    g_ctx.total_mem = p.totalGlobalMem;
    MPL_CUDA_CHECK(cudaStreamCreateWithFlags(&g_ctx.stream, cudaStreamNonBlocking));
    g_ctx.ready = true;
    return MPL_OK;
}

int mpl_ensure_ctx_locked()
{
    const int device = g_device.load(std::memory_order_relaxed);
    if (g_ctx.ready && g_ctx.device == device) { MPL_CUDA_CHECK(cudaSetDevice(device)); return MPL_OK; }
    return init_locked(device);
}

extern "C" int mpl_init(int device)
{
    std::lock_guard<std::mutex> g(g_mu);
    int rc = init_locked(device);
    if (rc == MPL_OK) g_device.store(device, std::memory_order_relaxed);
    return rc;
}

extern "C" int mpl_shutdown(void)
{
    std::lock_guard<std::mutex> g(g_mu);
    if (!g_ctx.ready) return MPL_OK;
    cudaSetDevice(g_ctx.device);
    ctx_release();
    return MPL_OK;
}

// ---- huge-page pinned host memory -----------------------------------------------------------------
static std::mutex g_host_mu;
static std::map<void*, std::pair<void*, size_t>> g_host_allocs;      // user pointer -> (mmap base, mmap size)

extern "C" void* mpl_host_alloc(size_t bytes)
{
    if (bytes == 0) { mpl_set_error("mpl_host_alloc: zero bytes"); return nullptr; }
    {
        std::lock_guard<std::mutex> g(g_mu);
        if (mpl_ensure_ctx_locked()) return nullptr;
    }
    const size_t H = (size_t)2 << 20;
    const size_t size = ((bytes + H - 1) / H) * H + H;
    void* base = mmap(nullptr, size, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (base == MAP_FAILED) { mpl_set_error("mpl_host_alloc: mmap of %zu bytes failed", size); return nullptr; }
    char* al = (char*)((((uintptr_t)base + H - 1) / H) * H);
    const size_t usable = ((bytes + H - 1) / H) * H;
    (void)madvise(al, usable, MADV_HUGEPAGE);                      // advisory: THP off just means 4 KiB pages
    for (size_t o = 0; o < usable; o += 4096) al[o] = 0;           // first touch after the advice
    cudaError_t e = cudaHostRegister(al, usable, cudaHostRegisterDefault);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        munmap(base, size);
        mpl_set_error("mpl_host_alloc: cudaHostRegister failed (%s)", cudaGetErrorString(e));
        return nullptr;
    }
    std::lock_guard<std::mutex> g(g_host_mu);
    g_host_allocs[al] = std::make_pair(base, size);
    return al;
}

// True when [p, p + bytes) lies inside one mpl_host_alloc region: pinned, registered with cudaHostRegisterDefault (mapped and portable under
// unified addressing), i.e. kernels may read it through the same pointer.  The registration covers whole 2 MiB pages, so a 16-byte read that
// starts inside the range never leaves mapped memory.
bool mpl_host_is_mapped(const void* p, size_t bytes)
{
    std::lock_guard<std::mutex> g(g_host_mu);
    auto it = g_host_allocs.upper_bound(const_cast<void*>(p));
    if (it == g_host_allocs.begin()) return false;
    --it;
    const char* al = (const char*)it->first;
    const size_t H = (size_t)2 << 20;
    const char* reg_end = al + (it->second.second - H);                         // mpl_host_alloc registers (mmap size - 2 MiB) bytes from the aligned start
    return (const char*)p >= al && (const char*)p + bytes <= reg_end;
}

extern "C" int mpl_host_free(void* p)
{
    if (!p) return MPL_OK;
    std::pair<void*, size_t> rec;
    {
        std::lock_guard<std::mutex> g(g_host_mu);
        auto it = g_host_allocs.find(p);
        if (it == g_host_allocs.end()) { mpl_set_error("mpl_host_free: unknown pointer"); return MPL_ERR_BAD_ARG; }
        rec = it->second;
        g_host_allocs.erase(it);
    }
    (void)cudaHostUnregister(p);
    (void)cudaGetLastError();
    munmap(rec.first, rec.second);
    return MPL_OK;
}

extern "C" const char* mpl_last_error(void) { return g_err; }
extern "C" long long mpl_kernel_launch_count(void) { return g_launches.load(); }

extern "C" int mpl_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem)
{
    std::lock_guard<std::mutex> g(g_mu);
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    if (sm_count) *sm_count = g_ctx.sm_count;
    if (cc_major) *cc_major = g_ctx.cc_major;
    if (cc_minor) *cc_minor = g_ctx.cc_minor;
    if (total_mem) *total_mem = g_ctx.total_mem;
    return MPL_OK;
}

// ------------------------------------------------------------------------------------------------
// Hamming
// ------------------------------------------------------------------------------------------------
static int round_w16(int nbytes_used)
{
    int w = (nbytes_used + 15) / 16;
    int r = 1;
    while (r < w) r *= 2;
    return r;     // 1,2,4,8,16
}

// Hamming k-NN path: 0 = auto, 1 = popc scan (hamming_top2_kernel), 2 = tensor cores (bits as e4m3 bytes, knn_l2_tc.cu), 3 = tensor cores
// with bf16 operands (the first form; kept for comparison).
// auto rule of the Hamming path, from measurements (tools/hamming_tc_time.py): the tensor-core path is 1.4x
// (5000 x 5000 x 64 B) to 3.4x (10 000 x 12.5 M x 32 B) faster once the scan hides its four launches and fills the 256-row query
// blocks; few queries (a tile computes 256 rows whatever nq is) and small scans stay on the popc kernel
#ifndef MPL_HAMMING_TC_AUTO
#define MPL_HAMMING_TC_AUTO(nq, nt, w16) ((nq) >= 128 && (long long)(nq) * (nt) >= (1ll << 24))
#endif
static std::atomic<int> g_hamming_path{0};
extern "C" int mpl_set_hamming_path(int path) { if (path < 0 || path > 3) return MPL_ERR_BAD_ARG; g_hamming_path.store(path, std::memory_order_relaxed); return MPL_OK; }
int mpl_hamming_path() { return g_hamming_path.load(std::memory_order_relaxed); }
static bool hamming_wants_tc(int nq, int nt, int w16)
{
    const int path = g_hamming_path.load(std::memory_order_relaxed);
    if (path == 1 || !mpl_hamming_tc_supported(nq, nt, w16)) return false;
    if (path >= 2) return true;
    return MPL_HAMMING_TC_AUTO(nq, nt, w16);
}

extern "C" size_t mpl_knn_hamming_scratch_bytes(int nq, int nt, int w16)
{
    if (nq <= 0 || nt <= 0) return 16;
    int sm = g_ctx.ready ? g_ctx.sm_count : 148;
    int splits = mpl_hamming_pick_splits(nq, nt, w16, sm);
    const size_t popc = (size_t)splits * nq * 2 * sizeof(mpl_key64);
    // sized for either path, so that a buffer obtained here serves whatever mpl_set_hamming_path selects later
    const size_t tc = mpl_hamming_tc_supported(nq, nt, w16) ? mpl_hamming_tc_scratch_bytes(nq, nt, w16, sm) : 0;
    return popc > tc ? popc : tc;
}

extern "C" int mpl_pack_rows_u8_dev(const void* d_src, int n, size_t src_step, int nbytes_used, void* d_dst, int w16, void* stream)
{
    return mpl_launch_pack_rows_u8((const uint8_t*)d_src, n, src_step, nbytes_used, (uint8_t*)d_dst, w16, (cudaStream_t)stream);
}

extern "C" int mpl_knn2_hamming_dev(const void* d_q, int nq, const void* d_t, int nt, int w16, long long train_offset,
                                    void* d_scratch, size_t scratch_bytes, void* d_keys_out, void* stream)
{
    if (nq <= 0) return MPL_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (nt <= 0) { MPL_CUDA_CHECK(cudaMemsetAsync(d_keys_out, 0xFF, (size_t)nq * 2 * sizeof(mpl_key64), st)); return MPL_OK; }
    int sm = g_ctx.ready ? g_ctx.sm_count : 148;
    if (hamming_wants_tc(nq, nt, w16) && d_scratch && ((uintptr_t)d_scratch & 255) == 0 &&
        scratch_bytes >= mpl_hamming_tc_scratch_bytes(nq, nt, w16, sm))
        return mpl_hamming_tc_top2((const uint8_t*)d_q, nq, (const uint8_t*)d_t, nt, w16, train_offset, d_scratch, scratch_bytes,
                                   (mpl_key64*)d_keys_out, sm, st, mpl_hamming_path() == 3 ? 1 : 2);
    if (g_hamming_path.load(std::memory_order_relaxed) >= 2) { mpl_set_error("hamming_dev: tensor-core path forced but shape / scratch do not allow it (w16=%d, scratch %zu bytes)", w16, scratch_bytes); return MPL_ERR_BAD_ARG; }
    const int splits = mpl_hamming_pick_splits(nq, nt, w16, sm);
    if (splits == 1)
        return mpl_launch_hamming_top2((const uint8_t*)d_q, nq, (const uint8_t*)d_t, nt, w16, 1, 1, train_offset,
                                       (mpl_key64*)d_keys_out, st);
    const size_t need = (size_t)splits * nq * 2 * sizeof(mpl_key64);
    if (scratch_bytes < need || !d_scratch) { mpl_set_error("hamming_dev: scratch %zu < %zu bytes", scratch_bytes, need); return MPL_ERR_BAD_ARG; }
    int rc = mpl_launch_hamming_top2((const uint8_t*)d_q, nq, (const uint8_t*)d_t, nt, w16, splits, 1, train_offset,
                                     (mpl_key64*)d_scratch, st);
    if (rc) return rc;
    return mpl_launch_merge_top2((const mpl_key64*)d_scratch, splits, nq, 2, nullptr, nullptr, -1, (mpl_key64*)d_keys_out, st);
}

extern "C" int mpl_knn2_hamming_batch_dev(const void* d_q, const void* d_t, int n_pairs, int nq, int nt, int w16,
                                          void* d_keys_out, void* stream)
{
    if (nq <= 0 || n_pairs <= 0) return MPL_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (nt <= 0) { MPL_CUDA_CHECK(cudaMemsetAsync(d_keys_out, 0xFF, (size_t)n_pairs * nq * 2 * sizeof(mpl_key64), st)); return MPL_OK; }
    if (nt > (1 << 20)) { mpl_set_error("hamming_batch: nt=%d exceeds 2^20 rows per pair", nt); return MPL_ERR_LIMIT; }
    return mpl_launch_hamming_top2((const uint8_t*)d_q, nq, (const uint8_t*)d_t, nt, w16, 1, n_pairs, 0,
                                   (mpl_key64*)d_keys_out, st);
}

extern "C" int mpl_merge_top2_dev(const void* d_keys_parts, int n_parts, int nq, int k, int is_float,
                                  void* d_idx, void* d_dist, void* d_keys_merged, void* stream)
{
    if (k != 1 && k != 2) { mpl_set_error("merge: k must be 1 or 2"); return MPL_ERR_BAD_ARG; }
    return mpl_launch_merge_top2((const mpl_key64*)d_keys_parts, n_parts, nq, k, (int32_t*)d_idx, d_dist,
                                 is_float ? 0 : -1, (mpl_key64*)d_keys_merged, (cudaStream_t)stream);
}

// ---- peer memory for the sharded database regime ---------------------------------------------------
extern "C" size_t mpl_peer_gather_bytes(int nq, int world) { return (size_t)2 * (world > 0 ? world : 1) * (nq > 0 ? nq : 0) * 2 * sizeof(mpl_key64) + 256; }

extern "C" int mpl_peer_alloc(size_t bytes, void** d_ptr, unsigned char* ipc_handle64)
{
    if (!d_ptr || !ipc_handle64 || bytes == 0) { mpl_set_error("mpl_peer_alloc: bad argument"); return MPL_ERR_BAD_ARG; }
    {
        std::lock_guard<std::mutex> g(g_mu);
        int rc = mpl_ensure_ctx_locked();
        if (rc) return rc;
    }
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    MPL_CUDA_CHECK(cudaMalloc(d_ptr, bytes));
    MPL_CUDA_CHECK(cudaMemset(*d_ptr, 0, bytes));
    cudaIpcMemHandle_t h;
    MPL_CUDA_CHECK(cudaIpcGetMemHandle(&h, *d_ptr));
    memcpy(ipc_handle64, &h, 64);
    return MPL_OK;
}
extern "C" int mpl_peer_open(const unsigned char* ipc_handle64, void** d_ptr)
{
    if (!d_ptr || !ipc_handle64) { mpl_set_error("mpl_peer_open: bad argument"); return MPL_ERR_BAD_ARG; }
    {
        std::lock_guard<std::mutex> g(g_mu);
        int rc = mpl_ensure_ctx_locked();
        if (rc) return rc;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle64, 64);
    MPL_CUDA_CHECK(cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess));
    return MPL_OK;
}
extern "C" int mpl_peer_close(void* d_ptr) { if (d_ptr) MPL_CUDA_CHECK(cudaIpcCloseMemHandle(d_ptr)); return MPL_OK; }
extern "C" int mpl_peer_free(void* d_ptr) { if (d_ptr) MPL_CUDA_CHECK(cudaFree(d_ptr)); return MPL_OK; }

extern "C" int mpl_peer_status(const void* d_sync_own, int* timed_out, void* stream)
{
    if (!d_sync_own || !timed_out) { mpl_set_error("mpl_peer_status: null argument"); return MPL_ERR_BAD_ARG; }
    unsigned w[4] = {0, 0, 0, 0};
    MPL_CUDA_CHECK(cudaStreamSynchronize((cudaStream_t)stream));
    MPL_CUDA_CHECK(cudaMemcpy(w, d_sync_own, sizeof(w), cudaMemcpyDeviceToHost));
    *timed_out = w[3] ? 1 : 0;
    if (w[3]) mpl_set_error("allgather_merge: timed out waiting for the peer ranks (step %u)", w[0]);
    return MPL_OK;
}

extern "C" int mpl_allgather_merge_top2_dev(const void* d_keys_local, int nq, int rank, int world, void* const* peer_gather,
                                            void* const* peer_sync, int k, int is_float, void* d_idx, void* d_dist, void* d_keys_merged,
                                            void* stream)
{
    if (k != 1 && k != 2) { mpl_set_error("allgather_merge: k must be 1 or 2"); return MPL_ERR_BAD_ARG; }
    if (!d_keys_local || !peer_gather || !peer_sync) { mpl_set_error("allgather_merge: null argument"); return MPL_ERR_BAD_ARG; }
    return mpl_launch_allgather_merge((const mpl_key64*)d_keys_local, nq, rank, world, peer_gather, peer_sync, k, (int32_t*)d_idx, d_dist,
                                      is_float ? 0 : -1, (mpl_key64*)d_keys_merged, (cudaStream_t)stream);
}

// upload a strided host matrix (rows x row_bytes, stride step) into a dense device buffer
static int upload_rows(MplBuf& buf, const void* h, int rows, size_t row_bytes, size_t step, size_t dst_pitch, cudaStream_t st)
{
    int rc = buf.ensure((size_t)rows * dst_pitch + 16);
    if (rc) return rc;
    if (rows == 0) return MPL_OK;
    if (step == row_bytes && dst_pitch == row_bytes)
        MPL_CUDA_CHECK(cudaMemcpyAsync(buf.ptr, h, (size_t)rows * row_bytes, cudaMemcpyHostToDevice, st));
    else
        MPL_CUDA_CHECK(cudaMemcpy2DAsync(buf.ptr, dst_pitch, h, step, row_bytes, rows, cudaMemcpyHostToDevice, st));
    return MPL_OK;
}

// Upload both descriptor sets, scan, and leave the merged per-query top-2 keys on the device (context buffer KEYS).
// Caller holds the context mutex; nq, nt > 0.
int mpl_hamming_keys_locked(const uint8_t* q, int nq, size_t q_step, const uint8_t* t, int nt, size_t t_step, int nbytes,
                            int ignore_tail_bytes, mpl_key64** d_keys)
{
    int rc;
    cudaStream_t st = g_ctx.stream;
    const int used = nbytes - ignore_tail_bytes;
    const int w16 = round_w16(used);
    const size_t rowb = (size_t)w16 * 16;
    MplBuf &bq = g_ctx.bufs[MPL_BUF_Q], &bt = g_ctx.bufs[MPL_BUF_T], &braw = g_ctx.bufs[MPL_BUF_RAW];
    // queries and train rows -> packed device layout
    const uint8_t* src[2] = {q, t}; const int n[2] = {nq, nt}; const size_t stp[2] = {q_step, t_step};
    MplBuf* dst[2] = {&bq, &bt};
    for (int s = 0; s < 2; ++s) {
        if ((size_t)used == rowb) {                       // already in device layout: strided copy only
            rc = upload_rows(*dst[s], src[s], n[s], rowb, stp[s], rowb, st);
            if (rc) return rc;
        } else {
            rc = upload_rows(braw, src[s], n[s], (size_t)nbytes, stp[s], (size_t)nbytes, st);
            if (rc) return rc;
            rc = dst[s]->ensure((size_t)n[s] * rowb);
            if (rc) return rc;
            rc = mpl_launch_pack_rows_u8((const uint8_t*)braw.ptr, n[s], (size_t)nbytes, used, (uint8_t*)dst[s]->ptr, w16, st);
            if (rc) return rc;
        }
    }
    const int splits = mpl_hamming_pick_splits(nq, nt, w16, g_ctx.sm_count);
    MplBuf &bpart = g_ctx.bufs[MPL_BUF_PART], &bkeys = g_ctx.bufs[MPL_BUF_KEYS];
    if ((rc = bkeys.ensure((size_t)nq * 2 * sizeof(mpl_key64)))) return rc;
    if (hamming_wants_tc(nq, nt, w16)) {
        const size_t need = mpl_hamming_tc_scratch_bytes(nq, nt, w16, g_ctx.sm_count);
        if ((rc = bpart.ensure(need))) return rc;
        rc = mpl_hamming_tc_top2((const uint8_t*)bq.ptr, nq, (const uint8_t*)bt.ptr, nt, w16, 0, bpart.ptr, need, (mpl_key64*)bkeys.ptr, g_ctx.sm_count, st, mpl_hamming_path() == 3 ? 1 : 2);
        if (rc) return rc;
        *d_keys = (mpl_key64*)bkeys.ptr;
        return MPL_OK;
    }
    if ((rc = bpart.ensure((size_t)splits * nq * 2 * sizeof(mpl_key64)))) return rc;
    rc = mpl_launch_hamming_top2((const uint8_t*)bq.ptr, nq, (const uint8_t*)bt.ptr, nt, w16, splits, 1, 0, (mpl_key64*)bpart.ptr, st);
    if (rc) return rc;
    rc = mpl_launch_merge_top2((const mpl_key64*)bpart.ptr, splits, nq, 2, nullptr, nullptr, -1, (mpl_key64*)bkeys.ptr, st);
    if (rc) return rc;
    *d_keys = (mpl_key64*)bkeys.ptr;
    return MPL_OK;
}

// idx/dist columns of merged device keys -> caller's host arrays through the pinned staging buffer
static int keys_to_host_locked(const mpl_key64* d_keys, int nq, int k, int float_mode, int32_t* idx, void* dist)
{
    int rc;
    cudaStream_t st = g_ctx.stream;
    MplBuf &bidx = g_ctx.bufs[MPL_BUF_IDX], &bdist = g_ctx.bufs[MPL_BUF_DIST];
    const size_t bytes = (size_t)nq * k * 4;
    if ((rc = bidx.ensure(bytes))) return rc;
    if ((rc = bdist.ensure(bytes))) return rc;
    if ((rc = stage_ensure(2 * bytes))) return rc;
    rc = mpl_launch_merge_top2(d_keys, 1, nq, k, (int32_t*)bidx.ptr, bdist.ptr, float_mode, nullptr, st);
    if (rc) return rc;
    MPL_CUDA_CHECK(cudaMemcpyAsync(g_stage, bidx.ptr, bytes, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaMemcpyAsync((char*)g_stage + bytes, bdist.ptr, bytes, cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    memcpy(idx, g_stage, bytes);
    memcpy(dist, (char*)g_stage + bytes, bytes);
    return MPL_OK;
}

extern "C" int mpl_knn2_hamming(const uint8_t* q, int nq, size_t q_step, const uint8_t* t, int nt, size_t t_step,
                                int nbytes, int ignore_tail_bytes, int k, int32_t* idx, int32_t* dist)
{
    if (nq < 0 || nt < 0 || nbytes <= 0 || (k != 1 && k != 2) || ignore_tail_bytes < 0 || ignore_tail_bytes >= nbytes ||
        (nq > 0 && (!q || !idx || !dist)) || (nt > 0 && !t) || q_step < (size_t)nbytes || t_step < (size_t)nbytes) {
        mpl_set_error("mpl_knn2_hamming: bad argument"); return MPL_ERR_BAD_ARG;
    }
    if (nbytes - ignore_tail_bytes > 256) { mpl_set_error("mpl_knn2_hamming: descriptors longer than 256 bytes are not supported"); return MPL_ERR_LIMIT; }
    std::lock_guard<std::mutex> g(g_mu);
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    if (nq == 0) return MPL_OK;
    if (nt == 0) { for (int i = 0; i < nq * k; ++i) { idx[i] = -1; dist[i] = INT_MAX; } return MPL_OK; }
    mpl_key64* d_keys = nullptr;
    if ((rc = mpl_hamming_keys_locked(q, nq, q_step, t, nt, t_step, nbytes, ignore_tail_bytes, &d_keys))) return rc;
    return keys_to_host_locked(d_keys, nq, k, -1, idx, dist);
}

// ------------------------------------------------------------------------------------------------
// float L2
// ------------------------------------------------------------------------------------------------
static std::atomic<int> g_l2_path{0};
extern "C" int mpl_set_l2_path(int path) { if (path < 0 || path > 2) return MPL_ERR_BAD_ARG; g_l2_path.store(path, std::memory_order_relaxed); return MPL_OK; }

extern "C" size_t mpl_knn_l2_scratch_bytes(int nq, int nt, int dim)
{
    (void)dim;
    if (nq <= 0 || nt <= 0) return 16;
    int sm = g_ctx.ready ? g_ctx.sm_count : 148;
    size_t exact = (size_t)mpl_l2_pick_splits(nq, nt, sm) * nq * 2 * sizeof(mpl_key64);
    size_t tc = mpl_l2tc_scratch_bytes(nq, nt, dim, sm);
    return exact > tc ? exact : tc;
}

extern "C" int mpl_knn2_l2f32_dev(const void* d_q, int nq, size_t q_ld, const void* d_t, int nt, size_t t_ld, int dim,
                                  int sqrt_out, long long train_offset, void* d_scratch, size_t scratch_bytes,
                                  void* d_keys_out, void* stream)
{
    if (nq <= 0) return MPL_OK;
    cudaStream_t st = (cudaStream_t)stream;
    if (nt <= 0) { MPL_CUDA_CHECK(cudaMemsetAsync(d_keys_out, 0xFF, (size_t)nq * 2 * sizeof(mpl_key64), st)); return MPL_OK; }
    int sm = g_ctx.ready ? g_ctx.sm_count : 148;
    // auto: the tensor-core path pays off once the scan is big enough to hide its 5 launches (prep, pack, rank, certify, list)
    const int l2_path = g_l2_path.load(std::memory_order_relaxed);
    const bool want_tc = (l2_path == 2) || (l2_path == 0 && mpl_l2tc_supported(nq, nt, dim) && (long long)nq * nt >= (1ll << 22));
    if (want_tc) {
        if (!mpl_l2tc_supported(nq, nt, dim)) { mpl_set_error("l2: tensor-core path forced but shape unsupported (dim=%d)", dim); return MPL_ERR_BAD_ARG; }
        return mpl_l2tc_top2((const float*)d_q, nq, q_ld, (const float*)d_t, nt, t_ld, dim, sqrt_out, train_offset,
                             d_scratch, scratch_bytes, (mpl_key64*)d_keys_out, sm, st);
    }
    const int splits = mpl_l2_pick_splits(nq, nt, sm);
    const size_t need = (size_t)splits * nq * 2 * sizeof(mpl_key64);
    if (splits == 1)
        return mpl_launch_l2_exact_top2((const float*)d_q, nq, q_ld, (const float*)d_t, nt, t_ld, dim, sqrt_out ? 1 : 0, 1,
                                        train_offset, (mpl_key64*)d_keys_out, st);
    if (scratch_bytes < need || !d_scratch) { mpl_set_error("l2_dev: scratch %zu < %zu bytes", scratch_bytes, need); return MPL_ERR_BAD_ARG; }
    int rc = mpl_launch_l2_exact_top2((const float*)d_q, nq, q_ld, (const float*)d_t, nt, t_ld, dim, sqrt_out ? 1 : 0, splits,
                                      train_offset, (mpl_key64*)d_scratch, st);
    if (rc) return rc;
    return mpl_launch_merge_top2((const mpl_key64*)d_scratch, splits, nq, 2, nullptr, nullptr, 0, (mpl_key64*)d_keys_out, st);
}

extern "C" int mpl_knn_l2_last_stats(const void* d_scratch, unsigned* fallbacks, unsigned* tf32_mode, void* stream)
{
    if (!d_scratch) { mpl_set_error("knn_l2_last_stats: null scratch"); return MPL_ERR_BAD_ARG; }
    return mpl_l2tc_last_fallbacks(d_scratch, fallbacks, tf32_mode, (cudaStream_t)stream);
}
// same for the host-pointer entry point (library-owned scratch)
extern "C" int mpl_knn_l2_last_stats_host(unsigned* fallbacks, unsigned* tf32_mode)
{
    std::lock_guard<std::mutex> g(g_mu);
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    if (!g_ctx.bufs[MPL_BUF_PART].ptr) { mpl_set_error("knn_l2_last_stats_host: no float k-NN call yet"); return MPL_ERR_BAD_ARG; }
    return mpl_l2tc_last_fallbacks(g_ctx.bufs[MPL_BUF_PART].ptr, fallbacks, tf32_mode, g_ctx.stream);
}

// Upload both float descriptor sets, scan, and leave the per-query top-2 keys on the device (context buffer KEYS).
// Caller holds the context mutex; nq, nt > 0.
static int l2_keys_locked(const float* q, int nq, size_t q_step, const float* t, int nt, size_t t_step, int dim, int sqrt_out,
                          mpl_key64** d_keys)
{
    int rc;
    cudaStream_t st = g_ctx.stream;
    const size_t ld = (size_t)((dim + 3) / 4) * 4;           // device row pitch in floats (16-byte aligned rows)
    MplBuf &bq = g_ctx.bufs[MPL_BUF_Q], &bt = g_ctx.bufs[MPL_BUF_T];
    if (ld != (size_t)dim) {                                  // pad columns are never read as payload, but keep them clean
        if ((rc = bq.ensure((size_t)nq * ld * 4 + 16))) return rc;
        if ((rc = bt.ensure((size_t)nt * ld * 4 + 16))) return rc;
        MPL_CUDA_CHECK(cudaMemsetAsync(bq.ptr, 0, (size_t)nq * ld * 4, st));
        MPL_CUDA_CHECK(cudaMemsetAsync(bt.ptr, 0, (size_t)nt * ld * 4, st));
    }
    MplBuf &bscr = g_ctx.bufs[MPL_BUF_PART], &bkeys = g_ctx.bufs[MPL_BUF_KEYS];
    const size_t scr = mpl_knn_l2_scratch_bytes(nq, nt, dim);
    if ((rc = bscr.ensure(scr))) return rc;
    if ((rc = bkeys.ensure((size_t)nq * 2 * sizeof(mpl_key64)))) return rc;
    if ((rc = upload_rows(bq, q, nq, (size_t)dim * 4, q_step, ld * 4, st))) return rc;
    if ((rc = upload_rows(bt, t, nt, (size_t)dim * 4, t_step, ld * 4, st))) return rc;
    rc = mpl_knn2_l2f32_dev(bq.ptr, nq, ld, bt.ptr, nt, ld, dim, sqrt_out, 0, bscr.ptr, scr, bkeys.ptr, (void*)st);
    if (rc) return rc;
    *d_keys = (mpl_key64*)bkeys.ptr;
    return MPL_OK;
}

extern "C" int mpl_knn2_l2f32(const float* q, int nq, size_t q_step, const float* t, int nt, size_t t_step,
                              int dim, int sqrt_out, int k, int32_t* idx, float* dist)
{
    if (nq < 0 || nt < 0 || dim <= 0 || (k != 1 && k != 2) || (nq > 0 && (!q || !idx || !dist)) || (nt > 0 && !t) ||
        q_step < (size_t)dim * 4 || t_step < (size_t)dim * 4) {
        mpl_set_error("mpl_knn2_l2f32: bad argument"); return MPL_ERR_BAD_ARG;
    }
    std::lock_guard<std::mutex> g(g_mu);
    int rc = mpl_ensure_ctx_locked();
    if (rc) return rc;
    if (nq == 0) return MPL_OK;
    if (nt == 0) { for (int i = 0; i < nq * k; ++i) { idx[i] = -1; dist[i] = FLT_MAX; } return MPL_OK; }
    mpl_key64* d_keys = nullptr;
    if ((rc = l2_keys_locked(q, nq, q_step, t, nt, t_step, dim, sqrt_out, &d_keys))) return rc;
    return keys_to_host_locked(d_keys, nq, k, 0, idx, dist);
}

// ------------------------------------------------------------------------------------------------
// ratio test + getMatches mirror (host logic, O(nq))
// ------------------------------------------------------------------------------------------------
extern "C" int mpl_ratio_matches_i32(const int32_t* idx, const int32_t* dist, int nq, int k, int ratio_test,
                                     mpl_dmatch* out, int* n_out)
{
    if (nq < 0 || (k != 1 && k != 2) || (ratio_test && k != 2) || !n_out) { mpl_set_error("ratio_matches: bad argument"); return MPL_ERR_BAD_ARG; }
    int n = 0;
    for (int q = 0; q < nq; ++q) {
        // matchers.cpp:603 "dists[q][0] < (0.75f * dists[q][1])": int -> float promotion on both sides
        if (ratio_test && !((float)dist[(size_t)q * k] < 0.75f * (float)dist[(size_t)q * k + 1])) continue;
        out[n].queryIdx = q; out[n].trainIdx = idx[(size_t)q * k]; out[n].imgIdx = -1;   // cv::DMatch() default imgIdx
        out[n].distance = (float)dist[(size_t)q * k];
        ++n;
    }
    *n_out = n;
    return MPL_OK;
}

extern "C" int mpl_ratio_matches_f32(const int32_t* idx, const float* dist, int nq, int k, int ratio_test,
                                     mpl_dmatch* out, int* n_out)
{
    if (nq < 0 || (k != 1 && k != 2) || (ratio_test && k != 2) || !n_out) { mpl_set_error("ratio_matches: bad argument"); return MPL_ERR_BAD_ARG; }
    int n = 0;
    for (int q = 0; q < nq; ++q) {
        if (ratio_test && !(dist[(size_t)q * k] < 0.75f * dist[(size_t)q * k + 1])) continue;
        out[n].queryIdx = q; out[n].trainIdx = idx[(size_t)q * k]; out[n].imgIdx = -1;
        out[n].distance = dist[(size_t)q * k];
        ++n;
    }
    *n_out = n;
    return MPL_OK;
}

extern "C" int mpl_get_matches_api(int n_keypoints1, int n_keypoints2,
                                   const void* desc1, int rows1, int cols1, size_t step1, int cv_type1,
                                   const void* desc2, int rows2, int cols2, size_t step2, int cv_type2,
                                   const char* matcher_name, int ratio_test, int nr_threads,
                                   mpl_dmatch* out, int* n_out)
{
    (void)nr_threads;   // matchers.cpp:121 NMSLIB worker threads: meaningless on the device path
    const int CV_8U_ = 0, CV_32F_ = 5;
    if (n_out) *n_out = 0;
    if (!matcher_name || !n_out) { mpl_set_error("getMatches: null argument"); return -1; }
    // matchers.cpp:119 CV_Assert(descriptors1.type() == descriptors2.type()) -- the C++ shim throws
    if (cv_type1 != cv_type2) { mpl_set_error("CV_Assert(descriptors1.type() == descriptors2.type())"); return MPL_ERR_ASSERT; }
    if (n_keypoints1 < 15 || n_keypoints2 < 15) { mpl_set_error("Too less keypoits!"); return -4; }             // :123-127
    if (n_keypoints1 != rows1 || n_keypoints2 != rows2) {                                                        // :129-133
        mpl_set_error("Number of descriptors must be equal to the number of keypoints!"); return -1;
    }
    const std::string m(matcher_name);
    const bool linear = (m == "LINEAR"), nms = (m == "BRUTEFORCENMS");
    if (!linear && !nms) { mpl_set_error("Matcher %s is not supported.", matcher_name); return -2; }            // :715-719 (no CPU fallback)
    if (cv_type1 != CV_8U_ && cv_type1 != CV_32F_) {
        // LINEAR :540-547 "Format of descriptors not supported!"; BRUTEFORCENMS CV_64F (:502-513) is out of scope
        mpl_set_error("Format of descriptors not supported!"); return -1;
    }
    if (cols1 != cols2) { mpl_set_error("descriptor lengths differ (%d vs %d)", cols1, cols2); return -1; }
    // :529-538 nn = ratioTest ? 2 : 1 (LINEAR); nmslib_matchers.h:180 K = 2: the scan always keeps two, the emission below
    // applies the ratio test (or not) on the device and the compacted cv::DMatch list comes back in one copy.
    const int ignore = (cv_type1 == CV_8U_ && nms) ? ((cols1 & 1) ? 1 : 2) : 0;   // last packed 16-bit word never compared
    if (cv_type1 == CV_8U_ && cols1 - ignore <= 0) {        // nothing to compare: every distance 0, neighbours 0 and 1
        int n = 0;
        for (int qi = 0; qi < rows1; ++qi) {
            if (ratio_test) continue;                       // 0 < 0.75 * 0 never holds
            out[n].queryIdx = qi; out[n].trainIdx = 0; out[n].imgIdx = -1; out[n].distance = 0.f; ++n;
        }
        *n_out = n;
    } else {
        if (cv_type1 == CV_8U_ && cols1 - ignore > 256) { mpl_set_error("descriptors longer than 256 bytes are not supported"); return -3; }
        if (!desc1 || !desc2 || !out || step1 < (size_t)cols1 * (cv_type1 == CV_8U_ ? 1 : 4) || step2 < (size_t)cols2 * (cv_type1 == CV_8U_ ? 1 : 4)) {
            mpl_set_error("getMatches: bad descriptor matrix"); return -3;
        }
        std::lock_guard<std::mutex> g(g_mu);
        if (mpl_ensure_ctx_locked()) return -3;
        cudaStream_t st = g_ctx.stream;
        mpl_key64* d_keys = nullptr;
        int rc;
        if (cv_type1 == CV_8U_) rc = mpl_hamming_keys_locked((const uint8_t*)desc1, rows1, step1, (const uint8_t*)desc2, rows2, step2, cols1, ignore, &d_keys);
        else                    rc = l2_keys_locked((const float*)desc1, rows1, step1, (const float*)desc2, rows2, step2, cols1, nms ? 1 : 0, &d_keys);
        if (rc) return -3;
        MplBuf& bm = g_ctx.bufs[MPL_BUF_IDX];
        const size_t bytes = 16 + (size_t)rows1 * sizeof(mpl_dmatch);
        if (bm.ensure(bytes) || stage_ensure(bytes)) return -3;
        if (mpl_launch_ratio_dmatch(d_keys, rows1, ratio_test, cv_type1 == CV_32F_, bm.ptr, st)) return -3;
        if (cudaMemcpyAsync(g_stage, bm.ptr, bytes, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
            mpl_set_error("getMatches: device copy failed (%s)", cudaGetErrorString(cudaGetLastError())); return -3;
        }
        const int n = *(const int*)g_stage;
        memcpy(out, (const char*)g_stage + 16, (size_t)n * sizeof(mpl_dmatch));
        *n_out = n;
    }
    if (linear && *n_out < 2) { mpl_set_error("Too less remaining matches using the LINEAR matcher."); return -3; }  // :709-713
    return 0;
}
