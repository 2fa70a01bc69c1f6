// knn_l2_tc.cu -- float-descriptor brute-force k-NN (k = 2) on the 5th-generation tensor cores, sm_100a.
//
// Replaces the hot loop of the reference's LINEAR matcher for CV_32F descriptors
//   cvflann::Index<L2<float>>(LinearIndexParams).knnSearch      (matchers.cpp:634-664; OpenCV 4.2.0 flann/dist.h)
// and of BRUTEFORCENMS (NMSLIB seq_search + l2, nmslib_matchers.h:159-424, distcomp_lp.cc:399-466).
//
// The reference evaluates d(q,t) = sum (q_i - t_i)^2 pair by pair in fp32.  Here the Q x T^T contraction runs
// as a dense GEMM on tcgen05 and only RANKS: score(q,t) = |t|^2 - 2 q.t (the |q|^2 term is constant per
// query).  Every distance that leaves the library is then recomputed in exact fp32, in the reference's own
// accumulation order, for a short certified candidate list:
//
//   1. l2_prep_scan_kernel   |x|^2 per row, max |t|^2, and whether every value is exactly representable in
//                            bf16 (true for OpenCV SIFT: integers 0..255).  No host round trip: the flag
//                            lives in device memory and the later kernels branch on it.
//   2. l2_prep_pack_kernel   operand copies: bf16 rows (exact data, kind::f16, full tensor rate, products and
//                            fp32 accumulation exact for integer data) or fp32 rows read as tf32 (kind::tf32).
//   3. l2_tc_rank_kernel     persistent, warp specialised: warp 0 = TMA producer (128B-swizzled K-major slabs,
//                            cp.async.bulk.tensor), warp 1 = tcgen05.mma issuer (one elected thread),
//                            warp 2 = TMEM allocator, warps 4-11 = epilogue.  CTA tile = 256 queries (two
//                            UMMA M=128 blocks, resident in smem) x 128 train rows per step; accumulators
//                            double-buffered in TMEM (2 x 2 x 128 columns = all 512).  Epilogue thread =
//                            one query row: tcgen05.ld 32 columns at a time, score = fma(-2, acc, |t|^2),
//                            register-resident sorted top-4 per thread (strict '<' keeps the lower train
//                            index on ties).  Work = (query block, train tile) units split evenly over the
//                            CTAs (grid = SM count), so a query block may be covered by several CTAs; each
//                            writes its own candidate list ("slot").
//   4. l2_certify_kernel     one warp per query: exact fp32 distances (cvflann / NMSLIB order) of the
//                            candidates that can still be among the two nearest, top-2 by (dist, trainIdx),
//                            and the certificate  d2 < tau + |q|^2 - eps  (tau = smallest "4th best" score
//                            over the slots: every train row that is NOT a candidate scores >= tau; eps bounds
//                            |score + |q|^2 - d|).  Certified queries are final and bit-exact.
//   5. the rare uncertified query (near-duplicate clusters) goes through the exact fp32 scan
//      (knn_l2_exact.cu, list-driven persistent kernel) -- still on the device, no host decision.
//
// Roofline: tensor pipe, 2*D flop per pair.
#include "mpl_common.cuh"
#include "mpl_ctx.h"

#include <cuda.h>
#include <cuda_bf16.h>
#include <math.h>

namespace {

// ---- geometry -------------------------------------------------------------------------------------
constexpr int BM = 256;              // queries per CTA tile (2 x UMMA_M)
constexpr int UM = 128;              // UMMA M
constexpr int BN = 128;              // train rows per tile = UMMA N
constexpr int SLAB_BYTES = 128;      // K extent of one smem slab in bytes (= swizzle width)
constexpr int A_SLAB = BM * SLAB_BYTES;      // 32 KB
constexpr int B_SLAB = BN * SLAB_BYTES;      // 16 KB
constexpr int MAX_A_SLABS = 4;       // query block of BM rows: dim <= 128 floats (tf32) / 256 bf16
constexpr int MAX_DIM = 256;         // wider rows (128 < dim <= 256) run with query blocks of UM rows: 8 tf32 slabs of 16 KB, one
                                     // UMMA M block (issuer warp 3 and the m = 1 epilogue warps idle), same rings and lists
constexpr int AUG_BYTES = 32;        // one extra K step per row: |t|^2 split in three exact pieces (B), ones (A)
constexpr int A_AUG = BM * AUG_BYTES;                        //  8 KB
constexpr int STAGE_BYTES = B_SLAB + BN * AUG_BYTES;         // 20 KB: one B slab (+ the tile's aug slab beside slab 0)
constexpr int MAX_STAGES = 8;
constexpr int SMEM_DATA = 197 * 1024;                        // A slabs | A aug | B ring (as many stages as fit)
constexpr int NC = 4;                // candidates kept per list
constexpr int LPS = 2;               // lists per (CTA, query block) slot: one per half of the tile's columns
constexpr int EPI_WARPS = 16;        // 4 warpgroups: (UMMA M block m) x (column half); thread = one query row x 64 columns
constexpr int TC_THREADS = 128 + 32 * EPI_WARPS;             // warps 0..3 control, 4..19 epilogue
constexpr int PEND_CAP = 6;          // pending (score, train index) entries per epilogue thread, drained into the sorted list
constexpr int PEND_STRIDE = 32 * EPI_WARPS * 8;              // bytes between consecutive entries of one thread (lane-interleaved)
constexpr int SMEM_SHARE = LPS * BM * 8;                     // published list thresholds: (NC-th best score, query block)
constexpr int SMEM_TAIL = 256 + SMEM_SHARE + PEND_CAP * PEND_STRIDE;      // barriers + TMEM slot | thresholds | pending buffers
constexpr int SMEM_BYTES = 1024 + SMEM_DATA + SMEM_TAIL;
static_assert(SMEM_BYTES <= 227 * 1024, "dynamic shared memory limit of sm_100");
static_assert((SMEM_DATA - MAX_A_SLABS * A_SLAB - A_AUG) / STAGE_BYTES >= 3, "B ring needs at least 3 stages in tf32 mode");

struct TcHeader {                    // first 256 bytes of the scratch buffer (zeroed per call)
    unsigned not_bf16_exact;         // 0 -> bf16 mode, else tf32 mode
    unsigned max_tnorm_bits;         // max |t|^2 (float bits; non-negative floats order like uints)
    unsigned fallback_count;         // queries that failed certification
    unsigned not_integer;            // some value is not an integer: the exact-integer mode (top-2 lists, no certificate) is off
    unsigned scan_ticket;            // CTAs of the fallback scan that have finished (the last one merges)
    unsigned pad[59];
};

// ---- PTX wrappers -----------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
// one lane of a converged warp (keeps the surrounding code warp-uniform, so descriptors and addresses stay in
// uniform registers instead of going through per-lane waterfall loops)
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after()  { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(uint64_t* bar) {   // arrives on bar when every MMA issued so far has completed
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_mma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tc_mma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// e4m3 x e4m3 -> f32, K = 32 per instruction (32 bytes of a K-major row, like the other kinds)
__device__ __forceinline__ void tc_mma_f8(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// 32 lanes x 32 consecutive columns of 32-bit accumulators -> 32 registers per thread (thread = lane = row)
__device__ __forceinline__ void tc_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
}

// K-major operand slab, 128-byte swizzle: 8-row groups of 1024 bytes (SBO), descriptor version 1 (sm_100)
[[maybe_unused]] __device__ __forceinline__ uint64_t umma_desc(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// K-major 32-byte rows, 32-byte swizzle: 8-row groups of 256 bytes (the aug slab)
[[maybe_unused]] __device__ __forceinline__ uint64_t umma_desc_sw32(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)(256 >> 4) << 32) | (1ull << 46) | (6ull << 61);
}
// instruction descriptor: D = f32, A/B format (kind::f16: 1 = bf16; kind::tf32: 2; kind::f8f6f4: 0 = e4m3), both K-major, N = 128, M = 128
__host__ __device__ constexpr uint32_t umma_idesc(uint32_t fmt) {
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(UM >> 4) << 24);
}

// ---- 1 + 2. operand preparation -------------------------------------------------------------------------
// The "aug" K step folds |t|^2 into the contraction, so that the accumulator IS the score |t|^2 - 2 q.t: train
// aug row = |t|^2 split into three pieces that are each exactly representable in the operand type (3 x 8 bits
// bf16, 3 x 11 bits tf32 >= the 24-bit fp32 mantissa), query aug row = (1, 1, 1, 0, ...).  Train pad rows (up to a
// whole tile) get +inf, so they never enter a list.  Query rows are scaled by -2 (exact).
__device__ __forceinline__ void split3(float n, unsigned keep_mask, float (&o)[3])
{
    if (!(n < INFINITY)) { o[0] = INFINITY; o[1] = 0.f; o[2] = 0.f; return; }
    o[0] = __uint_as_float(__float_as_uint(n) & keep_mask);
    const float r1 = n - o[0];                                   // exact
    o[1] = __uint_as_float(__float_as_uint(r1) & keep_mask);
    o[2] = r1 - o[1];                                            // exact, and fits the remaining bits
}

__device__ __forceinline__ void write_aug(uint4* dst, bool is_t, float norm, bool bf16_mode)
{
    float o[3];
    if (is_t) split3(norm, bf16_mode ? 0xFFFF0000u : 0xFFFFE000u, o);
    else { o[0] = 1.f; o[1] = 1.f; o[2] = 1.f; }
    uint4 w0 = make_uint4(0, 0, 0, 0);
    if (bf16_mode) {
        __nv_bfloat162 a = __floats2bfloat162_rn(o[0], o[1]), b2 = __floats2bfloat162_rn(o[2], 0.f);
        w0.x = *reinterpret_cast<uint32_t*>(&a); w0.y = *reinterpret_cast<uint32_t*>(&b2);
    } else {
        w0.x = __float_as_uint(o[0]); w0.y = __float_as_uint(o[1]); w0.z = __float_as_uint(o[2]);
    }
    dst[0] = w0; dst[1] = make_uint4(0, 0, 0, 0);
}

// One pass, 8 lanes per row (rows [0, nt_pad) train, then [nt_pad, nt_pad + nq) queries): |x|^2, max |t|^2, whether
// every value is exactly representable in bf16 (true for OpenCV SIFT: integers 0..255), and -- speculatively -- the
// bf16 operand rows (kp16 = dim rounded up to 64, zero padded) and bf16 aug rows.  No host round trip: the flag lives
// in device memory; if it ends up set, l2_prep_tf32_kernel overwrites the operands with fp32 rows read as tf32.
__global__ void __launch_bounds__(256)
l2_prep_bf16_kernel(const float* __restrict__ t, int nt, size_t t_ld, const float* __restrict__ q, int nq, size_t q_ld, int dim,
                    void* __restrict__ tp, void* __restrict__ qp, int kp16, float* __restrict__ tnorm, int nt_pad,
                    float* __restrict__ qnorm, uint4* __restrict__ taug, uint4* __restrict__ qaug, TcHeader* hdr,
                    uint32_t* __restrict__ gthr, int n_gthr)
{
    const int lane = threadIdx.x & 31;
    const int warps_per_grid = gridDim.x * (blockDim.x >> 5);
    const int total = nt_pad + nq;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_gthr; i += gridDim.x * blockDim.x) gthr[i] = 0xFFFFFFFFu;   // "nothing published"
    const bool vec_ok = (((uintptr_t)q | (uintptr_t)t) & 15) == 0 && ((q_ld | t_ld) & 3) == 0;
    unsigned inexact = 0;
    bool nonint = false;
    float tmax = 0.f;
    // 8 lanes per row, 4 rows per warp and iteration: 128-byte coalesced reads per row segment, a 3-step reduction, and
    // four group leaders instead of one lane doing the per-row bookkeeping
    const int sub = lane >> 3, sl = lane & 7;
    for (int base = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 4; base < total; base += warps_per_grid * 4) {
        const int row = base + sub;
        const bool live = row < total;
        const bool is_t = row < nt_pad;
        const int r = is_t ? row : row - nt_pad;
        const bool pad = live && is_t && r >= nt;                // train pad row of the last tile
        float s = 0.f;
        if (live && !pad) {
            const float* x = is_t ? t + (size_t)r * t_ld : q + (size_t)r * q_ld;
            const float sc = is_t ? 1.f : -2.f;
            char* dst = (char*)(is_t ? tp : qp) + (size_t)r * kp16 * 2;
            for (int i = 4 * sl; i < kp16; i += 32) {
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (i + 3 < dim && vec_ok) v = *reinterpret_cast<const float4*>(x + i);
                else {
                    if (i < dim) v.x = x[i];
                    if (i + 1 < dim) v.y = x[i + 1];
                    if (i + 2 < dim) v.z = x[i + 2];
                    if (i + 3 < dim) v.w = x[i + 3];
                }
                inexact |= (__float_as_uint(v.x) | __float_as_uint(v.y) | __float_as_uint(v.z) | __float_as_uint(v.w)) & 0xFFFFu;
                nonint |= (v.x != truncf(v.x)) | (v.y != truncf(v.y)) | (v.z != truncf(v.z)) | (v.w != truncf(v.w));   // NaN/inf: true
                s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
                __nv_bfloat162 lo = __floats2bfloat162_rn(sc * v.x, sc * v.y), hi = __floats2bfloat162_rn(sc * v.z, sc * v.w);
                uint2 o; o.x = *reinterpret_cast<uint32_t*>(&lo); o.y = *reinterpret_cast<uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(dst + 2 * i) = o;
            }
        }
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (live && sl == 0) {
            if (pad) { tnorm[r] = INFINITY; write_aug(taug + 2 * (size_t)r, true, INFINITY, true); }
            else {
                (is_t ? tnorm : qnorm)[r] = s;
                write_aug((is_t ? taug : qaug) + 2 * (size_t)r, is_t, s, true);
            }
        }
        if (live && is_t && !pad) tmax = fmaxf(tmax, s);
    }
    // one atomic per CTA (a same-address atomic per warp serialises in L2 and dominated this kernel)
    __shared__ unsigned s_inexact, s_tmax;
    if (threadIdx.x == 0) { s_inexact = 0; s_tmax = 0; }
    __syncthreads();
    inexact = (__any_sync(0xffffffffu, inexact != 0) ? 1u : 0u) | (__any_sync(0xffffffffu, nonint) ? 2u : 0u);
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) tmax = fmaxf(tmax, __shfl_xor_sync(0xffffffffu, tmax, o));
    if (lane == 0) {
        if (inexact) atomicOr(&s_inexact, inexact);
        if (tmax > 0.f) atomicMax(&s_tmax, __float_as_uint(tmax));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (s_inexact & 1u) atomicOr(&hdr->not_bf16_exact, 1u);
        if (s_inexact & 2u) atomicOr(&hdr->not_integer, 1u);
        if (s_tmax) atomicMax(&hdr->max_tnorm_bits, s_tmax);
    }
}

// Only when some value is not bf16-exact: fp32 operand rows (kp32 = dim rounded up to 32) and tf32 aug rows.
__global__ void __launch_bounds__(256)
l2_prep_tf32_kernel(const float* __restrict__ t, int nt, size_t t_ld, const float* __restrict__ q, int nq, size_t q_ld, int dim,
                    void* __restrict__ tp, void* __restrict__ qp, int kp32, const float* __restrict__ tnorm, int nt_pad,
                    uint4* __restrict__ taug, uint4* __restrict__ qaug, const TcHeader* __restrict__ hdr)
{
    if (hdr->not_bf16_exact == 0) return;
    const int groups = kp32 / 4;
    const long long total = (long long)(nt + nq) * groups;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (long long i = tid; i < total; i += stride) {
        const int row = (int)(i / groups), g = (int)(i - (long long)row * groups);
        const bool is_t = row < nt;
        const int r = is_t ? row : row - nt;
        const float* x = is_t ? t + (size_t)r * t_ld : q + (size_t)r * q_ld;
        const float sc = is_t ? 1.f : -2.f;
        float v[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = (4 * g + e < dim) ? sc * x[4 * g + e] : 0.f;
        char* dst = (char*)(is_t ? tp : qp);
        *reinterpret_cast<float4*>(dst + ((size_t)r * kp32 + 4 * g) * 4) = make_float4(v[0], v[1], v[2], v[3]);
    }
    for (long long i = tid; i < (long long)nt_pad + nq; i += stride) {
        const bool is_t = i < nt_pad;
        write_aug(is_t ? taug + 2 * i : qaug + 2 * (i - nt_pad), is_t, is_t ? tnorm[i] : 0.f, false);
    }
}

// Binary descriptors (Hamming distance as a contraction, section "Hamming on the tensor cores" below): rows of w16 x 16 bytes ->
// operand rows of 128 w16 elements, train bits as 0 / 1, query bits as 0 / -2, in e4m3 (one byte per bit, kind::f8f6f4) or bf16;
// |x|^2 = popcount.  8 lanes per row, lane sl takes bytes sl, sl + 8, ...: the 8 lanes of a row write 64 (e4m3) / 128 (bf16)
// contiguous bytes per step.  Element 8 b + i = bit i of byte b.
template <bool FP8>
__device__ __forceinline__ unsigned unpack_bits_row(const uint8_t* __restrict__ x, int row_bytes, int sl, bool is_t, bool real, char* __restrict__ dst)
{
    unsigned pc = 0;
    for (int b = sl; b < row_bytes; b += 8) {
        const unsigned v = real ? x[b] : 0u;                                     // padding rows: finite (zero) operands
        pc += __popc(v);
        if (FP8) {
            const uint32_t one = is_t ? 0x38u : 0xC0u;                           // e4m3 1.0 / -2.0
            uint2 o;
            o.x = (((v & 0xFu) * 0x00204081u) & 0x01010101u) * one;
            o.y = (((v >> 4) * 0x00204081u) & 0x01010101u) * one;
            reinterpret_cast<uint2*>(dst)[b] = o;
        } else {
            const uint32_t one = is_t ? 0x3F80u : 0xC000u;                       // bf16 1.0 / -2.0
            uint4 o;
            o.x = ((v & 1u) ? one : 0u) | ((v & 2u) ? one << 16 : 0u);
            o.y = ((v & 4u) ? one : 0u) | ((v & 8u) ? one << 16 : 0u);
            o.z = ((v & 16u) ? one : 0u) | ((v & 32u) ? one << 16 : 0u);
            o.w = ((v & 64u) ? one : 0u) | ((v & 128u) ? one << 16 : 0u);
            reinterpret_cast<uint4*>(dst)[b] = o;
        }
    }
    return pc;
}

// e4m3 code of an integer with at most four significant bits (<= 448)
__device__ __forceinline__ uint32_t e4m3_int(unsigned v)
{
    if (v == 0) return 0;
    const int e = 31 - __clz(v);
    const unsigned m = e >= 3 ? (v >> (e - 3)) & 7u : (v << (3 - e)) & 7u;
    return ((uint32_t)(e + 7) << 3) | m;
}

// aug row (32 bytes) of a bit-vector row with popcount pc (<= 512).  bf16: write_aug.  e4m3 has four significant bits, a maximum of
// 448 and no infinity: |t|^2 = pc goes into six exact pieces (bit 9 as 256 + 256, bits 5-8 as two halves of 16 m, bits 1-4, bit 0)
// against eight ones in the query row; a padding row gets 8 x 448 = 3584, above every real score (<= 512 + 0).
template <bool FP8>
__device__ __forceinline__ void write_aug_bits(uint4* dst, bool is_t, unsigned pc, bool pad)
{
    if (!FP8) { write_aug(dst, is_t, pad ? INFINITY : (float)pc, true); return; }
    uint4 w = make_uint4(0, 0, 0, 0);
    if (!is_t) { w.x = 0x38383838u; w.y = 0x38383838u; }
    else if (pad) { w.x = 0x7E7E7E7Eu; w.y = 0x7E7E7E7Eu; }
    else {
        const uint32_t h9 = (pc & 0x200u) ? e4m3_int(256) : 0u, h5 = e4m3_int((pc & 0x1E0u) >> 1);
        w.x = h9 | (h9 << 8) | (h5 << 16) | (h5 << 24);
        w.y = e4m3_int(pc & 0x1Eu) | (e4m3_int(pc & 1u) << 8);
    }
    dst[0] = w; dst[1] = make_uint4(0, 0, 0, 0);
}

template <bool FP8>
__global__ void __launch_bounds__(256)
hamming_prep_kernel(const uint8_t* __restrict__ t, int nt, const uint8_t* __restrict__ q, int nq, int w16, void* __restrict__ tp,
                    void* __restrict__ qp, float* __restrict__ tnorm, int nt_pad, float* __restrict__ qnorm, uint4* __restrict__ taug,
                    uint4* __restrict__ qaug, TcHeader* hdr, uint32_t* __restrict__ gthr, int n_gthr)
{
    const int lane = threadIdx.x & 31, sub = lane >> 3, sl = lane & 7;
    const int warps_per_grid = gridDim.x * (blockDim.x >> 5);
    const int total = nt_pad + nq;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_gthr; i += gridDim.x * blockDim.x) gthr[i] = 0xFFFFFFFFu;
    const int row_bytes = 16 * w16;
    const size_t out_row = (size_t)row_bytes * (FP8 ? 8 : 16);
    unsigned tmax = 0;
    for (int base = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 4; base < total; base += warps_per_grid * 4) {
        const int row = base + sub;
        const bool live = row < total, is_t = row < nt_pad;
        const int r = is_t ? row : row - nt_pad;
        const bool pad = live && is_t && r >= nt;
        unsigned pc = 0;
        if (live && !pad) pc = unpack_bits_row<FP8>((is_t ? t : q) + (size_t)r * row_bytes, row_bytes, sl, is_t, true, (char*)(is_t ? tp : qp) + (size_t)r * out_row);
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) pc += __shfl_xor_sync(0xffffffffu, pc, o);
        if (live && sl == 0) {
            if (is_t) tnorm[r] = pad ? INFINITY : (float)pc; else qnorm[r] = (float)pc;
            write_aug_bits<FP8>((is_t ? taug : qaug) + 2 * (size_t)r, is_t, pc, pad);
        }
        if (live && is_t && !pad) tmax = max(tmax, pc);
    }
    tmax = __reduce_max_sync(0xffffffffu, tmax);
    if (lane == 0 && tmax) atomicMax(&hdr->max_tnorm_bits, __float_as_uint((float)tmax));
}

// ---- 3. tensor-core ranking kernel ----------------------------------------------------------------------
struct TcParams {
    int nq, nt, NT;                  // NT = train tiles of BN rows
    long long U;                     // work units = query blocks x train tiles
    int max_slots;                   // candidate slots per query block
    int rq;                          // query rows per block: BM, or UM for 128 < dim <= MAX_DIM (one UMMA M block per CTA)
    // batch of independent pairs in one launch (Hamming path): query block qb belongs to pair qb / NQBp; a pair's operand rows start
    // at pair * q_pair_rows / pair * t_pair_rows, its aug rows are padded per block / per tile.  NQBp = 0: one problem.
    int NQBp, q_pair_rows, t_pair_rows;
    int slabs16, slabs32;            // K slabs of 128 bytes in bf16 / tf32 mode
    int slabs8;                      // > 0: e4m3 operands (bits of binary descriptors as 0 / 1 bytes), K slabs of 128 elements; exact-integer mode
    uint4* cand;                     // [qblock][slot][LPS][rq rows][2 x uint4] = 4 scores + 4 train indices
    uint32_t* gthr;                  // [NQB * rq] per query: smallest list threshold any CTA has published (order-preserving uint)
    const TcHeader* hdr;
};

// CTA c owns units [c*U/G, (c+1)*U/G); owner(u) = the CTA whose range holds unit u
__host__ __device__ inline int tc_owner(long long u, long long U, int G) { return (int)(((u + 1) * G - 1) / U); }

// Sorted candidate lists in registers (ascending scores; ties keep arrival = train index order).  TopList4 serves the
// general case (the certificate needs the 4th best as tau); TopList2 the exact-integer mode (see l2_certify_kernel).
struct TopList4 {
    float s0, s1, s2, s3; int i0, i1, i2, i3;
    __device__ __forceinline__ void reset() { s0 = s1 = s2 = s3 = INFINITY; i0 = i1 = i2 = i3 = -1; }
    __device__ __forceinline__ float last() const { return s3; }
    // branch-free sorted insertion, active only where p (p implies s < last())
    __device__ __forceinline__ void insert_if(bool p, float s, int i) {
        const bool c2 = p && (s < s2), c1 = c2 && (s < s1), c0 = c1 && (s < s0);
        s3 = p ? (c2 ? s2 : s) : s3;  i3 = p ? (c2 ? i2 : i) : i3;
        s2 = c2 ? (c1 ? s1 : s) : s2; i2 = c2 ? (c1 ? i1 : i) : i2;
        s1 = c1 ? (c0 ? s0 : s) : s1; i1 = c1 ? (c0 ? i0 : i) : i1;
        s0 = c0 ? s : s0;             i0 = c0 ? i : i0;
    }
    __device__ __forceinline__ void store(uint4* dst) const {
        dst[0] = make_uint4(__float_as_uint(s0), __float_as_uint(s1), __float_as_uint(s2), __float_as_uint(s3));
        dst[1] = make_uint4((uint32_t)i0, (uint32_t)i1, (uint32_t)i2, (uint32_t)i3);
    }
};
struct TopList2 {
    float s0, s1; int i0, i1;
    __device__ __forceinline__ void reset() { s0 = s1 = INFINITY; i0 = i1 = -1; }
    __device__ __forceinline__ float last() const { return s1; }
    __device__ __forceinline__ void insert_if(bool p, float s, int i) {
        const bool c0 = p && (s < s0);
        s1 = p ? (c0 ? s0 : s) : s1;  i1 = p ? (c0 ? i0 : i) : i1;
        s0 = c0 ? s : s0;             i0 = c0 ? i : i0;
    }
    __device__ __forceinline__ void store(uint4* dst) const {
        dst[0] = make_uint4(__float_as_uint(s0), __float_as_uint(s1), 0x7F800000u, 0x7F800000u);
        dst[1] = make_uint4((uint32_t)i0, (uint32_t)i1, 0xFFFFFFFFu, 0xFFFFFFFFu);
    }
};

__device__ __forceinline__ void tc_wait_ld_regs(uint32_t (&r)[32]) {
    // the loaded registers become valid here: tie them to the wait so no use can be scheduled above it
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                   "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                 :: "memory");
}

// Epilogue state of one thread (= one query row): the sorted top-NC list in registers, a stale copy of its NC-th
// best score (thr) used as the filter, and a small pending buffer in shared memory.
//
// Fast path (per 32-column chunk): minimum of every group of 4 accumulators, one sign bit per group from
// (min - thr) (FADD on the fma pipe + one funnel shift; the alu pipe, which FMNMX/SEL/ISETP share, is what bounds this
// kernel), OR over the warp.  Slow path: ONE compact loop per call site handles a flagged group for the whole warp:
// re-read its 4 accumulators from TMEM at a dynamic column (registers cannot be indexed dynamically) and APPEND the
// elements below thr to the thread's pending buffer -- no sorted insertion here.  Pending entries are drained into the
// sorted list, lane-parallel, only when some lane's buffer could overflow on the next group (and at the end of the
// segment); thr is refreshed then.  A stale thr is >= the final NC-th best, so every element that is rejected by the
// filter or evicted by the drain scores >= the list's final NC-th best: the certificate's tau stays valid.  Entries are
// drained in arrival (= ascending train index) order with strict '<', i.e. the list is exactly what one-by-one
// insertion would have produced (lower train index wins ties).
// float <-> uint whose unsigned order is the float order (scores can be negative): atomicMin target
__device__ __forceinline__ uint32_t ord_of(float f) { const uint32_t b = __float_as_uint(f); return b ^ ((b >> 31) ? 0xFFFFFFFFu : 0x80000000u); }
__device__ __forceinline__ float ord_to_float(uint32_t o) { return __uint_as_float((o & 0x80000000u) ? (o ^ 0x80000000u) : ~o); }   // 0xFFFFFFFF -> NaN (ignored by fminf)

template <class List>
struct EpiState {
    List L;
    uint32_t* gq;                    // this query's entry of TcParams::gthr
    float gpub;                      // what this thread last published there
    float thr;                       // filter threshold: min(own NC-th best, partner list's published NC-th best), possibly stale
    float ext;                       // the partner's published value as read at the start of the tile
    uint32_t pbase, pptr;            // shared-memory byte addresses: first entry of this thread / next free entry
    uint32_t pub;                    // shared-memory address where this list's threshold is published
    int qb;                          // query block of the current segment (tag of the published threshold)
};

template <class List>
__device__ __forceinline__ void epi_drain(EpiState<List>& S)
{
    const int cnt = (int)((S.pptr - S.pbase) / PEND_STRIDE);
    const int n = __reduce_max_sync(0xffffffffu, cnt);
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
        float sc = INFINITY; int id = -1;
        if (i < cnt) asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=f"(sc), "=r"(id) : "r"(S.pbase + i * PEND_STRIDE) : "memory");
        S.L.insert_if(sc < S.L.last(), sc, id);
    }
    S.pptr = S.pbase;
    S.thr = fminf(S.L.last(), S.ext);
    asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(S.pub), "r"(__float_as_uint(S.L.last())), "r"(S.qb) : "memory");
    // ... and to the CTAs that cover other train ranges of the same query block (and the other UMMA half): every list's
    // last entry bounds the query's final k-th best from above, so all lists may filter with the smallest one published
    if (S.L.last() < S.gpub) { S.gpub = S.L.last(); atomicMin(S.gq, ord_of(S.gpub)); }
}

#ifndef MPL_TC_REG_EVENTS
#define MPL_TC_REG_EVENTS 1          // 0 = round 1's slow path (TMEM re-read of the flagged groups), kept for A/B builds
#endif
#if !MPL_TC_REG_EVENTS
// one bit per group of 4 columns whose minimum is below thr: sign(min - thr) (FADD on the fma pipe) funnelled into the mask
__device__ __forceinline__ unsigned group_mask(const uint32_t (&r)[32], float thr)
{
    unsigned lo = 0, hi = 0;                                                      // two independent chains of 4
#pragma unroll
    for (int g = 3; g >= 0; --g) {
        const float m0 = fminf(fminf(__uint_as_float(r[4 * g]), __uint_as_float(r[4 * g + 1])),
                               fminf(__uint_as_float(r[4 * g + 2]), __uint_as_float(r[4 * g + 3])));
        const float m1 = fminf(fminf(__uint_as_float(r[16 + 4 * g]), __uint_as_float(r[16 + 4 * g + 1])),
                               fminf(__uint_as_float(r[16 + 4 * g + 2]), __uint_as_float(r[16 + 4 * g + 3])));
        // the sign of (m - thr) is set iff m < thr (inf - inf gives the positive canonical NaN: pad rows never flag)
        lo = __funnelshift_l(__float_as_uint(__fsub_rn(m0, thr)), lo, 1);
        hi = __funnelshift_l(__float_as_uint(__fsub_rn(m1, thr)), hi, 1);
    }
    return (lo & 0xFu) | ((hi & 0xFu) << 4);
}

__device__ __forceinline__ void tc_ld4(uint32_t taddr, uint32_t (&v)[4]) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr) : "memory");
}

// Slow path: the flagged groups of a tile half (bit g of mask = columns 4g..4g+3 from col0 / taddr), for the whole warp.
// Up to EVB groups are re-read from TMEM with one round trip, then appended one after the other.
constexpr int EVB = 2;
template <class List>
__device__ __forceinline__ void epi_events(unsigned mask, int col0, uint32_t taddr, EpiState<List>& S)
{
#pragma unroll 1
    while (mask) {
        uint32_t v[EVB][4]; int g[EVB];
#pragma unroll
        for (int j = 0; j < EVB; ++j) {
            g[j] = -1;
            if (mask) {                                                          // warp-uniform
                g[j] = __ffs(mask) - 1;
                mask &= mask - 1;
                tc_ld4(taddr + 4 * g[j], v[j]);
            }
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
        for (int j = 0; j < EVB; ++j) {
            if (g[j] >= 0) {
                // the loaded registers become valid at the wait above: tie them to this point
                asm volatile("" : "+r"(v[j][0]), "+r"(v[j][1]), "+r"(v[j][2]), "+r"(v[j][3]) :: "memory");
                const int col = col0 + 4 * g[j];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    if (__uint_as_float(v[j][e]) < S.thr) {
                        asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(S.pptr), "r"(v[j][e]), "r"(col + e) : "memory");
                        S.pptr += PEND_STRIDE;
                    }
                }
                if (__any_sync(0xffffffffu, S.pptr > S.pbase + (PEND_CAP - 4) * PEND_STRIDE)) epi_drain(S);
            }
        }
    }
}

#endif  // !MPL_TC_REG_EVENTS

// Slow path without the TMEM round trip (round 2): the flagged group index is warp-uniform, so a uniform switch picks the
// group's four accumulators straight out of the registers the fast path has just scanned (no dynamic register indexing, no
// tcgen05.ld + wait: ~40 instead of ~400 clocks per event).  That needs the 32 loaded registers alive during the events,
// which the register budget (96 at 640 threads) only allows for ONE 32-column chunk at a time: the epilogue therefore
// handles the two chunks of its 64 columns one after the other (load, scan, events; load, scan, events).
// Why it matters: an accumulator buffer goes back to the MMA issuer only when the SLOWEST of its 8 epilogue warps is done
// with it; with ~1.3 events per warp and tile the slowest warp had 3 - 4 events, i.e. the old slow path (not the scan)
// set the tile time.
template <class List>
__device__ __forceinline__ void epi_events_reg(unsigned mask, int col0, const uint32_t (&r)[32], EpiState<List>& S)
{
#pragma unroll 1
    while (mask) {                                                               // warp-uniform
        const int g = __ffs(mask) - 1;
        mask &= mask - 1;
        uint32_t v0, v1, v2, v3;
        switch (g) {
            case 0: v0 = r[0]; v1 = r[1]; v2 = r[2]; v3 = r[3]; break;
            case 1: v0 = r[4]; v1 = r[5]; v2 = r[6]; v3 = r[7]; break;
            case 2: v0 = r[8]; v1 = r[9]; v2 = r[10]; v3 = r[11]; break;
            case 3: v0 = r[12]; v1 = r[13]; v2 = r[14]; v3 = r[15]; break;
            case 4: v0 = r[16]; v1 = r[17]; v2 = r[18]; v3 = r[19]; break;
            case 5: v0 = r[20]; v1 = r[21]; v2 = r[22]; v3 = r[23]; break;
            case 6: v0 = r[24]; v1 = r[25]; v2 = r[26]; v3 = r[27]; break;
            default: v0 = r[28]; v1 = r[29]; v2 = r[30]; v3 = r[31]; break;
        }
        const int col = col0 + 4 * g;
        const uint32_t v[4] = {v0, v1, v2, v3};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            if (__uint_as_float(v[e]) < S.thr) {
                asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(S.pptr), "r"(v[e]), "r"(col + e) : "memory");
                S.pptr += PEND_STRIDE;
            }
        }
        if (__any_sync(0xffffffffu, S.pptr > S.pbase + (PEND_CAP - 4) * PEND_STRIDE)) epi_drain(S);
    }
}

// group flags of ONE 32-column chunk: bit g = columns 4g .. 4g+3 hold something below thr
__device__ __forceinline__ unsigned group_mask32(const uint32_t (&r)[32], float thr)
{
    unsigned lo = 0, hi = 0;                                                      // two independent chains of 4
#pragma unroll
    for (int g = 3; g >= 0; --g) {
        const float m0 = fminf(fminf(__uint_as_float(r[4 * g]), __uint_as_float(r[4 * g + 1])),
                               fminf(__uint_as_float(r[4 * g + 2]), __uint_as_float(r[4 * g + 3])));
        const float m1 = fminf(fminf(__uint_as_float(r[16 + 4 * g]), __uint_as_float(r[16 + 4 * g + 1])),
                               fminf(__uint_as_float(r[16 + 4 * g + 2]), __uint_as_float(r[16 + 4 * g + 3])));
        lo = __funnelshift_l(__float_as_uint(__fsub_rn(m0, thr)), lo, 1);
        hi = __funnelshift_l(__float_as_uint(__fsub_rn(m1, thr)), hi, 1);
    }
    return (lo & 0xFu) | ((hi & 0xFu) << 4);
}

// MMA issuer of ONE UMMA M block (two issuer warps per CTA, m = 0 and m = 1: the scalar work between two tcgen05.mma --
// barrier waits, descriptor arithmetic, commits -- costs a single warp more time per tile than the tensor pipe needs for
// the tile's MMAs, so it is split).  The whole warp runs the loop (uniform control flow), one elected lane issues.
// Accumulators are handed over per M block: t_full / t_empty index = accumulator buffer * 2 + m; a B slab is released
// when BOTH issuers' MMAs on it have retired (b_empty counts 2 commits), likewise the resident A block (a_empty).
template <int MODE, int SPS>   // MODE 0 = bf16 (kind::f16), 1 = tf32, 2 = e4m3 (kind::f8f6f4); SPS = K slabs per B ring stage (one barrier round trip per stage)
__device__ __forceinline__ void mma_issue_loop(int m, long long u0, long long u1, int NT, int n_slabs, int n_stages, uint32_t a_slab16, uint32_t a_addr, uint32_t aa_addr,
                                               uint32_t b_addr, uint32_t tmem_base, uint64_t* a_full, uint64_t* a_empty, uint64_t* b_full,
                                               uint64_t* b_empty, uint64_t* t_full, uint64_t* t_empty)
{
    constexpr uint32_t idesc = umma_idesc(MODE == 1 ? 2u : MODE == 2 ? 0u : 1u);
    constexpr uint64_t hi128 = ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);     // see umma_desc()
    constexpr uint64_t hi32 = ((uint64_t)(256 >> 4) << 32) | (1ull << 46) | (6ull << 61);       // see umma_desc_sw32()
    const uint64_t ada = hi32 | (uint64_t)((aa_addr + m * (UM * AUG_BYTES)) >> 4);
    const uint64_t ad0 = hi128 | (uint64_t)((a_addr + m * (UM * SLAB_BYTES)) >> 4);
    int stage = 0; uint32_t sphase = 0, a_phase = 0; unsigned it = 0;
    int tile = (int)(u0 - (u0 / NT) * NT);                                       // tile index inside the current query block
    bool new_qb = true;
#ifdef MPL_TC_LAG
    if (m == 1) { const long long t_go = clock64() + MPL_TC_LAG; while (clock64() < t_go) { } }   // experiment: de-phase the two issuers
#endif
#ifdef MPL_TC_TRACE
    long long w_te = 0, w_bf = 0, w_af = 0, w_is = 0, t_start = clock64(), tq;
#define MPL_TWAIT(acc, stmt) do { tq = clock64(); stmt; acc += clock64() - tq; } while (0)
#else
#define MPL_TWAIT(acc, stmt) stmt
#endif
    for (long long u = u0; u < u1; ++u, ++it) {
        if (new_qb) { MPL_TWAIT(w_af, mbar_wait(a_full, a_phase)); a_phase ^= 1; new_qb = false; }
        const bool last_of_qb = (++tile == NT);
        if (last_of_qb) { tile = 0; new_qb = true; }
        const uint32_t b = it & 1;
        const uint32_t d = tmem_base + b * 256 + m * 128;
        MPL_TWAIT(w_te, mbar_wait(&t_empty[b * 2 + m], ((it >> 1) & 1) ^ 1));    // epilogue drained this accumulator block
        tc_fence_after();
        for (int s = 0; s < n_slabs; s += SPS) {
            MPL_TWAIT(w_bf, mbar_wait(&b_full[stage], sphase));
            tc_fence_after();
#ifdef MPL_TC_TRACE
            tq = clock64();
#endif
            if (elect_one()) {
                const uint32_t sb = b_addr + stage * (SPS * B_SLAB + BN * AUG_BYTES);
                if (s == 0) {                                                    // aug K step first: D = |t|^2 (overwrites)
                    const uint64_t bda = hi32 | (uint64_t)((sb + SPS * B_SLAB) >> 4);
                    if (MODE == 1) tc_mma_tf32(d, ada, bda, idesc, 0u); else if (MODE == 2) tc_mma_f8(d, ada, bda, idesc, 0u); else tc_mma_f16(d, ada, bda, idesc, 0u);
                }
#pragma unroll
                for (int j = 0; j < SPS; ++j) {
                    const uint64_t bd = hi128 | (uint64_t)((sb + j * B_SLAB) >> 4);
                    const uint64_t ad = ad0 + (uint64_t)((uint32_t)(s + j) * a_slab16);
#pragma unroll
                    for (int k = 0; k < 4; ++k) {                                // 4 x 32 bytes of K per slab row
                        if (MODE == 1) tc_mma_tf32(d, ad + 2 * k, bd + 2 * k, idesc, 1u);
                        else if (MODE == 2) tc_mma_f8(d, ad + 2 * k, bd + 2 * k, idesc, 1u);
                        else tc_mma_f16(d, ad + 2 * k, bd + 2 * k, idesc, 1u);
                    }
                }
                tc_commit(&b_empty[stage]);                                      // stage free once both issuers' MMAs retire
                if (s + SPS >= n_slabs) {
                    tc_commit(&t_full[b * 2 + m]);
                    if (u + 1 < u1 && last_of_qb) tc_commit(a_empty);            // only when the producer will wait for it
                }
            }
            __syncwarp();
#ifdef MPL_TC_TRACE
            w_is += clock64() - tq;
#endif
            if (++stage == n_stages) { stage = 0; sphase ^= 1; }
        }
    }
#ifdef MPL_TC_TRACE
    if (blockIdx.x == 3 && (threadIdx.x & 31) == 0)
        printf("mma warp m=%d: total %lld cycles for %lld tiles; waits: t_empty %lld  b_full %lld  a_full %lld; issue sections %lld\n",
               m, clock64() - t_start, u1 - u0, w_te, w_bf, w_af, w_is);
#endif
#undef MPL_TWAIT
}

// Epilogue warps 4..19: thread = one query row of UMMA block m, one half (64 columns) of the tile.
template <class List, bool EXACT>
__device__ __forceinline__ void epilogue_loop(const TcParams& p, long long u0, long long u1, int G, uint32_t tmem_base,
                                              uint32_t share_s, uint32_t pend_s, uint64_t* t_full, uint64_t* t_empty)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int wg = (warp - 4) >> 2, m = wg & 1, hf = wg >> 1, quad = warp & 3, row = quad * 32 + lane;
    EpiState<List> S;
    S.L.reset(); S.thr = INFINITY; S.ext = INFINITY; S.gpub = INFINITY;
    S.pbase = S.pptr = pend_s + (uint32_t)(threadIdx.x - 128) * 8u;
    S.pub = share_s + 8u * (uint32_t)(hf * BM + m * UM + row);
    const uint32_t partner = share_s + 8u * (uint32_t)((hf ^ 1) * BM + m * UM + row);
    int tile = (int)(u0 - (u0 / p.NT) * p.NT);
    S.qb = (int)(u0 / p.NT);
    S.gq = p.gthr + (size_t)S.qb * p.rq + m * UM + row;
    unsigned it = 0;
    auto flush = [&]() {
        epi_drain(S);
        const int slot = (int)blockIdx.x - tc_owner((long long)S.qb * p.NT, p.U, G);
        S.L.store(p.cand + ((((size_t)S.qb * p.max_slots + slot) * LPS + hf) * p.rq + m * UM + row) * 2);
        S.L.reset(); S.thr = INFINITY; S.ext = INFINITY; S.gpub = INFINITY;
        S.gq += p.rq;                                                            // flush() ends a query block
    };
    const uint32_t tbase = tmem_base + ((uint32_t)(quad * 32) << 16) + m * 128 + hf * 64;
#ifdef MPL_TC_TRACE
    long long e_wait = 0, e_fast = 0, e_slow = 0, e_start = clock64(), e_q; int e_groups = 0;
#endif
    for (long long u = u0; u < u1; ++u, ++it) {
        const uint32_t b = it & 1;
        const uint32_t taddr = tbase + b * 256;
        const uint32_t gv = __ldcg(S.gq);                                        // issued before the wait: latency hidden
#ifdef MPL_TC_TRACE
        e_q = clock64();
#endif
        mbar_wait(&t_full[b * 2 + m], (it >> 1) & 1);
#ifdef MPL_TC_TRACE
        e_wait += clock64() - e_q; e_q = clock64();
#endif
        tc_fence_after();
#if defined(MPL_TC_DEBUG) && MPL_TC_DEBUG == 1
        const int col0 = tile * BN + hf * 64; (void)col0; (void)gv; (void)partner; (void)taddr;
#elif MPL_TC_REG_EVENTS
        uint32_t ra[32];
        tc_ld32(taddr, ra);
        const int col0 = tile * BN + hf * 64;
        float ox; int oq;
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=f"(ox), "=r"(oq) : "r"(partner) : "memory");
        float gx = ord_to_float(gv);                                             // NaN while nothing is published
        if (EXACT) { ox = __fadd_rn(ox, 1.0f); gx = __fadd_rn(gx, 1.0f); }
        S.ext = fminf((oq == S.qb) ? ox : INFINITY, gx);
        S.thr = fminf(S.thr, S.ext);
        tc_wait_ld_regs(ra);
        {
            const unsigned mask = __reduce_or_sync(0xffffffffu, group_mask32(ra, S.thr));
#ifdef MPL_TC_TRACE
            e_groups += __popc(mask);
#endif
            epi_events_reg(mask, col0, ra, S);
        }
        tc_ld32(taddr + 32, ra);
        tc_wait_ld_regs(ra);
        {
            const unsigned mask = __reduce_or_sync(0xffffffffu, group_mask32(ra, S.thr));
#ifdef MPL_TC_TRACE
            e_groups += __popc(mask);
#endif
            epi_events_reg(mask, col0 + 32, ra, S);
        }
#else
        uint32_t ra[32], rb[32];
        tc_ld32(taddr, ra);
        tc_ld32(taddr + 32, rb);
        const int col0 = tile * BN + hf * 64;
        // The other column half keeps its own list for the same query: filter with the minimum of both lists' last
        // entries.  General mode: any threshold >= the final tau keeps the certificate valid.  Exact mode: an element
        // may only be dropped if two others are STRICTLY better or equal with a lower train index; the partner's
        // columns are not ordered against ours, so its threshold is relaxed by one (scores are integers there).  The
        // query-block tag guards against a stale value from the partner's previous segment.
        float ox; int oq;
        asm volatile("ld.shared.v2.b32 {%0, %1}, [%2];" : "=f"(ox), "=r"(oq) : "r"(partner) : "memory");
        float gx = ord_to_float(gv);                                             // NaN while nothing is published
        if (EXACT) { ox = __fadd_rn(ox, 1.0f); gx = __fadd_rn(gx, 1.0f); }
        S.ext = fminf((oq == S.qb) ? ox : INFINITY, gx);
        S.thr = fminf(S.thr, S.ext);
        tc_wait_ld_regs(ra); tc_wait_ld_regs(rb);
        unsigned mask = group_mask(ra, S.thr) | (group_mask(rb, S.thr) << 8);
        mask = __reduce_or_sync(0xffffffffu, mask);                               // warp-uniform: tcgen05.ld is .aligned
#ifdef MPL_TC_TRACE
        e_fast += clock64() - e_q; e_q = clock64(); e_groups += __popc(mask);
#endif
        epi_events(mask, col0, taddr, S);
#ifdef MPL_TC_TRACE
        e_slow += clock64() - e_q;
#endif
#endif
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&t_empty[b * 2 + m]);
        if (++tile == p.NT) { flush(); tile = 0; ++S.qb; }
    }
    if (tile != 0) flush();
#ifdef MPL_TC_TRACE
    if (blockIdx.x == 3 && lane == 0 && (warp == 4 || warp == 19))
        printf("epilogue warp %d: total %lld cycles for %lld tiles; wait t_full %lld  fast path %lld  slow path %lld  flagged groups %d\n",
               warp, clock64() - e_start, u1 - u0, e_wait, e_fast, e_slow, e_groups);
#endif
}

struct TcMaps { CUtensorMap a16, b16, a32, b32, aa16, ba16, aa32, ba32, a8, b8, aa8, ba8; };   // main slabs (128B swizzle) and aug slabs (32B swizzle)

__global__ void __launch_bounds__(TC_THREADS, 1)
l2_tc_rank_kernel(const __grid_constant__ TcMaps maps, const TcParams p)
{
    extern __shared__ unsigned char smem_dyn[];
    unsigned char* smem = (unsigned char*)(((uintptr_t)smem_dyn + 1023) & ~(uintptr_t)1023);   // swizzle-128B atoms: 1024-byte aligned
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + SMEM_DATA);
    uint64_t* a_full = bars + 0; uint64_t* a_empty = bars + 1;
    uint64_t* b_full = bars + 2; uint64_t* b_empty = bars + 2 + MAX_STAGES;
    uint64_t* t_full = bars + 2 + 2 * MAX_STAGES; uint64_t* t_empty = t_full + 4;      // [accumulator buffer][UMMA M block]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(t_empty + 4);
    const uint32_t share_s = smem_u32(smem + SMEM_DATA + 256);       // [LPS][BM] x (NC-th best score, query block)
    const uint32_t pend_s = share_s + SMEM_SHARE;                    // [PEND_CAP][epilogue threads] x (score, train index)

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool fp8 = p.slabs8 > 0;
    const bool tf32 = !fp8 && p.hdr->not_bf16_exact != 0;
    const int n_slabs = fp8 ? p.slabs8 : tf32 ? p.slabs32 : p.slabs16;
    const int KS = fp8 ? 128 : tf32 ? 32 : 64;                   // elements per 128-byte slab row
    const CUtensorMap* mapA = fp8 ? &maps.a8 : tf32 ? &maps.a32 : &maps.a16;
    const CUtensorMap* mapB = fp8 ? &maps.b8 : tf32 ? &maps.b32 : &maps.b16;
    const CUtensorMap* mapAa = fp8 ? &maps.aa8 : tf32 ? &maps.aa32 : &maps.aa16;
    const CUtensorMap* mapBa = fp8 ? &maps.ba8 : tf32 ? &maps.ba32 : &maps.ba16;
    const bool wide = p.rq == UM;                                // 128 < dim <= 256: query blocks of one UMMA M block
    const int a_slab = p.rq * SLAB_BYTES, a_aug = p.rq * AUG_BYTES;
    unsigned char* sA = smem;                                    // n_slabs x 32 KB (wide: 16 KB)
    unsigned char* sAa = smem + n_slabs * a_slab;                // 8 KB (wide: 4 KB)
    unsigned char* sB = sAa + a_aug;                             // ring of n_stages x 20 KB
    // B ring stage = `sps` K slabs of one train tile + the tile's aug slab; two slabs per stage (a whole tile at dim <= 128
    // in bf16 mode) when at least 3 such stages fit: half the barrier round trips and commits on the issuers' critical path
    const int ring_bytes = SMEM_DATA - n_slabs * a_slab - a_aug;
    const int sps = ((n_slabs & 1) == 0 && ring_bytes / (2 * B_SLAB + BN * AUG_BYTES) >= 3) ? 2 : 1;
    const int stage_bytes = sps * B_SLAB + BN * AUG_BYTES;
    int n_stages = ring_bytes / stage_bytes;
    if (n_stages > MAX_STAGES) n_stages = MAX_STAGES;
    const int G = gridDim.x;
    const long long u0 = (long long)blockIdx.x * p.U / G, u1 = (long long)(blockIdx.x + 1) * p.U / G;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(mapA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(mapB) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(mapAa) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(mapBa) : "memory");
    }
    if (warp == 1 && lane == 0) {
        const uint32_t n_issuers = wide ? 1u : 2u;
        mbar_init(a_full, 1); mbar_init(a_empty, n_issuers);         // *_empty: one commit per MMA issuer warp
        for (int s = 0; s < MAX_STAGES; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], n_issuers); }
        for (int b = 0; b < 4; ++b) { mbar_init(&t_full[b], 1); mbar_init(&t_empty[b], EPI_WARPS / 2); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {                                             // whole warp: allocate all 512 TMEM columns
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    for (int i = threadIdx.x; i < LPS * BM; i += TC_THREADS)
        asm volatile("st.shared.v2.b32 [%0], {%1, %2};" ::"r"(share_s + 8 * i), "r"(0x7F800000u), "r"(-1) : "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===== TMA producer: the whole warp runs the loop (uniform control flow), one elected lane issues =====
        int stage = 0; uint32_t sphase = 0; int a_cnt = 0, cur_qb = -1, b_row0 = 0, ba_row0 = 0;
        int qb = (int)(u0 / p.NT), tile = (int)(u0 - (long long)qb * p.NT) - 1;
        for (long long u = u0; u < u1; ++u) {
            if (++tile == p.NT) { tile = 0; ++qb; }
            if (qb != cur_qb) {
                mbar_wait(a_empty, (uint32_t)((a_cnt & 1) ^ 1));                 // MMAs of the previous query block are done
                int a_row = qb * p.rq;
                if (p.NQBp) { const int pair = qb / p.NQBp; a_row = pair * p.q_pair_rows + (qb - pair * p.NQBp) * p.rq; b_row0 = pair * p.t_pair_rows; ba_row0 = pair * p.NT * BN; }
                if (elect_one()) {
                    mbar_expect_tx(a_full, (uint32_t)(n_slabs * a_slab + a_aug));
                    for (int s = 0; s < n_slabs; ++s) tma_load_2d(sA + s * a_slab, mapA, s * KS, a_row, a_full);
                    tma_load_2d(sAa, mapAa, 0, qb * p.rq, a_full);
                }
                __syncwarp();
                ++a_cnt; cur_qb = qb;
            }
            for (int s = 0; s < n_slabs; s += sps) {
                mbar_wait(&b_empty[stage], sphase ^ 1);
                if (elect_one()) {
                    unsigned char* dst = sB + stage * stage_bytes;
                    mbar_expect_tx(&b_full[stage], (uint32_t)(sps * B_SLAB + (s == 0 ? BN * AUG_BYTES : 0)));
                    for (int j = 0; j < sps; ++j) tma_load_2d(dst + j * B_SLAB, mapB, (s + j) * KS, b_row0 + tile * BN, &b_full[stage]);
                    if (s == 0) tma_load_2d(dst + sps * B_SLAB, mapBa, 0, ba_row0 + tile * BN, &b_full[stage]);
                }
                __syncwarp();
                if (++stage == n_stages) { stage = 0; sphase ^= 1; }
            }
        }
    } else if (warp == 1 || (warp == 3 && !wide)) {
        // ===== MMA issuers: warp 1 -> UMMA M block 0, warp 3 -> M block 1 =====
        const int m = warp >> 1;
#define MPL_ISSUE(MODE, SPS) mma_issue_loop<MODE, SPS>(m, u0, u1, p.NT, n_slabs, n_stages, (uint32_t)(a_slab >> 4), smem_u32(sA), smem_u32(sAa), smem_u32(sB), tmem_base, \
                                                       a_full, a_empty, b_full, b_empty, t_full, t_empty)
        if (fp8)       { if (sps == 2) MPL_ISSUE(2, 2); else MPL_ISSUE(2, 1); }
        else if (tf32) { if (sps == 2) MPL_ISSUE(1, 2); else MPL_ISSUE(1, 1); }
        else           { if (sps == 2) MPL_ISSUE(0, 2); else MPL_ISSUE(0, 1); }
#undef MPL_ISSUE
    } else if (warp >= 4 && !(wide && (((warp - 4) >> 2) & 1))) {
        // ===== epilogue: thread = one query row of UMMA block m, one half (64 columns) of the tile (wide: m = 0 only) =====
        // Exact-integer mode (bf16-exact integer data, e.g. OpenCV SIFT): the scores are exact, so sorted top-2 lists give
        // the exact answer and no certificate margin is needed -- half as many insertions as the general top-4 lists.
        if (!tf32 && p.hdr->not_integer == 0)
            epilogue_loop<TopList2, true>(p, u0, u1, G, tmem_base, share_s, pend_s, t_full, t_empty);
        else
            epilogue_loop<TopList4, false>(p, u0, u1, G, tmem_base, share_s, pend_s, t_full, t_empty);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
}

// ---- 4. exact re-rank + certificate --------------------------------------------------------------------
// cvflann::L2<float> order (MODE 0): result += ((d0^2 + d1^2) + d2^2) + d3^2 per group of 4, scalar tail;
// NMSLIB L2NormSIMD order (MODE 1): 4 lanes, ((l0 + l1) + l2) + l3, scalar tail, sqrtf.
// The whole warp evaluates ONE pair: lane l owns the group of four dims 4l..4l+3 of each 128-dim round (coalesced
// 512-byte row reads), then every lane replays the reference's sequential accumulation over the groups in ascending
// order with warp shuffles -- same operations, same order, same roundings as the reference's scalar loop.
template <int MODE>
__device__ __forceinline__ float warp_exact_dist(const float* __restrict__ a /* train row */, const float* __restrict__ b /* query row */,
                                                 int dim, int lane, bool vec_ok)
{
    const int dim4 = dim & ~3;
    float r = 0.f, l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
    for (int base = 0; base < dim4; base += 128) {
        const int i = base + 4 * lane;
        float p0 = 0.f, p1 = 0.f, p2 = 0.f, p3 = 0.f;
        if (i < dim4) {
            float4 x, y;
            if (vec_ok) { x = *reinterpret_cast<const float4*>(a + i); y = *reinterpret_cast<const float4*>(b + i); }
            else { x = make_float4(a[i], a[i + 1], a[i + 2], a[i + 3]); y = make_float4(b[i], b[i + 1], b[i + 2], b[i + 3]); }
            const float d0 = __fsub_rn(x.x, y.x), d1 = __fsub_rn(x.y, y.y), d2 = __fsub_rn(x.z, y.z), d3 = __fsub_rn(x.w, y.w);
            p0 = __fmul_rn(d0, d0); p1 = __fmul_rn(d1, d1); p2 = __fmul_rn(d2, d2); p3 = __fmul_rn(d3, d3);
        }
        const int groups = min(32, (dim4 - base) >> 2);
        if (MODE == 0) {
            const float s = __fadd_rn(__fadd_rn(__fadd_rn(p0, p1), p2), p3);
            for (int g = 0; g < groups; ++g) r = __fadd_rn(r, __shfl_sync(0xffffffffu, s, g));
        } else {
            for (int g = 0; g < groups; ++g) {
                l0 = __fadd_rn(l0, __shfl_sync(0xffffffffu, p0, g)); l1 = __fadd_rn(l1, __shfl_sync(0xffffffffu, p1, g));
                l2 = __fadd_rn(l2, __shfl_sync(0xffffffffu, p2, g)); l3 = __fadd_rn(l3, __shfl_sync(0xffffffffu, p3, g));
            }
        }
    }
    if (MODE == 1) r = __fadd_rn(__fadd_rn(__fadd_rn(l0, l1), l2), l3);
    for (int i = dim4; i < dim; ++i) { const float d = __fsub_rn(a[i], b[i]); r = __fadd_rn(r, __fmul_rn(d, d)); }
    return MODE == 1 ? sqrtf(r) : r;
}

template <int MODE>
__global__ void __launch_bounds__(256)
l2_certify_kernel(const float* __restrict__ q, int nq, size_t q_ld, const float* __restrict__ t, int nt, size_t t_ld, int dim,
                  const float* __restrict__ qnorm, const uint4* __restrict__ cand, int NT, long long U, int G, int max_slots,
                  long long train_offset, TcHeader* hdr, mpl_key64* __restrict__ keys_out, int* __restrict__ fallback_list, int rq)
{
    if (hdr->not_bf16_exact == 0 && hdr->not_integer == 0) {
        // Exact-integer mode (the ranking kernel kept top-2 lists): every value is an integer that bf16 holds exactly.  If
        // (|q| + |t|max)^2 < 2^24 as well, every product, partial sum and distance on both sides is an integer below 2^24,
        // i.e. exact in fp32 in any order: score + |q|^2 IS the reference distance, the union of the lists' top-2 (ties
        // kept by lowest train index) contains the two nearest, no margin is needed and no descriptor row is read again.
        // One THREAD per query (the first ceil(nq / 256) blocks; the others have nothing to do), coalesced 32-byte
        // records.  A query outside the bound cannot be certified from top-2 lists and takes the exact scan.
        const int q1 = blockIdx.x * blockDim.x + threadIdx.x;
        if (q1 >= nq) return;
        const int qb1 = q1 / rq, row1 = q1 - qb1 * rq;
        const int f1 = tc_owner((long long)qb1 * NT, U, G), l1 = tc_owner((long long)qb1 * NT + NT - 1, U, G);
        const float qq1 = qnorm[q1];
        const float sq1 = sqrtf(qq1), st1 = sqrtf(__uint_as_float(hdr->max_tnorm_bits));
        if (!((sq1 + st1) * (sq1 + st1) < 16777216.f * 0.999f)) { fallback_list[atomicAdd(&hdr->fallback_count, 1u)] = q1; return; }
        mpl_key64 k0 = MPL_KEY_INVALID, k1 = MPL_KEY_INVALID;
        for (int sl = 0; sl < (l1 - f1 + 1) * LPS; ++sl) {
            const uint4* rec = cand + (((size_t)qb1 * max_slots * LPS + sl) * rq + row1) * 2;
            const uint4 sc = rec[0], id = rec[1];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int ti = (int)(e == 0 ? id.x : id.y);
                if (ti < 0 || ti >= nt) continue;
                float d = __fadd_rn(__uint_as_float(e == 0 ? sc.x : sc.y), qq1);
                if (MODE == 1) d = sqrtf(d);
                // MODE 2 (binary descriptors): d = |t| + |q| - 2 q.t is the Hamming distance, an exact integer; integer key like the popc kernel's
                const uint32_t dbits = MODE == 2 ? __float2uint_rn(d) : __float_as_uint(d);
                const mpl_key64 key = ((mpl_key64)dbits << 32) | (mpl_key64)(uint32_t)(train_offset + ti);
                const mpl_key64 hi = max(k0, key);
                k0 = min(k0, key); k1 = min(k1, hi);
            }
        }
        reinterpret_cast<ulonglong2*>(keys_out)[q1] = make_ulonglong2(k0, k1);
        return;
    }
    if (MODE == 2) return;                                       // 0 / 1 data is always in exact-integer mode
    const int lane = threadIdx.x & 31;
    const int qi = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (qi >= nq) return;
    const int qb = qi / rq, row = qi - qb * rq;
    const int first = tc_owner((long long)qb * NT, U, G), last = tc_owner((long long)qb * NT + NT - 1, U, G);
    const int n_cand = (last - first + 1) * LPS * NC;          // lists of this query block are contiguous: (slot, half)
    const float qq = qnorm[qi];
    const float tmax = __uint_as_float(hdr->max_tnorm_bits);
    // eps >= |score + |q|^2 - d_ref|: operand quantisation (tf32 truncation, none in bf16-exact mode) plus a generous
    // bound on fp32 accumulation error on both sides
    const float sq = sqrtf(qq), st = sqrtf(tmax);
    float eps = 1.2207031e-4f * (sq + st) * (sq + st);                            // 2^-13 (|q| + |t|max)^2
    if (hdr->not_bf16_exact) eps += 3.90625e-3f * sq * st;                        // 2^-8 |q| |t|max
    constexpr int nc = NC;                                        // general mode: top-4 lists, certificate below
    // pass A: the two smallest scores over all candidates, and tau = min over slots of the slot's 4th best
    float s_a = INFINITY, s_b = INFINITY, tau = INFINITY;
    for (int c = lane; c < n_cand; c += 32) {
        const int slot = c / NC, e = c - slot * NC;
        const uint4* rec = cand + (((size_t)qb * max_slots * LPS + slot) * rq + row) * 2;
        const uint4 sc = rec[0];
        const float s = __uint_as_float(e == 0 ? sc.x : e == 1 ? sc.y : e == 2 ? sc.z : sc.w);
        if (e == nc - 1) tau = fminf(tau, s);
        const float hi = fmaxf(s_a, s); s_a = fminf(s_a, s); s_b = fminf(s_b, hi);
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const float oa = __shfl_xor_sync(0xffffffffu, s_a, off), ob = __shfl_xor_sync(0xffffffffu, s_b, off);
        const float hi = fmaxf(s_a, oa);
        s_a = fminf(s_a, oa); s_b = fminf(fminf(s_b, ob), hi);
        tau = fminf(tau, __shfl_xor_sync(0xffffffffu, tau, off));
    }
    // pass B: exact distance of every candidate that can still be one of the two nearest
    const float cut = s_b + 2.f * eps;
    mpl_key64 b0 = MPL_KEY_INVALID, b1 = MPL_KEY_INVALID;
    const float* qrow = q + (size_t)qi * q_ld;
    const bool vec_ok = (((uintptr_t)q | (uintptr_t)t) & 15) == 0 && ((q_ld | t_ld) & 3) == 0;
    for (int c0 = 0; c0 < n_cand; c0 += 32) {
        const int c = c0 + lane;
        int ti = -1; float s = INFINITY;
        if (c < n_cand) {
            const int slot = c / NC, e = c - slot * NC;
            const uint4* rec = cand + (((size_t)qb * max_slots * LPS + slot) * rq + row) * 2;
            const uint4 sc = rec[0], id = rec[1];
            s = __uint_as_float(e == 0 ? sc.x : e == 1 ? sc.y : e == 2 ? sc.z : sc.w);
            ti = (int)(e == 0 ? id.x : e == 1 ? id.y : e == 2 ? id.z : id.w);
        }
        unsigned todo = __ballot_sync(0xffffffffu, ti >= 0 && ti < nt && s <= cut);
        while (todo) {                                                            // warp-uniform: one pair at a time
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const int tj = __shfl_sync(0xffffffffu, ti, src);
            const float d = warp_exact_dist<MODE>(t + (size_t)tj * t_ld, qrow, dim, lane, vec_ok);
            const mpl_key64 key = ((mpl_key64)__float_as_uint(d) << 32) | (mpl_key64)(uint32_t)(train_offset + tj);
            const mpl_key64 hi = max(b0, key);
            b0 = min(b0, key); b1 = min(b1, hi);                                  // identical in every lane
        }
    }
    if (lane == 0) {
        keys_out[2 * (size_t)qi] = b0; keys_out[2 * (size_t)qi + 1] = b1;
        // certificate: every non-candidate has score >= tau, hence reference distance >= tau + |q|^2 - eps
        bool ok;
        if (b1 == MPL_KEY_INVALID) ok = (tau == INFINITY);                         // fewer than 2 train rows scanned in total
        else {
            float d2 = __uint_as_float((uint32_t)(b1 >> 32));
            if (MODE == 1) d2 = __fmul_rn(d2, d2) * 1.000002f;                     // back to the squared domain, rounding margin
            ok = d2 < tau + qq - eps;
        }
        if (!ok) fallback_list[atomicAdd(&hdr->fallback_count, 1u)] = qi;
    }
}

// ---- host side ------------------------------------------------------------------------------------------
struct TcPlan {
    int NQB, NT, G, max_slots, kp16, kp32, rq;
    long long U;
    size_t off_tnorm, off_qnorm, off_tp, off_qp, off_taug, off_qaug, off_cand, off_list, off_part, off_gthr, total;
};

inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

TcPlan make_plan(int nq, int nt, int dim, int sm_count, int bits_fmt = 0)   // bits_fmt: 0 float rows; binary rows as 1 = bf16, 2 = e4m3
{
    TcPlan P;
    P.rq = dim > (bits_fmt == 2 ? 512 : bits_fmt == 1 ? 256 : 128) ? UM : BM;     // resident A block: 4 slabs of BM rows, or 8 of UM rows
    P.NQB = (nq + P.rq - 1) / P.rq;
    P.NT = (nt + BN - 1) / BN;
    P.U = (long long)P.NQB * P.NT;
    P.G = (int)(P.U < sm_count ? P.U : sm_count);
    P.kp16 = ((dim + 63) / 64) * 64;
    P.kp32 = ((dim + 31) / 32) * 32;
    P.max_slots = 1;
    for (int qb = 0; qb < P.NQB; ++qb) {
        const int s = tc_owner((long long)qb * P.NT + P.NT - 1, P.U, P.G) - tc_owner((long long)qb * P.NT, P.U, P.G) + 1;
        if (s > P.max_slots) P.max_slots = s;
    }
    size_t o = 256;
    P.off_tnorm = o; o = align256(o + (size_t)P.NT * BN * 4);
    P.off_qnorm = o; o = align256(o + (size_t)P.NQB * P.rq * 4);
    const size_t row_bytes = bits_fmt == 2 ? (size_t)dim : bits_fmt == 1 ? (size_t)P.kp16 * 2 : (size_t)P.kp32 * 4;
    P.off_tp = o;    o = align256(o + (size_t)nt * row_bytes);
    P.off_qp = o;    o = align256(o + (size_t)nq * row_bytes);
    P.off_taug = o;  o = align256(o + (size_t)P.NT * BN * AUG_BYTES);
    P.off_qaug = o;  o = align256(o + (size_t)nq * AUG_BYTES);
    P.off_cand = o;  o = align256(o + (size_t)P.NQB * P.max_slots * LPS * P.rq * 32);
    P.off_list = o;  o = align256(o + (size_t)nq * 4);
    P.off_part = o;  o = align256(o + mpl_l2_list_scratch_entries(nq, sm_count) * 16);
    P.off_gthr = o;  o = align256(o + (size_t)P.NQB * P.rq * 4);
    P.total = o;
    return P;
}

int encode_map(CUtensorMap* map, CUtensorMapDataType dt, int elt_bytes, void* base, int kp, int rows, int box_k, int box_rows,
               CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B)
{
    cuuint64_t dims[2] = {(cuuint64_t)kp, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)kp * elt_bytes};
    cuuint32_t box[2] = {(cuuint32_t)box_k, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    // libcuda is reached through the runtime (no link-time dependency: the library must still load on a box
    // without a driver so that it can fail loudly with MPL_ERR_NO_DEVICE instead of a dlopen error)
    typedef CUresult (*encode_fn_t)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static encode_fn_t encode = nullptr;
    if (!encode) {
        void* fn = nullptr;
        cudaDriverEntryPointQueryResult qres;
        MPL_CUDA_CHECK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres));
        if (!fn || qres != cudaDriverEntryPointSuccess) { mpl_set_error("cuTensorMapEncodeTiled not available from the driver"); return MPL_ERR_CUDA; }
        encode = (encode_fn_t)fn;
    }
    const CUresult r = encode(map, dt, 2, base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                              swz, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { mpl_set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return MPL_ERR_CUDA; }
    return MPL_OK;
}

}  // namespace

bool mpl_l2tc_supported(int nq, int nt, int dim) { return nq >= 1 && nt >= 2 && dim >= 1 && dim <= MAX_DIM; }

size_t mpl_l2tc_scratch_bytes(int nq, int nt, int dim, int sm_count)
{
    if (!mpl_l2tc_supported(nq, nt, dim)) return 16;
    return make_plan(nq, nt, dim, sm_count).total;
}

int mpl_l2tc_top2(const float* d_q, int nq, size_t q_ld, const float* d_t, int nt, size_t t_ld, int dim, int sqrt_out,
                  long long train_offset, void* d_scratch, size_t scratch_bytes, mpl_key64* d_keys_out, int sm_count, cudaStream_t st)
{
    if (!mpl_l2tc_supported(nq, nt, dim)) { mpl_set_error("l2 tensor-core path: unsupported shape (dim=%d)", dim); return MPL_ERR_BAD_ARG; }
    const TcPlan P = make_plan(nq, nt, dim, sm_count);
    if (!d_scratch || scratch_bytes < P.total) { mpl_set_error("l2 tensor-core path: scratch %zu < %zu bytes", scratch_bytes, P.total); return MPL_ERR_BAD_ARG; }
    if (((uintptr_t)d_scratch & 255) != 0) { mpl_set_error("l2 tensor-core path: scratch must be 256-byte aligned"); return MPL_ERR_BAD_ARG; }
    char* base = (char*)d_scratch;
    TcHeader* hdr = (TcHeader*)base;
    float* tnorm = (float*)(base + P.off_tnorm); float* qnorm = (float*)(base + P.off_qnorm);
    void* tp = base + P.off_tp; void* qp = base + P.off_qp;
    uint4* cand = (uint4*)(base + P.off_cand);
    int* flist = (int*)(base + P.off_list);
    mpl_key64* part = (mpl_key64*)(base + P.off_part);
    uint32_t* gthr = (uint32_t*)(base + P.off_gthr);

    // per launch, not once per process: the attribute belongs to the CURRENT device's copy of the function (mpl_init may
    // re-target, per-thread contexts call concurrently); the call costs about a microsecond
    MPL_CUDA_CHECK(cudaFuncSetAttribute(l2_tc_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    uint4* taug = (uint4*)(base + P.off_taug); uint4* qaug = (uint4*)(base + P.off_qaug);
    alignas(64) TcMaps maps;
    int rc;
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, F32 = CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    if ((rc = encode_map(&maps.a16, BF, 2, qp, P.kp16, nq, 64, P.rq))) return rc;
    if ((rc = encode_map(&maps.b16, BF, 2, tp, P.kp16, nt, 64, BN))) return rc;
    if ((rc = encode_map(&maps.a32, F32, 4, qp, P.kp32, nq, 32, P.rq))) return rc;
    if ((rc = encode_map(&maps.b32, F32, 4, tp, P.kp32, nt, 32, BN))) return rc;
    if ((rc = encode_map(&maps.aa16, BF, 2, qaug, 16, nq, 16, P.rq, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
    if ((rc = encode_map(&maps.ba16, BF, 2, taug, 16, P.NT * BN, 16, BN, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
    if ((rc = encode_map(&maps.aa32, F32, 4, qaug, 8, nq, 8, P.rq, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
    if ((rc = encode_map(&maps.ba32, F32, 4, taug, 8, P.NT * BN, 8, BN, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;

    MPL_CUDA_CHECK(cudaMemsetAsync(hdr, 0, sizeof(TcHeader), st));
    {
        MplProfScope prof(MPL_PROF_L2_PREP, st);
        const int rows = P.NT * BN + nq;
        int blocks = (rows + 31) / 32; if (blocks > sm_count * 16) blocks = sm_count * 16;      // 4 rows per warp and iteration
        l2_prep_bf16_kernel<<<blocks, 256, 0, st>>>(d_t, nt, t_ld, d_q, nq, q_ld, dim, tp, qp, P.kp16, tnorm, P.NT * BN, qnorm, taug, qaug, hdr, gthr, P.NQB * P.rq);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        l2_prep_tf32_kernel<<<sm_count * 8, 256, 0, st>>>(d_t, nt, t_ld, d_q, nq, q_ld, dim, tp, qp, P.kp32, tnorm, P.NT * BN, taug, qaug, hdr);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    }
    {
        MplProfScope prof(MPL_PROF_L2_TC, st);
        TcParams prm;
        prm.nq = nq; prm.nt = nt; prm.NT = P.NT; prm.U = P.U; prm.max_slots = P.max_slots; prm.rq = P.rq; prm.NQBp = 0; prm.q_pair_rows = 0; prm.t_pair_rows = 0;
        prm.slabs16 = P.kp16 / 64; prm.slabs32 = P.kp32 / 32; prm.slabs8 = 0; prm.cand = cand; prm.gthr = gthr; prm.hdr = hdr;
        l2_tc_rank_kernel<<<P.G, TC_THREADS, SMEM_BYTES, st>>>(maps, prm);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    }
    {
        MplProfScope prof(MPL_PROF_L2_RERANK, st);
        const int blocks = (nq + 7) / 8;
        if (sqrt_out) l2_certify_kernel<1><<<blocks, 256, 0, st>>>(d_q, nq, q_ld, d_t, nt, t_ld, dim, qnorm, cand, P.NT, P.U, P.G, P.max_slots, train_offset, hdr, d_keys_out, flist, P.rq);
        else          l2_certify_kernel<0><<<blocks, 256, 0, st>>>(d_q, nq, q_ld, d_t, nt, t_ld, dim, qnorm, cand, P.NT, P.U, P.G, P.max_slots, train_offset, hdr, d_keys_out, flist, P.rq);
        MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
    }
    return mpl_launch_l2_exact_list(d_q, flist, &hdr->fallback_count, q_ld, d_t, nt, t_ld, dim, sqrt_out ? 1 : 0, train_offset, part,
                                    d_keys_out, sm_count, &hdr->scan_ticket, st);
}

// Number of queries of the last call on this scratch buffer that took the exact-scan fallback (diagnostics / tests).
int mpl_l2tc_last_fallbacks(const void* d_scratch, unsigned* count, unsigned* tf32_mode, cudaStream_t st)
{
    TcHeader h;
    MPL_CUDA_CHECK(cudaMemcpyAsync(&h, d_scratch, sizeof(h), cudaMemcpyDeviceToHost, st));
    MPL_CUDA_CHECK(cudaStreamSynchronize(st));
    if (count) *count = h.fallback_count;
    if (tf32_mode) *tf32_mode = h.not_bf16_exact;
    return MPL_OK;
}

// ---- Hamming on the tensor cores ------------------------------------------------------------------------------------------
// For bit vectors q, t:  hamming(q, t) = |q| + |t| - 2 q.t  (popcounts and a dot product of 0 / 1 values).  With the bits unpacked
// to bf16 (0 / 1 train, 0 / -2 query) the ranking kernel above computes score = |t| - 2 q.t exactly (integers far below 2^24) in
// its exact-integer mode: sorted top-2 lists, ties kept by lowest train index, no certificate and no re-read of any row; the
// certify kernel adds |q| and emits the same integer keys (distance << 32 | train index) as hamming_top2_kernel.  256 bits are
// K = 256 (four bf16 slabs, BM = 256 rows resident), 512 bits K = 512 (eight slabs, 128-row query blocks).
// A train set whose bf16 image exceeds the L2 is walked in chunks of ~48 MB (every chunk's tiles are then read from HBM once and
// re-read by the other query blocks out of L2), each chunk leaving partial keys that one merge launch folds.
bool mpl_hamming_tc_supported(int nq, int nt, int w16) { return nq >= 1 && nt >= 2 && (w16 == 1 || w16 == 2 || w16 == 4); }

static int hamming_tc_chunk_rows(int w16, int fmt) { return std::max(BN, (int)(((size_t)48 << 20) / ((size_t)w16 * (fmt == 2 ? 128 : 256))) / BN * BN); }

static size_t hamming_tc_scratch_fmt(int nq, int nt, int w16, int sm_count, int fmt)
{
    const int chunk = std::min(nt, hamming_tc_chunk_rows(w16, fmt));
    const int n_chunks = (nt + chunk - 1) / chunk;
    const size_t plan = make_plan(nq, chunk, w16 * 128, sm_count, fmt).total;
    return plan + (n_chunks > 1 ? align256((size_t)n_chunks * nq * 2 * sizeof(mpl_key64)) : 0);
}
// sized for either operand format (fmt 1 = bf16, 2 = e4m3)
size_t mpl_hamming_tc_scratch_bytes(int nq, int nt, int w16, int sm_count)
{
    if (!mpl_hamming_tc_supported(nq, nt, w16)) return 16;
    return std::max(hamming_tc_scratch_fmt(nq, nt, w16, sm_count, 1), hamming_tc_scratch_fmt(nq, nt, w16, sm_count, 2));
}

int mpl_hamming_tc_top2(const uint8_t* d_q, int nq, const uint8_t* d_t, int nt, int w16, long long train_offset, void* d_scratch,
                        size_t scratch_bytes, mpl_key64* d_keys_out, int sm_count, cudaStream_t st, int fmt)
{
    if (fmt != 1) fmt = 2;
    if (!mpl_hamming_tc_supported(nq, nt, w16)) { mpl_set_error("hamming tensor-core path: unsupported shape (w16=%d, nt=%d)", w16, nt); return MPL_ERR_BAD_ARG; }
    const size_t need = hamming_tc_scratch_fmt(nq, nt, w16, sm_count, fmt);
    if (!d_scratch || scratch_bytes < need) { mpl_set_error("hamming tensor-core path: scratch %zu < %zu bytes", scratch_bytes, need); return MPL_ERR_BAD_ARG; }
    if (((uintptr_t)d_scratch & 255) != 0) { mpl_set_error("hamming tensor-core path: scratch must be 256-byte aligned"); return MPL_ERR_BAD_ARG; }
    const int dim = w16 * 128;
    const int chunk = std::min(nt, hamming_tc_chunk_rows(w16, fmt));
    const int n_chunks = (nt + chunk - 1) / chunk;
    const size_t plan_bytes = make_plan(nq, chunk, dim, sm_count, fmt).total;
    char* base = (char*)d_scratch;
    mpl_key64* parts = n_chunks > 1 ? (mpl_key64*)(base + plan_bytes) : d_keys_out;
    MPL_CUDA_CHECK(cudaFuncSetAttribute(l2_tc_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, U8 = CU_TENSOR_MAP_DATA_TYPE_UINT8;
    for (int c = 0; c < n_chunks; ++c) {
        const int t0 = c * chunk, ntc = std::min(chunk, nt - t0);
        const TcPlan P = make_plan(nq, ntc, dim, sm_count, fmt);                      // <= the first chunk's plan
        TcHeader* hdr = (TcHeader*)base;
        float* tnorm = (float*)(base + P.off_tnorm); float* qnorm = (float*)(base + P.off_qnorm);
        void* tp = base + P.off_tp; void* qp = base + P.off_qp;
        uint4* cand = (uint4*)(base + P.off_cand); int* flist = (int*)(base + P.off_list);
        uint32_t* gthr = (uint32_t*)(base + P.off_gthr);
        uint4* taug = (uint4*)(base + P.off_taug); uint4* qaug = (uint4*)(base + P.off_qaug);
        alignas(64) TcMaps maps;
        memset(&maps, 0, sizeof(maps));
        int rc;
        if (fmt == 2) {
            if ((rc = encode_map(&maps.a8, U8, 1, qp, dim, nq, 128, P.rq))) return rc;
            if ((rc = encode_map(&maps.b8, U8, 1, tp, dim, ntc, 128, BN))) return rc;
            if ((rc = encode_map(&maps.aa8, U8, 1, qaug, 32, nq, 32, P.rq, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
            if ((rc = encode_map(&maps.ba8, U8, 1, taug, 32, P.NT * BN, 32, BN, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
        } else {
            if ((rc = encode_map(&maps.a16, BF, 2, qp, P.kp16, nq, 64, P.rq))) return rc;
            if ((rc = encode_map(&maps.b16, BF, 2, tp, P.kp16, ntc, 64, BN))) return rc;
            if ((rc = encode_map(&maps.aa16, BF, 2, qaug, 16, nq, 16, P.rq, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
            if ((rc = encode_map(&maps.ba16, BF, 2, taug, 16, P.NT * BN, 16, BN, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
        }
        MPL_CUDA_CHECK(cudaMemsetAsync(hdr, 0, sizeof(TcHeader), st));
        {
            MplProfScope prof(MPL_PROF_L2_PREP, st);
            const int rows = P.NT * BN + nq;
            int blocks = (rows + 31) / 32; if (blocks > sm_count * 16) blocks = sm_count * 16;
            auto prep = fmt == 2 ? hamming_prep_kernel<true> : hamming_prep_kernel<false>;
            prep<<<blocks, 256, 0, st>>>(d_t + (size_t)t0 * 16 * w16, ntc, d_q, nq, w16, tp, qp, tnorm, P.NT * BN, qnorm, taug, qaug,
                                                        hdr, gthr, P.NQB * P.rq);
            MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        }
        {
            MplProfScope prof(MPL_PROF_L2_TC, st);
            TcParams prm;
            prm.nq = nq; prm.nt = ntc; prm.NT = P.NT; prm.U = P.U; prm.max_slots = P.max_slots; prm.rq = P.rq; prm.NQBp = 0; prm.q_pair_rows = 0; prm.t_pair_rows = 0;
            prm.slabs16 = P.kp16 / 64; prm.slabs32 = 0; prm.slabs8 = fmt == 2 ? dim / 128 : 0; prm.cand = cand; prm.gthr = gthr; prm.hdr = hdr;
            l2_tc_rank_kernel<<<P.G, TC_THREADS, SMEM_BYTES, st>>>(maps, prm);
            MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        }
        {
            MplProfScope prof(MPL_PROF_L2_RERANK, st);
            l2_certify_kernel<2><<<(nq + 255) / 256, 256, 0, st>>>(                     // exact-integer branch only: one THREAD per query
                nullptr, nq, 0, nullptr, ntc, 0, dim, qnorm, cand, P.NT, P.U, P.G, P.max_slots,
                                                               train_offset + t0, hdr, parts + (size_t)c * nq * 2, flist, P.rq);
            MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        }
    }
    if (n_chunks > 1) return mpl_launch_merge_top2(parts, n_chunks, nq, 2, nullptr, nullptr, -1, d_keys_out, st);
    return MPL_OK;
}

// ---- a batch of independent pairs in ONE ranking launch (the pair-batch pipeline, C4) --------------------------------------
// Pairs share the geometry (strides nq_stride / nt_stride rows, real sizes in dims[pair] = (n1, n2) like the popc batch kernel):
// work units run over (pair, query block, train tile), operand rows keep the pipeline's pair-strided layout, aug rows / norms /
// thresholds / lists are padded per query block and per train tile, and a train row beyond n2 -- padding or, when the stride is not
// a multiple of the tile, the next pair's first rows -- carries |t|^2 = +inf in its aug row so that it never enters a list.
namespace {

// rows [0, nb * NTp * BN): padded train rows, then nb * NQBp * rq padded query rows; 8 lanes per row
template <bool FP8>
__global__ void __launch_bounds__(256)
hamming_prep_batch_kernel(const uint8_t* __restrict__ t, int nt_stride, const uint8_t* __restrict__ q, int nq_stride, int w16, int nb,
                          const int2* __restrict__ dims, int NTp, int NQBp, int rq, void* __restrict__ tp, void* __restrict__ qp,
                          float* __restrict__ qnorm, uint4* __restrict__ taug, uint4* __restrict__ qaug, TcHeader* hdr,
                          uint32_t* __restrict__ gthr, int n_gthr)
{
    const int lane = threadIdx.x & 31, sub = lane >> 3, sl = lane & 7;
    const int warps_per_grid = gridDim.x * (blockDim.x >> 5);
    const int t_rows = NTp * BN, q_rows = NQBp * rq;
    const long long total_t = (long long)nb * t_rows, total = total_t + (long long)nb * q_rows;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_gthr; i += gridDim.x * blockDim.x) gthr[i] = 0xFFFFFFFFu;
    const int row_bytes = 16 * w16;
    const size_t out_row = (size_t)row_bytes * (FP8 ? 8 : 16);
    unsigned tmax = 0;
    for (long long base = (long long)(blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 4; base < total; base += (long long)warps_per_grid * 4) {
        const long long R = base + sub;
        const bool live = R < total, is_t = R < total_t;
        const long long Rl = is_t ? R : R - total_t;
        const int per = is_t ? t_rows : q_rows, stride = is_t ? nt_stride : nq_stride;
        const int pr = live ? (int)(Rl / per) : 0, r = live ? (int)(Rl - (long long)pr * per) : 0;
        const int n_real = live ? (is_t ? dims[pr].y : dims[pr].x) : 0;
        const size_t row = (size_t)pr * stride + r;                              // operand row in the pair-strided layout
        unsigned pc = 0;
        if (live && r < stride)                                                  // rows beyond the stride belong to the next pair (or to nobody)
            pc = unpack_bits_row<FP8>((is_t ? t : q) + row * row_bytes, row_bytes, sl, is_t, r < n_real, (char*)(is_t ? tp : qp) + row * out_row);
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) pc += __shfl_xor_sync(0xffffffffu, pc, o);
        if (live && sl == 0) {
            if (!is_t) qnorm[Rl] = (float)pc;
            write_aug_bits<FP8>((is_t ? taug : qaug) + 2 * (size_t)Rl, is_t, pc, is_t && r >= n_real);
        }
        if (live && is_t && r < n_real) tmax = max(tmax, pc);
    }
    tmax = __reduce_max_sync(0xffffffffu, tmax);
    if (lane == 0 && tmax) atomicMax(&hdr->max_tnorm_bits, __float_as_uint((float)tmax));
}

// thread per (pair, query): union of the block's exact top-2 lists -> integer Hamming keys, local train indices
__global__ void __launch_bounds__(256)
hamming_certify_batch_kernel(int nb, int nq_stride, const int2* __restrict__ dims, int NQBp, int rq, const float* __restrict__ qnorm,
                             const uint4* __restrict__ cand, int NT, long long U, int G, int max_slots, mpl_key64* __restrict__ keys_out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)nb * nq_stride) return;
    const int pr = (int)(i / nq_stride), q = (int)(i - (long long)pr * nq_stride);
    const int2 d = dims[pr];
    mpl_key64 k0 = MPL_KEY_INVALID, k1 = MPL_KEY_INVALID;
    if (q < d.x) {
        const int qb = pr * NQBp + q / rq, row = q % rq;
        const int f1 = tc_owner((long long)qb * NT, U, G), l1 = tc_owner((long long)qb * NT + NT - 1, U, G);
        const float qq = qnorm[(size_t)qb * rq + row];
        for (int sl = 0; sl < (l1 - f1 + 1) * LPS; ++sl) {
            const uint4* rec = cand + (((size_t)qb * max_slots * LPS + sl) * rq + row) * 2;
            const uint4 sc = rec[0], id = rec[1];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
                const int ti = (int)(e == 0 ? id.x : id.y);
                if (ti < 0 || ti >= d.y) continue;
                const float dd = __fadd_rn(__uint_as_float(e == 0 ? sc.x : sc.y), qq);
                const mpl_key64 key = ((mpl_key64)__float2uint_rn(dd) << 32) | (mpl_key64)(uint32_t)ti;
                const mpl_key64 hi = max(k0, key);
                k0 = min(k0, key); k1 = min(k1, hi);
            }
        }
    }
    reinterpret_cast<ulonglong2*>(keys_out)[i] = make_ulonglong2(k0, k1);
}

struct TcBatchPlan { int rq, NQBp, NTp, NQB, G, max_slots, kp16; long long U; size_t off_qnorm, off_tp, off_qp, off_taug, off_qaug, off_cand, off_gthr, total; };

TcBatchPlan make_batch_plan(int nq_stride, int nt_stride, int w16, int nb, int sm_count, int fmt)
{
    TcBatchPlan P;
    const int dim = w16 * 128;
    P.kp16 = dim;
    P.rq = dim > (fmt == 2 ? 512 : 256) ? UM : BM;
    const size_t elt_row = fmt == 2 ? (size_t)dim : (size_t)dim * 2;
    P.NQBp = (nq_stride + P.rq - 1) / P.rq; P.NTp = (nt_stride + BN - 1) / BN;
    P.NQB = nb * P.NQBp;
    P.U = (long long)P.NQB * P.NTp;
    P.G = (int)(P.U < sm_count ? P.U : sm_count);
    P.max_slots = 1;
    for (int qb = 0; qb < P.NQB; ++qb) {
        const int s = tc_owner((long long)qb * P.NTp + P.NTp - 1, P.U, P.G) - tc_owner((long long)qb * P.NTp, P.U, P.G) + 1;
        if (s > P.max_slots) P.max_slots = s;
    }
    size_t o = 256;
    P.off_qnorm = o; o = align256(o + (size_t)P.NQB * P.rq * 4);
    P.off_tp = o;    o = align256(o + (size_t)nb * nt_stride * elt_row);
    P.off_qp = o;    o = align256(o + (size_t)nb * nq_stride * elt_row);
    P.off_taug = o;  o = align256(o + (size_t)nb * P.NTp * BN * AUG_BYTES);
    P.off_qaug = o;  o = align256(o + (size_t)P.NQB * P.rq * AUG_BYTES);
    P.off_cand = o;  o = align256(o + (size_t)P.NQB * P.max_slots * LPS * P.rq * 32);
    P.off_gthr = o;  o = align256(o + (size_t)P.NQB * P.rq * 4);
    P.total = o;
    return P;
}

// pairs per ranking launch: the train operands of the pairs in flight should stay L2-resident (every query block of a pair
// re-reads the pair's train tiles), and one launch should still hold a few thousand work units
int hamming_tc_group(int nt_stride, int w16, int nb, int fmt)
{
    const size_t per_pair = (size_t)nt_stride * w16 * (fmt == 2 ? 128 : 256);
    int g = (int)std::max<size_t>(1, ((size_t)80 << 20) / std::max<size_t>(per_pair, 1));
    return std::min(std::min(g, 64), nb);
}

}  // namespace

bool mpl_hamming_tc_batch_supported(int nq_stride, int nt_stride, int w16)
{
    return (w16 == 1 || w16 == 2 || w16 == 4) && nq_stride >= 1 && nt_stride >= 2;
}

size_t mpl_hamming_tc_batch_scratch_bytes(int nq_stride, int nt_stride, int w16, int nb, int sm_count)
{
    if (!mpl_hamming_tc_batch_supported(nq_stride, nt_stride, w16) || nb <= 0) return 16;
    return std::max(make_batch_plan(nq_stride, nt_stride, w16, hamming_tc_group(nt_stride, w16, nb, 1), sm_count, 1).total,
                    make_batch_plan(nq_stride, nt_stride, w16, hamming_tc_group(nt_stride, w16, nb, 2), sm_count, 2).total);
}

int mpl_hamming_tc_top2_batch(const uint8_t* d_q, int nq_stride, const uint8_t* d_t, int nt_stride, int w16, int nb, const int2* d_dims,
                              void* d_scratch, size_t scratch_bytes, mpl_key64* d_keys_out, int sm_count, cudaStream_t st, int fmt)
{
    if (nb <= 0) return MPL_OK;
    if (fmt != 1) fmt = 2;
    if (!mpl_hamming_tc_batch_supported(nq_stride, nt_stride, w16)) { mpl_set_error("hamming tensor-core batch: unsupported shape (w16=%d)", w16); return MPL_ERR_BAD_ARG; }
    const int group = hamming_tc_group(nt_stride, w16, nb, fmt);
    const size_t need = make_batch_plan(nq_stride, nt_stride, w16, group, sm_count, fmt).total;
    if (!d_scratch || scratch_bytes < need || ((uintptr_t)d_scratch & 255) != 0) { mpl_set_error("hamming tensor-core batch: scratch %zu < %zu bytes or not 256-byte aligned", scratch_bytes, need); return MPL_ERR_BAD_ARG; }
    MPL_CUDA_CHECK(cudaFuncSetAttribute(l2_tc_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    const CUtensorMapDataType BF = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, U8 = CU_TENSOR_MAP_DATA_TYPE_UINT8;
    char* base = (char*)d_scratch;
    const size_t rowb = (size_t)w16 * 16;
    for (int g0 = 0; g0 < nb; g0 += group) {
        const int gn = std::min(group, nb - g0);
        const TcBatchPlan P = make_batch_plan(nq_stride, nt_stride, w16, gn, sm_count, fmt);
        TcHeader* hdr = (TcHeader*)base;
        float* qnorm = (float*)(base + P.off_qnorm);
        void* tp = base + P.off_tp; void* qp = base + P.off_qp;
        uint4* taug = (uint4*)(base + P.off_taug); uint4* qaug = (uint4*)(base + P.off_qaug);
        uint4* cand = (uint4*)(base + P.off_cand); uint32_t* gthr = (uint32_t*)(base + P.off_gthr);
        alignas(64) TcMaps maps;
        memset(&maps, 0, sizeof(maps));
        int rc;
        if (fmt == 2) {
            if ((rc = encode_map(&maps.a8, U8, 1, qp, P.kp16, gn * nq_stride, 128, P.rq))) return rc;
            if ((rc = encode_map(&maps.b8, U8, 1, tp, P.kp16, gn * nt_stride, 128, BN))) return rc;
            if ((rc = encode_map(&maps.aa8, U8, 1, qaug, 32, P.NQB * P.rq, 32, P.rq, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
            if ((rc = encode_map(&maps.ba8, U8, 1, taug, 32, gn * P.NTp * BN, 32, BN, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
        } else {
            if ((rc = encode_map(&maps.a16, BF, 2, qp, P.kp16, gn * nq_stride, 64, P.rq))) return rc;
            if ((rc = encode_map(&maps.b16, BF, 2, tp, P.kp16, gn * nt_stride, 64, BN))) return rc;
            if ((rc = encode_map(&maps.aa16, BF, 2, qaug, 16, P.NQB * P.rq, 16, P.rq, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
            if ((rc = encode_map(&maps.ba16, BF, 2, taug, 16, gn * P.NTp * BN, 16, BN, CU_TENSOR_MAP_SWIZZLE_32B))) return rc;
        }
        MPL_CUDA_CHECK(cudaMemsetAsync(hdr, 0, sizeof(TcHeader), st));
        {
            MplProfScope prof(MPL_PROF_L2_PREP, st);
            const long long rows = (long long)gn * P.NTp * BN + (long long)P.NQB * P.rq;
            long long blocks = (rows + 31) / 32; if (blocks > sm_count * 16) blocks = sm_count * 16;
            auto prep = fmt == 2 ? hamming_prep_batch_kernel<true> : hamming_prep_batch_kernel<false>;
            prep<<<(unsigned)blocks, 256, 0, st>>>(d_t + (size_t)g0 * nt_stride * rowb, nt_stride, d_q + (size_t)g0 * nq_stride * rowb, nq_stride,
                                                                        w16, gn, d_dims + g0, P.NTp, P.NQBp, P.rq, tp, qp, qnorm, taug, qaug, hdr, gthr, P.NQB * P.rq);
            MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        }
        {
            MplProfScope prof(MPL_PROF_L2_TC, st);
            TcParams prm;
            prm.nq = gn * nq_stride; prm.nt = nt_stride; prm.NT = P.NTp; prm.U = P.U; prm.max_slots = P.max_slots; prm.rq = P.rq;
            prm.NQBp = P.NQBp; prm.q_pair_rows = nq_stride; prm.t_pair_rows = nt_stride;
            prm.slabs16 = P.kp16 / 64; prm.slabs32 = 0; prm.slabs8 = fmt == 2 ? P.kp16 / 128 : 0; prm.cand = cand; prm.gthr = gthr; prm.hdr = hdr;
            l2_tc_rank_kernel<<<P.G, TC_THREADS, SMEM_BYTES, st>>>(maps, prm);
            MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        }
        {
            MplProfScope prof(MPL_PROF_L2_RERANK, st);
            const long long n = (long long)gn * nq_stride;
            hamming_certify_batch_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(gn, nq_stride, d_dims + g0, P.NQBp, P.rq, qnorm, cand, P.NTp, P.U, P.G,
                                                                                      P.max_slots, d_keys_out + (size_t)g0 * nq_stride * 2);
            MPL_CUDA_CHECK(cudaGetLastError()); mpl_count_launch();
        }
    }
    return MPL_OK;
}
