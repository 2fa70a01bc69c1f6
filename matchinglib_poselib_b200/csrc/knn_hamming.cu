// knn_hamming.cu -- brute-force Hamming k-NN (k = 2) for binary descriptors, sm_100a.
//
// Replaces the hot loop of the reference's LINEAR matcher for CV_8U descriptors
//   cvflann::Index<HammingLUT>(LinearIndexParams).knnSearch   (matchers.cpp:567-588, OpenCV 4.2.0 flann)
// and of BRUTEFORCENMS (NMSLIB seq_search + bit_hamming, nmslib_matchers.h:159-424,
// seqsearch.cc:146-152, space_bit_hamming.cc:33-42).  The BRUTEFORCENMS "last 16-bit word is
// ignored" quirk is applied by the row packer (nbytes_used < nbytes), not by the scan kernel.
//
// Data layout in HBM: every descriptor row is w16 x 16 bytes, zero padded (ORB 32 B -> w16 = 2;
// BRISK/FREAK 64 B and AKAZE 61 B -> w16 = 4).  Zero padding adds 0 to every XOR/popcount.
//
// Scan kernel (hamming_top2_kernel):
//   * one thread owns RQ query descriptors in registers (RQ*W16*4 32-bit words);
//   * the CTA's train split is streamed through a 4-stage shared-memory ring, each 8 KB tile
//     fetched by ONE elected thread with a 1-D TMA bulk copy (cp.async.bulk ... mbarrier
//     complete_tx; SASS: UBLKCP) -- 128-bit coalesced by construction;
//   * every thread reads the same train row (LDS.128 broadcast), XORs against its RQ queries,
//     __popc, and keeps a register-resident top-2 as two 32-bit keys (dist << 20 | row-in-split)
//     updated with 3 integer min/max per pair; key order == (dist, trainIdx) lexicographic,
//     i.e. cvflann's KNNUniqueResultSet order (ties -> lowest train index);
//   * grid = (query blocks) x (train splits) sized to a multiple of the SM count; per-split
//     results are 64-bit keys (dist << 32 | global trainIdx) merged by merge_top2_kernel,
//     which is also the merge applied after the NCCL all-gather of the train-sharded path.
//
// Roofline: POPC issue rate (16 lanes/clk/SM), W16*4 POPC per pair; HBM traffic is compulsory
// bytes x (query blocks), irrelevant except for nq of a few rows.
#include "mpl_common.cuh"

namespace {

constexpr int kThreads   = 256;
constexpr int kTileBytes = 8192;      // bytes per TMA stage
constexpr int kStages    = 4;
constexpr int kIdxBits   = 20;        // rows per split <= 2^20 (key = dist << 20 | row)
constexpr unsigned kKeyInvalid = 0xFFFFFFFFu;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
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
// 1-D TMA bulk copy global -> shared, completion signalled on an mbarrier (bytes % 16 == 0).
__device__ __forceinline__ void tma_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// Hamming distance of NW 32-bit words.  POPC issues at a quarter of the LOP3 rate (16 vs 64 lanes/clk/SM, measured:
// profiles/r1_microbench.json), so part of the counting moves to the logic pipe: carry-save adders (Harley-Seal) fold
// three XOR words into a sum word and a carry word (one LOP3 each).  Three adders per 8 words balance the two pipes:
//   8 words -> popc(s3) + popc(x7) + 2 (popc(c1) + popc(c2) + popc(c3)) = 5 POPC + 14 LOP3   (plain: 8 POPC + 8 LOP3)
// (the full tree, 4 POPC + 22 LOP3, is logic-pipe bound and measured slower).
__device__ __forceinline__ void csa(uint32_t& h, uint32_t& l, uint32_t a, uint32_t b, uint32_t c)
{
    // explicit LOP3s: left to itself the compiler folds the operand XORs into the sum and then needs two LOP3 per carry
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(l) : "r"(a), "r"(b), "r"(c));     // a ^ b ^ c
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(h) : "r"(a), "r"(b), "r"(c));     // majority
}
template <int NW>
__device__ __forceinline__ unsigned hamming_words(const uint32_t (&q)[NW], const uint32_t (&t)[NW])
{
    unsigned ones = 0, twos = 0;
    if constexpr (NW % 8 == 0) {
#pragma unroll
        for (int g = 0; g < NW; g += 8) {
            uint32_t x[8];
#pragma unroll
            for (int w = 0; w < 8; ++w) x[w] = q[g + w] ^ t[g + w];
            uint32_t c1, s1, c2, s2, c3, s3;
            csa(c1, s1, x[0], x[1], x[2]);
            csa(c2, s2, x[3], x[4], x[5]);
            csa(c3, s3, s1, s2, x[6]);
            ones += __popc(s3) + __popc(x[7]);
            twos += __popc(c1) + __popc(c2) + __popc(c3);
        }
    } else {
        static_assert(NW == 4, "rows are multiples of 16 bytes");
        uint32_t c, s;
        csa(c, s, q[0] ^ t[0], q[1] ^ t[1], q[2] ^ t[2]);
        ones = __popc(s) + __popc(q[3] ^ t[3]);
        twos = __popc(c);
    }
    return ones + 2 * twos;
}

template <int W16> struct HammingCfg {
    static constexpr int RQ = (W16 <= 2) ? 4 : (W16 <= 4 ? 2 : 1);   // queries per thread
    static constexpr int kTileRows = kTileBytes / (16 * W16);
    static constexpr int kQueriesPerCta = kThreads * RQ;
};

template <int W16>
__global__ void __launch_bounds__(kThreads)
hamming_top2_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ t, int nt,
                    int rows_per_split, long long train_offset, mpl_key64* __restrict__ partial, const int2* __restrict__ dims)
{
    using Cfg = HammingCfg<W16>;
    constexpr int RQ = Cfg::RQ;
    constexpr int TR = Cfg::kTileRows;
    __shared__ __align__(128) uint4 tile[kStages][kTileBytes / 16];
    __shared__ __align__(8) uint64_t full_bar[kStages];

    // blockIdx.z = image pair of a batch: independent problems stored at the common strides nq / nt rows; their own
    // sizes come from dims[pair] = (rows of queries, rows of train) when given, else every pair is nq x nt
    q += (size_t)blockIdx.z * nq * W16;
    t += (size_t)blockIdx.z * nt * W16;
    partial += (size_t)blockIdx.z * gridDim.y * nq * 2;
    const int nq_stride = nq;
    if (dims) {
        const int2 d = dims[blockIdx.z];
        nq = d.x; nt = d.y;
        if ((int)blockIdx.x * Cfg::kQueriesPerCta >= nq) return;
    }
    const int tid   = threadIdx.x;
    const int split = blockIdx.y;
    const int row0  = split * rows_per_split;                        // first train row of this split
    const int rows  = max(0, min(rows_per_split, nt - row0));        // > 0 by construction except for an empty pair
    const int n_tiles = (rows + TR - 1) / TR;

    if (tid == 0) {
        for (int s = 0; s < kStages; ++s) mbar_init(&full_bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    auto issue = [&](int tile_i) {
        const int s = tile_i % kStages;
        const int r = min(TR, rows - tile_i * TR);
        const uint32_t bytes = (uint32_t)r * 16u * W16;
        mbar_expect_tx(&full_bar[s], bytes);
        tma_bulk_g2s(&tile[s][0], t + (size_t)(row0 + tile_i * TR) * W16, bytes, &full_bar[s]);
    };
    if (tid == 0)
        for (int i = 0; i < kStages && i < n_tiles; ++i) issue(i);

    // ---- this thread's RQ query descriptors -> registers ----
    uint32_t qr[RQ][W16 * 4];
    const int qbase = blockIdx.x * Cfg::kQueriesPerCta;
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        const int qi = qbase + j * kThreads + tid;
#pragma unroll
        for (int w = 0; w < W16; ++w) {
            uint4 v = make_uint4(0, 0, 0, 0);
            if (qi < nq) v = __ldg(q + (size_t)qi * W16 + w);
            qr[j][4 * w + 0] = v.x; qr[j][4 * w + 1] = v.y; qr[j][4 * w + 2] = v.z; qr[j][4 * w + 3] = v.w;
        }
    }
    unsigned k1[RQ], k2[RQ];
#pragma unroll
    for (int j = 0; j < RQ; ++j) { k1[j] = kKeyInvalid; k2[j] = kKeyInvalid; }

    for (int ti = 0; ti < n_tiles; ++ti) {
        const int s = ti % kStages;
        mbar_wait(&full_bar[s], (uint32_t)((ti / kStages) & 1));
        const int r_in_tile = min(TR, rows - ti * TR);
        const uint4* srow = &tile[s][0];
        unsigned lrow = (unsigned)(ti * TR);
#pragma unroll 2
        for (int r = 0; r < r_in_tile; ++r, ++lrow) {
            uint32_t tr[W16 * 4];
#pragma unroll
            for (int w = 0; w < W16; ++w) {
                const uint4 v = srow[r * W16 + w];                     // LDS.128, warp-wide broadcast
                tr[4 * w + 0] = v.x; tr[4 * w + 1] = v.y; tr[4 * w + 2] = v.z; tr[4 * w + 3] = v.w;
            }
#pragma unroll
            for (int j = 0; j < RQ; ++j) {
                const unsigned d = hamming_words<W16 * 4>(qr[j], tr);
                const unsigned key = (d << kIdxBits) | lrow;
                const unsigned hi  = max(k1[j], key);
                k1[j] = min(k1[j], key);
                k2[j] = min(k2[j], hi);
            }
        }
        __syncthreads();                                               // every warp is done with stage s
        if (tid == 0 && ti + kStages < n_tiles) issue(ti + kStages);
    }

    // ---- per-split result: 64-bit keys with GLOBAL train indices ----
#pragma unroll
    for (int j = 0; j < RQ; ++j) {
        const int qi = qbase + j * kThreads + tid;
        if (qi >= nq) continue;
        const unsigned kk[2] = {k1[j], k2[j]};
        ulonglong2 out;
        mpl_key64 o[2];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            if (kk[c] == kKeyInvalid) o[c] = MPL_KEY_INVALID;
            else {
                const unsigned d = kk[c] >> kIdxBits;
                const long long gi = train_offset + row0 + (long long)(kk[c] & ((1u << kIdxBits) - 1));
                o[c] = ((mpl_key64)d << 32) | (mpl_key64)(uint32_t)gi;
            }
        }
        out.x = o[0]; out.y = o[1];
        reinterpret_cast<ulonglong2*>(partial)[(size_t)split * nq_stride + qi] = out;
    }
}

// Pack rows of `nbytes_used` payload bytes (src stride src_step) into w16*16-byte zero-padded rows.
__global__ void pack_rows_u8_kernel(const uint8_t* __restrict__ src, int n, size_t src_step, int nbytes_used,
                                    uint8_t* __restrict__ dst, int row_bytes)
{
    const size_t total = (size_t)n * row_bytes;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const size_t r = i / row_bytes;
        const int c = (int)(i - r * row_bytes);
        dst[i] = c < nbytes_used ? src[r * src_step + c] : (uint8_t)0;
    }
}

// Top-2 of n_parts candidate pairs per query.  One thread per query, coalesced 16-byte loads.
// Used (a) across the train splits of one GPU, (b) across ranks after the NCCL all-gather.
__global__ void merge_top2_kernel(const ulonglong2* __restrict__ partial, int n_parts, int nq, int k,
                                  int32_t* __restrict__ idx, uint32_t* __restrict__ dist, int float_sqrt,
                                  ulonglong2* __restrict__ merged)
{
    const int qi = blockIdx.x * blockDim.x + threadIdx.x;
    if (qi >= nq) return;
    mpl_key64 b0 = MPL_KEY_INVALID, b1 = MPL_KEY_INVALID;
    for (int p = 0; p < n_parts; ++p) {
        const ulonglong2 v = partial[(size_t)p * nq + qi];
        // v.x <= v.y inside one part
        mpl_key64 hi = max(b0, v.x);
        b0 = min(b0, v.x);
        b1 = min(b1, hi);
        b1 = min(b1, v.y);
    }
    if (merged) merged[qi] = make_ulonglong2(b0, b1);
    if (idx) {
        const mpl_key64 kk[2] = {b0, b1};
        for (int c = 0; c < k; ++c) {
            int32_t id; uint32_t dv;
            if (kk[c] == MPL_KEY_INVALID) { id = -1; dv = float_sqrt >= 0 ? 0x7F7FFFFFu /*FLT_MAX*/ : 0x7FFFFFFFu /*INT_MAX*/; }
            else {
                id = (int32_t)(uint32_t)(kk[c] & 0xFFFFFFFFull);
                dv = (uint32_t)(kk[c] >> 32);
                if (float_sqrt == 1) dv = __float_as_uint(sqrtf(__uint_as_float(dv)));
            }
            idx[(size_t)qi * k + c] = id;
            dist[(size_t)qi * k + c] = dv;
        }
    }
}

// Lowe ratio test + cv::DMatch emission on merged keys (matchers.cpp:601-625; nmslib_matchers.h:399-421), order-preserving:
// one CTA, every warp owns a contiguous chunk of queries (ballot counts, block scan, write).  out = 16-byte header (count) followed
// by the matches in ascending queryIdx, so one D2H copy brings back both.  ratio_test == 0: every query emits its nearest.
// An absent second neighbour compares as FLT_MAX / INT_MAX, like the idx/dist arrays of merge_top2_kernel would.
__global__ void __launch_bounds__(1024)
ratio_dmatch_kernel(const ulonglong2* __restrict__ keys, int nq, int ratio_test, int is_float, int4* __restrict__ out)
{
    __shared__ int warp_sums[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    // warp w owns the contiguous query range [w * per, (w + 1) * per), per a multiple of 32: lanes read consecutive keys
    // (coalesced 512 B), ballots give the order inside the warp
    const int per = ((nq + 1023) / 1024) * 32;
    const int start = min(nq, wid * per), end = min(nq, start + per);
    auto dist_f = [&](mpl_key64 k) {
        if (k == MPL_KEY_INVALID) return is_float ? 3.402823466e+38f : (float)0x7FFFFFFF;
        const uint32_t d = (uint32_t)(k >> 32);
        return is_float ? __uint_as_float(d) : (float)(int32_t)d;
    };
    auto passes = [&](const ulonglong2 k) { return !ratio_test || dist_f(k.x) < 0.75f * dist_f(k.y); };
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
            const int ti = k.x == MPL_KEY_INVALID ? -1 : (int)(uint32_t)(k.x & 0xFFFFFFFFull);
            // queryIdx, trainIdx, imgIdx = -1, distance
            out[1 + pos + __popc(b & ((1u << lane) - 1u))] = make_int4(i + lane, ti, -1, __float_as_int(dist_f(k.x)));
        }
        pos += __popc(b);
    }
    if (tid == 1023) out[0] = make_int4(pos, 0, 0, 0);
}

// All-gather of every rank's per-query top-2 keys over NVLink peer memory, fused with the merge (the database regime,
// DESIGN.md section 6): each rank stores its local keys into slot [parity][rank] of EVERY rank's gather buffer (plain
// 16-byte stores through the peer mappings), publishes with a system-scope fence + one atomic per CTA on every rank's
// arrival counter, waits until all world x gridDim.x arrivals of this step have reached its own counter, then merges the
// world parts out of its own buffer.  One kernel instead of an NCCL all-gather (latency-bound at 16 B x Q per rank) plus a
// merge launch.  Steps alternate between two buffers: a rank can run at most one step ahead of a peer (it cannot pass the
// wait of step e+1 before that peer has finished merging step e and arrived in e+1), so its early stores never land in a
// buffer a peer is still reading.  sync words per rank: [0] step counter, [1] arrivals (monotone), [2] CTA ticket, [3] time-out flag.
// Contract: every rank launches the same nq (hence the same grid: min(ceil(nq / 256), 64) CTAs, all co-resident -- 64 CTAs
// of 256 threads always fit on the 148 SMs).
struct MplPeerSet { ulonglong2* gather[8]; unsigned* sync[8]; };
constexpr unsigned long long kPeerTimeoutNs = 4000000000ull;     // 4 s: far above any rank start skew inside one step

__global__ void __launch_bounds__(256)
allgather_merge_kernel(const ulonglong2* __restrict__ local_keys, int nq, int rank, int world, MplPeerSet P, int k,
                       int32_t* __restrict__ idx, uint32_t* __restrict__ dist, int float_sqrt, ulonglong2* __restrict__ merged)
{
    unsigned* my = P.sync[rank];
    __shared__ unsigned s_epoch;
    if (threadIdx.x == 0) s_epoch = *reinterpret_cast<volatile unsigned*>(&my[0]);
    __syncthreads();
    const unsigned e = s_epoch, par = e & 1u;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
    for (int q = gtid; q < nq; q += gsize) {
        const ulonglong2 v = local_keys[q];
        for (int p = 0; p < world; ++p) P.gather[p][((size_t)par * world + rank) * nq + q] = v;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int p = 0; p < world; ++p) atomicAdd_system(&P.sync[p][1], 1u);
        const unsigned target = (e + 1u) * (unsigned)world * gridDim.x;
        // bounded wait: a peer that never launches (failed rank, different nq / grid) must not hang this GPU.  After
        // kPeerTimeoutNs the CTA records the failure in sync word [3] (read by mpl_peer_status) and goes on; the merged
        // result of that step is then undefined and the host path falls back to NCCL (sharded.py) or fails.
        unsigned long long t0, t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        while ((int)(*reinterpret_cast<volatile unsigned*>(&my[1]) - target) < 0) {
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > kPeerTimeoutNs) { atomicExch(&my[3], 1u); break; }
            __nanosleep(64);
        }
        __threadfence_system();
    }
    __syncthreads();
    const ulonglong2* mine = P.gather[rank] + (size_t)par * world * nq;
    for (int q = gtid; q < nq; q += gsize) {
        mpl_key64 b0 = MPL_KEY_INVALID, b1 = MPL_KEY_INVALID;
        for (int p = 0; p < world; ++p) {
            const ulonglong2 v = __ldcg(mine + (size_t)p * nq + q);                 // written by peer p: not through L1
            const mpl_key64 hi = max(b0, v.x);
            b0 = min(b0, v.x);
            b1 = min(min(b1, hi), v.y);
        }
        if (merged) merged[q] = make_ulonglong2(b0, b1);
        if (idx) {
            const mpl_key64 kk[2] = {b0, b1};
            for (int c = 0; c < k; ++c) {
                int32_t id; uint32_t dv;
                if (kk[c] == MPL_KEY_INVALID) { id = -1; dv = float_sqrt >= 0 ? 0x7F7FFFFFu : 0x7FFFFFFFu; }
                else {
                    id = (int32_t)(uint32_t)(kk[c] & 0xFFFFFFFFull);
                    dv = (uint32_t)(kk[c] >> 32);
                    if (float_sqrt == 1) dv = __float_as_uint(sqrtf(__uint_as_float(dv)));
                }
                idx[(size_t)q * k + c] = id;
                dist[(size_t)q * k + c] = dv;
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {                                                         // the last CTA closes the step
        __threadfence();
        if (atomicAdd(&my[2], 1u) == gridDim.x - 1) { my[2] = 0; __threadfence(); my[0] = e + 1u; }
    }
}

template <int W16>
int launch_w(const uint8_t* q, int nq, const uint8_t* t, int nt, int n_splits, int n_pairs, long long train_offset,
             mpl_key64* partial, const int2* dims, cudaStream_t st)
{
    using Cfg = HammingCfg<W16>;
    const int rows_per_split = (nt + n_splits - 1) / n_splits;    // the last TMA tile of a split may be partial
    if (rows_per_split > (1 << kIdxBits)) { mpl_set_error("hamming: split of %d rows exceeds 2^20", rows_per_split); return MPL_ERR_LIMIT; }
    const int eff_splits = (nt + rows_per_split - 1) / rows_per_split;
    if (eff_splits != n_splits) { mpl_set_error("hamming: internal split mismatch %d vs %d", eff_splits, n_splits); return MPL_ERR_BAD_ARG; }
    dim3 grid((nq + Cfg::kQueriesPerCta - 1) / Cfg::kQueriesPerCta, n_splits, n_pairs);
    MplProfScope prof(MPL_PROF_HAMMING, st);
    hamming_top2_kernel<W16><<<grid, kThreads, 0, st>>>((const uint4*)q, nq, (const uint4*)t, nt, rows_per_split,
                                                         train_offset, partial, dims);
    MPL_CUDA_CHECK(cudaGetLastError());
    mpl_count_launch();
    return MPL_OK;
}

int queries_per_cta(int w16) { return w16 <= 2 ? kThreads * 4 : (w16 <= 4 ? kThreads * 2 : kThreads); }

}  // namespace

// Number of train splits.  CTAs that share an SM share its integer pipes, so the scan time is the largest number of train
// rows any SM is handed: ceil(q-blocks x splits / SMs) x (rows per split + a per-CTA prologue worth ~24 rows).  Pick the
// split count that minimises it with at least two CTAs per SM when the problem allows (latency hiding); C1: 8 q-blocks x
// 37 splits = 2 x 148 CTAs of 222 rows.  Every split holds <= 2^20 rows (index bits of the in-kernel key).
int mpl_hamming_pick_splits(int nq, int nt, int w16, int sm_count)
{
    const int qblocks = (nq + queries_per_cta(w16) - 1) / queries_per_cta(w16);
    const int min_splits = (nt + (1 << kIdxBits) - 1) / (1 << kIdxBits);
    int max_splits = nt / 32;                                   // >= 32 rows per split
    if (max_splits < min_splits) max_splits = min_splits;
    if (max_splits < 1) max_splits = 1;
    const int cap = (8 * sm_count + qblocks - 1) / qblocks;       // no point beyond 8 CTAs per SM
    if (max_splits > cap && cap >= min_splits) max_splits = cap;
    long long best_cost = -1; int best = min_splits > 0 ? min_splits : 1;
    for (int sp = best; sp <= max_splits; ++sp) {
        const int rps = (nt + sp - 1) / sp;
        if ((nt + rps - 1) / rps != sp) continue;                 // not a reachable split count
        const long long ctas = (long long)qblocks * sp;
        const long long per_sm = (ctas + sm_count - 1) / sm_count;
        long long cost = per_sm * (rps + 24);
        if (per_sm < 2 && sp < max_splits) cost += cost / 2;      // a lone CTA per SM cannot hide its own latencies
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = sp; }
    }
    return best;
}

int mpl_launch_pack_rows_u8(const uint8_t* src, int n, size_t src_step, int nbytes_used, uint8_t* dst, int w16,
                            cudaStream_t st)
{
    if (n <= 0) return MPL_OK;
    const size_t total = (size_t)n * w16 * 16;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 16) blocks = 148 * 16;
    pack_rows_u8_kernel<<<blocks, 256, 0, st>>>(src, n, src_step, nbytes_used, dst, w16 * 16);
    MPL_CUDA_CHECK(cudaGetLastError());
    mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_hamming_top2(const uint8_t* q, int nq, const uint8_t* t, int nt, int w16, int n_splits, int n_pairs,
                            long long train_offset, mpl_key64* partial, cudaStream_t st, const void* pair_dims)
{
    const int2* dims = (const int2*)pair_dims;
    if (nq <= 0 || nt <= 0 || n_pairs <= 0) return MPL_OK;
    if (n_pairs > 65535) { mpl_set_error("hamming: at most 65535 pairs per batch launch"); return MPL_ERR_LIMIT; }
    switch (w16) {
        case 1:  return launch_w<1>(q, nq, t, nt, n_splits, n_pairs, train_offset, partial, dims, st);
        case 2:  return launch_w<2>(q, nq, t, nt, n_splits, n_pairs, train_offset, partial, dims, st);
        case 4:  return launch_w<4>(q, nq, t, nt, n_splits, n_pairs, train_offset, partial, dims, st);
        case 8:  return launch_w<8>(q, nq, t, nt, n_splits, n_pairs, train_offset, partial, dims, st);
        case 16: return launch_w<16>(q, nq, t, nt, n_splits, n_pairs, train_offset, partial, dims, st);
        default: mpl_set_error("hamming: unsupported row width w16=%d (use 1,2,4,8,16)", w16); return MPL_ERR_BAD_ARG;
    }
}

int mpl_launch_merge_top2(const mpl_key64* partial, int n_parts, int nq, int k, int32_t* idx, void* dist,
                          int float_sqrt, mpl_key64* merged_keys, cudaStream_t st)
{
    if (nq <= 0) return MPL_OK;
    MplProfScope prof(MPL_PROF_MERGE, st);
    merge_top2_kernel<<<(nq + 127) / 128, 128, 0, st>>>((const ulonglong2*)partial, n_parts, nq, k, idx,
                                                        (uint32_t*)dist, float_sqrt, (ulonglong2*)merged_keys);
    MPL_CUDA_CHECK(cudaGetLastError());
    mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_ratio_dmatch(const mpl_key64* keys, int nq, int ratio_test, int is_float, void* out, cudaStream_t st)
{
    MplProfScope prof(MPL_PROF_MERGE, st);
    ratio_dmatch_kernel<<<1, 1024, 0, st>>>((const ulonglong2*)keys, nq, ratio_test, is_float, (int4*)out);
    MPL_CUDA_CHECK(cudaGetLastError());
    mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_allgather_merge(const mpl_key64* local_keys, int nq, int rank, int world, void* const* peer_gather, void* const* peer_sync,
                               int k, int32_t* idx, void* dist, int float_sqrt, mpl_key64* merged, cudaStream_t st)
{
    if (nq <= 0) return MPL_OK;
    if (world < 1 || world > 8 || rank < 0 || rank >= world) { mpl_set_error("allgather_merge: world %d / rank %d out of range (1..8)", world, rank); return MPL_ERR_BAD_ARG; }
    MplPeerSet P;
    for (int p = 0; p < 8; ++p) { P.gather[p] = p < world ? (ulonglong2*)peer_gather[p] : nullptr; P.sync[p] = p < world ? (unsigned*)peer_sync[p] : nullptr; }
    int grid = (nq + 255) / 256; if (grid > 64) grid = 64;                          // co-resident: the CTAs wait on each other's peers
    MplProfScope prof(MPL_PROF_MERGE, st);
    allgather_merge_kernel<<<grid, 256, 0, st>>>((const ulonglong2*)local_keys, nq, rank, world, P, k, idx, (uint32_t*)dist, float_sqrt, (ulonglong2*)merged);
    MPL_CUDA_CHECK(cudaGetLastError());
    mpl_count_launch();
    return MPL_OK;
}
