// knn_l2_exact.cu -- exact fp32 L2 k-NN (k = 2) in the reference's accumulation order, sm_100a.
//
// Two jobs:
//  (1) mpl_launch_l2_exact_top2: full brute-force scan, bit-exact with
//        cvflann::L2<float> (OpenCV 4.2.0 flann/dist.h; call site matchers.cpp:634-664):
//          result += ((d0*d0 + d1*d1) + d2*d2) + d3*d3   per group of 4 dims, then a scalar tail,
//        all fp32, no FMA contraction (order_mode 0, LINEAR, squared distance), or with
//        NMSLIB's 4-lane order (distcomp_lp.cc:399-462: lane l accumulates dims 4g+l, lanes added
//        ((l0+l1)+l2)+l3, scalar tail, sqrtf) for BRUTEFORCENMS (order_mode 1).
//      It serves descriptors the tensor-core kernel (knn_l2_tc.cu) does not take and is the
//      on-device cross-check of that kernel in the parity tests.
//  (2) mpl_launch_l2_rerank: exact distances of a short candidate list per query (the tensor-core
//      kernel only RANKS; every distance that leaves the library is produced here).
//
// Scan layout: CTA = 16 x 16 threads, tile = 64 queries x 64 train rows, 4 x 4 pairs per thread,
// K staged through shared memory 16 dims at a time.  Per-thread top-2 keys
// (float bits << 32 | trainIdx; non-negative floats order like uints) are merged across the 16
// threads that share a query row with warp shuffles, then written per train split.
#include "mpl_common.cuh"

namespace {

constexpr int TQ = 64, TT = 64, KC = 16, PAD = 4;

__device__ __forceinline__ void top2_push(mpl_key64& b0, mpl_key64& b1, mpl_key64 key) {
    const mpl_key64 hi = max(b0, key);
    b0 = min(b0, key);
    b1 = min(b1, hi);
}

// One (query block, train split) tile job.  Query positions [q0, q0 + TQ) of a list of n_pos positions; position
// p reads query row qmap[p] (or p itself when qmap is null) and writes partial[split * n_pos + p].
template <int MODE>
__device__ __forceinline__ void
l2_exact_tile(const float* __restrict__ q, int n_pos, const int* __restrict__ qmap, size_t q_ld, const float* __restrict__ t, int nt,
              size_t t_ld, int dim, int rows_per_split, long long train_offset, int qblock, int split,
              mpl_key64* __restrict__ partial)
{
    __shared__ __align__(16) float sq[TQ][KC + PAD];
    __shared__ __align__(16) float st[TT][KC + PAD];
    const int nq = n_pos;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int q0 = qblock * TQ;
    const int row0 = split * rows_per_split;
    const int row_end = min(nt, row0 + rows_per_split);
    const int dim4 = (dim / 4) * 4;
    constexpr int L = MODE == 1 ? 4 : 1;   // accumulator lanes per pair

    mpl_key64 best0[4], best1[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { best0[i] = MPL_KEY_INVALID; best1[i] = MPL_KEY_INVALID; }

    for (int tb = row0; tb < row_end; tb += TT) {
        float acc[4][4][L];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j)
#pragma unroll
                for (int l = 0; l < L; ++l) acc[i][j][l] = 0.f;
        float res[4][4];

        for (int k0 = 0; k0 < dim; k0 += KC) {
            // stage 64 x 16 floats of each operand (zero filled past the matrix edges)
            {
                const int r = tid >> 2, c = (tid & 3) * 4;
                const int qrow = (q0 + r < nq) ? (qmap ? qmap[q0 + r] : q0 + r) : 0;
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int kk = k0 + c + e;
                    sq[r][c + e] = (q0 + r < nq && kk < dim) ? q[(size_t)qrow * q_ld + kk] : 0.f;
                    st[r][c + e] = (tb + r < row_end && kk < dim) ? t[(size_t)(tb + r) * t_ld + kk] : 0.f;
                }
            }
            __syncthreads();
#pragma unroll
            for (int g = 0; g < KC / 4; ++g) {
                const int kg = k0 + 4 * g;
                if (kg >= dim) break;
                float4 a[4], b[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(&sq[ty * 4 + i][4 * g]);
#pragma unroll
                for (int j = 0; j < 4; ++j) b[j] = *reinterpret_cast<const float4*>(&st[tx + 16 * j][4 * g]);
                if (kg + 4 <= dim4) {
                    // full group of 4
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            // cvflann calls distance_(dataset_row, query): diff = t - q (same square)
                            const float d0 = __fsub_rn(b[j].x, a[i].x), d1 = __fsub_rn(b[j].y, a[i].y);
                            const float d2 = __fsub_rn(b[j].z, a[i].z), d3 = __fsub_rn(b[j].w, a[i].w);
                            if (MODE == 0) {
                                float s = __fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1));
                                s = __fadd_rn(s, __fmul_rn(d2, d2));
                                s = __fadd_rn(s, __fmul_rn(d3, d3));
                                acc[i][j][0] = __fadd_rn(acc[i][j][0], s);
                            } else {
                                acc[i][j][0] = __fadd_rn(acc[i][j][0], __fmul_rn(d0, d0));
                                acc[i][j][1 % L] = __fadd_rn(acc[i][j][1 % L], __fmul_rn(d1, d1));
                                acc[i][j][2 % L] = __fadd_rn(acc[i][j][2 % L], __fmul_rn(d2, d2));
                                acc[i][j][3 % L] = __fadd_rn(acc[i][j][3 % L], __fmul_rn(d3, d3));
                            }
                        }
                } else {
                    // scalar tail (dim % 4 leftover dims): collapse lanes first (MODE 1), then add one by one
                    const int ntail = dim - kg;   // 1..3
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            float r;
                            if (MODE == 1) {
                                r = __fadd_rn(acc[i][j][0], acc[i][j][1 % L]);
                                r = __fadd_rn(r, acc[i][j][2 % L]);
                                r = __fadd_rn(r, acc[i][j][3 % L]);
                            } else r = acc[i][j][0];
                            const float av[3] = {a[i].x, a[i].y, a[i].z};
                            const float bv[3] = {b[j].x, b[j].y, b[j].z};
#pragma unroll
                            for (int e = 0; e < 3; ++e)
                                if (e < ntail) { const float d = __fsub_rn(bv[e], av[e]); r = __fadd_rn(r, __fmul_rn(d, d)); }
                            acc[i][j][0] = r;
                            if (MODE == 1) { acc[i][j][1 % L] = 0.f; acc[i][j][2 % L] = 0.f; acc[i][j][3 % L] = 0.f; }
                        }
                }
            }
            __syncthreads();
        }
        // final value per pair
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float r;
                if (MODE == 1) {
                    // lanes are already collapsed (and zeroed) when a tail was processed; adding +0 is exact
                    r = __fadd_rn(acc[i][j][0], acc[i][j][1 % L]);
                    r = __fadd_rn(r, acc[i][j][2 % L]);
                    r = __fadd_rn(r, acc[i][j][3 % L]);
                    r = sqrtf(r);     // compare in the sqrt domain like KNNQueue does (knnqueue.h:57-66)
                } else r = acc[i][j][0];
                res[i][j] = r;
            }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int tr = tb + tx + 16 * j;
                if (tr < row_end) {
                    const mpl_key64 key = ((mpl_key64)__float_as_uint(res[i][j]) << 32) | (mpl_key64)(uint32_t)(train_offset + tr);
                    top2_push(best0[i], best1[i], key);
                }
            }
    }
    // merge the 16 tx-threads of each query row (lanes 0-15 / 16-31 of a warp) with shuffles
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int off = 8; off >= 1; off >>= 1) {
            const mpl_key64 o0 = __shfl_xor_sync(0xffffffffu, best0[i], off);
            const mpl_key64 o1 = __shfl_xor_sync(0xffffffffu, best1[i], off);
            top2_push(best0[i], best1[i], o0);
            best1[i] = min(best1[i], o1);
        }
        const int qi = q0 + ty * 4 + i;
        if (tx == 0 && qi < nq)
            reinterpret_cast<ulonglong2*>(partial)[(size_t)split * nq + qi] = make_ulonglong2(best0[i], best1[i]);
    }
}

template <int MODE>
__global__ void __launch_bounds__(256)
l2_exact_top2_kernel(const float* __restrict__ q, int nq, size_t q_ld, const float* __restrict__ t, int nt, size_t t_ld,
                     int dim, int rows_per_split, long long train_offset, mpl_key64* __restrict__ partial)
{
    l2_exact_tile<MODE>(q, nq, nullptr, q_ld, t, nt, t_ld, dim, rows_per_split, train_offset, blockIdx.x, blockIdx.y, partial);
}

// List-driven persistent variant (queries the tensor-core path could not certify): *count positions in qlist.
// The split count is chosen on the device so that even one 64-query block spreads over the whole grid.
__device__ __forceinline__ int list_splits(int n, int nt, int grid)
{
    const int nblocks = (n + TQ - 1) / TQ;
    int want = (grid + nblocks - 1) / nblocks;
    const int max_splits = (nt + TT - 1) / TT;
    if (want > max_splits) want = max_splits;
    if (want < 1) want = 1;
    int rps = (nt + want - 1) / want;
    rps = ((rps + TT - 1) / TT) * TT;
    return rps;                                   // rows per split
}

// Exact distance of ONE pair by ONE thread, in the reference's accumulation order (cvflann L2<float>, MODE 0:
// result += ((d0^2 + d1^2) + d2^2) + d3^2 per group of four, scalar tail; NMSLIB L2NormSIMD, MODE 1: four running lanes,
// ((l0 + l1) + l2) + l3, scalar tail, sqrtf).
template <int MODE>
__device__ __forceinline__ float thread_pair_dist(const float* __restrict__ a, const float* __restrict__ b, int dim, bool vec_ok)
{
    const int dim4 = dim & ~3;
    float r = 0.f, l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
#pragma unroll 8                                                                 // eight independent row loads in flight (each lane reads its own row)
    for (int i = 0; i < dim4; i += 4) {
        float4 x, y;
        if (vec_ok) { x = *reinterpret_cast<const float4*>(a + i); y = *reinterpret_cast<const float4*>(b + i); }
        else { x = make_float4(a[i], a[i + 1], a[i + 2], a[i + 3]); y = make_float4(b[i], b[i + 1], b[i + 2], b[i + 3]); }
        const float d0 = __fsub_rn(x.x, y.x), d1 = __fsub_rn(x.y, y.y), d2 = __fsub_rn(x.z, y.z), d3 = __fsub_rn(x.w, y.w);
        const float p0 = __fmul_rn(d0, d0), p1 = __fmul_rn(d1, d1), p2 = __fmul_rn(d2, d2), p3 = __fmul_rn(d3, d3);
        if (MODE == 0) r = __fadd_rn(r, __fadd_rn(__fadd_rn(__fadd_rn(p0, p1), p2), p3));
        else { l0 = __fadd_rn(l0, p0); l1 = __fadd_rn(l1, p1); l2 = __fadd_rn(l2, p2); l3 = __fadd_rn(l3, p3); }
    }
    if (MODE == 1) r = __fadd_rn(__fadd_rn(__fadd_rn(l0, l1), l2), l3);
    for (int i = dim4; i < dim; ++i) { const float d = __fsub_rn(a[i], b[i]); r = __fadd_rn(r, __fmul_rn(d, d)); }
    return MODE == 1 ? sqrtf(r) : r;
}

constexpr int kFewQueries = 16;      // at most this many uncertified queries: one warp per (query, row range) instead of 64-query tiles

template <int MODE>
__global__ void __launch_bounds__(256)
l2_exact_list_kernel(const float* __restrict__ q, const int* __restrict__ qlist, const unsigned* __restrict__ count, size_t q_ld,
                     const float* __restrict__ t, int nt, size_t t_ld, int dim, long long train_offset, mpl_key64* __restrict__ partial,
                     unsigned* __restrict__ ticket, ulonglong2* __restrict__ keys_out)
{
    const int n = (int)*count;
    if (n == 0) return;                                            // the usual case: one empty launch, nothing else
    int splits;
    if (n <= kFewQueries) {
        // The usual non-empty case (a handful of near-duplicate clusters): a 64 x 64 tile kernel would run with 4 live rows of
        // 64 (measured 0.059 ms for 4 queries x 20 000 rows).  Here every warp of the grid takes one (query, row range) item, lane = train row.
        const int warps = gridDim.x * (blockDim.x >> 5);
        splits = max(1, min(warps / n, (nt + 31) / 32));
        const int rps = (((nt + splits - 1) / splits + 31) / 32) * 32;          // whole warps of rows
        splits = (nt + rps - 1) / rps;
        const int lane = threadIdx.x & 31;
        const bool vec_ok = (((uintptr_t)q | (uintptr_t)t) & 15) == 0 && ((q_ld | t_ld) & 3) == 0;
        for (int w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); w < n * splits; w += warps) {
            const int p = w / splits, sp = w - p * splits;
            const float* qrow = q + (size_t)qlist[p] * q_ld;
            const int r0 = sp * rps, r1 = min(nt, r0 + rps);
            mpl_key64 b0 = MPL_KEY_INVALID, b1 = MPL_KEY_INVALID;
            for (int r = r0 + lane; r < r1; r += 32) {                          // lane = train row: the pair's own sequential order
                const float d = thread_pair_dist<MODE>(t + (size_t)r * t_ld, qrow, dim, vec_ok);
                const mpl_key64 key = ((mpl_key64)__float_as_uint(d) << 32) | (mpl_key64)(uint32_t)(train_offset + r);
                top2_push(b0, b1, key);
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                const mpl_key64 o0 = __shfl_xor_sync(0xffffffffu, b0, off), o1 = __shfl_xor_sync(0xffffffffu, b1, off);
                top2_push(b0, b1, o0);
                b1 = min(b1, o1);
            }
            if (lane == 0) reinterpret_cast<ulonglong2*>(partial)[(size_t)sp * n + p] = make_ulonglong2(b0, b1);
        }
    } else {
        const int rps = list_splits(n, nt, gridDim.x);
        splits = (nt + rps - 1) / rps;
        const int items = ((n + TQ - 1) / TQ) * splits;
        for (int w = blockIdx.x; w < items; w += gridDim.x) {
            l2_exact_tile<MODE>(q, n, qlist, q_ld, t, nt, t_ld, dim, rps, train_offset, w / splits, w % splits, partial);
            __syncthreads();
        }
    }
    // the last CTA to finish merges the per-split partial lists (no second launch)
    __shared__ bool is_last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(ticket, 1u) == gridDim.x - 1);
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    const ulonglong2* part2 = reinterpret_cast<const ulonglong2*>(partial);
    if (n <= kFewQueries) {
        // few queries, hundreds of row ranges each: one WARP per query walks the ranges lane-parallel (a single thread per
        // query would chain ~600 L2 round trips), then folds the 32 partial pairs
        const int lane = threadIdx.x & 31;
        for (int p = threadIdx.x >> 5; p < n; p += (int)(blockDim.x >> 5)) {
            mpl_key64 b0 = MPL_KEY_INVALID, b1 = MPL_KEY_INVALID;
            for (int s = lane; s < splits; s += 32) {
                const ulonglong2 v = __ldcg(part2 + (size_t)s * n + p);
                top2_push(b0, b1, v.x);
                b1 = min(b1, v.y);
            }
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) {
                const mpl_key64 o0 = __shfl_xor_sync(0xffffffffu, b0, off), o1 = __shfl_xor_sync(0xffffffffu, b1, off);
                top2_push(b0, b1, o0);
                b1 = min(b1, o1);
            }
            if (lane == 0) keys_out[qlist[p]] = make_ulonglong2(b0, b1);
        }
        return;
    }
    for (int p = threadIdx.x; p < n; p += blockDim.x) {
        mpl_key64 b0 = MPL_KEY_INVALID, b1 = MPL_KEY_INVALID;
        for (int s = 0; s < splits; ++s) {
            const ulonglong2 v = __ldcg(part2 + (size_t)s * n + p);
            const mpl_key64 hi = max(b0, v.x);
            b0 = min(b0, v.x);
            b1 = min(min(b1, hi), v.y);
        }
        keys_out[qlist[p]] = make_ulonglong2(b0, b1);
    }
}

// Exact distances of n_cand candidate train rows per query; output top-2 keys per query.
// One thread per (query, candidate): sequential cvflann/NMSLIB order over the dims.
template <int MODE>
__global__ void l2_rerank_kernel(const float* __restrict__ q, int nq, size_t q_ld, const float* __restrict__ t, size_t t_ld,
                                 int dim, const int32_t* __restrict__ cand, int n_cand, long long train_offset,
                                 mpl_key64* __restrict__ out)
{
    extern __shared__ mpl_key64 skeys[];
    const int qi = blockIdx.x;
    const int c = threadIdx.x;
    mpl_key64 key = MPL_KEY_INVALID;
    if (c < n_cand) {
        const int ti = cand[(size_t)qi * n_cand + c];
        if (ti >= 0) {
            const float* a = t + (size_t)ti * t_ld;
            const float* b = q + (size_t)qi * q_ld;
            float r;
            if (MODE == 0) {
                r = 0.f;
                int i = 0;
                for (; i + 3 < dim; i += 4) {
                    const float d0 = __fsub_rn(a[i], b[i]), d1 = __fsub_rn(a[i + 1], b[i + 1]);
                    const float d2 = __fsub_rn(a[i + 2], b[i + 2]), d3 = __fsub_rn(a[i + 3], b[i + 3]);
                    float s = __fadd_rn(__fmul_rn(d0, d0), __fmul_rn(d1, d1));
                    s = __fadd_rn(s, __fmul_rn(d2, d2));
                    s = __fadd_rn(s, __fmul_rn(d3, d3));
                    r = __fadd_rn(r, s);
                }
                for (; i < dim; ++i) { const float d = __fsub_rn(a[i], b[i]); r = __fadd_rn(r, __fmul_rn(d, d)); }
            } else {
                float l0 = 0.f, l1 = 0.f, l2 = 0.f, l3 = 0.f;
                int i = 0;
                for (; i + 3 < dim; i += 4) {
                    const float d0 = __fsub_rn(a[i], b[i]), d1 = __fsub_rn(a[i + 1], b[i + 1]);
                    const float d2 = __fsub_rn(a[i + 2], b[i + 2]), d3 = __fsub_rn(a[i + 3], b[i + 3]);
                    l0 = __fadd_rn(l0, __fmul_rn(d0, d0)); l1 = __fadd_rn(l1, __fmul_rn(d1, d1));
                    l2 = __fadd_rn(l2, __fmul_rn(d2, d2)); l3 = __fadd_rn(l3, __fmul_rn(d3, d3));
                }
                r = __fadd_rn(__fadd_rn(__fadd_rn(l0, l1), l2), l3);
                for (; i < dim; ++i) { const float d = __fsub_rn(a[i], b[i]); r = __fadd_rn(r, __fmul_rn(d, d)); }
                r = sqrtf(r);
            }
            key = ((mpl_key64)__float_as_uint(r) << 32) | (mpl_key64)(uint32_t)(train_offset + ti);
        }
    }
    skeys[c] = key;
    __syncthreads();
    if (c == 0) {
        mpl_key64 b0 = MPL_KEY_INVALID, b1 = MPL_KEY_INVALID;
        for (int i = 0; i < (int)blockDim.x; ++i) {
            const mpl_key64 kx = skeys[i];
            if (kx == b0 || kx == b1) continue;      // duplicate candidate (same dist & idx)
            top2_push(b0, b1, kx);
        }
        reinterpret_cast<ulonglong2*>(out)[qi] = make_ulonglong2(b0, b1);
    }
}

}  // namespace

int mpl_l2_pick_splits(int nq, int nt, int sm_count)
{
    const int qblocks = (nq + TQ - 1) / TQ;
    int want = (2 * sm_count + qblocks - 1) / qblocks;
    const int max_splits = (nt + TT - 1) / TT;
    if (want > max_splits) want = max_splits;
    if (want < 1) want = 1;
    int rps = (nt + want - 1) / want;
    rps = ((rps + TT - 1) / TT) * TT;
    return (nt + rps - 1) / rps;
}

int mpl_launch_l2_exact_top2(const float* q, int nq, size_t q_ld, const float* t, int nt, size_t t_ld, int dim,
                             int order_mode, int n_splits, long long train_offset, mpl_key64* partial,
                             cudaStream_t st)
{
    int rps = (nt + n_splits - 1) / n_splits;
    rps = ((rps + TT - 1) / TT) * TT;
    if ((nt + rps - 1) / rps != n_splits) { mpl_set_error("l2 exact: internal split mismatch"); return MPL_ERR_BAD_ARG; }
    dim3 grid((nq + TQ - 1) / TQ, n_splits);
    MplProfScope prof(MPL_PROF_L2_EXACT, st);
    if (order_mode == 0)
        l2_exact_top2_kernel<0><<<grid, 256, 0, st>>>(q, nq, q_ld, t, nt, t_ld, dim, rps, train_offset, partial);
    else
        l2_exact_top2_kernel<1><<<grid, 256, 0, st>>>(q, nq, q_ld, t, nt, t_ld, dim, rps, train_offset, partial);
    MPL_CUDA_CHECK(cudaGetLastError());
    mpl_count_launch();
    return MPL_OK;
}

int mpl_launch_l2_rerank(const float* q, int nq, size_t q_ld, const float* t, size_t t_ld, int dim,
                         const int32_t* cand, int n_cand, int order_mode, long long train_offset,
                         mpl_key64* out, cudaStream_t st)
{
    if (nq <= 0) return MPL_OK;
    int threads = 32;
    while (threads < n_cand) threads *= 2;
    if (threads > 1024) { mpl_set_error("l2 rerank: too many candidates (%d)", n_cand); return MPL_ERR_LIMIT; }
    const size_t smem = threads * sizeof(mpl_key64);
    MplProfScope prof(MPL_PROF_L2_RERANK, st);
    if (order_mode == 0)
        l2_rerank_kernel<0><<<nq, threads, smem, st>>>(q, nq, q_ld, t, t_ld, dim, cand, n_cand, train_offset, out);
    else
        l2_rerank_kernel<1><<<nq, threads, smem, st>>>(q, nq, q_ld, t, t_ld, dim, cand, n_cand, train_offset, out);
    MPL_CUDA_CHECK(cudaGetLastError());
    mpl_count_launch();
    return MPL_OK;
}

// Exact scan of the *count queries listed in qlist (device-side count: nothing is launched conditionally on the host).
// partial needs mpl_l2_list_scratch_entries(nq, sm_count) 16-byte entries.
size_t mpl_l2_list_scratch_entries(int nq, int sm_count) { return (size_t)2 * sm_count * TQ + 2 * (size_t)nq + TQ; }

int mpl_launch_l2_exact_list(const float* q, const int* qlist, const unsigned* count, size_t q_ld, const float* t, int nt, size_t t_ld,
                             int dim, int order_mode, long long train_offset, mpl_key64* partial, mpl_key64* keys_out, int sm_count,
                             unsigned* ticket, cudaStream_t st)
{
    const int grid = 2 * sm_count;
    MplProfScope prof(MPL_PROF_L2_EXACT, st);
    if (order_mode == 0) l2_exact_list_kernel<0><<<grid, 256, 0, st>>>(q, qlist, count, q_ld, t, nt, t_ld, dim, train_offset, partial, ticket, (ulonglong2*)keys_out);
    else                 l2_exact_list_kernel<1><<<grid, 256, 0, st>>>(q, qlist, count, q_ld, t, nt, t_ld, dim, train_offset, partial, ticket, (ulonglong2*)keys_out);
    MPL_CUDA_CHECK(cudaGetLastError());
    mpl_count_launch();
    return MPL_OK;
}
