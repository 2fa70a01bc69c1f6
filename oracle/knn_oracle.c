/*
 * knn_oracle.c -- CPU restatement of the reference's brute-force k-NN + ratio test.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (matchinglib_poselib_b200/)
 * may link, import or execute this file; it is the checker used by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 *
 * What is restated (reference = /root/reference/matchinglib_poselib/source, "MP/"):
 *
 *  LINEAR matcher  (MP/matchinglib/source/matchers.cpp:525-714)
 *    The arithmetic lives in OpenCV 4.2.0's bundled FLANN (third party, NOT vendored
 *    in the reference; pinned by ci/make_opencv.sh:6 and find_package(OpenCV 4.2.0)):
 *      flann/linear_index.h  LinearIndex::findNeighbors  -- scan every train row
 *      flann/result_set.h    KNNUniqueResultSet           -- k best by (dist, index)
 *      flann/dist.h          HammingLUT / L2<float>       -- distance functors
 *    Call sites: matchers.cpp:584-588 (CV_8U), :660-664 (CV_32F).
 *    Hamming = sum of popcount over ALL bytes (int).  L2 = SQUARED distance, fp32
 *    accumulator, "result += d0*d0 + d1*d1 + d2*d2 + d3*d3" per group of 4, scalar tail.
 *    Parity pin: this file is checked against python cv2.flann_Index(algorithm=0)
 *    (the same cvflann linear index, OpenCV 4.13) by oracle/gen_golden.py and
 *    tests/test_oracle_knn.py; golden vectors are committed under tests/golden/.
 *
 *  BRUTEFORCENMS matcher (matchers.cpp:476-519 -> MP/matchinglib/include/nmslib/nmslib_matchers.h:159-424)
 *    NMSLIB seq_search (similarity_search/src/method/seqsearch.cc:146-152), K = 2 heap
 *    with strict '<' replacement (include/knnqueue.h:57-66, src/knnquery.cc:69-82),
 *    bit_hamming over packed 16-bit words with the LAST WORD DROPPED
 *    (nmslib_matchers.h:212-231 packing; src/space/space_bit_hamming.cc:33-42 "length - 1"),
 *    l2 = sqrtf of a 4-lane SSE sum (src/distcomp_lp.cc:399-466).
 *    Parity pin: oracle/ref_nmslib/ compiles the reference's own NMSLIB sources into
 *    oracle/_ref/ and tests/test_oracle_knn.py compares both on random + adversarial inputs.
 *
 *  Ratio test + DMatch emission (matchers.cpp:601-625, 677-701; nmslib_matchers.h:399-421)
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (see oracle/Makefile).
 * -ffp-contract=off matters: the reference x86-64 build has no FMA contraction.
 */
#include <stdint.h>
#include <stddef.h>
#include <math.h>
#include <limits.h>
#include <float.h>

typedef struct { int32_t queryIdx, trainIdx, imgIdx; float distance; } orc_dmatch; /* cv::DMatch, 16 B */

/* ---- distance functors -------------------------------------------------- */

/* cvflann::HammingLUT (flann/dist.h): per-byte table popcount, summed into int. */
static inline int hamming_bytes(const uint8_t *a, const uint8_t *b, int nbytes)
{
    int d = 0;
    for (int i = 0; i < nbytes; ++i) d += __builtin_popcount((unsigned)(a[i] ^ b[i]));
    return d;
}

/* nmslib_matchers.h:212-231 packing + space_bit_hamming.cc:33-42 + distcomp.h:222-231.
 * Row of B bytes -> ceil(B/2) ints, int j = (b[2j] << 8) | b[2j+1] (odd tail byte alone);
 * the distance loop runs over n_ints - 1 words, so the last word never contributes. */
static inline int hamming_nmslib(const uint8_t *a, const uint8_t *b, int nbytes)
{
    int n_ints = (nbytes + 1) / 2;
    int d = 0;
    for (int j = 0; j < n_ints - 1; ++j) {
        uint32_t wa = ((uint32_t)a[2 * j] << 8) | a[2 * j + 1];
        uint32_t wb = ((uint32_t)b[2 * j] << 8) | b[2 * j + 1];
        d += __builtin_popcount(wa ^ wb);
    }
    return d;
}

/* cvflann::L2<float>::operator() (flann/dist.h) without worst_dist (LinearIndex passes none).
 * a = train row, b = query row (linear_index.h calls distance_(dataset_[i], vec, cols)). */
static inline float l2sqr_cvflann(const float *a, const float *b, int n)
{
    float result = 0.f;
    int i = 0;
    for (; i + 3 < n; i += 4) {
        float d0 = a[i] - b[i], d1 = a[i + 1] - b[i + 1];
        float d2 = a[i + 2] - b[i + 2], d3 = a[i + 3] - b[i + 3];
        result += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
    }
    for (; i < n; ++i) { float d0 = a[i] - b[i]; result += d0 * d0; }
    return result;
}

/* NMSLIB L2SqrSIMD (distcomp_lp.cc:399-456): four SSE lanes, lane sums added
 * T0+T1+T2+T3, scalar tail, then sqrt (L2NormSIMD :458-462).  pVect1 = query-side
 * object in KNNQuery::CheckAndAddToResult -> DistanceObjLeft(object) ... the order of the
 * subtraction does not change diff*diff. */
static inline float l2_nmslib(const float *a, const float *b, int n)
{
    float lane[4] = {0.f, 0.f, 0.f, 0.f};
    int n4 = (n / 4) * 4, i = 0;
    for (; i < n4; i += 4)
        for (int l = 0; l < 4; ++l) { float d = a[i + l] - b[i + l]; lane[l] += d * d; }
    float res = lane[0] + lane[1] + lane[2] + lane[3];
    for (; i < n; ++i) { float d = a[i] - b[i]; res += d * d; }
    return sqrtf(res);
}

/* ---- k = 1|2 selection by (dist, index) lexicographic order ------------- */
/* KNNUniqueResultSet keeps a std::set<DistIndex> ordered by (dist, index): ties go to the
 * lowest train index.  NMSLIB's K=2 max-heap with strict '<' replacement scanned in index
 * order keeps the same SET of two; its order inside a tie is unobservable under the ratio
 * test (d0 < 0.75*d1 is false for d0 == d1) and is defined as lowest index here. */

#define KNN_LOOP(DTYPE, DISTEXPR, DMAX)                                            \
    for (int q = q0; q < q1; ++q) {                                                \
        DTYPE b0 = DMAX, b1 = DMAX; int i0 = -1, i1 = -1;                          \
        for (int t = 0; t < nt; ++t) {                                             \
            DTYPE d = DISTEXPR;                                                    \
            if (i0 < 0 || d < b0) { b1 = b0; i1 = i0; b0 = d; i0 = t; }            \
            else if (i1 < 0 || d < b1) { b1 = d; i1 = t; }                         \
        }                                                                          \
        idx[(size_t)q * k] = i0; dist[(size_t)q * k] = b0;                         \
        if (k == 2) { idx[(size_t)q * 2 + 1] = i1; dist[(size_t)q * 2 + 1] = b1; } \
    }

/* mode: 0 = LINEAR (all bytes), 1 = BRUTEFORCENMS (packed words, last word dropped).
 * q_step / t_step are row strides in bytes (cv::Mat::step).  Queries [q0,q1) only, so the
 * caller can stripe queries over threads the way nmslib_matchers.h:112-113 does. */
void orc_knn_hamming(const uint8_t *qd, int nq, size_t q_step, const uint8_t *td, int nt, size_t t_step,
                     int nbytes, int mode, int k, int q0, int q1, int32_t *idx, int32_t *dist)
{
    (void)nq;
    if (mode == 0) {
        KNN_LOOP(int32_t, hamming_bytes(td + (size_t)t * t_step, qd + (size_t)q * q_step, nbytes), INT_MAX)
    } else {
        KNN_LOOP(int32_t, hamming_nmslib(qd + (size_t)q * q_step, td + (size_t)t * t_step, nbytes), INT_MAX)
    }
}

/* mode: 0 = LINEAR (squared L2, cvflann order), 1 = BRUTEFORCENMS (sqrt L2, 4-lane order). */
void orc_knn_l2(const float *qd, int nq, size_t q_step, const float *td, int nt, size_t t_step,
                int dim, int mode, int k, int q0, int q1, int32_t *idx, float *dist)
{
    (void)nq;
#define QROW ((const float *)((const char *)qd + (size_t)q * q_step))
#define TROW ((const float *)((const char *)td + (size_t)t * t_step))
    if (mode == 0) {
        KNN_LOOP(float, l2sqr_cvflann(TROW, QROW, dim), FLT_MAX)
    } else {
        KNN_LOOP(float, l2_nmslib(QROW, TROW, dim), FLT_MAX)
    }
}

/* Lowe ratio + DMatch emission.  matchers.cpp:601-625 (int dists: int -> float promotion in
 * "dists[q][0] < (0.75f * dists[q][1])"), :677-701 (float), nmslib_matchers.h:399-415.
 * k == 1 (ratioTest=false with LINEAR): every query is emitted.  Returns the match count. */
int orc_ratio_matches_i32(const int32_t *idx, const int32_t *dist, int nq, int k, int ratio_test,
                          orc_dmatch *out)
{
    int n = 0;
    for (int q = 0; q < nq; ++q) {
        if (ratio_test) {
            if (!((float)dist[(size_t)q * k] < 0.75f * (float)dist[(size_t)q * k + 1])) continue;
        }
        out[n].queryIdx = q; out[n].trainIdx = idx[(size_t)q * k]; out[n].imgIdx = -1;
        out[n].distance = (float)dist[(size_t)q * k];
        ++n;
    }
    return n;
}

int orc_ratio_matches_f32(const int32_t *idx, const float *dist, int nq, int k, int ratio_test,
                          orc_dmatch *out)
{
    int n = 0;
    for (int q = 0; q < nq; ++q) {
        if (ratio_test) {
            if (!(dist[(size_t)q * k] < 0.75f * dist[(size_t)q * k + 1])) continue;
        }
        out[n].queryIdx = q; out[n].trainIdx = idx[(size_t)q * k]; out[n].imgIdx = -1;
        out[n].distance = dist[(size_t)q * k];
        ++n;
    }
    return n;
}
