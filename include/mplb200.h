/*
 * mplb200.h -- C ABI of libmplb200.so: the B200 (sm_100a) implementation of the data-parallel
 * hot path of josefmaierfl/matchinglib_poselib (brute-force descriptor k-NN + Lowe ratio test;
 * RANSAC 5-point essential matrix hypothesis generation + Sampson scoring).
 *
 * Plain pointers and sizes only: this is what the reference's C++ API functions
 * (matchinglib::getMatches, poselib::estimateEssentialMat, poselib::getPoseTriangPts) bind to in
 * the drop-in shim include/matchinglib_poselib_b200.hpp, and what the Python host mirror
 * (matchinglib_poselib_b200/*.py) calls through ctypes.  Reference paths below are relative to
 * /root/reference/matchinglib_poselib/source ("MP/").
 *
 * Conventions
 *   - every function returns 0 (MPL_OK) or a negative status; mpl_last_error() gives the text.
 *     The *_api functions return the REFERENCE's own codes (documented per function).
 *   - "host" entry points take host pointers, copy to the device, run the kernels on the
 *     library's stream, copy back and synchronise: they are synchronous like the reference API.
 *   - "*_dev" entry points take DEVICE pointers and a cudaStream_t (passed as void*), enqueue
 *     work and return without synchronising; scratch memory comes from the caller.
 *   - There is no CPU fallback anywhere: without a usable CUDA device every call fails with
 *     MPL_ERR_NO_DEVICE / MPL_ERR_CUDA.
 */
#ifndef MPLB200_H_
#define MPLB200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPL_OK              0
#define MPL_ERR_BAD_ARG    -1
#define MPL_ERR_CUDA      -100
#define MPL_ERR_NO_DEVICE -101
#define MPL_ERR_LIMIT     -102
#define MPL_ERR_ASSERT    -103   /* a CV_Assert of the reference would have thrown */

/* cv::DMatch memory layout (16 bytes): queryIdx, trainIdx, imgIdx, distance. */
typedef struct mpl_dmatch { int32_t queryIdx, trainIdx, imgIdx; float distance; } mpl_dmatch;

/* ---- library / device --------------------------------------------------------------------- */
/* Contexts are per HOST THREAD (own CUDA stream, scratch, pinned staging), created on first use on the device the last
 * mpl_init() named (default 0) and freed at thread exit or mpl_shutdown() (calling thread's only).  Entry points are
 * synchronous for their caller; calls from different threads overlap on the GPU -- the way to run a batch of image pairs. */
int         mpl_init(int device);            /* select device, create the caller's stream + scratch; idempotent */
int         mpl_shutdown(void);
const char* mpl_last_error(void);            /* thread-local text of the last failure */
int         mpl_device_info(int* sm_count, int* cc_major, int* cc_minor, size_t* total_mem);
long long   mpl_kernel_launch_count(void);   /* kernels launched by this library since mpl_init */
/* Pinned host memory for descriptor matrices and keypoints (what cv::Mat::data should point at for the fastest calls):
 * anonymous 2 MiB-aligned pages with MADV_HUGEPAGE, registered with cudaHostRegister.  Huge pages matter: with 4 KiB pages
 * (cudaHostAlloc in an already fragmented process) the same host-to-device copy was measured at 25-40 GB/s, with 2 MiB
 * pages at the PCIe link rate (52 GB/s).  Any host pointer is accepted by the entry points; this is the fast one. */
void*       mpl_host_alloc(size_t bytes);    /* NULL on failure (mpl_last_error) */
int         mpl_host_free(void* p);
/* Device timers per kernel class (CUDA events on the launching stream), off by default.  Class ids:
 * 0 Hamming scan, 1 exact L2 scan, 2 tensor-core L2 ranking, 3 L2 re-rank/certify, 4 top-2 merge,
 * 5 five-point solve, 6 Sampson scoring, 7 L2 operand preparation.  mpl_prof_read waits for the recorded
 * launches of one class, returns their summed duration and count, and resets that class. */
int         mpl_prof_enable(int on);
int         mpl_prof_read(int kernel_class, double* total_ms, long long* launches);

/* ---- brute-force k-NN, host pointers -------------------------------------------------------
 * mpl_knn2_hamming replaces cvflann::Index<HammingLUT>(LinearIndexParams).knnSearch
 *   (MP/matchinglib/source/matchers.cpp:567-588; OpenCV 4.2.0 flann linear_index.h/result_set.h)
 *   with ignore_tail_bytes = 0, and the NMSLIB seq_search/bit_hamming scan of
 *   nmslibMatching<int> (MP/matchinglib/include/nmslib/nmslib_matchers.h:159-424,
 *   similarity_search/src/space/space_bit_hamming.cc:33-42) with ignore_tail_bytes = 2 (even
 *   nbytes) or 1 (odd nbytes) -- the reference's "last packed word is dropped" behaviour.
 *   q/t: row-major uint8, row strides q_step/t_step bytes (cv::Mat::step). k = 1 or 2.
 *   idx/dist: [nq*k] int32; neighbours ordered by (dist, trainIdx); missing neighbour = (-1, INT_MAX).
 * mpl_knn2_l2f32 replaces cvflann::Index<L2<float>>(LinearIndexParams).knnSearch
 *   (matchers.cpp:634-664; SQUARED distances, cvflann fp32 accumulation order; sqrt_out = 0) and
 *   nmslibMatching<float> seq_search/l2 (sqrt distances, NMSLIB 4-lane order; sqrt_out = 1).
 *   Every returned distance is bit-exact fp32 in the stated order: the tensor-core kernel only ranks
 *   candidates, the exact re-rank certifies each query and uncertified ones are rescanned exactly.
 */
int mpl_knn2_hamming(const uint8_t* q, int nq, size_t q_step, const uint8_t* t, int nt, size_t t_step,
                     int nbytes, int ignore_tail_bytes, int k, int32_t* idx, int32_t* dist);
int mpl_knn2_l2f32(const float* q, int nq, size_t q_step, const float* t, int nt, size_t t_step,
                   int dim, int sqrt_out, int k, int32_t* idx, float* dist);
/* force a float path: 0 = auto, 1 = exact SIMT scan only, 2 = tensor-core ranking + exact re-rank */
int mpl_set_l2_path(int path);
/* Hamming k-NN path of mpl_knn2_hamming / mpl_knn2_hamming_dev / getMatches / mpl_match_pose_batch on CV_8U descriptors: 0 = auto
 * (tensor cores for large scans), 1 = popc scan (hamming_top2_kernel), 2 = tensor cores: bits unpacked to e4m3 bytes (0 / 1 train,
 * 0 / -2 query), hamming = |q| + |t| - 2 q.t on the float ranking kernel in its exact-integer mode (tcgen05 kind::f8f6f4; rows of
 * 16 / 32 / 64 bytes only), 3 = the same with bf16 operands (kind::f16; kept for comparison).  All give the same keys bit for bit;
 * process-wide. */
int mpl_set_hamming_path(int path);

/* Lowe ratio test + cv::DMatch emission (matchers.cpp:601-625, 677-701; nmslib_matchers.h:399-415):
 * keep q iff d0 < 0.75f*d1 (ratio_test != 0; needs k == 2), emit every query otherwise.
 * out must hold nq entries; *n_out receives the count; matches are in ascending queryIdx. */
int mpl_ratio_matches_i32(const int32_t* idx, const int32_t* dist, int nq, int k, int ratio_test,
                          mpl_dmatch* out, int* n_out);
int mpl_ratio_matches_f32(const int32_t* idx, const float* dist, int nq, int k, int ratio_test,
                          mpl_dmatch* out, int* n_out);

/* matchinglib::getMatches for matcher_name in {"LINEAR","BRUTEFORCENMS"} over raw cv::Mat fields
 * (MP/matchinglib/source/matchers.cpp:115-135, 476-519, 525-714).  cv_type: 0 = CV_8U, 5 = CV_32F.
 * Returns the REFERENCE codes: 0 ok, -1 wrong input, -2 matcher not supported (every other
 * matcher_name: no CPU fallback), -3 fewer than 2 matches (LINEAR only), -4 fewer than 15
 * keypoints on either side; device failures map to -3 with mpl_last_error() set.
 * out must hold rows1 entries. */
int mpl_get_matches_api(int n_keypoints1, int n_keypoints2,
                        const void* desc1, int rows1, int cols1, size_t step1, int cv_type1,
                        const void* desc2, int rows2, int cols2, size_t step2, int cv_type2,
                        const char* matcher_name, int ratio_test, int nr_threads /*ignored*/,
                        mpl_dmatch* out, int* n_out);

/* ---- GMS post-filter ---------------------------------------------------------------
 * mpl_gms_filter replaces filterMatchesGMS(keypoints1, imageSize1, keypoints2, imageSize2, matches, inlierMask, useScale,
 *   useRotation) (MP/matchinglib/source/gms.cpp:54-60 -> MatchGMS, MP/matchinglib/thirdparty/gms-1.0/src/MatchGMS.cpp;
 *   called by getCorrespondences with GMSrefine, correspondences.cpp:386, useScale = useRotation = false).
 *   kp1 / kp2: arrays of records whose first 8 bytes are the pixel position (float x, float y) -- pass &keypoints[0] with
 *   kp_stride = sizeof(cv::KeyPoint) = 28, or packed float pairs with kp_stride = 8.  inlier_mask[n_matches] 0/1,
 *   *n_inliers = the reference's return value.  Bit-exact against the reference's grid arithmetic. */
int mpl_gms_filter(const void* kp1, int n1, size_t kp_stride1, int width1, int height1, const void* kp2, int n2, size_t kp_stride2,
                   int width2, int height2, const mpl_dmatch* matches, int n_matches, int use_scale, int use_rotation,
                   uint8_t* inlier_mask, int* n_inliers);

/* ---- sharded database regime: all-gather over NVLink peer memory fused with the merge -----------------
 * Every rank (one process per GPU) allocates one gather buffer of mpl_peer_gather_bytes(nq, world) and one 256-byte sync
 * block with mpl_peer_alloc (zeroed, returns a 64-byte CUDA IPC handle), exchanges the handles (e.g. torch.distributed
 * all_gather_object) and maps the other ranks' allocations with mpl_peer_open.  mpl_allgather_merge_top2_dev then
 * replaces "NCCL all-gather of every rank's keys_out + mpl_merge_top2_dev": one kernel stores the local keys into every
 * rank's buffer, signals, waits for the world's arrivals and merges (same (dist, idx) key order: the result equals the
 * one-GPU scan).  peer_gather[r] / peer_sync[r]: rank r's allocations as seen from this process (own ones at [rank]).
 * Every rank must call it the same number of times with the same nq; world <= 8.  Capturable into a CUDA graph.  The wait
 * for the peers is bounded (mpl_peer_status). */
size_t mpl_peer_gather_bytes(int nq, int world);
int mpl_peer_alloc(size_t bytes, void** d_ptr, unsigned char* ipc_handle64);
int mpl_peer_open(const unsigned char* ipc_handle64, void** d_ptr);
int mpl_peer_close(void* d_ptr);
int mpl_peer_free(void* d_ptr);
/* *timed_out = 1 when some step of mpl_allgather_merge_top2_dev gave up waiting for its peers (4 s; a rank that never
 * launched, or launched with another nq); the merged result of that step is undefined.  Synchronises `stream` first. */
int mpl_peer_status(const void* d_sync_own, int* timed_out, void* stream);
int mpl_allgather_merge_top2_dev(const void* d_keys_local, int nq, int rank, int world, void* const* peer_gather,
                                 void* const* peer_sync, int k, int is_float, void* d_idx, void* d_dist, void* d_keys_merged,
                                 void* stream);

/* ---- brute-force k-NN, device pointers (bench `value`, multi-GPU plumbing) ------------------
 * Descriptor rows must be in the library's device layout: row stride = w16*16 bytes, zero padded
 * (mpl_pack_rows_u8_dev produces it; 32-byte ORB rows already are, with w16 = 2).
 * keys_out[nq*2]: 64-bit keys (dist << 32 | global trainIdx) of the two nearest train rows of this
 * shard, global trainIdx = train_offset + local row.  scratch: mpl_knn_hamming_scratch_bytes() -- sized for either Hamming
 * path (mpl_set_hamming_path): the popc scan's per-split partial keys or the tensor-core path's operand images of one train chunk,
 * lists and per-chunk partial keys (C5 shard: 80 MB); 256-byte aligned, or the call stays on the popc scan.
 * mpl_merge_top2_dev: merge n_parts key lists laid out [part][nq][2] (e.g. the NCCL all-gather of
 * every rank's keys_out) into idx/dist (is_float: 0 int32 Hamming, 1 float bits) and/or merged keys. */
size_t mpl_knn_hamming_scratch_bytes(int nq, int nt, int w16);
int    mpl_pack_rows_u8_dev(const void* d_src, int n, size_t src_step, int nbytes_used, void* d_dst, int w16, void* stream);
int    mpl_knn2_hamming_dev(const void* d_q, int nq, const void* d_t, int nt, int w16, long long train_offset,
                            void* d_scratch, size_t scratch_bytes, void* d_keys_out, void* stream);
int    mpl_merge_top2_dev(const void* d_keys_parts, int n_parts, int nq, int k, int is_float,
                          void* d_idx, void* d_dist, void* d_keys_merged, void* stream);
/* Same problem for a batch of independent image pairs (config C4): pair p uses queries
 * d_q + p*nq*w16*16 and train rows d_t + p*nt*w16*16; keys_out [n_pairs][nq][2], train indices local
 * to the pair.  No scratch needed (no train split). */
int    mpl_knn2_hamming_batch_dev(const void* d_q, const void* d_t, int n_pairs, int nq, int nt, int w16,
                                  void* d_keys_out, void* stream);

/* mpl_knn2_hamming_static_dev: mpl_knn2_hamming_dev without a scratch argument and WITHOUT ANY ALLOCATION.  It has no
 * counterpart in the reference; it was added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The partial keys of the train-split
 * path live in a file-scope device array of MPL_KNN_STATIC_SCRATCH_BYTES (4 MiB) that every GPU loading the library holds
 * until the process exits (csrc/knn_hamming_static.cu).  Same inputs, outputs and results as mpl_knn2_hamming_dev.  When the
 * preferred split count x nq x 16 bytes exceeds the array, fewer splits are used (same keys, less parallelism); when even the
 * minimum, ceil(nt / 2^20) >= 2 splits, does not fit, the call returns MPL_ERR_LIMIT with a message and launches nothing.  All calls
 * on a device share the array; nothing in it needs clearing between calls, but calls that OVERLAP on the GPU (two streams,
 * two threads) race on it -- serialise them or use mpl_knn2_hamming_dev with a scratch buffer per stream. */
#define MPL_KNN_STATIC_SCRATCH_BYTES (4u << 20)
size_t mpl_knn_hamming_static_max_bytes(void);
int    mpl_knn2_hamming_static_dev(const void* d_q, int nq, const void* d_t, int nt, int w16, long long train_offset,
                                   void* d_keys_out, void* stream);

/* d_scratch of the float path must be 256-byte aligned.  mpl_knn_l2_last_stats reads back (synchronising the stream) how
 * many queries of the last tensor-core call on that scratch buffer failed certification and went through the exact
 * scan, and whether the operands were bf16-exact (tf32_mode = 0) or ranked in tf32 (1).  *_host: same for the last
 * mpl_knn2_l2f32 / getMatches call (meaningful only when that call took the tensor-core path). */
size_t mpl_knn_l2_scratch_bytes(int nq, int nt, int dim);
int    mpl_knn_l2_last_stats(const void* d_scratch, unsigned* fallbacks, unsigned* tf32_mode, void* stream);
int    mpl_knn_l2_last_stats_host(unsigned* fallbacks, unsigned* tf32_mode);
int    mpl_knn2_l2f32_dev(const void* d_q, int nq, size_t q_ld, const void* d_t, int nt, size_t t_ld, int dim,
                          int sqrt_out, long long train_offset, void* d_scratch, size_t scratch_bytes,
                          void* d_keys_out, void* stream);

/* ---- RANSAC 5-point essential matrix ---------------------------------------------------------
 * mpl_ransac5pt replaces CvModelEstimator3::runRANSAC with CvEMEstimator kernels
 *   (MP/poselib/source/five-point-nister/modelest.cpp:343-474; five-point.cpp:366-503):
 *   p1, p2: n x 2 doubles (camera-normalised coordinates, row-major, contiguous);
 *   samples: h x 5 int32 PRE-DRAWN minimal samples (NULL: max_iters samples are drawn inside like getSubset does,
 *   modelest.cpp:567-610 -- five distinct indices; checkSubset never rejects in the reference, :622-649 -- from an
 *   internal generator, mpl_ransac_set_seed);
 *   th: Sampson threshold (inlier iff (float)err <= th*th); max_iters (reference: 1000) and conf
 *   (reference: 0.999) drive the same sequential selection + adaptive stop as modelest.cpp:377-416, replayed
 *   over per-model (count, sum err) computed on the device; refine != 0 adds the least-squares refit on all
 *   inliers (modelest.cpp:420-464; device Givens QR of the n x 9 constraint rows + one-sided Jacobi SVD of the triangle).
 *   Outputs: E[9] row-major (unit Frobenius norm), mask[n] 0/1, *n_inl (0 = runRANSAC returned false -- no model, or a
 *   refit without a model, modelest.cpp:443-445; E and mask are then untouched), *iters_run.  Optional debug outputs (NULL to skip; when any is
 *   given all h samples are solved and scored): models [<=10h][9] dense in sample-major order,
 *   n_models[h], counts[<=10h], errsums[<=10h].
 * mpl_estimate_essential_api mirrors poselib::estimateEssentialMat(E, p1, p2, "RANSAC", threshold, refine, mask)
 *   (MP/poselib/source/pose_estim.cpp:857-890 -> findEssentialMat five-point.cpp:69-148): returns 1/0 like the
 *   reference's bool (negative MPL_ERR_* on failure; MPL_ERR_ASSERT where CV_Assert would throw).  method "RANSAC",
 *   "ARRSAC" (the reference's default, mpl_arrsac5pt) or "LMEDS" (mpl_lmeds5pt; refine is ignored like in the reference,
 *   pose_estim.cpp:874-877); "USAC" (the reference exits, pose_estim.cpp:878-887; use mpl_usac5pt) / anything else is
 *   refused -- no CPU fallback.  n == 5: solver only, *n_E models stacked in E_out.
 * mpl_lmeds5pt replaces CvModelEstimator3::runLMeDS (modelest.cpp:483-564) as findEssentialMat calls it
 *   (five-point.cpp:124-128: prob 0.999, maxIters 2000): a fixed number of minimal samples (round(log(1 - conf) /
 *   log(1 - 0.55^5)) in [3, max_iters] = 134; samples: h x 5 PRE-DRAWN, or NULL to draw that many inside like getSubset),
 *   per model the median of the n float Sampson errors (sorted through their bit patterns as ints, even n: float sum of
 *   the two middle values times 0.5), strict '<' keeps the first minimal model, sigma = 2.5 * 1.4826 * (1 + 5 / (n - 5)) *
 *   sqrt(min median), at least 0.001, and findInliers with sigma.  No refit.  Outputs: E[9], mask[n], *n_inl (the reference's
 *   result is n_inl >= 5; 0 with *min_median = DBL_MAX: no model had a finite median), optional *min_median, *sigma,
 *   medians[<= 10 h] in evaluation order, n_models[h].  n > 5 (n == 5 is findEssentialMat's solver-only branch). */
int mpl_ransac_set_seed(unsigned long long seed);
/* Experiment switch, default 0 = the reference's 300 Durand-Kerner sweeps (cv::solvePoly, five-point.cpp:430).  1: the
 * minimal solver stops once a sweep moved every root by less than 2^-40 relative (DESIGN.md 4.4 reports the measured
 * root-set / inlier-count differences and the speed-up); process-wide. */
int mpl_set_solver_early_exit(int on);
int mpl_ransac_draw_samples(const double* p1, const double* p2, int n, int h, int32_t* samples, int* n_drawn);
int mpl_ransac5pt(const double* p1, const double* p2, int n, const int32_t* samples, int h, double th,
                  int max_iters, double conf, int refine, double* E, uint8_t* mask, int* n_inl, int* iters_run,
                  double* models, int32_t* n_models, int32_t* counts, double* errsums);
int mpl_estimate_essential_api(const double* p1, const double* p2, int n, const char* method, double threshold,
                               int refine, double* E_out, int* n_E, uint8_t* mask);
int mpl_lmeds5pt(const double* p1, const double* p2, int n, const int32_t* samples, int h, double conf, int max_iters,
                 double* E, uint8_t* mask, int* n_inl, double* min_median, double* sigma, double* medians, int32_t* n_models);
/* mpl_arrsac5pt replaces CvModelEstimator3::runARRSAC (MP/poselib/source/five-point-nister/modelest.cpp:197-341) -- what
 *   poselib::estimateEssentialMat runs for its DEFAULT method "ARRSAC" (pose_estim.cpp:866-869): theia::Arrsac
 *   (MP/poselib/include/arrsac/arrsac.h:206-545: PROSAC-ordered initial hypotheses verified by SPRT on the first block of
 *   100 correspondences, inner RANSAC on the best support, preemptive breadth-first scoring in blocks of 100 / sub-blocks
 *   of 20) with parameters (5, th^2, 500, 100, 14, 8), then findInliers, the acceptance rule (< 15 inliers, or < 50 of
 *   more than 200 -> false) and, with refine, robustEssentialRefine (pose_estim.cpp:337-837: pseudo-Huber IRLS, threshold
 *   th / 50).  Hypothesis generation (5-point / 8-point solves, ValidModel) and verification (inlier bit matrix) run on
 *   the device in speculative batches; the schedule is the reference's, walked on the host.  Parity is statistical (own
 *   solver roundings; see csrc/arrsac.cu).  *n_inl = 0: the reference returned false.  n_hypotheses / n_batches: optional
 *   diagnostics.  mpl_arrsac_set_seed sets the two sampler states (0 = cv::RNG's default 0xffffffff); like the reference's
 *   function-local static generators they keep running across calls of a thread. */
int mpl_arrsac_set_seed(unsigned long long prosac_state, unsigned long long random_state);
int mpl_arrsac5pt(const double* p1, const double* p2, int n, double th, int refine, double* E, uint8_t* mask, int* n_inl,
                  int* n_hypotheses, int* n_batches);
/* mpl_vfc_filter replaces filterWithVFC (MP/matchinglib/source/vfcMatches.cpp:49-86 -> class VFC, source/vfc.cpp, SparseVFC), the
 *   post-filter matchinglib::getMatches applies with VFCrefine = true (source/matchers.cpp:721-733).  kp1 / kp2: keypoint records
 *   with the cv::Point2f pt at offset 0 (cv::KeyPoint: stride 28); matches: cv::DMatch records.  Outputs: keep[n_matches] (1 = the
 *   match is in matches_out), *n_keep, *vfc_code = filterWithVFC's return value (0 ok, -1 fewer than 5 matches -- nothing kept,
 *   -2 fewer than 10 % kept: "Maybe VFC failed"), *n_iterations (may be NULL) = EM iterations run.  The set-up (normalisation,
 *   the rand()-driven choice of 16 control points) runs on the host as written, the kernel matrix and the EM iterations in one
 *   kernel launch that does the reference's float arithmetic in the reference's order (sequential sums, glibc's expf restated):
 *   the kept set equals the reference's for the same rand() state (csrc/vfc.cu, tests/test_vfc_gpu.py).  mpl_vfc_set_seed
 *   restarts the calling thread's rand() stream (the reference draws from the process' state; a fresh process is srand(1)). */
int mpl_vfc_set_seed(unsigned seed);
int mpl_vfc_filter(const void* kp1, int n1, size_t kp_stride1, const void* kp2, int n2, size_t kp_stride2, const mpl_dmatch* matches,
                   int n_matches, uint8_t* keep, int* n_keep, int* vfc_code, int* n_iterations);
/* mpl_usac5pt replaces estimateEssentialMatUsac (MP/poselib/source/usac/usac_estimations.cpp:283-735) -- the robust
 *   estimator behind poselib::estimateEssentialOrPoseUSAC (pose_estim.cpp:1737), the default RobMethod of poselib-test:
 *   USAC<EssentialMatEstimator>::solve (MP/poselib/include/usac/estimators/USAC.h:345-612; uniform or PROSAC sampling,
 *   SPRT verification over a random evaluation pool, LOSAC local optimisation, standard / PROSAC / SPRT stopping) with the
 *   estimator of EssentialMatEstimator.h (5-point minimal solver, collinearity pre-test of the sample, oriented-epipolar
 *   pre-test of the model, plain-double Sampson error, 8-point refits with Torr weights).  Minimal solves, model tests, the
 *   inlier bit matrix and the whole local optimisation run on the device; the schedule is the reference's, walked on the host
 *   over those bits with glibc's rand() restated (csrc/usac.cu).  Parity is statistical.
 *   p1, p2: n x 2 camera-normalised points.  sorted_idx: n correspondence indices by increasing match distance (PROSAC,
 *   what getSortedMatchIdx hands over) or NULL (uniform sampling).  mpl_usac_params_default fills the values
 *   estimateEssentialMatUsac uses (conf 0.99, 50000 hypotheses, LOSAC 5 x 14 / 2.0 / 4, PROSAC 1000 / 20 / 0.99, beta 0.09,
 *   SPRT delta 0.05 epsilon 0.15, tM 2314 mS 8.5 for the Nister estimator).  estimator: 0 NISTER, 2 STEWENIUS (same device
 *   solver, tM / mS are the caller's); refine must be 0 (REFINE_WEIGHTS) and degeneracy 0 -- other values return
 *   MPL_ERR_BAD_ARG, there is no CPU fallback.  Outputs: E[9] (zeros when no model was found, like final_model_params_),
 *   mask[n], res (may be NULL).  res->ok = 0: solve() returned false (fewer than 5 points, or fewer than minStopLen with
 *   PROSAC) -- the reference then returns EXIT_FAILURE.
 *   lo_mode: the reference's local optimisation is data dependent in a way its code does not state -- FTools::singulF
 *   (FundmatrixFunctions.cpp:334-360) goes through the ccmath port MathTools::svduv, which for SOME matrices with two nearly
 *   equal singular values (essential-like ones) returns wrong singular triplets (measured: D = (0.651, 1.1e-5, 0.697) for a
 *   matrix whose singular values are (0.717, 0.697, 1.0e-5), |U D V^T - F|^2 = 0.09).  In scenes where that happens every
 *   refitted model is mangled, fails its SPRT, and the reference's verbose mode prints "Inlier count before: N, after: N"
 *   for every local optimisation; in other scenes its local optimisation works.  That 200-line third-party routine is not
 *   restated here, so there are two modes and the reference behaves like one or the other, scene by scene
 *   (tests/test_usac_gpu.py reports which):  lo_mode 0 (default) runs the local optimisation as USAC.h:948-1060 states it,
 *   on the device (one launch, a CTA per repetition);  lo_mode 1 draws the inner samples (the random stream stays the
 *   reference's) and counts their models as rejected -- what the reference does where its singulF fails. */
typedef struct mpl_usac_params {
    double th, conf;
    int max_hypotheses, prevalidate_sample;
    double prosac_beta, prosac_non_rand_conf; int prosac_max_samples, prosac_min_stop_len;
    double sprt_delta, sprt_epsilon, sprt_tM, sprt_mS;
    int lo_inner_sample_size, lo_inner_reps, lo_num_steps; double lo_threshold_multiplier;
    int estimator, refine, degeneracy;
    int lo_mode;                                /* 0 (default): local optimisation as stated; 1: its models count as rejected -- see above */
    unsigned seed;                              /* srand(seed); the reference uses time(NULL) */
} mpl_usac_params;
typedef struct mpl_usac_result {
    int ok;
    unsigned hyp_count, model_count, rejected_sample_count, rejected_model_count, total_points_verified,
             num_local_optimizations, best_inlier_count;
    double sprt_delta, sprt_epsilon;            /* usac_results_.sprt_delta_/epsilon_: the last stored test (0 if none) */
    int n_batches, n_launches;
} mpl_usac_result;
void mpl_usac_params_default(mpl_usac_params* p, double th);
int mpl_usac5pt(const double* p1, const double* p2, int n, const unsigned* sorted_idx, const mpl_usac_params* params,
                double* E, uint8_t* mask, mpl_usac_result* res);
/* Host logic exposed for the CPU tests: the first `count` values of rand() after srand(seed) as csrc/usac.cu restates them. */
int mpl_usac_rand_stream(unsigned seed, int count, int* out);
/* mpl_get_pose_triang_pts_api mirrors poselib::getPoseTriangPts(E, p1, p2, R, t, Q, mask, dist, translatE)
 *   (MP/poselib/source/pose_estim.cpp:913-946 -> recoverPose five-point.cpp:150-338, decomposeEssentialMat :340-352):
 *   E 3x3 row-major, p1/p2 n x 2 doubles (camera-normalised).  mask (may be NULL): n bytes; mask_given != 0 -> its
 *   content restricts the points that count (the RANSAC inlier mask) and is AND-ed in place, mask_given == 0 -> treated
 *   as all ones and overwritten.  Outputs R[9], t[3], Q[n*3] (triangulated points of the chosen combination, all n).
 *   Returns the number of valid 3-D points (>= 0), -1 if R, t or Q is missing (the reference's code), or MPL_ERR_*.
 *   The per-point triangulation / cheirality / distance tests of all four (R, +-t) combinations run on the device. */
int mpl_get_pose_triang_pts_api(const double* E, const double* p1, const double* p2, int n, uint8_t* mask, int mask_given,
                                double dist, int translatE, double* R, double* t, double* Q);
/* ---- both halves of the hot path for a batch of image pairs, device resident in between --------------------------
 * (SURVEY.md 8(e) "batch of image pairs", 8(f) N3; the reference driver's per-pair sequence tests/poselib-test/main.cpp:
 *  1461-1599, 1717 with host vectors between the steps.)  For every pair: descriptors (uint8 rows of `cols` bytes, or float32
 *  rows of `cols` values) -> brute-force k = 2 (matcher_name "LINEAR" or "BRUTEFORCENMS") -> ratio test + cv::DMatch list in
 *  ascending queryIdx (matchers.cpp:601-625) -> correspondences (pts1[queryIdx], pts2[trainIdx]) -> h minimal samples drawn
 *  on the device (getSubset, modelest.cpp:567-610: five distinct indices; counter-based generator from the pair's seed) ->
 *  5-point RANSAC as in mpl_ransac5pt with max_iters = h, selection replayed on the device (modelest.cpp:394-415), refit
 *  (:420-464).  pts_mode 0: pts are n x 2 doubles, camera-normalised by the caller.  pts_mode 1: pts are n x 2 floats in
 *  pixels (cv::KeyPoint::pt); the library applies ImgToCamCoordTrans (pose_helper.cpp:1100-1109) with K0 / K1 (row-major
 *  3 x 3) and Remove_LensDist (pose_helper.cpp:1169-1279) with dist0 / dist1 (8 OpenCV coefficients; all ~0 = no
 *  undistortion, as the reference decides) on the device, float -> double as the driver does (main.cpp:1588-1596); a pair
 *  with fewer than 16 surviving correspondences gets status 1 (the driver skips it).  th is in camera units
 *  (main.cpp:1597-1599).  Pairs are processed in sub-batches on two internal streams (uploads overlap kernels); one call
 *  from one host thread keeps the GPU busy.  Host buffers in mpl_host_alloc memory upload at the PCIe link rate. */
typedef struct mpl_pair_io {
    const void* desc1; const void* desc2;      /* cv::Mat::data of the two descriptor matrices, n1 / n2 rows */
    size_t step1, step2;                       /* cv::Mat::step (bytes per row) */
    const void* pts1; const void* pts2;        /* keypoint coordinates, n1 x 2 / n2 x 2 (see pts_mode) */
    int n1, n2;
    unsigned long long seed;                   /* minimal-sample generator seed of this pair */
    mpl_dmatch* matches;                       /* out: >= n1 entries */
    uint8_t* mask;                             /* out (may be NULL): >= n1 bytes, RANSAC inlier mask over the match list */
    int32_t* samples;                          /* out (may be NULL): h x 5, the samples that were drawn */
} mpl_pair_io;
typedef struct mpl_pair_params {
    int cv_type;                               /* 0 = CV_8U (Hamming), 5 = CV_32F (L2) */
    int cols;                                  /* descriptor length (bytes or floats) */
    const char* matcher_name;                  /* "LINEAR" | "BRUTEFORCENMS" */
    int ratio_test;
    int pts_mode;                              /* 0: double camera-normalised, 1: float pixels + K0 / K1 / dist0 / dist1 */
    double K0[9], K1[9], dist0[8], dist1[8];
    double th; int h; double conf; int refine; /* as mpl_ransac5pt (h = sample budget = max_iters) */
} mpl_pair_params;
typedef struct mpl_pair_result {
    int n_matches, n_corr, n_samples, n_inl, iters_run, status;   /* n_inl = 0: runRANSAC returned false; status 1: undistortion failed */
    double E[9];
} mpl_pair_result;
int mpl_match_pose_batch(const mpl_pair_io* pairs, int n_pairs, const mpl_pair_params* prm, mpl_pair_result* results);
/* mpl_match_pose_pair: the same for ONE pair of binary descriptors with camera-normalised double keypoints (a batch of
 *   one: same kernels, same results).  Outputs: matches[<= n1] / *n_matches, samples_out[h x 5] (optional) / *n_samples,
 *   E[9], mask[*n_matches] over the match list, *n_inl, *iters_run. */
int mpl_match_pose_pair(const uint8_t* desc1, int n1, size_t step1, const uint8_t* desc2, int n2, size_t step2, int nbytes,
                        const double* pts1, const double* pts2, const char* matcher_name, int ratio_test,
                        double th, int h, double conf, int refine, unsigned long long seed,
                        mpl_dmatch* matches, int* n_matches, int32_t* samples_out, int* n_samples,
                        double* E, uint8_t* mask, int* n_inl, int* iters_run);
/* device-pointer stages (enqueue only): corr = n x 32-byte records (x1,y1,x2,y2); models h x 90 doubles;
 * offsets h+1 int32 (offsets[h] = number of dense models); dense up to 10h x 9 doubles. */
int mpl_ransac_pack_corr_dev(const void* d_p1, const void* d_p2, int n, void* d_corr, void* stream);
int mpl_ransac_solve_dev(const void* d_corr, int n, const void* d_samples, int h, void* d_models, void* d_n_models,
                         void* d_offsets, void* d_dense, void* stream);
int mpl_ransac_score_dev(const void* d_dense, int m, const void* d_corr, int n, double th, void* d_counts,
                         void* d_errsums, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MPLB200_H_ */
