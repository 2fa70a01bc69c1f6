// matchinglib_poselib_b200.hpp -- drop-in C++ shim: the reference's own API names on top of libmplb200.so.
//
// A maintainer of josefmaierfl/matchinglib_poselib replaces the bodies of
//   matchinglib::getMatches            (matchinglib/include/matchinglib/matchinglib_matchers.h:61-64,
//                                       matchinglib/source/matchers.cpp:115-736)
//   poselib::estimateEssentialMat      (poselib/include/poselib/pose_estim.h:204-210,
//                                       poselib/source/pose_estim.cpp:857-890)
// with the inline functions below (or includes this header in place of those translation units) and links
// -lmplb200.  Signatures, defaults, return codes, exceptions and output ownership are the reference's.  Only the
// selectors the B200 hot path serves are accepted -- matcher_name in {"LINEAR", "BRUTEFORCENMS"}, method ==
// "RANSAC" -- everything else returns the reference's "not supported" result; there is no CPU fallback.
//
// Needs OpenCV headers (cv::Mat, cv::KeyPoint, cv::DMatch); the build image of this repository has none, so the
// header is syntax-checked against a minimal mock in tests/cv_mock and exercised functionally through the
// identical C ABI by the Python mirror (numpy arrays share cv::Mat's row-major layout).
#pragma once
#if defined(__has_include)
#  if !__has_include(<opencv2/core.hpp>)
#    error "matchinglib_poselib_b200.hpp needs OpenCV's <opencv2/core.hpp>"
#  endif
#endif
#include <opencv2/core.hpp>
#include <opencv2/imgproc.hpp>      // convexHull, contourArea (estimateSprtDeltaInit)

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <iostream>
#include <string>
#include <cstring>
#include <vector>

#include "mplb200.h"

// matchinglib/include/matchinglib/vfcMatches.h (global namespace in the reference), source/vfcMatches.cpp:49-86 -> class VFC
// (source/vfc.cpp, SparseVFC): 0 ok, -1 too few matches for refinement, -2 "maybe VFC failed" (fewer than 10 % kept).
inline int filterWithVFC(std::vector<cv::KeyPoint> const& keypL, std::vector<cv::KeyPoint> const& keypR,
                         std::vector<cv::DMatch> const& matches_in, std::vector<cv::DMatch>& matches_out)
{
    static_assert(sizeof(cv::DMatch) == sizeof(mpl_dmatch), "cv::DMatch layout");
    matches_out.clear();
    std::vector<uchar> keep(matches_in.size() + 1);
    int n_keep = 0, code = 0;
    const int rc = mpl_vfc_filter(keypL.data(), (int)keypL.size(), sizeof(cv::KeyPoint), keypR.data(), (int)keypR.size(), sizeof(cv::KeyPoint),
                                  reinterpret_cast<const mpl_dmatch*>(matches_in.data()), (int)matches_in.size(), keep.data(), &n_keep, &code, nullptr);
    if (rc < 0) CV_Error(cv::Error::StsError, mpl_last_error());
    if (code == -1) { std::cout << "Too less matches for refinement with VFC!" << std::endl; return -1; }
    for (size_t i = 0; i < matches_in.size(); ++i) if (keep[i]) matches_out.push_back(matches_in[i]);
    return code;
}

namespace matchinglib {

// matchers.cpp:115-736, LINEAR (:525-714) and BRUTEFORCENMS (:476-519) branches.
// Returns 0 ok, -1 wrong input, -2 matcher not supported, -3 matching failed, -4 too few keypoints.
inline int getMatches(const std::vector<cv::KeyPoint>& keypoints1, const std::vector<cv::KeyPoint>& keypoints2,
                      const cv::Mat& descriptors1, const cv::Mat& descriptors2, cv::Size imgSi,
                      std::vector<cv::DMatch>& finalMatches, const std::string& matcher_name = "GMBSOF",
                      bool VFCrefine = false, bool ratioTest = true, const std::string& descriptor_name = "",
                      std::string idxPars_NMSLIB = "", std::string queryPars_NMSLIB = "", const size_t nr_threads = 0)
{
    (void)imgSi; (void)descriptor_name; (void)idxPars_NMSLIB; (void)queryPars_NMSLIB;
    CV_Assert(descriptors1.type() == descriptors2.type());                      // matchers.cpp:119
    finalMatches.clear();                                                        // matchers.cpp:135
    static_assert(sizeof(cv::DMatch) == sizeof(mpl_dmatch), "cv::DMatch layout");
    std::vector<cv::DMatch> out(descriptors1.rows > 0 ? (size_t)descriptors1.rows : 1);
    int n_out = 0;
    // the reference reads descriptorsN.data with cols as the row length (:567, :634): pass step so that
    // non-continuous Mats work too
    const int err = mpl_get_matches_api((int)keypoints1.size(), (int)keypoints2.size(),
                                        descriptors1.data, descriptors1.rows, descriptors1.cols, (size_t)descriptors1.step[0], descriptors1.depth(),
                                        descriptors2.data, descriptors2.rows, descriptors2.cols, (size_t)descriptors2.step[0], descriptors2.depth(),
                                        matcher_name.c_str(), ratioTest ? 1 : 0, (int)nr_threads,
                                        reinterpret_cast<mpl_dmatch*>(out.data()), &n_out);
    if (err == MPL_ERR_ASSERT) CV_Error(cv::Error::StsAssert, mpl_last_error());
    if (err != 0) std::cout << mpl_last_error() << std::endl;                    // the reference prints its diagnostics to cout
    if (err == 0 || err == -3) finalMatches.assign(out.begin(), out.begin() + n_out);
    if (err == 0 && VFCrefine) {                                                 // matchers.cpp:721-733 (the early returns skip it there too)
        std::vector<cv::DMatch> vfcfilteredMatches;
        if (!filterWithVFC(keypoints1, keypoints2, finalMatches, vfcfilteredMatches))
            if (vfcfilteredMatches.size() > 8 || finalMatches.size() < 24) finalMatches = vfcfilteredMatches;
    }
    return err;
}

}  // namespace matchinglib

// matchinglib/include/matchinglib/gms.h (global namespace in the reference), source/gms.cpp:54-82 -> MatchGMS
// (thirdparty/gms-1.0): the grid statistics run on the GPU (mpl_gms_filter); getCorrespondences (correspondences.cpp:386)
// calls the first overload with useScale = useRotation = false.
inline int filterMatchesGMS(const std::vector<cv::KeyPoint>& keypoints1, const cv::Size imageSize1,
                            const std::vector<cv::KeyPoint>& keypoints2, const cv::Size imageSize2,
                            const std::vector<cv::DMatch>& matches, std::vector<bool>& inlierMask,
                            const bool useScale = false, const bool useRotation = false)
{
    static_assert(sizeof(cv::DMatch) == sizeof(mpl_dmatch), "cv::DMatch layout");
    std::vector<uchar> m(matches.size() + 1);
    int n = 0;
    const int rc = mpl_gms_filter(keypoints1.data(), (int)keypoints1.size(), sizeof(cv::KeyPoint), imageSize1.width, imageSize1.height,
                                  keypoints2.data(), (int)keypoints2.size(), sizeof(cv::KeyPoint), imageSize2.width, imageSize2.height,
                                  reinterpret_cast<const mpl_dmatch*>(matches.data()), (int)matches.size(), useScale ? 1 : 0, useRotation ? 1 : 0,
                                  m.data(), &n);
    if (rc != 0) { std::cout << mpl_last_error() << std::endl; return 0; }
    if (n > 0) inlierMask.assign(m.begin(), m.begin() + (std::ptrdiff_t)matches.size());     // MatchGMS.cpp:92-96: assigned only by a run with inliers
    return n;
}
inline int filterMatchesGMS(const std::vector<cv::KeyPoint>& keypoints1, const cv::Size imageSize1,
                            const std::vector<cv::KeyPoint>& keypoints2, const cv::Size imageSize2,
                            const std::vector<cv::DMatch>& matches, std::vector<cv::DMatch>& matches_filtered,
                            const bool useScale = false, const bool useRotation = false)
{
    std::vector<bool> inlierMask;
    const int n = filterMatchesGMS(keypoints1, imageSize1, keypoints2, imageSize2, matches, inlierMask, useScale, useRotation);
    matches_filtered.clear();                                                    // gms.cpp:71-80
    matches_filtered.resize((size_t)n);
    if (n) for (size_t i = 0, j = 0; i < inlierMask.size(); i++) if (inlierMask[i]) matches_filtered[j++] = matches[i];
    return n;
}

namespace poselib {

// pose_estim.cpp:857-890 -> findEssentialMat (five-point-nister/five-point.cpp:69-148), methods "ARRSAC" (the default;
// runARRSAC + robustEssentialRefine on the GPU, mpl_arrsac5pt), "RANSAC" and "LMEDS" (runLMeDS, mpl_lmeds5pt; no refinement, :874-877).
// p1, p2: n x 2 (or n x 1 2-channel) camera-normalised points of any depth; E: 3 x 3 CV_64F (stacked 3k x 3 when
// n == 5); mask: 1 x n CV_8U.
inline bool estimateEssentialMat(cv::OutputArray E, cv::InputArray p1, cv::InputArray p2, const std::string& method = "ARRSAC",
                                 double threshold = 1.6, bool refine = true, cv::OutputArray mask = cv::noArray())
{
    if (method == "USAC") {                                                      // pose_estim.cpp:878-881
        std::cout << "USAC must be executed by function estimateEssentialOrPoseUSAC as it needs additional paramters! Exiting." << std::endl;
        std::exit(1);
    }
    if (method != "RANSAC" && method != "ARRSAC" && method != "LMEDS") {         // :883-887
        std::cout << "Either there is a typo in the specified robust estimation method or the method is not supported. Exiting." << std::endl;
        std::exit(1);
    }
    cv::Mat points1 = p1.getMat().clone(), points2 = p2.getMat().clone();        // five-point.cpp:77-78 deep copies
    const int npoints = points1.checkVector(2);
    CV_Assert(npoints >= 5 && points2.checkVector(2) == npoints && points1.type() == points2.type());   // :81-82
    points1.convertTo(points1, CV_64F);                                          // :89-90
    points2.convertTo(points2, CV_64F);
    points1 = points1.reshape(1, npoints);                                       // n x 2, continuous after clone/convert
    points2 = points2.reshape(1, npoints);
    if (!E.needed()) return false;                                               // :137-144
    double e_buf[90];
    int n_E = 0;
    std::vector<uchar> m((size_t)npoints);
    const int rc = mpl_estimate_essential_api(points1.ptr<double>(), points2.ptr<double>(), npoints, method.c_str(), threshold,
                                              refine ? 1 : 0, e_buf, &n_E, m.data());
    if (rc == MPL_ERR_ASSERT) CV_Error(cv::Error::StsAssert, mpl_last_error());
    if (rc < 0) { std::cout << mpl_last_error() << std::endl; return false; }
    if (rc != 1) return false;                                                   // :117-118: on failure neither E nor mask is written
    if (mask.needed()) cv::Mat(1, npoints, CV_8U, m.data()).copyTo(mask);        // :130-135
    cv::Mat(3 * (npoints == 5 ? n_E : 1), 3, CV_64F, e_buf).copyTo(E);           // :137-141
    return true;
}

// poselib::estimateEssentialOrPoseUSAC (pose_estim.h:212-223; pose_estim.cpp:1737-2262 -> estimateEssentialMatUsac,
// usac/usac_estimations.cpp:283-735) for the configurations mpl_usac5pt serves: estimator POSE_NISTER / POSE_STEWENIUS, refinealg
// REF_WEIGHTS, degeneracyCheck DEGEN_NO_CHECK (poselib-test: cfgUSAC digits 4-6 = 0, 0|2, 0).  Any other value prints and returns
// -1 like the reference does for values it does not know -- nothing falls back to the CPU.  The SPRT / PROSAC start values follow
// the reference's first call (:1810-1976); its cross-call running statistics (function statics) are not kept.
enum UsacChkDegenType { DEGEN_NO_CHECK, DEGEN_QDEGSAC, DEGEN_USAC_INTERNAL };                    // pose_estim.h:61
enum PoseEstimator { POSE_NISTER, POSE_EIG_KNEIP, POSE_STEWENIUS };                              // :64
enum RefineAlg { REF_WEIGHTS, REF_8PT_PSEUDOHUBER, REF_EIG_KNEIP, REF_EIG_KNEIP_WEIGHTS, REF_STEWENIUS, REF_STEWENIUS_WEIGHTS, REF_NISTER, REF_NISTER_WEIGHTS };   // :67-76
enum SprtInit { SPRT_DEFAULT_INIT = 0x0, SPRT_DELTA_AUTOM_INIT = 0x1, SPRT_EPSILON_AUTOM_INIT = 0x2 };       // :91
struct ConfigUSAC {                                                                              // :94-132, same defaults
    double focalLength = 800; double th_pixels = 0.8;
    UsacChkDegenType degeneracyCheck = DEGEN_USAC_INTERNAL; PoseEstimator estimator = POSE_STEWENIUS; RefineAlg refinealg = REF_STEWENIUS_WEIGHTS;
    bool prevalidateSample = false, noAutomaticProsacParamters = false;
    int automaticSprtInit = SPRT_DELTA_AUTOM_INIT | SPRT_EPSILON_AUTOM_INIT;
    std::vector<cv::DMatch>* matches = nullptr; std::vector<cv::KeyPoint>* keypoints1 = nullptr; std::vector<cv::KeyPoint>* keypoints2 = nullptr;
    unsigned int nrMatchesVfcFiltered = 0; cv::Size imgSize = cv::Size(800, 600); double degenDecisionTh = 0.85;
    unsigned seed = 1; int lo_mode = 0;                                                          // this library's additions (mpl_usac_params)
};

inline int estimateEssentialOrPoseUSAC(const cv::Mat& p1, const cv::Mat& p2, cv::OutputArray E, double th, ConfigUSAC& cfg, bool& isDegenerate,
                                       cv::OutputArray inliers = cv::noArray(), cv::OutputArray R_degenerate = cv::noArray(),
                                       cv::OutputArray inliers_degenerate_R = cv::noArray(), cv::OutputArray R = cv::noArray(),
                                       cv::OutputArray t = cv::noArray(), bool verbose = false)
{
    if (cfg.estimator != POSE_NISTER && cfg.estimator != POSE_STEWENIUS) { std::cout << "Estimator not supported!" << std::endl; return -1; }     // :1776-1778
    if (cfg.refinealg != REF_WEIGHTS) { std::cout << "Refinement algorithm not supported!" << std::endl; return -1; }                            // :1806-1808
    if (cfg.degeneracyCheck != DEGEN_NO_CHECK) { std::cout << "Mothod for checking degeneracy not available!" << std::endl; return -1; }         // :2218-2222
    if (cfg.automaticSprtInit < 0 || cfg.automaticSprtInit > 3) { std::cout << "Method for SPRT initialization not supported!" << std::endl; return -1; }
    CV_Assert(p1.cols == 2 && p2.cols == 2 && p1.rows == p2.rows && p1.type() == CV_64FC1 && p2.type() == CV_64FC1);   // usac_estimations.cpp:315
    CV_Assert(!cfg.matches || p1.rows == (int)cfg.matches->size());
    mpl_usac_params prm;
    mpl_usac_params_default(&prm, th);
    if (cfg.automaticSprtInit & SPRT_DELTA_AUTOM_INIT) {                           // estimateSprtDeltaInit, pose_helper.cpp:2830-2867
        CV_Assert(cfg.matches && cfg.keypoints1 && cfg.keypoints2);
        std::vector<cv::Point2f> pts1, pts2, h1, h2;
        for (const cv::DMatch& m : *cfg.matches) { pts1.push_back((*cfg.keypoints1)[m.queryIdx].pt); pts2.push_back((*cfg.keypoints2)[m.trainIdx].pt); }
        cv::convexHull(pts1, h1); cv::convexHull(pts2, h2);
        double area = std::min(cv::contourArea(h1), cv::contourArea(h2));
        const double maxEpi = std::sqrt((double)(cfg.imgSize.width * cfg.imgSize.width + cfg.imgSize.height * cfg.imgSize.height)) * 2 * cfg.th_pixels;
        area = std::max(area, 6 * maxEpi);
        prm.sprt_delta = std::max(maxEpi / area, 0.001);
    }
    if (cfg.automaticSprtInit & SPRT_EPSILON_AUTOM_INIT) {                         // estimateSprtEpsilonInit, pose_helper.cpp:2881-2889
        CV_Assert(cfg.matches);
        prm.sprt_epsilon = std::min(std::max(0.8 * (double)cfg.nrMatchesVfcFiltered / (double)cfg.matches->size(), 0.1), 0.4);
    }
    if (!cfg.noAutomaticProsacParamters) prm.prosac_beta = prm.sprt_delta;         // pose_estim.cpp:1972-1975
    std::vector<unsigned> sorted;
    if (cfg.matches) {                                                             // getSortedMatchIdx, pose_helper.cpp:2902-2926
        sorted.resize(cfg.matches->size());
        for (size_t i = 0; i < sorted.size(); ++i) sorted[i] = (unsigned)i;
        const std::vector<cv::DMatch>& mm = *cfg.matches;
        std::stable_sort(sorted.begin(), sorted.end(), [&mm](unsigned a, unsigned b) { return mm[a].distance < mm[b].distance; });
    }
    prm.estimator = cfg.estimator == POSE_STEWENIUS ? 2 : 0;
    if (cfg.estimator == POSE_STEWENIUS) { prm.sprt_tM = 2736.0; prm.sprt_mS = 6.0; }                // usac_estimations.cpp:392-421, first call
    prm.seed = cfg.seed; prm.lo_mode = cfg.lo_mode;
    cv::Mat a = p1.isContinuous() ? p1 : p1.clone(), b = p2.isContinuous() ? p2 : p2.clone();
    double e_buf[9];
    std::vector<uchar> m((size_t)std::max(a.rows, 1));
    mpl_usac_result res;
    const int rc = mpl_usac5pt(a.ptr<double>(), b.ptr<double>(), a.rows, cfg.matches ? sorted.data() : nullptr, &prm, e_buf, m.data(), &res);
    if (rc < 0) { std::cout << mpl_last_error() << std::endl; return -2; }
    if (!res.ok) { std::cout << "USAC failed!" << std::endl; return -2; }                           // pose_estim.cpp:2018-2022
    if (verbose) std::cout << "Number of hypotheses/models: " << res.hyp_count << "/" << res.model_count << std::endl
                           << "Max inliers/total points: " << res.best_inlier_count << "/" << a.rows << std::endl;
    isDegenerate = false;                                                          // :2023
    if (E.needed()) cv::Mat(3, 3, CV_64F, e_buf).copyTo(E);
    if (inliers.needed()) cv::Mat(1, a.rows, CV_8U, m.data()).copyTo(inliers);
    if (R.needed() && !R.empty()) R.clear();                                       // :2036-2042: R, t only come from Kneip's solver
    if (t.needed() && !t.empty()) t.clear();
    (void)R_degenerate; (void)inliers_degenerate_R;
    return 0;
}

// poselib::getPoseTriangPts (pose_estim.h:192-200; pose_estim.cpp:913-946 -> recoverPose five-point.cpp:150-338)
inline int getPoseTriangPts(cv::InputArray E, cv::InputArray p1, cv::InputArray p2, cv::OutputArray R, cv::OutputArray t, cv::OutputArray Q,
                            cv::InputOutputArray mask = cv::noArray(), const double dist = 50.0, bool translatE = false)
{
    if (!R.needed() || !t.needed() || !Q.needed()) return -1;                    // :925-926
    cv::Mat Em, points1, points2;
    E.getMat().convertTo(Em, CV_64F);
    p1.getMat().copyTo(points1);                                                 // five-point.cpp:154-156 deep copies
    p2.getMat().copyTo(points2);
    const int npoints = points1.checkVector(2);
    CV_Assert(npoints >= 0 && points2.checkVector(2) == npoints && points1.type() == points2.type());   // :157-159
    points1.convertTo(points1, CV_64F);
    points2.convertTo(points2, CV_64F);
    points1 = points1.reshape(1, npoints);
    points2 = points2.reshape(1, npoints);
    Em = Em.clone();
    bool given = false;
    std::vector<uchar> m((size_t)npoints, 1);
    if (mask.needed()) {
        cv::Mat mm = mask.getMat();
        if (!mm.empty()) { CV_Assert((int)mm.total() == npoints && mm.type() == CV_8U); std::memcpy(m.data(), mm.clone().data, (size_t)npoints); given = true; }
    }
    double r_buf[9], t_buf[3];
    std::vector<double> q((size_t)npoints * 3 + 3);
    const int rc = mpl_get_pose_triang_pts_api(Em.ptr<double>(), points1.ptr<double>(), points2.ptr<double>(), npoints,
                                               mask.needed() ? m.data() : nullptr, given ? 1 : 0, dist, translatE ? 1 : 0, r_buf, t_buf, q.data());
    if (rc == MPL_ERR_ASSERT) CV_Error(cv::Error::StsAssert, mpl_last_error());
    if (rc < -1) CV_Error(cv::Error::StsError, mpl_last_error());
    cv::Mat(3, 3, CV_64F, r_buf).copyTo(R);                                      // :930-940
    cv::Mat(3, 1, CV_64F, t_buf).copyTo(t);
    cv::Mat(npoints, 3, CV_64F, q.data()).copyTo(Q);
    if (mask.needed()) cv::Mat(1, npoints, CV_8U, m.data()).copyTo(mask);        // five-point.cpp:284-301
    return rc;
}

}  // namespace poselib
