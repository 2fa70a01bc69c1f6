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

#include <cstdlib>
#include <iostream>
#include <string>
#include <cstring>
#include <vector>

#include "mplb200.h"

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
    if (VFCrefine) {                                                             // :722-733 optional post-filter, not on the hot path
        std::cout << "VFC refinement is not served by the B200 hot path." << std::endl;
        return -2;
    }
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
// runARRSAC + robustEssentialRefine on the GPU, mpl_arrsac5pt) and "RANSAC".
// p1, p2: n x 2 (or n x 1 2-channel) camera-normalised points of any depth; E: 3 x 3 CV_64F (stacked 3k x 3 when
// n == 5); mask: 1 x n CV_8U.
inline bool estimateEssentialMat(cv::OutputArray E, cv::InputArray p1, cv::InputArray p2, const std::string& method = "ARRSAC",
                                 double threshold = 1.6, bool refine = true, cv::OutputArray mask = cv::noArray())
{
    if (method == "USAC") {                                                      // pose_estim.cpp:878-881
        std::cout << "USAC must be executed by function estimateEssentialOrPoseUSAC as it needs additional paramters! Exiting." << std::endl;
        std::exit(1);
    }
    if (method != "RANSAC" && method != "ARRSAC") {
        // :883-887 exits for unknown strings; LMEDS is served by the reference but not by this library -- fail the call
        // instead of terminating the caller's process
        std::cout << "Method " << method << " is not served by the B200 hot path (ARRSAC, RANSAC)." << std::endl;
        return false;
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
