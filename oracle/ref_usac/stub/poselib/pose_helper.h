// Stand-in for poselib/pose_helper.h + pose_estim.h: the six poselib:: functions EssentialMatEstimator.h names.  nearZero is
// the reference's definition (pose_helper.h:82-87).  The others are reached only from the Kneip-eigensolver, degeneracy-upgrade
// and pseudo-Huber refinement branches, which the ref_usac driver never configures: they abort if called.
#pragma once
#include "opencv2/core.hpp"
#include <Eigen/Core>
namespace poselib {
inline bool nearZero(double d) { const double EPSILON = 1e-3; return (d < EPSILON) && (d > -EPSILON); }
[[noreturn]] inline void usac_stub_abort(const char* f) { std::fprintf(stderr, "ref_usac: poselib::%s is outside this build's configuration\n", f); std::abort(); }
inline bool isMatRoationMat(Eigen::Matrix3d) { usac_stub_abort("isMatRoationMat"); }
inline cv::Mat getEfromRT(const cv::Mat&, const cv::Mat&) { usac_stub_abort("getEfromRT"); }
inline void getAnglesRotMat(const cv::Mat&, double&, double&, double&, bool = true) { usac_stub_abort("getAnglesRotMat"); }
int robustEssentialRefine(const cv::Mat& p1, const cv::Mat& p2, const cv::Mat& E_init, cv::Mat& E_refined, double th, unsigned iters,
                          bool makeClosestE, double* sumSqrErr_init, double* sumSqrErr, const cv::Mat& errors, const cv::Mat& mask,
                          int model, bool tryOrientedEpipolar, bool normalizeCorrs);
}  // namespace poselib
