// oracle/ref_usac/driver.cc -- C entry over the reference's OWN USAC essential-matrix estimator.  TEST INFRASTRUCTURE ONLY.
//
// Compiled by oracle/Makefile (target ref_usac) together with, where they lie under /root/reference:
//   poselib/include/usac/estimators/{USAC.h,EssentialMatEstimator.h}      the schedule and the estimator, unmodified
//   poselib/source/usac/utils/{FundmatrixFunctions,MathFunctions,HomographyFunctions,PoseFunctions}.cpp
//   poselib/thirdparty/opengv (fivept_nister, fivept_stewenius, eightpt, ...) against the Eigen 3.2 vendored beside it
// and against oracle/ref_usac/stub (a typed cv::Mat, eigen2cv/cv2eigen, six poselib:: helpers).  This file restates only the
// CONFIGURATION that estimateEssentialMatUsac fills (poselib/source/usac/usac_estimations.cpp:283-475) -- that translation unit
// also pulls in the F / H / R estimators and cannot be compiled here -- with the two differences a checker needs: the seed is
// an argument (the reference calls srand(time(nullptr)), :328) and mS is an argument (the reference keeps a running
// models-per-sample average in function statics, :392-409).
//
// Configurations served: estimator NISTER or STEWENIUS, refinement REFINE_WEIGHTS (8-point with Torr weights, the 6th cfgUSAC
// digit's default 0), no degeneracy test (4th digit 0).  REFINE_8PT_PSEUDOHUBER needs poselib::robustEssentialRefine, which
// lives in pose_estim.cpp with real OpenCV: the stub aborts if it is reached.
#include <stdlib.h>
#include <math.h>
#include "usac/estimators/EssentialMatEstimator.h"

// weightingEssential.cpp needs the SBA bundle-adjustment driver; its functions are only reached from the *_WEIGHTS refinements
// of the Stewenius / Nister kind, which this build never configures.
opengv::complexEssentials_t fivept_stewenius_weight(const opengv::relative_pose::CentralRelativeWeightingAdapter&) { poselib::usac_stub_abort("fivept_stewenius_weight"); }
opengv::essentials_t fivept_nister_weight(const opengv::relative_pose::CentralRelativeWeightingAdapter&) { poselib::usac_stub_abort("fivept_nister_weight"); }
double computePseudoHuberWeight(const opengv::bearingVector_t&, const opengv::bearingVector_t&, const opengv::essential_t&, const double&) { poselib::usac_stub_abort("computePseudoHuberWeight"); }
// LAPACK's dgeqp3 is named by FTools::nullspaceQR7x9 (the 7-point F solver, FundmatrixFunctions.cpp:187), never on the E path.
extern "C" int dgeqp3_(const int*, const int*, double*, const int*, int*, double*, double*, int*, int*) { poselib::usac_stub_abort("dgeqp3_"); }
namespace poselib {
int robustEssentialRefine(const cv::Mat&, const cv::Mat&, const cv::Mat&, cv::Mat&, double, unsigned, bool, double*, double*, const cv::Mat&,
                          const cv::Mat&, int, bool, bool) { usac_stub_abort("robustEssentialRefine"); }
}

extern "C" {

// p1, p2: n x 2 camera-normalised points.  sorted_idx: n indices by increasing match distance (PROSAC) or NULL (uniform).
// estimator: 0 NISTER, 2 STEWENIUS (USACConfig::EssentialMatEstimatorUsed).  mS <= 0: the reference's first-call value.
// stats[8]: hyp_count, model_count, rejected_sample_count, rejected_model_count, total_points_verified, num_local_optimizations,
// best_inlier_count, solve() ok.  sprt_out[2]: the last test's delta, epsilon (usac_results_).  Returns 0 / 1 like the reference.
int usac_ref_essential(const double* p1, const double* p2, int n, double th, double focal_length, double th_pixels,
                       double prosac_beta, double sprt_delta, double sprt_epsilon, double mS, int estimator,
                       const unsigned* sorted_idx, unsigned seed, double* E_out, unsigned char* mask, unsigned* stats,
                       double* sprt_out) {
    srand(seed);
    std::vector<double> pointData(6 * (size_t)n);                                  // usac_estimations.cpp:331-341
    for (int i = 0; i < n; ++i) {
        pointData[6 * i] = p1[2 * i]; pointData[6 * i + 1] = p1[2 * i + 1]; pointData[6 * i + 2] = 1.0;
        pointData[6 * i + 3] = p2[2 * i]; pointData[6 * i + 4] = p2[2 * i + 1]; pointData[6 * i + 5] = 1.0;
    }
    USACConfig::Common c_com; USACConfig::Losac c_lo; USACConfig::Prosac c_pro; USACConfig::Sprt c_sprt; USACConfig::EssMat c_ess;
    const USACConfig::EssentialMatEstimatorUsed est = estimator == 2 ? USACConfig::ESTIM_STEWENIUS : USACConfig::ESTIM_NISTER;
    c_com.confThreshold = 0.99; c_com.minSampleSize = 5; c_com.inlierThreshold = th; c_com.maxHypotheses = 50000;   // :351-355
    c_com.maxSolutionsPerSample = 10; c_com.numDataPoints = n;                                                       // :359-360
    c_com.prevalidateSample = true; c_com.prevalidateModel = true; c_com.testDegeneracy = false; c_com.testDegeneracyLOSAC = false;
    c_com.randomSamplingMethod = sorted_idx ? USACConfig::SAMP_PROSAC : USACConfig::SAMP_UNIFORM;                  // :372-375
    c_com.verifMethod = USACConfig::VERIF_SPRT; c_com.localOptMethod = USACConfig::LO_LOSAC;                        // :376-377
    c_lo.innerRansacRepetitions = 5; c_lo.innerSampleSize = 14; c_lo.thresholdMultiplier = 2.0; c_lo.numStepsIterative = 4;   // :380-386
    std::vector<unsigned> sorted;
    if (sorted_idx) {                                                                                               // :389-396
        sorted.assign(sorted_idx, sorted_idx + n);
        c_pro.beta = prosac_beta; c_pro.maxSamples = 1000; c_pro.minStopLen = 20; c_pro.nonRandConf = 0.99;
        c_pro.sortedPointIndices = sorted.data();
    }
    c_sprt.delta = sprt_delta; c_sprt.epsilon = sprt_epsilon;                                                      // :399-400
    c_sprt.mS = mS > 0 ? mS : (est == USACConfig::ESTIM_NISTER ? 8.5 : 6.0);                                        // :401-416
    c_sprt.tM = est == USACConfig::ESTIM_STEWENIUS ? 2736.0 : 2314.0;                                               // :417-440
    c_ess.focalLength = focal_length; c_ess.hDegenThreshold = 2.0 * th; c_ess.maxUpgradeSamples = 8000;             // :444-446
    c_ess.refineMethod = USACConfig::REFINE_WEIGHTS; c_ess.rotDegenThesholdPix = th_pixels; c_ess.used_estimator = est;
    c_ess.ransacLikeUpgradeDegenerate = true; c_ess.enableHDegen = true; c_ess.enableUpgradeDegenPose = true;       // :450-456
    ConfigParamsEssential cfg(c_com, c_pro, c_sprt, c_lo, c_ess, getenv("USAC_REF_VERBOSE") != nullptr);     // verbose: the reference's own progress lines

    std::unique_ptr<EssentialMatEstimator> fund(new EssentialMatEstimator);                                         // :461-470
    fund->initParamsUSAC(cfg);
    fund->initDataUSAC(cfg);
    fund->initProblem(cfg, pointData.data());
    const bool ok = fund->solve();
    const UsacResults& r = fund->usac_results_;
    stats[0] = r.hyp_count_; stats[1] = r.model_count_; stats[2] = r.rejected_sample_count_; stats[3] = r.rejected_model_count_;
    stats[4] = r.total_points_verified_; stats[5] = r.num_local_optimizations_; stats[6] = r.best_inlier_count_; stats[7] = ok;
    sprt_out[0] = r.sprt_delta_; sprt_out[1] = r.sprt_epsilon_;
    if (!ok) return 1;
    for (int i = 0; i < 9; ++i) E_out[i] = fund->final_model_params_.empty() ? 0.0 : fund->final_model_params_[i];
    for (int i = 0; i < n; ++i) mask[i] = r.inlier_flags_[i] ? 1 : 0;
    return 0;
}

// FTools::singulF (FundmatrixFunctions.cpp:334-360) on a caller's 3 x 3 matrix: lets tests/test_oracle_usac.py pin the defect that
// makes the reference's REFINE_WEIGHTS local optimisation a no-op (its ccmath svduv port returns wrong singular triplets when two
// singular values are nearly equal, i.e. for essential-like matrices).
void usac_ref_singulF(double* F) { FTools::singulF(F); }

}  // extern "C"
