"""CPU: the C++ drop-in shim (include/matchinglib_poselib_b200.hpp) parses and type-checks against the C ABI header.
The image has no OpenCV C++ SDK, so <opencv2/core.hpp> is a declarations-only mock (tests/cv_mock)."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shim_compiles(tmp_path):
    src = tmp_path / "use_shim.cpp"
    src.write_text('#include "matchinglib_poselib_b200.hpp"\n'
                   'int f(const std::vector<cv::KeyPoint>& k, const cv::Mat& d, std::vector<cv::DMatch>& m) {\n'
                   '  return matchinglib::getMatches(k, k, d, d, cv::Size(), m, "LINEAR"); }\n'
                   'bool g(cv::OutputArray E, cv::InputArray p) { return poselib::estimateEssentialMat(E, p, p, "RANSAC", 0.001, true); }\n'
                   'int h(cv::InputArray E, cv::InputArray p, cv::OutputArray R, cv::OutputArray t, cv::OutputArray Q, cv::InputOutputArray m) {\n'
                   '  return poselib::getPoseTriangPts(E, p, p, R, t, Q, m, 130.0); }\n')
    r = subprocess.run(["g++", "-std=c++14", "-fsyntax-only", "-Wall", "-I", os.path.join(ROOT, "include"),
                        "-I", os.path.join(ROOT, "tests", "cv_mock"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
