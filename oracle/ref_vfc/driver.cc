// oracle/ref_vfc/driver.cc -- C entry over the reference's OWN VFC filter: matchinglib/source/vfcMatches.cpp (filterWithVFC) and
// source/vfc.cpp (class VFC, SparseVFC), compiled unmodified against oracle/ref_vfc/stub.  TEST INFRASTRUCTURE ONLY.
// The reference draws SparseVFC's control points with rand() (vfc.cpp:136) from whatever state the process has; here the
// caller seeds it (srand(1) is a fresh process).
#include <cstdlib>
#include <vector>
#include "matchinglib/vfcMatches.h"

extern "C" int vfc_ref_filter(const float* kpL, int nL, const float* kpR, int nR, const int* matches /* n x 2: queryIdx, trainIdx */, int n,
                              unsigned seed, unsigned char* keep, int* n_keep) {
    std::vector<cv::KeyPoint> L(nL), R(nR);
    for (int i = 0; i < nL; ++i) L[i].pt = cv::Point2f(kpL[2 * i], kpL[2 * i + 1]);
    for (int i = 0; i < nR; ++i) R[i].pt = cv::Point2f(kpR[2 * i], kpR[2 * i + 1]);
    std::vector<cv::DMatch> in(n), out;
    for (int i = 0; i < n; ++i) { in[i].queryIdx = matches[2 * i]; in[i].trainIdx = matches[2 * i + 1]; in[i].imgIdx = i; }   // imgIdx carries the position
    srand(seed);
    const int rc = filterWithVFC(L, R, in, out);
    for (int i = 0; i < n; ++i) keep[i] = 0;
    for (const cv::DMatch& m : out) keep[m.imgIdx] = 1;
    *n_keep = (int)out.size();
    return rc;
}
