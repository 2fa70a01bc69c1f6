// oracle/ref_gms/driver.cc -- C entry point around the reference's OWN MatchGMS class
// (matchinglib/thirdparty/gms-1.0/src/MatchGMS.cpp, compiled unmodified from /root/reference against the declarations in
// cv_stub/).  What it restates is only the 10-line wrapper filterMatchesGMS (matchinglib/source/gms.cpp:54-60).
// TEST INFRASTRUCTURE ONLY.
#include "MatchGMS.h"
#include <cstdint>

extern "C" int ref_gms_filter(const float* kp1_xy, int n1, int w1, int h1, const float* kp2_xy, int n2, int w2, int h2,
                              const int32_t* matches_qt /* n x 2 */, int n_matches, int use_scale, int use_rotation, uint8_t* mask)
{
    std::vector<cv::KeyPoint> k1(n1), k2(n2);
    for (int i = 0; i < n1; ++i) { k1[i].pt.x = kp1_xy[2 * i]; k1[i].pt.y = kp1_xy[2 * i + 1]; }
    for (int i = 0; i < n2; ++i) { k2[i].pt.x = kp2_xy[2 * i]; k2[i].pt.y = kp2_xy[2 * i + 1]; }
    std::vector<cv::DMatch> m(n_matches);
    for (int i = 0; i < n_matches; ++i) { m[i].queryIdx = matches_qt[2 * i]; m[i].trainIdx = matches_qt[2 * i + 1]; }
    std::vector<bool> inl;
    MatchGMS gms(k1, cv::Size(w1, h1), k2, cv::Size(w2, h2), m);            // gms.cpp:59
    const int n = gms.getInlierMask(inl, use_scale != 0, use_rotation != 0); // gms.cpp:60
    for (int i = 0; i < n_matches; ++i) mask[i] = (i < (int)inl.size() && inl[i]) ? 1 : 0;
    return n;
}
