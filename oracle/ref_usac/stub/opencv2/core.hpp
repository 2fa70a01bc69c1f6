// Typed stand-in for the few OpenCV declarations poselib/include/usac/estimators/EssentialMatEstimator.h touches (cv::Mat of
// doubles / bytes, eigen2cv / cv2eigen, noArray) so that header can be compiled UNMODIFIED where it lies under /root/reference
// (oracle/Makefile, target ref_usac).  This directory shadows oracle/cv_stub for that target only.  TEST INFRASTRUCTURE ONLY.
#pragma once
#include <cassert>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iostream>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_64F 6
#define CV_64FC1 6
#define CV_Assert(x) assert(x)

namespace cv {

class Mat {
public:
    int rows = 0, cols = 0, type_ = CV_64F;
    Mat() {}
    Mat(int r, int c, int t) { create(r, c, t); }
    void create(int r, int c, int t) {
        rows = r; cols = c; type_ = t;
        buf.reset(new std::vector<unsigned char>((size_t)r * c * esz(), 0));
    }
    static Mat zeros(int r, int c, int t) { return Mat(r, c, t); }
    static Mat eye(int r, int c, int t) { Mat m(r, c, t); for (int i = 0; i < std::min(r, c); ++i) m.at<double>(i, i) = 1.0; return m; }
    bool empty() const { return !buf || rows * cols == 0; }
    int type() const { return type_; }
    size_t esz() const { return type_ == CV_64F ? 8 : 1; }
    template <class T> T& at(int r, int c) { return reinterpret_cast<T*>(buf->data())[(size_t)r * cols + c]; }
    template <class T> const T& at(int r, int c) const { return reinterpret_cast<const T*>(buf->data())[(size_t)r * cols + c]; }
    template <class T> T& at(int i) { return reinterpret_cast<T*>(buf->data())[i]; }
    template <class T> const T& at(int i) const { return reinterpret_cast<const T*>(buf->data())[i]; }
    Mat clone() const { Mat m(rows, cols, type_); if (buf) *m.buf = *buf; return m; }
    void copyTo(Mat& d) const { d = clone(); }
private:
    std::shared_ptr<std::vector<unsigned char>> buf;
};

inline Mat noArray() { return Mat(); }

}  // namespace cv
