// Minimal stand-in for <opencv2/core.hpp>: ONLY what the reference sources compiled into oracle/_ref touch -- thirdparty/
// gms-1.0/src/MatchGMS.cpp (Mat, Size, KeyPoint, DMatch, sum) and poselib/include/arrsac/*.h (RNG) -- so that those files
// can be compiled unmodified where they lie under /root/reference (oracle/Makefile, targets ref_gms, ref_arrsac).  The image
// has no OpenCV C++ SDK.  TEST INFRASTRUCTURE ONLY.  Semantics follow OpenCV's documented behaviour for these members:
// Mat::zeros / ptr / at / row (a view sharing the data) / setTo, cv::sum (sum of all elements; count of true for vector<bool>),
// RNG (multiply-with-carry, default state 0xffffffff, uniform(a, b) = a + next() % (b - a)).
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <cstring>
#include <memory>
#include <vector>

#define CV_32SC1 4

namespace cv {

struct Point2f { float x = 0, y = 0; Point2f() {} Point2f(float x_, float y_) : x(x_), y(y_) {} };
struct Size { int width = 0, height = 0; Size() {} Size(int w, int h) : width(w), height(h) {} };
struct KeyPoint { Point2f pt; float size = 0, angle = -1, response = 0; int octave = 0, class_id = -1; };
struct DMatch { int queryIdx = -1, trainIdx = -1, imgIdx = -1; float distance = 0; };
struct Scalar { double val[4] = {0, 0, 0, 0}; double operator[](int i) const { return val[i]; } };

class Mat {
public:
    int rows = 0, cols = 0;
    Mat() {}
    static Mat zeros(int r, int c, int /*type*/) { Mat m; m.rows = r; m.cols = c; m.buf.reset(new std::vector<int>((size_t)r * c, 0)); m.base = m.buf->data(); return m; }
    template <class T> T* ptr(int i) { return reinterpret_cast<T*>(base + (size_t)i * cols); }
    template <class T> const T* ptr(int i) const { return reinterpret_cast<const T*>(base + (size_t)i * cols); }
    template <class T> T& at(int r, int c) { return reinterpret_cast<T*>(base + (size_t)r * cols)[c]; }
    Mat row(int i) const { Mat m; m.rows = 1; m.cols = cols; m.buf = buf; m.base = base + (size_t)i * cols; return m; }
    Mat& setTo(int v) { for (size_t i = 0; i < (size_t)rows * cols; ++i) base[i] = v; return *this; }
    const int* data_i() const { return base; }
private:
    std::shared_ptr<std::vector<int>> buf;
    int* base = nullptr;
};

// cv::Ptr: a shared_ptr (sample_consensus_estimator.h holds its sampler / quality measurement in one)
template <class T> using Ptr = std::shared_ptr<T>;

class RNG {
public:
    unsigned long long state;
    RNG() : state(0xffffffffULL) {}
    explicit RNG(unsigned long long s) : state(s ? s : 0xffffffffULL) {}
    unsigned next() { state = (unsigned long long)(unsigned)state * 4164903690U + (unsigned)(state >> 32); return (unsigned)state; }
    int uniform(int a, int b) { return a == b ? a : (int)(next() % (unsigned)(b - a) + a); }
};

inline Scalar sum(const Mat& m) { Scalar s; const int* p = m.data_i(); for (size_t i = 0; i < (size_t)m.rows * m.cols; ++i) s.val[0] += p[i]; return s; }
inline Scalar sum(const std::vector<bool>& v) { Scalar s; for (bool b : v) s.val[0] += b ? 1 : 0; return s; }

}  // namespace cv
