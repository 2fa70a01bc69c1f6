// Stand-in for the OpenCV declarations matchinglib/source/vfc.cpp and vfcMatches.cpp touch (a float cv::Mat with create / ptr /
// at / copyTo / t / * / -, cv::solve(DECOMP_LU), cv::eigen, Point2f, KeyPoint, DMatch), so those two files compile UNMODIFIED
// where they lie under /root/reference (oracle/Makefile, target ref_vfc).  TEST INFRASTRUCTURE ONLY.
// cv::solve follows OpenCV's LU for small float systems (modules/core/src/matrix_decomp.cpp, LUImpl: partial pivoting by largest
// |a|, eps = 10 * FLT_EPSILON, row updates with alpha = a[j][i] * (-1 / a[i][i]), back substitution) -- the same arithmetic order
// csrc/vfc.cu uses.  cv::eigen is only reached from FastVFC, which the reference never selects (vfc.cpp:41): it aborts.
#pragma once
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <memory>
#include <vector>

#define CV_32FC1 5
namespace cv {
enum { DECOMP_LU = 0 };
struct Point2f { float x = 0, y = 0; Point2f() {} Point2f(float x_, float y_) : x(x_), y(y_) {} };
struct KeyPoint { Point2f pt; float size = 0, angle = -1, response = 0; int octave = 0, class_id = -1; };
struct DMatch { int queryIdx = -1, trainIdx = -1, imgIdx = -1; float distance = 0; };

class Mat {
public:
    int rows = 0, cols = 0;
    Mat() {}
    Mat(int r, int c, int) { create(r, c, CV_32FC1); }
    void create(int r, int c, int) { rows = r; cols = c; buf.reset(new std::vector<float>((size_t)r * c, 0.0f)); }
    template <class T> T* ptr(int i) { return buf->data() + (size_t)i * cols; }
    template <class T> const T* ptr(int i) const { return buf->data() + (size_t)i * cols; }
    template <class T> T& at(int i, int j) { return (*buf)[(size_t)i * cols + j]; }
    template <class T> const T& at(int i, int j) const { return (*buf)[(size_t)i * cols + j]; }
    void copyTo(Mat& d) const { d.create(rows, cols, CV_32FC1); if (buf) *d.buf = *buf; }
    Mat t() const { Mat m(cols, rows, CV_32FC1); for (int i = 0; i < rows; ++i) for (int j = 0; j < cols; ++j) m.at<float>(j, i) = at<float>(i, j); return m; }
private:
    std::shared_ptr<std::vector<float>> buf;
};
inline Mat operator*(const Mat& a, const Mat& b) {
    Mat m(a.rows, b.cols, CV_32FC1);
    for (int i = 0; i < a.rows; ++i) for (int j = 0; j < b.cols; ++j) { float s = 0; for (int k = 0; k < a.cols; ++k) s += a.at<float>(i, k) * b.at<float>(k, j); m.at<float>(i, j) = s; }
    return m;
}
inline Mat operator-(const Mat& a, const Mat& b) { Mat m; a.copyTo(m); for (int i = 0; i < a.rows; ++i) for (int j = 0; j < a.cols; ++j) m.at<float>(i, j) -= b.at<float>(i, j); return m; }
inline Mat operator/(const Mat& a, float d) { Mat m; a.copyTo(m); for (int i = 0; i < a.rows; ++i) for (int j = 0; j < a.cols; ++j) m.at<float>(i, j) /= d; return m; }
inline bool eigen(const Mat&, Mat&, Mat&) { std::fprintf(stderr, "ref_vfc: cv::eigen is outside this build (FastVFC)\n"); std::abort(); }

inline bool solve(const Mat& A_, const Mat& B_, Mat& X, int) {
    Mat A, b; A_.copyTo(A); B_.copyTo(b);
    const int m = A.rows, n = b.cols; const float eps = FLT_EPSILON * 10;
    for (int i = 0; i < m; i++) {
        int k = i;
        for (int j = i + 1; j < m; j++) if (std::abs(A.at<float>(j, i)) > std::abs(A.at<float>(k, i))) k = j;
        if (std::abs(A.at<float>(k, i)) < eps) { X.create(m, n, CV_32FC1); return false; }
        if (k != i) { for (int j = i; j < m; j++) std::swap(A.at<float>(i, j), A.at<float>(k, j)); for (int j = 0; j < n; j++) std::swap(b.at<float>(i, j), b.at<float>(k, j)); }
        const float d = -1 / A.at<float>(i, i);
        for (int j = i + 1; j < m; j++) {
            const float alpha = A.at<float>(j, i) * d;
            for (int kk = i + 1; kk < m; kk++) A.at<float>(j, kk) += alpha * A.at<float>(i, kk);
            for (int kk = 0; kk < n; kk++) b.at<float>(j, kk) += alpha * b.at<float>(i, kk);
        }
    }
    for (int i = m - 1; i >= 0; i--)
        for (int j = 0; j < n; j++) {
            float s = b.at<float>(i, j);
            for (int k = i + 1; k < m; k++) s -= A.at<float>(i, k) * b.at<float>(k, j);
            b.at<float>(i, j) = s / A.at<float>(i, i);
        }
    b.copyTo(X);
    return true;
}
}  // namespace cv
