// Minimal stand-in for <opencv2/core.hpp>: ONLY for syntax-checking include/matchinglib_poselib_b200.hpp on an image
// without OpenCV C++ headers (tests/test_shim_syntax.py).  Declarations only; nothing here is linked or run.
#pragma once
#include <cstddef>
#include <stdexcept>
#include <string>
#include <vector>
typedef unsigned char uchar;
#define CV_8U 0
#define CV_32F 5
#define CV_64F 6
#define CV_64FC1 6
#define CV_Assert(expr) do { if (!(expr)) throw std::runtime_error(#expr); } while (0)
#define CV_Error(code, msg) throw std::runtime_error(msg)
namespace cv {
namespace Error { enum { StsAssert = -215, StsError = -2 }; }
struct Size { int width = 0, height = 0; Size() {} Size(int w, int h) : width(w), height(h) {} };
struct Point2f { float x, y; };
struct KeyPoint { Point2f pt; float size, angle, response; int octave, class_id; };
struct DMatch { int queryIdx, trainIdx, imgIdx; float distance; };
class Mat;
class _OutputArray;
class Mat {
public:
    int rows = 0, cols = 0; uchar* data = nullptr;
    struct Step { size_t v[2]; size_t operator[](int i) const { return v[i]; } } step;
    Mat(); Mat(int rows, int cols, int type, void* data);
    int type() const; int depth() const; Mat clone() const; int checkVector(int ch) const;
    void convertTo(Mat& dst, int type) const; Mat reshape(int cn, int rows) const;
    template <class T> T* ptr(int r = 0);
    template <class T> const T* ptr(int r = 0) const;
    bool isContinuous() const;
    void copyTo(const _OutputArray& dst) const;
    void copyTo(Mat& dst) const;
    bool empty() const; size_t total() const;
};
class _InputArray { public: Mat getMat() const; };
class _OutputArray : public _InputArray { public: bool needed() const; bool empty() const; void clear() const; };
class _InputOutputArray : public _OutputArray {};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;
typedef const _InputOutputArray& InputOutputArray;
const _InputOutputArray& noArray();
// <opencv2/imgproc.hpp>: what estimateSprtDeltaInit uses
void convexHull(const std::vector<Point2f>& points, std::vector<Point2f>& hull);
double contourArea(const std::vector<Point2f>& contour);
}
