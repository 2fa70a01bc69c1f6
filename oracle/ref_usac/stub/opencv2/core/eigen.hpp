#pragma once
#include "../core.hpp"
#include <Eigen/Core>
namespace cv {
template <class T, int R, int C, int O, int MR, int MC>
inline void eigen2cv(const Eigen::Matrix<T, R, C, O, MR, MC>& s, Mat& d) {
    d.create((int)s.rows(), (int)s.cols(), CV_64F);
    for (int i = 0; i < (int)s.rows(); ++i) for (int j = 0; j < (int)s.cols(); ++j) d.at<double>(i, j) = (double)s(i, j);
}
template <class T, int R, int C, int O, int MR, int MC>
inline void cv2eigen(const Mat& s, Eigen::Matrix<T, R, C, O, MR, MC>& d) {
    for (int i = 0; i < s.rows; ++i) for (int j = 0; j < s.cols; ++j) d(i, j) = (T)s.at<double>(i, j);
}
}  // namespace cv
