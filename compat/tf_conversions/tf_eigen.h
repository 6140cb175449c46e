// Stand-in for tf_conversions/tf_eigen.h: element-wise copies between tf::Matrix3x3 and Eigen::Matrix3d.
#pragma once
#include <Eigen/Core>
#include <tf/tf.h>
namespace tf {
inline void matrixTFToEigen(const Matrix3x3& t, Eigen::Matrix3d& e) {
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) e(i, j) = t[i][j];
}
inline void matrixEigenToTF(const Eigen::Matrix3d& e, Matrix3x3& t) {
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) t[i][j] = e(i, j);
}
}  // namespace tf
