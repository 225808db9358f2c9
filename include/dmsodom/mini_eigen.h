/* Minimal stand-in for the few Eigen types that appear in RGBDOdometry's public surface
 * (Utils/RGBDOdometry.h:32-153), used ONLY when <Eigen/Core> is not installed (it is not on this box).
 * With Eigen present, dmsodom/RGBDOdometry.h includes the real headers and this file is unused. */
#ifndef DMSODOM_MINI_EIGEN_H_
#define DMSODOM_MINI_EIGEN_H_
#include <cstddef>
#include <vector>

namespace Eigen {
enum { ColMajor = 0, RowMajor = 1, Dynamic = -1 };

template <class T, int R, int C, int Opt = ColMajor>
class Matrix {
 public:
  Matrix() { for (int i = 0; i < R * C; ++i) v_[i] = T(0); }
  static Matrix Zero() { return Matrix(); }
  static Matrix Identity() { Matrix m; for (int i = 0; i < (R < C ? R : C); ++i) m(i, i) = T(1); return m; }
  T& operator()(int r, int c) { return v_[(Opt & RowMajor) ? r * C + c : c * R + r]; }
  const T& operator()(int r, int c) const { return v_[(Opt & RowMajor) ? r * C + c : c * R + r]; }
  T& operator()(int i) { return v_[i]; }
  const T& operator()(int i) const { return v_[i]; }
  T* data() { return v_; }
  const T* data() const { return v_; }
  int rows() const { return R; }
  int cols() const { return C; }
 private:
  T v_[R * C];
};
typedef Matrix<float, 3, 1> Vector3f;
typedef Matrix<float, 4, 4> Matrix4f;

class MatrixXd {
 public:
  MatrixXd() : r_(0), c_(0) {}
  MatrixXd(int r, int c) : r_(r), c_(c), v_((size_t)r * c, 0.0) {}
  double& operator()(int r, int c) { return v_[(size_t)c * r_ + r]; }
  const double& operator()(int r, int c) const { return v_[(size_t)c * r_ + r]; }
  int rows() const { return r_; }
  int cols() const { return c_; }
 private:
  int r_, c_;
  std::vector<double> v_;
};
}  // namespace Eigen
#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW
#endif
