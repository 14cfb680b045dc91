// mini_geometry.h -- Translation / AngleAxis / Transform and the three decompositions of the Eigen
// stand-in.  TEST INFRASTRUCTURE ONLY; see mini_core.h.  The arithmetic is oracle/thirdparty_math.h.
#pragma once
#include "mini_matrix.h"

namespace Eigen {

template <class T, int N>
struct Translation {
  Matrix<T, N, 1> v;
  Translation(const T& x, const T& y, const T& z) { v[0] = x; v[1] = y; v[2] = z; }
};

template <class T>
struct AngleAxis {
  T angle_;
  Matrix<T, 3, 1> axis_;
  AngleAxis(const T& a, const Matrix<T, 3, 1>& ax) : angle_(a), axis_(ax) {}
  // Eigen's Rodrigues formula in Eigen's operation order (thirdparty_math.h); sin/cos = the shared
  // deterministic definition (float)sin((double)a), (float)cos((double)a)
  Matrix<T, 3, 3> toRotationMatrix() const {
    static_assert(std::is_same<T, float>::value, "mini Eigen: AngleAxis<float> only");
    const float ax[3] = {axis_[0], axis_[1], axis_[2]};
    float R[3][3];
    tpm::angle_axis_matrix_v(angle_, ax, R);
    Matrix<T, 3, 3> m;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) m(i, j) = R[i][j];
    return m;
  }
};

template <class T, int Dim, int Mode, int Options = ColMajor>
class Transform {
  static_assert(Dim == 3, "mini Eigen: 3-D transforms only");
  Matrix<T, 4, 4> m_;

 public:
  Transform() { m_.setIdentity(); }
  Matrix<T, 4, 4>& matrix() { return m_; }
  const Matrix<T, 4, 4>& matrix() const { return m_; }
  Matrix<T, 3, 3> linear() const { return Matrix<T, 3, 3>(m_.template block<3, 3>(0, 0)); }
  Matrix<T, 3, 1> translation() const { return Matrix<T, 3, 1>(m_.template block<3, 1>(0, 3)); }
  // Eigen's Affine-mode rotation() extracts U V^T of the linear part (an SVD); for the rigid matrices
  // this path feeds it (NDT_I:96-102) that equals the linear part to float rounding.  Disclosed decision
  // shared with oracle/ndt_oracle.cpp: rotation() := linear().
  Matrix<T, 3, 3> rotation() const { return linear(); }
  void setLinear(const Matrix<T, 3, 3>& l) { m_.template block<3, 3>(0, 0) = l; }
  void setTranslation(const Matrix<T, 3, 1>& t) { m_.template block<3, 1>(0, 3) = t; }
  // Transform * rotation: linear part right-multiplied, translation kept (Eigen: res.rotate(r))
  Transform operator*(const AngleAxis<T>& r) const {
    Transform out(*this);
    out.setLinear(linear() * r.toRotationMatrix());
    return out;
  }
};
template <class T>
Transform<T, 3, Affine> operator*(const Translation<T, 3>& t, const AngleAxis<T>& r) {
  Transform<T, 3, Affine> out;
  out.setLinear(r.toRotationMatrix());
  out.setTranslation(t.v);
  return out;
}
typedef Transform<float, 3, Affine> Affine3f;

// JacobiSVD<Matrix<double,6,6>>(H, FullU|FullV).solve(b): pseudo-inverse solve (thirdparty_math.h)
template <class M>
class JacobiSVD {
  M h_;

 public:
  JacobiSVD(const M& h, unsigned int = 0) : h_(h) {}
  template <class B>
  Matrix<double, 6, 1> solve(const DenseBase<B>& b) const {
    static_assert(traits<M>::Rows == 6 && traits<M>::Cols == 6 && std::is_same<typename traits<M>::Scalar, double>::value,
                  "mini Eigen: JacobiSVD is provided for Matrix<double,6,6> only");
    double H[36], bb[6], x[6];
    for (int i = 0; i < 6; ++i) {
      for (int j = 0; j < 6; ++j) H[i * 6 + j] = h_(i, j);
      bb[i] = b.coeffLin(i);
    }
    tpm::svd_solve6(H, bb, x);
    Matrix<double, 6, 1> out;
    for (int i = 0; i < 6; ++i) out[i] = x[i];
    return out;
  }
};

// SelfAdjointEigenSolver<Matrix3d>: eigenvalues ascending, eigenvectors in columns (thirdparty_math.h)
template <class M>
class SelfAdjointEigenSolver {
  Matrix<double, 3, 1> w_;
  Matrix<double, 3, 3> v_;

 public:
  SelfAdjointEigenSolver() {}
  SelfAdjointEigenSolver& compute(const M& m) {
    static_assert(traits<M>::Rows == 3 && traits<M>::Cols == 3 && std::is_same<typename traits<M>::Scalar, double>::value,
                  "mini Eigen: SelfAdjointEigenSolver is provided for Matrix3d only");
    tpm::M3 a, V;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) a.a[i][j] = m(i, j);
    double w[3];
    tpm::eigh3(a, w, V);
    for (int i = 0; i < 3; ++i) {
      w_[i] = w[i];
      for (int j = 0; j < 3; ++j) v_(i, j) = V.a[i][j];
    }
    return *this;
  }
  const Matrix<double, 3, 1>& eigenvalues() const { return w_; }
  const Matrix<double, 3, 3>& eigenvectors() const { return v_; }
};

// LLT<Matrix3d>: only getDisplayCloud (a viewer helper, never called on the path) uses it
template <class M>
class LLT {
  M l_;

 public:
  LLT() {}
  LLT& compute(const M& a) {
    const int n = a.rows();
    l_ = a;
    l_.setZero();
    for (int j = 0; j < n; ++j) {
      double s = a(j, j);
      for (int k = 0; k < j; ++k) s -= l_(j, k) * l_(j, k);
      l_(j, j) = std::sqrt(s);
      for (int i = j + 1; i < n; ++i) {
        double t = a(i, j);
        for (int k = 0; k < j; ++k) t -= l_(i, k) * l_(j, k);
        l_(i, j) = t / l_(j, j);
      }
    }
    return *this;
  }
  const M& matrixL() const { return l_; }
};

}  // namespace Eigen
