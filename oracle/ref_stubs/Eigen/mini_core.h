// mini_core.h -- stand-in for the subset of Eigen 3.3 that the reference's NDT path uses
// (/root/reference/xchu_mapping/include/pclomp/{ndt_omp.h, ndt_omp_impl.hpp,
// voxel_grid_covariance_omp.h, voxel_grid_covariance_omp_impl.hpp}).
//
// TEST INFRASTRUCTURE ONLY.  Eigen is not in this image; this header lets the reference's own
// NDT sources compile UNMODIFIED into oracle/_ref/libndt_ref.so.  It is not Eigen and not a copy
// of it: every operation is evaluated EAGERLY, coefficient-wise, inner index ascending, with no
// vectorisation -- the same disclosed decisions as oracle/ndt_oracle.cpp (see its header), and
// the decompositions are the shared restatements of oracle/thirdparty_math.h.
#pragma once
#include <cassert>
#include <cmath>
#include <cstddef>
#include <cstring>
#include <memory>
#include <type_traits>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../../thirdparty_math.h"

#define EIGEN_MAKE_ALIGNED_OPERATOR_NEW

namespace Eigen {

const int Dynamic = -1;
enum { ColMajor = 0, RowMajor = 1 };
enum TransformTraits { Isometry = 1, Affine = 2, AffineCompact = 0x12, Projective = 0x20 };
enum { ComputeFullU = 0x04, ComputeThinU = 0x08, ComputeFullV = 0x10, ComputeThinV = 0x20 };

template <class T> using aligned_allocator = std::allocator<T>;

template <class D> struct traits;
template <class T, int R, int C> class Matrix;
template <class T, int R, int C> class Array;
template <class M, int R, int C> class Block;
template <class D> class CommaInitializer;

template <class S> struct is_scalar : std::integral_constant<bool, std::is_arithmetic<S>::value> {};

// ---------------------------------------------------------------------------------------------
// DenseBase: everything a Matrix, an Array and a Block have in common.  Derived provides
// rows(), cols(), coeff(i, j) and coeffRef(i, j).
// ---------------------------------------------------------------------------------------------
template <class D>
class DenseBase {
 public:
  typedef typename traits<D>::Scalar Scalar;
  enum { RowsAtCompileTime = traits<D>::Rows, ColsAtCompileTime = traits<D>::Cols };
  enum { VecSize = (RowsAtCompileTime == 1 ? ColsAtCompileTime : RowsAtCompileTime) };
  typedef Matrix<Scalar, RowsAtCompileTime, ColsAtCompileTime> PlainObject;

  const D& derived() const { return *static_cast<const D*>(this); }
  D& derived() { return *static_cast<D*>(this); }
  int rows() const { return derived().rows(); }
  int cols() const { return derived().cols(); }
  int size() const { return rows() * cols(); }

  // linear (vector) access: a column or a row vector, else column-major
  Scalar coeffLin(int k) const {
    if (cols() == 1) return derived().coeff(k, 0);
    if (rows() == 1) return derived().coeff(0, k);
    return derived().coeff(k % rows(), k / rows());
  }
  Scalar& coeffRefLin(int k) {
    if (cols() == 1) return derived().coeffRef(k, 0);
    if (rows() == 1) return derived().coeffRef(0, k);
    return derived().coeffRef(k % rows(), k / rows());
  }
  Scalar operator()(int i, int j) const { return derived().coeff(i, j); }
  Scalar& operator()(int i, int j) { return derived().coeffRef(i, j); }
  Scalar operator()(int k) const { return coeffLin(k); }
  Scalar& operator()(int k) { return coeffRefLin(k); }
  Scalar operator[](int k) const { return coeffLin(k); }
  Scalar& operator[](int k) { return coeffRefLin(k); }

  PlainObject eval() const { return PlainObject(*this); }
  D& noalias() { return derived(); }

  D& setZero() {
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i) derived().coeffRef(i, j) = Scalar(0);
    return derived();
  }
  D& setOnes() {
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i) derived().coeffRef(i, j) = Scalar(1);
    return derived();
  }
  D& setIdentity() {
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i) derived().coeffRef(i, j) = Scalar(i == j ? 1 : 0);
    return derived();
  }

  Matrix<Scalar, ColsAtCompileTime, RowsAtCompileTime> transpose() const {
    Matrix<Scalar, ColsAtCompileTime, RowsAtCompileTime> r;
    r.resize(cols(), rows());
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i) r.coeffRef(j, i) = derived().coeff(i, j);
    return r;
  }
  template <class U>
  Matrix<U, RowsAtCompileTime, ColsAtCompileTime> cast() const {
    Matrix<U, RowsAtCompileTime, ColsAtCompileTime> r;
    r.resize(rows(), cols());
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i) r.coeffRef(i, j) = static_cast<U>(derived().coeff(i, j));
    return r;
  }
  Array<Scalar, RowsAtCompileTime, ColsAtCompileTime> array() const {
    return Array<Scalar, RowsAtCompileTime, ColsAtCompileTime>(*this);
  }
  PlainObject matrix() const { return eval(); }

  // reductions: sequential, index ascending
  template <class O>
  Scalar dot(const DenseBase<O>& o) const {
    const int n = size();
    assert(n == o.size());
    Scalar s = coeffLin(0) * o.coeffLin(0);
    for (int k = 1; k < n; ++k) s += coeffLin(k) * o.coeffLin(k);
    return s;
  }
  Scalar squaredNorm() const {
    const int n = size();
    Scalar s = coeffLin(0) * coeffLin(0);
    for (int k = 1; k < n; ++k) s += coeffLin(k) * coeffLin(k);
    return s;
  }
  Scalar norm() const { return std::sqrt(squaredNorm()); }
  Scalar sum() const {
    const int n = size();
    if (n == 0) return Scalar(0);
    Scalar s = coeffLin(0);
    for (int k = 1; k < n; ++k) s += coeffLin(k);
    return s;
  }
  Scalar mean() const { return sum() / Scalar(size()); }   // Eigen: sum() / size()
  void normalize() {
    const Scalar z = squaredNorm();
    if (z > Scalar(0)) *this /= std::sqrt(z);
  }
  Scalar maxCoeff() const {
    Scalar m = coeffLin(0);
    for (int k = 1; k < size(); ++k) { const Scalar v = coeffLin(k); if (v > m) m = v; }
    return m;
  }
  Scalar minCoeff() const {
    Scalar m = coeffLin(0);
    for (int k = 1; k < size(); ++k) { const Scalar v = coeffLin(k); if (v < m) m = v; }
    return m;
  }

  template <class O>
  D& operator+=(const DenseBase<O>& o) {
    if (o.rows() == rows() && o.cols() == cols()) {
      for (int j = 0; j < cols(); ++j)
        for (int i = 0; i < rows(); ++i) derived().coeffRef(i, j) += o.derived().coeff(i, j);
    } else {
      assert(size() == o.size());
      for (int k = 0; k < size(); ++k) coeffRefLin(k) += o.coeffLin(k);
    }
    return derived();
  }
  template <class O>
  D& operator-=(const DenseBase<O>& o) {
    if (o.rows() == rows() && o.cols() == cols()) {
      for (int j = 0; j < cols(); ++j)
        for (int i = 0; i < rows(); ++i) derived().coeffRef(i, j) -= o.derived().coeff(i, j);
    } else {
      assert(size() == o.size());
      for (int k = 0; k < size(); ++k) coeffRefLin(k) -= o.coeffLin(k);
    }
    return derived();
  }
  template <class S>
  typename std::enable_if<is_scalar<S>::value, D&>::type operator*=(const S& s) {
    const Scalar f = static_cast<Scalar>(s);
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i) derived().coeffRef(i, j) *= f;
    return derived();
  }
  template <class S>
  typename std::enable_if<is_scalar<S>::value, D&>::type operator/=(const S& s) {
    const Scalar f = static_cast<Scalar>(s);
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i) derived().coeffRef(i, j) /= f;
    return derived();
  }

  template <class O>
  bool operator==(const DenseBase<O>& o) const {
    if (rows() != o.rows() || cols() != o.cols()) return false;
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i)
        if (!(derived().coeff(i, j) == o.derived().coeff(i, j))) return false;
    return true;
  }
  template <class O>
  bool operator!=(const DenseBase<O>& o) const { return !(*this == o); }

  // sub-blocks (views)
  template <int BR, int BC> Block<D, BR, BC> block(int i, int j) { return Block<D, BR, BC>(derived(), i, j, BR, BC); }
  template <int BR, int BC> Block<D, BR, BC> block(int i, int j) const {
    return Block<D, BR, BC>(const_cast<D&>(derived()), i, j, BR, BC);
  }
  Block<D, Dynamic, Dynamic> block(int i, int j, int r, int c) const {
    return Block<D, Dynamic, Dynamic>(const_cast<D&>(derived()), i, j, r, c);
  }
  Block<D, Dynamic, Dynamic> topLeftCorner(int r, int c) const { return block(0, 0, r, c); }
  Block<D, RowsAtCompileTime, 1> col(int j) const {
    return Block<D, RowsAtCompileTime, 1>(const_cast<D&>(derived()), 0, j, rows(), 1);
  }
  Block<D, 1, ColsAtCompileTime> row(int i) const {
    return Block<D, 1, ColsAtCompileTime>(const_cast<D&>(derived()), i, 0, 1, cols());
  }
  template <int N> Block<D, (ColsAtCompileTime == 1 ? N : 1), (ColsAtCompileTime == 1 ? 1 : N)> head() const {
    typedef Block<D, (ColsAtCompileTime == 1 ? N : 1), (ColsAtCompileTime == 1 ? 1 : N)> B;
    return cols() == 1 ? B(const_cast<D&>(derived()), 0, 0, N, 1) : B(const_cast<D&>(derived()), 0, 0, 1, N);
  }
  Block<D, Dynamic, Dynamic> head(int n) const { return cols() == 1 ? block(0, 0, n, 1) : block(0, 0, 1, n); }

  Matrix<Scalar, (RowsAtCompileTime < ColsAtCompileTime ? RowsAtCompileTime : ColsAtCompileTime), 1> diagonal() const {
    Matrix<Scalar, (RowsAtCompileTime < ColsAtCompileTime ? RowsAtCompileTime : ColsAtCompileTime), 1> d;
    const int n = rows() < cols() ? rows() : cols();
    d.resize(n, 1);
    for (int i = 0; i < n; ++i) d.coeffRef(i, 0) = derived().coeff(i, i);
    return d;
  }
  Matrix<Scalar, VecSize, VecSize> asDiagonal() const {
    Matrix<Scalar, VecSize, VecSize> m;
    const int n = size();
    m.resize(n, n);
    m.setZero();
    for (int i = 0; i < n; ++i) m.coeffRef(i, i) = coeffLin(i);
    return m;
  }

  CommaInitializer<D> operator<<(const Scalar& s);
  template <class O> CommaInitializer<D> operator<<(const DenseBase<O>& o);

  // 3x3 double only: Eigen's cofactor inverse (thirdparty_math.h inverse3)
  PlainObject inverse() const {
    static_assert(RowsAtCompileTime == 3 && ColsAtCompileTime == 3 && std::is_same<Scalar, double>::value,
                  "mini Eigen: inverse() is provided for Matrix3d only");
    tpm::M3 m;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) m.a[i][j] = derived().coeff(i, j);
    const tpm::M3 r = tpm::inverse3(m);
    PlainObject out;
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j) out.coeffRef(i, j) = r.a[i][j];
    return out;
  }
  // Matrix3f only, axes (0, 1, 2): thirdparty_math.h euler_012
  Matrix<Scalar, 3, 1> eulerAngles(int a0, int a1, int a2) const {
    static_assert(RowsAtCompileTime == 3 && ColsAtCompileTime == 3 && std::is_same<Scalar, float>::value,
                  "mini Eigen: eulerAngles() is provided for Matrix3f only");
    assert(a0 == 0 && a1 == 1 && a2 == 2);
    (void)a0; (void)a1; (void)a2;
    float T[16] = {0};
    for (int c = 0; c < 3; ++c)
      for (int r = 0; r < 3; ++r) T[c * 4 + r] = derived().coeff(r, c);
    float e[3];
    tpm::euler_012(T, e);
    Matrix<Scalar, 3, 1> out;
    out.coeffRef(0, 0) = e[0]; out.coeffRef(1, 0) = e[1]; out.coeffRef(2, 0) = e[2];
    return out;
  }
};

}  // namespace Eigen
