// mini_matrix.h -- Matrix / Array / Block / comma initialiser / free operators of the Eigen stand-in.
// TEST INFRASTRUCTURE ONLY; see mini_core.h.
#pragma once
#include "mini_core.h"

namespace Eigen {

template <class T, int R, int C> struct traits<Matrix<T, R, C>> { typedef T Scalar; enum { Rows = R, Cols = C }; };
template <class T, int R, int C> struct traits<Array<T, R, C>> { typedef T Scalar; enum { Rows = R, Cols = C }; };
template <class M, int R, int C> struct traits<Block<M, R, C>> { typedef typename traits<M>::Scalar Scalar; enum { Rows = R, Cols = C }; };

// column-major storage, fixed or dynamic
template <class T, int R, int C, bool Dyn = (R == Dynamic || C == Dynamic)>
struct DenseStorage {
  T d[R * C];
  DenseStorage() { for (int k = 0; k < R * C; ++k) d[k] = T(0); }
  int rows() const { return R; }
  int cols() const { return C; }
  void resize(int r, int c) { assert(r == R && c == C); (void)r; (void)c; }
  T* data() { return d; }
  const T* data() const { return d; }
};
template <class T, int R, int C>
struct DenseStorage<T, R, C, true> {
  std::vector<T> d;
  int r_, c_;
  DenseStorage() : r_(R == Dynamic ? 0 : R), c_(C == Dynamic ? 0 : C) {}
  int rows() const { return r_; }
  int cols() const { return c_; }
  void resize(int r, int c) {
    if (r == r_ && c == c_) return;
    r_ = r; c_ = c;
    d.assign((size_t)r * c, T(0));
  }
  T* data() { return d.data(); }
  const T* data() const { return d.data(); }
};

template <class Derived, class T, int R, int C>
class PlainBase : public DenseBase<Derived> {
 protected:
  DenseStorage<T, R, C> st_;

 public:
  typedef T Scalar;
  int rows() const { return st_.rows(); }
  int cols() const { return st_.cols(); }
  T coeff(int i, int j) const { return st_.data()[(size_t)j * st_.rows() + i]; }
  T& coeffRef(int i, int j) { return st_.data()[(size_t)j * st_.rows() + i]; }
  T* data() { return st_.data(); }
  const T* data() const { return st_.data(); }
  void resize(int r, int c) { st_.resize(r, c); }
  void resize(int n) {
    if (C == 1) st_.resize(n, 1);
    else if (R == 1) st_.resize(1, n);
    else assert(false && "resize(n) on a matrix");
  }
  // copy with Eigen's implicit vector transposition (1xN <-> Nx1)
  template <class O>
  void assignFrom(const DenseBase<O>& o) {
    const bool fixed = (R != Dynamic && C != Dynamic);
    if (fixed && (o.rows() != R || o.cols() != C)) {
      assert((R == 1 || C == 1) && o.size() == R * C && "size mismatch");
      for (int k = 0; k < R * C; ++k) st_.data()[k] = static_cast<T>(o.coeffLin(k));
      return;
    }
    if (!fixed) {
      if ((R == 1 && o.cols() == 1) || (C == 1 && o.rows() == 1 && o.cols() != 1)) {   // vector transposition
        resize(o.size());
        for (int k = 0; k < o.size(); ++k) st_.data()[k] = static_cast<T>(o.coeffLin(k));
        return;
      }
      st_.resize(o.rows(), o.cols());
    }
    for (int j = 0; j < cols(); ++j)
      for (int i = 0; i < rows(); ++i) coeffRef(i, j) = static_cast<T>(o.derived().coeff(i, j));
  }
};

template <class T, int R, int C>
class Matrix : public PlainBase<Matrix<T, R, C>, T, R, C> {
  typedef PlainBase<Matrix<T, R, C>, T, R, C> Base;

 public:
  Matrix() {}
  Matrix(const Matrix& o) = default;
  Matrix& operator=(const Matrix& o) = default;
  template <class O> Matrix(const DenseBase<O>& o) { this->assignFrom(o); }
  template <class O> Matrix& operator=(const DenseBase<O>& o) {
    Matrix tmp; tmp.assignFrom(o);   // evaluate before writing (aliasing-safe)
    *static_cast<Base*>(this) = tmp;
    return *this;
  }
  // VectorXf(n)
  template <class A, class = typename std::enable_if<std::is_integral<A>::value && (R == Dynamic || C == Dynamic)>::type>
  explicit Matrix(A n) { this->resize((int)n); }
  // MatrixXi(rows, cols)  /  Vector2(x, y)
  template <class A, class B, class = typename std::enable_if<is_scalar<A>::value && is_scalar<B>::value>::type>
  Matrix(const A& a, const B& b) {
    if (R == Dynamic || C == Dynamic) this->resize((int)a, (int)b);
    else { assert(R * C == 2); this->data()[0] = static_cast<T>(a); this->data()[1] = static_cast<T>(b); }
  }
  template <class A, class B, class D3, class = typename std::enable_if<is_scalar<A>::value && is_scalar<B>::value && is_scalar<D3>::value>::type>
  Matrix(const A& a, const B& b, const D3& c) {
    static_assert(R * C == 3, "3-coefficient constructor on a non-3-vector");
    this->data()[0] = static_cast<T>(a); this->data()[1] = static_cast<T>(b); this->data()[2] = static_cast<T>(c);
  }
  template <class A, class B, class D3, class E, class = typename std::enable_if<is_scalar<A>::value && is_scalar<B>::value && is_scalar<D3>::value && is_scalar<E>::value>::type>
  Matrix(const A& a, const B& b, const D3& c, const E& d) {
    static_assert(R * C == 4, "4-coefficient constructor on a non-4-vector");
    this->data()[0] = static_cast<T>(a); this->data()[1] = static_cast<T>(b); this->data()[2] = static_cast<T>(c); this->data()[3] = static_cast<T>(d);
  }

  static Matrix Zero() { Matrix m; m.setZero(); return m; }
  static Matrix Zero(int n) { Matrix m; m.resize(n); m.setZero(); return m; }
  static Matrix Zero(int r, int c) { Matrix m; m.resize(r, c); m.setZero(); return m; }
  static Matrix Ones() { Matrix m; m.setOnes(); return m; }
  static Matrix Ones(int r, int c) { Matrix m; m.resize(r, c); m.setOnes(); return m; }
  static Matrix Identity() { Matrix m; m.setIdentity(); return m; }
  static Matrix Constant(const T& v) { Matrix m; for (int k = 0; k < m.size(); ++k) m.data()[k] = v; return m; }
  static Matrix Unit(int k) { Matrix m; m.setZero(); m.data()[k] = T(1); return m; }
  static Matrix UnitX() { return Unit(0); }
  static Matrix UnitY() { return Unit(1); }
  static Matrix UnitZ() { return Unit(2); }
};

template <class T, int R, int C>
class Array : public PlainBase<Array<T, R, C>, T, R, C> {
  typedef PlainBase<Array<T, R, C>, T, R, C> Base;

 public:
  Array() {}
  Array(const Array&) = default;
  Array& operator=(const Array&) = default;
  template <class O> Array(const DenseBase<O>& o) { this->assignFrom(o); }
  template <class O> Array& operator=(const DenseBase<O>& o) { Array tmp; tmp.assignFrom(o); *static_cast<Base*>(this) = tmp; return *this; }
  static Array Ones() { Array m; m.setOnes(); return m; }
  static Array Zero() { Array m; m.setZero(); return m; }
  bool all() const { for (int k = 0; k < this->size(); ++k) if (!this->data()[k]) return false; return true; }
  bool any() const { for (int k = 0; k < this->size(); ++k) if (this->data()[k]) return true; return false; }
};
#define MINI_EIGEN_ARRAY_CMP(OP)                                                             \
  template <class T, int R, int C>                                                           \
  Array<bool, R, C> operator OP(const Array<T, R, C>& a, const Array<T, R, C>& b) {          \
    Array<bool, R, C> r;                                                                     \
    r.resize(a.rows(), a.cols());                                                            \
    for (int k = 0; k < a.size(); ++k) r.data()[k] = a.data()[k] OP b.data()[k];             \
    return r;                                                                                \
  }
MINI_EIGEN_ARRAY_CMP(<=)
MINI_EIGEN_ARRAY_CMP(>=)
MINI_EIGEN_ARRAY_CMP(<)
MINI_EIGEN_ARRAY_CMP(>)
#undef MINI_EIGEN_ARRAY_CMP

// A view into R x C coefficients of another object (rows/cols are run-time when Dynamic).
template <class M, int R, int C>
class Block : public DenseBase<Block<M, R, C>> {
  M* m_;
  int i0_, j0_, r_, c_;

 public:
  typedef typename traits<M>::Scalar Scalar;
  Block(M& m, int i0, int j0, int r, int c) : m_(&m), i0_(i0), j0_(j0), r_(r), c_(c) {}
  Block(const Block&) = default;
  int rows() const { return r_; }
  int cols() const { return c_; }
  Scalar coeff(int i, int j) const { return static_cast<const M*>(m_)->coeff(i0_ + i, j0_ + j); }
  Scalar& coeffRef(int i, int j) { return m_->coeffRef(i0_ + i, j0_ + j); }
  template <class O> Block& operator=(const DenseBase<O>& o) {
    const Matrix<Scalar, R, C> tmp(o);
    if (tmp.rows() == r_ && tmp.cols() == c_) {
      for (int j = 0; j < c_; ++j)
        for (int i = 0; i < r_; ++i) coeffRef(i, j) = tmp.coeff(i, j);
    } else {
      assert(tmp.size() == r_ * c_);
      for (int k = 0; k < r_ * c_; ++k) this->coeffRefLin(k) = tmp.coeffLin(k);
    }
    return *this;
  }
  Block& operator=(const Block& o) { return this->template operator=<Block>(o); }
};

// m << a, b, c ...   (coefficients in row-major order; a vector block fills the current column run)
template <class D>
class CommaInitializer {
  D& m_;
  int row_, col_, block_rows_;

 public:
  typedef typename traits<D>::Scalar Scalar;
  CommaInitializer(D& m) : m_(m), row_(0), col_(0), block_rows_(1) {}
  CommaInitializer& add(const Scalar& s) {
    if (col_ == m_.cols()) { row_ += block_rows_; col_ = 0; block_rows_ = 1; }
    m_.coeffRef(row_, col_++) = s;
    return *this;
  }
  template <class O> CommaInitializer& add(const DenseBase<O>& o) {
    if (col_ == m_.cols()) { row_ += block_rows_; col_ = 0; }
    block_rows_ = o.rows();
    for (int j = 0; j < o.cols(); ++j)
      for (int i = 0; i < o.rows(); ++i) m_.coeffRef(row_ + i, col_ + j) = o.derived().coeff(i, j);
    col_ += o.cols();
    return *this;
  }
  template <class S> typename std::enable_if<is_scalar<S>::value, CommaInitializer&>::type operator,(const S& s) { return add(static_cast<Scalar>(s)); }
  template <class O> CommaInitializer& operator,(const DenseBase<O>& o) { return add(o); }
  D& finished() { return m_; }
};
template <class D> CommaInitializer<D> DenseBase<D>::operator<<(const Scalar& s) {
  CommaInitializer<D> c(derived());
  c.add(s);
  return c;
}
template <class D> template <class O> CommaInitializer<D> DenseBase<D>::operator<<(const DenseBase<O>& o) {
  CommaInitializer<D> c(derived());
  c.add(o);
  return c;
}

// ---- free operators: eager, coefficient-wise, inner index ascending ---------------------------
template <class A, class B>
Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> operator+(const DenseBase<A>& a, const DenseBase<B>& b) {
  Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> r(a);
  r += b;
  return r;
}
template <class A, class B>
Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> operator-(const DenseBase<A>& a, const DenseBase<B>& b) {
  Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> r(a);
  r -= b;
  return r;
}
template <class A>
Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> operator-(const DenseBase<A>& a) {
  Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> r(a);
  for (int j = 0; j < r.cols(); ++j)
    for (int i = 0; i < r.rows(); ++i) r.coeffRef(i, j) = -r.coeff(i, j);
  return r;
}
template <class A, class B>
Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<B>::Cols> operator*(const DenseBase<A>& a, const DenseBase<B>& b) {
  typedef typename traits<A>::Scalar S;
  static_assert(std::is_same<S, typename traits<B>::Scalar>::value, "mixed scalar types in a product");
  assert(a.cols() == b.rows());
  Matrix<S, traits<A>::Rows, traits<B>::Cols> r;
  r.resize(a.rows(), b.cols());
  const int K = a.cols();
  for (int j = 0; j < b.cols(); ++j)
    for (int i = 0; i < a.rows(); ++i) {
      S s = a.derived().coeff(i, 0) * b.derived().coeff(0, j);
      for (int k = 1; k < K; ++k) s += a.derived().coeff(i, k) * b.derived().coeff(k, j);
      r.coeffRef(i, j) = s;
    }
  return r;
}
template <class S, class A>
typename std::enable_if<is_scalar<S>::value, Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols>>::type
operator*(const S& s, const DenseBase<A>& a) {
  Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> r(a);
  const typename traits<A>::Scalar f = static_cast<typename traits<A>::Scalar>(s);
  for (int j = 0; j < r.cols(); ++j)
    for (int i = 0; i < r.rows(); ++i) r.coeffRef(i, j) = f * r.coeff(i, j);
  return r;
}
template <class S, class A>
typename std::enable_if<is_scalar<S>::value, Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols>>::type
operator*(const DenseBase<A>& a, const S& s) {
  Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> r(a);
  r *= s;
  return r;
}
template <class S, class A>
typename std::enable_if<is_scalar<S>::value, Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols>>::type
operator/(const DenseBase<A>& a, const S& s) {
  Matrix<typename traits<A>::Scalar, traits<A>::Rows, traits<A>::Cols> r(a);
  r /= s;
  return r;
}

typedef Matrix<double, 3, 3> Matrix3d;
typedef Matrix<float, 3, 3> Matrix3f;
typedef Matrix<float, 4, 4> Matrix4f;
typedef Matrix<double, 4, 4> Matrix4d;
typedef Matrix<double, 3, 1> Vector3d;
typedef Matrix<float, 3, 1> Vector3f;
typedef Matrix<float, 4, 1> Vector4f;
typedef Matrix<double, 4, 1> Vector4d;
typedef Matrix<int, 3, 1> Vector3i;
typedef Matrix<int, 4, 1> Vector4i;
typedef Matrix<float, Dynamic, 1> VectorXf;
typedef Matrix<double, Dynamic, 1> VectorXd;
typedef Matrix<int, Dynamic, Dynamic> MatrixXi;
typedef Matrix<float, Dynamic, Dynamic> MatrixXf;
typedef Matrix<double, Dynamic, Dynamic> MatrixXd;
typedef Array<int, 4, 1> Array4i;
typedef Array<float, 4, 1> Array4f;

}  // namespace Eigen
