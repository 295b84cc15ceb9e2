/*
 * mini_eigen.h — a small, eager (no expression templates) stand-in for the subset of Eigen 3
 * that the reference's convex-MPC sources use.  TEST INFRASTRUCTURE ONLY (see ../Makefile,
 * target `ref`): it exists so that the UNMODIFIED reference files
 *     src/legged_ctrl/src/mpc_ctrl/convex_mpc/{ConvexQPSolver,ConvexMpc}.cpp
 *     src/legged_ctrl/src/mpc_ctrl/LeggedMPC.cpp, src/LeggedState.cpp
 *     src/legged_ctrl/src/utils/{LeggedContactFSM,Utils,A1Kinematics}.cpp
 * compile with g++ in an image that has no Eigen, and the QP they would hand to OSQP can be
 * captured.  Nothing here is derived from Eigen's sources; it implements the documented
 * semantics of the calls those files make (column-major storage, row-wise comma initialiser,
 * cofactor 3x3 inverse, k-ascending dot products).  The product never includes this file.
 */
#ifndef LMPC_MINI_EIGEN_H_
#define LMPC_MINI_EIGEN_H_

#include <algorithm>
#include <cassert>
#include <cmath>
#include <math.h> /* the reference calls unqualified isnan (ConvexQPSolver.cpp:321) */
#include <cstddef>
#include <iostream>
#include <limits>
#include <type_traits>
#include <vector>

namespace Eigen {

typedef std::ptrdiff_t Index;
const int Dynamic = -1;
const int Infinity = -1; /* Eigen's norm-type tag; A1Kinematics.cpp:312 passes it as a coefficient */
enum StorageOptions { ColMajor = 0, RowMajor = 1 };
enum DecompositionOptions { ComputeFullU = 0x04, ComputeThinU = 0x08, ComputeFullV = 0x10, ComputeThinV = 0x20 };

namespace detail {
constexpr int mul_dim(int a, int b) { return (a == Dynamic || b == Dynamic) ? Dynamic : a * b; }

template <class T, int R, int C, bool Dyn = (R == Dynamic || C == Dynamic)>
struct Storage {
  T v[R * C];
  Index rows() const { return R; }
  Index cols() const { return C; }
  void resize(Index r, Index c) { assert(r == R && c == C); (void)r; (void)c; }
  T* data() { return v; }
  const T* data() const { return v; }
};
template <class T, int R, int C>
struct Storage<T, R, C, true> {
  std::vector<T> v;
  Index r_ = (R == Dynamic ? 0 : R), c_ = (C == Dynamic ? 0 : C);
  Index rows() const { return r_; }
  Index cols() const { return c_; }
  void resize(Index r, Index c) {
    assert((R == Dynamic || r == R) && (C == Dynamic || c == C));
    r_ = r; c_ = c; v.resize((size_t)(r * c));
  }
  T* data() { return v.data(); }
  const T* data() const { return v.data(); }
};
}  // namespace detail

template <class T, int R, int C> class Matrix;
template <class T, int BR, int BC> class Block;
template <class T, int N> class DiagonalMatrix;
template <class T, int R, int C> class Array;

/* Row-wise comma initialiser: `m << a, b, c, ...;` fills row by row; a matrix argument occupies
 * a block (only whole-object blocks are needed here). */
template <class M>
class CommaInitializer {
 public:
  typedef typename M::Scalar T;
  CommaInitializer(M& m) : m_(m), r_(0), c_(0) {}
  template <class S, typename std::enable_if<std::is_arithmetic<S>::value, int>::type = 0>
  CommaInitializer& operator,(S s) { put((T)s); return *this; }
  template <int R2, int C2>
  CommaInitializer& operator,(const Matrix<T, R2, C2>& b) { put_block(b); return *this; }
  template <class S, typename std::enable_if<std::is_arithmetic<S>::value, int>::type = 0>
  void put(S s) {
    assert(r_ < m_.rows());
    m_(r_, c_) = (T)s;
    if (++c_ == m_.cols()) { c_ = 0; ++r_; }
  }
  template <int R2, int C2>
  void put_block(const Matrix<T, R2, C2>& b) {
    for (Index i = 0; i < b.rows(); ++i)
      for (Index j = 0; j < b.cols(); ++j) m_(r_ + i, c_ + j) = b(i, j);
    c_ += b.cols();
    if (c_ >= m_.cols()) { c_ = 0; r_ += b.rows(); }
  }
 private:
  M& m_;
  Index r_, c_;
};

template <class T, int R, int C>
class Matrix {
 public:
  typedef T Scalar;
  enum { RowsAtCompileTime = R, ColsAtCompileTime = C };
  static const bool kVector = (R == 1 || C == 1);

  Matrix() {}
  /* size constructors of dynamic objects / coefficient constructors of small fixed vectors */
  explicit Matrix(Index n) {
    if (R == Dynamic && C == 1) s_.resize(n, 1);
    else if (C == Dynamic && R == 1) s_.resize(1, n);
    else if (R == 1 && C == 1) s_.v[0] = (T)n;
    else assert(false && "unsupported 1-arg ctor");
  }
  Matrix(T a, T b) {
    if (R == Dynamic || C == Dynamic) s_.resize((Index)a, (Index)b);
    else { assert(R * C == 2); s_.data()[0] = a; s_.data()[1] = b; }
  }
  Matrix(T a, T b, T c) { static_assert(detail::mul_dim(R, C) == 3, "3-coefficient ctor"); T* d = s_.data(); d[0] = a; d[1] = b; d[2] = c; }
  Matrix(T a, T b, T c, T e) { static_assert(detail::mul_dim(R, C) == 4, "4-coefficient ctor"); T* d = s_.data(); d[0] = a; d[1] = b; d[2] = c; d[3] = e; }

  template <int R2, int C2>
  Matrix(const Matrix<T, R2, C2>& o) { assign(o); }
  template <int BR, int BC>
  Matrix(const Block<T, BR, BC>& b);
  template <int R2, int C2>
  Matrix& operator=(const Matrix<T, R2, C2>& o) { assign(o); return *this; }
  template <int BR, int BC>
  Matrix& operator=(const Block<T, BR, BC>& b);

  Index rows() const { return s_.rows(); }
  Index cols() const { return s_.cols(); }
  Index size() const { return rows() * cols(); }
  Index outerSize() const { return cols(); }
  T* data() { return s_.data(); }
  const T* data() const { return s_.data(); }
  void resize(Index n) { if (C == 1) s_.resize(n, 1); else s_.resize(1, n); }
  void resize(Index r, Index c) { s_.resize(r, c); }

  T& operator()(Index i, Index j) { assert(i >= 0 && i < rows() && j >= 0 && j < cols()); return s_.data()[i + j * rows()]; }
  const T& operator()(Index i, Index j) const { assert(i >= 0 && i < rows() && j >= 0 && j < cols()); return s_.data()[i + j * rows()]; }
  T& operator()(Index i) { assert(i >= 0 && i < size()); return s_.data()[i]; }
  const T& operator()(Index i) const { assert(i >= 0 && i < size()); return s_.data()[i]; }
  T& operator[](Index i) { return (*this)(i); }
  const T& operator[](Index i) const { return (*this)(i); }
  T& x() { return (*this)(0); } T& y() { return (*this)(1); } T& z() { return (*this)(2); } T& w() { return (*this)(3); }
  const T& x() const { return (*this)(0); } const T& y() const { return (*this)(1); }
  const T& z() const { return (*this)(2); } const T& w() const { return (*this)(3); }

  Matrix& setZero() { std::fill(data(), data() + size(), T(0)); return *this; }
  Matrix& setConstant(T v) { std::fill(data(), data() + size(), v); return *this; }
  Matrix& setIdentity() { setZero(); for (Index i = 0; i < std::min(rows(), cols()); ++i) (*this)(i, i) = T(1); return *this; }
  static Matrix Zero() { Matrix m; m.setZero(); return m; }
  static Matrix Zero(Index n) { Matrix m(n); m.setZero(); return m; }
  static Matrix Zero(Index r, Index c) { Matrix m; m.resize(r, c); m.setZero(); return m; }
  static Matrix Identity() { Matrix m; m.setIdentity(); return m; }
  static Matrix Identity(Index r, Index c) { Matrix m; m.resize(r, c); m.setIdentity(); return m; }
  static Matrix Ones() { Matrix m; m.setConstant(T(1)); return m; }
  static Matrix Constant(T v) { Matrix m; m.setConstant(v); return m; }

  template <class S, typename std::enable_if<std::is_arithmetic<S>::value, int>::type = 0>
  CommaInitializer<Matrix> operator<<(S s) { CommaInitializer<Matrix> ci(*this); ci.put((T)s); return ci; }
  template <int R2, int C2>
  CommaInitializer<Matrix> operator<<(const Matrix<T, R2, C2>& b) { CommaInitializer<Matrix> ci(*this); ci.put_block(b); return ci; }

  /* views */
  template <int BR, int BC> Block<T, BR, BC> block(Index i, Index j) { return Block<T, BR, BC>(&(*this)(i, j), rows(), BR, BC); }
  template <int BR, int BC> Matrix<T, BR, BC> block(Index i, Index j) const { return copy_block<BR, BC>(i, j, BR, BC); }
  Block<T, Dynamic, Dynamic> block(Index i, Index j, Index r, Index c) { return Block<T, Dynamic, Dynamic>(&(*this)(i, j), rows(), r, c); }
  Matrix<T, Dynamic, Dynamic> block(Index i, Index j, Index r, Index c) const { return copy_block<Dynamic, Dynamic>(i, j, r, c); }
  template <int N> Block<T, (C == 1 ? N : 1), (C == 1 ? 1 : N)> segment(Index i) {
    return C == 1 ? Block<T, (C == 1 ? N : 1), (C == 1 ? 1 : N)>(data() + i, rows(), N, 1)
                  : Block<T, (C == 1 ? N : 1), (C == 1 ? 1 : N)>(data() + i * rows(), rows(), 1, N);
  }
  template <int N> Matrix<T, (C == 1 ? N : 1), (C == 1 ? 1 : N)> segment(Index i) const {
    Matrix<T, (C == 1 ? N : 1), (C == 1 ? 1 : N)> m;
    for (Index k = 0; k < N; ++k) m(k) = (*this)(i + k);
    return m;
  }
  Block<T, (C == 1 ? Dynamic : 1), (C == 1 ? 1 : Dynamic)> segment(Index i, Index n) {
    return C == 1 ? Block<T, (C == 1 ? Dynamic : 1), (C == 1 ? 1 : Dynamic)>(data() + i, rows(), n, 1)
                  : Block<T, (C == 1 ? Dynamic : 1), (C == 1 ? 1 : Dynamic)>(data() + i * rows(), rows(), 1, n);
  }
  Matrix<T, (C == 1 ? Dynamic : 1), (C == 1 ? 1 : Dynamic)> segment(Index i, Index n) const {
    Matrix<T, (C == 1 ? Dynamic : 1), (C == 1 ? 1 : Dynamic)> m;
    if (C == 1) m.resize(n, 1); else m.resize(1, n);
    for (Index k = 0; k < n; ++k) m(k) = (*this)(i + k);
    return m;
  }
  template <int N> Block<T, (C == 1 ? N : 1), (C == 1 ? 1 : N)> head() { return segment<N>(0); }
  template <int N> Block<T, (C == 1 ? N : 1), (C == 1 ? 1 : N)> tail() { return segment<N>(size() - N); }
  Block<T, R, 1> col(Index j) { return Block<T, R, 1>(&(*this)(0, j), rows(), rows(), 1); }
  Matrix<T, R, 1> col(Index j) const { Matrix<T, R, 1> m; m.resize(rows(), 1); for (Index i = 0; i < rows(); ++i) m(i) = (*this)(i, j); return m; }
  Block<T, 1, C> row(Index i) { return Block<T, 1, C>(&(*this)(i, 0), rows(), 1, cols()); }
  Matrix<T, 1, C> row(Index i) const { Matrix<T, 1, C> m; m.resize(1, cols()); for (Index j = 0; j < cols(); ++j) m(j) = (*this)(i, j); return m; }

  /* arithmetic (eager) */
  Matrix<T, C, R> transpose() const {
    Matrix<T, C, R> t; t.resize(cols(), rows());
    for (Index i = 0; i < rows(); ++i) for (Index j = 0; j < cols(); ++j) t(j, i) = (*this)(i, j);
    return t;
  }
  Matrix<T, C, R> adjoint() const { return transpose(); }
  T squaredNorm() const { T s = T(0); for (Index k = 0; k < size(); ++k) s += (*this)(k) * (*this)(k); return s; }
  T norm() const { return std::sqrt(squaredNorm()); }
  T sum() const { T s = T(0); for (Index k = 0; k < size(); ++k) s += (*this)(k); return s; }
  T dot(const Matrix& o) const { T s = T(0); for (Index k = 0; k < size(); ++k) s += (*this)(k) * o(k); return s; }
  T determinant() const;
  Matrix inverse() const;
  Matrix normalized() const { return *this / norm(); }
  Matrix<T, 3, 1> cross(const Matrix<T, 3, 1>& o) const {
    const Matrix& a = *this;
    return Matrix<T, 3, 1>(a(1) * o(2) - a(2) * o(1), a(2) * o(0) - a(0) * o(2), a(0) * o(1) - a(1) * o(0));
  }
  Matrix operator-() const { Matrix m = *this; for (Index k = 0; k < size(); ++k) m(k) = -m(k); return m; }
  Matrix& operator+=(const Matrix& o) { for (Index k = 0; k < size(); ++k) (*this)(k) += o(k); return *this; }
  Matrix& operator-=(const Matrix& o) { for (Index k = 0; k < size(); ++k) (*this)(k) -= o(k); return *this; }
  Matrix& operator*=(T s) { for (Index k = 0; k < size(); ++k) (*this)(k) *= s; return *this; }
  Matrix& operator/=(T s) { for (Index k = 0; k < size(); ++k) (*this)(k) /= s; return *this; }
  Matrix operator/(T s) const { Matrix m = *this; m /= s; return m; }
  Matrix operator*(T s) const { Matrix m = *this; m *= s; return m; }

  Array<T, R, C> array() const;
  DiagonalMatrix<T, (C == 1 ? R : C)> asDiagonal() const;
  Matrix<T, (R < C ? R : C) == Dynamic ? Dynamic : (R < C ? R : C), 1> diagonal() const {
    Matrix<T, (R < C ? R : C) == Dynamic ? Dynamic : (R < C ? R : C), 1> d;
    Index n = std::min(rows(), cols()); d.resize(n, 1);
    for (Index i = 0; i < n; ++i) d(i) = (*this)(i, i);
    return d;
  }
  bool hasNaN() const { for (Index k = 0; k < size(); ++k) if ((*this)(k) != (*this)(k)) return true; return false; }
  Matrix eval() const { return *this; }

 private:
  template <int R2, int C2>
  void assign(const Matrix<T, R2, C2>& o) {
    static_assert((R == Dynamic || R2 == Dynamic || R == R2) && (C == Dynamic || C2 == Dynamic || C == C2),
                  "mini_eigen: incompatible fixed sizes");
    s_.resize(o.rows(), o.cols());
    std::copy(o.data(), o.data() + o.size(), data());
  }
  template <int BR, int BC>
  Matrix<T, BR, BC> copy_block(Index i, Index j, Index r, Index c) const {
    Matrix<T, BR, BC> m; m.resize(r, c);
    for (Index a = 0; a < r; ++a) for (Index b = 0; b < c; ++b) m(a, b) = (*this)(i + a, j + b);
    return m;
  }
  detail::Storage<T, R, C> s_;
};

/* Writable view into a column-major parent (base pointer + leading dimension). */
template <class T, int BR, int BC>
class Block {
 public:
  typedef T Scalar;
  Block(T* base, Index ld, Index r, Index c) : p_(base), ld_(ld), r_(r), c_(c) {}
  Index rows() const { return r_; }
  Index cols() const { return c_; }
  Index size() const { return r_ * c_; }
  T& operator()(Index i, Index j) const { return p_[i + j * ld_]; }
  T& operator()(Index i) const { return (BC == 1 || c_ == 1) ? p_[i] : p_[i * ld_]; }
  T& operator[](Index i) const { return (*this)(i); }
  Block& operator=(const Block& o) { return copy_from(o); }
  template <int R2, int C2>
  Block& operator=(const Block<T, R2, C2>& o) { return copy_from(o); }
  template <int R2, int C2>
  Block& operator=(const Matrix<T, R2, C2>& o) { return copy_from(o); }
  Block& setZero() { for (Index j = 0; j < c_; ++j) for (Index i = 0; i < r_; ++i) (*this)(i, j) = T(0); return *this; }
  Matrix<T, BR, BC> eval() const {
    Matrix<T, BR, BC> m; m.resize(r_, c_);
    for (Index j = 0; j < c_; ++j) for (Index i = 0; i < r_; ++i) m(i, j) = (*this)(i, j);
    return m;
  }
  T norm() const { return eval().norm(); }
  T squaredNorm() const { return eval().squaredNorm(); }
  Matrix<T, BC, BR> transpose() const { return eval().transpose(); }
  template <class S, typename std::enable_if<std::is_arithmetic<S>::value, int>::type = 0>
  CommaInitializer<Block> operator<<(S s) { CommaInitializer<Block> ci(*this); ci.put((T)s); return ci; }
 private:
  template <class O>
  Block& copy_from(const O& o) {
    assert(o.rows() * o.cols() == r_ * c_);
    if (o.rows() == r_) { for (Index j = 0; j < c_; ++j) for (Index i = 0; i < r_; ++i) (*this)(i, j) = o(i, j); }
    else { for (Index k = 0; k < r_ * c_; ++k) (*this)(k) = o(k); } /* vector <-> transposed vector shapes */
    return *this;
  }
  T* p_;
  Index ld_, r_, c_;
};

template <class T, int R, int C>
template <int BR, int BC>
Matrix<T, R, C>::Matrix(const Block<T, BR, BC>& b) { assign(b.eval()); }
template <class T, int R, int C>
template <int BR, int BC>
Matrix<T, R, C>& Matrix<T, R, C>::operator=(const Block<T, BR, BC>& b) { assign(b.eval()); return *this; }

/* ---- free operators ---- */
template <class T, int R, int C>
Matrix<T, R, C> operator+(const Matrix<T, R, C>& a, const Matrix<T, R, C>& b) { Matrix<T, R, C> m = a; m += b; return m; }
template <class T, int R, int C>
Matrix<T, R, C> operator-(const Matrix<T, R, C>& a, const Matrix<T, R, C>& b) { Matrix<T, R, C> m = a; m -= b; return m; }
template <class T, int R, int C, int BR, int BC>
Matrix<T, R, C> operator+(const Matrix<T, R, C>& a, const Block<T, BR, BC>& b) { return a + Matrix<T, R, C>(b); }
template <class T, int R, int C, int BR, int BC>
Matrix<T, R, C> operator-(const Matrix<T, R, C>& a, const Block<T, BR, BC>& b) { return a - Matrix<T, R, C>(b); }
template <class T, int R, int C, int BR, int BC>
Matrix<T, R, C> operator+(const Block<T, BR, BC>& b, const Matrix<T, R, C>& a) { return Matrix<T, R, C>(b) + a; }
template <class T, int R, int C, int BR, int BC>
Matrix<T, R, C> operator-(const Block<T, BR, BC>& b, const Matrix<T, R, C>& a) { return Matrix<T, R, C>(b) - a; }
template <class T, int R, int C>
Matrix<T, R, C> operator*(double s, const Matrix<T, R, C>& a) { Matrix<T, R, C> m = a; m *= (T)s; return m; }
/* product: coefficient (i,j) is the k-ascending sum of a(i,k)*b(k,j) */
template <class T, int R, int K, int K2, int C>
Matrix<T, R, C> operator*(const Matrix<T, R, K>& a, const Matrix<T, K2, C>& b) {
  static_assert(K == Dynamic || K2 == Dynamic || K == K2, "mini_eigen: inner dimensions differ");
  assert(a.cols() == b.rows());
  Matrix<T, R, C> m; m.resize(a.rows(), b.cols());
  for (Index j = 0; j < b.cols(); ++j)
    for (Index i = 0; i < a.rows(); ++i) {
      T s = T(0);
      for (Index k = 0; k < a.cols(); ++k) s += a(i, k) * b(k, j);
      m(i, j) = s;
    }
  return m;
}
template <class T, int R, int K, int BR, int BC>
Matrix<T, R, BC> operator*(const Matrix<T, R, K>& a, const Block<T, BR, BC>& b) { return a * b.eval(); }
template <class T, int R, int C>
std::ostream& operator<<(std::ostream& os, const Matrix<T, R, C>& m) {
  for (Index i = 0; i < m.rows(); ++i) { for (Index j = 0; j < m.cols(); ++j) os << (j ? " " : "") << m(i, j); if (i + 1 < m.rows()) os << "\n"; }
  return os;
}

template <class T, int R, int C>
T Matrix<T, R, C>::determinant() const {
  assert(rows() == cols());
  const Matrix& m = *this;
  if (rows() == 1) return m(0, 0);
  if (rows() == 2) return m(0, 0) * m(1, 1) - m(0, 1) * m(1, 0);
  if (rows() == 3)
    return m(0, 0) * (m(1, 1) * m(2, 2) - m(1, 2) * m(2, 1)) - m(0, 1) * (m(1, 0) * m(2, 2) - m(1, 2) * m(2, 0)) +
           m(0, 2) * (m(1, 0) * m(2, 1) - m(1, 1) * m(2, 0));
  Matrix a = *this; T det = T(1); Index n = rows();
  for (Index k = 0; k < n; ++k) {
    Index p = k; for (Index i = k + 1; i < n; ++i) if (std::abs(a(i, k)) > std::abs(a(p, k))) p = i;
    if (a(p, k) == T(0)) return T(0);
    if (p != k) { for (Index j = 0; j < n; ++j) std::swap(a(k, j), a(p, j)); det = -det; }
    det *= a(k, k);
    for (Index i = k + 1; i < n; ++i) { T f = a(i, k) / a(k, k); for (Index j = k; j < n; ++j) a(i, j) -= f * a(k, j); }
  }
  return det;
}

/* inverse: closed-form cofactors for n <= 3 (the size the path uses: ConvexQPSolver.cpp:205),
 * partial-pivot Gauss-Jordan otherwise. */
template <class T, int R, int C>
Matrix<T, R, C> Matrix<T, R, C>::inverse() const {
  assert(rows() == cols());
  const Matrix& m = *this; Matrix r; r.resize(rows(), cols());
  if (rows() == 1) { r(0, 0) = T(1) / m(0, 0); return r; }
  if (rows() == 2) {
    T id = T(1) / determinant();
    r(0, 0) = m(1, 1) * id; r(0, 1) = -m(0, 1) * id; r(1, 0) = -m(1, 0) * id; r(1, 1) = m(0, 0) * id;
    return r;
  }
  if (rows() == 3) {
    T c00 = m(1, 1) * m(2, 2) - m(1, 2) * m(2, 1), c10 = m(1, 2) * m(2, 0) - m(1, 0) * m(2, 2), c20 = m(1, 0) * m(2, 1) - m(1, 1) * m(2, 0);
    T det = m(0, 0) * c00 + m(0, 1) * c10 + m(0, 2) * c20, id = T(1) / det;
    r(0, 0) = c00 * id; r(1, 0) = c10 * id; r(2, 0) = c20 * id;
    r(0, 1) = (m(0, 2) * m(2, 1) - m(0, 1) * m(2, 2)) * id;
    r(1, 1) = (m(0, 0) * m(2, 2) - m(0, 2) * m(2, 0)) * id;
    r(2, 1) = (m(0, 1) * m(2, 0) - m(0, 0) * m(2, 1)) * id;
    r(0, 2) = (m(0, 1) * m(1, 2) - m(0, 2) * m(1, 1)) * id;
    r(1, 2) = (m(0, 2) * m(1, 0) - m(0, 0) * m(1, 2)) * id;
    r(2, 2) = (m(0, 0) * m(1, 1) - m(0, 1) * m(1, 0)) * id;
    return r;
  }
  Index n = rows(); Matrix a = *this; r.setIdentity();
  for (Index k = 0; k < n; ++k) {
    Index p = k; for (Index i = k + 1; i < n; ++i) if (std::abs(a(i, k)) > std::abs(a(p, k))) p = i;
    if (p != k) for (Index j = 0; j < n; ++j) { std::swap(a(k, j), a(p, j)); std::swap(r(k, j), r(p, j)); }
    T d = T(1) / a(k, k);
    for (Index j = 0; j < n; ++j) { a(k, j) *= d; r(k, j) *= d; }
    for (Index i = 0; i < n; ++i) if (i != k) { T f = a(i, k); if (f != T(0)) for (Index j = 0; j < n; ++j) { a(i, j) -= f * a(k, j); r(i, j) -= f * r(k, j); } }
  }
  return r;
}

/* ---- DiagonalMatrix ---- */
template <class T, int N>
class DiagonalMatrix {
 public:
  DiagonalMatrix() {}
  explicit DiagonalMatrix(const Matrix<T, N, 1>& d) : d_(d) {}
  Matrix<T, N, 1>& diagonal() { return d_; }
  const Matrix<T, N, 1>& diagonal() const { return d_; }
  Index rows() const { return d_.size(); }
  Index cols() const { return d_.size(); }
  Matrix<T, N, N> toDenseMatrix() const { Matrix<T, N, N> m; m.resize(rows(), rows()); m.setZero(); for (Index i = 0; i < rows(); ++i) m(i, i) = d_(i); return m; }
 private:
  Matrix<T, N, 1> d_;
};
template <class T, int N>
DiagonalMatrix<T, N> operator*(double s, const DiagonalMatrix<T, N>& d) { DiagonalMatrix<T, N> r = d; r.diagonal() *= (T)s; return r; }
template <class T, int N, int R, int C>
Matrix<T, R, C> operator*(const DiagonalMatrix<T, N>& d, const Matrix<T, R, C>& m) {
  assert(d.cols() == m.rows());
  Matrix<T, R, C> r = m;
  for (Index j = 0; j < m.cols(); ++j) for (Index i = 0; i < m.rows(); ++i) r(i, j) = d.diagonal()(i) * m(i, j);
  return r;
}
template <class T, int N, int R, int C>
Matrix<T, R, C> operator*(const Matrix<T, R, C>& m, const DiagonalMatrix<T, N>& d) {
  assert(d.rows() == m.cols());
  Matrix<T, R, C> r = m;
  for (Index j = 0; j < m.cols(); ++j) for (Index i = 0; i < m.rows(); ++i) r(i, j) = m(i, j) * d.diagonal()(j);
  return r;
}
template <class T, int R, int C>
DiagonalMatrix<T, (C == 1 ? R : C)> Matrix<T, R, C>::asDiagonal() const {
  DiagonalMatrix<T, (C == 1 ? R : C)> d; d.diagonal().resize(size(), 1);
  for (Index k = 0; k < size(); ++k) d.diagonal()(k) = (*this)(k);
  return d;
}

/* ---- Array (coefficient-wise view; only what Utils::pseudo_inverse needs) ---- */
template <class T, int R, int C>
class Array {
 public:
  Array() {}
  explicit Array(const Matrix<T, R, C>& m) : m_(m) {}
  Index size() const { return m_.size(); }
  T operator()(Index i) const { return m_(i); }
  T& operator()(Index i) { return m_(i); }
  Array abs() const { Array a = *this; for (Index k = 0; k < size(); ++k) a.m_(k) = std::abs(m_(k)); return a; }
  Array inverse() const { Array a = *this; for (Index k = 0; k < size(); ++k) a.m_(k) = T(1) / m_(k); return a; }
  Array square() const { Array a = *this; for (Index k = 0; k < size(); ++k) a.m_(k) = m_(k) * m_(k); return a; }
  Array sqrt() const { Array a = *this; for (Index k = 0; k < size(); ++k) a.m_(k) = std::sqrt(m_(k)); return a; }
  Matrix<T, R, C> matrix() const { return m_; }
  struct Mask {
    std::vector<char> b;
    Array select(const Array& then_, T else_) const { Array a = then_; for (Index k = 0; k < a.size(); ++k) if (!b[(size_t)k]) a.m_(k) = else_; return a; }
  };
  Mask operator>(T t) const { Mask m; m.b.resize((size_t)size()); for (Index k = 0; k < size(); ++k) m.b[(size_t)k] = m_(k) > t; return m; }
  Mask operator<(T t) const { Mask m; m.b.resize((size_t)size()); for (Index k = 0; k < size(); ++k) m.b[(size_t)k] = m_(k) < t; return m; }
 private:
  Matrix<T, R, C> m_;
};
template <class T, int R, int C>
Array<T, R, C> Matrix<T, R, C>::array() const { return Array<T, R, C>(*this); }

/* ---- Quaternion (coefficients stored x,y,z,w as Eigen documents) ---- */
template <class T>
class Quaternion {
 public:
  Quaternion() {}
  Quaternion(T w, T x, T y, T z) { c_(0) = x; c_(1) = y; c_(2) = z; c_(3) = w; }
  T& x() { return c_(0); } T& y() { return c_(1); } T& z() { return c_(2); } T& w() { return c_(3); }
  const T& x() const { return c_(0); } const T& y() const { return c_(1); } const T& z() const { return c_(2); } const T& w() const { return c_(3); }
  Matrix<T, 4, 1>& coeffs() { return c_; }
  const Matrix<T, 4, 1>& coeffs() const { return c_; }
  Matrix<T, 3, 1> vec() const { return Matrix<T, 3, 1>(c_(0), c_(1), c_(2)); }
  T norm() const { return c_.norm(); }
  void normalize() { c_ /= c_.norm(); }
  Matrix<T, 3, 3> toRotationMatrix() const {
    T x = c_(0), y = c_(1), z = c_(2), w = c_(3);
    Matrix<T, 3, 3> r;
    r(0, 0) = 1 - 2 * (y * y + z * z); r(0, 1) = 2 * (x * y - w * z); r(0, 2) = 2 * (x * z + w * y);
    r(1, 0) = 2 * (x * y + w * z); r(1, 1) = 1 - 2 * (x * x + z * z); r(1, 2) = 2 * (y * z - w * x);
    r(2, 0) = 2 * (x * z - w * y); r(2, 1) = 2 * (y * z + w * x); r(2, 2) = 1 - 2 * (x * x + y * y);
    return r;
  }
 private:
  Matrix<T, 4, 1> c_;
};

/* ---- JacobiSVD (one-sided Jacobi; only Utils::pseudo_inverse uses it, off the hot path) ---- */
template <class M>
class JacobiSVD {
 public:
  typedef typename M::Scalar T;
  JacobiSVD(const M& a, unsigned int = 0) {
    Index m = a.rows(), n = a.cols();
    u_ = a; v_.resize(n, n); v_.setIdentity(); s_.resize(n, 1);
    for (int sweep = 0; sweep < 60; ++sweep) {
      T off = 0;
      for (Index p = 0; p < n; ++p)
        for (Index q = p + 1; q < n; ++q) {
          T al = 0, be = 0, ga = 0;
          for (Index i = 0; i < m; ++i) { al += u_(i, p) * u_(i, p); be += u_(i, q) * u_(i, q); ga += u_(i, p) * u_(i, q); }
          if (ga == T(0)) continue;
          off = std::max(off, std::abs(ga) / std::sqrt(al * be + std::numeric_limits<T>::min()));
          T zeta = (be - al) / (2 * ga);
          T t = (zeta >= 0 ? T(1) : T(-1)) / (std::abs(zeta) + std::sqrt(1 + zeta * zeta));
          T c = 1 / std::sqrt(1 + t * t), s = c * t;
          for (Index i = 0; i < m; ++i) { T x = u_(i, p), y = u_(i, q); u_(i, p) = c * x - s * y; u_(i, q) = s * x + c * y; }
          for (Index i = 0; i < n; ++i) { T x = v_(i, p), y = v_(i, q); v_(i, p) = c * x - s * y; v_(i, q) = s * x + c * y; }
        }
      if (off < 1e-15) break;
    }
    for (Index j = 0; j < n; ++j) {
      T nrm = 0; for (Index i = 0; i < m; ++i) nrm += u_(i, j) * u_(i, j);
      nrm = std::sqrt(nrm); s_(j) = nrm;
      if (nrm > 0) for (Index i = 0; i < m; ++i) u_(i, j) /= nrm;
    }
    for (Index a0 = 0; a0 < n; ++a0) { /* descending singular values */
      Index b = a0; for (Index j = a0 + 1; j < n; ++j) if (s_(j) > s_(b)) b = j;
      if (b != a0) { std::swap(s_(a0), s_(b)); for (Index i = 0; i < m; ++i) std::swap(u_(i, a0), u_(i, b)); for (Index i = 0; i < n; ++i) std::swap(v_(i, a0), v_(i, b)); }
    }
  }
  const M& matrixU() const { return u_; }
  const M& matrixV() const { return v_; }
  const Matrix<T, M::ColsAtCompileTime, 1>& singularValues() const { return s_; }
 private:
  M u_, v_;
  Matrix<T, M::ColsAtCompileTime, 1> s_;
};

typedef Matrix<double, 2, 1> Vector2d;
typedef Matrix<double, 3, 1> Vector3d;
typedef Matrix<double, 4, 1> Vector4d;
typedef Matrix<double, Dynamic, 1> VectorXd;
typedef Matrix<double, 2, 2> Matrix2d;
typedef Matrix<double, 3, 3> Matrix3d;
typedef Matrix<double, 4, 4> Matrix4d;
typedef Matrix<double, Dynamic, Dynamic> MatrixXd;
typedef Matrix<float, 3, 1> Vector3f;
typedef Matrix<int, Dynamic, 1> VectorXi;
typedef Quaternion<double> Quaterniond;

}  // namespace Eigen
#endif
