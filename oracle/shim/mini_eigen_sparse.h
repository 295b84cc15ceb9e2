/*
 * mini_eigen_sparse.h — the slice of Eigen's SparseMatrix / Triplet API that
 * ConvexQPSolver.{h,cpp} uses (insert, makeCompressed, InnerIterator with valueRef, outerSize).
 * TEST INFRASTRUCTURE ONLY; see mini_eigen.h.  Column-major, inner (row) indices kept sorted,
 * which is the order Eigen's InnerIterator visits them in.
 */
#ifndef LMPC_MINI_EIGEN_SPARSE_H_
#define LMPC_MINI_EIGEN_SPARSE_H_
#include "mini_eigen.h"
#include <utility>
#include <vector>

namespace Eigen {

template <class T, class I = int>
class Triplet {
 public:
  Triplet() : r_(0), c_(0), v_(0) {}
  Triplet(const I& r, const I& c, const T& v = T(0)) : r_(r), c_(c), v_(v) {}
  const I& row() const { return r_; }
  const I& col() const { return c_; }
  const T& value() const { return v_; }
 private:
  I r_, c_;
  T v_;
};

template <class T, int Options = ColMajor, class I = int>
class SparseMatrix {
  static_assert(Options == ColMajor, "mini_eigen: only column-major sparse matrices");
 public:
  typedef T Scalar;
  SparseMatrix() : rows_(0) {}
  SparseMatrix(Index r, Index c) : rows_(0) { resize(r, c); }
  void resize(Index r, Index c) { rows_ = r; cols_.assign((size_t)c, std::vector<std::pair<I, T> >()); }
  Index rows() const { return rows_; }
  Index cols() const { return (Index)cols_.size(); }
  Index outerSize() const { return cols(); }
  Index innerSize() const { return rows(); }
  Index nonZeros() const { Index n = 0; for (size_t k = 0; k < cols_.size(); ++k) n += (Index)cols_[k].size(); return n; }
  /* insert(i,j): the entry must not exist yet; returns a reference to the new (zero) value */
  T& insert(Index i, Index j) {
    assert(i >= 0 && i < rows_ && j >= 0 && j < cols());
    std::vector<std::pair<I, T> >& col = cols_[(size_t)j];
    size_t p = 0;
    while (p < col.size() && col[p].first < (I)i) ++p;
    assert(p == col.size() || col[p].first != (I)i);
    col.insert(col.begin() + (std::ptrdiff_t)p, std::make_pair((I)i, T(0)));
    return col[p].second;
  }
  T coeff(Index i, Index j) const {
    const std::vector<std::pair<I, T> >& col = cols_[(size_t)j];
    for (size_t p = 0; p < col.size(); ++p) if (col[p].first == (I)i) return col[p].second;
    return T(0);
  }
  T& coeffRef(Index i, Index j) {
    std::vector<std::pair<I, T> >& col = cols_[(size_t)j];
    for (size_t p = 0; p < col.size(); ++p) if (col[p].first == (I)i) return col[p].second;
    return insert(i, j);
  }
  void makeCompressed() {}
  bool isCompressed() const { return true; }
  void setZero() { for (size_t k = 0; k < cols_.size(); ++k) cols_[k].clear(); }
  template <class It>
  void setFromTriplets(It b, It e) { setZero(); for (; b != e; ++b) coeffRef(b->row(), b->col()) += b->value(); }

  class InnerIterator {
   public:
    InnerIterator(SparseMatrix& m, Index outer) : col_(&m.cols_[(size_t)outer]), outer_(outer), p_(0) {}
    InnerIterator(const SparseMatrix& m, Index outer) : col_(const_cast<std::vector<std::pair<I, T> >*>(&m.cols_[(size_t)outer])), outer_(outer), p_(0) {}
    InnerIterator& operator++() { ++p_; return *this; }
    operator bool() const { return p_ < col_->size(); }
    Index row() const { return (Index)(*col_)[p_].first; }
    Index col() const { return outer_; }
    Index index() const { return row(); }
    const T& value() const { return (*col_)[p_].second; }
    T& valueRef() { return (*col_)[p_].second; }
   private:
    std::vector<std::pair<I, T> >* col_;
    Index outer_;
    size_t p_;
  };

 private:
  Index rows_;
  std::vector<std::vector<std::pair<I, T> > > cols_;
};

}  // namespace Eigen
#endif
