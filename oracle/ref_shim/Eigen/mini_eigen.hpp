// mini_eigen.hpp -- a minimal, eagerly evaluated stand-in for the part of the Eigen 3.4 API that
// the reference's HOST hot path uses (nlsolvers/common/include/laplacians.hpp,
// nlsolvers/host/include/eigen_krylov_{real,complex}.hpp, *_solver*.hpp, boundaries*.hpp).
//
// TEST INFRASTRUCTURE ONLY.  Eigen3 is an un-vendored dependency of the reference and is not in
// this image, so the reference's sources cannot be compiled as they are.  With this header on the
// include path they compile UNMODIFIED (oracle/Makefile, target _ref/libnls_ref.so), which pins the
// CPU restatement (nls_oracle.cpp) against the reference's own source text: operation order of the
// Lanczos process, function tables, step formulas, boundary copies, matrix assembly.  What it does
// NOT pin is Eigen's own arithmetic: products are plain loops here and SelfAdjointEigenSolver is an
// implicit-shift QL on the tridiagonal part (the published algorithm Eigen also uses after its
// tridiagonalisation), not Eigen's code.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <complex>
#include <cstdint>
#include <tuple>
#include <type_traits>
#include <vector>

namespace Eigen {

constexpr int Dynamic = -1;
enum NoChange_t { NoChange };

template <class T> struct NumTraits { using Real = T; };
template <class T> struct NumTraits<std::complex<T>> { using Real = T; };
template <class A, class B> using Common = decltype(std::declval<A>() * std::declval<B>());
inline double conj_(double a) { return a; }
inline std::complex<double> conj_(std::complex<double> a) { return std::conj(a); }
inline double abs2_(double a) { return a * a; }
inline double abs2_(std::complex<double> a) { return a.real() * a.real() + a.imag() * a.imag(); }

template <class T> class VectorX;
template <class T> class MatrixX;

template <class T> struct Diag { std::vector<T> d; };
template <class T> struct RowVec { std::vector<T> d; }; // adjoint()/transpose() of a column vector

// strided view (matrix rows/columns, segments): assignable, readable
template <class T> struct Strided {
  T *p;
  long n, stride;
  Strided segment(long a, long len) const { return Strided{p + a * stride, len, stride}; }
  Strided &operator=(const Strided &o) {
    assert(n == o.n);
    std::vector<T> tmp(n);
    for (long i = 0; i < n; ++i) tmp[i] = o.p[i * o.stride];
    for (long i = 0; i < n; ++i) p[i * stride] = tmp[i];
    return *this;
  }
  Strided &operator=(const VectorX<T> &v);
  void setZero() { for (long i = 0; i < n; ++i) p[i * stride] = T(0); }
  VectorX<T> vec() const;
  RowVec<T> adjoint() const { RowVec<T> r; r.d.resize(n); for (long i = 0; i < n; ++i) r.d[i] = conj_(p[i * stride]); return r; }
  double norm() const { double s = 0; for (long i = 0; i < n; ++i) s += abs2_(p[i * stride]); return std::sqrt(s); }
};

template <class T> class VectorX {
public:
  using Scalar = T;
  using Real = typename NumTraits<T>::Real;
  std::vector<T> v;
  VectorX() {}
  explicit VectorX(long n) : v(n) {}
  template <class U, class = typename std::enable_if<!std::is_same<U, T>::value && std::is_convertible<U, T>::value>::type>
  VectorX(const VectorX<U> &o) : v(o.v.begin(), o.v.end()) {} // e.g. real expression assigned to a complex vector
  VectorX(const Strided<T> &s) : v(s.n) { for (long i = 0; i < s.n; ++i) v[i] = s.p[i * s.stride]; }
  static VectorX Zero(long n) { VectorX r(n); std::fill(r.v.begin(), r.v.end(), T(0)); return r; }
  static VectorX Ones(long n) { VectorX r(n); std::fill(r.v.begin(), r.v.end(), T(1)); return r; }
  long size() const { return (long)v.size(); }
  long rows() const { return (long)v.size(); }
  T *data() { return v.data(); }
  const T *data() const { return v.data(); }
  T &operator[](long i) { return v[i]; }
  const T &operator[](long i) const { return v[i]; }
  T &operator()(long i) { return v[i]; }
  const T &operator()(long i) const { return v[i]; }
  double norm() const { double s = 0; for (auto &x : v) s += abs2_(x); return std::sqrt(s); }
  // Eigen: a.dot(b) = sum conj(a_i) b_i
  T dot(const VectorX &o) const { T s = T(0); for (long i = 0; i < size(); ++i) s += conj_(v[i]) * o.v[i]; return s; }
  T dot(const Strided<T> &o) const { T s = T(0); for (long i = 0; i < size(); ++i) s += conj_(v[i]) * o.p[i * o.stride]; return s; }
  RowVec<T> adjoint() const { RowVec<T> r; r.d.resize(v.size()); for (size_t i = 0; i < v.size(); ++i) r.d[i] = conj_(v[i]); return r; }
  RowVec<T> transpose() const { RowVec<T> r; r.d = v; return r; }
  VectorX<Real> real() const { VectorX<Real> r(size()); for (long i = 0; i < size(); ++i) r.v[i] = std::real(v[i]); return r; }
  VectorX<Real> imag() const { VectorX<Real> r(size()); for (long i = 0; i < size(); ++i) r.v[i] = std::imag(v[i]); return r; }
  template <class U> VectorX<Common<T, U>> cwiseProduct(const VectorX<U> &o) const {
    VectorX<Common<T, U>> r(size());
    for (long i = 0; i < size(); ++i) r.v[i] = v[i] * o.v[i];
    return r;
  }
  VectorX cwiseInverse() const { VectorX r(size()); for (long i = 0; i < size(); ++i) r.v[i] = T(1) / v[i]; return r; }
  template <class F> auto unaryExpr(F f) const -> VectorX<decltype(f(std::declval<T>()))> {
    VectorX<decltype(f(std::declval<T>()))> r(size());
    for (long i = 0; i < size(); ++i) r.v[i] = f(v[i]);
    return r;
  }
  // array()/matrix() views are the vector itself (coefficient-wise API below)
  const VectorX &array() const { return *this; }
  const VectorX &matrix() const { return *this; }
  VectorX &noalias() { return *this; }
  VectorX<Real> abs() const { VectorX<Real> r(size()); for (long i = 0; i < size(); ++i) r.v[i] = std::abs(v[i]); return r; }
  VectorX sqrt() const { VectorX r(size()); for (long i = 0; i < size(); ++i) r.v[i] = std::sqrt(v[i]); return r; }
  VectorX cos() const { VectorX r(size()); for (long i = 0; i < size(); ++i) r.v[i] = std::cos(v[i]); return r; }
  VectorX square() const { VectorX r(size()); for (long i = 0; i < size(); ++i) r.v[i] = v[i] * v[i]; return r; }
  Diag<T> asDiagonal() const { return Diag<T>{v}; }
  VectorX &operator+=(const VectorX &o) { for (long i = 0; i < size(); ++i) v[i] += o.v[i]; return *this; }
  VectorX &operator-=(const VectorX &o) { for (long i = 0; i < size(); ++i) v[i] -= o.v[i]; return *this; }
  VectorX operator-() const { VectorX r(size()); for (long i = 0; i < size(); ++i) r.v[i] = -v[i]; return r; }
};

template <class T> Strided<T> &Strided<T>::operator=(const VectorX<T> &o) {
  assert(n == o.size());
  for (long i = 0; i < n; ++i) p[i * stride] = o.v[i];
  return *this;
}
template <class T> VectorX<T> Strided<T>::vec() const { return VectorX<T>(*this); }

template <class A, class B> VectorX<Common<A, B>> operator+(const VectorX<A> &a, const VectorX<B> &b) {
  VectorX<Common<A, B>> r(a.size());
  for (long i = 0; i < a.size(); ++i) r.v[i] = a.v[i] + b.v[i];
  return r;
}
template <class A, class B> VectorX<Common<A, B>> operator-(const VectorX<A> &a, const VectorX<B> &b) {
  VectorX<Common<A, B>> r(a.size());
  for (long i = 0; i < a.size(); ++i) r.v[i] = a.v[i] - b.v[i];
  return r;
}
template <class S> struct is_scalar : std::integral_constant<bool, std::is_arithmetic<S>::value> {};
template <class S> struct is_scalar<std::complex<S>> : std::true_type {};
// scalar (of any arithmetic/complex type) times vector, both orders; int/float scalars act as double
template <class S> using AsNum = typename std::conditional<std::is_arithmetic<S>::value, double, S>::type;
template <class S, class T, class = typename std::enable_if<is_scalar<S>::value>::type>
VectorX<Common<AsNum<S>, T>> operator*(const S &s, const VectorX<T> &a) {
  VectorX<Common<AsNum<S>, T>> r(a.size());
  for (long i = 0; i < a.size(); ++i) r.v[i] = AsNum<S>(s) * a.v[i];
  return r;
}
template <class S, class T, class = typename std::enable_if<is_scalar<S>::value>::type>
VectorX<Common<AsNum<S>, T>> operator*(const VectorX<T> &a, const S &s) { return s * a; }
template <class S, class T, class = typename std::enable_if<is_scalar<S>::value>::type>
VectorX<Common<AsNum<S>, T>> operator/(const VectorX<T> &a, const S &s) {
  VectorX<Common<AsNum<S>, T>> r(a.size());
  for (long i = 0; i < a.size(); ++i) r.v[i] = a.v[i] / AsNum<S>(s);
  return r;
}
template <class S, class T, class = typename std::enable_if<is_scalar<S>::value>::type>
VectorX<Common<AsNum<S>, T>> operator*(const S &s, const Strided<T> &a) { return s * a.vec(); }
// row-vector times column vector = scalar (V.col(j).adjoint() * w)
template <class T> T operator*(const RowVec<T> &r, const VectorX<T> &c) {
  T s = T(0);
  for (size_t i = 0; i < r.d.size(); ++i) s += r.d[i] * c.v[i];
  return s;
}

// ---------------------------------------------------------------- dense matrix (column major)
template <class T> class MatrixX {
public:
  using Scalar = T;
  std::vector<T> a;
  long r = 0, c = 0;
  MatrixX() {}
  MatrixX(long rr, long cc) : a((size_t)rr * cc), r(rr), c(cc) {}
  static MatrixX Zero(long rr, long cc) { MatrixX m(rr, cc); std::fill(m.a.begin(), m.a.end(), T(0)); return m; }
  long rows() const { return r; }
  long cols() const { return c; }
  T *data() { return a.data(); }
  T &operator()(long i, long j) { return a[(size_t)j * r + i]; }
  const T &operator()(long i, long j) const { return a[(size_t)j * r + i]; }
  Strided<T> col(long j) { return Strided<T>{a.data() + (size_t)j * r, r, 1}; }
  Strided<T> col(long j) const { return Strided<T>{const_cast<T *>(a.data()) + (size_t)j * r, r, 1}; }
  Strided<T> row(long i) { return Strided<T>{a.data() + i, c, r}; }
  MatrixX adjoint() const { MatrixX m(c, r); for (long i = 0; i < r; ++i) for (long j = 0; j < c; ++j) m(j, i) = conj_((*this)(i, j)); return m; }
  MatrixX transpose() const { MatrixX m(c, r); for (long i = 0; i < r; ++i) for (long j = 0; j < c; ++j) m(j, i) = (*this)(i, j); return m; }
  template <class A, class B> void conservativeResize(A, B) { assert(false && "breakdown branch is commented out in the reference"); }
};
template <class T, class D> MatrixX<Common<T, D>> operator*(const MatrixX<T> &m, const Diag<D> &d) {
  MatrixX<Common<T, D>> o(m.r, m.c);
  for (long j = 0; j < m.c; ++j) for (long i = 0; i < m.r; ++i) o(i, j) = m(i, j) * d.d[j];
  return o;
}
template <class A, class B> MatrixX<Common<A, B>> operator*(const MatrixX<A> &x, const MatrixX<B> &y) {
  assert(x.c == y.r);
  MatrixX<Common<A, B>> o = MatrixX<Common<A, B>>::Zero(x.r, y.c);
  for (long j = 0; j < y.c; ++j)
    for (long k = 0; k < x.c; ++k) {
      const B ykj = y(k, j);
      for (long i = 0; i < x.r; ++i) o(i, j) += x(i, k) * ykj;
    }
  return o;
}
template <class A, class B> VectorX<Common<A, B>> operator*(const MatrixX<A> &x, const VectorX<B> &y) {
  assert(x.c == y.size());
  VectorX<Common<A, B>> o = VectorX<Common<A, B>>::Zero(x.r);
  for (long k = 0; k < x.c; ++k) for (long i = 0; i < x.r; ++i) o.v[i] += x(i, k) * y.v[k];
  return o;
}
template <class S, class T, class = typename std::enable_if<is_scalar<S>::value>::type>
MatrixX<Common<AsNum<S>, T>> operator*(const S &s, const MatrixX<T> &m) {
  MatrixX<Common<AsNum<S>, T>> o(m.r, m.c);
  for (size_t i = 0; i < m.a.size(); ++i) o.a[i] = AsNum<S>(s) * m.a[i];
  return o;
}

// Eigen::Matrix<T, Dynamic, Dynamic> and Eigen::Map of it (boundaries.hpp)
template <class T, int R, int C> struct Matrix { using Scalar = T; };
template <class M> class Map {
public:
  using T = typename M::Scalar;
  T *p;
  long r, c;
  Map(T *d, long rr, long cc) : p(d), r(rr), c(cc) {}
  Strided<T> row(long i) { return Strided<T>{p + i, c, r}; }
  Strided<T> col(long j) { return Strided<T>{p + (size_t)j * r, r, 1}; }
  struct Block { T *p; long r0, c0, nr, nc, ld; void setZero() { for (long j = 0; j < nc; ++j) for (long i = 0; i < nr; ++i) p[(c0 + j) * ld + r0 + i] = T(0); } };
  Block block(long i, long j, long nr, long nc) { return Block{p, i, j, nr, nc, r}; }
};

// ---------------------------------------------------------------- sparse matrix
template <class T> struct Triplet {
  long r, c;
  T v;
  Triplet(long rr, long cc, const T &vv) : r(rr), c(cc), v(vv) {}
  long row() const { return r; }
  long col() const { return c; }
  const T &value() const { return v; }
};
// compressed COLUMN storage like Eigen's default; duplicates summed by setFromTriplets
template <class T> class SparseMatrix {
public:
  using Scalar = T;
  long nr = 0, nc = 0;
  std::vector<long> outer; // nc+1
  std::vector<long> inner; // row indices, sorted within a column
  std::vector<T> val;
  SparseMatrix() {}
  SparseMatrix(long r, long c) : nr(r), nc(c), outer(c + 1, 0) {}
  long rows() const { return nr; }
  long cols() const { return nc; }
  long nonZeros() const { return (long)val.size(); }
  template <class It> void setFromTriplets(It b, It e) {
    std::vector<std::tuple<long, long, T>> t;
    for (It i = b; i != e; ++i) t.emplace_back(i->col(), i->row(), i->value());
    std::stable_sort(t.begin(), t.end(), [](const auto &x, const auto &y) {
      return std::get<0>(x) != std::get<0>(y) ? std::get<0>(x) < std::get<0>(y) : std::get<1>(x) < std::get<1>(y);
    });
    outer.assign(nc + 1, 0);
    inner.clear();
    val.clear();
    for (size_t k = 0; k < t.size();) {
      size_t q = k;
      T s = T(0);
      while (q < t.size() && std::get<0>(t[q]) == std::get<0>(t[k]) && std::get<1>(t[q]) == std::get<1>(t[k])) s += std::get<2>(t[q++]);
      inner.push_back(std::get<1>(t[k]));
      val.push_back(s);
      outer[std::get<0>(t[k]) + 1]++;
      k = q;
    }
    for (long j = 0; j < nc; ++j) outer[j + 1] += outer[j];
  }
  void makeCompressed() {}
  SparseMatrix &operator*=(const T &s) { for (auto &x : val) x *= s; return *this; }
  // y = A x, column by column (Eigen's column-major sparse * dense kernel)
  template <class X> VectorX<T> times(const X &x) const {
    VectorX<T> y = VectorX<T>::Zero(nr);
    for (long j = 0; j < nc; ++j) {
      const T xj = x(j);
      for (long k = outer[j]; k < outer[j + 1]; ++k) y.v[inner[k]] += val[k] * xj;
    }
    return y;
  }
};
template <class T> struct StridedAt { const Strided<T> &s; T operator()(long i) const { return s.p[i * s.stride]; } };
template <class T> VectorX<T> operator*(const SparseMatrix<T> &A, const VectorX<T> &x) { return A.times(x); }
template <class T> VectorX<T> operator*(const SparseMatrix<T> &A, const Strided<T> &x) { return A.times(StridedAt<T>{x}); }
template <class T> SparseMatrix<T> operator*(const T &s, const SparseMatrix<T> &A) { SparseMatrix<T> B = A; B *= s; return B; }
template <class M> class SparseLU { // only named by an un-instantiated helper (phi1m_multiply)
public:
  void compute(const M &) {}
  template <class V> V solve(const V &v) const { assert(false && "SparseLU is not part of the hot path"); return v; }
};

// ---------------------------------------------------------------- SelfAdjointEigenSolver
// Implicit-shift QL on the symmetric tridiagonal part of the (lower triangle of the) input.
// The reference only ever passes the Lanczos matrix T, which is tridiagonal; anything else asserts.
template <class M> class SelfAdjointEigenSolver {
public:
  using T = typename M::Scalar;
  using Real = typename NumTraits<T>::Real;
  explicit SelfAdjointEigenSolver(const M &A) {
    const long m = A.rows();
    std::vector<double> d(m), e(m, 0.0), Z((size_t)m * m, 0.0);
    for (long i = 0; i < m; ++i) {
      d[i] = std::real(A(i, i));
      Z[(size_t)i * m + i] = 1.0;
      if (i + 1 < m) e[i] = std::real(A(i + 1, i));
      for (long j = 0; j + 1 < i; ++j) assert(std::abs(A(i, j)) == 0.0 && "tridiagonal input expected");
    }
    tql2(m, d, e, Z);
    // ascending order like Eigen
    std::vector<long> idx(m);
    for (long i = 0; i < m; ++i) idx[i] = i;
    std::sort(idx.begin(), idx.end(), [&](long x, long y) { return d[x] < d[y]; });
    lam = VectorX<Real>(m);
    Q = M(m, m);
    for (long k = 0; k < m; ++k) {
      lam.v[k] = d[idx[k]];
      for (long i = 0; i < m; ++i) Q(i, k) = T(Z[(size_t)i * m + idx[k]]);
    }
  }
  const VectorX<Real> &eigenvalues() const { return lam; }
  const M &eigenvectors() const { return Q; }

private:
  VectorX<Real> lam;
  M Q;
  static void tql2(long n, std::vector<double> &d, std::vector<double> &e, std::vector<double> &z) {
    for (long l = 0; l < n; ++l) {
      int iter = 0;
      long m;
      do {
        for (m = l; m < n - 1; ++m) {
          const double dd = std::abs(d[m]) + std::abs(d[m + 1]);
          if (std::abs(e[m]) <= 2.220446049250313e-16 * dd) break;
        }
        if (m != l) {
          assert(iter++ < 60);
          double g = (d[l + 1] - d[l]) / (2.0 * e[l]);
          double r = std::hypot(g, 1.0);
          g = d[m] - d[l] + e[l] / (g + std::copysign(r, g));
          double s = 1.0, c = 1.0, p = 0.0;
          long i;
          for (i = m - 1; i >= l; --i) {
            double f = s * e[i];
            const double b = c * e[i];
            e[i + 1] = (r = std::hypot(f, g));
            if (r == 0.0) { d[i + 1] -= p; e[m] = 0.0; break; }
            s = f / r; c = g / r;
            g = d[i + 1] - p;
            r = (d[i] - g) * s + 2.0 * c * b;
            d[i + 1] = g + (p = s * r);
            g = c * r - b;
            for (long k = 0; k < n; ++k) {
              f = z[(size_t)k * n + i + 1];
              z[(size_t)k * n + i + 1] = s * z[(size_t)k * n + i] + c * f;
              z[(size_t)k * n + i] = c * z[(size_t)k * n + i] - s * f;
            }
          }
          if (r == 0.0 && i >= l) continue;
          d[l] -= p; e[l] = g; e[m] = 0.0;
        }
      } while (m != l);
    }
  }
};

using VectorXd = VectorX<double>;
using VectorXcd = VectorX<std::complex<double>>;
using MatrixXd = MatrixX<double>;
using MatrixXcd = MatrixX<std::complex<double>>;

} // namespace Eigen
