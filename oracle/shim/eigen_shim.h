// eigen_shim.h — TEST INFRASTRUCTURE.  A stand-in for the slice of Eigen 3 that the reference's hot-path
// headers use (include/sspp.h, Collision.h, utility.h, sspp/tsp_*.h), so that those headers compile
// UNMODIFIED into oracle/_ref/libref.so in a container that has no Eigen.  Written from Eigen's public API
// (names, argument meaning, column-major storage); no Eigen source is involved.  Dense arithmetic only, no
// expression templates: every operator returns a value.
//
// Covered: Matrix<double,R,C> with R,C fixed or Dynamic; Ones/Zero/UnitX/Y/Z; (i), (i,j), [i]; + - unary-,
// scalar products, += -= *=; cwiseMax/cwiseMin/cwiseSqrt/cwiseProduct; norm; size/rows/cols/resize/setZero;
// comma initialiser; col(i) = v; Map<VectorXd>; Quaterniond, AngleAxisd, toRotationMatrix, eulerAngles.
#pragma once
#include <array>
#include <cmath>
#include <cstddef>
#include <type_traits>
#include <vector>

namespace Eigen {

using Index = std::ptrdiff_t;
using DenseIndex = std::ptrdiff_t;
constexpr int Dynamic = -1;

template <class T, int R, int C>
class Matrix {
  static constexpr bool kFixed = (R != Dynamic && C != Dynamic);
  using Store = std::conditional_t<kFixed, std::array<T, (kFixed ? R * C : 1)>, std::vector<T>>;
  Store a_{};
  Index r_ = (R == Dynamic ? 0 : R), c_ = (C == Dynamic ? 0 : C);

 public:
  using Scalar = T;
  Matrix() { if constexpr (kFixed) a_.fill(T(0)); }
  // (rows, cols) for matrices; Eigen also accepts (n, 1) on a column vector type
  Matrix(Index r, Index c) { resize(r, c); }
  explicit Matrix(Index n) { if constexpr (C == 1) resize(n, 1); else resize(1, n); }
  template <int R2, int C2>
  Matrix(const Matrix<T, R2, C2>& o) { resize(o.rows(), o.cols()); for (Index i = 0; i < size(); ++i) a_[i] = o.data()[i]; }
  Matrix(T x, T y, T z) { static_assert(kFixed && R * C == 3, "3-vector constructor"); a_[0] = x; a_[1] = y; a_[2] = z; }

  void resize(Index r, Index c) {
    if constexpr (kFixed) { (void)r; (void)c; }
    else { r_ = (R == Dynamic ? r : R); c_ = (C == Dynamic ? c : C); a_.assign((size_t)(r_ * c_), T(0)); }
  }
  void resize(Index n) { if constexpr (C == 1) resize(n, 1); else resize(1, n); }
  Index rows() const { return r_; }
  Index cols() const { return c_; }
  Index size() const { return r_ * c_; }
  T* data() { return a_.data(); }
  const T* data() const { return a_.data(); }
  void setZero() { for (Index i = 0; i < size(); ++i) a_[i] = T(0); }

  T& operator()(Index i) { return a_[i]; }
  const T& operator()(Index i) const { return a_[i]; }
  T& operator[](Index i) { return a_[i]; }
  const T& operator[](Index i) const { return a_[i]; }
  T& operator()(Index i, Index j) { return a_[i + j * r_]; }          // column-major, like Eigen's default
  const T& operator()(Index i, Index j) const { return a_[i + j * r_]; }
  T x() const { return a_[0]; }
  T y() const { return a_[1]; }
  T z() const { return a_[2]; }

  static Matrix Constant(T v) { Matrix m; for (Index i = 0; i < m.size(); ++i) m.a_[i] = v; return m; }
  static Matrix Ones() { return Constant(T(1)); }
  static Matrix Zero() { return Constant(T(0)); }
  static Matrix Zero(Index r, Index c) { Matrix m(r, c); return m; }
  static Matrix Unit(Index k) { Matrix m = Zero(); m.a_[k] = T(1); return m; }
  static Matrix UnitX() { return Unit(0); }
  static Matrix UnitY() { return Unit(1); }
  static Matrix UnitZ() { return Unit(2); }

  Matrix operator-() const { Matrix m(*this); for (Index i = 0; i < size(); ++i) m.a_[i] = -a_[i]; return m; }
  template <int R2, int C2> Matrix& operator+=(const Matrix<T, R2, C2>& o) { for (Index i = 0; i < size(); ++i) a_[i] += o.data()[i]; return *this; }
  template <int R2, int C2> Matrix& operator-=(const Matrix<T, R2, C2>& o) { for (Index i = 0; i < size(); ++i) a_[i] -= o.data()[i]; return *this; }
  Matrix& operator*=(T s) { for (Index i = 0; i < size(); ++i) a_[i] *= s; return *this; }
  Matrix& operator/=(T s) { for (Index i = 0; i < size(); ++i) a_[i] /= s; return *this; }

  template <class F> Matrix zip(const Matrix& o, F f) const { Matrix m(*this); for (Index i = 0; i < size(); ++i) m.a_[i] = f(a_[i], o.a_[i]); return m; }
  Matrix cwiseMax(const Matrix& o) const { return zip(o, [](T p, T q) { return p < q ? q : p; }); }
  Matrix cwiseMin(const Matrix& o) const { return zip(o, [](T p, T q) { return q < p ? q : p; }); }
  Matrix cwiseProduct(const Matrix& o) const { return zip(o, [](T p, T q) { return p * q; }); }
  Matrix cwiseSqrt() const { Matrix m(*this); for (Index i = 0; i < size(); ++i) m.a_[i] = std::sqrt(a_[i]); return m; }
  T squaredNorm() const { T s = T(0); for (Index i = 0; i < size(); ++i) s += a_[i] * a_[i]; return s; }
  T norm() const { return std::sqrt(squaredNorm()); }
  T mean() const { T s = T(0); for (Index i = 0; i < size(); ++i) s += a_[i]; return s / T(size()); }

  // v << a, b, c;
  struct Comma {
    Matrix* m; Index at;
    Comma& operator,(T v) { m->a_[at++] = v; return *this; }
  };
  Comma operator<<(T v) { a_[0] = v; return Comma{this, 1}; }

  // M.col(j) = v   /   v = M.col(j)
  struct ColRef {
    Matrix* m; Index j;
    template <int R2, int C2> ColRef& operator=(const Matrix<T, R2, C2>& v) { for (Index i = 0; i < m->r_; ++i) (*m)(i, j) = v.data()[i]; return *this; }
    operator Matrix<T, R, 1>() const { Matrix<T, R, 1> v; v.resize(m->r_, 1); for (Index i = 0; i < m->r_; ++i) v(i) = (*m)(i, j); return v; }
  };
  ColRef col(Index j) { return ColRef{this, j}; }
  Matrix<T, R, 1> col(Index j) const { Matrix<T, R, 1> v; v.resize(r_, 1); for (Index i = 0; i < r_; ++i) v(i) = (*this)(i, j); return v; }

  // rotation matrix -> Euler angles about axes (a0, a1, a2), Eigen's MatrixBase::eulerAngles convention:
  // first angle in [0, pi], the other two in [-pi, pi]
  Matrix<T, 3, 1> eulerAngles(Index a0, Index a1, Index a2) const {
    static_assert(kFixed && R == 3 && C == 3, "eulerAngles is a 3x3 operation");
    const Matrix& M = *this;
    Matrix<T, 3, 1> res;
    const Index odd = ((a0 + 1) % 3 == a1) ? 0 : 1;
    const Index i = a0, j = (a0 + 1 + odd) % 3, k = (a0 + 2 - odd) % 3;
    const T pi = T(3.14159265358979323846);
    if (a0 == a2) {
      res[0] = std::atan2(M(j, i), M(k, i));
      if ((odd && res[0] < T(0)) || (!odd && res[0] > T(0))) {
        res[0] = (res[0] > T(0)) ? res[0] - pi : res[0] + pi;
        const T s2 = std::sqrt(M(j, i) * M(j, i) + M(k, i) * M(k, i));
        res[1] = -std::atan2(s2, M(i, i));
      } else {
        const T s2 = std::sqrt(M(j, i) * M(j, i) + M(k, i) * M(k, i));
        res[1] = std::atan2(s2, M(i, i));
      }
      const T s1 = std::sin(res[0]), c1 = std::cos(res[0]);
      res[2] = std::atan2(c1 * M(j, k) - s1 * M(k, k), c1 * M(j, j) - s1 * M(k, j));
    } else {
      res[0] = std::atan2(M(j, k), M(k, k));
      const T c2 = std::sqrt(M(i, i) * M(i, i) + M(i, j) * M(i, j));
      if ((odd && res[0] < T(0)) || (!odd && res[0] > T(0))) {
        res[0] = (res[0] > T(0)) ? res[0] - pi : res[0] + pi;
        res[1] = std::atan2(-M(i, k), -c2);
      } else res[1] = std::atan2(-M(i, k), c2);
      const T s1 = std::sin(res[0]), c1 = std::cos(res[0]);
      res[2] = std::atan2(s1 * M(k, i) - c1 * M(j, i), c1 * M(j, j) - s1 * M(k, j));
    }
    if (!odd) res = -res;
    return res;
  }
};

template <class T, int R, int C, int R2, int C2>
Matrix<T, R, C> operator+(const Matrix<T, R, C>& a, const Matrix<T, R2, C2>& b) { Matrix<T, R, C> m(a); m += b; return m; }
template <class T, int R, int C, int R2, int C2>
Matrix<T, R, C> operator-(const Matrix<T, R, C>& a, const Matrix<T, R2, C2>& b) { Matrix<T, R, C> m(a); m -= b; return m; }
template <class T, int R, int C> Matrix<T, R, C> operator*(const Matrix<T, R, C>& a, T s) { Matrix<T, R, C> m(a); m *= s; return m; }
template <class T, int R, int C> Matrix<T, R, C> operator*(T s, const Matrix<T, R, C>& a) { Matrix<T, R, C> m(a); m *= s; return m; }
template <class T, int R, int C> Matrix<T, R, C> operator/(const Matrix<T, R, C>& a, T s) { Matrix<T, R, C> m(a); m /= s; return m; }
// int * matrix (the reference writes `(1 - t) * start` with t double, but `-Point::Ones()*2.0` style also occurs)
template <class T, int R, int C> Matrix<T, R, C> operator*(int s, const Matrix<T, R, C>& a) { return T(s) * a; }
template <class T, int R, int C> Matrix<T, R, C> operator*(const Matrix<T, R, C>& a, int s) { return a * T(s); }

using MatrixXd = Matrix<double, Dynamic, Dynamic>;
using VectorXd = Matrix<double, Dynamic, 1>;
using Vector3d = Matrix<double, 3, 1>;
using Vector4d = Matrix<double, 4, 1>;
using Matrix3d = Matrix<double, 3, 3>;

// Map<VectorXd>(ptr, n): a view; only element access and size are needed by the call sites
template <class V>
class Map {
  using T = typename std::remove_const_t<V>::Scalar;
  const T* p_; Index n_;
 public:
  Map(const T* p, Index n) : p_(p), n_(n) {}
  Index size() const { return n_; }
  const T& operator()(Index i) const { return p_[i]; }
  const T& operator[](Index i) const { return p_[i]; }
};

// ---- Geometry ----
template <class T> class Quaternion;
template <class T>
class AngleAxis {
 public:
  T angle; Matrix<T, 3, 1> axis;
  AngleAxis(T a, const Matrix<T, 3, 1>& ax) : angle(a), axis(ax) {}
  Quaternion<T> toQuat() const;
};
template <class T>
class Quaternion {
  T w_, x_, y_, z_;
 public:
  Quaternion() : w_(1), x_(0), y_(0), z_(0) {}
  Quaternion(T w, T x, T y, T z) : w_(w), x_(x), y_(y), z_(z) {}
  Quaternion(const AngleAxis<T>& aa) { *this = aa.toQuat(); }
  T w() const { return w_; }
  T x() const { return x_; }
  T y() const { return y_; }
  T z() const { return z_; }
  Quaternion operator*(const Quaternion& b) const {
    return Quaternion(w_ * b.w_ - x_ * b.x_ - y_ * b.y_ - z_ * b.z_, w_ * b.x_ + x_ * b.w_ + y_ * b.z_ - z_ * b.y_,
                      w_ * b.y_ + y_ * b.w_ + z_ * b.x_ - x_ * b.z_, w_ * b.z_ + z_ * b.w_ + x_ * b.y_ - y_ * b.x_);
  }
  Quaternion operator*(const AngleAxis<T>& b) const { return *this * b.toQuat(); }
  Matrix<T, 3, 3> toRotationMatrix() const {
    Matrix<T, 3, 3> R;
    const T tx = 2 * x_, ty = 2 * y_, tz = 2 * z_;
    const T twx = tx * w_, twy = ty * w_, twz = tz * w_, txx = tx * x_, txy = ty * x_, txz = tz * x_, tyy = ty * y_, tyz = tz * y_, tzz = tz * z_;
    R(0, 0) = 1 - (tyy + tzz); R(0, 1) = txy - twz; R(0, 2) = txz + twy;
    R(1, 0) = txy + twz; R(1, 1) = 1 - (txx + tzz); R(1, 2) = tyz - twx;
    R(2, 0) = txz - twy; R(2, 1) = tyz + twx; R(2, 2) = 1 - (txx + tyy);
    return R;
  }
};
template <class T>
Quaternion<T> AngleAxis<T>::toQuat() const {
  const T h = angle / 2, s = std::sin(h);
  return Quaternion<T>(std::cos(h), s * axis[0], s * axis[1], s * axis[2]);
}
template <class T> Quaternion<T> operator*(const AngleAxis<T>& a, const AngleAxis<T>& b) { return a.toQuat() * b.toQuat(); }
using Quaterniond = Quaternion<double>;
using AngleAxisd = AngleAxis<double>;

}  // namespace Eigen
