// sspp_spline.cpp — see sspp_spline.h
#include "sspp_spline.h"

#include <algorithm>
#include <cmath>
#include <stdexcept>

namespace sspp {

std::vector<double> knot_averaging(const std::vector<double>& u, int p) {
  const int n = (int)u.size();
  if (n < p + 1) throw std::runtime_error("spline needs at least degree+1 points");
  std::vector<double> k(n + p + 1, 0.0);
  for (int j = 1; j < n - p; ++j) {
    double s = 0.0;
    for (int i = 0; i < p; ++i) s += u[j + i];
    k[j + p] = s / p;
  }
  for (int i = 0; i <= p; ++i) { k[i] = 0.0; k[k.size() - 1 - i] = 1.0; }
  return k;
}

int find_span(double u, int p, const std::vector<double>& knots) {
  if (u <= knots[0]) return p;
  const double* b = knots.data() + p - 1;
  const double* e = knots.data() + knots.size() - p - 1;
  const double* pos = std::upper_bound(b, e, u);
  return (int)(pos - knots.data()) - 1;
}

std::vector<double> basis_row(double u, int p, const std::vector<double>& knots) {
  const int n = (int)knots.size() - p - 1;
  const int span = find_span(u, p, knots);
  std::vector<double> N(p + 1, 0.0), left(p + 1, 0.0), right(p + 1, 0.0);
  N[0] = 1.0;
  for (int j = 1; j <= p; ++j) {
    left[j] = u - knots[span + 1 - j];
    right[j] = knots[span + j] - u;
    double saved = 0.0;
    for (int r = 0; r < j; ++r) {
      const double tmp = N[r] / (right[r + 1] + left[j - r]);
      N[r] = saved + right[r + 1] * tmp;
      saved = left[j - r] * tmp;
    }
    N[j] = saved;
  }
  std::vector<double> row(n, 0.0);
  for (int j = 0; j <= p; ++j) row[span - p + j] = N[j];
  return row;
}

std::vector<double> interpolation_matrix(const std::vector<double>& u, int p, const std::vector<double>& knots) {
  const int n = (int)u.size();
  std::vector<double> A((size_t)n * n, 0.0);
  for (int i = 1; i < n - 1; ++i) {
    const std::vector<double> r = basis_row(u[i], p, knots);
    std::copy(r.begin(), r.end(), A.begin() + (size_t)i * n);
  }
  A[0] = 1.0;
  A[(size_t)n * n - 1] = 1.0;
  return A;
}

std::vector<double> invert(const std::vector<double>& Ain, int n) {
  std::vector<double> A(Ain), I((size_t)n * n, 0.0);
  for (int i = 0; i < n; ++i) I[(size_t)i * n + i] = 1.0;
  for (int c = 0; c < n; ++c) {
    int piv = c;
    for (int r = c + 1; r < n; ++r) if (std::fabs(A[(size_t)r * n + c]) > std::fabs(A[(size_t)piv * n + c])) piv = r;
    if (std::fabs(A[(size_t)piv * n + c]) < 1e-14) throw std::runtime_error("singular interpolation matrix");
    if (piv != c) for (int k = 0; k < n; ++k) { std::swap(A[(size_t)piv * n + k], A[(size_t)c * n + k]); std::swap(I[(size_t)piv * n + k], I[(size_t)c * n + k]); }
    const double d = A[(size_t)c * n + c];
    for (int k = 0; k < n; ++k) { A[(size_t)c * n + k] /= d; I[(size_t)c * n + k] /= d; }
    for (int r = 0; r < n; ++r) {
      if (r == c) continue;
      const double f = A[(size_t)r * n + c];
      if (f == 0.0) continue;
      for (int k = 0; k < n; ++k) { A[(size_t)r * n + k] -= f * A[(size_t)c * n + k]; I[(size_t)r * n + k] -= f * I[(size_t)c * n + k]; }
    }
  }
  return I;
}

SplineSetup make_setup(int K, int p) {
  if (K < p + 1) throw std::runtime_error("init_points must be at least degree + 1");
  SplineSetup s;
  s.K = K; s.p = p;
  s.u.resize(K);
  for (int i = 0; i < K; ++i) s.u[i] = (double)i / (K - 1);
  s.knots = knot_averaging(s.u, p);
  s.Ainv = invert(interpolation_matrix(s.u, p, s.knots), K);
  return s;
}

std::vector<double> collocation_matrix(const SplineSetup& s, const std::vector<double>& grid, bool via_form) {
  const int M = (int)grid.size(), K = s.K;
  std::vector<double> Psi((size_t)M * K, 0.0);
  for (int m = 0; m < M; ++m) {
    const std::vector<double> b = basis_row(grid[m], s.p, s.knots);
    if (!via_form) { std::copy(b.begin(), b.end(), Psi.begin() + (size_t)m * K); continue; }
    for (int j = 0; j < K; ++j) {
      double acc = 0.0;
      for (int k = 0; k < K; ++k) acc += b[k] * s.Ainv[(size_t)k * K + j];
      Psi[(size_t)m * K + j] = acc;
    }
  }
  return Psi;
}

std::vector<double> uniform_knots(int n_ctrl, int k) {
  if (k < 0 || n_ctrl < k + 1) throw std::runtime_error("knot_vector: need at least k + 1 control points");
  const int inner = n_ctrl + k + 1 - 2 * k;               // np.linspace(0, 1, n_knots - 2k)
  std::vector<double> t;
  t.reserve((size_t)n_ctrl + k + 1);
  for (int i = 0; i < k; ++i) t.push_back(0.0);
  const double step = inner > 1 ? 1.0 / (inner - 1) : 0.0;     // numpy.linspace: i * step, the last value set to the stop exactly
  for (int i = 0; i < inner; ++i) t.push_back((inner > 1 && i == inner - 1) ? 1.0 : (double)i * step);
  for (int i = 0; i < k; ++i) t.push_back(1.0);
  return t;
}

namespace {
// values of the degree-q functions B_{0,q} .. B_{len-q-2,q} at theta, raised degree by degree from the half-open
// indicator row; a vanishing denominator drops its term (BSplines.py:20-27)
std::vector<double> degree_row(double theta, int q, const std::vector<double>& t) {
  const int m = (int)t.size();
  std::vector<double> row((size_t)std::max(m - 1, 0), 0.0);
  for (int i = 0; i + 1 < m; ++i) row[i] = (t[i] <= theta && theta < t[i + 1]) ? 1.0 : 0.0;
  for (int j = 1; j <= q; ++j) {
    std::vector<double> nxt((size_t)std::max(m - j - 1, 0), 0.0);
    for (int i = 0; i < (int)nxt.size(); ++i) {
      const double d1 = t[i + j] - t[i], d2 = t[i + j + 1] - t[i + 1];
      const double c1 = d1 == 0.0 ? 0.0 : (theta - t[i]) / d1 * row[i];
      const double c2 = d2 == 0.0 ? 0.0 : (t[i + j + 1] - theta) / d2 * row[i + 1];
      nxt[i] = c1 + c2;
    }
    row.swap(nxt);
  }
  return row;
}
}  // namespace

std::vector<double> half_open_basis_row(double theta, int k, const std::vector<double>& t, bool deriv) {
  const int n = (int)t.size() - k - 1;
  if (n < 1) throw std::runtime_error("knot vector too short for the order");
  if (!deriv) { std::vector<double> r = degree_row(theta, k, t); r.resize(n); return r; }
  std::vector<double> out((size_t)n, 0.0);
  if (k == 0) return out;
  const std::vector<double> low = degree_row(theta, k - 1, t);
  for (int i = 0; i < n; ++i) {
    const double d1 = t[i + k] - t[i], d2 = t[i + k + 1] - t[i + 1];
    const double c1 = d1 == 0.0 ? 0.0 : k / d1 * low[i];
    const double c2 = d2 == 0.0 ? 0.0 : -k / d2 * low[i + 1];
    out[i] = c1 + c2;
  }
  return out;
}

std::vector<double> curve_row(double theta, int k, const std::vector<double>& t, bool deriv) {
  const int n = (int)t.size() - k - 1;
  if (deriv) return half_open_basis_row(theta, k, t, true);
  std::vector<double> row((size_t)n, 0.0);
  if (theta < 0.0) { row[0] = half_open_basis_row(0.0, k, t, false)[0]; return row; }
  if (theta >= 1.0) { row[n - 1] = 1.0; return row; }
  return half_open_basis_row(theta, k, t, false);
}

std::vector<double> fit_uniform(const std::vector<double>& via, int n, int d, int k, std::vector<double>* knots_out) {
  if (n < 2 || d < 1 || (int)via.size() != n * d) throw std::runtime_error("compute_control_points: via points must be [n >= 2][d]");
  const std::vector<double> t = uniform_knots(n, k);
  std::vector<double> A((size_t)n * n, 0.0);
  for (int i = 0; i < n; ++i) {
    const std::vector<double> r = half_open_basis_row((double)i / (n - 1), k, t, false);
    std::copy(r.begin(), r.end(), A.begin() + (size_t)i * n);
  }
  A[0] = 1.0;
  A[(size_t)n * n - 1] = 1.0;                             // the half-open basis vanishes at theta = 1: the reference pins the corner
  const std::vector<double> Ai = invert(A, n);            // square and well conditioned: the least-squares solution is the solution
  std::vector<double> c((size_t)n * d, 0.0);
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      const double a = Ai[(size_t)i * n + j];
      if (a == 0.0) continue;
      for (int e = 0; e < d; ++e) c[(size_t)i * d + e] += a * via[(size_t)j * d + e];
    }
  if (knots_out) *knots_out = t;
  return c;
}

}  // namespace sspp
