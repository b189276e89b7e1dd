// sspp_spline.h — host FP64 B-spline algebra used once per planner to build the collocation
// matrices the kernel applies (q_m = Psi[m,:] . X).  Restates the parts of Eigen's unsupported
// Spline module the reference calls (third party, unpinned, not in the reference tree):
//   Spline::Span / BasisFunctions   <- /root/reference/include/sspp.h:101,137; tsp_evaluator.h:21,24
//   KnotAveraging + SplineFitting::Interpolate <- include/sspp.h:95; include/sspp/tsp_path_model.h:25-28,39-42
// (Piegl & Tiller, The NURBS Book, A2.1 FindSpan, A2.2 BasisFuns, eq. 9.8 knot averaging.)
#pragma once
#include <vector>

namespace sspp {

// knots of Eigen::KnotAveraging for interpolation parameters u[0..n) and degree p: n + p + 1 values
std::vector<double> knot_averaging(const std::vector<double>& u, int p);

// span index i with knots[i] <= u < knots[i+1] (last span for u == knots.back()), as Eigen::Spline::Span
int find_span(double u, int p, const std::vector<double>& knots);

// full basis row N_0(u)..N_{n-1}(u), n = knots.size() - p - 1
std::vector<double> basis_row(double u, int p, const std::vector<double>& knots);

// n x n collocation matrix of SplineFitting::Interpolate (rows 0 and n-1 are unit rows), row-major
std::vector<double> interpolation_matrix(const std::vector<double>& u, int p, const std::vector<double>& knots);

// dense inverse by Gauss-Jordan with partial pivoting; throws std::runtime_error when singular
std::vector<double> invert(const std::vector<double>& A, int n);

struct SplineSetup {
  int K = 0, p = 0;
  std::vector<double> u;        // interpolation parameters i/(K-1)
  std::vector<double> knots;    // K + p + 1
  std::vector<double> Ainv;     // K x K : control points = Ainv * interpolation points
};
SplineSetup make_setup(int K, int p);

// Psi for evaluation parameters `grid`:
//   control-point form (planner A): Psi = Basis(grid)           [M x K]
//   via-point form     (planner B): Psi = Basis(grid) * Ainv    [M x K]
std::vector<double> collocation_matrix(const SplineSetup& s, const std::vector<double>& grid, bool via_form);

// ---- the reference's Python spline twins (sspp/BSplines.py, sspp/CubicPath.py), host FP64 ----
// clamped UNIFORM knot vector of BSplines.py:58-62: k zeros, linspace(0, 1, n - k + 1), k ones  (n + k + 1 values)
std::vector<double> uniform_knots(int n_ctrl, int k);
// Cox-de Boor basis with HALF-OPEN spans as BSplines.py:11-29 defines it (B_{i,0} = 1 iff t[i] <= theta < t[i+1], so every
// value is 0 at theta = t.back()); deriv: BSplines.py:31-42.  Full row of n = knots.size() - k - 1 values.
std::vector<double> half_open_basis_row(double theta, int k, const std::vector<double>& knots, bool deriv);
// row that gives bspline(theta, t, c, k) = row . c with the end rules of BSplines.py:44-51 (theta < 0: c[0] B_0(0);
// theta >= 1: c[n-1]); deriv: bspline_derivative, BSplines.py:53-55 (no end rules)
std::vector<double> curve_row(double theta, int k, const std::vector<double>& knots, bool deriv);
// compute_control_points, BSplines.py:65-106: via [n][d] row-major at theta_i = i/(n-1) on uniform knots -> ctrl [n][d]
std::vector<double> fit_uniform(const std::vector<double>& via, int n, int d, int k, std::vector<double>* knots_out);

}  // namespace sspp
