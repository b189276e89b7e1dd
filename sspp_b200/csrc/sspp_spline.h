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

}  // namespace sspp
