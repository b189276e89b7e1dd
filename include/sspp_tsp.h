// sspp_tsp.h — host-side mirror of the reference's task-space planner (planner B) over the C ABI.
//
//   tsp::TaskSpacePlanner   <- /root/reference/include/sspp/tsp.h:7-109 (the adapter the benchmarks and the UI use)
//   tsp::PlannerConfig      <- include/sspp/tsp_planner.h:14-29
//   tsp::PathCandidate, SolverStatus, Point, ViaSet, Spline(4, degree 2)   <- include/sspp/tsp_types.h:8-34
//
// plan(start, end, iterate) keeps the reference's contract — one CES iteration per call, successes
// returned by value, distribution and forwarded best kept between calls when iterate == true — but the
// whole iteration (seed set = mean, forwarded best, samples; spline fit; collision + floor cost; elites;
// distribution update; sigma adaptation) is ONE fused CUDA launch, and the state lives on the device.
// Additive: plan_iterations() runs n iterations inside one launch (no host round trip per iteration),
// BatchPlanner plans many independent start/goal pairs at once.
//
// The adapter's effective values are reproduced when faithful_quirks is true (default):
// tsp.h:53 passes stddev_initial where Planner expects z_min (so dist z_min = stddev_initial and the
// initial sigma stays 0.3, tsp_planner.h:177), and Planner::evaluator_ is never configured, so the floor
// penalty uses z_min = 0, margin = 0.01, scale = 10 (tsp_evaluator.h:8) whatever the caller passes.
#ifndef SSPP_TSP_H
#define SSPP_TSP_H

#include <algorithm>
#include <array>
#include <cmath>
#include <cstdint>
#include <memory>
#include <optional>
#include <stdexcept>
#include <string>
#include <vector>

#include "sspp_planner.h"

namespace tsp {

constexpr int kDOF = 4;
constexpr int kSplineDegree = 2;
using Point = std::array<double, kDOF>;
using Spline = sspp::Spline<kDOF, kSplineDegree>;
using ViaSet = std::vector<Point>;

enum class SolverStatus { Converged, Failed, BelowFloor, MaxIter, Unknown };
inline const char* SolverStatustoString(SolverStatus s) {
  switch (s) {
    case SolverStatus::Converged: return "Converged";
    case SolverStatus::Failed: return "Failed";
    case SolverStatus::BelowFloor: return "BelowFloor";
    case SolverStatus::MaxIter: return "MaxIter";
    default: return "Unknown";
  }
}

struct PathCandidate {
  ViaSet via;
  std::optional<ViaSet> refined;   // unused (CES only), kept for source compatibility
  SolverStatus status{SolverStatus::Failed};
  double L{-1.0}, C_nf{-1.0}, C_wf{-1.0};
};

struct PlannerConfig {   // tsp_planner.h:14-29
  int samples{50}, checks{40}, total_points{3};
  int num_vias() const { return std::max(0, total_points - 2); }
  double w_collision{1.0};
  double elite_fraction{0.3};
  double inc{1.5}, dec{0.95};
  double sigma_floor{0.0}, var_beta{0.2}, mean_lr{0.5};
  double stddev_min{1e-3}, stddev_max{0.5};
  double z_min{0.0}, floor_margin{0.01}, floor_scale{10.0};
  bool verbose{false};
};

inline Point ones(double v) { return Point{v, v, v, v}; }

class TaskSpacePlanner {
 public:
  using GradientStepType = int;   // the reference's gradient types are unused by the CES planner

  // `model`: a loaded scene shared between planners (the reference takes an mjModel*)
  TaskSpacePlanner(std::shared_ptr<sspp::CudaModel> model, std::string body_name, double stddev_initial = 0.3,
                   double stddev_min = 0.01, double stddev_max = 0.5, double stddev_increase_factor = 1.5,
                   double stddev_decay_factor = 0.95, double elite_fraction = 0.3, int sample_count = 50, int check_points = 50,
                   int gd_iterations = 0, int init_points = 3, double collision_weight = 1.0, double z_min = 0.0,
                   Point limits_min = ones(-2.0), Point limits_max = ones(2.0), bool enable_gradient_descent = false,
                   double sigma_floor = 0.0, double var_ema_beta = 0.2, double mean_lr = 0.5, double max_step_norm = 0.1,
                   double floor_margin = 0.01, double floor_penalty_scale = 10.0, bool faithful_quirks = true, int device = 0,
                   int n_queries = 1)
      : model_(std::move(model)), body_(std::move(body_name)), limits_min_(limits_min), limits_max_(limits_max), n_queries_(n_queries) {
    (void)gd_iterations; (void)enable_gradient_descent; (void)max_step_norm;
    cfg_.samples = sample_count; cfg_.checks = check_points; cfg_.total_points = init_points;
    cfg_.w_collision = collision_weight; cfg_.elite_fraction = elite_fraction;
    cfg_.inc = stddev_increase_factor; cfg_.dec = stddev_decay_factor;
    cfg_.sigma_floor = sigma_floor; cfg_.var_beta = var_ema_beta; cfg_.mean_lr = mean_lr;
    cfg_.stddev_min = stddev_min; cfg_.stddev_max = stddev_max;
    cfg_.z_min = z_min; cfg_.floor_margin = floor_margin; cfg_.floor_scale = floor_penalty_scale;
    SsppPlannerConfig c;
    ssppcu_default_config(SSPPCU_MODE_TASK, &c);
    c.dof = kDOF; c.degree = kSplineDegree; c.n_points = init_points; c.check_points = check_points; c.n_samples = sample_count;
    c.include_mean = 1; c.include_best = 1; c.do_update = 1; c.n_queries = n_queries; c.device = device;
    c.body_name = body_.c_str();
    c.w_collision = collision_weight; c.elite_fraction = elite_fraction; c.inc = stddev_increase_factor; c.dec = stddev_decay_factor;
    c.sigma_floor = sigma_floor; c.var_beta = var_ema_beta; c.mean_lr = mean_lr; c.stddev_min = stddev_min; c.stddev_max = stddev_max;
    c.sample_zmin = z_min;
    if (faithful_quirks) { sigma0_ = 0.3; c.dist_zmin = stddev_initial; c.eval_zmin = 0.0; c.floor_margin = 0.01; c.floor_scale = 10.0; }
    else { sigma0_ = stddev_initial; c.dist_zmin = z_min; c.eval_zmin = z_min; c.floor_margin = floor_margin; c.floor_scale = floor_penalty_scale; }
    for (int d = 0; d < kDOF; ++d) { c.lo[d] = limits_min[d]; c.hi[d] = limits_max[d]; }
    char err[1024] = {0};
    if (ssppcu_create(model_->get(), &c, &h_, err, sizeof(err)) != SSPPCU_OK) throw std::runtime_error(std::string("sspp CUDA path: ") + err);
    ssppcu_dims(h_, &dims_);
    std::vector<double> knots((size_t)init_points + kSplineDegree + 1);
    ainv_.resize((size_t)init_points * init_points);
    ssppcu_spline_setup(init_points, kSplineDegree, knots.data(), ainv_.data());
    knots_ = knots;
  }
  ~TaskSpacePlanner() { ssppcu_destroy(h_); }
  TaskSpacePlanner(const TaskSpacePlanner&) = delete;
  TaskSpacePlanner& operator=(const TaskSpacePlanner&) = delete;

  // tsp.h:58-60 / tsp_planner.h:72-145: one CES iteration; iterate == false resets the distribution first
  std::vector<PathCandidate> plan(const Point& start, const Point& end, bool iterate_flag = false) {
    run(start, end, iterate_flag, 1);
    fetch();
    return successes_;
  }
  // additive: n CES iterations inside ONE launch (the first one cold unless iterate_flag); returns the
  // successes of the last iteration.  any_success() tells whether any iteration found a path.
  std::vector<PathCandidate> plan_iterations(const Point& start, const Point& end, int n_iters, bool iterate_flag = false) {
    run(start, end, iterate_flag, n_iters);
    fetch();
    return successes_;
  }
  bool any_success() const { return any_success_ != 0; }

  std::vector<PathCandidate> get_succesful_path_candidates() { return successes_; }
  std::vector<PathCandidate> get_failed_path_candidates() { return failures_; }
  const std::vector<ViaSet>& get_sampled_via_sets() const { return sampled_sets_; }
  std::vector<Point> get_sampled_via_pts() const {
    std::vector<Point> out;
    for (const auto& S : sampled_sets_) out.push_back(S.empty() ? Point{} : S.front());
    return out;
  }
  const std::vector<Point>& get_via_pts() const { return initial_via_; }
  Point get_current_mean() const { return mean_.empty() ? Point{} : mean_.front(); }
  Point get_current_stddev() const { return sigma_.empty() ? Point{} : sigma_.front(); }
  const ViaSet& mean_vias() const { return mean_; }
  const ViaSet& sigma_vias() const { return sigma_; }
  Point get_limits_min() const { return limits_min_; }
  Point get_limits_max() const { return limits_max_; }

  static Point evaluate(double u, const Spline& s) { return s(u); }
  Point evaluate(double u) const { return path_(u); }
  std::vector<Point> get_path_pts(int n = 10) const {
    std::vector<Point> pts;
    for (int i = 0; i < n; ++i) pts.push_back(path_((double)i / (n - 1)));
    return pts;
  }
  const sspp::ControlPoints<kDOF>& get_ctrl_pts() const { return path_.ctrls(); }
  const std::vector<double>& get_knot_vector() const { return path_.knots(); }
  void set_verbose(bool) {}
  void set_seed(uint64_t seed) { seed_ = seed; iteration_ = 0; }
  void reset() {}   // back-compat no-op, as in the reference (tsp.h:101)

  // tsp_path_model.h:32-43: quadratic spline interpolating start, vias, end at uniform parameters
  Spline spline_from_vias(const ViaSet& vias) const {
    const int K = cfg_.total_points;
    if ((int)vias.size() != K - 2) throw std::invalid_argument("spline_from_vias: expected " + std::to_string(K - 2) + " via points");
    sspp::ControlPoints<kDOF> c;
    c.resize(K);
    for (int i = 0; i < K; ++i)
      for (int d = 0; d < kDOF; ++d) {
        double acc = 0.0;
        for (int k = 0; k < K; ++k) {
          const double v = (k == 0) ? start_[d] : (k == K - 1) ? end_[d] : vias[k - 1][d];
          acc += ainv_[(size_t)i * K + k] * v;
        }
        c(d, i) = acc;
      }
    return Spline(knots_, std::move(c));
  }
  Spline spline_from_via(const Point& via) const { return spline_from_vias(ViaSet{via}); }
  const Spline& spline() const { return path_; }
  ssppcu_t* handle() { return h_; }
  int iterations_done() const { return (int)iteration_; }

 private:
  static void check(ssppcu_t* h, int rc) {
    if (rc < 0) throw std::runtime_error(std::string("sspp CUDA path: ") + ssppcu_last_error(h));
  }
  void run(const Point& start, const Point& end, bool iterate_flag, int n_iters) {
    if (n_queries_ != 1) throw std::logic_error("TaskSpacePlanner::plan is the single-query interface; use BatchPlanner");
    const bool same = have_query_ && start == start_ && end == end_;
    start_ = start; end_ = end;
    if (!same) { check(h_, ssppcu_set_queries(h_, start.data(), end.data())); have_query_ = true; }
    if (!iterate_flag) check(h_, ssppcu_reset_distribution(h_, sigma0_));     // Planner::reset (tsp_planner.h:54-69)
    const int K = cfg_.total_points;
    initial_via_.clear();
    for (int i = 0; i < K; ++i) {
      const double t = (double)i / (K - 1);
      Point p;
      for (int d = 0; d < kDOF; ++d) p[d] = (1 - t) * start[d] + t * end[d];
      initial_via_.push_back(p);
    }
    check(h_, ssppcu_iterate(h_, seed_, iteration_, n_iters));
    iteration_ += (uint32_t)n_iters;
  }
  void fetch() {
    const int N = dims_.n_cand, nv = dims_.n_free;
    L_.resize(N); cnf_.resize(N); cwf_.resize(N); st_.resize(N); params_.resize((size_t)N * dims_.nfd);
    int32_t ns = 0, bi = -1, any = 0;
    check(h_, ssppcu_get_results(h_, L_.data(), cnf_.data(), cwf_.data(), nullptr, st_.data(), &ns, nullptr, &bi, nullptr, &any));
    check(h_, ssppcu_get_params(h_, params_.data()));
    any_success_ = any;
    successes_.clear(); failures_.clear(); sampled_sets_.clear();
    for (int i = 0; i < N; ++i) {
      if (st_[i] == SSPPCU_ST_ABSENT) continue;            // the forwarded-best slot of a cold iteration
      PathCandidate c;
      c.via.resize(nv);
      for (int r = 0; r < nv; ++r)
        for (int d = 0; d < kDOF; ++d) c.via[r][d] = params_[((size_t)i * nv + r) * kDOF + d];
      c.L = L_[i]; c.C_nf = cnf_[i]; c.C_wf = cwf_[i];
      c.status = (st_[i] == SSPPCU_ST_CONVERGED) ? SolverStatus::Converged : SolverStatus::Failed;
      sampled_sets_.push_back(c.via);
      (c.status == SolverStatus::Converged ? successes_ : failures_).push_back(std::move(c));
    }
    std::vector<double> m((size_t)dims_.nfd), s((size_t)dims_.nfd);
    check(h_, ssppcu_get_distribution(h_, m.data(), s.data()));
    mean_.assign(nv, Point{}); sigma_.assign(nv, Point{});
    for (int r = 0; r < nv; ++r)
      for (int d = 0; d < kDOF; ++d) { mean_[r][d] = m[(size_t)r * kDOF + d]; sigma_[r][d] = s[(size_t)r * kDOF + d]; }
    // path_: every plan() call starts from the straight start-end line (tsp_planner.h:74-76) and replaces it with the best
    // path only when THIS iteration had successes (tsp_planner.h:122, 131-138) - a best forwarded from an earlier query that
    // fails now (iterate = true with a moved start / end) must not leak into path_
    std::vector<float> best((size_t)dims_.nfd);
    int32_t valid = 0;
    check(h_, ssppcu_get_best(h_, best.data(), &valid));
    const bool use_best = valid && ns > 0;
    ViaSet v(nv);
    for (int r = 0; r < nv; ++r)
      for (int d = 0; d < kDOF; ++d) v[r][d] = use_best ? (double)best[(size_t)r * kDOF + d] : initial_via_[1 + r][d];
    path_ = spline_from_vias(v);
  }

  std::shared_ptr<sspp::CudaModel> model_;
  std::string body_;
  PlannerConfig cfg_;
  Point limits_min_, limits_max_, start_{}, end_{};
  bool have_query_ = false;
  int n_queries_ = 1;
  double sigma0_ = 0.3;
  ssppcu_t* h_ = nullptr;
  SsppDims dims_{};
  std::vector<double> knots_, ainv_;
  std::vector<float> L_, cnf_, cwf_, params_;
  std::vector<uint8_t> st_;
  std::vector<PathCandidate> successes_, failures_;
  std::vector<ViaSet> sampled_sets_;
  std::vector<Point> initial_via_;
  ViaSet mean_, sigma_;
  Spline path_;
  int any_success_ = 0;
  uint64_t seed_ = 0x53535050ull;
  uint32_t iteration_ = 0;
};

}  // namespace tsp

#endif  // SSPP_TSP_H
