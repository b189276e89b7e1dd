// sspp_planner.h — host-side planner core over the C ABI (include/sspp_cuda.h), mirroring the
// reference's C++ planner interface so callers switch by changing an include:
//
//   sspp::SamplingPathPlanner<kDOF>   <- /root/reference/include/sspp.h:20-245   (planner A)
//   sspp::Spline<kDOF>                <- Eigen::Spline<double,kDOF,3> as used there (knots(), ctrls(), operator())
//
// Same method names, argument meaning and error behaviour (std::runtime_error with the reference's
// message prefix when the model cannot be loaded).  What changed underneath: plan() keeps the
// reference's loop shape (initializePath -> N x (sampleWithNoise, checkCollision) -> findBestPath)
// but the N candidates are sampled, evaluated, collision-checked, measured and ranked by ONE fused
// CUDA launch (ssppcu_iterate); the host only turns the returned rows into Spline objects.
// There is no CPU fallback: without a CUDA device plan() throws.
//
// Header-only, C++17, no Eigen (absent from this image): Point is std::array<double,kDOF>,
// control points are a column-major kDOF x K matrix like Eigen's ControlPointVectorType.
#ifndef SSPP_PLANNER_H
#define SSPP_PLANNER_H

#include <array>
#include <cmath>
#include <cstdint>
#include <iostream>
#include <limits>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>
#include <vector>

#include "sspp_cuda.h"

namespace sspp {

// kDOF x K control points, column-major (column j = control point j), as Eigen stores them
template <int kDOF>
struct ControlPoints {
  int cols = 0;
  std::vector<double> data;
  int rows() const { return kDOF; }
  double& operator()(int i, int j) { return data[(size_t)j * kDOF + i]; }
  double operator()(int i, int j) const { return data[(size_t)j * kDOF + i]; }
  void resize(int k) { cols = k; data.assign((size_t)k * kDOF, 0.0); }
};

// What the successful candidates of one plan() share: the initial spline (knots + control points, sspp.h:82-97) and the
// block of their sampled control-point rows as it came back from the device.  A Spline that refers to a row of the block
// builds its own knots / control points on first use, so handing a few thousand successes back to the caller
// (sspp.h:222 copies every spline) costs one small object each instead of two heap vectors each.
template <int kDOF>
struct SplineBlock {
  std::vector<double> knots;
  ControlPoints<kDOF> init;          // the unperturbed control points
  int first_free = 0, n_free = 0;    // columns [first_free, first_free + n_free) are the sampled ones (sspp.h:123)
  std::vector<float> rows;           // [n][n_free][kDOF]
};

template <int kDOF, int kDegree = 3>
class Spline {
 public:
  using Point = std::array<double, kDOF>;
  Spline() = default;
  Spline(std::vector<double> knots, ControlPoints<kDOF> ctrls) : knots_(std::move(knots)), ctrls_(std::move(ctrls)) {}
  Spline(std::shared_ptr<const SplineBlock<kDOF>> block, int row) : block_(std::move(block)), row_(row) {}
  // point an existing object at row `row` of `block` (what assigning Spline(block, row) does, without the temporary)
  void rebind(const std::shared_ptr<const SplineBlock<kDOF>>& block, int row) {
    knots_.clear(); ctrls_.data.clear(); ctrls_.cols = 0;
    block_ = block; row_ = row;
  }
  const std::vector<double>& knots() const { materialise(); return knots_; }
  const ControlPoints<kDOF>& ctrls() const { materialise(); return ctrls_; }
  ControlPoints<kDOF>& ctrls() { materialise(); return ctrls_; }
  int degree() const { return kDegree; }

  // Eigen::Spline::operator(): span by upper_bound on the knots (last span at u = 1), Cox-de Boor basis
  Point operator()(double u) const {
    materialise();
    Point out{};
    const int n = ctrls_.cols, p = kDegree;
    if (n < p + 1 || (int)knots_.size() != n + p + 1) return out;
    int span = p;
    if (u > knots_[0]) {
      int lo = p, hi = n;                         // first index in [p, n) with knots[idx] > u
      while (lo < hi) { const int mid = (lo + hi) / 2; if (knots_[mid] > u) hi = mid; else lo = mid + 1; }
      span = lo - 1;
    }
    double N[kDegree + 1], left[kDegree + 1], right[kDegree + 1];
    N[0] = 1.0;
    for (int j = 1; j <= p; ++j) {
      left[j] = u - knots_[span + 1 - j];
      right[j] = knots_[span + j] - u;
      double saved = 0.0;
      for (int r = 0; r < j; ++r) {
        const double tmp = N[r] / (right[r + 1] + left[j - r]);
        N[r] = saved + right[r + 1] * tmp;
        saved = left[j - r] * tmp;
      }
      N[j] = saved;
    }
    for (int j = 0; j <= p; ++j)
      for (int i = 0; i < kDOF; ++i) out[i] += N[j] * ctrls_(i, span - p + j);
    return out;
  }

 private:
  void materialise() const {
    if (!block_) return;
    const SplineBlock<kDOF>& b = *block_;
    knots_ = b.knots;
    ctrls_ = b.init;
    const float* r = b.rows.data() + (size_t)row_ * b.n_free * kDOF;
    for (int j = 0; j < b.n_free; ++j)
      for (int d = 0; d < kDOF; ++d) ctrls_(d, b.first_free + j) = r[(size_t)j * kDOF + d];
    block_.reset();
  }
  mutable std::vector<double> knots_;
  mutable ControlPoints<kDOF> ctrls_;
  mutable std::shared_ptr<const SplineBlock<kDOF>> block_;   // non-null: knots_ / ctrls_ not built yet
  int row_ = -1;
};

// RAII wrappers of the C handles
class CudaModel {
 public:
  explicit CudaModel(const std::string& path) {
    char err[1024] = {0};
    if (ssppcu_model_load(path.c_str(), &m_, err, sizeof(err)) != SSPPCU_OK) throw std::runtime_error(err);
  }
  ~CudaModel() { ssppcu_model_free(m_); }
  CudaModel(const CudaModel&) = delete;
  CudaModel& operator=(const CudaModel&) = delete;
  const ssppcu_model_t* get() const { return m_; }
 private:
  ssppcu_model_t* m_ = nullptr;
};

// everything one plan() produced, still addressable on the device through `view`
struct PlanBuffers {
  int n_cand = 0, n_free = 0;
  std::vector<float> L, cost;        // per candidate: arc length, L + w*Cwf
  std::vector<uint8_t> status;       // SSPPCU_ST_*
  std::vector<float> params;         // [n_cand][n_free][kDOF] sampled control points of EVERY candidate: filled on demand (all_params)
  const float* succ_rows = nullptr;  // [n_success][n_free][kDOF] rows of the collision-free candidates, candidate order (pinned view,
  const int32_t* succ_idx = nullptr; //  valid until the next plan) and their candidate indices
  int n_success = 0, best_idx = -1;
  float best_cost = std::numeric_limits<float>::infinity();
  SsppDeviceView view{};             // raw device pointers (valid until the next plan on this shape)
};

template <int kDOF>
class SamplingPathPlanner {
 public:
  using Point = std::array<double, kDOF>;
  static constexpr int kSplineDegree = 3;
  using Spline = sspp::Spline<kDOF, kSplineDegree>;

  // `xml_string` is a FILE NAME, as in the reference (sspp.h:47 hands it to mj_loadXML)
  explicit SamplingPathPlanner(const std::string& xml_string, int device = 0) : device_(device) {
    try { model_ = new CudaModel(xml_string); }
    catch (const std::exception& e) {
      const std::string w = e.what();
      throw std::runtime_error(w.rfind("Failed to load", 0) == 0 ? w : "Failed to load MuJoCo model from XML: " + w);
    }
  }
  ~SamplingPathPlanner() {
    for (auto& kv : handles_) ssppcu_destroy(kv.second);
    delete model_;
  }
  SamplingPathPlanner(const SamplingPathPlanner&) = delete;
  SamplingPathPlanner& operator=(const SamplingPathPlanner&) = delete;

  // sspp.h:82-97: cubic spline interpolating num_points linearly spaced points (uniform parameters)
  bool initializePath(const Point& start, const Point& end, Spline& init_spline, int num_points = 10) {
    if (num_points < kSplineDegree + 1) throw std::invalid_argument("initializePath: num_points must be at least degree + 1");
    std::vector<double> knots((size_t)num_points + kSplineDegree + 1), ainv((size_t)num_points * num_points);
    if (ssppcu_spline_setup(num_points, kSplineDegree, knots.data(), ainv.data()) != SSPPCU_OK)
      throw std::runtime_error("initializePath: spline setup failed");
    ControlPoints<kDOF> c;
    c.resize(num_points);
    for (int i = 0; i < num_points; ++i)
      for (int d = 0; d < kDOF; ++d) {
        double acc = 0.0;
        for (int k = 0; k < num_points; ++k) {
          const double t = (double)k / (num_points - 1);
          acc += ainv[(size_t)i * num_points + k] * ((1 - t) * start[d] + t * end[d]);
        }
        c(d, i) = acc;
      }
    init_spline = Spline(std::move(knots), std::move(c));
    return true;
  }

  Point evaluate(double u) const { return path_spline_(u); }
  Point evaluate(const Spline& spline, double u) const { return spline(u); }
  const ControlPoints<kDOF>& get_ctrl_pts() const { return path_spline_.ctrls(); }
  const Spline& path() const { return path_spline_; }

  // sspp.h:152-169 (single spline, host FP64: a utility outside the batched path)
  double computeArcLength(const Spline& spline, int check_points) const {
    double total = 0.0;
    for (int i = 1; i < check_points; ++i) {
      const Point a = spline((double)(i - 1) / (check_points - 1)), b = spline((double)i / (check_points - 1));
      double s2 = 0.0;
      for (int d = 0; d < kDOF; ++d) s2 += (b[d] - a[d]) * (b[d] - a[d]);
      total += std::sqrt(s2);
    }
    return total;
  }

  // sspp.h:171-192; ties resolve to the first path (the reference's order is thread arrival)
  bool findBestPath(const std::vector<Spline>& successful_paths, Spline& best_spline, int check_points = 10) {
    double min_cost = std::numeric_limits<double>::infinity();
    bool found = false;
    for (const Spline& s : successful_paths) {
      const double cost = computeArcLength(s, check_points);
      if (cost < min_cost) { min_cost = cost; best_spline = s; found = true; }
    }
    return found;
  }

  // sspp.h:194-232.  One fused launch scores all sample_count candidates; successes come back in
  // candidate order (the reference's order is nondeterministic).  The best path = shortest arc length
  // among the collision-free candidates is stored like the reference's path_spline_.
  // plan() up to the point where the reference starts copying splines out: the successes of the scored population as ONE block
  // (initial spline + their sampled rows, candidate order); row s of it is the s-th returned path.  `found` as plan()'s result.
  struct PlannedBlock { std::shared_ptr<const SplineBlock<kDOF>> block; int n_success = 0; bool found = false; };
  PlannedBlock plan_block(const Point& start, const Point& end, double sigma, const Point& limits, int sample_count = 50,
                          int check_points = 50, int init_points = 10) {
    if (sample_count <= 0) {        // sspp.h:207-221: the loop body never runs, the summary line is printed, findBestPath finds nothing
      initializePath(start, end, init_spline_, init_points);
      if (verbose_) std::cout << "Sampled " << sample_count << " splines. Successful paths found: 0" << std::endl;
      return PlannedBlock{};
    }
    PlanBuffers& B = plan_device(start, end, sigma, limits, sample_count, check_points, init_points);
    auto block = std::make_shared<SplineBlock<kDOF>>();
    block->knots = init_spline_.knots(); block->init = init_spline_.ctrls();
    block->first_free = kSplineDegree; block->n_free = B.n_free;
    block->rows.assign(B.succ_rows, B.succ_rows + (size_t)B.n_success * B.n_free * kDOF);
    int best_row = -1;
    if (B.best_idx >= 0)
      for (int s = 0; s < B.n_success; ++s) if (B.succ_idx[s] == B.best_idx) { best_row = s; break; }
    if (verbose_) std::cout << "Sampled " << sample_count << " splines. Successful paths found: " << B.n_success << std::endl;
    PlannedBlock r;
    r.n_success = B.n_success; r.found = B.best_idx >= 0 && best_row >= 0;
    r.block = std::move(block);
    if (r.found) path_spline_ = Spline(r.block, best_row);
    return r;
  }
  bool plan(const Point& start, const Point& end, double sigma, const Point& limits, std::vector<Spline>& ret_success_paths,
            int sample_count = 50, int check_points = 50, int init_points = 10) {
    const PlannedBlock r = plan_block(start, end, sigma, limits, sample_count, check_points, init_points);
    // the successes share one block; each returned Spline is a (block, row) reference that builds its control points on first use
    ret_success_paths.clear();
    ret_success_paths.reserve(r.n_success);
    for (int s = 0; s < r.n_success; ++s) ret_success_paths.emplace_back(r.block, s);
    return r.found;
  }
  bool plan(const Point& start, const Point& end, double sigma, const Point& limits, int sample_count = 50, int check_points = 50,
            int init_points = 10) {
    std::vector<Spline> successful_paths;
    return plan(start, end, sigma, limits, successful_paths, sample_count, check_points, init_points);
  }

  // ---- additive: the device-buffer path.  Scores the population and leaves every result on the
  // device (B.view) besides the host copies; no Spline objects are built.
  PlanBuffers& plan_device(const Point& start, const Point& end, double sigma, const Point& limits, int sample_count,
                           int check_points, int init_points) {
    ssppcu_t* h = handle_for(sample_count, check_points, init_points);
    // ONE call, ONE scoring launch: the query and its noise scale ride in the kernel's parameter block; the compaction
    // kernel behind it posts the successful rows to pinned host memory; one stream synchronisation reads everything back
    check(h, ssppcu_plan_step(h, start.data(), end.data(), sigma, limits.data(), seed_, iteration_++, 1));
    PlanBuffers& B = buffers_;
    SsppDims d;
    ssppcu_dims(h, &d);
    B.n_cand = d.n_cand; B.n_free = d.n_free;
    B.L.resize(d.n_cand); B.cost.resize(d.n_cand); B.status.resize(d.n_cand);
    B.params.clear();
    int32_t ns = 0, bi = -1; float bc = 0;
    check(h, ssppcu_get_success_rows(h, 0, &B.succ_rows, &B.succ_idx, &ns));
    int32_t ns2 = 0;
    check(h, ssppcu_get_results(h, B.L.data(), nullptr, nullptr, B.cost.data(), B.status.data(), &ns2, nullptr, &bi, &bc, nullptr));
    if (ns2 != ns) throw std::runtime_error("sspp CUDA path: success rows and result block disagree");
    B.n_success = ns; B.best_idx = bi; B.best_cost = bc;
    last_handle_ = h;
    ssppcu_device_view(h, &B.view);
    // the spline pieces shared by every candidate of this plan
    initializePath(start, end, init_spline_, init_points);
    return B;
  }

  void set_seed(uint64_t seed) { seed_ = seed; iteration_ = 0; }
  void set_verbose(bool v) { verbose_ = v; }
  int device() const { return device_; }

  // sampled control points of EVERY candidate of the last plan [n_cand][n_free][kDOF] (failed ones included): one extra
  // device-to-host copy, made on demand
  const std::vector<float>& all_params() {
    PlanBuffers& B = buffers_;
    if (B.params.empty() && last_handle_) {
      B.params.resize((size_t)B.n_cand * B.n_free * kDOF);
      check(last_handle_, ssppcu_get_params(last_handle_, B.params.data()));
    }
    return B.params;
  }

  // Spline of candidate i of the last plan (any status): init spline with the candidate's control points written in,
  // fetched from the device row by row (ssppcu_candidate_ctrl)
  Spline candidate_spline(const PlanBuffers& B, int i) const {
    Spline s = init_spline_;
    std::vector<float> pts((size_t)s.ctrls().cols * kDOF);
    check(last_handle_, ssppcu_candidate_ctrl(last_handle_, 0, (uint32_t)i, pts.data()));
    for (int j = 0; j < s.ctrls().cols; ++j)
      for (int d = 0; d < kDOF; ++d) s.ctrls()(d, j) = pts[(size_t)j * kDOF + d];
    (void)B;
    return s;
  }

 private:
  static void check(ssppcu_t* h, int rc) {
    if (rc < 0) throw std::runtime_error(std::string("sspp CUDA path: ") + ssppcu_last_error(h));
  }
  ssppcu_t* handle_for(int sample_count, int check_points, int init_points) {
    const auto key = std::make_tuple(sample_count, check_points, init_points);
    auto it = handles_.find(key);
    if (it != handles_.end()) return it->second;
    SsppPlannerConfig cfg;
    ssppcu_default_config(SSPPCU_MODE_JOINT, &cfg);
    cfg.dof = kDOF; cfg.degree = kSplineDegree; cfg.n_points = init_points;
    cfg.check_points = check_points; cfg.n_samples = sample_count; cfg.device = device_;
    char err[1024] = {0};
    ssppcu_t* h = nullptr;
    const int rc = ssppcu_create(model_->get(), &cfg, &h, err, sizeof(err));
    if (rc != SSPPCU_OK) throw std::runtime_error(std::string("sspp CUDA path: ") + err);
    check(h, ssppcu_set_success_rows(h, 1));
    handles_[key] = h;
    return h;
  }

  CudaModel* model_ = nullptr;
  int device_ = 0;
  Spline path_spline_, init_spline_;
  std::map<std::tuple<int, int, int>, ssppcu_t*> handles_;   // one device-resident scene per planner shape
  PlanBuffers buffers_;
  ssppcu_t* last_handle_ = nullptr;
  uint64_t seed_ = 0x53535050ull;
  uint32_t iteration_ = 0;
  bool verbose_ = true;
};

}  // namespace sspp

#endif  // SSPP_PLANNER_H
