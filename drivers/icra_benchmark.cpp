// icra_benchmark.cpp — the reference's anytime benchmark driver re-hosted on the B200 planner core.
//
// Mirrors /root/reference/src/main_icra_benchmark.cpp: same command line
//   icra_benchmark <model.xml> <collision_body> [N=50] [start_body=block_green/] [end_body=block_orange/] [num_vias=1]
//                  [--max_iter=60] [--budgets_ms=10,20,50]
// same fixed planner parameters (:152-160), same start/goal rule (body points lifted by 0.02, :176-178), same
// runners (run_converged :48-64, run_anytime :66-89), cold / warm trials (:92-118) and the same report lines.
// Additive: --fused runs the converged benchmark's max_iter iterations inside ONE launch (plan_iterations).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <iostream>
#include <limits>
#include <numeric>
#include <string>
#include <tuple>
#include <vector>

#include "sspp_tsp.h"

using Point = tsp::Point;
using Clock = std::chrono::steady_clock;

struct Stats { double mean_ms{0}, std_ms{0}, min_ms{0}, max_ms{0}; };
static Stats stats(const std::vector<double>& ms) {
  Stats s{};
  if (ms.empty()) return s;
  const double n = (double)ms.size();
  s.mean_ms = std::accumulate(ms.begin(), ms.end(), 0.0) / n;
  double sq = 0;
  for (double v : ms) sq += v * v;
  s.std_ms = std::sqrt(std::max(0.0, sq / n - s.mean_ms * s.mean_ms));
  s.min_ms = *std::min_element(ms.begin(), ms.end());
  s.max_ms = *std::max_element(ms.begin(), ms.end());
  return s;
}
static double path_len_xyz(const tsp::TaskSpacePlanner& p, int samples = 60) {
  const auto pts = p.get_path_pts(samples);
  double L = 0;
  for (size_t i = 1; i < pts.size(); ++i) {
    double s2 = 0;
    for (int d = 0; d < 3; ++d) s2 += (pts[i][d] - pts[i - 1][d]) * (pts[i][d] - pts[i - 1][d]);
    L += std::sqrt(s2);
  }
  return L;
}
static std::vector<int> parse_csv_ints(const std::string& s) {
  std::vector<int> v;
  size_t i = 0;
  while (i < s.size()) {
    const size_t j = s.find(',', i);
    const std::string t = (j == std::string::npos) ? s.substr(i) : s.substr(i, j - i);
    if (!t.empty()) v.push_back(std::max(1, std::atoi(t.c_str())));
    if (j == std::string::npos) break;
    i = j + 1;
  }
  return v;
}

static bool g_fused = false;

// k cumulative calls: 1x cold, (k-1)x warm; returns (time_ms, ok, path_len, iters_done)
static std::tuple<double, bool, double, int> run_converged(tsp::TaskSpacePlanner& plan, const Point& q0, const Point& qT, int max_iter) {
  const auto t0 = Clock::now();
  int iters = 0;
  bool ok;
  if (g_fused) {
    plan.plan_iterations(q0, qT, max_iter, false);
    ok = plan.any_success();
    iters = max_iter;
  } else {
    ok = !plan.plan(q0, qT, false).empty();
    ++iters;
    for (int k = 1; k < max_iter; ++k) { ok |= !plan.plan(q0, qT, true).empty(); ++iters; }
  }
  const double ms = std::chrono::duration<double, std::milli>(Clock::now() - t0).count();
  return {ms, ok, ok ? path_len_xyz(plan) : 0.0, iters};
}

// Anytime: run until the wall-clock budget; returns (used_ms, ok, best_len, iters_done)
static std::tuple<double, bool, double, int> run_anytime(tsp::TaskSpacePlanner& plan, const Point& q0, const Point& qT, double budget_ms) {
  const auto t0 = Clock::now();
  const auto deadline = t0 + std::chrono::duration<double, std::milli>(budget_ms);
  int iters = 0;
  bool ok = !plan.plan(q0, qT, false).empty();
  ++iters;
  double best = ok ? path_len_xyz(plan) : std::numeric_limits<double>::infinity();
  while (Clock::now() < deadline) {
    const bool now_ok = !plan.plan(q0, qT, true).empty();
    ++iters;
    if (now_ok) { ok = true; best = std::min(best, path_len_xyz(plan)); }
  }
  const double used = std::chrono::duration<double, std::milli>(Clock::now() - t0).count();
  return {used, ok, ok ? best : 0.0, iters};
}

template <class Runner, class Maker, class... Args>
static std::tuple<Stats, int, double, double> trials(int N, bool warm, Runner run, Maker make, Args&&... args) {
  std::vector<double> times;
  int succ = 0; double sumL = 0.0; long long sum_iters = 0;
  std::unique_ptr<tsp::TaskSpacePlanner> shared;
  if (warm) shared = make();
  for (int i = 0; i < N; ++i) {
    std::unique_ptr<tsp::TaskSpacePlanner> fresh;
    if (!warm) fresh = make();
    auto [ms, ok, L, iters] = run(warm ? *shared : *fresh, args...);
    times.push_back(ms);
    if (ok) { ++succ; sumL += L; }
    sum_iters += iters;
  }
  return {stats(times), succ, succ ? sumL / succ : 0.0, (double)sum_iters / N};
}

int main(int argc, char** argv) {
  if (argc < 3) {
    std::cerr << "Usage: " << argv[0] << " <model.xml> <collision_body> [N=50] [start_body=block_green/] [end_body=block_orange/] [num_vias=1]"
              << " [--max_iter=60] [--budgets_ms=10,20,50] [--fused] [--converged]\n";
    return 1;
  }
  const std::string model_xml = argv[1], coll_body = argv[2];
  int N = (argc >= 4 && argv[3][0] != '-') ? std::max(1, std::atoi(argv[3])) : 50;
  std::string start_body = (argc >= 5 && argv[4][0] != '-') ? argv[4] : "block_green/";
  std::string end_body = (argc >= 6 && argv[5][0] != '-') ? argv[5] : "block_orange/";
  int num_vias = (argc >= 7 && argv[6][0] != '-') ? std::max(0, std::atoi(argv[6])) : 1;
  int max_iter = 60;
  bool converged = false;
  std::vector<int> budgets = {10, 20, 50};
  for (int i = 3; i < argc; ++i) {
    const std::string a = argv[i];
    if (a.rfind("--max_iter=", 0) == 0) max_iter = std::max(1, std::atoi(a.substr(11).c_str()));
    else if (a.rfind("--budgets_ms=", 0) == 0) budgets = parse_csv_ints(a.substr(13));
    else if (a == "--fused") g_fused = true;
    else if (a == "--converged") converged = true;
  }
  std::shared_ptr<sspp::CudaModel> m;
  try { m = std::make_shared<sspp::CudaModel>(model_xml); }
  catch (const std::exception& e) { std::cerr << "MuJoCo load error: " << e.what() << "\n"; return 2; }

  // planner setup (main_icra_benchmark.cpp:152-160)
  const double stddev_initial = 0.2, stddev_min = 1e-4, stddev_max = 0.5, inc = 1.5, dec = 0.9, elite = 0.3;
  const int samples = 15, checks = 40, gd_iters = 0, total_points = num_vias + 2;
  const double w_col = 1.0, z_min = 0.1;
  const double sigma_floor = 0.005, var_ema_beta = 0.2, mean_lr = 0.5, max_step_norm = 0.1, floor_margin = 0.01, floor_scale = 10.0;
  const Point lim_hi{0.7, 0.7, 0.6, 1.6}, lim_lo{0.0, -0.7, 0.1, -1.6};
  auto make_planner = [&]() {
    return std::make_unique<tsp::TaskSpacePlanner>(m, coll_body, stddev_initial, stddev_min, stddev_max, inc, dec, elite, samples, checks,
                                                   gd_iters, total_points, w_col, z_min, lim_lo, lim_hi, false, sigma_floor,
                                                   var_ema_beta, mean_lr, max_step_norm, floor_margin, floor_scale);
  };
  Point q0, qT;
  if (ssppcu_model_body_point(m->get(), start_body.c_str(), q0.data()) != SSPPCU_OK ||
      ssppcu_model_body_point(m->get(), end_body.c_str(), qT.data()) != SSPPCU_OK) {
    std::cerr << "start/end body is not a free body of the model\n";
    return 3;
  }
  q0[2] += 0.02; qT[2] += 0.02;
  try {
    if (converged) {
      const auto [Sc, nc, Lc, Ic] = trials(N, false, run_converged, make_planner, q0, qT, max_iter);
      const auto [Sw, nw, Lw, Iw] = trials(N, true, run_converged, make_planner, q0, qT, max_iter);
      std::cout << "\n=== Converged (max_iter=" << max_iter << ", N=" << N << (g_fused ? ", fused launch" : "") << ") ===\n";
      std::cout << "[Cold]  succ " << nc << "/" << N << "  mean " << Sc.mean_ms << " ms  std " << Sc.std_ms << "  min " << Sc.min_ms
                << "  max " << Sc.max_ms << "  avg iters " << Ic << "\n";
      if (nc) std::cout << "        avg length " << Lc << " m\n";
      std::cout << "[Warm]  succ " << nw << "/" << N << "  mean " << Sw.mean_ms << " ms  std " << Sw.std_ms << "  min " << Sw.min_ms
                << "  max " << Sw.max_ms << "  avg iters " << Iw << "\n";
      if (nw) std::cout << "        avg length " << Lw << " m\n";
    }
    if (!budgets.empty()) {
      std::cout << "\n=== Anytime (budgets ms: ";
      for (size_t i = 0; i < budgets.size(); ++i) std::cout << budgets[i] << (i + 1 < budgets.size() ? "," : "");
      std::cout << ", N=" << N << ") ===\n";
      for (int B : budgets) {
        const auto [Sc, nc, Lc, Ic] = trials(N, false, run_anytime, make_planner, q0, qT, (double)B);
        const auto [Sw, nw, Lw, Iw] = trials(N, true, run_anytime, make_planner, q0, qT, (double)B);
        std::cout << "[B=" << B << "] Cold: succ " << nc << "/" << N << "  mean " << Sc.mean_ms << " ms  std " << Sc.std_ms << "  min "
                  << Sc.min_ms << "  max " << Sc.max_ms << "  avg iters " << Ic;
        if (nc) std::cout << "  avgL " << Lc << " m";
        std::cout << "\n";
        std::cout << "       Warm: succ " << nw << "/" << N << "  mean " << Sw.mean_ms << " ms  std " << Sw.std_ms << "  min " << Sw.min_ms
                  << "  max " << Sw.max_ms << "  avg iters " << Iw;
        if (nw) std::cout << "  avgL " << Lw << " m";
        std::cout << "\n";
      }
    }
  } catch (const std::exception& e) { std::cerr << "error: " << e.what() << "\n"; return 4; }
  return 0;
}
