// tsp_benchmark.cpp — the reference's cold / warm single-iteration benchmark re-hosted on the B200 planner core.
//
// Mirrors /root/reference/src/main_tsp_benchmark.cpp (not built by the reference's own CMake, CMakeLists.txt:95):
//   tsp_benchmark <model.xml> <collision_body> [N=50] [start_body=block_green/] [end_body=block_orange/] [num_vias=1]
// N x plan(start, end, iterate=false) then N x plan(..., iterate=true) on ONE planner (:122-151), timed per call,
// same planner parameters (:73-93), same start / goal rule (:111-115) and the same report (:157-172).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <iostream>
#include <numeric>
#include <string>
#include <vector>

#include "sspp_tsp.h"

using Point = tsp::Point;

struct Stats { double mean_ms{0}, std_ms{0}, min_ms{0}, max_ms{0}; };
static Stats compute_stats(const std::vector<double>& v) {
  Stats s;
  if (v.empty()) return s;
  s.mean_ms = std::accumulate(v.begin(), v.end(), 0.0) / v.size();
  double var = 0;
  for (double x : v) var += (x - s.mean_ms) * (x - s.mean_ms);
  s.std_ms = std::sqrt(var / v.size());
  s.min_ms = *std::min_element(v.begin(), v.end());
  s.max_ms = *std::max_element(v.begin(), v.end());
  return s;
}
static double compute_path_length(const tsp::TaskSpacePlanner& p, int n = 60) {
  const auto pts = p.get_path_pts(n);
  double L = 0;
  for (size_t i = 1; i < pts.size(); ++i) {
    double s2 = 0;
    for (int d = 0; d < 3; ++d) s2 += (pts[i][d] - pts[i - 1][d]) * (pts[i][d] - pts[i - 1][d]);
    L += std::sqrt(s2);
  }
  return L;
}

int main(int argc, char** argv) {
  if (argc < 3 || argc > 7) {
    std::cerr << "Usage: " << argv[0] << " <model.xml> <collision_body> [N=50] [start_body=block_green/] [end_body=block_orange/] [num_vias=1]\n";
    return 1;
  }
  const std::string modelFile = argv[1], collisionBody = argv[2];
  const int N = (argc >= 4) ? std::max(1, std::atoi(argv[3])) : 50;
  const std::string start_body_name = (argc >= 5) ? argv[4] : "block_green/";
  const std::string end_body_name = (argc >= 6) ? argv[5] : "block_orange/";
  const int num_vias = (argc >= 7) ? std::max(0, std::atoi(argv[6])) : 1;
  std::shared_ptr<sspp::CudaModel> m;
  try { m = std::make_shared<sspp::CudaModel>(modelFile); }
  catch (const std::exception& e) { std::cerr << "Failed to load MuJoCo model: " << e.what() << "\n"; return 2; }
  const Point limit_max{0.7, 0.7, 0.6, 1.6}, limit_min{0.0, -0.7, 0.1, -1.6};
  try {
    tsp::TaskSpacePlanner planner(m, collisionBody, 0.2, 0.0001, 0.5, 1.5, 0.9, 0.3, 15, 40, 0, num_vias + 2, 1.0, 0.1, limit_min, limit_max,
                                  false, 0.005, 0.2, 0.5, 0.1, 0.01, 10.0);
    Point start_pt, end_pt;
    if (ssppcu_model_body_point(m->get(), start_body_name.c_str(), start_pt.data()) != SSPPCU_OK ||
        ssppcu_model_body_point(m->get(), end_body_name.c_str(), end_pt.data()) != SSPPCU_OK) {
      std::cerr << "start/end body is not a free body of the model\n";
      return 3;
    }
    start_pt[2] += 0.02; end_pt[2] += 0.02;
    std::vector<double> cold_ms, warm_ms;
    double len_cold = 0, len_warm = 0;
    int succ_cold = 0, succ_warm = 0;
    for (int pass = 0; pass < 2; ++pass)
      for (int i = 0; i < N; ++i) {
        const auto t0 = std::chrono::high_resolution_clock::now();
        const auto succ = planner.plan(start_pt, end_pt, /*iterate=*/pass == 1);
        const double dt = std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t0).count();
        (pass ? warm_ms : cold_ms).push_back(dt);
        if (!succ.empty()) { (pass ? succ_warm : succ_cold)++; (pass ? len_warm : len_cold) += compute_path_length(planner); }
      }
    const Stats cold = compute_stats(cold_ms), warm = compute_stats(warm_ms);
    std::cout << "\n=== Benchmark Results (N=" << N << ") ===\n[Cold start] iterate=false\n  Successes: " << succ_cold << " / " << N << "\n"
              << "  Mean time: " << cold.mean_ms << " ms,  Std: " << cold.std_ms << " ms,  Min: " << cold.min_ms << " ms,  Max: " << cold.max_ms << " ms\n";
    if (succ_cold) std::cout << "  Avg path length: " << len_cold / succ_cold << " m\n";
    std::cout << "\n[Warm start] iterate=true\n  Successes: " << succ_warm << " / " << N << "\n"
              << "  Mean time: " << warm.mean_ms << " ms,  Std: " << warm.std_ms << " ms,  Min: " << warm.min_ms << " ms,  Max: " << warm.max_ms << " ms\n";
    if (succ_warm) std::cout << "  Avg path length: " << len_warm / succ_warm << " m\n";
  } catch (const std::exception& e) { std::cerr << "error: " << e.what() << "\n"; return 4; }
  return 0;
}
