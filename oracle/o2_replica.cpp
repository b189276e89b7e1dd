// o2_replica.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Oracle O2: the FP32 bit-replica of one
// fused iteration.  It compiles the product's single-source arithmetic header
// (sspp_b200/csrc/sspp_math.h) with g++ -ffp-contract=off and restates, on the host, the parts
// that live in the kernel: the work decomposition is irrelevant to the result, the reduction
// order (lane-strided partial sums, then an xor butterfly over 32 lanes) is restated here.
// Its purpose is the north-star's bit-exact check of elite indices, statuses and collision
// flags on seeded inputs; the independent FP64 statement of the semantics is oracle O1.
// Only tests/, __graft_entry__.smoke() and bench.py may load it.
//
// PARITY STATUS: as O1 — unpinned against MuJoCo (absent); pinned only through O1 agreement.
#include <algorithm>
#include <cstring>
#include <string>
#include <vector>

#include "../include/sspp_cuda.h"
#include "../sspp_b200/csrc/sspp_math.h"
#include "../sspp_b200/csrc/sspp_model.h"

using namespace ssppm;

namespace {

struct Replica {
  DevModel dev;
  ScoreConfig cfg;
  std::vector<float> nodes;   // SDF mode: the node grid, baked with the same function as the device kernel
  bool count_work = false;    // o2_work: accumulate the work counters of sspp_math.h over the points scored
  unsigned long long work[WC_COUNT] = {};
};

struct HostXF {
  float v[MAX_MBODY * 12];
  void put(int b, int k, float x) { v[b * 12 + k] = x; }
  float get(int b, int k) const { return v[b * 12 + k]; }
};
struct HostQ {
  const float* q;
  float get(int d) const { return q[d]; }
};

// the kernel's warp_strided_sum / warp_butterfly
float butterfly(float part[32]) {
  for (int m = 16; m > 0; m >>= 1) {
    float nxt[32];
    for (int l = 0; l < 32; ++l) nxt[l] = part[l] + part[l ^ m];
    std::memcpy(part, nxt, sizeof(nxt));
  }
  return part[0];
}
float strided_sum(const float* v, int n) {
  float part[32];
  for (int l = 0; l < 32; ++l) { float acc = 0.0f; for (int i = l; i < n; i += 32) acc += v[i]; part[l] = acc; }
  return butterfly(part);
}
// the kernel's wide sums (elite reductions): 8 groups of 32 virtual lanes, virtual lane v sums its terms
// j = v, v + 256, ... in ascending order (term(j, acc) -> acc'), butterfly per group, group sums added in
// ascending group order, empty groups skipped
template <class Term>
float wide_sum(int k, Term term) {
  float tot = 0.0f;
  for (int g = 0; g < 8 && (g == 0 || 32 * g < k); ++g) {
    float part[32];
    for (int l = 0; l < 32; ++l) { float acc = 0.0f; for (int j = 32 * g + l; j < k; j += 256) acc = term(j, acc); part[l] = acc; }
    const float sg = butterfly(part);
    tot = (g == 0) ? sg : tot + sg;
  }
  return tot;
}

}  // namespace

extern "C" {

void* o2_create(const char* mjcf_path, const SsppPlannerConfig* c, char* err, size_t errlen) {
  try {
    sspp::FlatModel fm = sspp::load_mjcf(mjcf_path);
    Replica* r = new Replica;
    const bool task = c->mode == SSPPCU_MODE_TASK;
    sspp::CompileOptions opt;
    opt.sdf = c->collision_mode == SSPPCU_COLLIDE_SDF;
    opt.sphere_ratio = c->sdf_sphere_ratio > 0 ? c->sdf_sphere_ratio : 0.5;
    opt.use_center_dist = c->use_center_dist != 0;
    r->dev = sspp::compile_model(fm, task ? sspp::PLAN_TASK : sspp::PLAN_JOINT, c->dof, c->body_name ? c->body_name : "", nullptr, opt);
    if (r->dev.sdf_mode) {      // same grid geometry arithmetic as ssppcu_create
      SdfGrid& g = r->dev.grid;
      g.nx = c->sdf_dims[0]; g.ny = c->sdf_dims[1]; g.nz = c->sdf_dims[2];
      const int dims[3] = {g.nx, g.ny, g.nz};
      for (int a = 0; a < 3; ++a) {
        g.lo[a] = (float)c->sdf_lo[a];
        g.h[a] = (float)((c->sdf_hi[a] - c->sdf_lo[a]) / (dims[a] - 1));
        g.inv_h[a] = 1.0f / g.h[a];
      }
      r->nodes.resize((size_t)g.nx * g.ny * g.nz);
#pragma omp parallel for schedule(static)
      for (int k = 0; k < g.nz; ++k)
        for (int j = 0; j < g.ny; ++j)
          for (int i = 0; i < g.nx; ++i) r->nodes[((size_t)k * g.ny + j) * g.nx + i] = sdf_node_value(r->dev, i, j, k);
      g.nodes = r->nodes.data(); g.cells = nullptr;
    }
    ScoreConfig& s = r->cfg;
    std::memset(&s, 0, sizeof(s));
    s.mode = task ? MODE_TASK : MODE_JOINT;
    s.D = c->dof; s.K = c->n_points;
    s.n_free = task ? c->n_points - 2 : c->n_points - 2 * c->degree;
    s.free_begin = task ? 1 : c->degree;
    const int M = c->check_points;
    s.Mc = M + 1; s.col_first = task ? 1 : 0; s.Ma = task ? M + 1 : M; s.arc_same = task ? 1 : 0;
    s.include_mean = c->include_mean ? 1 : 0; s.include_best = c->include_best ? 1 : 0; s.do_update = c->do_update ? 1 : 0;
    s.n_cand = c->n_samples + ((c->shard_global > 0) ? 0 : s.include_mean + s.include_best);
    s.w_collision = (float)c->w_collision;
    s.eval_zmin = (float)c->eval_zmin; s.floor_margin = (float)c->floor_margin; s.floor_scale = (float)c->floor_scale;
    s.limit_scale = (float)c->limit_scale; s.sample_zmin = (float)c->sample_zmin; s.dist_zmin = (float)c->dist_zmin;
    s.inc = (float)c->inc; s.dec = (float)c->dec; s.sigma_min = (float)c->stddev_min; s.sigma_max = (float)c->stddev_max;
    s.sigma_floor = (float)c->sigma_floor; s.var_beta = (float)c->var_beta; s.mean_lr = (float)c->mean_lr;
    for (int d = 0; d < SSPPCU_MAX_DOF; ++d) { s.lo[d] = (float)c->lo[d]; s.hi[d] = (float)c->hi[d]; }
    s.elite_frac = c->elite_fraction;
    return r;
  } catch (const std::exception& e) {
    if (err && errlen) { std::strncpy(err, e.what(), errlen - 1); err[errlen - 1] = 0; }
    return nullptr;
  }
}
void o2_destroy(void* h) { delete static_cast<Replica*>(h); }
const float* o2_sdf_nodes(void* h, int32_t dims[3]) {
  Replica* r = static_cast<Replica*>(h);
  dims[0] = r->dev.grid.nx; dims[1] = r->dev.grid.ny; dims[2] = r->dev.grid.nz;
  return r->nodes.data();
}
int o2_spheres(void* h, float* out, int cap) {   // [n][5] like ssppcu_get_spheres
  const DevModel& d = static_cast<Replica*>(h)->dev;
  for (int i = 0; out && i < d.n_sphere && i < cap; ++i) {
    out[5 * i] = d.sph[i].c[0]; out[5 * i + 1] = d.sph[i].c[1]; out[5 * i + 2] = d.sph[i].c[2]; out[5 * i + 3] = d.sph[i].r; out[5 * i + 4] = (float)d.sph[i].body;
  }
  return d.n_sphere;
}
float o2_sdf_sample(void* h, const float* x) { return sdf_sample_nodes(static_cast<Replica*>(h)->dev.grid, mk(x[0], x[1], x[2])); }
int o2_n_cand(void* h) { return static_cast<Replica*>(h)->cfg.n_cand; }

float o2_point_cost(void* h, const float* q) {
  Replica* r = static_cast<Replica*>(h);
  HostXF xf; HostQ qa{q};
  return point_collision_cost(r->dev, qa, xf);
}

// work counters: enable / disable, read (and clear) the totals accumulated by o2_iterate since
int o2_work(void* h, int enable, unsigned long long* out, int cap) {
  Replica* r = static_cast<Replica*>(h);
  for (int e = 0; out && e < WC_COUNT && e < cap; ++e) out[e] = r->work[e];
  std::memset(r->work, 0, sizeof(r->work));
  r->count_work = enable != 0;
  return WC_COUNT;
}

void o2_philox(uint32_t ctr[4], uint32_t k0, uint32_t k1) { philox4x32_10(ctr, k0, k1); }
float o2_normal(uint64_t seed, uint32_t iter, uint32_t query, uint32_t cand, uint32_t stream, uint32_t j) { return normal_at(seed, iter, query, cand, stream, j); }
void o2_sincos(float x, float* s, float* c) { sincos_poly(x, s, c); }
float o2_log(float x) { return log_poly(x); }

// One iteration of one query, restating sspp_iterate_kernel.  In/out: mean, sigma, best [nfd],
// best_valid.  cand_offset / n_local select a shard of the population (0, n_cand otherwise).
// Outputs per local candidate: params [n][nfd], L, Cnf, Cwf, cost, status; keys [n] sorted
// ascending; weights [k]; pts (optional) [n][Mc][D].
int o2_iterate(void* h, const float* psi_col, const float* psi_arc, const float* base, float* mean, float* sigma, float* best,
               int32_t* best_valid, uint64_t seed, uint32_t iter, uint32_t query, int cand_offset, int n_local, int do_elites,
               float* params, float* L, float* Cnf, float* Cwf, float* cost, uint8_t* status, unsigned long long* keys,
               float* weights, int32_t* n_success, int32_t* best_idx, float* pts) {
  Replica* r = static_cast<Replica*>(h);
  const ScoreConfig& c = r->cfg;
  const int D = c.D, K = c.K, Mc = c.Mc, Ma = c.Ma, nfd = c.n_free * D;
  const int Marc = c.arc_same ? Mc : Ma;
  const int best_ok = c.include_best ? *best_valid : 0;
  std::vector<float> X((size_t)K * D), qcol((size_t)Mc * D), qarc((size_t)Marc * D), cc(Mc), cw(Mc), seg(Marc);
  // banded rows exactly as the C ABI packs them for the kernel
  std::vector<int> first_col(Mc), first_arc(Ma > 0 ? Ma : 1);
  const int w_col = compute_band(psi_col, Mc, K, first_col.data());
  std::vector<float> band_col((size_t)Mc * w_col), band_arc;
  pack_band(psi_col, Mc, K, w_col, first_col.data(), band_col.data());
  int w_arc = 1;
  if (!c.arc_same) {
    w_arc = compute_band(psi_arc, Ma, K, first_arc.data());
    band_arc.resize((size_t)Ma * w_arc);
    pack_band(psi_arc, Ma, K, w_arc, first_arc.data(), band_arc.data());
  }
  for (int i = 0; i < n_local; ++i) {
    const int gc = cand_offset + i;
    const int kind = candidate_kind(c, gc, best_ok);
    std::copy(base, base + K * D, X.begin());
    for (int t = 0; t < nfd; ++t) {
      const float v = candidate_param(c, seed, iter, query, gc, kind, t, mean, sigma, best);
      X[c.free_begin * D + t] = v;
      params[(size_t)i * nfd + t] = v;
    }
    for (int m = 0; m < Mc; ++m) {
      for (int d = 0; d < D; ++d) {
        const float acc = spline_coord(&band_col[(size_t)m * w_col], first_col[m], w_col, X.data(), D, d);
        qcol[m * D + d] = acc;
        if (pts) pts[((size_t)i * Mc + m) * D + d] = acc;
      }
      float a = 0.0f, b = 0.0f;
      if (m >= c.col_first) {
        HostXF xf; HostQ qa{&qcol[m * D]};
        if (r->count_work) {
          uint32_t wc[WC_COUNT] = {};
          a = point_collision_cost(r->dev, qa, xf, wc);
          for (int e = 0; e < WC_COUNT; ++e) r->work[e] += wc[e];
        } else a = point_collision_cost(r->dev, qa, xf);
        b = a + point_penalty(c, qa);
      }
      cc[m] = a; cw[m] = b;
    }
    const float* qa_ = qcol.data();
    if (!c.arc_same) {
      for (int m = 0; m < Ma; ++m)
        for (int d = 0; d < D; ++d) {
          qarc[m * D + d] = spline_coord(&band_arc[(size_t)m * w_arc], first_arc[m], w_arc, X.data(), D, d);
        }
      qa_ = qarc.data();
    }
    for (int m = 1; m < Marc; ++m) {
      float s2 = 0.0f;
      for (int d = 0; d < D; ++d) { const float df = qa_[m * D + d] - qa_[(m - 1) * D + d]; s2 = fmaf(df, df, s2); }
      seg[m - 1] = sqrtf(s2);
    }
    L[i] = strided_sum(seg.data(), Marc - 1);
    Cnf[i] = strided_sum(cc.data() + c.col_first, Mc - c.col_first);
    Cwf[i] = strided_sum(cw.data() + c.col_first, Mc - c.col_first);
    const int st = (kind == CK_ABSENT) ? ST_ABSENT : ((Cnf[i] == 0.0f) ? ST_CONVERGED : ST_FAILED);
    cost[i] = fmaf(c.w_collision, Cwf[i], L[i]);
    status[i] = (uint8_t)st;
    keys[i] = sort_key(cost[i], st, (uint32_t)gc);
  }
  std::sort(keys, keys + n_local);
  if (!do_elites) return 0;
  // ---- elites_and_update ----
  int n_succ = 0;
  for (int i = 0; i < n_local; ++i) if ((uint32_t)(keys[i] >> 32) != 0xFFFFFFFFu) ++n_succ;
  *n_success = n_succ;
  const int k = elite_count(n_succ, c.elite_frac);
  if (n_succ == 0) {
    *best_idx = -1;
    if (c.do_update) for (int t = 0; t < nfd; ++t) sigma[t] = clamp_sigma(c, sigma[t] * c.inc);
    return 0;
  }
  *best_idx = (int)(uint32_t)(keys[0] & 0xFFFFFFFFu);
  const float lk = log_poly((float)k + 0.5f);
  for (int j = 0; j < k; ++j) weights[j] = ces_log_weight(k, lk, j);
  const float sumw = wide_sum(k, [&](int j, float acc) { return acc + weights[j]; });
  for (int j = 0; j < k; ++j) weights[j] = weights[j] / sumw;
  auto P = [&](int j, int t) { return params[(size_t)((uint32_t)(keys[j] & 0xFFFFFFFFu) - cand_offset) * nfd + t]; };
  std::vector<float> bestn(nfd);
  for (int t = 0; t < nfd; ++t) bestn[t] = P(0, t);
  if (c.do_update) {
    for (int t = 0; t < nfd; ++t) {
      const int d = t % D;
      const float em = wide_sum(k, [&](int j, float acc) { return fmaf(weights[j], P(j, t), acc); });
      const float m0 = mean[t];
      float nm = fmaf(c.mean_lr, em - m0, m0);
      if (c.mode == MODE_TASK && d == 2) nm = fmaxf(nm, c.dist_zmin);
      nm = fminf(fmaxf(nm, c.lo[d]), c.hi[d]);
      const bool wrap = (c.mode == MODE_TASK && d == 3 && c.lo[3] != c.hi[3]);
      const float var = wide_sum(k, [&](int j, float acc) {
        const float v = P(j, t);
        const float df = wrap ? wrap_angle_diff(v, nm, c.lo[3], c.hi[3]) : v - nm;
        return fmaf(weights[j], df * df, acc);
      });
      const float s0 = sigma[t];
      const float blend = fmaf(c.var_beta, var, (1.0f - c.var_beta) * (s0 * s0));
      float sg = clamp_sigma(c, sqrtf(blend));
      sg = clamp_sigma(c, sg * c.dec);
      mean[t] = nm; sigma[t] = sg;
    }
  }
  std::copy(bestn.begin(), bestn.end(), best);
  *best_valid = 1;
  return k;
}

// Population-sharded path: restates sspp_merge_rank_kernel + sspp_merge_finish_kernel.  keys: the
// gathered keys of every shard (any order: the merged order of unique keys is their sorted order).
// In/out: mean, sigma, best [nfd], best_valid.  Elites' rows are regenerated from the Philox counters
// with the PRE-update state, as the kernel does.  Returns k; keys come back sorted.
int o2_merge_update(void* h, unsigned long long* keys, int n_keys, uint64_t seed, uint32_t iter, float* mean, float* sigma,
                    float* best, int32_t* best_valid, float* weights, int32_t* n_success, int32_t* best_idx) {
  Replica* r = static_cast<Replica*>(h);
  const ScoreConfig& c = r->cfg;
  const int D = c.D, nfd = c.n_free * D;
  std::sort(keys, keys + n_keys);
  int n_succ = 0;
  for (int i = 0; i < n_keys; ++i) if ((uint32_t)(keys[i] >> 32) != 0xFFFFFFFFu) ++n_succ;
  *n_success = n_succ;
  const int k = elite_count(n_succ, c.elite_frac);
  if (n_succ == 0) {
    *best_idx = -1;
    if (c.do_update) for (int t = 0; t < nfd; ++t) sigma[t] = clamp_sigma(c, sigma[t] * c.inc);
    return 0;
  }
  const std::vector<float> mean0(mean, mean + nfd), sigma0(sigma, sigma + nfd), best0(best, best + nfd);
  const int best_ok0 = c.include_best ? *best_valid : 0;
  auto P = [&](int j, int t) {
    const int idx = (int)(uint32_t)(keys[j] & 0xFFFFFFFFu);
    return candidate_param(c, seed, iter, 0u, idx, candidate_kind(c, idx, best_ok0), t, mean0.data(), sigma0.data(), best0.data());
  };
  *best_idx = (int)(uint32_t)(keys[0] & 0xFFFFFFFFu);
  const float lk = log_poly((float)k + 0.5f);
  for (int j = 0; j < k; ++j) weights[j] = ces_log_weight(k, lk, j);
  const float sumw = wide_sum(k, [&](int j, float acc) { return acc + weights[j]; });
  for (int j = 0; j < k; ++j) weights[j] = weights[j] / sumw;
  for (int t = 0; t < nfd; ++t) best[t] = P(0, t);
  *best_valid = 1;
  if (c.do_update) {
    for (int t = 0; t < nfd; ++t) {
      const int d = t % D;
      const float em = wide_sum(k, [&](int j, float acc) { return fmaf(weights[j], P(j, t), acc); });
      const float m0 = mean[t];
      float nm = fmaf(c.mean_lr, em - m0, m0);
      if (c.mode == MODE_TASK && d == 2) nm = fmaxf(nm, c.dist_zmin);
      nm = fminf(fmaxf(nm, c.lo[d]), c.hi[d]);
      const bool wrap = (c.mode == MODE_TASK && d == 3 && c.lo[3] != c.hi[3]);
      const float var = wide_sum(k, [&](int j, float acc) {
        const float v = P(j, t);
        const float df = wrap ? wrap_angle_diff(v, nm, c.lo[3], c.hi[3]) : v - nm;
        return fmaf(weights[j], df * df, acc);
      });
      const float s0 = sigma[t];
      const float blend = fmaf(c.var_beta, var, (1.0f - c.var_beta) * (s0 * s0));
      float sg = clamp_sigma(c, sqrtf(blend));
      sg = clamp_sigma(c, sg * c.dec);
      mean[t] = nm; sigma[t] = sg;
    }
  }
  return k;
}

}  // extern "C"
