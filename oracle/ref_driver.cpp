// ref_driver.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Builds oracle/_ref/libref.so: the REFERENCE'S OWN
// hot-path headers, compiled unmodified from where they lie under /root/reference/include
//     sspp.h  Collision.h  utility.h  sspp/tsp.h  sspp/tsp_planner.h  sspp/tsp_sampler.h  sspp/tsp_path_model.h
//     sspp/tsp_evaluator.h  sspp/tsp_elites.h  sspp/tsp_distribution.h  sspp/tsp_collision_world.h  sspp/tsp_types.h
// behind a thin extern "C" driver.  Their third-party dependencies are absent from this container, so they are
// compiled against the stand-ins in oracle/shim/ (see each header there):
//     Eigen (Core / Dense / Geometry / unsupported Splines)  -> oracle/shim/eigen_shim.h, unsupported/Eigen/Splines
//     MuJoCo C API                                           -> oracle/shim/mujoco/mujoco.h, implemented below on top of
//                                                               oracle O1's restatement of mj_kinematics + mj_collision
//     <random>                                               -> oracle/shim/random: the engines hand the reference's
//                                                               sampler the Philox normals that the GPU and O1 use
// What this pins: every line of the reference's own code on the path — loop bounds, evaluation grids, the cost and
// status rules, elite count / partial_sort / CES weights, the distribution update and its clamp order, best forwarding,
// the adapter's argument quirks (tsp.h:53).  What stays restated (third-party arithmetic): the Eigen spline algorithms
// and MuJoCo's kinematics / narrow phase.
//
// Only tests/ and bench.py's CPU legs load this library.  It is built in THIS container only (the GPU box has no
// /root/reference); the built .so travels with the snapshot.
#include "o1_oracle.cpp"   // O1's FP64 kinematics / narrow phase / Philox (anonymous namespace: this TU only)

#include <omp.h>

#include <map>
#include <memory>
#include <sstream>
#include <iostream>
#include <cstring>

#include "mujoco/mujoco.h"
#include "sspp.h"
#include "sspp/tsp.h"

// ------------------------------------------------------------------------------------------
// the <random> tape
// ------------------------------------------------------------------------------------------
namespace refshim {
struct Tape {
  int mode = 0;                 // 0 JOINT (a stream per candidate), 1 TASK (a stream per (row, dim) of a candidate)
  uint64_t seed = 0; uint32_t iter = 0, query = 0;
  uint32_t cand0 = 0;           // candidate index of the first stream opened after arming
  int streams_per_row = 4;      // TASK: 3 when the yaw limits coincide (no yaw distribution is constructed)
  int rows = 1;                 // TASK: via points per candidate
  long opened = 0;              // streams opened since arming
  uint32_t cand = 0, stream = 0, draw = 0;
  long normals = 0, uniforms = 0;
};
static thread_local Tape g_tape;
void open_normal_stream() {
  Tape& t = g_tape;
  const long n = t.opened++;
  if (t.mode == 0) { t.cand = t.cand0 + (uint32_t)n; t.stream = 0; }
  else {
    const long per_cand = (long)t.streams_per_row * t.rows;
    const long k = n % per_cand;
    t.cand = t.cand0 + (uint32_t)(n / per_cand);
    t.stream = (uint32_t)((k / t.streams_per_row) * 4 + (k % t.streams_per_row));
  }
  t.draw = 0;
}
double draw_normal() { Tape& t = g_tape; ++t.normals; return normal_at(t.seed, t.iter, t.query, t.cand, t.stream, t.draw++); }
double draw_uniform() {
  Tape& t = g_tape; ++t.uniforms;
  uint32_t c[4] = {t.cand, t.query, t.iter, (t.stream << 8) | 25u};
  philox(c, (uint32_t)t.seed, (uint32_t)(t.seed >> 32));
  return uni(c[0]);
}
}  // namespace refshim

// ------------------------------------------------------------------------------------------
// MuJoCo stand-in
// ------------------------------------------------------------------------------------------
namespace {
struct ShimModel {
  Model m;                                  // O1's flat model
  std::vector<std::string> body_names, geom_names;
  std::vector<int> body_jntadr, jnt_type, jnt_qposadr, jnt_bodyid;
  mjModel api;
};
struct ShimData {
  std::vector<double> qpos, xpos, xquat, gxpos, gxmat;
  std::vector<mjContact> contact;
  std::vector<V3> bp; std::vector<Quat> bq; std::vector<Geom> G;    // scratch reused by every mj_forward
  mjData api;
};
const ShimModel& SM(const mjModel* m) { return *static_cast<const ShimModel*>(m->shim); }
void wire(ShimData* d) {
  d->api.qpos = d->qpos.data(); d->api.xpos = d->xpos.data(); d->api.xquat = d->xquat.data();
  d->api.geom_xpos = d->gxpos.data(); d->api.geom_xmat = d->gxmat.data(); d->api.contact = d->contact.data();
  d->api.shim = d;
}
}  // namespace

extern "C" {

mjModel* mj_loadXML(const char* filename, const void*, char* error, int error_sz) {
  if (error && error_sz > 0)
    std::snprintf(error, (size_t)error_sz, "oracle stand-in for MuJoCo cannot parse MJCF ('%s'): models come from a flat view", filename ? filename : "");
  return nullptr;
}
void mj_deleteModel(mjModel* m) { if (m) delete static_cast<ShimModel*>(m->shim); }
mjData* mj_makeData(const mjModel* m) {
  const ShimModel& s = SM(m);
  auto* d = new ShimData;
  d->qpos = s.m.qpos0;
  d->xpos.assign(3 * s.m.nbody, 0.0); d->xquat.assign(4 * s.m.nbody, 0.0);
  d->gxpos.assign(3 * s.m.ngeom, 0.0); d->gxmat.assign(9 * s.m.ngeom, 0.0);
  d->contact.resize(s.m.pairs.size() + 1);
  d->api.ncon = 0;
  wire(d);
  return &d->api;
}
mjData* mj_copyData(mjData* dest, const mjModel* m, const mjData* src) {
  mjData* d = dest ? dest : mj_makeData(m);
  auto* sd = static_cast<ShimData*>(d->shim);
  const auto* ss = static_cast<const ShimData*>(src->shim);
  sd->qpos = ss->qpos; sd->xpos = ss->xpos; sd->xquat = ss->xquat; sd->gxpos = ss->gxpos; sd->gxmat = ss->gxmat;
  sd->contact = ss->contact; sd->api.ncon = ss->api.ncon;
  wire(sd);
  return d;
}
void mj_deleteData(mjData* d) { if (d) delete static_cast<ShimData*>(d->shim); }

// positions + contacts (what the reference reads after mj_forward: sspp.h:143-144, Collision.h:86-101)
void mj_forward(const mjModel* mm, mjData* dd) {
  const Model& m = SM(mm).m;
  auto* d = static_cast<ShimData*>(dd->shim);
  kinematics(m, d->qpos, d->bp, d->bq);
  geom_frames(m, d->bp, d->bq, d->G);
  for (int b = 0; b < m.nbody; ++b) {
    d->xpos[3 * b] = d->bp[b].x; d->xpos[3 * b + 1] = d->bp[b].y; d->xpos[3 * b + 2] = d->bp[b].z;
    d->xquat[4 * b] = d->bq[b].w; d->xquat[4 * b + 1] = d->bq[b].x; d->xquat[4 * b + 2] = d->bq[b].y; d->xquat[4 * b + 3] = d->bq[b].z;
  }
  for (int g = 0; g < m.ngeom; ++g) {
    const Geom& G = d->G[g];
    d->gxpos[3 * g] = G.p.x; d->gxpos[3 * g + 1] = G.p.y; d->gxpos[3 * g + 2] = G.p.z;
    for (int j = 0; j < 3; ++j) { d->gxmat[9 * g + j] = G.ax[j].x; d->gxmat[9 * g + 3 + j] = G.ax[j].y; d->gxmat[9 * g + 6 + j] = G.ax[j].z; }
  }
  int ncon = 0;
  for (const auto& pr : m.pairs) {
    bool ok;
    const double dist = geom_dist(d->G[pr.first], d->G[pr.second], &ok);
    if (!ok) continue;
    const double margin = std::max(m.gmargin[pr.first], m.gmargin[pr.second]);
    if (!(dist < margin)) continue;
    mjContact& c = d->contact[ncon++];
    c.dist = dist; c.geom1 = c.geom[0] = pr.first; c.geom2 = c.geom[1] = pr.second;
  }
  dd->ncon = ncon;
}
int mj_name2id(const mjModel* m, int type, const char* name) {
  const ShimModel& s = SM(m);
  const auto& v = (type == mjOBJ_BODY) ? s.body_names : s.geom_names;
  if (type != mjOBJ_BODY && type != mjOBJ_GEOM) return -1;
  for (size_t i = 0; i < v.size(); ++i) if (v[i] == name) return (int)i;
  return -1;
}
const char* mj_id2name(const mjModel* m, int type, int id) {
  const ShimModel& s = SM(m);
  const auto& v = (type == mjOBJ_BODY) ? s.body_names : s.geom_names;
  if ((type != mjOBJ_BODY && type != mjOBJ_GEOM) || id < 0 || id >= (int)v.size() || v[id].empty()) return nullptr;
  return v[id].c_str();
}

// ------------------------------------------------------------------------------------------
// driver
// ------------------------------------------------------------------------------------------
void* ref_model_create(const FlatView* v, const char* const* body_names, const char* const* geom_names) {
  auto* s = new ShimModel;
  Model* tmp = static_cast<Model*>(o1_create(v, 0, 0, -1));
  s->m = *tmp;
  delete tmp;
  for (int b = 0; b < v->nbody; ++b) s->body_names.emplace_back(body_names && body_names[b] ? body_names[b] : "");
  for (int g = 0; g < v->ngeom; ++g) s->geom_names.emplace_back(geom_names && geom_names[g] ? geom_names[g] : "");
  s->body_jntadr.assign(v->nbody, -1); s->jnt_type.assign(v->nbody, 0); s->jnt_qposadr.assign(v->nbody, -1); s->jnt_bodyid.assign(v->nbody, -1);
  int nj = 0;
  for (int b = 0; b < v->nbody; ++b)
    if (v->body_jtype[b] >= 0) { s->body_jntadr[b] = nj; s->jnt_type[nj] = v->body_jtype[b]; s->jnt_qposadr[nj] = v->body_qadr[b]; s->jnt_bodyid[nj] = b; ++nj; }
  s->api.nq = v->nq; s->api.nbody = v->nbody; s->api.njnt = nj; s->api.ngeom = v->ngeom;
  s->api.body_jntadr = s->body_jntadr.data(); s->api.jnt_type = s->jnt_type.data(); s->api.jnt_qposadr = s->jnt_qposadr.data();
  s->api.jnt_bodyid = s->jnt_bodyid.data();
  s->api.shim = s;
  return &s->api;
}
void ref_model_destroy(void* m) { mj_deleteModel(static_cast<mjModel*>(m)); }
int ref_model_ncon(void* mm, const double* qpos_full) {   // ncon of the stand-in at a full qpos (unit check)
  mjModel* m = static_cast<mjModel*>(mm);
  mjData* d = mj_makeData(m);
  for (int i = 0; i < m->nq; ++i) d->qpos[i] = qpos_full[i];
  mj_forward(m, d);
  const int n = d->ncon;
  mj_deleteData(d);
  return n;
}

void ref_set_threads(int n) { omp_set_num_threads(n < 1 ? 1 : n); }
int ref_max_threads(void) { return omp_get_max_threads(); }

void ref_tape_arm(int mode, uint64_t seed, uint32_t iter, uint32_t query, uint32_t cand0, int streams_per_row, int rows) {
  refshim::Tape& t = refshim::g_tape;
  t = refshim::Tape();
  t.mode = mode; t.seed = seed; t.iter = iter; t.query = query; t.cand0 = cand0;
  t.streams_per_row = streams_per_row; t.rows = rows < 1 ? 1 : rows;
}
void ref_tape_counts(long* opened, long* normals, long* uniforms) {
  const refshim::Tape& t = refshim::g_tape;
  if (opened) *opened = t.opened;
  if (normals) *normals = t.normals;
  if (uniforms) *uniforms = t.uniforms;
}

// ---- the CES pieces, called directly (tsp_elites.h, tsp_distribution.h, tsp_sampler.h, tsp_evaluator.h, tsp_path_model.h) ----
static tsp::Point P4(const double* p) { tsp::Point q; for (int d = 0; d < 4; ++d) q[d] = p[d]; return q; }

// EliteSelector::select + weights over n successful candidates with total costs `cost`; returns k
int ref_elites(const double* cost, int n, double frac, int64_t* idx_out, double* w_out) {
  std::vector<tsp::PathCandidate> C((size_t)n);
  for (int i = 0; i < n; ++i) { C[i].L = cost[i]; C[i].C_wf = 0.0; C[i].C_nf = 0.0; C[i].status = tsp::SolverStatus::Converged; }
  tsp::EliteSelector sel;
  const auto idx = sel.select(C, frac, [](const tsp::PathCandidate& c) { return c.L + 1.0 * c.C_wf; });
  const auto w = sel.weights((int)idx.size());
  for (size_t i = 0; i < idx.size(); ++i) { idx_out[i] = (int64_t)idx[i]; w_out[i] = w[i]; }
  return (int)idx.size();
}

struct RefDistCfg { double sigma_floor, var_beta, mean_lr, stddev_min, stddev_max, z_min, lo[4], hi[4]; };
static tsp::Distribution make_dist(const RefDistCfg* c) {
  tsp::Distribution d;
  d.sigma_floor = c->sigma_floor; d.var_ema_beta = c->var_beta; d.mean_lr = c->mean_lr; d.stddev_min = c->stddev_min;
  d.stddev_max = c->stddev_max; d.z_min = c->z_min; d.lo = P4(c->lo); d.hi = P4(c->hi);
  return d;
}
// Distribution::reset(mean0, s0): mean/sigma [Kv][4] in/out
void ref_dist_reset(const RefDistCfg* c, int Kv, const double* mean0, double s0, double* mean, double* sigma) {
  tsp::Distribution d = make_dist(c);
  tsp::ViaSet m0;
  for (int i = 0; i < Kv; ++i) m0.push_back(P4(mean0 + 4 * i));
  d.reset(m0, s0);
  for (int i = 0; i < Kv; ++i) for (int k = 0; k < 4; ++k) { mean[4 * i + k] = d.mean_vias[i][k]; sigma[4 * i + k] = d.sigma_vias[i][k]; }
}
// Distribution::update(C, elite_idx, w, get_vias) then adapt(success, inc, dec) exactly as Planner::plan does
// (tsp_planner.h:122-141).  vias [n][Kv][4] of the n successes; elite_idx [k]; mean/sigma [Kv][4] in/out.
// k == 0: adapt(false) only.
void ref_dist_update(const RefDistCfg* c, int Kv, int n, const double* vias, int k, const int64_t* elite_idx, const double* w,
                     double inc, double dec, double* mean, double* sigma) {
  tsp::Distribution d = make_dist(c);
  for (int i = 0; i < Kv; ++i) { d.mean_vias.push_back(P4(mean + 4 * i)); d.sigma_vias.push_back(P4(sigma + 4 * i)); }
  if (k > 0) {
    std::vector<tsp::PathCandidate> C((size_t)n);
    for (int i = 0; i < n; ++i) for (int j = 0; j < Kv; ++j) C[i].via.push_back(P4(vias + ((size_t)i * Kv + j) * 4));
    std::vector<size_t> idx(elite_idx, elite_idx + k);
    std::vector<double> ww(w, w + k);
    d.update(C, idx, ww, [](const tsp::PathCandidate& pc) { return pc.refined ? *pc.refined : pc.via; });
    d.adapt(true, inc, dec);
  } else d.adapt(false, inc, dec);
  for (int i = 0; i < Kv; ++i) for (int kk = 0; kk < 4; ++kk) { mean[4 * i + kk] = d.mean_vias[i][kk]; sigma[4 * i + kk] = d.sigma_vias[i][kk]; }
}
double ref_wrap_angle_diff(double a, double b, double lo, double hi) { return tsp::Distribution::wrap_angle_diff(a, b, lo, hi); }

// Sampler::sample_set for n candidates in sequence (arm the tape first): out [n][Kv][4]
void ref_sample_sets(const double* lo, const double* hi, int Kv, const double* mean, const double* sigma, double z_min, int n, double* out) {
  tsp::Sampler s(P4(lo), P4(hi));
  std::vector<tsp::Point> m, sg;
  for (int i = 0; i < Kv; ++i) { m.push_back(P4(mean + 4 * i)); sg.push_back(P4(sigma + 4 * i)); }
  for (int c = 0; c < n; ++c) {
    const tsp::ViaSet v = s.sample_set(m, sg, z_min);
    for (int i = 0; i < Kv; ++i) for (int k = 0; k < 4; ++k) out[((size_t)c * Kv + i) * 4 + k] = v[i][k];
  }
}
double ref_floor_penalty(double z_min, double margin, double scale, const double* p) {
  tsp::Evaluator e; e.z_min = z_min; e.floor_margin = margin; e.floor_scale = scale;
  return e.floorPenalty(P4(p));
}
// PathModel::initLinear + fromVias: control points [4][total] (column-major as Eigen stores them), knots, and the
// spline at n_u parameters
void ref_from_vias(const double* start, const double* end, int total_points, const double* vias, int n_u, const double* u,
                   double* ctrl_out, double* knots_out, double* pts_out) {
  tsp::PathModel pm;
  pm.initLinear(P4(start), P4(end), total_points);
  tsp::ViaSet vs;
  for (int i = 0; i < total_points - 2; ++i) vs.push_back(P4(vias + 4 * i));
  const tsp::Spline s = pm.fromVias(vs);
  if (ctrl_out) for (int j = 0; j < total_points; ++j) for (int d = 0; d < 4; ++d) ctrl_out[j * 4 + d] = s.ctrls()(d, j);
  if (knots_out) for (int i = 0; i < (int)s.knots().size(); ++i) knots_out[i] = s.knots()(i);
  for (int i = 0; i < n_u; ++i) { const tsp::Point p = s(u[i]); for (int d = 0; d < 4; ++d) pts_out[4 * i + d] = p[d]; }
}
// Collision<Point>::collision_point_cost / check_collision_point at n task-space points
void ref_point_costs(void* model, const char* body, const double* pts, int n, int use_center, double* cost_out) {
  Collision<tsp::Point> col(body, static_cast<mjModel*>(model));
  for (int i = 0; i < n; ++i) cost_out[i] = col.collision_point_cost(P4(pts + 4 * i), use_center != 0);
}

// Collision<Point>::check_collision_point (Collision.h:51-82) at n task-space points.  The function reports through std::cout;
// its text is captured and parsed: hit[i] = its return value, g1 / g2 / dist = the ids and distance it printed.
void ref_check_points(void* model, const char* body, const double* pts, int n, int32_t* hit, int32_t* g1, int32_t* g2, double* dist) {
  Collision<tsp::Point> col(body, static_cast<mjModel*>(model));
  for (int i = 0; i < n; ++i) {
    std::ostringstream cap;
    std::streambuf* old = std::cout.rdbuf(cap.rdbuf());
    const bool h = col.check_collision_point(P4(pts + 4 * i));
    std::cout.rdbuf(old);
    hit[i] = h ? 1 : 0; g1[i] = g2[i] = -1; dist[i] = 0.0;
    if (!h) continue;
    const std::string t = cap.str();
    auto after = [&](const char* key) { const size_t p = t.find(key); return p == std::string::npos ? std::string() : t.substr(p + std::strlen(key)); };
    dist[i] = std::atof(after("(col_dist): ").c_str());
    g1[i] = std::atoi(after("Geom 1 ID: ").c_str());
    g2[i] = std::atoi(after("Geom 2 ID: ").c_str());
  }
}

// ---- Planner B whole: tsp::TaskSpacePlanner (tsp.h adapter over tsp_planner.h) ----
struct RefTspArgs {
  double stddev_initial, stddev_min, stddev_max, inc, dec, elite_fraction;
  int sample_count, check_points, init_points;
  double collision_weight, z_min, lo[4], hi[4], sigma_floor, var_ema_beta, mean_lr, floor_margin, floor_penalty_scale;
};
struct RefTsp { std::unique_ptr<tsp::TaskSpacePlanner> p; RefTspArgs a; };
void* ref_tsp_create(void* model, const char* body, const RefTspArgs* a) {
  auto* h = new RefTsp;
  h->a = *a;
  h->p = std::make_unique<tsp::TaskSpacePlanner>(static_cast<mjModel*>(model), std::string(body), a->stddev_initial, a->stddev_min, a->stddev_max,
                                                 a->inc, a->dec, a->elite_fraction, a->sample_count, a->check_points, 0, a->init_points,
                                                 a->collision_weight, a->z_min, P4(a->lo), P4(a->hi), false, a->sigma_floor, a->var_ema_beta,
                                                 a->mean_lr, 0.1, a->floor_margin, a->floor_penalty_scale);
  return h;
}
void ref_tsp_destroy(void* h) { delete static_cast<RefTsp*>(h); }
// One plan(start, end, iterate) call.  Outputs per seed, in seed order (tsp_planner.h:78-91: mean, forwarded best if any,
// samples): vias [n][Kv][4], L, Cnf, Cwf, status (0 converged, 1 failed); mean / sigma of the FIRST via after the update
// (all the adapter exposes); the planner's best path at n_u parameters.  Returns the number of seeds.
int ref_tsp_plan(void* hh, const double* start, const double* end, int iterate, int cap, double* vias, double* L, double* Cnf, double* Cwf,
                 int32_t* status, double* mean0, double* sigma0, int n_u, const double* u, double* path_pts) {
  RefTsp* h = static_cast<RefTsp*>(hh);
  h->p->plan(P4(start), P4(end), iterate != 0);
  const auto& seeds = h->p->get_sampled_via_sets();
  const auto succ = h->p->get_succesful_path_candidates();
  const auto fail = h->p->get_failed_path_candidates();
  const int n = (int)seeds.size();
  if (n > cap) return -n;
  const int Kv = h->a.init_points - 2;
  // with one OpenMP thread both lists keep seed order: walk them beside the seeds
  size_t is = 0, jf = 0;
  auto same = [&](const tsp::ViaSet& a, const tsp::ViaSet& b) {
    if (a.size() != b.size()) return false;
    for (size_t i = 0; i < a.size(); ++i) for (int d = 0; d < 4; ++d) if (a[i][d] != b[i][d]) return false;
    return true;
  };
  for (int i = 0; i < n; ++i) {
    const tsp::PathCandidate* c = nullptr;
    if (is < succ.size() && same(succ[is].via, seeds[i])) c = &succ[is++];
    else if (jf < fail.size() && same(fail[jf].via, seeds[i])) c = &fail[jf++];
    if (!c) return -1000000 - i;                      // order assumption broken (more than one OpenMP thread?)
    for (int j = 0; j < Kv; ++j) for (int d = 0; d < 4; ++d) vias[((size_t)i * Kv + j) * 4 + d] = seeds[i][j][d];
    L[i] = c->L; Cnf[i] = c->C_nf; Cwf[i] = c->C_wf; status[i] = (c->status == tsp::SolverStatus::Converged) ? 0 : 1;
  }
  const tsp::Point m = h->p->get_current_mean(), s = h->p->get_current_stddev();
  for (int d = 0; d < 4; ++d) { mean0[d] = m[d]; sigma0[d] = s[d]; }
  for (int i = 0; i < n_u; ++i) { const tsp::Point p = h->p->evaluate(u[i]); for (int d = 0; d < 4; ++d) path_pts[4 * i + d] = p[d]; }
  return n;
}
// timing leg: iterations of plan() with whatever thread count is set; returns scored points
long ref_tsp_run(void* hh, const double* start, const double* end, int iters, int cold) {
  RefTsp* h = static_cast<RefTsp*>(hh);
  long pts = 0;
  for (int it = 0; it < iters; ++it) {
    h->p->plan(P4(start), P4(end), !(cold && it == 0));
    pts += (long)h->p->get_sampled_via_sets().size() * h->a.check_points;
  }
  return pts;
}

}  // extern "C"

// ---- Planner A whole: sspp::SamplingPathPlanner<kDOF> (sspp.h) ----
struct RefSsppBase {
  virtual ~RefSsppBase() {}
  virtual int dof() const = 0;
  virtual int candidates(const double* start, const double* end, double sigma, const double* limits, int n, int M, int K,
                         double* ctrl, int32_t* collides, double* arc) = 0;
  virtual int plan(const double* start, const double* end, double sigma, const double* limits, int n, int M, int K, int cap,
                   double* succ_ctrl, double* best_ctrl, int* found) = 0;
  virtual void init_path(const double* start, const double* end, int K, double* ctrl, double* knots) = 0;
};
template <int D>
struct RefSspp : RefSsppBase {
  using PL = sspp::SamplingPathPlanner<D>;
  using Point = typename PL::Point;
  using Spline = typename PL::Spline;
  mjModel* model; mjData* data; std::unique_ptr<PL> p;
  explicit RefSspp(mjModel* m) : model(m), data(mj_makeData(m)), p(std::make_unique<PL>(m, data)) {}
  ~RefSspp() override { p.reset(); mj_deleteData(data); }
  int dof() const override { return D; }
  static Point PD(const double* v) { Point q; for (int d = 0; d < D; ++d) q[d] = v[d]; return q; }
  void init_path(const double* start, const double* end, int K, double* ctrl, double* knots) override {
    Spline s;
    p->initializePath(PD(start), PD(end), s, K);
    for (int j = 0; j < K; ++j) for (int d = 0; d < D; ++d) ctrl[j * D + d] = s.ctrls()(d, j);
    if (knots) for (int i = 0; i < (int)s.knots().size(); ++i) knots[i] = s.knots()(i);
  }
  // the body of plan()'s loop (sspp.h:208-216), candidate by candidate, plus computeArcLength of every candidate:
  // ctrl [n][K][D], collides [n], arc [n]
  int candidates(const double* start, const double* end, double sigma, const double* limits, int n, int M, int K,
                 double* ctrl, int32_t* collides, double* arc) override {
    Spline init;
    p->initializePath(PD(start), PD(end), init, K);
    std::default_random_engine gen(0);
    const Point lim = PD(limits);
    for (int i = 0; i < n; ++i) {
      const Spline s = p->sampleWithNoise(init, sigma, lim, gen);
      for (int j = 0; j < K; ++j) for (int d = 0; d < D; ++d) ctrl[((size_t)i * K + j) * D + d] = s.ctrls()(d, j);
      collides[i] = p->checkCollision(s, M, data) ? 1 : 0;
      arc[i] = p->computeArcLength(s, M);
    }
    return n;
  }
  int plan(const double* start, const double* end, double sigma, const double* limits, int n, int M, int K, int cap,
           double* succ_ctrl, double* best_ctrl, int* found) override {
    std::vector<Spline> succ;
    const bool ok = p->plan(PD(start), PD(end), sigma, PD(limits), succ, n, M, K);
    *found = ok ? 1 : 0;
    const int ns = (int)succ.size();
    for (int i = 0; i < ns && i < cap; ++i)
      for (int j = 0; j < K; ++j) for (int d = 0; d < D; ++d) succ_ctrl[((size_t)i * K + j) * D + d] = succ[i].ctrls()(d, j);
    if (ok) { const auto c = p->get_ctrl_pts(); for (int j = 0; j < K; ++j) for (int d = 0; d < D; ++d) best_ctrl[j * D + d] = c(d, j); }
    return ns;
  }
};
extern "C" {
void* ref_sspp_create(void* model, int dof) {
  mjModel* m = static_cast<mjModel*>(model);
  switch (dof) {                                  // the instantiations the reference binds (src/sspp_bindings.cpp:58-61)
    case 3: return static_cast<RefSsppBase*>(new RefSspp<3>(m));
    case 6: return static_cast<RefSsppBase*>(new RefSspp<6>(m));
    case 7: return static_cast<RefSsppBase*>(new RefSspp<7>(m));
    case 9: return static_cast<RefSsppBase*>(new RefSspp<9>(m));
    default: return nullptr;
  }
}
void ref_sspp_destroy(void* h) { delete static_cast<RefSsppBase*>(h); }
void ref_sspp_init_path(void* h, const double* start, const double* end, int K, double* ctrl, double* knots) {
  static_cast<RefSsppBase*>(h)->init_path(start, end, K, ctrl, knots);
}
int ref_sspp_candidates(void* h, const double* start, const double* end, double sigma, const double* limits, int n, int M, int K,
                        double* ctrl, int32_t* collides, double* arc) {
  return static_cast<RefSsppBase*>(h)->candidates(start, end, sigma, limits, n, M, K, ctrl, collides, arc);
}
int ref_sspp_plan(void* h, const double* start, const double* end, double sigma, const double* limits, int n, int M, int K, int cap,
                  double* succ_ctrl, double* best_ctrl, int* found) {
  return static_cast<RefSsppBase*>(h)->plan(start, end, sigma, limits, n, M, K, cap, succ_ctrl, best_ctrl, found);
}

}  // extern "C"
