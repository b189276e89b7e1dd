// o1_oracle.cpp — TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Oracle O1: a plain FP64 CPU restatement
// of the reference's candidate-scoring path, written independently of sspp_b200/csrc (it shares
// no header with the product).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load it.
//
// PARITY STATUS: "parity unpinned" for kinematics / contacts / elites — the reference delegates
// those to MuJoCo's mj_forward (third party, pinned only as mujoco==3.2.5 in requirements.txt:3,
// absent from /root/reference and from this container) and holds no golden vector for them.
// The spline algebra IS pinned: against scripts/bspline_params.npy and sspp/tests/test_BSplines.py
// through oracle/bsplines_np.py and tests/test_oracle_spline.py.
//
// What follows the reference, loop for loop (paths under /root/reference):
//   o1_score_joint      include/sspp.h:132-169  checkCollision (i = 0..M inclusive, ncon > 0) + computeArcLength
//   o1_score_task       include/sspp/tsp_evaluator.h:17-32 eval_one_pass + floorPenalty :10-14
//   point cost (TASK)   include/Collision.h:84-103 (sum over contacts with dist < -1e-3 of -1/(centre dist + 1e-4))
//   set point           include/utility.h:149-175,198-206 (x,y,z,yaw -> free-joint qpos)
//   o1_elites           include/sspp/tsp_elites.h:13-32
//   o1_update           include/sspp/tsp_distribution.h:31-83, include/sspp/tsp_planner.h:122-141
//   o1_sample_task      include/sspp/tsp_sampler.h:12-51;  o1_sample_joint  include/sspp.h:114-130
//   spline              Eigen::Spline / SplineFitting semantics (Piegl-Tiller A2.1/A2.2, knot averaging)
// MuJoCo semantics restated: mj_kinematics (body tree, one joint per body), geom frames,
// collision filtering (contype/conaffinity, same weld body, parent-child, <exclude>), contact kept
// iff dist < max(margin1, margin2).  One contact per geom pair (MuJoCo's multi-contact manifolds
// are not reproduced; they only scale the cost of candidates that already failed).
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <numeric>
#include <string>
#include <vector>

namespace {

struct FlatView {   // must mirror SsppFlatModelView (include/sspp_cuda.h)
  int32_t nbody, ngeom, nq, nexclude;
  const int32_t *body_parent, *body_weld, *body_jtype, *body_qadr;
  const double *body_pos, *body_quat, *jnt_axis, *jnt_pos, *jnt_range, *qpos0;
  const int32_t *geom_body, *geom_type, *geom_contype, *geom_conaffinity;
  const double *geom_size, *geom_pos, *geom_quat, *geom_margin;
  const int32_t* exclude;
};

struct V3 { double x, y, z; };
V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
V3 operator*(double s, V3 a) { return {s * a.x, s * a.y, s * a.z}; }
double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
double norm(V3 a) { return std::sqrt(dot(a, a)); }

struct Quat { double w, x, y, z; };
Quat qmul(Quat a, Quat b) {
  return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z, a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
          a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x, a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w};
}
Quat qnorm(Quat q) { double n = std::sqrt(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z); return {q.w / n, q.x / n, q.y / n, q.z / n}; }
V3 qrot(Quat q, V3 v) {   // v' = q v q*
  V3 u{q.x, q.y, q.z};
  V3 t = 2.0 * cross(u, v);
  return v + q.w * t + cross(u, t);
}

struct Geom { int type; V3 p; V3 ax[3]; V3 s; };   // world frame: centre, axes, size

struct Model {
  int nbody, ngeom, nq;
  std::vector<int> parent, weld, jtype, qadr, gbody, gtype, gct, gca, excl;
  std::vector<double> bpos, bquat, jaxis, jpos, qpos0, gsize, gpos, gquat, gmargin;
  int mode, dof, plan_body;
  std::vector<std::pair<int, int>> pairs;   // geom pairs that pass MuJoCo's filters
  // SDF collision mode (build-defined, not in the reference): FP64 grid of the static partners + sample spheres
  bool sdf = false;
  int nx = 0, ny = 0, nz = 0;
  double glo[3], gh[3];
  std::vector<double> nodes;
  struct Sph { V3 c; double r; int body; };
  std::vector<Sph> spheres;
  double sdf_margin = 0.0;
};

std::vector<char> moving_bodies(const Model& m) {
  std::vector<char> mv(m.nbody, 0);
  for (int b = 1; b < m.nbody; ++b) {
    const bool planned = (m.mode == 0) ? ((m.jtype[b] == 2 || m.jtype[b] == 3) && m.qadr[b] >= 0 && m.qadr[b] < m.dof) : (b == m.plan_body);
    mv[b] = planned || mv[m.parent[b]];
  }
  return mv;
}

bool filter_pass(const Model& m, int g1, int g2) {
  const int b1 = m.gbody[g1], b2 = m.gbody[g2];
  const int w1 = m.weld[b1], w2 = m.weld[b2];
  if (w1 == w2) return false;                                            // same (welded) body
  if (w1 != 0 && w2 != 0) {                                              // parent-child, world excepted
    if (m.weld[m.parent[w1]] == w2 || m.weld[m.parent[w2]] == w1) return false;
  }
  for (size_t k = 0; k + 1 < m.excl.size(); k += 2) {
    if ((m.excl[k] == b1 && m.excl[k + 1] == b2) || (m.excl[k] == b2 && m.excl[k + 1] == b1)) return false;
  }
  if (m.gtype[g1] == 0 && m.gtype[g2] == 0) return false;
  return ((m.gct[g1] & m.gca[g2]) != 0) || ((m.gct[g2] & m.gca[g1]) != 0);
}

// mj_kinematics for the supported joint types; qpos is the full coordinate vector
void kinematics(const Model& m, const std::vector<double>& qpos, std::vector<V3>& xpos, std::vector<Quat>& xquat) {
  xpos.assign(m.nbody, V3{0, 0, 0});
  xquat.assign(m.nbody, Quat{1, 0, 0, 0});
  for (int b = 1; b < m.nbody; ++b) {
    const int p = m.parent[b];
    const V3 lp{m.bpos[3 * b], m.bpos[3 * b + 1], m.bpos[3 * b + 2]};
    const Quat lq = qnorm(Quat{m.bquat[4 * b], m.bquat[4 * b + 1], m.bquat[4 * b + 2], m.bquat[4 * b + 3]});
    if (m.jtype[b] == 0) {   // free joint: qpos holds the world pose
      const double* q = &qpos[m.qadr[b]];
      xpos[b] = V3{q[0], q[1], q[2]};
      xquat[b] = qnorm(Quat{q[3], q[4], q[5], q[6]});
      continue;
    }
    xpos[b] = xpos[p] + qrot(xquat[p], lp);
    xquat[b] = qmul(xquat[p], lq);
    if (m.jtype[b] == 3) {   // hinge about axis through the anchor
      const double th = qpos[m.qadr[b]];
      const V3 a{m.jaxis[3 * b], m.jaxis[3 * b + 1], m.jaxis[3 * b + 2]};
      const V3 anchor_local{m.jpos[3 * b], m.jpos[3 * b + 1], m.jpos[3 * b + 2]};
      const V3 anchor_world = xpos[b] + qrot(xquat[b], anchor_local);
      const Quat jq{std::cos(th / 2), a.x * std::sin(th / 2), a.y * std::sin(th / 2), a.z * std::sin(th / 2)};
      xquat[b] = qnorm(qmul(xquat[b], jq));
      xpos[b] = anchor_world - qrot(xquat[b], anchor_local);
    } else if (m.jtype[b] == 2) {   // slide
      const V3 a{m.jaxis[3 * b], m.jaxis[3 * b + 1], m.jaxis[3 * b + 2]};
      xpos[b] = xpos[b] + qpos[m.qadr[b]] * qrot(xquat[b], a);
    }
  }
}

void geom_frames(const Model& m, const std::vector<V3>& xpos, const std::vector<Quat>& xquat, std::vector<Geom>& out) {
  out.resize(m.ngeom);
  for (int g = 0; g < m.ngeom; ++g) {
    const int b = m.gbody[g];
    const Quat gq = qnorm(Quat{m.gquat[4 * g], m.gquat[4 * g + 1], m.gquat[4 * g + 2], m.gquat[4 * g + 3]});
    const Quat wq = qmul(xquat[b], gq);
    Geom& G = out[g];
    G.type = m.gtype[g];
    G.p = xpos[b] + qrot(xquat[b], V3{m.gpos[3 * g], m.gpos[3 * g + 1], m.gpos[3 * g + 2]});
    G.ax[0] = qrot(wq, V3{1, 0, 0}); G.ax[1] = qrot(wq, V3{0, 1, 0}); G.ax[2] = qrot(wq, V3{0, 0, 1});
    G.s = V3{m.gsize[3 * g], m.gsize[3 * g + 1], m.gsize[3 * g + 2]};
  }
}

// full qpos from the planned coordinates
void make_qpos(const Model& m, const double* q, std::vector<double>& qp) {
  qp = m.qpos0;
  if (m.mode == 0) { for (int d = 0; d < m.dof; ++d) qp[d] = q[d]; }
  else {   // utility.h:149-175: x,y,z then quaternion of a pure yaw
    const int a = m.qadr[m.plan_body];
    qp[a] = q[0]; qp[a + 1] = q[1]; qp[a + 2] = q[2];
    qp[a + 3] = std::cos(q[3] / 2); qp[a + 4] = 0; qp[a + 5] = 0; qp[a + 6] = std::sin(q[3] / 2);
  }
}
std::vector<double> make_qpos(const Model& m, const double* q) { std::vector<double> qp; make_qpos(m, q, qp); return qp; }

// ---- narrow phase (signed distance, negative = penetration depth) ----
V3 to_local(const Geom& B, V3 w) { V3 d = w - B.p; return {dot(d, B.ax[0]), dot(d, B.ax[1]), dot(d, B.ax[2])}; }
double point_box(V3 x, V3 h) {
  const double qx = std::fabs(x.x) - h.x, qy = std::fabs(x.y) - h.y, qz = std::fabs(x.z) - h.z;
  const double ox = std::max(qx, 0.0), oy = std::max(qy, 0.0), oz = std::max(qz, 0.0);
  const double out = std::sqrt(ox * ox + oy * oy + oz * oz);
  return out > 0 ? out : std::max(qx, std::max(qy, qz));
}
double point_segment(V3 x, V3 a, V3 b) {
  const V3 d = b - a;
  const double dd = dot(d, d);
  double t = dd > 0 ? dot(x - a, d) / dd : 0.0;
  t = std::min(1.0, std::max(0.0, t));
  return norm(x - (a + t * d));
}
double segment_segment(V3 a0, V3 a1, V3 b0, V3 b1) {
  // the minimum is attained at an endpoint of one segment or at the interior stationary point
  double best = std::min(std::min(point_segment(a0, b0, b1), point_segment(a1, b0, b1)),
                         std::min(point_segment(b0, a0, a1), point_segment(b1, a0, a1)));
  const V3 u = a1 - a0, v = b1 - b0, w = a0 - b0;
  const double A = dot(u, u), B = dot(u, v), C = dot(v, v), D = dot(u, w), E = dot(v, w);
  const double den = A * C - B * B;
  if (den > 1e-18) {
    const double s = (B * E - C * D) / den, t = (A * E - B * D) / den;
    if (s >= 0 && s <= 1 && t >= 0 && t <= 1) best = std::min(best, norm((a0 + s * u) - (b0 + t * v)));
  }
  return best;
}
double segment_box(V3 p0, V3 p1, V3 h) {   // box frame; convex in t: golden-section search
  auto f = [&](double t) {
    const V3 x = p0 + t * (p1 - p0);
    const double ex = std::max(std::fabs(x.x) - h.x, 0.0), ey = std::max(std::fabs(x.y) - h.y, 0.0), ez = std::max(std::fabs(x.z) - h.z, 0.0);
    return std::sqrt(ex * ex + ey * ey + ez * ez);
  };
  double lo = 0, hi = 1;
  const double gr = (std::sqrt(5.0) - 1) / 2;
  double c = hi - gr * (hi - lo), d = lo + gr * (hi - lo), fc = f(c), fd = f(d);
  for (int it = 0; it < 200 && hi - lo > 1e-14; ++it) {   // the bracket shrinks by 0.618 per step: 67 steps reach 1e-14
    if (fc <= fd) { hi = d; d = c; fd = fc; c = hi - gr * (hi - lo); fc = f(c); }
    else { lo = c; c = d; fc = fd; d = lo + gr * (hi - lo); fd = f(d); }
  }
  return std::min(std::min(f(0.0), f(1.0)), f(0.5 * (lo + hi)));
}
double box_support(const Geom& B, V3 L) { return B.s.x * std::fabs(dot(L, B.ax[0])) + B.s.y * std::fabs(dot(L, B.ax[1])) + B.s.z * std::fabs(dot(L, B.ax[2])); }
double cyl_support(const Geom& C, V3 L) {   // L need not be unit
  const double la = dot(L, C.ax[2]);
  return C.s.y * std::fabs(la) + C.s.x * std::sqrt(std::max(dot(L, L) - la * la, 0.0));
}
double box_box(const Geom& A, const Geom& B) {   // 15-axis SAT: max separation / min penetration
  const V3 t = B.p - A.p;
  double best = -std::numeric_limits<double>::infinity();
  auto axis = [&](V3 L) {
    const double n = norm(L);
    if (n * n <= 1e-6) return;
    const V3 U = (1.0 / n) * L;
    best = std::max(best, std::fabs(dot(t, U)) - box_support(A, U) - box_support(B, U));
  };
  for (int i = 0; i < 3; ++i) axis(A.ax[i]);
  for (int j = 0; j < 3; ++j) axis(B.ax[j]);
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) axis(cross(A.ax[i], B.ax[j]));
  return best;
}
double cyl_box(const Geom& C, const Geom& B) {   // support-function SAT over the documented direction set
  const V3 t = C.p - B.p;
  double best = -std::numeric_limits<double>::infinity();
  auto axis = [&](V3 L, double min_n2) {
    const double n2 = dot(L, L);
    if (n2 <= min_n2) return;
    const double n = std::sqrt(n2);
    best = std::max(best, (std::fabs(dot(t, L)) - box_support(B, L) - cyl_support(C, L)) / n);
  };
  for (int i = 0; i < 3; ++i) axis(B.ax[i], 0.0);
  axis(C.ax[2], 0.0);
  for (int i = 0; i < 3; ++i) axis(cross(C.ax[2], B.ax[i]), 1e-6);
  const V3 x = to_local(B, C.p);
  const V3 cp{std::min(std::max(x.x, -B.s.x), B.s.x), std::min(std::max(x.y, -B.s.y), B.s.y), std::min(std::max(x.z, -B.s.z), B.s.z)};
  const V3 dl = x - cp;
  axis(dl.x * B.ax[0] + dl.y * B.ax[1] + dl.z * B.ax[2], 1e-12);
  return best;
}
double point_cyl(V3 x, const Geom& C) {
  const V3 w = x - C.p;
  const double z = dot(w, C.ax[2]);
  const double rho = norm(w - z * C.ax[2]);
  const double dz = std::fabs(z) - C.s.y, dr = rho - C.s.x;
  const double oz = std::max(dz, 0.0), orr = std::max(dr, 0.0);
  const double out = std::sqrt(oz * oz + orr * orr);
  return out > 0 ? out : std::max(dz, dr);
}
double geom_dist(const Geom& A0, const Geom& B0, bool* ok) {
  *ok = true;
  const Geom& A = (A0.type <= B0.type) ? A0 : B0;
  const Geom& B = (A0.type <= B0.type) ? B0 : A0;
  const int ta = A.type, tb = B.type;
  if (ta == 0) {
    const V3 n = A.ax[2];
    const double d0 = dot(n, B.p - A.p);
    if (tb == 2) return d0 - B.s.x;
    if (tb == 3) return d0 - B.s.y * std::fabs(dot(n, B.ax[2])) - B.s.x;
    if (tb == 5) { const double na = dot(n, B.ax[2]); return d0 - B.s.y * std::fabs(na) - B.s.x * std::sqrt(std::max(1 - na * na, 0.0)); }
    if (tb == 6) return d0 - box_support(B, n);
  }
  if (ta == 2) {
    if (tb == 2) return norm(B.p - A.p) - A.s.x - B.s.x;
    if (tb == 3) return point_segment(A.p, B.p - B.s.y * B.ax[2], B.p + B.s.y * B.ax[2]) - A.s.x - B.s.x;
    if (tb == 5) return point_cyl(A.p, B) - A.s.x;
    if (tb == 6) return point_box(to_local(B, A.p), B.s) - A.s.x;
  }
  if (ta == 3) {
    const V3 a0 = A.p - A.s.y * A.ax[2], a1 = A.p + A.s.y * A.ax[2];
    if (tb == 3) return segment_segment(a0, a1, B.p - B.s.y * B.ax[2], B.p + B.s.y * B.ax[2]) - A.s.x - B.s.x;
    if (tb == 6) return segment_box(to_local(B, a0), to_local(B, a1), B.s) - A.s.x;
  }
  if (ta == 5 && tb == 6) return cyl_box(A, B);
  if (ta == 6 && tb == 6) return box_box(A, B);
  *ok = false;
  return std::numeric_limits<double>::infinity();
}

double point_geom(V3 x, const Geom& G) {   // exact signed distance of a point to a primitive
  switch (G.type) {
    case 0: return dot(G.ax[2], x - G.p);
    case 2: return norm(x - G.p) - G.s.x;
    case 3: return point_segment(x, G.p - G.s.y * G.ax[2], G.p + G.s.y * G.ax[2]) - G.s.x;
    case 5: return point_cyl(x, G);
    default: return point_box(to_local(G, x), G.s);
  }
}
double sdf_value(const Model& m, V3 x) {   // trilinear interpolation, coordinates clamped to the grid
  const double g[3] = {std::min(std::max((x.x - m.glo[0]) / m.gh[0], 0.0), (double)(m.nx - 1)),
                       std::min(std::max((x.y - m.glo[1]) / m.gh[1], 0.0), (double)(m.ny - 1)),
                       std::min(std::max((x.z - m.glo[2]) / m.gh[2], 0.0), (double)(m.nz - 1))};
  const int i = std::min((int)g[0], m.nx - 2), j = std::min((int)g[1], m.ny - 2), k = std::min((int)g[2], m.nz - 2);
  const double fx = g[0] - i, fy = g[1] - j, fz = g[2] - k;
  auto N = [&](int a, int b, int c) { return m.nodes[((size_t)(k + c) * m.ny + (j + b)) * m.nx + (i + a)]; };
  const double c00 = N(0, 0, 0) * (1 - fx) + N(1, 0, 0) * fx, c10 = N(0, 1, 0) * (1 - fx) + N(1, 1, 0) * fx;
  const double c01 = N(0, 0, 1) * (1 - fx) + N(1, 0, 1) * fx, c11 = N(0, 1, 1) * (1 - fx) + N(1, 1, 1) * fx;
  return (c00 * (1 - fy) + c10 * fy) * (1 - fz) + (c01 * (1 - fy) + c11 * fy) * fz;
}

// collision cost of one configuration: the whole scene, as mj_forward + the reference's read-out
// band_out (optional): number of tested pairs / sphere samples whose signed distance lies within EPS_BAND of a decision
// threshold (planner A: the pair's margin, sspp.h:144 through MuJoCo's dist < margin; planner B: -1e-3, Collision.h:93) -
// where FP32 and FP64 may legitimately decide differently
constexpr double EPS_BAND = 1e-4;
double point_cost(const Model& m, const double* q, int* ncon_out, double* min_dist_out, int* band_out = nullptr) {
  static thread_local std::vector<V3> xpos; static thread_local std::vector<Quat> xquat; static thread_local std::vector<Geom> G;
  static thread_local std::vector<double> qp;      // scratch reused across calls (one set per OpenMP thread)
  make_qpos(m, q, qp);
  kinematics(m, qp, xpos, xquat);
  geom_frames(m, xpos, xquat, G);
  double cost = 0.0, mind = std::numeric_limits<double>::infinity();
  int ncon = 0;
  if (m.sdf) {   // sphere samples against the trilinear grid; cost rule of Collision.h:97-98 (use_center_dist = false)
    for (const auto& sp : m.spheres) {
      const V3 c = xpos[sp.body] + qrot(xquat[sp.body], sp.c);
      const double d = sdf_value(m, c) - sp.r;
      mind = std::min(mind, d);
      if (band_out && std::fabs(d - (m.mode == 0 ? m.sdf_margin : -1e-3)) <= EPS_BAND) ++*band_out;
      if (!(d < m.sdf_margin)) continue;
      ++ncon;
      if (m.mode == 0) cost += 1.0;
      else if (d < -1e-3) cost += -d;
    }
    if (ncon_out) *ncon_out = ncon;
    if (min_dist_out) *min_dist_out = mind;
    return cost;
  }
  for (const auto& pr : m.pairs) {
    bool ok;
    const double d = geom_dist(G[pr.first], G[pr.second], &ok);
    if (!ok) continue;
    mind = std::min(mind, d);
    const double margin = std::max(m.gmargin[pr.first], m.gmargin[pr.second]);
    if (band_out && std::fabs(d - (m.mode == 0 ? margin : std::min(margin, -1e-3))) <= EPS_BAND) ++*band_out;
    if (!(d < margin)) continue;   // MuJoCo keeps a contact iff dist < margin
    ++ncon;
    if (m.mode == 0) cost += 1.0;                                   // sspp.h:144  ncon > 0
    else if (d < -1e-3) cost += -1.0 / (norm(G[pr.second].p - G[pr.first].p) + 1e-4);   // Collision.h:92-96
  }
  if (ncon_out) *ncon_out = ncon;
  if (min_dist_out) *min_dist_out = mind;
  return cost;
}

// ---- spline algebra (Eigen semantics), written from the Cox-de Boor recursion ----
double bfun(double u, int p, int i, const std::vector<double>& t, bool last_span_closed) {
  if (p == 0) {
    if (t[i] <= u && u < t[i + 1]) return 1.0;
    if (last_span_closed && u == t[i + 1] && t[i] < t[i + 1] && t[i + 1] == t.back()) return 1.0;
    return 0.0;
  }
  double c1 = 0, c2 = 0;
  if (t[i + p] != t[i]) c1 = (u - t[i]) / (t[i + p] - t[i]) * bfun(u, p - 1, i, t, last_span_closed);
  if (t[i + p + 1] != t[i + 1]) c2 = (t[i + p + 1] - u) / (t[i + p + 1] - t[i + 1]) * bfun(u, p - 1, i + 1, t, last_span_closed);
  return c1 + c2;
}
std::vector<double> averaged_knots(int K, int p) {
  std::vector<double> u(K), t(K + p + 1, 0.0);
  for (int i = 0; i < K; ++i) u[i] = (double)i / (K - 1);
  for (int j = 1; j < K - p; ++j) { double s = 0; for (int i = j; i < j + p; ++i) s += u[i]; t[j + p] = s / p; }
  for (int i = 0; i <= p; ++i) t[K + p - i] = 1.0;
  return t;
}
std::vector<double> solve(std::vector<double> A, std::vector<double> B, int n, int nrhs) {   // Gaussian elimination
  for (int c = 0; c < n; ++c) {
    int piv = c;
    for (int r = c + 1; r < n; ++r) if (std::fabs(A[r * n + c]) > std::fabs(A[piv * n + c])) piv = r;
    for (int k = 0; k < n; ++k) std::swap(A[c * n + k], A[piv * n + k]);
    for (int k = 0; k < nrhs; ++k) std::swap(B[c * nrhs + k], B[piv * nrhs + k]);
    for (int r = c + 1; r < n; ++r) {
      const double f = A[r * n + c] / A[c * n + c];
      for (int k = c; k < n; ++k) A[r * n + k] -= f * A[c * n + k];
      for (int k = 0; k < nrhs; ++k) B[r * nrhs + k] -= f * B[c * nrhs + k];
    }
  }
  for (int c = n - 1; c >= 0; --c)
    for (int k = 0; k < nrhs; ++k) {
      double s = B[c * nrhs + k];
      for (int j = c + 1; j < n; ++j) s -= A[c * n + j] * B[j * nrhs + k];
      B[c * nrhs + k] = s / A[c * n + c];
    }
  return B;
}
// control points of the spline interpolating V (K x D) at u_i = i/(K-1)  (SplineFitting::Interpolate)
std::vector<double> interpolate(const std::vector<double>& V, int K, int D, int p, const std::vector<double>& t) {
  std::vector<double> A(K * K, 0.0);
  for (int i = 1; i < K - 1; ++i) for (int j = 0; j < K; ++j) A[i * K + j] = bfun((double)i / (K - 1), p, j, t, true);
  A[0] = 1.0; A[K * K - 1] = 1.0;
  return solve(A, V, K, D);
}
void spline_point(const std::vector<double>& ctrl, int K, int D, int p, const std::vector<double>& t, double u, double* out) {
  for (int d = 0; d < D; ++d) out[d] = 0.0;
  for (int j = 0; j < K; ++j) {
    const double b = bfun(u, p, j, t, true);
    if (b != 0.0) for (int d = 0; d < D; ++d) out[d] += b * ctrl[j * D + d];
  }
}

struct Cfg {   // mirrors the fields of SsppPlannerConfig the path consumes (doubles)
  int mode, D, K, p, M, n_free, free_begin;
  double w_collision, eval_zmin, floor_margin, floor_scale, limit_scale, sample_zmin, dist_zmin;
  double elite_frac, inc, dec, sigma_min, sigma_max, sigma_floor, var_beta, mean_lr;
  double lo[9], hi[9];
};

// ---- Philox4x32-10 (integer part must equal the product's stream; normals in double) ----
void philox(uint32_t c[4], uint32_t k0, uint32_t k1) {
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = 0xD2511F53ull * c[0], p1 = 0xCD9E8D57ull * c[2];
    const uint32_t n[4] = {(uint32_t)(p1 >> 32) ^ c[1] ^ k0, (uint32_t)p1, (uint32_t)(p0 >> 32) ^ c[3] ^ k1, (uint32_t)p0};
    std::memcpy(c, n, sizeof(n));
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}
double uni(uint32_t x) { return ((double)(x >> 9) + 0.5) / 8388608.0; }
double normal_at(uint64_t seed, uint32_t iter, uint32_t query, uint32_t cand, uint32_t stream, uint32_t j) {
  uint32_t c[4] = {cand, query, iter, (stream << 8) | ((j >> 2) & 0xFF)};
  philox(c, (uint32_t)seed, (uint32_t)(seed >> 32));
  const uint32_t a = (j & 2) ? c[2] : c[0], b = (j & 2) ? c[3] : c[1];
  const double r = std::sqrt(-2.0 * std::log(uni(a))), th = 6.283185307179586476925 * uni(b);
  return (j & 1) ? r * std::sin(th) : r * std::cos(th);
}

}  // namespace

extern "C" {

void* o1_create(const FlatView* v, int mode, int dof, int plan_body) {
  Model* m = new Model;
  m->nbody = v->nbody; m->ngeom = v->ngeom; m->nq = v->nq; m->mode = mode; m->dof = dof; m->plan_body = plan_body;
  auto ci = [](const int32_t* p, int n) { return std::vector<int>(p, p + n); };
  auto cd = [](const double* p, int n) { return std::vector<double>(p, p + n); };
  m->parent = ci(v->body_parent, v->nbody); m->weld = ci(v->body_weld, v->nbody); m->jtype = ci(v->body_jtype, v->nbody);
  m->qadr = ci(v->body_qadr, v->nbody); m->bpos = cd(v->body_pos, 3 * v->nbody); m->bquat = cd(v->body_quat, 4 * v->nbody);
  m->jaxis = cd(v->jnt_axis, 3 * v->nbody); m->jpos = cd(v->jnt_pos, 3 * v->nbody); m->qpos0 = cd(v->qpos0, v->nq);
  m->gbody = ci(v->geom_body, v->ngeom); m->gtype = ci(v->geom_type, v->ngeom); m->gct = ci(v->geom_contype, v->ngeom);
  m->gca = ci(v->geom_conaffinity, v->ngeom); m->gsize = cd(v->geom_size, 3 * v->ngeom); m->gpos = cd(v->geom_pos, 3 * v->ngeom);
  m->gquat = cd(v->geom_quat, 4 * v->ngeom); m->gmargin = cd(v->geom_margin, v->ngeom); m->excl = ci(v->exclude, 2 * v->nexclude);
  for (int g1 = 0; g1 < m->ngeom; ++g1) for (int g2 = g1 + 1; g2 < m->ngeom; ++g2) if (filter_pass(*m, g1, g2)) m->pairs.emplace_back(g1, g2);
  return m;
}
void o1_destroy(void* h) { delete static_cast<Model*>(h); }
int o1_num_pairs(void* h) { return (int)static_cast<Model*>(h)->pairs.size(); }
void o1_get_pairs(void* h, int32_t* out) { int k = 0; for (auto& p : static_cast<Model*>(h)->pairs) { out[k++] = p.first; out[k++] = p.second; } }

// SDF mode: bake the FP64 grid from the static partners of the moving geoms (at the model's initial
// pose) and take the sample spheres [n][5] = centre (frame of the k-th planned body), radius, k
void o1_set_sdf(void* h, const int32_t* dims, const double* lo, const double* hi, const float* spheres, int n_spheres, double margin) {
  Model& m = *static_cast<Model*>(h);
  m.nx = dims[0]; m.ny = dims[1]; m.nz = dims[2];
  const int d3[3] = {m.nx, m.ny, m.nz};
  for (int a = 0; a < 3; ++a) {   // the product defines the cell size in FP32; the node positions are inputs like the samples
    const float lo32 = (float)lo[a], h32 = (float)((hi[a] - lo[a]) / (d3[a] - 1));
    m.glo[a] = lo32; m.gh[a] = h32;
  }
  const std::vector<char> mv = moving_bodies(m);
  std::vector<char> obstacle(m.ngeom, 0);
  for (const auto& pr : m.pairs) {
    const bool m1 = mv[m.gbody[pr.first]], m2 = mv[m.gbody[pr.second]];
    if (m1 && !m2) obstacle[pr.second] = 1;
    if (m2 && !m1) obstacle[pr.first] = 1;
  }
  std::vector<V3> xpos; std::vector<Quat> xquat; std::vector<Geom> G;
  kinematics(m, m.qpos0, xpos, xquat);
  geom_frames(m, xpos, xquat, G);
  m.nodes.assign((size_t)m.nx * m.ny * m.nz, 0.0);
#pragma omp parallel for schedule(static)
  for (int k = 0; k < m.nz; ++k)
    for (int j = 0; j < m.ny; ++j)
      for (int i = 0; i < m.nx; ++i) {
        const V3 x{m.glo[0] + i * m.gh[0], m.glo[1] + j * m.gh[1], m.glo[2] + k * m.gh[2]};
        double best = std::numeric_limits<double>::infinity();
        for (int g = 0; g < m.ngeom; ++g) if (obstacle[g]) best = std::min(best, point_geom(x, G[g]));
        m.nodes[((size_t)k * m.ny + j) * m.nx + i] = best;
      }
  std::vector<int> planned;   // the k-th planned body in id order (the product's moving-body numbering)
  for (int b = 1; b < m.nbody; ++b)
    if ((m.mode == 0) ? ((m.jtype[b] == 2 || m.jtype[b] == 3) && m.qadr[b] >= 0 && m.qadr[b] < m.dof) : (b == m.plan_body)) planned.push_back(b);
  m.spheres.clear();
  for (int s = 0; s < n_spheres; ++s)
    m.spheres.push_back({V3{spheres[5 * s], spheres[5 * s + 1], spheres[5 * s + 2]}, spheres[5 * s + 3], planned[(int)spheres[5 * s + 4]]});
  m.sdf_margin = margin;
  m.sdf = true;
}
void o1_get_sdf(void* h, double* nodes) { const Model& m = *static_cast<Model*>(h); std::copy(m.nodes.begin(), m.nodes.end(), nodes); }

double o1_point_cost(void* h, const double* q, int32_t* ncon, double* min_dist) {
  int nc; double md;
  const double c = point_cost(*static_cast<Model*>(h), q, &nc, &md);
  if (ncon) *ncon = nc;
  if (min_dist) *min_dist = md;
  return c;
}

// world frames of every geom at q: xpos [ngeom][3], xmat [ngeom][9] row-major (columns = axes)
void o1_geom_frames(void* h, const double* q, double* xpos, double* xmat) {
  const Model& m = *static_cast<Model*>(h);
  std::vector<V3> bp; std::vector<Quat> bq; std::vector<Geom> G;
  kinematics(m, make_qpos(m, q), bp, bq);
  geom_frames(m, bp, bq, G);
  for (int g = 0; g < m.ngeom; ++g) {
    xpos[3 * g] = G[g].p.x; xpos[3 * g + 1] = G[g].p.y; xpos[3 * g + 2] = G[g].p.z;
    for (int j = 0; j < 3; ++j) { xmat[9 * g + j] = G[g].ax[j].x; xmat[9 * g + 3 + j] = G[g].ax[j].y; xmat[9 * g + 6 + j] = G[g].ax[j].z; }
  }
}
double o1_pair_dist(void* h, const double* q, int g1, int g2) {
  const Model& m = *static_cast<Model*>(h);
  std::vector<V3> bp; std::vector<Quat> bq; std::vector<Geom> G;
  kinematics(m, make_qpos(m, q), bp, bq);
  geom_frames(m, bp, bq, G);
  bool ok;
  return geom_dist(G[g1], G[g2], &ok);
}

// Eigen-style spline pieces for the tests
void o1_knots(int K, int p, double* out) { auto t = averaged_knots(K, p); std::copy(t.begin(), t.end(), out); }
void o1_interpolate(const double* V, int K, int D, int p, double* ctrl) {
  auto t = averaged_knots(K, p);
  auto c = interpolate(std::vector<double>(V, V + K * D), K, D, p, t);
  std::copy(c.begin(), c.end(), ctrl);
}
void o1_spline_eval(const double* ctrl, int K, int D, int p, double u, double* out) {
  auto t = averaged_knots(K, p);
  spline_point(std::vector<double>(ctrl, ctrl + K * D), K, D, p, t, u, out);
}

// spline evaluation on an explicit knot vector (pins the Cox-de Boor statement against the
// reference's Python twin sspp/BSplines.py:44-51 through tests/golden/spline_golden.npz)
void o1_spline_eval_knots(const double* ctrl, int K, int D, int p, const double* knots, double u, double* out) {
  std::vector<double> t(knots, knots + K + p + 1);
  spline_point(std::vector<double>(ctrl, ctrl + K * D), K, D, p, t, u, out);
}
void o1_basis_row(int K, int p, const double* knots, double u, double* out) {
  std::vector<double> t(knots, knots + K + p + 1);
  for (int j = 0; j < K; ++j) out[j] = bfun(u, p, j, t, true);
}

// Score candidates given their sampled rows (samples are an input).
//   base [K][D] point matrix (control points for JOINT, interpolation points for TASK)
//   params [n][n_free*D] -> L, Cnf, Cwf, cost, status per candidate; pts_out (optional) [n][M+1][D]
void o1_score(void* h, const Cfg* c, const double* base, const double* params, int n, double* L, double* Cnf, double* Cwf,
              double* cost, int32_t* status, double* min_dist_out, double* pts_out, int32_t* band_out) {
  const Model& m = *static_cast<Model*>(h);
  const int K = c->K, D = c->D, p = c->p, M = c->M;
  const std::vector<double> t = averaged_knots(K, p);
#pragma omp parallel for schedule(dynamic, 1)
  for (int i = 0; i < n; ++i) {
    std::vector<double> X(base, base + K * D);
    for (int e = 0; e < c->n_free * D; ++e) X[c->free_begin * D + e] = params[(size_t)i * c->n_free * D + e];
    std::vector<double> ctrl = (c->mode == 0) ? X : interpolate(X, K, D, p, t);   // tsp_path_model.h:32-43
    double pt[9], prev[9];
    double len = 0, cnf = 0, cwf = 0, md_all = std::numeric_limits<double>::infinity();
    int band = 0;
    if (c->mode == 0) {
      for (int k = 0; k <= M; ++k) {                          // sspp.h:134-148 (no early exit: same flag)
        spline_point(ctrl, K, D, p, t, (double)k / M, pt);
        if (pts_out) std::copy(pt, pt + D, pts_out + ((size_t)i * (M + 1) + k) * D);
        int ncon; double md;
        const double cc = point_cost(m, pt, &ncon, &md, &band);
        md_all = std::min(md_all, md);
        double pen = 0;
        if (c->limit_scale > 0) for (int d = 0; d < D; ++d) { const double ex = std::fabs(pt[d]) - c->hi[d]; if (ex > 0) pen += c->limit_scale * ex * ex; }
        cnf += cc; cwf += cc + pen;
      }
      for (int k = 1; k < M; ++k) {                           // sspp.h:157-166
        spline_point(ctrl, K, D, p, t, (double)(k - 1) / (M - 1), prev);
        spline_point(ctrl, K, D, p, t, (double)k / (M - 1), pt);
        double s2 = 0; for (int d = 0; d < D; ++d) s2 += (pt[d] - prev[d]) * (pt[d] - prev[d]);
        len += std::sqrt(s2);
      }
    } else {
      const double du = 1.0 / M;                              // tsp_evaluator.h:19-31
      spline_point(ctrl, K, D, p, t, 0.0, prev);
      if (pts_out) std::copy(prev, prev + D, pts_out + ((size_t)i * (M + 1)) * D);
      for (int k = 1; k <= M; ++k) {
        spline_point(ctrl, K, D, p, t, k * du, pt);
        if (pts_out) std::copy(pt, pt + D, pts_out + ((size_t)i * (M + 1) + k) * D);
        double s2 = 0; for (int d = 0; d < D; ++d) s2 += (pt[d] - prev[d]) * (pt[d] - prev[d]);
        len += std::sqrt(s2);
        int ncon; double md;
        const double cc = point_cost(m, pt, &ncon, &md, &band);
        md_all = std::min(md_all, md);
        const double deficit = (c->eval_zmin + c->floor_margin) - pt[2];
        const double fl = deficit > 0 ? c->floor_scale * deficit * deficit : 0.0;
        cnf += cc; cwf += cc + fl;
        std::copy(pt, pt + D, prev);
      }
    }
    L[i] = len; Cnf[i] = cnf; Cwf[i] = cwf; cost[i] = len + c->w_collision * cwf;
    status[i] = (cnf == 0.0) ? 0 : 1;                         // tsp_planner.h:110
    if (min_dist_out) min_dist_out[i] = md_all;
    if (band_out) band_out[i] = band;
  }
}

// top-k of the successes by (cost, index) and CES log-weights; returns k
int o1_elites(const double* cost, const int32_t* status, int n, double frac, int32_t* idx_out, double* w_out, int32_t* n_success) {
  std::vector<int> succ;
  for (int i = 0; i < n; ++i) if (status[i] == 0) succ.push_back(i);
  *n_success = (int)succ.size();
  if (succ.empty()) return 0;
  const int k = std::max(1, (int)(succ.size() * frac));
  std::sort(succ.begin(), succ.end(), [&](int a, int b) { return cost[a] < cost[b] || (cost[a] == cost[b] && a < b); });
  double sum = 0;
  for (int i = 0; i < k; ++i) { idx_out[i] = succ[i]; w_out[i] = std::log(k + 0.5) - std::log(i + 1.0); sum += w_out[i]; }
  for (int i = 0; i < k; ++i) w_out[i] /= sum;
  return k;
}

// Distribution::update followed by adapt(true) / adapt(false) when k == 0
void o1_update(const Cfg* c, double* mean, double* sigma, const double* elite_params, const double* w, int k) {
  const int D = c->D, nfd = c->n_free * D;
  auto clampS = [&](double s) { return std::max(std::min(std::max(s, c->sigma_min), c->sigma_max), c->sigma_floor); };
  if (k == 0) { for (int t = 0; t < nfd; ++t) sigma[t] = clampS(sigma[t] * c->inc); return; }
  for (int r = 0; r < c->n_free; ++r) {
    for (int d = 0; d < D; ++d) {
      const int t = r * D + d;
      double em = 0; for (int j = 0; j < k; ++j) em += w[j] * elite_params[(size_t)j * nfd + t];
      double nm = mean[t] + c->mean_lr * (em - mean[t]);
      if (c->mode == 1 && d == 2) nm = std::max(nm, c->dist_zmin);
      nm = std::min(std::max(nm, c->lo[d]), c->hi[d]);
      mean[t] = nm;
      double var = 0;
      for (int j = 0; j < k; ++j) {
        double df = elite_params[(size_t)j * nfd + t] - nm;
        if (c->mode == 1 && d == 3 && c->lo[3] != c->hi[3]) {
          const double range = c->hi[3] - c->lo[3];
          while (df > 0.5 * range) df -= range;
          while (df < -0.5 * range) df += range;
        }
        var += w[j] * df * df;
      }
      const double blend = (1.0 - c->var_beta) * sigma[t] * sigma[t] + c->var_beta * var;
      sigma[t] = clampS(std::sqrt(blend));
    }
  }
  for (int t = 0; t < nfd; ++t) sigma[t] = clampS(sigma[t] * c->dec);
}

// FP64 samplers on the same Philox counters as the product (for statistics tests and the CPU baseline)
void o1_philox(uint32_t ctr[4], uint32_t k0, uint32_t k1) { philox(ctr, k0, k1); }
double o1_normal(uint64_t seed, uint32_t iter, uint32_t query, uint32_t cand, uint32_t stream, uint32_t j) { return normal_at(seed, iter, query, cand, stream, j); }
void o1_sample(const Cfg* c, uint64_t seed, uint32_t iter, uint32_t query, int cand0, int n, const double* mean, const double* sigma, double* params) {
  const int D = c->D, nfd = c->n_free * D;
  for (int i = 0; i < n; ++i) {
    const uint32_t cand = (uint32_t)(cand0 + i);
    for (int r = 0; r < c->n_free; ++r)
      for (int d = 0; d < D; ++d) {
        const int t = r * D + d;
        double v;
        if (c->mode == 0) v = mean[t] + normal_at(seed, iter, query, cand, 0, (uint32_t)(d * c->n_free + r)) * sigma[t];
        else if (d < 3) {
          const uint32_t stream = (uint32_t)(r * 4 + d);
          int tries = 0; bool inside = false; v = 0;
          while (!inside && tries < 99) { v = mean[t] + normal_at(seed, iter, query, cand, stream, (uint32_t)tries) * sigma[t]; ++tries; inside = !(v < c->lo[d] || v > c->hi[d]); }
          if (!inside) { uint32_t cc[4] = {cand, query, iter, (stream << 8) | 25u}; philox(cc, (uint32_t)seed, (uint32_t)(seed >> 32)); v = c->lo[d] + uni(cc[0]) * (c->hi[d] - c->lo[d]); }
          if (d == 2) v = std::max(v, c->sample_zmin);
        } else if (c->lo[3] == c->hi[3]) v = mean[t];
        else {
          v = mean[t] + normal_at(seed, iter, query, cand, (uint32_t)(r * 4 + 3), 0) * sigma[t];
          const double range = c->hi[3] - c->lo[3];
          while (v < c->lo[3]) v += range;
          while (v > c->hi[3]) v -= range;
        }
        params[(size_t)i * nfd + t] = v;
      }
  }
}

}  // extern "C"
