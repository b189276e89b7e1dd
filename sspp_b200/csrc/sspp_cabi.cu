// sspp_cabi.cu — implementation of include/sspp_cuda.h: handle bookkeeping, device buffers and
// launches.  No arithmetic of the scoring path lives here (that is sspp_kernels.cu) and nothing
// here can run the path on the CPU: every compute entry point ends in a kernel launch.
#include "../../include/sspp_cuda.h"

#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "sspp_kernels.h"
#include "sspp_model.h"
#include "sspp_spline.h"

using namespace sspp;
using ssppm::DevModel;
using ssppm::ScoreConfig;

struct ssppcu_model {
  FlatModel m;
};

struct ssppcu_planner {
  SsppPlannerConfig cfg;
  std::string body_name;
  DevModel dev;            // host copy of the scene (tables + scalars)
  ssppm::ModelView view;   // the same scalars with the tables in device global memory
  size_t smem_model = 0;   // bytes of the tables staged at the front of the kernel's shared memory
  IterParams P;
  int threads = 256;
  int variant = 0;   // index into kIterVariants
  size_t smem = 0;
  int n_sm = 148, grid = 1;
  cudaStream_t own_stream = nullptr, stream = nullptr;
  std::string err;
  SplineSetup spl;
  std::vector<float> psi_col, psi_arc, base_host;
  std::vector<void*> allocs;
  float* d_eval_pts = nullptr; float* d_eval_out = nullptr; int eval_cap = 0;
  std::vector<int> mgeom_src, sgeom_src;   // FlatModel geom ids behind dev.mg / dev.sg (ssppcu_check_points)
  int* d_geom_src = nullptr;               // the two maps on the device, uploaded by the first ssppcu_check_points
  unsigned long long* d_merge_keys = nullptr; int merge_cap = 0, merged_n = 0;
  float* d_live = nullptr;       // mean | sigma | best | best_valid | any_success (ints), one block
  float* d_state = nullptr;      // snapshot of d_live
  size_t state_words = 0; bool state_saved = false;
  unsigned long long* d_probe = nullptr;   // phase probe stamps [grid][6], allocated on demand
  // host transfers: every per-candidate / per-query result lives in ONE device block mirrored by a pinned host
  // block (one D2H copy per read-back), and host inputs go through a pinned staging ring (asynchronous H2D)
  unsigned char* d_res = nullptr; unsigned char* h_res = nullptr; size_t res_bytes = 0;
  size_t o_L = 0, o_Cnf = 0, o_Cwf = 0, o_cost = 0, o_status = 0, o_nsucc = 0, o_nelite = 0, o_bidx = 0, o_bcost = 0, o_extra = 0;
  unsigned char* h_stage = nullptr; size_t stage_bytes = 0, stage_at = 0;
  bool best_fresh = false;   // the pinned mirror holds best | best_valid of the current device state (set by get_results, cleared by any other call)
  // the pinned mirror is mapped: the cooperative kernel copies the result block into it itself (mirror_results, sspp_kernels.cu),
  // so a read-back right after such a launch is a stream synchronisation and no copy.  SSPPCU_RESULT_MIRROR=0 turns it off.
  unsigned char* h_res_dev = nullptr; bool mirror_on = true;
  bool res_fresh = false;    // the last call was a launch that mirrors its results (cleared by any other call)
  // success rows (ssppcu_set_success_rows): mapped pinned block rows [Q][N][nfd] | idx [Q][N] | count [Q], filled by a
  // compaction kernel enqueued behind every iteration launch
  bool rows_on = false;
  unsigned long long launch_epoch = 0, rows_epoch = ~0ull;   // the rows in the pinned block belong to launch rows_epoch
  unsigned char* h_rows = nullptr; unsigned char* h_rows_dev = nullptr; size_t o_ridx = 0, o_rcount = 0;
  // fused key exchange of a population shard (ssppcu_exchange_*): this rank's block, the peers' mappings
  struct Exchange {
    int world = 0, rank = 0;
    unsigned epoch = 0;
    bool attached = false;
    unsigned char* block = nullptr; size_t block_bytes = 0;                  // 16-byte lines [2 halves][world][shard_len]
    unsigned long long* d_gathered = nullptr;
    void* ipc_mapped[kMaxExchangeRanks] = {};                                // cudaIpcOpenMemHandle results (to close)
    ExchangeParams* d_params = nullptr;
    int* d_ticket = nullptr;
    int* h_status = nullptr; int* d_status = nullptr;                        // mapped pinned word
    float* d_vals = nullptr;
  } xc;
};

namespace {

void set_err(char* err, size_t n, const std::string& s) {
  if (err && n) { std::snprintf(err, n, "%s", s.c_str()); }
}
int fail(ssppcu_t* h, int code, const std::string& s) { if (h) h->err = s; return code; }
int cuda_fail(ssppcu_t* h, cudaError_t e, const char* what) {
  return fail(h, SSPPCU_ECUDA, std::string(what) + ": " + cudaGetErrorString(e));
}
#define CU(h, call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) return cuda_fail((h), e__, #call); } while (0)

template <class T>
int dalloc(ssppcu_t* h, T** p, size_t n) {
  void* q = nullptr;
  cudaError_t e = cudaMalloc(&q, std::max<size_t>(n, 1) * sizeof(T));
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMalloc");
  e = cudaMemset(q, 0, std::max<size_t>(n, 1) * sizeof(T));
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMemset");
  h->allocs.push_back(q);
  *p = static_cast<T*>(q);
  return 0;
}

int next_pow2(int n) { int p = 1; while (p < n) p <<= 1; return p; }

}  // namespace

extern "C" {

int ssppcu_version(void) { return 100; }

int ssppcu_spline_setup(int n_points, int degree, double* knots, double* Ainv) {
  if (degree < 1 || n_points < degree + 1) return SSPPCU_EINVAL;
  try {
    SplineSetup s = make_setup(n_points, degree);
    if (knots) std::copy(s.knots.begin(), s.knots.end(), knots);
    if (Ainv) std::copy(s.Ainv.begin(), s.Ainv.end(), Ainv);
    return SSPPCU_OK;
  } catch (const std::exception&) { return SSPPCU_EINVAL; }
}
int ssppcu_collocation_matrix(int n_points, int degree, const double* grid, int n_grid, int via_form, double* psi) {
  if (degree < 1 || n_points < degree + 1 || !grid || !psi || n_grid < 0) return SSPPCU_EINVAL;
  try {
    SplineSetup s = make_setup(n_points, degree);
    std::vector<double> P = collocation_matrix(s, std::vector<double>(grid, grid + n_grid), via_form != 0);
    std::copy(P.begin(), P.end(), psi);
    return SSPPCU_OK;
  } catch (const std::exception&) { return SSPPCU_EINVAL; }
}

// ---- sspp/BSplines.py + sspp/CubicPath.py as product functions ----
int ssppcu_bspline_knot_vector(int n_ctrl, int k, double* knots) {
  if (!knots) return SSPPCU_EINVAL;
  try { const std::vector<double> t = uniform_knots(n_ctrl, k); std::copy(t.begin(), t.end(), knots); return SSPPCU_OK; }
  catch (const std::exception&) { return SSPPCU_EINVAL; }
}
int ssppcu_bspline_rows(int n_ctrl, int k, const double* knots, const double* theta, int n_theta, int kind, double* out) {
  if (!knots || !theta || !out || n_theta < 0 || k < 0 || n_ctrl < 1 || kind < 0 || kind > 3) return SSPPCU_EINVAL;
  try {
    const std::vector<double> t(knots, knots + n_ctrl + k + 1);
    for (int m = 0; m < n_theta; ++m) {
      const std::vector<double> r = (kind < 2) ? half_open_basis_row(theta[m], k, t, kind == 1) : curve_row(theta[m], k, t, kind == 3);
      std::copy(r.begin(), r.end(), out + (size_t)m * n_ctrl);
    }
    return SSPPCU_OK;
  } catch (const std::exception&) { return SSPPCU_EINVAL; }
}
int ssppcu_bspline_fit(const double* via, int n, int d, int k, double* ctrl, double* knots) {
  if (!via || !ctrl || n < 2 || d < 1) return SSPPCU_EINVAL;
  try {
    std::vector<double> t;
    const std::vector<double> c = fit_uniform(std::vector<double>(via, via + (size_t)n * d), n, d, k, &t);
    std::copy(c.begin(), c.end(), ctrl);
    if (knots) std::copy(t.begin(), t.end(), knots);
    return SSPPCU_OK;
  } catch (const std::exception&) { return SSPPCU_EINVAL; }
}
int ssppcu_cubic_path(const double* start, const double* via, const double* end, int d, const double* u, int n_u, double* pos, double* vel,
                      double* acc) {
  if (!start || !via || !end || !u || d < 1 || n_u < 0) return SSPPCU_EINVAL;
  for (int e = 0; e < d; ++e) {
    const double a = 2.0 * (end[e] + 3.0 * start[e] - 4.0 * via[e]);       // CubicPath.py:23-26
    const double b = 4.0 * (via[e] - start[e] - a / 8.0);
    const double c = 0.0, dd = start[e];
    for (int m = 0; m < n_u; ++m) {
      const double x = std::min(std::max(u[m], 0.0), 1.0);                 // np.clip(u, 0, 1), CubicPath.py:39
      const double x3 = std::pow(x, 3.0), x2 = std::pow(x, 2.0);           // numpy's u**3, u**2; same operation order as :41, :56-58
      if (pos) pos[(size_t)m * d + e] = ((a * x3 + b * x2) + c * x) + dd;
      if (vel) vel[(size_t)m * d + e] = ((3.0 * a) * x2 + (2.0 * b) * x) + c;
      if (acc) acc[(size_t)m * d + e] = (6.0 * a) * x + 2.0 * b;
    }
  }
  return SSPPCU_OK;
}
int ssppcu_bspline_eval_batch(int device, const float* rows, int n_theta, int n_ctrl, const float* ctrl, int n_curves, int d, float* out,
                              char* err, size_t errlen) {
  if (!rows || !ctrl || !out || n_theta < 1 || n_ctrl < 1 || n_curves < 1 || d < 1) { set_err(err, errlen, "bad arguments"); return SSPPCU_EINVAL; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) { set_err(err, errlen, "no CUDA device (this library has no CPU path)"); return SSPPCU_ECUDA; }
  if (device < 0 || device >= ndev) { set_err(err, errlen, "device ordinal out of range"); return SSPPCU_EINVAL; }
  cudaDeviceProp prop;
  float *d_rows = nullptr, *d_ctrl = nullptr, *d_out = nullptr;
  const size_t nr = (size_t)n_theta * n_ctrl, nc = (size_t)n_curves * n_ctrl * d, no = (size_t)n_curves * n_theta * d;
  e = cudaSetDevice(device);
  if (e == cudaSuccess) e = cudaGetDeviceProperties(&prop, device);
  if (e == cudaSuccess) e = cudaMalloc(&d_rows, (nr + nc + no) * sizeof(float));
  if (e == cudaSuccess) { d_ctrl = d_rows + nr; d_out = d_ctrl + nc; e = cudaMemcpy(d_rows, rows, nr * sizeof(float), cudaMemcpyHostToDevice); }
  if (e == cudaSuccess) e = cudaMemcpy(d_ctrl, ctrl, nc * sizeof(float), cudaMemcpyHostToDevice);
  if (e == cudaSuccess) e = launch_bspline_eval(d_rows, n_theta, n_ctrl, d_ctrl, n_curves, d, d_out, prop.multiProcessorCount, 0);
  if (e == cudaSuccess) e = cudaMemcpy(out, d_out, no * sizeof(float), cudaMemcpyDeviceToHost);
  if (d_rows) cudaFree(d_rows);
  if (e != cudaSuccess) { set_err(err, errlen, std::string("ssppcu_bspline_eval_batch: ") + cudaGetErrorString(e)); return SSPPCU_ECUDA; }
  return SSPPCU_OK;
}

// No counterpart in the reference.  ssppcu_bspline_eval_batch without any
// allocation: the device staging is the file-scope 4 MiB array of sspp_kernels.cu (sspp::bspline_eval_static).  No
// cudaMalloc / cudaFree / allocating library call on this path; sizes beyond the array are refused before any CUDA call.
static_assert(SSPPCU_NOALLOC_MAX_FLOATS == sspp::kStaticScratchFloats, "include/sspp_cuda.h and sspp_kernels.h disagree on the static scratch size");
int ssppcu_bspline_eval_batch_noalloc(int device, const float* rows, int n_theta, int n_ctrl, const float* ctrl, int n_curves, int d, float* out,
                                      char* err, size_t errlen) {
  if (!rows || !ctrl || !out || n_theta < 1 || n_ctrl < 1 || n_curves < 1 || d < 1) { set_err(err, errlen, "bad arguments"); return SSPPCU_EINVAL; }
  const size_t need = sspp::static_scratch_floats_needed(n_theta, n_ctrl, n_curves, d);
  if (need > (size_t)SSPPCU_NOALLOC_MAX_FLOATS) {
    set_err(err, errlen, "ssppcu_bspline_eval_batch_noalloc: rows + control points + output need " + std::to_string(need) +
                             " floats (each region rounded up to 64), the static device scratch holds " +
                             std::to_string(SSPPCU_NOALLOC_MAX_FLOATS) + " (4 MiB); use ssppcu_bspline_eval_batch or split the batch");
    return SSPPCU_ELIMIT;
  }
  if ((size_t)n_theta * n_ctrl * sizeof(float) > 200 * 1024) {
    set_err(err, errlen, "ssppcu_bspline_eval_batch_noalloc: the row table n_theta x n_ctrl exceeds 200 KiB of shared memory");
    return SSPPCU_ELIMIT;
  }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) { set_err(err, errlen, "no CUDA device (this library has no CPU path)"); return SSPPCU_ECUDA; }
  if (device < 0 || device >= ndev) { set_err(err, errlen, "device ordinal out of range"); return SSPPCU_EINVAL; }
  e = sspp::bspline_eval_static(device, rows, n_theta, n_ctrl, ctrl, n_curves, d, out);
  if (e != cudaSuccess) { set_err(err, errlen, std::string("ssppcu_bspline_eval_batch_noalloc: ") + cudaGetErrorString(e)); return SSPPCU_ECUDA; }
  return SSPPCU_OK;
}

int ssppcu_model_load(const char* path, ssppcu_model_t** out, char* err, size_t errlen) {
  if (!path || !out) { set_err(err, errlen, "null argument"); return SSPPCU_EINVAL; }
  try { auto* m = new ssppcu_model; m->m = load_mjcf(path); *out = m; return SSPPCU_OK; }
  catch (const std::exception& e) { set_err(err, errlen, e.what()); return SSPPCU_EMODEL; }
}
int ssppcu_model_parse(const char* text, ssppcu_model_t** out, char* err, size_t errlen) {
  if (!text || !out) { set_err(err, errlen, "null argument"); return SSPPCU_EINVAL; }
  try { auto* m = new ssppcu_model; m->m = parse_mjcf(text); *out = m; return SSPPCU_OK; }
  catch (const std::exception& e) { set_err(err, errlen, std::string("Failed to load MuJoCo model from XML: ") + e.what()); return SSPPCU_EMODEL; }
}
void ssppcu_model_free(ssppcu_model_t* m) { delete m; }

int ssppcu_model_view(const ssppcu_model_t* mm, SsppFlatModelView* v) {
  if (!mm || !v) return SSPPCU_EINVAL;
  const FlatModel& m = mm->m;
  v->nbody = m.nbody; v->ngeom = m.ngeom; v->nq = m.nq; v->nexclude = (int)m.exclude.size() / 2;
  v->body_parent = m.body_parent.data(); v->body_weld = m.body_weld.data(); v->body_jtype = m.body_jtype.data();
  v->body_qadr = m.body_qadr.data(); v->body_pos = m.body_pos.data(); v->body_quat = m.body_quat.data();
  v->jnt_axis = m.jnt_axis.data(); v->jnt_pos = m.jnt_pos.data(); v->jnt_range = m.jnt_range.data();
  v->qpos0 = m.qpos0.data();
  v->geom_body = m.geom_body.data(); v->geom_type = m.geom_type.data(); v->geom_contype = m.geom_contype.data();
  v->geom_conaffinity = m.geom_conaffinity.data(); v->geom_size = m.geom_size.data(); v->geom_pos = m.geom_pos.data();
  v->geom_quat = m.geom_quat.data(); v->geom_margin = m.geom_margin.data(); v->exclude = m.exclude.data();
  return SSPPCU_OK;
}
int ssppcu_model_body_id(const ssppcu_model_t* m, const char* name) { return (m && name) ? m->m.body_id(name) : -1; }
const char* ssppcu_model_body_name(const ssppcu_model_t* m, int b) { return (m && b >= 0 && b < m->m.nbody) ? m->m.body_name[b].c_str() : nullptr; }
const char* ssppcu_model_geom_name(const ssppcu_model_t* m, int g) { return (m && g >= 0 && g < m->m.ngeom) ? m->m.geom_name[g].c_str() : nullptr; }

int ssppcu_model_body_point(const ssppcu_model_t* mm, const char* body, double out4[4]) {
  if (!mm || !body || !out4) return SSPPCU_EINVAL;
  const FlatModel& m = mm->m;
  const int b = m.body_id(body);
  if (b < 0 || m.body_jtype[b] != MJ_FREE) return SSPPCU_EMODEL;
  const double* q = &m.qpos0[m.body_qadr[b]];
  double R[9];
  quat_to_mat(q + 3, R);
  out4[0] = q[0]; out4[1] = q[1]; out4[2] = q[2];
  out4[3] = std::atan2(R[3], R[0]);   // ZYX yaw
  return SSPPCU_OK;
}

int ssppcu_model_pair_counts(const ssppcu_model_t* m, int mode, int dof, const char* body, int32_t* nms, int32_t* nmm,
                             int32_t* nss, char* err, size_t errlen) {
  if (!m) return SSPPCU_EINVAL;
  try {
    CompileInfo info;
    DevModel D = compile_model(m->m, mode == SSPPCU_MODE_JOINT ? PLAN_JOINT : PLAN_TASK, dof, body ? body : "", &info);
    if (nms) *nms = D.n_pair;
    if (nmm) *nmm = D.n_mmpair;
    if (nss) *nss = info.n_static_pairs;
    return SSPPCU_OK;
  } catch (const std::exception& e) { set_err(err, errlen, e.what()); return SSPPCU_EMODEL; }
}

void ssppcu_default_config(int mode, SsppPlannerConfig* c) {
  if (!c) return;
  std::memset(c, 0, sizeof(*c));
  c->mode = mode;
  c->n_queries = 1;
  c->elite_fraction = 0.3;
  c->collision_mode = SSPPCU_COLLIDE_PRIMITIVES; c->use_center_dist = 1; c->sdf_sphere_ratio = 0.5;
  c->sdf_dims[0] = c->sdf_dims[1] = c->sdf_dims[2] = 128;
  if (mode == SSPPCU_MODE_JOINT) {        // sspp.h:195-196 defaults; classic planner A does no CES update
    c->dof = 7; c->degree = 3; c->n_points = 10; c->check_points = 50; c->n_samples = 50;
    c->w_collision = 0.0; c->inc = 1.5; c->dec = 0.95; c->var_beta = 0.2; c->mean_lr = 0.5;
    c->stddev_min = 1e-3; c->stddev_max = 10.0;
    for (int d = 0; d < SSPPCU_MAX_DOF; ++d) { c->lo[d] = -3.14159265358979323846; c->hi[d] = 3.14159265358979323846; }
  } else {                                 // PlannerConfig defaults, tsp_planner.h:14-29
    c->dof = 4; c->degree = 2; c->n_points = 3; c->check_points = 40; c->n_samples = 50;
    c->include_mean = 1; c->include_best = 1; c->do_update = 1;
    c->w_collision = 1.0; c->inc = 1.5; c->dec = 0.95; c->sigma_floor = 0.0; c->var_beta = 0.2; c->mean_lr = 0.5;
    c->stddev_min = 1e-3; c->stddev_max = 0.5;
    c->eval_zmin = 0.0; c->floor_margin = 0.01; c->floor_scale = 10.0;
    for (int d = 0; d < SSPPCU_MAX_DOF; ++d) { c->lo[d] = -2.0; c->hi[d] = 2.0; }
  }
}

void ssppcu_destroy(ssppcu_t* h) {
  if (!h) return;
  cudaSetDevice(h->cfg.device);
  for (void* p : h->allocs) cudaFree(p);
  if (h->d_eval_pts) cudaFree(h->d_eval_pts);
  if (h->d_eval_out) cudaFree(h->d_eval_out);
  if (h->d_merge_keys) cudaFree(h->d_merge_keys);
  for (void* m : h->xc.ipc_mapped) if (m) cudaIpcCloseMemHandle(m);
  if (h->xc.block) cudaFree(h->xc.block);
  if (h->xc.d_params) cudaFree(h->xc.d_params);
  if (h->xc.d_ticket) cudaFree(h->xc.d_ticket);
  if (h->xc.d_vals) cudaFree(h->xc.d_vals);
  if (h->xc.d_gathered) cudaFree(h->xc.d_gathered);
  if (h->xc.h_status) cudaFreeHost(h->xc.h_status);
  if (h->d_probe) cudaFree(h->d_probe);
  if (h->d_res) cudaFree(h->d_res);
  if (h->h_res) cudaFreeHost(h->h_res);
  if (h->h_stage) cudaFreeHost(h->h_stage);
  if (h->h_rows) cudaFreeHost(h->h_rows);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  delete h;
}

const char* ssppcu_last_error(const ssppcu_t* h) { return h ? h->err.c_str() : "null handle"; }

int ssppcu_create(const ssppcu_model_t* mm, const SsppPlannerConfig* cfg, ssppcu_t** out, char* err, size_t errlen) {
  if (!mm || !cfg || !out) { set_err(err, errlen, "null argument"); return SSPPCU_EINVAL; }
  const SsppPlannerConfig& c = *cfg;
  const bool task = c.mode == SSPPCU_MODE_TASK;
  if (c.mode != SSPPCU_MODE_JOINT && !task) { set_err(err, errlen, "unknown mode"); return SSPPCU_EINVAL; }
  if (task && c.dof != 4) { set_err(err, errlen, "TASK mode plans (x,y,z,yaw): dof must be 4"); return SSPPCU_EINVAL; }
  if (c.dof < 1 || c.dof > SSPPCU_MAX_DOF) { set_err(err, errlen, "dof out of range"); return SSPPCU_EINVAL; }
  if (c.degree < 1 || c.degree > 5) { set_err(err, errlen, "degree out of range"); return SSPPCU_EINVAL; }
  if (c.n_points < c.degree + 1 || c.n_points > ssppm::MAX_K) { set_err(err, errlen, "n_points must be in [degree+1, 16]"); return SSPPCU_EINVAL; }
  if (c.check_points < 2 || c.check_points > 1024) { set_err(err, errlen, "check_points must be in [2, 1024]"); return SSPPCU_EINVAL; }
  if (c.n_samples < 1 || c.n_queries < 1) { set_err(err, errlen, "n_samples and n_queries must be positive"); return SSPPCU_EINVAL; }
  const int n_free = task ? c.n_points - 2 : c.n_points - 2 * c.degree;
  const int free_begin = task ? 1 : c.degree;
  if (n_free < 0) { set_err(err, errlen, "init_points leaves no room for the fixed boundary control points"); return SSPPCU_EINVAL; }
  int ndev = 0;
  cudaError_t ce = cudaGetDeviceCount(&ndev);
  if (ce != cudaSuccess || ndev == 0) {
    set_err(err, errlen, std::string("no CUDA device: ") + (ce != cudaSuccess ? cudaGetErrorString(ce) : "device count is 0") +
                             " (this library has no CPU path)");
    return SSPPCU_ECUDA;
  }
  if (c.device < 0 || c.device >= ndev) { set_err(err, errlen, "device ordinal out of range"); return SSPPCU_EINVAL; }
  auto* h = new ssppcu_planner;
  h->cfg = c;
  h->body_name = c.body_name ? c.body_name : "";
  h->cfg.body_name = nullptr;
  try {
    CompileOptions opt;
    opt.sdf = c.collision_mode == SSPPCU_COLLIDE_SDF;
    opt.sphere_ratio = c.sdf_sphere_ratio > 0 ? c.sdf_sphere_ratio : 0.5;
    opt.use_center_dist = c.use_center_dist != 0;
    CompileInfo info;
    h->dev = compile_model(mm->m, task ? PLAN_TASK : PLAN_JOINT, c.dof, h->body_name, &info, opt);
    h->mgeom_src = info.mgeom_src; h->sgeom_src = info.sgeom_src;
    h->spl = make_setup(c.n_points, c.degree);
  } catch (const std::exception& e) { set_err(err, errlen, e.what()); delete h; return SSPPCU_EMODEL; }
  if (c.collision_mode != SSPPCU_COLLIDE_PRIMITIVES && c.collision_mode != SSPPCU_COLLIDE_SDF) { set_err(err, errlen, "unknown collision_mode"); delete h; return SSPPCU_EINVAL; }
  if (h->dev.sdf_mode) {
    ssppm::SdfGrid& g = h->dev.grid;
    g.nx = c.sdf_dims[0]; g.ny = c.sdf_dims[1]; g.nz = c.sdf_dims[2];
    if (g.nx < 2 || g.ny < 2 || g.nz < 2 || (double)g.nx * g.ny * g.nz > 512.0 * 512 * 512) { set_err(err, errlen, "sdf_dims must be in [2, 512^3 nodes]"); delete h; return SSPPCU_EINVAL; }
    const int dims[3] = {g.nx, g.ny, g.nz};
    for (int a = 0; a < 3; ++a) {
      if (!(c.sdf_hi[a] > c.sdf_lo[a])) { set_err(err, errlen, "sdf_hi must exceed sdf_lo on every axis"); delete h; return SSPPCU_EINVAL; }
      g.lo[a] = (float)c.sdf_lo[a];
      g.h[a] = (float)((c.sdf_hi[a] - c.sdf_lo[a]) / (dims[a] - 1));
      g.inv_h[a] = 1.0f / g.h[a];
    }
  }
  // ---- collocation grids (sspp.h:134-137,157-160; tsp_evaluator.h:19-24) ----
  const int M = c.check_points;
  std::vector<double> gcol, garc;
  if (task) { for (int i = 0; i <= M; ++i) gcol.push_back((double)i * (1.0 / M)); }
  else {
    for (int i = 0; i <= M; ++i) gcol.push_back((double)i / M);
    for (int i = 0; i < M; ++i) garc.push_back((double)i / (M - 1));
  }
  {
    std::vector<double> pc = collocation_matrix(h->spl, gcol, task);
    h->psi_col.assign(pc.begin(), pc.end());
    if (!task) { std::vector<double> pa = collocation_matrix(h->spl, garc, false); h->psi_arc.assign(pa.begin(), pa.end()); }
  }
  IterParams& P = h->P;
  std::memset(&P, 0, sizeof(P));
  ScoreConfig& s = P.cfg;
  s.mode = task ? ssppm::MODE_TASK : ssppm::MODE_JOINT;
  s.D = c.dof; s.K = c.n_points; s.n_free = n_free; s.free_begin = free_begin;
  s.Mc = M + 1; s.col_first = task ? 1 : 0; s.Ma = task ? M + 1 : M; s.arc_same = task ? 1 : 0;
  s.include_mean = c.include_mean ? 1 : 0; s.include_best = c.include_best ? 1 : 0; s.do_update = c.do_update ? 1 : 0;
  const int n_local = c.n_samples + ((c.shard_global > 0) ? 0 : s.include_mean + s.include_best);
  s.n_cand = n_local;
  s.w_collision = (float)c.w_collision;
  s.eval_zmin = (float)c.eval_zmin; s.floor_margin = (float)c.floor_margin; s.floor_scale = (float)c.floor_scale;
  s.limit_scale = (float)c.limit_scale; s.sample_zmin = (float)c.sample_zmin; s.dist_zmin = (float)c.dist_zmin;
  s.inc = (float)c.inc; s.dec = (float)c.dec; s.sigma_min = (float)c.stddev_min; s.sigma_max = (float)c.stddev_max;
  s.sigma_floor = (float)c.sigma_floor; s.var_beta = (float)c.var_beta; s.mean_lr = (float)c.mean_lr;
  for (int d = 0; d < SSPPCU_MAX_DOF; ++d) { s.lo[d] = (float)c.lo[d]; s.hi[d] = (float)c.hi[d]; }
  s.elite_frac = c.elite_fraction;
  P.n_query = c.n_queries;
  P.cand_offset = (c.shard_global > 0) ? c.shard_offset : 0;
  P.sort_only = (c.shard_global > 0) ? 1 : 0;
  if (c.shard_global > 0 && (c.shard_offset < 0 || c.shard_offset + n_local > c.shard_global)) {
    set_err(err, errlen, "shard range exceeds the global population"); delete h; return SSPPCU_EINVAL;
  }
  if (c.shard_global > 0 && c.n_queries != 1) { set_err(err, errlen, "population sharding holds one query per handle"); delete h; return SSPPCU_EINVAL; }
  P.mq = std::max(s.Mc, s.Ma);
  P.n_pad = next_pow2(std::max(2, s.n_cand));
  // ---- banded collocation rows ----
  std::vector<int> first_col(s.Mc), first_arc(std::max(1, s.Ma));
  P.w_col = ssppm::compute_band(h->psi_col.data(), s.Mc, s.K, first_col.data());
  std::vector<float> band_col((size_t)s.Mc * P.w_col), band_arc;
  ssppm::pack_band(h->psi_col.data(), s.Mc, s.K, P.w_col, first_col.data(), band_col.data());
  P.w_arc = 1;
  if (!task) {
    P.w_arc = ssppm::compute_band(h->psi_arc.data(), s.Ma, s.K, first_arc.data());
    band_arc.resize((size_t)s.Ma * P.w_arc);
    ssppm::pack_band(h->psi_arc.data(), s.Ma, s.K, P.w_arc, first_arc.data(), band_arc.data());
  }
  if ((ce = cudaSetDevice(c.device)) != cudaSuccess) { set_err(err, errlen, std::string("cudaSetDevice: ") + cudaGetErrorString(ce)); delete h; return SSPPCU_ECUDA; }
  cudaDeviceProp prop;
  if ((ce = cudaGetDeviceProperties(&prop, c.device)) != cudaSuccess) { set_err(err, errlen, std::string("cudaGetDeviceProperties: ") + cudaGetErrorString(ce)); delete h; return SSPPCU_ECUDA; }
  h->n_sm = prop.multiProcessorCount;
  h->smem_model = model_smem_bytes(h->dev);
  P.model_bytes = (int)h->smem_model;
  if (h->smem_model + 16 * 1024 > std::min<size_t>(prop.sharedMemPerBlockOptin, 200 * 1024)) { set_err(err, errlen, "scene tables exceed the shared memory of a CTA"); delete h; return SSPPCU_ELIMIT; }
  const size_t smem_cap = std::min<size_t>(prop.sharedMemPerBlockOptin, 200 * 1024) - h->smem_model;   // h->smem excludes the tables
  // ---- tile shape ----
  // many small queries: one CTA per query, whole plans inside one launch (coop = 0);
  // otherwise a persistent cooperative grid over candidate tiles (coop = 1).
  // cfg.tile_hint (0 = automatic) forces the kernel variant / tile: bits 0-7 variant + 1, 8-19 candidates per
  // CTA, 20-31 threads per CTA.
  const int hint_variant = (c.tile_hint & 0xFF) - 1, hint_cpc = (c.tile_hint >> 8) & 0xFFF, hint_threads = (c.tile_hint >> 20) & 0xFFF;
  if (hint_variant >= kNumIterVariants || (hint_threads && (hint_threads % 32 || hint_threads > 1024))) {
    set_err(err, errlen, "tile_hint: unknown kernel variant or thread count"); delete h; return SSPPCU_EINVAL;
  }
  h->variant = hint_variant >= 0 ? hint_variant : 0;
  const bool auto_shape = hint_variant < 0 && !hint_threads;
  h->threads = ((long)s.n_cand * s.Mc <= 512) ? 128 : 256;
  if (auto_shape && h->threads == 256) h->variant = 1;   // one CTA per query: 3 x 256 threads per SM at <= 80 registers (measured, profiles/)
  if (hint_threads) h->threads = hint_threads;
  if (h->threads > kIterVariants[h->variant].max_threads) { set_err(err, errlen, "tile_hint: threads exceed the variant's launch bound"); delete h; return SSPPCU_EINVAL; }
  P.coop = 0; P.cand_per_cta = s.n_cand; P.ctas_per_query = 1;
  h->smem = iterate_smem_bytes(h->dev, P, h->threads);
  const bool one_cta_ok = h->smem <= std::min<size_t>(smem_cap, 64 * 1024) && (long)s.n_cand * s.Mc <= 4096 && c.n_queries >= h->n_sm;
  if (!one_cta_ok) {
    // one wave of tiles over the co-resident CTAs: guess the residency, shape the tiles, ask the
    // occupancy calculator, and reshape once if it answers differently
    P.coop = 1;
    if (auto_shape) { h->variant = 2; h->threads = 320; }   // 2 x 320 threads per SM at <= 96 registers: the measured optimum (profiles/)
    else if (!hint_threads) h->threads = std::min(256, kIterVariants[h->variant].max_threads);
    int per_sm = kIterVariants[h->variant].min_ctas;
    for (int round = 0; round < 3; ++round) {
      int ctas = std::max(1, std::min(s.n_cand, (per_sm * h->n_sm + c.n_queries - 1) / c.n_queries));
      ctas = std::max(ctas, (s.n_cand + 31) / 32);     // large populations: tiles of <= 32 candidates, several rounds per CTA
      {
        // More tiles than resident CTAs (several large queries in one launch, or one very large population): size the tiles so
        // that they fill a whole number of waves - 3 x 4096 candidates in 32-candidate tiles are 384 tiles on 296 CTAs, 1.3
        // waves that cost 2; in 21-candidate tiles they are 588, 1.99 waves.  Shapes that fit one wave are left as they were.
        const long long slots = (long long)per_sm * h->n_sm, total_cand = (long long)c.n_queries * s.n_cand;
        if ((long long)c.n_queries * ctas > slots) {
          const long long waves = (total_cand + slots * 32 - 1) / (slots * 32);
          const int cpc = (int)std::min<long long>(32, std::max<long long>(1, (total_cand + waves * slots - 1) / (waves * slots)));
          ctas = std::max(ctas, (s.n_cand + cpc - 1) / cpc);
        }
      }
      if (hint_cpc) ctas = (s.n_cand + hint_cpc - 1) / hint_cpc;
      for (;;) {
        P.cand_per_cta = (s.n_cand + ctas - 1) / ctas;
        P.ctas_per_query = (s.n_cand + P.cand_per_cta - 1) / P.cand_per_cta;
        h->smem = iterate_smem_bytes(h->dev, P, h->threads);
        if (h->smem <= std::min<size_t>(smem_cap, 100 * 1024) || P.cand_per_cta == 1) break;
        ctas = std::min(s.n_cand, ctas * 2);
      }
      if (h->smem > smem_cap) { set_err(err, errlen, "problem shape needs more shared memory than a CTA has"); delete h; return SSPPCU_ELIMIT; }
      P.n_tiles = P.ctas_per_query * c.n_queries;
      // ranking scratch (phase B reuses the scoring shared memory): a chunk of keys + this CTA's own keys and counters
      const long long total = (long long)c.n_queries * s.n_cand;
      const int grid_guess = std::min(P.n_tiles, per_sm * h->n_sm);
      const long long per = (total + grid_guess - 1) / grid_guess;
      const size_t fixed = (size_t)per * 16 + 64;
      if (fixed + 256 * 8 > smem_cap) { set_err(err, errlen, "population too large for the ranking scratch"); delete h; return SSPPCU_ELIMIT; }
      h->smem = std::max(h->smem, std::min<size_t>(std::min<size_t>((size_t)s.n_cand + 1, 8192) * 8 + fixed, smem_cap));
      int got = 0;
      if ((ce = iterate_occupancy(h->variant, h->threads, h->smem + h->smem_model, &got)) != cudaSuccess || got < 1) {
        set_err(err, errlen, std::string("occupancy query failed: ") + cudaGetErrorString(ce)); delete h; return SSPPCU_ECUDA;
      }
      h->grid = std::min(P.n_tiles, got * h->n_sm);
      const long long per2 = (total + h->grid - 1) / h->grid;
      const size_t fixed2 = (size_t)per2 * 16 + 64;
      if (fixed2 + 256 * 8 > h->smem) { per_sm = std::max(1, got - 1); continue; }
      P.rank_chunk = (int)std::min<size_t>((size_t)s.n_cand + 1, (h->smem - fixed2) / 8) & ~1;   // even: 16-byte chunk rows
      if (got == per_sm || round == 2) break;
      per_sm = got;
    }
    {   // whatever round the loop left on: the ranking scratch of the FINAL grid must fit, and a chunk must hold keys
      const long long total = (long long)c.n_queries * s.n_cand;
      const size_t fixed = (size_t)((total + h->grid - 1) / h->grid) * 16 + 64;
      if (fixed + 256 * 8 > h->smem) {
        h->smem = std::min<size_t>(fixed + std::min<size_t>((size_t)s.n_cand + 1, 8192) * 8, smem_cap);
        if (fixed + 256 * 8 > h->smem) { set_err(err, errlen, "population too large for the ranking scratch of the final grid"); delete h; return SSPPCU_ELIMIT; }
        int got = 0;
        if ((ce = iterate_occupancy(h->variant, h->threads, h->smem + h->smem_model, &got)) != cudaSuccess || got < 1 ||
            (long long)got * h->n_sm < h->grid) {
          set_err(err, errlen, "the grid does not stay co-resident with the ranking scratch it needs"); delete h; return SSPPCU_ELIMIT;
        }
      }
      P.rank_chunk = (int)std::min<size_t>((size_t)s.n_cand + 1, (h->smem - fixed) / 8) & ~1;
      if (P.rank_chunk < 2) { set_err(err, errlen, "ranking chunk too small"); delete h; return SSPPCU_ELIMIT; }
    }
  } else {
    if (h->smem > smem_cap) { set_err(err, errlen, "problem shape needs more shared memory than a CTA has"); delete h; return SSPPCU_ELIMIT; }
    P.n_tiles = c.n_queries;
    h->grid = P.n_tiles;
  }
  P.smem_words = (int)(h->smem / 4);
  if ((ce = cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking)) != cudaSuccess) { set_err(err, errlen, std::string("cudaStreamCreate: ") + cudaGetErrorString(ce)); delete h; return SSPPCU_ECUDA; }
  h->stream = h->own_stream;
  // ---- device buffers ----
  const size_t Q = c.n_queries, N = s.n_cand, nfd = (size_t)n_free * s.D;
  float *d_psi_col = nullptr, *d_psi_arc = nullptr, *d_base = nullptr;
  int *d_first_col = nullptr, *d_first_arc = nullptr;
  int rc = 0;
  rc |= dalloc(h, &d_psi_col, band_col.size());
  rc |= dalloc(h, &d_psi_arc, band_arc.size());
  rc |= dalloc(h, &d_first_col, first_col.size());
  rc |= dalloc(h, &d_first_arc, first_arc.size());
  rc |= dalloc(h, &d_base, Q * s.K * s.D);
  // sampling state in ONE block (mean | sigma | best | best_valid | any_success) + its snapshot, so that
  // save / restore is a single device-to-device copy
  h->state_words = 3 * Q * nfd + 2 * Q;
  rc |= dalloc(h, &h->d_live, h->state_words); rc |= dalloc(h, &h->d_state, h->state_words);
  if (!rc) {
    P.mean = h->d_live; P.sigma = P.mean + Q * nfd; P.best = P.sigma + Q * nfd;
    P.best_valid = reinterpret_cast<int*>(P.best + Q * nfd); P.any_success = P.best_valid + Q;
  }
  rc |= dalloc(h, &P.params, Q * N * nfd);
  rc |= dalloc(h, &P.keys, Q * P.n_pad); rc |= dalloc(h, &P.keys_raw, Q * P.n_pad); rc |= dalloc(h, &P.elite_w, Q * P.n_pad);
  rc |= dalloc(h, &P.succ_acc, Q);
  rc |= dalloc(h, &P.tile_counter, 1);
  {
    size_t off = 0;
    auto take = [&](size_t bytes) { const size_t o = off; off += (bytes + 15) & ~(size_t)15; return o; };
    h->o_L = take(Q * N * 4); h->o_Cnf = take(Q * N * 4); h->o_Cwf = take(Q * N * 4); h->o_cost = take(Q * N * 4);
    h->o_status = take(Q * N); h->o_nsucc = take(Q * 4); h->o_nelite = take(Q * 4); h->o_bidx = take(Q * 4); h->o_bcost = take(Q * 4);
    h->res_bytes = off;
    h->o_extra = take(Q * nfd * 4 + 2 * Q * 4);    // host mirror only: best | best_valid | any_success (they live in the state block)
    if (!rc) {
      cudaError_t e1 = cudaMalloc(&h->d_res, h->res_bytes);
      if (e1 == cudaSuccess) e1 = cudaMemset(h->d_res, 0, h->res_bytes);
      if (e1 == cudaSuccess) e1 = cudaHostAlloc(&h->h_res, off, cudaHostAllocMapped);
      if (e1 == cudaSuccess) e1 = cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->h_res_dev), h->h_res, 0);
      if (e1 == cudaSuccess) std::memset(h->h_res, 0, off);
      if (const char* m = std::getenv("SSPPCU_RESULT_MIRROR")) h->mirror_on = std::atoi(m) != 0;
      h->stage_bytes = std::max<size_t>(64 * 1024, 4 * (Q * s.K * s.D + 2 * Q * nfd) * sizeof(float));
      if (e1 == cudaSuccess) e1 = cudaMallocHost(&h->h_stage, h->stage_bytes);
      if (e1 != cudaSuccess) rc |= cuda_fail(h, e1, "result / staging blocks");
      else {
        P.L = reinterpret_cast<float*>(h->d_res + h->o_L); P.Cnf = reinterpret_cast<float*>(h->d_res + h->o_Cnf);
        P.Cwf = reinterpret_cast<float*>(h->d_res + h->o_Cwf); P.cost = reinterpret_cast<float*>(h->d_res + h->o_cost);
        P.status = h->d_res + h->o_status;
        P.n_success = reinterpret_cast<int*>(h->d_res + h->o_nsucc); P.n_elite = reinterpret_cast<int*>(h->d_res + h->o_nelite);
        P.best_idx = reinterpret_cast<int*>(h->d_res + h->o_bidx); P.best_cost = reinterpret_cast<float*>(h->d_res + h->o_bcost);
      }
    }
  }
  if (rc) { set_err(err, errlen, h->err); ssppcu_destroy(h); return SSPPCU_ECUDA; }
  P.psi_col = d_psi_col; P.psi_arc = d_psi_arc; P.base = d_base; P.first_col = d_first_col; P.first_arc = d_first_arc;
  ce = cudaMemcpy(d_psi_col, band_col.data(), band_col.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess) ce = cudaMemcpy(d_first_col, first_col.data(), first_col.size() * sizeof(int), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess && !band_arc.empty()) ce = cudaMemcpy(d_psi_arc, band_arc.data(), band_arc.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (ce == cudaSuccess && !band_arc.empty()) ce = cudaMemcpy(d_first_arc, first_arc.data(), first_arc.size() * sizeof(int), cudaMemcpyHostToDevice);
  if (ce != cudaSuccess) { set_err(err, errlen, std::string("cudaMemcpy: ") + cudaGetErrorString(ce)); ssppcu_destroy(h); return SSPPCU_ECUDA; }
  // ---- scene tables in global memory (the kernels stage them into shared memory) ----
  {
    const DevModel& D = h->dev;
    ssppm::DevBody* d_body = nullptr; ssppm::DevMGeom* d_mg = nullptr; ssppm::DevSGeom* d_sg = nullptr;
    ssppm::DevPair* d_pair = nullptr; ssppm::DevMMPair* d_mm = nullptr; ssppm::DevSphere* d_sph = nullptr;
    int rt = 0;
    rt |= dalloc(h, &d_body, D.n_mbody); rt |= dalloc(h, &d_mg, D.n_mgeom); rt |= dalloc(h, &d_sg, D.n_sgeom);
    rt |= dalloc(h, &d_pair, D.n_pair + D.n_bpair); rt |= dalloc(h, &d_mm, D.n_mmpair); rt |= dalloc(h, &d_sph, D.n_sphere);
    if (rt) { set_err(err, errlen, h->err); ssppcu_destroy(h); return SSPPCU_ECUDA; }
    ce = cudaMemcpy(d_body, D.body, D.n_mbody * sizeof(ssppm::DevBody), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(d_mg, D.mg, D.n_mgeom * sizeof(ssppm::DevMGeom), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(d_sg, D.sg, D.n_sgeom * sizeof(ssppm::DevSGeom), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(d_pair, D.pair, (D.n_pair + D.n_bpair) * sizeof(ssppm::DevPair), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(d_mm, D.mm, D.n_mmpair * sizeof(ssppm::DevMMPair), cudaMemcpyHostToDevice);
    if (ce == cudaSuccess) ce = cudaMemcpy(d_sph, D.sph, D.n_sphere * sizeof(ssppm::DevSphere), cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) { set_err(err, errlen, std::string("cudaMemcpy (scene tables): ") + cudaGetErrorString(ce)); ssppcu_destroy(h); return SSPPCU_ECUDA; }
    static_cast<ssppm::DevModelHead&>(h->view) = D;
    h->view.body = d_body; h->view.mg = d_mg; h->view.sg = d_sg; h->view.pair = d_pair; h->view.mm = d_mm; h->view.sph = d_sph;
  }
  if (h->dev.sdf_mode) {
    ssppm::SdfGrid& g = h->dev.grid;
    const size_t n_nodes = (size_t)g.nx * g.ny * g.nz, n_cells = (size_t)(g.nx - 1) * (g.ny - 1) * (g.nz - 1);
    float *d_nodes = nullptr, *d_cells = nullptr;
    if (dalloc(h, &d_nodes, n_nodes) || dalloc(h, &d_cells, n_cells * 8)) { set_err(err, errlen, h->err); ssppcu_destroy(h); return SSPPCU_ECUDA; }
    g.nodes = d_nodes; g.cells = d_cells;
    h->view.grid = g;
    ce = launch_sdf_bake(h->view, d_nodes, d_cells, h->n_sm, h->stream);
    if (ce == cudaSuccess) ce = cudaStreamSynchronize(h->stream);
    if (ce != cudaSuccess) { set_err(err, errlen, std::string("SDF bake: ") + cudaGetErrorString(ce)); ssppcu_destroy(h); return SSPPCU_ECUDA; }
  }
  h->base_host.assign(Q * s.K * s.D, 0.0f);
  *out = h;
  return SSPPCU_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// queries, distribution, iteration, read-back
// ------------------------------------------------------------------------------------------
namespace {

int check(ssppcu_t* h) {
  if (!h) return SSPPCU_EINVAL;
  h->best_fresh = false; h->res_fresh = false;
  cudaError_t e = cudaSetDevice(h->cfg.device);
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaSetDevice");
  return 0;
}

template <class T>
int d2h(ssppcu_t* h, T* dst, const T* src, size_t n) {
  if (!dst) return 0;
  cudaError_t e = cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, h->stream);
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMemcpyAsync D2H");
  return 0;
}
// host -> device through the pinned staging ring: asynchronous on the handle's stream; the ring is drained
// (one stream synchronisation) only when it wraps
template <class T>
int h2d(ssppcu_t* h, T* dst, const T* src, size_t n) {
  const size_t bytes = n * sizeof(T);
  if (bytes == 0) return 0;
  if (bytes > h->stage_bytes) {                       // larger than the ring: plain synchronous copy
    cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) return cuda_fail(h, e, "cudaMemcpyAsync H2D");
    return 0;
  }
  if (h->stage_at + bytes > h->stage_bytes) {
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (e != cudaSuccess) return cuda_fail(h, e, "cudaStreamSynchronize");
    h->stage_at = 0;
  }
  unsigned char* slot = h->h_stage + h->stage_at;
  std::memcpy(slot, src, bytes);
  h->stage_at += (bytes + 15) & ~(size_t)15;
  cudaError_t e = cudaMemcpyAsync(dst, slot, bytes, cudaMemcpyHostToDevice, h->stream);
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMemcpyAsync H2D");
  return 0;
}

}  // namespace

extern "C" {

int ssppcu_dims(const ssppcu_t* h, SsppDims* o) {
  if (!h || !o) return SSPPCU_EINVAL;
  const ScoreConfig& s = h->P.cfg;
  o->n_queries = h->P.n_query; o->n_cand = s.n_cand; o->n_free = s.n_free; o->dof = s.D; o->n_points = s.K;
  o->nfd = s.n_free * s.D; o->n_pad = h->P.n_pad; o->rows_collision = s.Mc; o->rows_arc = s.arc_same ? s.Mc : s.Ma;
  o->first_free_row = s.free_begin; o->threads = h->threads; o->cand_per_cta = h->P.cand_per_cta;
  o->ctas_per_query = h->P.ctas_per_query; o->smem_bytes = (int64_t)(h->smem + h->smem_model);
  o->variant = h->variant; o->grid = h->grid;
  o->n_mbody = h->dev.n_mbody; o->n_mgeom = h->dev.n_mgeom; o->n_sgeom = h->dev.n_sgeom; o->n_pair = h->dev.n_pair;
  o->n_mmpair = h->dev.n_mmpair; o->chain = h->dev.chain;
  return SSPPCU_OK;
}

int ssppcu_set_stream(ssppcu_t* h, void* s) {
  if (!h) return SSPPCU_EINVAL;
  if (h->stage_at) { cudaSetDevice(h->cfg.device); cudaStreamSynchronize(h->stream); h->stage_at = 0; }   // staged copies of the old stream
  h->stream = s ? static_cast<cudaStream_t>(s) : h->own_stream;
  return SSPPCU_OK;
}

}  // extern "C"

namespace {
// host FP64 set-up of a query (sspp.h:82-97 initializePath / tsp_path_model.h:11-19 setupLinear): fills h->base_host
void fill_base(ssppcu_t* h, const double* start, const double* end);
// initial sampling state of the planners: mean | sigma [Q][nfd] each
void fill_state_task(const ssppcu_t* h, double sigma0, float* mean, float* sigma);                       // Planner::reset, tsp_planner.h:54-69
void fill_state_joint(const ssppcu_t* h, double sg, const double* limits, float* mean, float* sigma);   // sspp.h:114-130 noise scale
}  // namespace

extern "C" {

int ssppcu_set_queries(ssppcu_t* h, const double* start, const double* end) {
  if (int rc = check(h)) return rc;
  if (!start || !end) return fail(h, SSPPCU_EINVAL, "null start/end");
  fill_base(h, start, end);
  return h2d(h, const_cast<float*>(h->P.base), h->base_host.data(), h->base_host.size());
}

}  // extern "C"

namespace {
void fill_base(ssppcu_t* h, const double* start, const double* end) {
  const ScoreConfig& s = h->P.cfg;
  const int Q = h->P.n_query, K = s.K, D = s.D;
  std::vector<double> V((size_t)K * D);
  for (int q = 0; q < Q; ++q) {
    for (int i = 0; i < K; ++i) {
      const double t = (double)i / (K - 1);
      for (int d = 0; d < D; ++d) V[(size_t)i * D + d] = (1 - t) * start[(size_t)q * D + d] + t * end[(size_t)q * D + d];
    }
    float* dst = &h->base_host[(size_t)q * K * D];
    if (s.mode == ssppm::MODE_TASK) {
      for (size_t e = 0; e < V.size(); ++e) dst[e] = (float)V[e];
    } else {   // control points of the interpolating spline: ctrl = Ainv * V (sspp.h:95)
      for (int i = 0; i < K; ++i)
        for (int d = 0; d < D; ++d) {
          double acc = 0.0;
          for (int k = 0; k < K; ++k) acc += h->spl.Ainv[(size_t)i * K + k] * V[(size_t)k * D + d];
          dst[(size_t)i * D + d] = (float)acc;
        }
    }
  }
}
void fill_state_task(const ssppcu_t* h, double sigma0, float* mean, float* sigma) {
  const ScoreConfig& s = h->P.cfg;
  const int Q = h->P.n_query, D = s.D, nfd = s.n_free * D;
  for (int q = 0; q < Q; ++q)
    for (int r = 0; r < s.n_free; ++r)
      for (int d = 0; d < D; ++d) {
        double v = h->base_host[((size_t)q * s.K + s.free_begin + r) * D + d];
        if (s.mode == ssppm::MODE_TASK && d == 2) { v = std::max(v, h->cfg.sample_zmin); v = std::max(v, h->cfg.dist_zmin); }
        v = std::min(std::max(v, h->cfg.lo[d]), h->cfg.hi[d]);
        mean[(size_t)q * nfd + r * D + d] = (float)v;
        double sg = std::max(std::min(std::max(sigma0, h->cfg.stddev_min), h->cfg.stddev_max), h->cfg.sigma_floor);
        sigma[(size_t)q * nfd + r * D + d] = (float)sg;
      }
}
void fill_state_joint(const ssppcu_t* h, double sg, const double* limits, float* mean, float* sigma) {
  const ScoreConfig& s = h->P.cfg;
  const int Q = h->P.n_query, D = s.D, nfd = s.n_free * D;
  for (int q = 0; q < Q; ++q)
    for (int r = 0; r < s.n_free; ++r)
      for (int d = 0; d < D; ++d) {
        mean[(size_t)q * nfd + r * D + d] = h->base_host[((size_t)q * s.K + s.free_begin + r) * D + d];
        sigma[(size_t)q * nfd + r * D + d] = (float)(sg * limits[d]);
      }
}
}  // namespace

extern "C" {

int ssppcu_reset_distribution(ssppcu_t* h, double sigma0) {
  if (int rc = check(h)) return rc;
  const size_t n = (size_t)h->P.n_query * h->P.cfg.n_free * h->P.cfg.D;
  const int Q = h->P.n_query;
  std::vector<float> ms(2 * n);                             // mean | sigma are neighbours in the live state block: one copy
  fill_state_task(h, sigma0, ms.data(), ms.data() + n);
  if (int rc = h2d(h, h->P.mean, ms.data(), ms.size())) return rc;
  CU(h, cudaMemsetAsync(h->P.best_valid, 0, 2 * Q * sizeof(int), h->stream));   // best_valid | any_success
  return SSPPCU_OK;
}

int ssppcu_set_distribution(ssppcu_t* h, const double* mean, const double* sigma) {
  if (int rc = check(h)) return rc;
  const size_t n = (size_t)h->P.n_query * h->P.cfg.n_free * h->P.cfg.D;
  std::vector<float> tmp(2 * n);
  if (mean) for (size_t i = 0; i < n; ++i) tmp[i] = (float)mean[i];
  if (sigma) for (size_t i = 0; i < n; ++i) tmp[n + i] = (float)sigma[i];
  if (mean && sigma) return h2d(h, h->P.mean, tmp.data(), 2 * n);            // neighbours in the live state block: one copy
  if (mean) return h2d(h, h->P.mean, tmp.data(), n);
  if (sigma) return h2d(h, h->P.sigma, tmp.data() + n, n);
  return SSPPCU_OK;
}

int ssppcu_get_distribution(ssppcu_t* h, double* mean, double* sigma) {
  if (int rc = check(h)) return rc;
  const size_t n = (size_t)h->P.n_query * h->P.cfg.n_free * h->P.cfg.D;
  std::vector<float> a(n), b(n);
  if (int rc = d2h(h, a.data(), h->P.mean, n)) return rc;
  if (int rc = d2h(h, b.data(), h->P.sigma, n)) return rc;
  CU(h, cudaStreamSynchronize(h->stream));
  if (mean) for (size_t i = 0; i < n; ++i) mean[i] = a[i];
  if (sigma) for (size_t i = 0; i < n; ++i) sigma[i] = b[i];
  return SSPPCU_OK;
}

int ssppcu_set_joint_noise(ssppcu_t* h, double sg, const double* limits) {
  if (int rc = check(h)) return rc;
  if (!limits) return fail(h, SSPPCU_EINVAL, "null limits");
  const size_t n = (size_t)h->P.n_query * h->P.cfg.n_free * h->P.cfg.D;
  const int Q = h->P.n_query;
  std::vector<float> ms(2 * n);
  fill_state_joint(h, sg, limits, ms.data(), ms.data() + n);
  CU(h, cudaMemsetAsync(h->P.best_valid, 0, 2 * Q * sizeof(int), h->stream));   // best_valid | any_success
  return h2d(h, h->P.mean, ms.data(), ms.size());
}

int ssppcu_save_state(ssppcu_t* h) {
  if (int rc = check(h)) return rc;
  CU(h, cudaMemcpyAsync(h->d_state, h->d_live, h->state_words * 4, cudaMemcpyDeviceToDevice, h->stream));
  h->state_saved = true;
  return SSPPCU_OK;
}
int ssppcu_restore_state(ssppcu_t* h) {
  if (int rc = check(h)) return rc;
  if (!h->state_saved) return fail(h, SSPPCU_EINVAL, "no saved state");
  CU(h, cudaMemcpyAsync(h->d_live, h->d_state, h->state_words * 4, cudaMemcpyDeviceToDevice, h->stream));
  return SSPPCU_OK;
}

int ssppcu_fp32_peak(int device, double* tflops, double* sm_clock_mhz) {
  if (!tflops) return SSPPCU_EINVAL;
  if (cudaSetDevice(device) != cudaSuccess) return SSPPCU_ECUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return SSPPCU_ECUDA;
  float* sink = nullptr;
  if (cudaMalloc(&sink, 4) != cudaSuccess) return SSPPCU_ECUDA;
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 4096;
  double best = 0.0;
  for (int rep = 0; rep < 6; ++rep) {
    cudaEventRecord(a, 0);
    launch_fma_probe(sink, blocks, threads, iters, 0);
    cudaEventRecord(b, 0);
    if (cudaEventSynchronize(b) != cudaSuccess) { cudaFree(sink); return SSPPCU_ECUDA; }
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, a, b);
    const double flop = 2.0 * 8 * 16 * (double)iters * threads * blocks;
    if (rep > 0) best = std::max(best, flop / (ms * 1e-3) * 1e-12);
  }
  cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(sink);
  *tflops = best;
  if (sm_clock_mhz) *sm_clock_mhz = prop.clockRate * 1e-3;
  return SSPPCU_OK;
}

// inline_state: NULL, or mean | sigma [Q][nfd] each of a cold start - carried with h->base_host in the parameter block
static int iterate_impl(ssppcu_t* h, uint64_t seed, uint32_t iter0, int n_iters, bool restored, const float* inline_state = nullptr) {
  if (int rc = check(h)) return rc;
  if (n_iters < 1) return fail(h, SSPPCU_EINVAL, "n_iters must be positive");
  if (restored && !h->state_saved) return fail(h, SSPPCU_EINVAL, "no saved state");
  IterParams P = h->P;
  P.seed = seed; P.iter0 = iter0; P.n_inner = n_iters;
  P.state_in = restored ? h->d_state : nullptr;
  if (inline_state) {
    const size_t nb = h->base_host.size(), n = (size_t)P.n_query * P.cfg.n_free * P.cfg.D;
    std::memcpy(P.inline_blk, h->base_host.data(), nb * sizeof(float));
    float* st = P.inline_blk + nb;                           // mean | sigma | best (zeros) | best_valid | any_success (zeros)
    std::memcpy(st, inline_state, 2 * n * sizeof(float));
    std::memset(st + 2 * n, 0, (n + 2 * (size_t)P.n_query) * sizeof(float));
    P.inline_words = (int)(nb + 3 * n + 2 * (size_t)P.n_query);
  }
  const bool mirror = h->mirror_on && P.coop && !P.sort_only && h->h_res_dev;
  if (mirror) {
    P.host_out = h->h_res_dev; P.res_dev = h->d_res;
    P.mir_bytes = (unsigned)h->res_bytes; P.mir_nsucc = (unsigned)h->o_nsucc; P.mir_nelite = (unsigned)h->o_nelite;
    P.mir_bidx = (unsigned)h->o_bidx; P.mir_extra = (unsigned)h->o_extra;
  }
  CU(h, launch_iterate(h->view, P, h->variant, h->threads, h->smem + h->smem_model, h->grid, h->stream));
  h->res_fresh = mirror;
  ++h->launch_epoch;
  if (h->rows_on) {
    CU(h, launch_success_rows(h->P, reinterpret_cast<float*>(h->h_rows_dev), reinterpret_cast<int32_t*>(h->h_rows_dev + h->o_ridx),
                              reinterpret_cast<int32_t*>(h->h_rows_dev + h->o_rcount), h->stream));
    h->rows_epoch = h->launch_epoch;
  }
  return SSPPCU_OK;
}

static_assert((int)SSPPCU_WC_COUNT == (int)ssppm::WC_COUNT, "include/sspp_cuda.h mirrors ssppm::WorkCounter");
int ssppcu_work_counters(ssppcu_t* h, uint64_t* out, int cap) {
  if (int rc = check(h)) return rc;
  if (!out || cap < 0) return fail(h, SSPPCU_EINVAL, "bad arguments");
  unsigned long long* d_tot = nullptr;
  CU(h, cudaMalloc(&d_tot, ssppm::WC_COUNT * sizeof(unsigned long long)));
  cudaError_t e = cudaMemsetAsync(d_tot, 0, ssppm::WC_COUNT * sizeof(unsigned long long), h->stream);
  if (e == cudaSuccess) e = launch_count(h->view, h->P, d_tot, h->n_sm, h->stream);
  unsigned long long tot[ssppm::WC_COUNT];
  if (e == cudaSuccess) e = cudaMemcpyAsync(tot, d_tot, sizeof(tot), cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  cudaFree(d_tot);
  if (e != cudaSuccess) return cuda_fail(h, e, "ssppcu_work_counters");
  for (int i = 0; i < cap && i < ssppm::WC_COUNT; ++i) out[i] = tot[i];
  return ssppm::WC_COUNT;
}

int ssppcu_phase_probe(ssppcu_t* h, int enable, unsigned long long* stamps, int cap_ctas) {
  if (int rc = check(h)) return rc;
  if (!h->P.coop) return fail(h, SSPPCU_EINVAL, "the phase probe exists in the cooperative kernel only");
  if (enable && !h->d_probe) {
    CU(h, cudaMalloc(&h->d_probe, (size_t)h->grid * kProbeSlots * sizeof(unsigned long long)));
    CU(h, cudaMemset(h->d_probe, 0, (size_t)h->grid * kProbeSlots * sizeof(unsigned long long)));
  }
  if (stamps && h->d_probe) {
    CU(h, cudaStreamSynchronize(h->stream));
    CU(h, cudaMemcpy(stamps, h->d_probe, (size_t)std::min(cap_ctas, h->grid) * kProbeSlots * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  }
  h->P.probe = enable ? h->d_probe : nullptr;
  return h->grid;
}

int ssppcu_iterate(ssppcu_t* h, uint64_t seed, uint32_t iter0, int n_iters) { return iterate_impl(h, seed, iter0, n_iters, false); }
int ssppcu_iterate_restored(ssppcu_t* h, uint64_t seed, uint32_t iter0, int n_iters) { return iterate_impl(h, seed, iter0, n_iters, true); }

int ssppcu_plan_step(ssppcu_t* h, const double* start, const double* end, double sigma, const double* limits, uint64_t seed,
                     uint32_t iter0, int n_iters) {
  if (!h) return SSPPCU_EINVAL;
  if (!start || !end) return fail(h, SSPPCU_EINVAL, "null start/end");
  const bool joint = h->P.cfg.mode == ssppm::MODE_JOINT;
  if (joint && !limits) return fail(h, SSPPCU_EINVAL, "null limits");
  const size_t n = (size_t)h->P.n_query * h->P.cfg.n_free * h->P.cfg.D;
  const size_t words = (size_t)h->P.n_query * h->P.cfg.K * h->P.cfg.D + 3 * n + 2 * (size_t)h->P.n_query;
  if (words > (size_t)kInlineWords) {                        // too many queries for the parameter block: the three-call path
    if (int rc = ssppcu_set_queries(h, start, end)) return rc;
    if (int rc = joint ? ssppcu_set_joint_noise(h, sigma, limits) : ssppcu_reset_distribution(h, sigma)) return rc;
    return iterate_impl(h, seed, iter0, n_iters, false);
  }
  fill_base(h, start, end);
  float st[kInlineWords];
  if (joint) fill_state_joint(h, sigma, limits, st, st + n); else fill_state_task(h, sigma, st, st + n);
  return iterate_impl(h, seed, iter0, n_iters, false, st);
}

int ssppcu_set_success_rows(ssppcu_t* h, int on) {
  if (int rc = check(h)) return rc;
  if (on && !h->h_rows) {
    const size_t Q = h->P.n_query, N = h->P.cfg.n_cand, nfd = (size_t)h->P.cfg.n_free * h->P.cfg.D;
    h->o_ridx = (Q * N * nfd * 4 + 15) & ~(size_t)15;
    h->o_rcount = h->o_ridx + ((Q * N * 4 + 15) & ~(size_t)15);
    CU(h, cudaHostAlloc(&h->h_rows, h->o_rcount + Q * 4, cudaHostAllocMapped));
    CU(h, cudaHostGetDevicePointer(reinterpret_cast<void**>(&h->h_rows_dev), h->h_rows, 0));
    std::memset(h->h_rows + h->o_rcount, 0, Q * 4);
  }
  h->rows_on = on != 0;
  return SSPPCU_OK;
}

int ssppcu_get_success_rows(ssppcu_t* h, int query, const float** rows, const int32_t** idx, int32_t* n) {
  if (int rc = check(h)) return rc;
  if (!h->rows_on || !h->h_rows) return fail(h, SSPPCU_EINVAL, "success rows are off (ssppcu_set_success_rows)");
  if (query < 0 || query >= h->P.n_query) return fail(h, SSPPCU_EINVAL, "query out of range");
  if (h->rows_epoch != h->launch_epoch) return fail(h, SSPPCU_EINVAL, "no iteration was launched since success rows were switched on");
  CU(h, cudaStreamSynchronize(h->stream));
  const size_t N = h->P.cfg.n_cand, nfd = (size_t)h->P.cfg.n_free * h->P.cfg.D;
  if (rows) *rows = reinterpret_cast<const float*>(h->h_rows) + (size_t)query * N * nfd;
  if (idx) *idx = reinterpret_cast<const int32_t*>(h->h_rows + h->o_ridx) + (size_t)query * N;
  if (n) *n = reinterpret_cast<const int32_t*>(h->h_rows + h->o_rcount)[query];
  return SSPPCU_OK;
}

int ssppcu_candidate_ctrl(ssppcu_t* h, int query, uint32_t idx, float* out) {
  if (int rc = check(h)) return rc;
  const ScoreConfig& s = h->P.cfg;
  if (!out || query < 0 || query >= h->P.n_query || idx >= (uint32_t)s.n_cand) return fail(h, SSPPCU_EINVAL, "bad arguments");
  const int D = s.D, nfd = s.n_free * D;
  std::memcpy(out, &h->base_host[(size_t)query * s.K * D], (size_t)s.K * D * sizeof(float));
  CU(h, cudaMemcpyAsync(out + (size_t)s.free_begin * D, h->P.params + ((size_t)query * s.n_cand + idx) * nfd, (size_t)nfd * sizeof(float),
                        cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return SSPPCU_OK;
}

int ssppcu_set_result_mirror(ssppcu_t* h, int on) {
  if (int rc = check(h)) return rc;
  h->mirror_on = on != 0;
  return SSPPCU_OK;
}

int ssppcu_sync(ssppcu_t* h) {
  if (int rc = check(h)) return rc;
  CU(h, cudaStreamSynchronize(h->stream));
  return SSPPCU_OK;
}

int ssppcu_get_results(ssppcu_t* h, float* L, float* Cnf, float* Cwf, float* cost, uint8_t* status, int32_t* n_success,
                       int32_t* n_elite, int32_t* best_idx, float* best_cost, int32_t* any_success) {
  const bool mirrored = h && h->res_fresh;                   // the launch just before this call wrote the pinned mirror itself
  if (int rc = check(h)) return rc;
  const size_t Q = h->P.n_query, QN = Q * h->P.cfg.n_cand;
  const size_t nfd = (size_t)h->P.cfg.n_free * h->P.cfg.D;
  int32_t* h_any = reinterpret_cast<int32_t*>(h->h_res + h->o_extra + Q * nfd * 4 + Q * 4);   // mirror tail: best | best_valid | any_success
  if (!mirrored) {
    // one copy of the result block into its pinned mirror (per-candidate arrays only when somebody wants them)
    const bool per_cand = L || Cnf || Cwf || cost || status;
    const size_t from = per_cand ? 0 : h->o_nsucc;
    CU(h, cudaMemcpyAsync(h->h_res + from, h->d_res + from, h->res_bytes - from, cudaMemcpyDeviceToHost, h->stream));
    // best | best_valid | any_success (the tail of the live state block) ride along: a get_best right after this call needs no second round trip
    CU(h, cudaMemcpyAsync(h->h_res + h->o_extra, h->P.best, Q * nfd * 4 + 2 * Q * 4, cudaMemcpyDeviceToHost, h->stream));
  }
  CU(h, cudaStreamSynchronize(h->stream));
  if (L) std::memcpy(L, h->h_res + h->o_L, QN * 4);
  if (Cnf) std::memcpy(Cnf, h->h_res + h->o_Cnf, QN * 4);
  if (Cwf) std::memcpy(Cwf, h->h_res + h->o_Cwf, QN * 4);
  if (cost) std::memcpy(cost, h->h_res + h->o_cost, QN * 4);
  if (status) std::memcpy(status, h->h_res + h->o_status, QN);
  if (n_success) std::memcpy(n_success, h->h_res + h->o_nsucc, Q * 4);
  if (n_elite) std::memcpy(n_elite, h->h_res + h->o_nelite, Q * 4);
  if (best_idx) std::memcpy(best_idx, h->h_res + h->o_bidx, Q * 4);
  if (best_cost) std::memcpy(best_cost, h->h_res + h->o_bcost, Q * 4);
  if (any_success) std::memcpy(any_success, h_any, Q * 4);
  h->best_fresh = true;
  return SSPPCU_OK;
}

int ssppcu_get_params(ssppcu_t* h, float* params) {
  if (int rc = check(h)) return rc;
  if (!params) return fail(h, SSPPCU_EINVAL, "null params");
  const size_t n = (size_t)h->P.n_query * h->P.cfg.n_cand * h->P.cfg.n_free * h->P.cfg.D;
  if (int rc = d2h(h, params, h->P.params, n)) return rc;
  CU(h, cudaStreamSynchronize(h->stream));
  return SSPPCU_OK;
}

int ssppcu_get_best(ssppcu_t* h, float* best, int32_t* valid) {
  const bool fresh = h && h->best_fresh;                     // fetched by the ssppcu_get_results call just before this one
  if (int rc = check(h)) return rc;
  const size_t Q = h->P.n_query, nfd = (size_t)h->P.cfg.n_free * h->P.cfg.D;
  unsigned char* m = h->h_res + h->o_extra;                  // mirror tail: best | best_valid | any_success
  if (!fresh) {
    CU(h, cudaMemcpyAsync(m, h->P.best, Q * nfd * 4 + Q * 4, cudaMemcpyDeviceToHost, h->stream));   // best | best_valid
    CU(h, cudaStreamSynchronize(h->stream));
  }
  h->best_fresh = true;
  if (best) std::memcpy(best, m, Q * nfd * 4);
  if (valid) std::memcpy(valid, m + Q * nfd * 4, Q * 4);
  return SSPPCU_OK;
}

int ssppcu_get_elites(ssppcu_t* h, int query, uint32_t* idx, float* cost, float* weight, int cap) {
  if (int rc = check(h)) return rc;
  if (query < 0 || query >= h->P.n_query || cap < 0) return fail(h, SSPPCU_EINVAL, "query out of range");
  const int n = std::min(cap, h->P.cfg.n_cand);
  std::vector<unsigned long long> keys(n);
  std::vector<float> w(n);
  if (int rc = d2h(h, keys.data(), h->P.keys + (size_t)query * h->P.n_pad, n)) return rc;
  if (int rc = d2h(h, w.data(), h->P.elite_w + (size_t)query * h->P.n_pad, n)) return rc;
  CU(h, cudaStreamSynchronize(h->stream));
  for (int i = 0; i < n; ++i) {
    if (idx) idx[i] = (uint32_t)(keys[i] & 0xFFFFFFFFull);
    if (cost) {
      uint32_t ob = (uint32_t)(keys[i] >> 32);
      if (ob == 0xFFFFFFFFu) cost[i] = INFINITY;
      else { ob = (ob & 0x80000000u) ? (ob & 0x7FFFFFFFu) : ~ob; std::memcpy(&cost[i], &ob, 4); }
    }
    if (weight) weight[i] = w[i];
  }
  return n;
}

int ssppcu_device_view(ssppcu_t* h, SsppDeviceView* o) {
  if (!h || !o) return SSPPCU_EINVAL;
  const IterParams& P = h->P;
  o->L = P.L; o->Cnf = P.Cnf; o->Cwf = P.Cwf; o->cost = P.cost; o->status = P.status; o->params = P.params;
  o->mean = P.mean; o->sigma = P.sigma; o->best = P.best; o->best_valid = P.best_valid; o->keys = P.keys;
  o->elite_w = P.elite_w; o->n_success = P.n_success; o->n_elite = P.n_elite; o->best_idx = P.best_idx;
  o->best_cost = P.best_cost; o->any_success = P.any_success; o->stream = h->stream;
  return SSPPCU_OK;
}

int ssppcu_eval_points(ssppcu_t* h, const float* pts, int n, float* out) {
  if (int rc = check(h)) return rc;
  if (!pts || !out || n < 0) return fail(h, SSPPCU_EINVAL, "bad arguments");
  if (n == 0) return SSPPCU_OK;
  const int D = h->P.cfg.D;
  if (n > h->eval_cap) {
    if (h->d_eval_pts) cudaFree(h->d_eval_pts);
    if (h->d_eval_out) cudaFree(h->d_eval_out);
    h->d_eval_pts = h->d_eval_out = nullptr; h->eval_cap = 0;
    CU(h, cudaMalloc(&h->d_eval_pts, (size_t)n * D * sizeof(float)));
    CU(h, cudaMalloc(&h->d_eval_out, (size_t)n * sizeof(float)));
    h->eval_cap = n;
  }
  CU(h, cudaMemcpyAsync(h->d_eval_pts, pts, (size_t)n * D * sizeof(float), cudaMemcpyHostToDevice, h->stream));
  CU(h, launch_eval_points(h->view, D, h->d_eval_pts, n, h->d_eval_out, h->stream));
  CU(h, cudaMemcpyAsync(out, h->d_eval_out, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
  CU(h, cudaStreamSynchronize(h->stream));
  return SSPPCU_OK;
}

int ssppcu_check_points(ssppcu_t* h, const float* pts, int n, int32_t* geom1, int32_t* geom2, float* dist) {
  if (int rc = check(h)) return rc;
  if (!pts || !geom1 || !geom2 || n < 0) return fail(h, SSPPCU_EINVAL, "bad arguments");
  if (h->dev.sdf_mode) return fail(h, SSPPCU_EINVAL, "ssppcu_check_points needs the primitives collision mode (an SDF scene has no geom pairs)");
  if (n == 0) return 0;
  const int D = h->P.cfg.D;
  const size_t nm = h->mgeom_src.size(), ns = h->sgeom_src.size();
  if (!h->d_geom_src) {
    std::vector<int> both(h->mgeom_src);
    both.insert(both.end(), h->sgeom_src.begin(), h->sgeom_src.end());
    if (int rc = dalloc(h, &h->d_geom_src, both.size())) return rc;
    CU(h, cudaMemcpy(h->d_geom_src, both.data(), both.size() * sizeof(int), cudaMemcpyHostToDevice));
  }
  (void)ns;
  float* d_pts = nullptr;
  cudaError_t e = cudaMalloc(&d_pts, ((size_t)n * D + 3 * (size_t)n) * sizeof(float));
  if (e != cudaSuccess) return cuda_fail(h, e, "cudaMalloc");
  int* d_g1 = reinterpret_cast<int*>(d_pts + (size_t)n * D); int* d_g2 = d_g1 + n; float* d_dist = reinterpret_cast<float*>(d_g2 + n);
  std::vector<float> hd((size_t)n);
  e = cudaMemcpyAsync(d_pts, pts, (size_t)n * D * sizeof(float), cudaMemcpyHostToDevice, h->stream);
  if (e == cudaSuccess) e = launch_check_points(h->view, D, d_pts, n, h->d_geom_src, h->d_geom_src + nm, d_g1, d_g2, d_dist, h->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(geom1, d_g1, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(geom2, d_g2, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(hd.data(), d_dist, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  cudaFree(d_pts);
  if (e != cudaSuccess) return cuda_fail(h, e, "ssppcu_check_points");
  if (dist) std::copy(hd.begin(), hd.end(), dist);
  int hits = 0;
  for (int i = 0; i < n; ++i) hits += geom1[i] >= 0;
  return hits;
}

int ssppcu_sdf_gradient(ssppcu_t* h, const float* pts, const float* weights, int rows, float* grad, float* dist, float* dirs) {
  if (int rc = check(h)) return rc;
  if (!h->dev.sdf_mode) return fail(h, SSPPCU_EINVAL, "handle is not in SDF collision mode");
  if (!pts || !grad) return fail(h, SSPPCU_EINVAL, "bad arguments");
  const int D = h->P.cfg.D, K = h->P.cfg.K;
  if (!weights) { weights = h->psi_col.data(); rows = h->P.cfg.Mc; }      // the handle's collision collocation matrix [Mc][K]
  if (rows < 1 || rows > 65536) return fail(h, SSPPCU_EINVAL, "rows out of range");
  const size_t nW = (size_t)rows * K, nP = (size_t)K * D, nO = (size_t)rows * 4 + (size_t)K * 3;
  float* d_buf = nullptr;
  CU(h, cudaMalloc(&d_buf, (nW + nP + nO) * sizeof(float)));
  float* d_W = d_buf; float* d_pts = d_W + nW; float* d_dist = d_pts + nP; float* d_dirs = d_dist + rows; float* d_G = d_dirs + 3 * (size_t)rows;
  std::vector<float> out(nO);
  cudaError_t e = cudaMemcpyAsync(d_W, weights, nW * sizeof(float), cudaMemcpyHostToDevice, h->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_pts, pts, nP * sizeof(float), cudaMemcpyHostToDevice, h->stream);
  if (e == cudaSuccess) e = launch_sdf_gradient(h->view, D, K, rows, d_W, d_pts, d_dist, d_dirs, d_G, h->stream);
  if (e == cudaSuccess) e = cudaMemcpyAsync(out.data(), d_dist, nO * sizeof(float), cudaMemcpyDeviceToHost, h->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
  cudaFree(d_buf);
  if (e != cudaSuccess) return cuda_fail(h, e, "ssppcu_sdf_gradient");
  if (dist) std::memcpy(dist, out.data(), (size_t)rows * sizeof(float));
  if (dirs) std::memcpy(dirs, out.data() + rows, (size_t)rows * 3 * sizeof(float));
  std::memcpy(grad, out.data() + (size_t)rows * 4, (size_t)K * 3 * sizeof(float));
  return SSPPCU_OK;
}

int ssppcu_get_sdf(ssppcu_t* h, float* nodes, int32_t dims[3], double lo[3], double cell[3]) {
  if (int rc = check(h)) return rc;
  if (!h->dev.sdf_mode) return fail(h, SSPPCU_EINVAL, "handle is not in SDF collision mode");
  const ssppm::SdfGrid& g = h->dev.grid;
  if (dims) { dims[0] = g.nx; dims[1] = g.ny; dims[2] = g.nz; }
  for (int a = 0; a < 3; ++a) { if (lo) lo[a] = g.lo[a]; if (cell) cell[a] = g.h[a]; }
  if (nodes) {
    CU(h, cudaMemcpyAsync(nodes, g.nodes, (size_t)g.nx * g.ny * g.nz * sizeof(float), cudaMemcpyDeviceToHost, h->stream));
    CU(h, cudaStreamSynchronize(h->stream));
  }
  return SSPPCU_OK;
}

int ssppcu_get_spheres(ssppcu_t* h, float* out, int cap) {
  if (!h) return SSPPCU_EINVAL;
  const int n = std::min(cap, h->dev.n_sphere);
  for (int i = 0; out && i < n; ++i) {
    const ssppm::DevSphere& S = h->dev.sph[i];
    out[5 * i] = S.c[0]; out[5 * i + 1] = S.c[1]; out[5 * i + 2] = S.c[2]; out[5 * i + 3] = S.r; out[5 * i + 4] = (float)S.body;
  }
  return h->dev.n_sphere;
}

int ssppcu_get_collocation(ssppcu_t* h, float* pc, float* pa) {
  if (!h) return SSPPCU_EINVAL;
  if (pc) std::memcpy(pc, h->psi_col.data(), h->psi_col.size() * sizeof(float));
  if (pa && !h->psi_arc.empty()) std::memcpy(pa, h->psi_arc.data(), h->psi_arc.size() * sizeof(float));
  return SSPPCU_OK;
}

int ssppcu_get_base(ssppcu_t* h, float* base) {
  if (!h || !base) return SSPPCU_EINVAL;
  std::memcpy(base, h->base_host.data(), h->base_host.size() * sizeof(float));
  return SSPPCU_OK;
}

int ssppcu_shard_keys(ssppcu_t* h, unsigned long long** keys, int32_t* n) {
  if (!h || !keys || !n) return SSPPCU_EINVAL;
  *keys = h->P.keys; *n = h->P.cfg.n_cand;
  return SSPPCU_OK;
}

static int ensure_merge_buffers(ssppcu_t* h, int n_keys) {
  const int n_pad = next_pow2(std::max(2, n_keys));
  const int nfd = h->P.cfg.n_free * h->P.cfg.D;
  if (n_pad > h->merge_cap) {
    if (h->d_merge_keys) cudaFree(h->d_merge_keys);
    h->d_merge_keys = nullptr; h->merge_cap = 0;
    CU(h, cudaMalloc(&h->d_merge_keys, (size_t)n_pad * (sizeof(unsigned long long) + sizeof(float)) + (3 * nfd + 2) * sizeof(float)));
    h->merge_cap = n_pad;
  }
  return SSPPCU_OK;
}

int ssppcu_merge_update(ssppcu_t* h, const unsigned long long* gathered, int n_keys, int n_shards, uint64_t seed, uint32_t iter) {
  if (int rc = check(h)) return rc;
  if (!gathered || n_keys < 1 || n_shards < 1 || n_shards > kMaxExchangeRanks || n_keys % n_shards) return fail(h, SSPPCU_EINVAL, "bad arguments: n_keys must be n_shards (at most 16) equal shards");
  if (h->cfg.shard_global <= 0) return fail(h, SSPPCU_EINVAL, "handle is not a population shard");
  if (h->xc.attached) return fail(h, SSPPCU_EINVAL, "this handle exchanges keys itself (ssppcu_iterate_exchange)");
  if (int rc = ensure_merge_buffers(h, n_keys)) return rc;
  const int n_pad = h->merge_cap;
  MergeParams M;
  std::memset(&M, 0, sizeof(M));
  M.cfg = h->P.cfg;
  M.cfg.n_cand = h->cfg.shard_global;
  M.n_keys = n_keys; M.n_shards = n_shards; M.shard_len = n_keys / n_shards; M.n_pad = n_pad; M.iter = iter; M.seed = seed;
  M.gathered = gathered;
  M.keys = h->d_merge_keys;
  M.elite_w = reinterpret_cast<float*>(h->d_merge_keys + n_pad);
  M.state0 = M.elite_w + n_pad;
  M.succ_acc = h->P.succ_acc;
  M.mean = h->P.mean; M.sigma = h->P.sigma; M.best = h->P.best; M.best_valid = h->P.best_valid;
  M.n_success = h->P.n_success; M.n_elite = h->P.n_elite; M.best_idx = h->P.best_idx; M.best_cost = h->P.best_cost;
  M.any_success = h->P.any_success;
  CU(h, cudaMemsetAsync(h->P.succ_acc, 0, sizeof(int), h->stream));
  CU(h, launch_merge_update(M, h->stream));
  h->merged_n = n_keys;
  return SSPPCU_OK;
}

// ---- fused key exchange over peer memory (population shards) ----
int ssppcu_exchange_export(ssppcu_t* h, int world, int rank, void* block_out, void* ipc_handle_out) {
  if (int rc = check(h)) return rc;
  if (h->cfg.shard_global <= 0) return fail(h, SSPPCU_EINVAL, "handle is not a population shard");
  if (!h->P.coop) return fail(h, SSPPCU_EINVAL, "the fused exchange needs the cooperative kernel (shard too small)");
  if (world < 1 || world > kMaxExchangeRanks || rank < 0 || rank >= world) return fail(h, SSPPCU_EINVAL, "bad world / rank");
  if ((long long)h->P.cfg.n_cand * world != h->cfg.shard_global) return fail(h, SSPPCU_EINVAL, "shards must be equal: n_local * world != shard_global");
  if (h->xc.block) return fail(h, SSPPCU_EINVAL, "exchange block already exported");
  auto& x = h->xc;
  x.world = world; x.rank = rank; x.epoch = 0; x.attached = false;
  x.block_bytes = (size_t)2 * world * h->P.cfg.n_cand * sizeof(uint4);
  CU(h, cudaMalloc(&x.d_gathered, (size_t)world * h->P.cfg.n_cand * sizeof(unsigned long long)));
  CU(h, cudaMalloc(&x.block, x.block_bytes));
  CU(h, cudaMemset(x.block, 0, x.block_bytes));
  CU(h, cudaMalloc(&x.d_params, sizeof(ExchangeParams)));
  CU(h, cudaMalloc(&x.d_ticket, sizeof(int)));
  CU(h, cudaMemset(x.d_ticket, 0, sizeof(int)));
  CU(h, cudaHostAlloc(&x.h_status, sizeof(int), cudaHostAllocMapped));
  *x.h_status = 0;
  CU(h, cudaHostGetDevicePointer(&x.d_status, x.h_status, 0));
  CU(h, cudaDeviceSynchronize());                      // the zeroed lines (epoch 0 = nothing published) exist before any peer can learn the address
  if (block_out) *reinterpret_cast<void**>(block_out) = x.block;
  if (ipc_handle_out) {
    cudaIpcMemHandle_t ih;
    CU(h, cudaIpcGetMemHandle(&ih, x.block));
    static_assert(sizeof(ih) == 64, "ipc handle size");
    std::memcpy(ipc_handle_out, &ih, sizeof(ih));
  }
  return SSPPCU_OK;
}

static int exchange_finish_attach(ssppcu_t* h, void* const* bases) {
  auto& x = h->xc;
  const int n_keys = x.world * h->P.cfg.n_cand;
  if (int rc = ensure_merge_buffers(h, n_keys)) return rc;
  const int nfd = h->P.cfg.n_free * h->P.cfg.D;
  ExchangeParams E;
  std::memset(&E, 0, sizeof(E));
  E.world = x.world; E.rank = x.rank; E.shard_len = h->P.cfg.n_cand;
  E.protocol = 1;
  if (const char* pr = std::getenv("SSPPCU_XCHG_PROTOCOL")) E.protocol = std::atoi(pr) != 0;
  E.n_global = h->cfg.shard_global; E.n_pad = h->merge_cap;
  for (int r = 0; r < x.world; ++r) E.peer_lines[r] = reinterpret_cast<uint4*>(bases[r]);
  E.lines_local = reinterpret_cast<const uint4*>(x.block);
  E.gathered = x.d_gathered;
  E.succ = x.d_ticket; E.status = x.d_status;
  E.timeout_ns = 2000000000ull;
  E.mkeys = h->d_merge_keys;
  E.elite_w = reinterpret_cast<float*>(h->d_merge_keys + h->merge_cap);
  E.state0 = E.elite_w + h->merge_cap;
  CU(h, cudaMalloc(&x.d_vals, (size_t)std::max(1, nfd) * h->merge_cap * sizeof(float)));
  E.vals = x.d_vals;
  CU(h, cudaMemcpy(x.d_params, &E, sizeof(E), cudaMemcpyHostToDevice));
  x.attached = true;
  return SSPPCU_OK;
}

int ssppcu_exchange_attach_ipc(ssppcu_t* h, const void* handles) {
  if (int rc = check(h)) return rc;
  auto& x = h->xc;
  if (!x.block || !handles) return fail(h, SSPPCU_EINVAL, "export the exchange block first");
  if (x.attached) return fail(h, SSPPCU_EINVAL, "already attached");
  void* bases[kMaxExchangeRanks] = {};
  for (int r = 0; r < x.world; ++r) {
    if (r == x.rank) { bases[r] = x.block; continue; }
    cudaIpcMemHandle_t ih;
    std::memcpy(&ih, reinterpret_cast<const unsigned char*>(handles) + (size_t)r * sizeof(ih), sizeof(ih));
    cudaError_t e = cudaIpcOpenMemHandle(&bases[r], ih, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) return cuda_fail(h, e, "cudaIpcOpenMemHandle (peer exchange block)");
    x.ipc_mapped[r] = bases[r];
  }
  return exchange_finish_attach(h, bases);
}

int ssppcu_exchange_attach_ptrs(ssppcu_t* h, void* const* blocks) {
  if (int rc = check(h)) return rc;
  auto& x = h->xc;
  if (!x.block || !blocks) return fail(h, SSPPCU_EINVAL, "export the exchange block first");
  if (x.attached) return fail(h, SSPPCU_EINVAL, "already attached");
  if (blocks[x.rank] != x.block) return fail(h, SSPPCU_EINVAL, "blocks[rank] is not this handle's block");
  return exchange_finish_attach(h, blocks);
}

int ssppcu_iterate_exchange(ssppcu_t* h, uint64_t seed, uint32_t iter) {
  if (int rc = check(h)) return rc;
  auto& x = h->xc;
  if (!x.attached) return fail(h, SSPPCU_EINVAL, "exchange blocks are not attached");
  if (*reinterpret_cast<volatile int*>(x.h_status)) return fail(h, SSPPCU_ECUDA, "key exchange timed out: a peer rank did not publish its keys");
  IterParams P = h->P;
  P.seed = seed; P.iter0 = iter; P.n_inner = 1; P.state_in = nullptr;
  P.xchg = x.d_params; P.xchg_epoch = ++x.epoch;
  CU(h, launch_iterate(h->view, P, h->variant, h->threads, h->smem + h->smem_model, h->grid, h->stream));
  ++h->launch_epoch;
  h->merged_n = x.world * h->P.cfg.n_cand;
  return SSPPCU_OK;
}

int ssppcu_exchange_status(ssppcu_t* h) {
  if (int rc = check(h)) return rc;
  if (!h->xc.h_status) return fail(h, SSPPCU_EINVAL, "no exchange block");
  CU(h, cudaStreamSynchronize(h->stream));
  if (*reinterpret_cast<volatile int*>(h->xc.h_status)) return fail(h, SSPPCU_ECUDA, "key exchange timed out: a peer rank did not publish its keys");
  return SSPPCU_OK;
}

/* merged global order of the last ssppcu_merge_update: up to cap entries, elites first with weights */
int ssppcu_get_merged(ssppcu_t* h, uint32_t* idx, float* cost, float* weight, int cap) {
  if (int rc = check(h)) return rc;
  if (!h->d_merge_keys || cap < 0) return fail(h, SSPPCU_EINVAL, "no merged keys");
  const int n = std::min(cap, h->merged_n);
  std::vector<unsigned long long> keys(n);
  std::vector<float> w(n);
  if (int rc = d2h(h, keys.data(), h->d_merge_keys, n)) return rc;
  if (int rc = d2h(h, w.data(), reinterpret_cast<float*>(h->d_merge_keys + h->merge_cap), n)) return rc;
  CU(h, cudaStreamSynchronize(h->stream));
  for (int i = 0; i < n; ++i) {
    if (idx) idx[i] = (uint32_t)(keys[i] & 0xFFFFFFFFull);
    if (cost) {
      uint32_t ob = (uint32_t)(keys[i] >> 32);
      if (ob == 0xFFFFFFFFu) cost[i] = INFINITY;
      else { ob = (ob & 0x80000000u) ? (ob & 0x7FFFFFFFu) : ~ob; std::memcpy(&cost[i], &ob, 4); }
    }
    if (weight) weight[i] = w[i];
  }
  return n;
}

}  // extern "C"
