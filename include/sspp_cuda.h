/* sspp_cuda.h — C ABI of the B200 candidate-scoring path (libsspp_cuda.so).
 *
 * Plain C: opaque handles, pointers and sizes only; no exceptions cross the boundary.  Every
 * function returns 0 on success or a negative SSPPCU_E* code, with text in ssppcu_last_error()
 * (or the caller's err buffer for the constructors).  There is NO CPU fallback behind this ABI:
 * without a CUDA device ssppcu_create fails with SSPPCU_ECUDA.
 *
 * What each entry point replaces in the reference (paths under /root/reference):
 *   ssppcu_model_load / _parse   mj_loadXML                       include/sspp.h:47, src/main_icra_benchmark.cpp:146
 *   ssppcu_create                per-thread mjData pools          include/sspp.h:235-244, include/sspp/tsp_collision_world.h:21-27
 *   ssppcu_set_queries           initializePath / initLinear      include/sspp.h:82-97, include/sspp/tsp_path_model.h:11-29
 *   ssppcu_reset_distribution    Planner::reset, Distribution::reset   include/sspp/tsp_planner.h:54-69, tsp_distribution.h:16-29
 *   ssppcu_set_distribution      sigma*limits of sampleWithNoise  include/sspp.h:116-125
 *   ssppcu_iterate               the candidate loop: sample -> spline -> mj_forward -> contacts -> costs -> elites -> update
 *                                include/sspp.h:203-224 and include/sspp/tsp_planner.h:78-142
 *   ssppcu_get_results/_params   successes_/failures_/sampled_sets_, ret_success_paths   tsp_planner.h:114-118, sspp.h:223
 *   ssppcu_get_elites            EliteSelector::select / weights  include/sspp/tsp_elites.h:13-32
 *   ssppcu_get_distribution      Distribution::mean_vias / sigma_vias   include/sspp/tsp_distribution.h:7-8
 *   ssppcu_eval_points           spline(u), qpos <- point, mj_forward, contact read-out
 *                                include/sspp.h:137-144, include/Collision.h:84-103
 */
#ifndef SSPP_CUDA_H
#define SSPP_CUDA_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSPPCU_OK 0
#define SSPPCU_EINVAL (-1)   /* bad argument */
#define SSPPCU_EMODEL (-2)   /* MJCF could not be loaded / specialised */
#define SSPPCU_ECUDA (-3)    /* CUDA runtime error (including: no device) */
#define SSPPCU_ELIMIT (-4)   /* a compiled-in limit is exceeded */

#define SSPPCU_MAX_DOF 9

#define SSPPCU_MODE_JOINT 0  /* planner A: sspp::SamplingPathPlanner<kDOF>, joint space, cubic */
#define SSPPCU_MODE_TASK 1   /* planner B: tsp::Planner, (x,y,z,yaw) of a free body, quadratic */

#define SSPPCU_COLLIDE_PRIMITIVES 0  /* analytic primitive pairs (MuJoCo's narrow phase restated) */
#define SSPPCU_COLLIDE_SDF 1         /* static scene baked into a dense signed-distance grid, moving geoms sphere-sampled */

#define SSPPCU_ST_CONVERGED 0
#define SSPPCU_ST_FAILED 1
#define SSPPCU_ST_ABSENT 2   /* the forwarded-best slot of a cold iteration */

typedef struct ssppcu_model ssppcu_model_t;
typedef struct ssppcu_planner ssppcu_t;

/* Read-only view of the flat FP64 model (an mjModel subset).  Pointers stay valid until
 * ssppcu_model_free.  body_jtype uses MuJoCo's enum: -1 none, 0 free, 2 slide, 3 hinge. */
typedef struct SsppFlatModelView {
  int32_t nbody, ngeom, nq, nexclude;
  const int32_t* body_parent; const int32_t* body_weld; const int32_t* body_jtype; const int32_t* body_qadr;
  const double* body_pos;  /* 3 */  const double* body_quat; /* 4 */
  const double* jnt_axis;  /* 3 */  const double* jnt_pos;   /* 3 */  const double* jnt_range; /* 2 */
  const double* qpos0;
  const int32_t* geom_body; const int32_t* geom_type; const int32_t* geom_contype; const int32_t* geom_conaffinity;
  const double* geom_size; /* 3 */  const double* geom_pos;  /* 3 */  const double* geom_quat; /* 4 */
  const double* geom_margin;
  const int32_t* exclude;  /* 2 per entry */
} SsppFlatModelView;

typedef struct SsppPlannerConfig {
  int32_t mode;            /* SSPPCU_MODE_* */
  int32_t dof;             /* JOINT: 1..9 planned scalar joints (first qpos entries); TASK: 4 */
  int32_t degree;          /* spline degree: 3 (sspp.h:25) or 2 (tsp_types.h:9) */
  int32_t n_points;        /* init_points (sspp.h:196) / total_points (tsp_planner.h:16) */
  int32_t check_points;    /* check_points / checks */
  int32_t n_samples;       /* random candidates per iteration (this shard's share when sharded) */
  int32_t include_mean;    /* TASK seeds: mean set first (tsp_planner.h:82-85) */
  int32_t include_best;    /* TASK seeds: forwarded best second (tsp_planner.h:87) */
  int32_t do_update;       /* CES update of the distribution each iteration */
  int32_t n_queries;       /* independent planning queries held by this handle */
  int32_t device;          /* CUDA device ordinal */
  int32_t shard_offset;    /* population sharding: global index of this shard's first candidate */
  int32_t shard_global;    /* population sharding: global candidate count, 0 = not sharded */
  const char* body_name;   /* TASK: the free body that is planned */
  double w_collision, elite_fraction, inc, dec;
  double sigma_floor, var_beta, mean_lr, stddev_min, stddev_max;
  double eval_zmin, floor_margin, floor_scale;   /* tsp_evaluator.h:8 */
  double limit_scale;      /* JOINT: penalty scale*max(0,|q|-hi)^2 per coordinate and point, 0 = off */
  double sample_zmin;      /* cfg_.z_min (tsp_sampler.h:47) */
  double dist_zmin;        /* dist_.z_min (tsp_distribution.h:64) */
  double lo[SSPPCU_MAX_DOF], hi[SSPPCU_MAX_DOF];   /* sampler limits / mean clamp */
  /* collision back end */
  int32_t collision_mode;  /* SSPPCU_COLLIDE_PRIMITIVES (default) or SSPPCU_COLLIDE_SDF */
  int32_t use_center_dist; /* TASK cost per contact: 1 = -1/(centre distance + 1e-4), 0 = -dist (Collision.h:84,93-98);
                              SDF mode always uses -dist (a grid has no geom centres) */
  int32_t sdf_dims[3];     /* grid nodes per axis (SDF mode), e.g. 128 128 128 */
  int32_t tile_hint;       /* 0 = automatic.  Tuning knob: bits 0-7 kernel launch-bound variant + 1, bits 8-19 candidates
                              per CTA, bits 20-31 threads per CTA (multiple of 32) */
  double sdf_lo[3], sdf_hi[3];   /* world box the grid spans */
  double sdf_sphere_ratio; /* sample-sphere radius / smallest half extent of a moving box (default 0.5) */
} SsppPlannerConfig;

/* sizes derived from a config, for sizing caller buffers */
typedef struct SsppDims {
  int32_t n_queries, n_cand, n_free, dof, n_points, nfd /* n_free*dof */, n_pad;
  int32_t rows_collision, rows_arc, first_free_row;
  int32_t threads, cand_per_cta, ctas_per_query; int64_t smem_bytes;
  int32_t n_mbody, n_mgeom, n_sgeom, n_pair, n_mmpair, chain;
  int32_t variant, grid;   /* kernel launch-bound variant and CTAs per launch */
} SsppDims;

/* raw device pointers of the handle-owned result buffers (device-buffer path; valid until the
 * next ssppcu_iterate on the handle's stream, layouts as in ssppcu_get_results / _get_params) */
typedef struct SsppDeviceView {
  float* L; float* Cnf; float* Cwf; float* cost; uint8_t* status;
  float* params; float* mean; float* sigma; float* best; int32_t* best_valid;
  unsigned long long* keys; float* elite_w;
  int32_t* n_success; int32_t* n_elite; int32_t* best_idx; float* best_cost; int32_t* any_success;
  void* stream;
} SsppDeviceView;

/* ---- model ---- */
int ssppcu_model_load(const char* mjcf_path, ssppcu_model_t** out, char* err, size_t errlen);
int ssppcu_model_parse(const char* mjcf_text, ssppcu_model_t** out, char* err, size_t errlen);
void ssppcu_model_free(ssppcu_model_t* m);
int ssppcu_model_view(const ssppcu_model_t* m, SsppFlatModelView* out);
int ssppcu_model_body_id(const ssppcu_model_t* m, const char* name);          /* -1 when absent */
const char* ssppcu_model_body_name(const ssppcu_model_t* m, int body);
const char* ssppcu_model_geom_name(const ssppcu_model_t* m, int geom);
/* (x,y,z,yaw) of a free body at the model's initial pose (Utility::get_body_point, utility.h:228-259) */
int ssppcu_model_body_point(const ssppcu_model_t* m, const char* body, double out4[4]);
/* filtered pair count for a planner: moving-vs-static, moving-vs-moving, never-moving */
int ssppcu_model_pair_counts(const ssppcu_model_t* m, int mode, int dof, const char* body,
                             int32_t* n_moving_static, int32_t* n_moving_moving, int32_t* n_static_static,
                             char* err, size_t errlen);

/* ---- the reference's Python spline twins as product functions (host FP64) ----
 * sspp/BSplines.py:58-62 knot_vector: clamped uniform knots, n_ctrl + k + 1 values. */
int ssppcu_bspline_knot_vector(int n_ctrl, int k, double* knots);
/* Rows [n_theta][n_ctrl] over a knot vector of n_ctrl + k + 1 values.  kind 0: B(theta, k, i, t) for every i
 * (BSplines.py:11-29, half-open spans); 1: dB (:31-42); 2: the row r with bspline(theta, t, c, k) = r . c including the
 * end rules theta < 0 / theta >= 1 (:44-51); 3: bspline_derivative (:53-55). */
int ssppcu_bspline_rows(int n_ctrl, int k, const double* knots, const double* theta, int n_theta, int kind, double* out);
/* sspp/BSplines.py:65-106 compute_control_points: via [n][d] at theta_i = i/(n-1) -> ctrl [n][d], knots [n + k + 1] (may be NULL) */
int ssppcu_bspline_fit(const double* via, int n, int d, int k, double* ctrl, double* knots);
/* sspp/CubicPath.py:8-60: the cubic through start, via (u = 1/2), end with zero start velocity, at u[n_u] (clipped to [0,1]):
 * pos / vel / acc [n_u][d], each may be NULL */
int ssppcu_cubic_path(const double* start, const double* via, const double* end, int d, const double* u, int n_u, double* pos, double* vel,
                      double* acc);
/* Batched evaluation on the GPU: out[n_curves][n_theta][d] = rows[n_theta][n_ctrl] . ctrl[n_curves][n_ctrl][d] in FP32 (rows from
 * ssppcu_bspline_rows kind 2 / 3, cast to float).  Host pointers; fails with SSPPCU_ECUDA without a device. */
int ssppcu_bspline_eval_batch(int device, const float* rows, int n_theta, int n_ctrl, const float* ctrl, int n_curves, int d, float* out,
                              char* err, size_t errlen);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments, same arithmetic and
 * bit-identical output as ssppcu_bspline_eval_batch, but this entry point allocates nothing: no cudaMalloc, cudaFree or
 * allocating library routine.  The device staging (rows | control points | output, each region rounded up to 64 floats)
 * lives in a file-scope `static __device__` array of SSPPCU_NOALLOC_MAX_FLOATS floats (4 MiB) that every GPU loading the
 * library holds until the process exits.  A call needing more returns SSPPCU_ELIMIT with the figures in `err` before any
 * CUDA call is made (also: n_theta * n_ctrl * 4 bytes <= 200 KiB, the kernel's shared-memory row table).  Host pointers;
 * synchronous.  All calls on a device share the array, which keeps the previous call's contents (nothing in it has to
 * start at zero: inputs are overwritten, every output element is stored before the read-back); calls overlapping on
 * the same GPU would race on it - inside one process a per-device mutex serialises them. */
#define SSPPCU_NOALLOC_MAX_FLOATS (1 << 20)
int ssppcu_bspline_eval_batch_noalloc(int device, const float* rows, int n_theta, int n_ctrl, const float* ctrl, int n_curves, int d,
                                      float* out, char* err, size_t errlen);

/* ---- planner handle ---- */
void ssppcu_default_config(int mode, SsppPlannerConfig* cfg);
int ssppcu_create(const ssppcu_model_t* m, const SsppPlannerConfig* cfg, ssppcu_t** out, char* err, size_t errlen);
void ssppcu_destroy(ssppcu_t* h);
const char* ssppcu_last_error(const ssppcu_t* h);
int ssppcu_dims(const ssppcu_t* h, SsppDims* out);
int ssppcu_set_stream(ssppcu_t* h, void* cuda_stream);   /* NULL restores the handle's own stream */

/* start/end: [n_queries][dof] host doubles.  Builds the point matrix of every query (JOINT: the
 * interpolating control points of the straight line; TASK: the linear interpolation points). */
int ssppcu_set_queries(ssppcu_t* h, const double* start, const double* end);
/* TASK reset(): mean <- interior points with z floor and clamp, sigma <- clamp(sigma0), best cleared */
int ssppcu_reset_distribution(ssppcu_t* h, double sigma0);
/* explicit distribution, [n_queries][nfd] host doubles; NULL leaves that part unchanged */
int ssppcu_set_distribution(ssppcu_t* h, const double* mean, const double* sigma);
int ssppcu_get_distribution(ssppcu_t* h, double* mean, double* sigma);
/* JOINT convenience: mean <- interior control points, sigma[r][d] <- sigma*limits[d] (sspp.h:116-125) */
int ssppcu_set_joint_noise(ssppcu_t* h, double sigma, const double* limits);

/* device-side snapshot / restore of the sampling state (mean, sigma, forwarded best): asynchronous
 * device-to-device copies on the handle's stream.  restore re-arms a planner for a new plan() of the
 * same query without touching the host (Planner::reset, tsp_planner.h:54-69). */
int ssppcu_save_state(ssppcu_t* h);
int ssppcu_restore_state(ssppcu_t* h);

/* n_iters fused iterations, asynchronous on the handle's stream; Philox counters (seed, iter0+i) */
int ssppcu_iterate(ssppcu_t* h, uint64_t seed, uint32_t iter0, int n_iters);
/* ssppcu_restore_state + ssppcu_iterate in ONE launch: the first iteration samples from the snapshot and re-arms
 * the live state with it (a planner that is reset before every plan(), tsp_planner.h:54-69, needs no separate
 * device-to-device copy) */
int ssppcu_iterate_restored(ssppcu_t* h, uint64_t seed, uint32_t iter0, int n_iters);
/* One cold-start planner step in ONE call and ONE launch: ssppcu_set_queries + ssppcu_set_joint_noise (JOINT: sigma,
 * limits - sspp.h:194-216) or ssppcu_reset_distribution (TASK: sigma = sigma0, limits ignored - tsp_planner.h:72-75 with
 * iterate = false) + ssppcu_iterate.  The query's point matrices and initial state travel in the kernel's parameter block
 * (no host-to-device copy precedes the launch) when they fit (1 KiB: one to a few queries); otherwise the three calls run
 * as they are.  Same results bit for bit either way. */
int ssppcu_plan_step(ssppcu_t* h, const double* start, const double* end, double sigma, const double* limits, uint64_t seed,
                     uint32_t iter0, int n_iters);
int ssppcu_sync(ssppcu_t* h);
/* Diagnostics of the cooperative kernel: with enable != 0 every CTA stamps %globaltimer (ns) at the phase
 * boundaries of each launch's first iteration: [0] start, [1] scored, [2] past grid barrier 1, [3] ranked, [4] past
 * grid barrier 2, [5] updated, [6] rank keys staged, [7] rank counted; ssppcu_iterate_exchange adds [8] shard ordered, [9] keys
 * published, [10] peers' flags seen, [11] merge ranked and stamps [5] when it is finished.  stamps (may be NULL)
 * receives [min(cap_ctas, grid)][12] values of the last launch.  Returns the grid
 * size or a negative error. */
int ssppcu_phase_probe(ssppcu_t* h, int enable, unsigned long long* stamps, int cap_ctas);
/* Diagnostics: the work the scoring path EXECUTED for the candidates of the last iteration (F_executed
 * beside F_nominal).  A separate, untimed kernel re-scores the stored candidates with counters switched on inside
 * the same arithmetic header and returns their totals: out[SSPPCU_WC_*], see the enum below; cap = entries of out.
 * Returns the number of counters or a negative error.  (The reference has no counterpart: it cannot see inside
 * mj_forward.) */
enum { SSPPCU_WC_POINTS = 0, SSPPCU_WC_FK_HINGE_AXIS, SSPPCU_WC_FK_HINGE_GEN, SSPPCU_WC_FK_SLIDE, SSPPCU_WC_FK_FREE4, SSPPCU_WC_FK_R0,
       SSPPCU_WC_FK_ANCHOR, SSPPCU_WC_GEOM_POS, SSPPCU_WC_GEOM_ROT_BOX, SSPPCU_WC_GEOM_ROT_AXIS, SSPPCU_WC_BROAD, SSPPCU_WC_BODY_BROAD, SSPPCU_WC_NP_PLANE,
       SSPPCU_WC_NP_SPHERE_BOX, SSPPCU_WC_NP_BOX_CULL, SSPPCU_WC_NP_CAPS_BOX_EARLY, SSPPCU_WC_NP_CAPS_BOX, SSPPCU_WC_NP_CAPS_NEWTON, SSPPCU_WC_NP_BOX_BOX_FACE,
       SSPPCU_WC_NP_BOX_BOX, SSPPCU_WC_NP_CYL_BOX_FACE, SSPPCU_WC_NP_CYL_BOX, SSPPCU_WC_NP_OTHER, SSPPCU_WC_CONTACTS, SSPPCU_WC_MM_TESTS,
       SSPPCU_WC_MM_NARROW, SSPPCU_WC_SDF_SAMPLES, SSPPCU_WC_COUNT };
int ssppcu_work_counters(ssppcu_t* h, uint64_t* out, int cap);

/* Result mirror (on by default; the environment variable SSPPCU_RESULT_MIRROR=0 sets the default to off): the
 * cooperative kernel copies its result block into the handle's mapped pinned host block itself, behind the grid
 * barrier that makes the values final, so that ssppcu_get_results / _get_best right after a launch are a stream
 * synchronisation and no D2H copy.  The device block is written as before (ssppcu_device_view stays valid).  A caller
 * that keeps the results on the device turns it off and saves the PCIe writes (about 70 KB per launch at 4096
 * candidates). */
int ssppcu_set_result_mirror(ssppcu_t* h, int on);
/* Success rows (off by default): with on != 0 every iteration launch is followed by a compaction kernel that writes the
 * sampled rows [n_free*dof] of the CONVERGED candidates, in candidate order, with their candidate indices, into a mapped
 * pinned host block - what SamplingPathPlanner::plan hands back as ret_success_paths (sspp.h:215-216, 222) without
 * copying the failed candidates' rows.  ssppcu_get_success_rows synchronises and returns views into that block for one
 * query (valid until the next launch): rows [n][n_free*dof], idx [n]. */
int ssppcu_set_success_rows(ssppcu_t* h, int on);
int ssppcu_get_success_rows(ssppcu_t* h, int query, const float** rows, const int32_t** idx, int32_t* n);
/* The K x dof point matrix (control points for JOINT, interpolation points for TASK) of candidate idx of the last
 * iteration of `query`: the query's fixed rows with the candidate's sampled rows written in (SURVEY 8b B-c: makes
 * returning list[Spline] lazy).  The sampled rows are kept in HBM, so this is a 100-byte read, not a regeneration. */
int ssppcu_candidate_ctrl(ssppcu_t* h, int query, uint32_t idx, float* out);

/* host copies of the last iteration; any pointer may be NULL.  Per-candidate arrays are
 * [n_queries][n_cand], per-query arrays [n_queries].  Synchronises the stream. */
int ssppcu_get_results(ssppcu_t* h, float* L, float* Cnf, float* Cwf, float* cost, uint8_t* status,
                       int32_t* n_success, int32_t* n_elite, int32_t* best_idx, float* best_cost,
                       int32_t* any_success);
int ssppcu_get_params(ssppcu_t* h, float* params /* [n_queries][n_cand][nfd] */);
int ssppcu_get_best(ssppcu_t* h, float* best /* [n_queries][nfd] */, int32_t* valid /* [n_queries] */);
/* sorted (cost,index) of one query: up to cap entries, elites first with their CES weights */
int ssppcu_get_elites(ssppcu_t* h, int query, uint32_t* idx, float* cost, float* weight, int cap);
int ssppcu_device_view(ssppcu_t* h, SsppDeviceView* out);

/* Score explicit spline points: pts [n][dof] host floats -> collision cost per point (the
 * contact rule of the handle's mode).  The unit check of FK + narrow phase. */
int ssppcu_eval_points(ssppcu_t* h, const float* pts, int n, float* cost_out);
/* include/Collision.h:51-82 check_collision_point, batched: for each of n configurations (pts [n][dof], float) the contact that
 * penetrates by more than 1 mm (contact dist < -1e-3, the reference's test) with the smallest geom pair: geom1[i] < geom2[i] are
 * model geom ids (names: ssppcu_model_geom_name), dist[i] its signed distance; geom1[i] = geom2[i] = -1 and dist[i] = 0 when the
 * configuration is free of such contacts (the reference returns false).  The reference walks mjData::contact and reports the first
 * hit; MuJoCo orders contacts by geom pair, which the smallest-pair rule restates.  dist may be NULL.  Returns the number of
 * colliding configurations or a negative error.  Primitives collision mode only.  Diagnostic: the planner loop does not use it. */
int ssppcu_check_points(ssppcu_t* h, const float* pts, int n, int32_t* geom1, int32_t* geom2, float* dist);

/* SDF mode: the baked node grid [nz][ny][nx] (host copy) and its geometry; nodes may be NULL to query sizes */
int ssppcu_get_sdf(ssppcu_t* h, float* nodes, int32_t dims[3], double lo[3], double cell[3]);
/* SDF mode, optional refinement step (restates scripts/sdf_gradient/main.py:37-81 on the baked grid): for every row m of
 * `weights` [rows][n_points] (NULL: the handle's collision collocation matrix, rows = check points) the configuration
 * q_m = sum_j weights[m][j] * pts[j] (pts [n_points][dof] host floats), the clearance of the closest sample sphere
 * (dist[m]) and dirs[m] = -clearance * unit gradient of the field there (clearance times the direction towards the
 * obstacle, main.py:60-61); grad[j] = (1/rows) * sum_m dirs[m] * weights[m][j] ([n_points][3], main.py:66-81).  With
 * weights = basis functions and pts = control points this is the reference's formula; with the collocation matrix and
 * the point matrix it is the same derivative in the planner's own parameters.  dist / dirs may be NULL.  Synchronises. */
int ssppcu_sdf_gradient(ssppcu_t* h, const float* pts, const float* weights, int rows, float* grad, float* dist, float* dirs);
/* SDF mode: sample spheres of the moving geoms, [n][5] = centre (body frame) x,y,z, radius, body index; returns n */
int ssppcu_get_spheres(ssppcu_t* h, float* out, int cap);

/* collocation matrices the handle applies (host copies, row-major [rows][n_points]) */
int ssppcu_get_collocation(ssppcu_t* h, float* psi_collision, float* psi_arc);
/* point matrix of every query, [n_queries][n_points][dof] */
int ssppcu_get_base(ssppcu_t* h, float* base);

/* ---- population sharding: each rank scores its shard with ssppcu_iterate, gathers the sorted
 * shard keys of all ranks (8 bytes per candidate) and applies the identical update ---- */
int ssppcu_shard_keys(ssppcu_t* h, unsigned long long** dev_keys, int32_t* n_keys);
/* dev_gathered_keys: n_shards equal-length ascending shards back to back (the all-gather output);
 * asynchronous on the handle's stream */
int ssppcu_merge_update(ssppcu_t* h, const unsigned long long* dev_gathered_keys, int n_keys, int n_shards,
                        uint64_t seed, uint32_t iter);

int ssppcu_get_merged(ssppcu_t* h, uint32_t* idx, float* cost, float* weight, int cap);

/* ---- population sharding with the key exchange fused into the iteration kernel ----
 * Instead of iterate + a collective + merge_update, every rank maps its peers' exchange blocks once;
 * ssppcu_iterate_exchange is then ONE cooperative launch per iteration that scores and ranks the shard,
 * stores the sorted keys into every peer's block over NVLink (8 bytes per candidate and peer), raises
 * a flag on every peer, waits for the peers' flags, merges and applies the update - bit-identical to
 * the unsharded handle on every rank, with no collective call and no host synchronisation.
 *   1. ssppcu_exchange_export: allocates this rank's block (world equal shards, rank in [0, world));
 *      *block_out (optional) = its device address (ranks that share a process), ipc_handle_out (optional) =
 *      64 bytes of cudaIpcMemHandle_t to hand to the other processes (any transport: MPI, torch.distributed ...).
 *   2. ssppcu_exchange_attach_ipc(handles[world][64]) or ssppcu_exchange_attach_ptrs(blocks[world]);
 *      every rank must have exported before any rank attaches.
 *   3. ssppcu_iterate_exchange once per iteration on every rank, in the same order everywhere.
 * A peer that never publishes trips a 2 s device-side timeout: the launch ends, the next call (or
 * ssppcu_exchange_status, which synchronises) returns SSPPCU_ECUDA; it never hangs the GPU. */
int ssppcu_exchange_export(ssppcu_t* h, int world, int rank, void* block_out, void* ipc_handle_out);
int ssppcu_exchange_attach_ipc(ssppcu_t* h, const void* handles);
int ssppcu_exchange_attach_ptrs(ssppcu_t* h, void* const* blocks);
int ssppcu_iterate_exchange(ssppcu_t* h, uint64_t seed, uint32_t iter);
int ssppcu_exchange_status(ssppcu_t* h);

/* ---- host-side spline algebra the handle applies (no device needed; FP64) ----
 * Restates Eigen's KnotAveraging + SplineFitting::Interpolate + Spline::operator() as the
 * reference calls them (include/sspp.h:95,137; include/sspp/tsp_path_model.h:25-28,39-42).
 *   knots [n_points+degree+1], Ainv [n_points][n_points]: control points = Ainv * interpolation points
 *   psi [n_grid][n_points]: via_form 0 -> Basis(grid) (acts on control points),
 *                           via_form 1 -> Basis(grid) * Ainv (acts on interpolation points) */
int ssppcu_spline_setup(int n_points, int degree, double* knots, double* Ainv);
int ssppcu_collocation_matrix(int n_points, int degree, const double* grid, int n_grid, int via_form, double* psi);

/* Measured FP32 FMA throughput of the device (register-resident FMA chains, all SMs), the
 * denominator of the roofline this path is bound by (it moves ~16 B per candidate). */
int ssppcu_fp32_peak(int device, double* tflops, double* sm_clock_mhz);

int ssppcu_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SSPP_CUDA_H */
