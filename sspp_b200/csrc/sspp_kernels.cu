// sspp_kernels.cu — the fused scoring iteration for sm_100a: one launch per planner iteration
// (or per batch of iterations).
//
//   K1 perturb     Philox4x32-10 + Box-Muller noise on the sampled rows      (sspp.h:114-130, tsp_sampler.h:12-51)
//   K2 spline      q_m = Psi[m,:] . X   (banded rows, X: K x D point matrix in shared memory)  (Eigen::Spline::operator(), sspp.h:137)
//   K3 kinematics  link frames from q, per thread                            (mj_kinematics via sspp.h:139-143, utility.h:149-175)
//   K4 collision   moving geoms against the scene primitives                 (mj_collision via sspp.h:144, Collision.h:84-103)
//   K5 reduce      arc length, collision and penalty sums per candidate      (sspp.h:152-169, tsp_evaluator.h:17-32)
//   K6 elites      order the (cost,index) keys, CES weights                  (sspp.h:171-192, tsp_elites.h:13-32)
//   K7 update      weighted mean / EMA variance / sigma adaptation           (tsp_distribution.h:31-83)
//
// Two shapes of the same kernel:
//   coop = 1  persistent cooperative grid (as many CTAs as are co-resident, looping over candidate
//             tiles).  Phase A scores the tiles; after a grid barrier phase B ranks every key against
//             all keys of its query (grid-wide counting rank: each CTA stages the query's keys in
//             shared memory chunk by chunk and counts, for the keys it owns, how many are smaller),
//             which yields the sorted order without a serial sort; after a second barrier phase C
//             updates the distribution with one WARP per (query, distribution element).
//   coop = 0  one CTA per query (the ICRA shape: 17 candidates x 40 points): bitonic sort in shared
//             memory and the update by the same CTA; all iterations of a plan run inside one launch.
// A CTA tile owns `cand_per_cta` consecutive candidates of one query; its threads stride over the
// flattened (candidate, collocation point) items, so a lane carries one full FK + collision
// evaluation whatever M is.  Per-point results go to shared memory; one warp per candidate reduces
// them in a fixed order (lane-strided partial sums, xor butterfly) that the oracles restate.
#include "sspp_kernels.h"

#include <cooperative_groups.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <mutex>

namespace cg = cooperative_groups;

namespace sspp {
using namespace ssppm;

namespace {

struct SmemXF {     // link frames staged in shared memory, [body*12 + k][thread]
  float* base; int T; int tid;
  __device__ __forceinline__ void put(int b, int k, float v) { base[(b * 12 + k) * T + tid] = v; }
  __device__ __forceinline__ float get(int b, int k) const { return base[(b * 12 + k) * T + tid]; }
};
struct SmemQ {      // planned coordinates of the point a thread is scoring, [d][thread]
  const float* base; int T; int tid;
  __device__ __forceinline__ float get(int d) const { return base[d * T + tid]; }
};

__device__ __forceinline__ float warp_butterfly(float acc) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, m);
  return acc;
}
__device__ __forceinline__ float warp_strided_sum(const float* v, int n, int lane) {
  float acc = 0.0f;
  for (int i = lane; i < n; i += 32) acc += v[i];
  return warp_butterfly(acc);
}

__device__ __forceinline__ float key_cost(unsigned long long key) {
  unsigned ob = (unsigned)(key >> 32);
  ob = (ob & 0x80000000u) ? (ob & 0x7FFFFFFFu) : ~ob;
  return __uint_as_float(ob);
}
__device__ __forceinline__ bool key_success(unsigned long long key) { return (unsigned)(key >> 32) != 0xFFFFFFFFu; }

// shared memory carve-up of the scoring phase (all 4-byte words; every segment starts on a 16-byte boundary).
// Banded collocation rows have a pitch of 4 floats when the band fits (degree <= 3), control-point rows a pitch
// of dp = D rounded up to 4 with zero padding, so a row is one / dp/4 128-bit reads.
__host__ __device__ __forceinline__ int row_pitch(int w) { return w <= 4 ? 4 : (w | 1); }
__host__ __device__ __forceinline__ int up4(int n) { return (n + 3) & ~3; }
struct Smem {
  unsigned long long* keys;   // coop == 0 only: n_pad sort keys
  float *psi_col, *psi_arc, *X, *cc, *cw, *seg, *qs, *xf;
  int *first_col, *first_arc;
  unsigned a_psi_col, a_psi_arc, a_X;   // shared-window byte addresses of the same segments
};
__device__ __forceinline__ Smem carve(unsigned char* raw, const ScoreConfig& c, const IterParams& P, int T) {
  Smem s;
  s.keys = reinterpret_cast<unsigned long long*>(raw);
  float* const sm0 = reinterpret_cast<float*>(raw);
  float* sm = sm0 + (P.coop ? 0 : 2 * P.n_pad);
  const int CT = P.cand_per_cta, wc = row_pitch(P.w_col), wa = row_pitch(P.w_arc), dp = up4(c.D);
  s.psi_col = sm;   sm += up4(c.Mc * wc);
  s.psi_arc = sm;   sm += c.arc_same ? 0 : up4(c.Ma * wa);
  s.first_col = reinterpret_cast<int*>(sm); sm += up4(c.Mc);
  s.first_arc = reinterpret_cast<int*>(sm); sm += c.arc_same ? 0 : up4(c.Ma);
  s.X = sm;         sm += CT * c.K * dp;
  s.cc = sm;        sm += CT * c.Mc;     // collision cost per point
  s.cw = sm;        sm += CT * c.Mc;     // collision cost + penalty per point
  s.seg = sm;       sm += CT * P.mq;     // segment lengths
  s.qs = sm;        sm += c.D * T;       // coordinates of the point in flight
  s.xf = sm;                              // link frames (only when !chain)
  const unsigned a0 = (unsigned)__cvta_generic_to_shared(raw);
  s.a_psi_col = a0 + (unsigned)((s.psi_col - sm0) * 4); s.a_psi_arc = a0 + (unsigned)((s.psi_arc - sm0) * 4);
  s.a_X = a0 + (unsigned)((s.X - sm0) * 4);
  return s;
}
// Scene tables staged at the front of dynamic shared memory (they stay there for the whole launch): every
// record is a multiple of 16 bytes, copied as 128-bit words.  The staged view addresses them by 32-bit
// offsets from the dynamic shared-memory base, so every access compiles to an LDS (warp-uniform index ->
// one broadcast wavefront) instead of an indexed constant load through the ADU or a generic load.
extern __shared__ uint4 smem_all4[];                    // the dynamic shared memory as 16-byte words
#define smem_all (reinterpret_cast<unsigned char*>(smem_all4))
// A staged table is addressed by its 32-bit shared-window byte address, formed ONCE per thread; every record
// read is one explicit ld.shared.v4 (the compiler's own shared addressing re-derives the window base from
// SR_CgaCtaId at every use inside these loops).
template <class T>
struct SmemTab { unsigned addr; };
struct SmemModel : ssppm::DevModelHead {
  SmemTab<DevBody> body; SmemTab<DevMGeom> mg; SmemTab<DevSGeom> sg; SmemTab<DevPair> pair; SmemTab<DevMMPair> mm; SmemTab<DevSphere> sph;
};
__device__ __forceinline__ uint4 lds128(unsigned addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
// overloads of the record accessors of sspp_math.h
template <class T>
__device__ __forceinline__ ssppm::f4 rec_f4(const SmemTab<T>& t, int i, int w) {
  const uint4 v = lds128(t.addr + (unsigned)i * (unsigned)sizeof(T) + (unsigned)(w << 2));
  ssppm::f4 r; r.x = __uint_as_float(v.x); r.y = __uint_as_float(v.y); r.z = __uint_as_float(v.z); r.w = __uint_as_float(v.w);
  return r;
}
template <class T>
__device__ __forceinline__ ssppm::i4 rec_i4(const SmemTab<T>& t, int i, int w) {
  const uint4 v = lds128(t.addr + (unsigned)i * (unsigned)sizeof(T) + (unsigned)(w << 2));
  ssppm::i4 r; r.x = (int)v.x; r.y = (int)v.y; r.z = (int)v.z; r.w = (int)v.w;
  return r;
}
__device__ __forceinline__ unsigned stage_table(const void* src, unsigned bytes, unsigned& at16, int tid, int T) {
  const uint4* s4 = reinterpret_cast<const uint4*>(src);
  for (int i = tid; i < (int)(bytes >> 4); i += T) smem_all4[at16 + i] = __ldg(s4 + i);
  const unsigned w16 = at16;
  at16 += bytes >> 4;
  return w16 << 4;
}
__device__ __forceinline__ SmemModel stage_model(const ModelView& G, int tid, int T) {
  SmemModel S;
  static_cast<ssppm::DevModelHead&>(S) = G;
  const unsigned base = (unsigned)__cvta_generic_to_shared(smem_all4);
  unsigned at = 0;
  S.body.addr = base + stage_table(G.body, G.n_mbody * (unsigned)sizeof(DevBody), at, tid, T);
  S.mg.addr = S.sg.addr = S.pair.addr = S.mm.addr = S.sph.addr = base;
  if (G.sdf_mode) {
    S.sph.addr = base + stage_table(G.sph, G.n_sphere * (unsigned)sizeof(DevSphere), at, tid, T);
  } else {
    S.mg.addr = base + stage_table(G.mg, G.n_mgeom * (unsigned)sizeof(DevMGeom), at, tid, T);
    S.sg.addr = base + stage_table(G.sg, G.n_sgeom * (unsigned)sizeof(DevSGeom), at, tid, T);
    S.pair.addr = base + stage_table(G.pair, (G.n_pair + G.n_bpair) * (unsigned)sizeof(DevPair), at, tid, T);
    S.mm.addr = base + stage_table(G.mm, G.n_mmpair * (unsigned)sizeof(DevMMPair), at, tid, T);
  }
  return S;
}

__device__ __forceinline__ void load_basis(const Smem& s, const ScoreConfig& c, const IterParams& P, int tid, int T) {
  const int wc = row_pitch(P.w_col), wa = row_pitch(P.w_arc);
  for (int e = tid; e < c.Mc * wc; e += T) { const int r = e / wc, k = e - r * wc; s.psi_col[e] = (k < P.w_col) ? P.psi_col[r * P.w_col + k] : 0.0f; }
  for (int e = tid; e < c.Mc; e += T) s.first_col[e] = P.first_col[e];
  if (!c.arc_same) {
    for (int e = tid; e < c.Ma * wa; e += T) { const int r = e / wa, k = e - r * wa; s.psi_arc[e] = (k < P.w_arc) ? P.psi_arc[r * P.w_arc + k] : 0.0f; }
    for (int e = tid; e < c.Ma; e += T) s.first_arc[e] = P.first_arc[e];
  }
}

// One spline point from 128-bit shared reads: q[d] = sum_k row[k] * X[first + k][d], k ascending from +0 — the
// chain of ssppm::spline_coord, so the same bits.  W: band width (<= 4, row pitch 4); NV: 4-float words per
// control-point row (pitch 4 NV, zero padded).
template <int W, int NV>
__device__ __forceinline__ void spline_point(unsigned a_rows, int m, int first, unsigned a_X, float* q) {
  const uint4 rv = lds128(a_rows + (unsigned)m * 16u);
  const float rc[4] = {__uint_as_float(rv.x), __uint_as_float(rv.y), __uint_as_float(rv.z), __uint_as_float(rv.w)};
#pragma unroll
  for (int d = 0; d < 4 * NV; ++d) q[d] = 0.0f;
#pragma unroll
  for (int k = 0; k < W; ++k)
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const uint4 x = lds128(a_X + (unsigned)((first + k) * NV + v) * 16u);
      q[4 * v + 0] = fmaf(rc[k], __uint_as_float(x.x), q[4 * v + 0]);
      q[4 * v + 1] = fmaf(rc[k], __uint_as_float(x.y), q[4 * v + 1]);
      q[4 * v + 2] = fmaf(rc[k], __uint_as_float(x.z), q[4 * v + 2]);
      q[4 * v + 3] = fmaf(rc[k], __uint_as_float(x.w), q[4 * v + 3]);
    }
}
// dispatch on the (band width, row words) of the problem; false: shape not covered, use the generic loop
__device__ __forceinline__ bool spline_point_any(int W, int NV, unsigned a_rows, int m, int first, unsigned a_X, float* q) {
  if (NV == 2) {
    if (W == 4) { spline_point<4, 2>(a_rows, m, first, a_X, q); return true; }
    if (W == 3) { spline_point<3, 2>(a_rows, m, first, a_X, q); return true; }
  } else if (NV == 1) {
    if (W == 3) { spline_point<3, 1>(a_rows, m, first, a_X, q); return true; }
    if (W == 4) { spline_point<4, 1>(a_rows, m, first, a_X, q); return true; }
  }
  return false;
}

// the snapshot a restoring launch starts from: device memory (ssppcu_iterate_restored) or the parameter block (ssppcu_plan_step)
__device__ __forceinline__ const float* snapshot_of(const IterParams& P) {
  return P.inline_words ? P.inline_blk + (size_t)P.n_query * P.cfg.K * P.cfg.D : P.state_in;
}
__device__ __forceinline__ bool has_snapshot(const IterParams& P) { return P.state_in != nullptr || P.inline_words != 0; }

// Phase A for one tile: candidates [c0, c0 + nc) of `query`.  Ends with the per-candidate outputs
// and raw keys in global memory; no trailing barrier.
__device__ __forceinline__ void score_tile(const SmemModel& M, const IterParams& P, const Smem& s, int query, int c0, int nc,
                                           uint32_t iter, bool from_saved) {
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = T >> 5;
  const int D = c.D, K = c.K, Mc = c.Mc, Ma = c.Ma, Mq = P.mq;
  const int nfd = c.n_free * D;
  const int Marc = c.arc_same ? Mc : Ma;
  const int wc = row_pitch(P.w_col), wa = row_pitch(P.w_arc), dp = up4(D), NV = dp >> 2;
  const bool fast_col = NV <= 2 && (P.w_col == 3 || P.w_col == 4);   // shapes spline_point_any covers
  const bool fast_arc = NV <= 2 && (P.w_arc == 3 || P.w_arc == 4);
  // sampling state: the live block, or (first iteration of a restoring launch) the snapshot it is being re-armed from
  const float* st_base = from_saved ? snapshot_of(P) : P.mean;
  const size_t st_q = (size_t)P.n_query * nfd;
  const float* mean = st_base + (size_t)query * nfd;
  const float* sigma = st_base + st_q + (size_t)query * nfd;
  const float* best = st_base + 2 * st_q + (size_t)query * nfd;
  const float* base = ((from_saved && P.inline_words) ? P.inline_blk : P.base) + (size_t)query * K * D;
  const size_t qn = (size_t)query * c.n_cand;
  SmemXF xf{s.xf, T, tid};
  SmemQ qa{s.qs, T, tid};
  const int best_ok = c.include_best ? reinterpret_cast<const int*>(st_base + 3 * st_q)[query] : 0;
  // ---- K1: candidate point matrices (rows padded to dp floats with zeros) ----
  for (int e = tid; e < nc * K * dp; e += T) { const int r = e / dp, d = e - r * dp; s.X[e] = (d < D) ? base[(r % K) * D + d] : 0.0f; }
  __syncthreads();
  for (int e = tid; e < nc * nfd; e += T) {
    const int cl = e / nfd, t = e - cl * nfd;
    const int cand = c0 + cl, gcand = P.cand_offset + cand;
    const int kind = candidate_kind(c, gcand, best_ok);
    const float v = candidate_param(c, P.seed, iter, (uint32_t)query, gcand, kind, t, mean, sigma, best);
    const int fr = t / D;
    s.X[(cl * K + c.free_begin + fr) * dp + (t - fr * D)] = v;
    P.params[(qn + cand) * nfd + t] = v;
  }
  __syncthreads();
  // ---- K2-K4: spline point, kinematics, collision per (candidate, point) item ----
  // When the arc-length grid is the collision grid (arc_same) the thread that owns a segment's end point
  // evaluates the start point a second time instead of reading it from a staging buffer: no per-point buffer,
  // and the same bits as any other evaluation of that point.
  for (int w = tid; w < nc * Mc; w += T) {
    const int cl = w / Mc, m = w - cl * Mc;
    const int first = s.first_col[m];
    float s2 = 0.0f;
    float q[8];
    if (fast_col && spline_point_any(P.w_col, NV, s.a_psi_col, m, first, s.a_X + (unsigned)(cl * K * dp) * 4u, q)) {
#pragma unroll
      for (int d = 0; d < 8; ++d) if (d < D) s.qs[d * T + tid] = q[d];
      if (c.arc_same && m > 0) {
        float q0[8];
        spline_point_any(P.w_col, NV, s.a_psi_col, m - 1, s.first_col[m - 1], s.a_X + (unsigned)(cl * K * dp) * 4u, q0);
#pragma unroll
        for (int d = 0; d < 8; ++d) if (d < D) { const float df = q[d] - q0[d]; s2 = fmaf(df, df, s2); }
      }
    } else {
      const float* row = s.psi_col + m * wc;
      const float* X = s.X + cl * K * dp;
      for (int d = 0; d < D; ++d) {
        const float acc = spline_coord(row, first, P.w_col, X, dp, d);
        s.qs[d * T + tid] = acc;
        if (c.arc_same && m > 0) {
          const float df = acc - spline_coord(row - wc, s.first_col[m - 1], P.w_col, X, dp, d);
          s2 = fmaf(df, df, s2);
        }
      }
    }
    if (c.arc_same && m > 0) s.seg[cl * Mq + m - 1] = sqrtf(s2);
    float cc = 0.0f, cw = 0.0f;
    if (m >= c.col_first) {
      cc = point_collision_cost(M, qa, xf);
      cw = cc + point_penalty(c, qa);
    }
    s.cc[cl * Mc + m] = cc;
    s.cw[cl * Mc + m] = cw;
  }
  if (!c.arc_same) {
    if (fast_arc) {
      // arc-length grid: a warp round takes 32 consecutive points of the flattened (candidate, arc point) list and
      // advances by 31, so every lane but 0 finds the start point of its segment in the lane below (shuffle): each
      // point is evaluated once
      const int total = nc * Ma;
      for (int pb = warp * 31; pb < total - 1; pb += nwarp * 31) {
        const int pt = min(pb + lane, total - 1);
        const int cl = pt / Ma, m = pt - cl * Ma;
        float q[8];
        spline_point_any(P.w_arc, NV, s.a_psi_arc, m, s.first_arc[m], s.a_X + (unsigned)(cl * K * dp) * 4u, q);
        float s2 = 0.0f;
#pragma unroll
        for (int d = 0; d < 8; ++d) {
          const float prev = __shfl_up_sync(0xFFFFFFFFu, q[d], 1);
          if (d < D) { const float df = q[d] - prev; s2 = fmaf(df, df, s2); }
        }
        if (lane > 0 && pb + lane < total && m > 0) s.seg[cl * Mq + m - 1] = sqrtf(s2);
      }
    } else {
      for (int w = tid; w < nc * (Ma - 1); w += T) {
        const int cl = w / (Ma - 1), m = 1 + (w - cl * (Ma - 1));
        const float* row = s.psi_arc + m * wa;
        const float* X = s.X + cl * K * dp;
        const int f1 = s.first_arc[m], f0 = s.first_arc[m - 1];
        float s2 = 0.0f;
        for (int d = 0; d < D; ++d) {
          const float df = spline_coord(row, f1, P.w_arc, X, dp, d) - spline_coord(row - wa, f0, P.w_arc, X, dp, d);
          s2 = fmaf(df, df, s2);
        }
        s.seg[cl * Mq + m - 1] = sqrtf(s2);
      }
    }
  }
  __syncthreads();
  for (int cl = warp; cl < nc; cl += nwarp) {
    const float L = warp_strided_sum(s.seg + cl * Mq, Marc - 1, lane);
    const float Cnf = warp_strided_sum(s.cc + cl * Mc + c.col_first, Mc - c.col_first, lane);
    const float Cwf = warp_strided_sum(s.cw + cl * Mc + c.col_first, Mc - c.col_first, lane);
    if (lane == 0) {
      const int cand = c0 + cl;
      const int kind = candidate_kind(c, P.cand_offset + cand, best_ok);
      const int st = (kind == CK_ABSENT) ? ST_ABSENT : ((Cnf == 0.0f) ? ST_CONVERGED : ST_FAILED);
      const float cost = fmaf(c.w_collision, Cwf, L);
      P.L[qn + cand] = L; P.Cnf[qn + cand] = Cnf; P.Cwf[qn + cand] = Cwf; P.cost[qn + cand] = cost;
      P.status[qn + cand] = (uint8_t)st;
      P.keys_raw[(size_t)query * P.n_pad + cand] = sort_key(cost, st, (uint32_t)(P.cand_offset + cand));
      // the ranking phase of OTHER CTAs reads these keys with a bulk asynchronous copy (async proxy): order this generic-proxy
      // store before it across the proxies; the grid barrier then orders it across the CTAs
      asm volatile("fence.proxy.async.global;" ::: "memory");
    }
  }
}

// ------------------------------------------------------------------------------------------
// K6/K7 arithmetic for one distribution element t of one query, executed by ONE WARP.
//   keys: sorted keys of the query; get_param(candidate index, t) -> sampled value.
// CES log-weights w_j = (ln(k + 1/2) - ln(j + 1)) / sum (tsp_elites.h:24-32) are recomputed on the
// fly: lane l owns j = l, l + 32, ... in every pass, partial sums meet in an xor butterfly.
// ------------------------------------------------------------------------------------------
// Wide sums.  The elite reductions are DEFINED over 256 virtual lanes = 8 groups of 32: virtual lane v sums its
// terms j = v, v + 256, ... in ascending order; the 32 partials of a group meet in an xor butterfly; the group
// sums are added in ascending group order, groups without terms (32 g >= k) skipped (adding +0 is exact).  One
// warp can walk the groups serially (small k: the one-CTA-per-query kernel) or the warps of a CTA can take one
// group each (coop phase C, merge kernel): the same bits either way, and what the replica oracle restates.
template <class Term>
__device__ __forceinline__ float group_partial(int g, int k, int lane, Term term) {
  float acc = 0.0f;
  for (int j = 32 * g + lane; j < k; j += 256) acc = term(j, acc);
  return warp_butterfly(acc);
}
template <class Term>
__device__ __forceinline__ float wide_sum_warp(int k, int lane, Term term) {
  float tot = group_partial(0, k, lane, term);
  for (int g = 1; g < 8 && 32 * g < k; ++g) tot += group_partial(g, k, lane, term);
  return tot;
}
// all threads of the CTA call this; s_part: 8 floats of shared scratch; two barriers inside
template <class Term>
__device__ __forceinline__ float wide_sum_cta(int k, Term term, float* s_part) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
  for (int g = warp; g < 8; g += nwarp)
    if (g == 0 || 32 * g < k) { const float pg = group_partial(g, k, lane, term); if (lane == 0) s_part[g] = pg; }
  __syncthreads();
  float tot = s_part[0];
  for (int g = 1; g < 8 && 32 * g < k; ++g) tot += s_part[g];
  __syncthreads();
  return tot;
}
struct WarpSum { int lane; template <class Term> __device__ __forceinline__ float operator()(int k, Term term) const { return wide_sum_warp(k, lane, term); } };
struct CtaSum { float* s_part; template <class Term> __device__ __forceinline__ float operator()(int k, Term term) const { return wide_sum_cta(k, term, s_part); } };

template <class Sum>
__device__ __forceinline__ float ces_weight_sum(int k, float lk, Sum sum) {
  return sum(k, [&](int j, float acc) { return acc + ces_log_weight(k, lk, j); });
}

// K7 for one distribution element t of one query.  value_of(j): the element's value in the j-th best candidate;
// weight_of(j): the CES weight (lk - ln(j + 1)) / sumw of rank j - recomputed on the fly (InlineWeight) or read from a
// table the caller filled with exactly that expression (the same bits either way).
// `sum` is WarpSum (one warp executes this) or CtaSum (the whole CTA does); `writer` is true in the thread that
// stores the result.
struct InlineWeight {
  int k; float lk, sumw;
  __device__ __forceinline__ float operator()(int j) const { return ces_log_weight(k, lk, j) / sumw; }
};
template <class ValueOf, class WeightOf, class Sum>
__device__ __forceinline__ void update_element_w(const ScoreConfig& c, int k, int t, float* mean, float* sigma,
                                                 ValueOf value_of, WeightOf weight_of, Sum sum, bool writer) {
  const int d = t % c.D;
  const float em = sum(k, [&](int j, float acc) { return fmaf(weight_of(j), value_of(j), acc); });
  const float m0 = mean[t];
  float nm = fmaf(c.mean_lr, em - m0, m0);
  if (c.mode == MODE_TASK && d == 2) nm = fmaxf(nm, c.dist_zmin);
  nm = fminf(fmaxf(nm, c.lo[d]), c.hi[d]);
  const bool wrap = (c.mode == MODE_TASK && d == 3 && c.lo[3] != c.hi[3]);
  const float var = sum(k, [&](int j, float acc) {
    const float v = value_of(j);
    const float df = wrap ? wrap_angle_diff(v, nm, c.lo[3], c.hi[3]) : v - nm;
    return fmaf(weight_of(j), df * df, acc);
  });
  const float s0 = sigma[t];
  const float blend = fmaf(c.var_beta, var, (1.0f - c.var_beta) * (s0 * s0));
  float sg = clamp_sigma(c, sqrtf(blend));
  sg = clamp_sigma(c, sg * c.dec);               // adapt(true)
  if (writer) { mean[t] = nm; sigma[t] = sg; }
}
template <class ValueOf, class Sum>
__device__ __forceinline__ void update_element(const ScoreConfig& c, int k, float lk, float sumw, int t, float* mean, float* sigma,
                                               ValueOf value_of, Sum sum, bool writer) {
  update_element_w(c, k, t, mean, sigma, value_of, InlineWeight{k, lk, sumw}, sum, writer);
}

struct StoredParams {
  const float* params; int nfd; unsigned offset;
  __device__ __forceinline__ float operator()(unsigned idx, int t) const { return params[(size_t)(idx - offset) * nfd + t]; }
};

// per-query scalars written once the order is known (thread-level; caller picks the thread)
__device__ __forceinline__ void write_best(const IterParams& P, int query, unsigned long long k0) {
  if (key_success(k0)) {
    P.best_idx[query] = (int)(unsigned)(k0 & 0xFFFFFFFFu);
    P.best_cost[query] = key_cost(k0);
    P.any_success[query] = 1;
  } else { P.best_idx[query] = -1; P.best_cost[query] = __int_as_float(0x7F800000); }
}

// ascending bitonic sort of n_pad (power of two) 64-bit keys by the whole CTA
__device__ void block_bitonic_sort(unsigned long long* keys, int n_pad) {
  const int tid = threadIdx.x, T = blockDim.x;
  for (int k2 = 2; k2 <= n_pad; k2 <<= 1) {
    for (int j = k2 >> 1; j > 0; j >>= 1) {
      for (int i = tid; i < n_pad; i += T) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long a = keys[i], b = keys[ixj];
          const bool up = (i & k2) == 0;
          if ((a > b) == up) { keys[i] = b; keys[ixj] = a; }
        }
      }
      __syncthreads();
    }
  }
}

// K6 + K7 of one query by one CTA (coop == 0, and the merge kernel): sorts `keys` in place
// (shared or global memory), then updates; get_param as in update_element.
template <class ParamGetter>
__device__ void cta_elites_and_update(const ScoreConfig& c, unsigned long long* keys, int n_pad, float* elite_w, float* mean,
                                      float* sigma, float* best, int* best_valid, int* n_success_out, int* n_elite_out,
                                      int* best_idx_out, float* best_cost_out, int* any_success, ParamGetter get_param,
                                      int* s_int) {
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = T >> 5;
  block_bitonic_sort(keys, n_pad);
  if (tid == 0) s_int[0] = 0;
  __syncthreads();
  int cnt = 0;
  for (int i = tid; i < n_pad; i += T) cnt += key_success(keys[i]) ? 1 : 0;
  cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
  if (lane == 0 && cnt) atomicAdd(&s_int[0], cnt);
  __syncthreads();
  const int n_succ = s_int[0];
  const int k = elite_count(n_succ, c.elite_frac);
  const int nfd = c.n_free * c.D;
  if (tid == 0) {
    *n_success_out = n_succ; *n_elite_out = k;
    const unsigned long long k0 = keys[0];
    if (n_succ > 0) { *best_idx_out = (int)(unsigned)(k0 & 0xFFFFFFFFu); *best_cost_out = key_cost(k0); *any_success = 1; }
    else { *best_idx_out = -1; *best_cost_out = __int_as_float(0x7F800000); }
  }
  if (n_succ == 0) {
    if (c.do_update) for (int t = tid; t < nfd; t += T) sigma[t] = clamp_sigma(c, sigma[t] * c.inc);   // adapt(false)
    __syncthreads();
    return;
  }
  const float lk = log_poly((float)k + 0.5f);
  const float sumw = ces_weight_sum(k, lk, WarpSum{lane});   // every warp computes the same number
  for (int j = tid; j < k; j += T) elite_w[j] = ces_log_weight(k, lk, j) / sumw;
  // the best candidate is read before any element is updated (its rows are not state, but keep the order of v1)
  const unsigned bi = (unsigned)(keys[0] & 0xFFFFFFFFu);
  if (c.do_update)
    for (int t = warp; t < nfd; t += nwarp)
      update_element(c, k, lk, sumw, t, mean, sigma, [&](int j) { return get_param((unsigned)(keys[j] & 0xFFFFFFFFu), t); }, WarpSum{lane}, lane == 0);
  for (int t = tid; t < nfd; t += T) best[t] = get_param(bi, t);
  if (tid == 0) *best_valid = 1;
  __syncthreads();
}

// ---- coop phase B: grid-wide counting rank -------------------------------------------------
// Items are the (query, candidate) keys, split evenly over the CTAs.  For its keys of one query a CTA stages the
// query's raw keys in shared memory, `chunk` at a time, with ONE bulk asynchronous copy (TMA, completion on an
// mbarrier), and ranks its own keys:
//   successful key: rank = number of smaller keys, counted against the staged keys — every warp scans its own slice
//                   of the chunk for up to 8 owned keys at a time (counters in registers);
//   failed key:     all failed keys share the top 32 bits and order by candidate index, so
//                   rank = n_success + failed keys at lower indices, two counters of the same scan.
// rank = position in the sorted order (keys are unique: they embed the candidate index).
__device__ __forceinline__ void tma_stage(unsigned long long* dst, const unsigned long long* src, int n, unsigned long long* mbar,
                                          unsigned& phase) {
  // all threads call; dst/src 16-byte aligned; n keys (8 bytes each)
  const unsigned bytes = (unsigned)(n & ~1) * 8u;
  if (threadIdx.x == 0) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // earlier generic-proxy accesses of dst are done
    if (bytes) {
      const unsigned d = (unsigned)__cvta_generic_to_shared(dst), m = (unsigned)__cvta_generic_to_shared(mbar);
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(m), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   :: "r"(d), "l"(src), "r"(bytes), "r"(m) : "memory");
    }
    if (n & 1) dst[n - 1] = __ldcg(src + n - 1);
  }
  if (bytes) {
    const unsigned m = (unsigned)__cvta_generic_to_shared(mbar);
    unsigned done = 0;
    while (!done) {
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                   : "=r"(done) : "r"(m), "r"(phase) : "memory");
    }
    phase ^= 1u;
  }
  __syncthreads();                                                   // the odd tail element, and a common point to go on from
}

template <class Stamp>
__device__ void rank_phase(const IterParams& P, unsigned char* smem_raw, unsigned long long* mbar, unsigned& mbar_phase, Stamp stamp) {
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = T >> 5;
  const int N = c.n_cand, nfd = c.n_free * c.D;
  const long long total = (long long)P.n_query * N;
  const long long per = (total + gridDim.x - 1) / gridDim.x;
  const long long lo_item = per * blockIdx.x, hi_item = min(total, lo_item + per);
  unsigned long long* s_chunk = reinterpret_cast<unsigned long long*>(smem_raw);
  unsigned long long* s_mine = s_chunk + P.rank_chunk;
  int* s_cnt = reinterpret_cast<int*>(s_mine + per);       // smaller keys found (successful own keys)
  int* s_oidx = s_cnt + per;                                // own successful keys, compacted: local index
  __shared__ int s_nos, s_fb, s_tf;                         // own successes; failed keys before this CTA's first key; failed keys
  for (long long item = lo_item; item < hi_item;) {
    const int query = (int)(item / N);
    const int i0 = (int)(item - (long long)query * N);
    const int nmine = (int)min((long long)(N - i0), hi_item - item);
    const unsigned long long* raw = P.keys_raw + (size_t)query * P.n_pad;
    unsigned long long* sorted = P.keys + (size_t)query * P.n_pad;
    __syncthreads();
    if (tid == 0) { s_nos = 0; s_fb = 0; s_tf = 0; }
    for (int a = tid; a < nmine; a += T) { s_mine[a] = __ldcg(raw + i0 + a); s_cnt[a] = 0; }
    __syncthreads();
    for (int a0 = warp * 32; a0 < nmine; a0 += nwarp * 32) {   // compact the successful own keys (any order)
      const int a = a0 + lane;
      const bool ok = a < nmine && key_success(s_mine[a]);
      const unsigned b = __ballot_sync(0xFFFFFFFFu, ok);
      int base = 0;
      if (lane == 0 && b) base = atomicAdd(&s_nos, __popc(b));
      base = __shfl_sync(0xFFFFFFFFu, base, 0);
      if (ok) s_oidx[base + __popc(b & ((1u << lane) - 1u))] = a;
    }
    for (int cs = 0; cs < N; cs += P.rank_chunk) {          // rank_chunk is even and the rows are 16-byte aligned
      const int cn = min(P.rank_chunk, N - cs);
      __syncthreads();                                       // the previous chunk has been consumed; s_nos / s_oidx are final
      tma_stage(s_chunk, raw + cs, cn, mbar, mbar_phase);
      stamp(6);                                              // chunk staged
      const int n_os = s_nos;
      const int slice = (cn + nwarp - 1) / nwarp;
      const int j0 = warp * slice, j1 = min(cn, j0 + slice);
      for (int u0 = 0; u0 < max(n_os, 1); u0 += 8) {
        unsigned long long key[8];
        int cnt[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) { key[u] = (u0 + u < n_os) ? s_mine[s_oidx[u0 + u]] : 0ull; cnt[u] = 0; }
        int fb = 0, tf = 0;
        for (int j = j0 + lane; j < j1; j += 32) {
          const unsigned long long x = s_chunk[j];
#pragma unroll
          for (int u = 0; u < 8; ++u)      // cnt += (x < key): compare, compare-extended, predicated add
            asm("{ .reg .pred p; setp.lt.u64 p, %1, %2; @p add.s32 %0, %0, 1; }" : "+r"(cnt[u]) : "l"(x), "l"(key[u]));
          if (u0 == 0) { const bool failed = !key_success(x); tf += failed; fb += failed && (cs + j < i0); }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int tot = __reduce_add_sync(0xFFFFFFFFu, cnt[u]);
          if (lane == 0 && tot && u0 + u < n_os) atomicAdd(&s_cnt[s_oidx[u0 + u]], tot);
        }
        if (u0 == 0) {
          fb = __reduce_add_sync(0xFFFFFFFFu, fb); tf = __reduce_add_sync(0xFFFFFFFFu, tf);
          if (lane == 0) { if (fb) atomicAdd(&s_fb, fb); if (tf) atomicAdd(&s_tf, tf); }
        }
      }
    }
    __syncthreads();
    stamp(7);                                                // counted
    const int n_succ = N - s_tf;
    for (int a = tid; a < nmine; a += T) {
      const unsigned long long key = s_mine[a];
      int r;
      if (key_success(key)) r = s_cnt[a];
      else { r = n_succ + s_fb; for (int b = 0; b < a; ++b) r += !key_success(s_mine[b]); }
      sorted[r] = key;
      if (r == 0 && !P.sort_only) {                       // best candidate of the query (last_best_, tsp_planner.h:131-138)
        write_best(P, query, key);
        if (key_success(key)) {
          const float* src = P.params + ((size_t)query * N + (i0 + a)) * nfd;
          float* dst = P.best + (size_t)query * nfd;
          for (int t = 0; t < nfd; ++t) dst[t] = src[t];
          P.best_valid[query] = 1;
        }
      }
    }
    if (i0 == 0) {
      for (int i = N + tid; i < P.n_pad; i += T) sorted[i] = ~0ull;
      if (tid == 0) P.succ_acc[query] = n_succ;            // read by phase C after the grid barrier
    }
    item += nmine;
  }
}

// ---- coop phase C: one CTA per (query, distribution element); element nfd = the query's scalars ----
template <bool MIRROR>
__device__ void update_phase(const IterParams& P, float* s_part, float* s_val, int s_val_cap, bool mirror_now) {
  const bool mirror = MIRROR && mirror_now;
  auto ustamp = [&](int slot) {      // diagnostics: sub-phase stamps of the update (probe slots 8-11, free outside the exchange path)
    if (P.probe && !P.xchg && threadIdx.x == 0) {
      unsigned long long tt;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tt));
      P.probe[(size_t)blockIdx.x * kProbeSlots + slot] = tt;
    }
  };
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x;
  const int nfd = c.n_free * c.D, N = c.n_cand;
  const int n_items = P.n_query * (nfd + 1);
  for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
    const int query = item / (nfd + 1), t = item - query * (nfd + 1);
    const int n_succ = __ldcg(P.succ_acc + query);
    const int k = elite_count(n_succ, c.elite_frac);
    const unsigned long long* keys = P.keys + (size_t)query * P.n_pad;
    float* mean = P.mean + (size_t)query * nfd;
    float* sigma = P.sigma + (size_t)query * nfd;
    // CES weights of the query, ONCE per CTA into shared memory: u_j = ln(k + 1/2) - ln(j + 1), their wide sum, w_j = u_j / sum -
    // the expressions update_element evaluates on the fly, so the table holds the same bits; the two reductions below then read
    // a weight instead of recomputing a logarithm and a division per term
    const bool table = n_succ > 0 && 2 * k <= s_val_cap;
    float* s_w = s_val + (table ? k : 0);              // [k] weights behind the [k] gathered values
    float lk = 0.0f, sumw = 1.0f;
    ustamp(8);
    if (n_succ > 0) {
      lk = log_poly((float)k + 0.5f);
      if (table) {
        __syncthreads();
        for (int j = tid; j < k; j += T) s_w[j] = ces_log_weight(k, lk, j);
        __syncthreads();
        sumw = CtaSum{s_part}(k, [&](int j, float acc) { return acc + s_w[j]; });
        for (int j = tid; j < k; j += T) s_w[j] = s_w[j] / sumw;
        __syncthreads();
      } else sumw = ces_weight_sum(k, lk, CtaSum{s_part});
    }
    ustamp(9);
    if (t == nfd) {                     // scalars + elite weights of the query
      if (tid == 0) {
        P.n_success[query] = n_succ; P.n_elite[query] = k;
        if (mirror) {                   // the two scalars mirror_scalars() leaves out (it runs beside this CTA)
          reinterpret_cast<int*>(P.host_out + P.mir_nsucc)[query] = n_succ;
          reinterpret_cast<int*>(P.host_out + P.mir_nelite)[query] = k;
        }
      }
      if (n_succ > 0) {
        float* ew = P.elite_w + (size_t)query * P.n_pad;
        for (int j = tid; j < k; j += T) ew[j] = table ? s_w[j] : ces_log_weight(k, lk, j) / sumw;
      }
      continue;
    }
    if (!c.do_update) continue;
    if (n_succ == 0) { if (tid == 0) sigma[t] = clamp_sigma(c, sigma[t] * c.inc); continue; }   // adapt(false)
    StoredParams gp{P.params + (size_t)query * N * nfd, nfd, (unsigned)P.cand_offset};
    if (table) {                    // gather the element's elite values once (keys -> sampled rows: two dependent reads)
      for (int j = tid; j < k; j += T) s_val[j] = gp((unsigned)(__ldcg(keys + j) & 0xFFFFFFFFu), t);
      __syncthreads();
      ustamp(10);
      update_element_w(c, k, t, mean, sigma, [&](int j) { return s_val[j]; }, [&](int j) { return s_w[j]; }, CtaSum{s_part}, tid == 0);
      ustamp(11);
    } else if (k <= s_val_cap) {
      __syncthreads();
      for (int j = tid; j < k; j += T) s_val[j] = gp((unsigned)(__ldcg(keys + j) & 0xFFFFFFFFu), t);
      __syncthreads();
      update_element(c, k, lk, sumw, t, mean, sigma, [&](int j) { return s_val[j]; }, CtaSum{s_part}, tid == 0);
    } else {
      update_element(c, k, lk, sumw, t, mean, sigma, [&](int j) { return gp((unsigned)(keys[j] & 0xFFFFFFFFu), t); }, CtaSum{s_part}, tid == 0);
    }
  }
}

// ---- coop: result block -> its mirror in mapped pinned host memory (IterParams::host_out) ----
// L / Cnf / Cwf / cost / status are final at grid barrier 1: they go out right behind it as 16-byte vectors, one per
// thread, by the CTAs at the top of the grid; the stores are posted, so they drain over PCIe while the keys are ranked
// and the distribution is updated.  best_idx / best_cost and best | best_valid | any_success are final at barrier 2
// and follow there; update_phase writes n_success / n_elite to both places.  ssppcu_get_results then needs no D2H copy.
__device__ __noinline__ void mirror_candidates(const IterParams& P) {
  const unsigned tid = threadIdx.x, T = blockDim.x;
  const unsigned total = P.mir_nsucc >> 4;                // [0, n_success): the per-candidate arrays
  const uint4* src = reinterpret_cast<const uint4*>(P.res_dev);
  uint4* dst = reinterpret_cast<uint4*>(P.host_out);
  for (unsigned v = (gridDim.x - 1 - blockIdx.x) * T + tid; v < total; v += gridDim.x * T) dst[v] = __ldcg(src + v);
}
__device__ __noinline__ void mirror_scalars(const IterParams& P) {
  if (blockIdx.x != gridDim.x - 1) return;
  const unsigned tid = threadIdx.x, T = blockDim.x;
  const unsigned w0 = P.mir_bidx >> 2, w1 = P.mir_bytes >> 2;     // [best_idx, end of the block)
  const unsigned* src = reinterpret_cast<const unsigned*>(P.res_dev);
  unsigned* dst = reinterpret_cast<unsigned*>(P.host_out);
  for (unsigned e = w0 + tid; e < w1; e += T) dst[e] = __ldcg(src + e);
  const unsigned words = (unsigned)P.n_query * (unsigned)(P.cfg.n_free * P.cfg.D + 2);   // best | best_valid | any_success
  const unsigned* s = reinterpret_cast<const unsigned*>(P.best);
  unsigned* d = reinterpret_cast<unsigned*>(P.host_out + P.mir_extra);
  for (unsigned e = tid; e < words; e += T) d[e] = __ldcg(s + e);
}

// ---- population shard with the fused key exchange (ExchangeParams, sspp_kernels.h) ----
__device__ __forceinline__ void merge_rank_body(const MergeParams& P, int block, int n_blocks, unsigned long long* s_keys, int cap_keys);
// regenerate a candidate's sampled element from the Philox counters (population-sharded path)
struct RegenParams {
  const ScoreConfig* c; uint64_t seed; uint32_t iter; const float* mean0; const float* sigma0; const float* best0; int best_ok;
  __device__ __forceinline__ float operator()(unsigned idx, int t) const {
    const int kind = candidate_kind(*c, (int)idx, best_ok);
    return candidate_param(*c, seed, iter, 0u, (int)idx, kind, t, mean0, sigma0, best0);
  }
};
__device__ __forceinline__ void merge_finish_body(const MergeParams& P, float* s_part, int block, int n_blocks, const float* vals = nullptr, int vals_pitch = 0);

// One exchange line = 16 bytes = two 8-byte halves {key word, epoch}.  A reader accepts a line when BOTH halves carry the
// epoch, so data and flag travel together and no fence or separate signal is needed - provided each 8-byte half becomes
// visible as a whole.  PTX promises that for a naturally aligned 8-byte SCALAR access (single-copy atomicity), not for a vector
// access.  protocol 1 (default) therefore writes / reads a line as two 8-byte scalar accesses, relaxed at system scope (a
// strong operation: not torn, not cached stale); protocol 0 is the first implementation - one 16-byte volatile vector access per
// line - which relies on the hardware never tearing an aligned 8-byte half of a vector store over NVLink.  Both were measured
// (profiles/r02_popshard_*): SSPPCU_XCHG_PROTOCOL selects.
__device__ __forceinline__ void st_line(uint4* p, unsigned long long key, unsigned epoch, int protocol) {
  if (protocol == 0) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};"
                 :: "l"(p), "r"((unsigned)key), "r"(epoch), "r"((unsigned)(key >> 32)), "r"(epoch) : "memory");
    return;
  }
  const unsigned long long lo = (unsigned long long)(unsigned)key | ((unsigned long long)epoch << 32);
  const unsigned long long hi = (key >> 32) | ((unsigned long long)epoch << 32);
  asm volatile("st.relaxed.sys.global.b64 [%0], %1;" :: "l"(p), "l"(lo) : "memory");
  asm volatile("st.relaxed.sys.global.b64 [%0 + 8], %1;" :: "l"(p), "l"(hi) : "memory");
}
__device__ __forceinline__ uint4 ld_line(const uint4* p, int protocol) {
  uint4 v;
  if (protocol == 0) {
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
  }
  unsigned long long lo, hi;
  asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(lo) : "l"(p) : "memory");
  asm volatile("ld.relaxed.sys.global.b64 %0, [%1 + 8];" : "=l"(hi) : "l"(p) : "memory");
  v.x = (unsigned)lo; v.y = (unsigned)(lo >> 32); v.z = (unsigned)hi; v.w = (unsigned)(hi >> 32);
  return v;
}

template <class Stamp>
__device__ void exchange_and_merge(const IterParams& P, cg::grid_group& grid, float* s_part, uint32_t iter, unsigned char* smem_raw, Stamp stamp) {
  __shared__ MergeParams s_mp;
  const ExchangeParams& X = *P.xchg;
  const int tid = threadIdx.x, T = blockDim.x;
  const int W = X.world, len = X.shard_len, me = X.rank;
  const unsigned epoch = P.xchg_epoch, half = epoch & 1u;
  int* succ = X.succ;
  if (blockIdx.x == 0 && tid == 0) *succ = 0;
  grid.sync();                                          // the shard's sorted order is complete
  stamp(8);
  // publish: key j of this shard -> line (half, this rank, j) of every rank's block, its own included
  const size_t slot = ((size_t)half * W + me) * len;
  const long long total = (long long)W * len;
  for (long long e = (long long)blockIdx.x * T + tid; e < total; e += (long long)gridDim.x * T) {
    const int j = (int)(e / W), r = (int)(e - (long long)j * W);      // neighbouring threads feed different links
    st_line(X.peer_lines[r] + slot + j, __ldcg(P.keys + j), epoch, X.protocol);
  }
  stamp(9);
  // receive: poll every line of this epoch's half until both flag words carry the epoch; keys go to a compact array
  const uint4* lines = X.lines_local + (size_t)half * W * len;
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  for (long long e = (long long)blockIdx.x * T + tid; e < total; e += (long long)gridDim.x * T) {
    uint4 v = ld_line(lines + e, X.protocol);
    while (v.y != epoch || v.w != epoch) {
      unsigned long long t1;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > X.timeout_ns) { *reinterpret_cast<volatile int*>(X.status) = 1; break; }
      __nanosleep(100);
      v = ld_line(lines + e, X.protocol);
    }
    X.gathered[e] = (unsigned long long)v.x | ((unsigned long long)v.z << 32);
  }
  grid.sync();
  stamp(10);
  if (tid == 0) {
    MergeParams& M = s_mp;
    M.cfg = P.cfg; M.cfg.n_cand = X.n_global;
    M.n_keys = W * len; M.n_shards = W; M.shard_len = len; M.n_pad = X.n_pad;
    M.iter = iter; M.seed = P.seed;
    M.gathered = X.gathered;
    M.keys = X.mkeys; M.elite_w = X.elite_w; M.state0 = X.state0; M.succ_acc = succ;
    M.mean = P.mean; M.sigma = P.sigma; M.best = P.best; M.best_valid = P.best_valid;
    M.n_success = P.n_success; M.n_elite = P.n_elite; M.best_idx = P.best_idx; M.best_cost = P.best_cost;
    M.any_success = P.any_success;
  }
  __syncthreads();
  merge_rank_body(s_mp, blockIdx.x, gridDim.x, reinterpret_cast<unsigned long long*>(smem_raw), P.smem_words / 2);
  stamp(11);
  grid.sync();
  // the elites' sampled values come back from the Philox counters (only keys crossed NVLink): regenerate them with the
  // whole grid into vals[element][elite], then the update reads them like the unsharded kernel reads its stored rows
  const ScoreConfig& cg = s_mp.cfg;
  const int nfd = cg.n_free * cg.D;
  const int n_succ = __ldcg(succ);
  const int k = elite_count(n_succ, cg.elite_frac);
  if (cg.do_update && n_succ > 0) {
    const int best_ok0 = cg.include_best ? __ldcg(reinterpret_cast<const int*>(X.state0) + 3 * nfd) : 0;
    RegenParams gp{&cg, P.seed, iter, X.state0, X.state0 + nfd, X.state0 + 2 * nfd, best_ok0};
    const long long items = (long long)nfd * k;
    for (long long e = (long long)blockIdx.x * T + tid; e < items; e += (long long)gridDim.x * T) {
      const int t = (int)(e / k), j = (int)(e - (long long)t * k);
      X.vals[(size_t)t * X.n_pad + j] = gp((unsigned)(__ldcg(X.mkeys + j) & 0xFFFFFFFFu), t);
    }
    grid.sync();
  }
  merge_finish_body(s_mp, s_part, blockIdx.x, gridDim.x, X.vals, X.n_pad);
  stamp(5);
}

// Launch-bound variants: (max threads, min CTAs per SM) fixes the register budget ptxas works with, and so
// the resident warps per SM; the handle picks one per problem shape (kIterVariants, sspp_kernels.h).
// MIRROR (cooperative shapes only): the instantiation that posts its results to the mapped pinned host block; the plain
// one carries none of that code, so a device-resident caller runs exactly the kernel that was tuned without it
template <bool COOP, int MAXT, int MINB, bool MIRROR = false>
__global__ void __launch_bounds__(MAXT, MINB)
sspp_iterate_kernel(const __grid_constant__ ModelView MG, const __grid_constant__ IterParams P) {
  __shared__ int s_int[4];
  __shared__ float s_part[8];
  __shared__ __align__(8) unsigned long long s_mbar;    // completion barrier of the bulk copies (rank phase)
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x;
  const int nfd = c.n_free * c.D;
  unsigned mbar_phase = 0;
  if (COOP && tid == 0) {
    const unsigned m = (unsigned)__cvta_generic_to_shared(&s_mbar);
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(m) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const SmemModel M = stage_model(MG, tid, T);      // visible after the first barrier below (load_basis / score_tile)
  unsigned char* smem_raw = smem_all + P.model_bytes;
  const Smem s = carve(smem_raw, c, P, T);
  if (!COOP) {
    // one CTA per query: every iteration of the launch stays inside this CTA
    const int query = blockIdx.x;
    load_basis(s, c, P, tid, T);
    for (int it = 0; it < P.n_inner; ++it) {
      const bool from_saved = has_snapshot(P) && it == 0;
      score_tile(M, P, s, query, 0, c.n_cand, P.iter0 + (uint32_t)it, from_saved);
      if (from_saved) {               // re-arm this query's live state from the snapshot (nobody else touches it)
        const float* snap = snapshot_of(P);
        const size_t sq = (size_t)P.n_query * nfd;
        for (int t = tid; t < nfd; t += T)
          for (int b = 0; b < 3; ++b) P.mean[b * sq + (size_t)query * nfd + t] = snap[b * sq + (size_t)query * nfd + t];
        if (tid == 0) {
          const int* si = reinterpret_cast<const int*>(snap + 3 * sq);
          P.best_valid[query] = si[query]; P.any_success[query] = si[P.n_query + query];
        }
        if (P.inline_words)
          for (int e = tid; e < c.K * c.D; e += T) const_cast<float*>(P.base)[(size_t)query * c.K * c.D + e] = P.inline_blk[(size_t)query * c.K * c.D + e];
      }
      __syncthreads();                                   // raw keys were written by this CTA: visible after the barrier
      unsigned long long* keys = s.keys;
      const unsigned long long* raw = P.keys_raw + (size_t)query * P.n_pad;
      unsigned long long* sorted = P.keys + (size_t)query * P.n_pad;
      for (int i = tid; i < P.n_pad; i += T) keys[i] = (i < c.n_cand) ? raw[i] : ~0ull;
      __syncthreads();
      if (P.sort_only) block_bitonic_sort(keys, P.n_pad);
      else {
        StoredParams gp{P.params + (size_t)query * c.n_cand * nfd, nfd, (unsigned)P.cand_offset};
        cta_elites_and_update(c, keys, P.n_pad, P.elite_w + (size_t)query * P.n_pad, P.mean + (size_t)query * nfd,
                              P.sigma + (size_t)query * nfd, P.best + (size_t)query * nfd, P.best_valid + query,
                              P.n_success + query, P.n_elite + query, P.best_idx + query, P.best_cost + query,
                              P.any_success + query, gp, s_int);
      }
      for (int i = tid; i < P.n_pad; i += T) sorted[i] = keys[i];
      __syncthreads();
    }
    return;
  }
  cg::grid_group grid = cg::this_grid();
  // optional phase probe (diagnostics): thread 0 of every CTA stamps %globaltimer at the phase boundaries of the
  // launch's first iteration: [cta][0] start, [1] scored, [2] past barrier 1, [3] ranked, [4] past barrier 2, [5] updated
  auto stamp = [&](int it, int slot) {
    if (P.probe && it == 0 && tid == 0) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      P.probe[(size_t)blockIdx.x * kProbeSlots + slot] = t;
    }
  };
  for (int it = 0; it < P.n_inner; ++it) {
    stamp(it, 0);
    const bool from_saved = has_snapshot(P) && it == 0;
    if (from_saved) {                 // re-arm the live state from the snapshot; phase A reads the snapshot itself
      const float* snap = snapshot_of(P);
      const size_t words = (size_t)P.n_query * (3 * nfd + 2);
      for (size_t e = (size_t)blockIdx.x * T + tid; e < words; e += (size_t)gridDim.x * T) P.mean[e] = snap[e];
      if (P.inline_words) {           // the point matrices too: later iterations of this launch and the read-back calls use them
        const size_t bw = (size_t)P.n_query * c.K * c.D;
        for (size_t e = (size_t)blockIdx.x * T + tid; e < bw; e += (size_t)gridDim.x * T) const_cast<float*>(P.base)[e] = P.inline_blk[e];
      }
    }
    load_basis(s, c, P, tid, T);                         // phase B reuses this shared memory
    // ---- phase A: score the tiles ----
    const bool handout = P.n_tiles > (int)gridDim.x;     // first wave by CTA index, the rest first come first served:
    for (int tile = blockIdx.x; tile < P.n_tiles;) {     // tiles differ in cost (culling), a fixed assignment leaves a tail
      const int query = tile / P.ctas_per_query;
      const int c0 = (tile - query * P.ctas_per_query) * P.cand_per_cta;
      const int nc = min(P.cand_per_cta, c.n_cand - c0);
      score_tile(M, P, s, query, c0, nc, P.iter0 + (uint32_t)it, from_saved);
      if (!handout) { __syncthreads(); break; }
      if (tid == 0) s_int[0] = (int)gridDim.x + atomicAdd(P.tile_counter, 1);
      __syncthreads();
      tile = s_int[0];
      __syncthreads();                                   // everybody has read the ticket before thread 0 can draw the next
    }
    stamp(it, 1);
    grid.sync();
    stamp(it, 2);
    const bool mirror = MIRROR && it + 1 == P.n_inner;
    if (mirror) mirror_candidates(P);
    if (handout && blockIdx.x == 0 && tid == 0) *P.tile_counter = 0;   // nobody draws again before the next iteration's barrier
    // ---- phase B: order the keys ----
    rank_phase(P, smem_raw, &s_mbar, mbar_phase, [&](int slot) { stamp(it, slot); });
    stamp(it, 3);
    if (P.sort_only) {
      if (P.xchg) exchange_and_merge(P, grid, s_part, P.iter0 + (uint32_t)it, smem_raw, [&](int slot) { stamp(it, slot); });
      if (it + 1 < P.n_inner) grid.sync();
      continue;
    }
    grid.sync();
    stamp(it, 4);
    // ---- phase C: distribution update ----
    if (mirror) mirror_scalars(P);
    update_phase<MIRROR>(P, s_part, reinterpret_cast<float*>(smem_raw), P.smem_words, mirror);
    stamp(it, 5);

    if (it + 1 < P.n_inner) grid.sync();
  }
}


// count of keys < key in the ascending run a[0..n)
__device__ __forceinline__ int lower_bound_cg(const unsigned long long* a, int n, unsigned long long key) {
  int lo = 0, hi = n;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (__ldcg(a + mid) < key) lo = mid + 1; else hi = mid; }
  return lo;
}
// the same by a whole warp: 32 probes per round cut the run 32-fold (3 rounds of L2 latency for 8192 keys instead of 13)
__device__ __forceinline__ int lower_bound_warp(const unsigned long long* a, int n, unsigned long long key, int lane) {
  int lo = 0, hi = n;
  while (hi > lo) {
    const int seg = (hi - lo + 31) >> 5;
    const int q = lo + lane * seg;
    const bool less = q < hi && __ldcg(a + q) < key;
    const int cnt = __popc(__ballot_sync(0xFFFFFFFFu, less));   // the probes are ascending: the first cnt of them are smaller
    if (cnt == 0) { hi = lo; break; }
    const int nlo = lo + (cnt - 1) * seg + 1;
    hi = min(hi, lo + cnt * seg);
    lo = nlo;
  }
  return lo;
}
__device__ __forceinline__ int lower_bound_smem(const unsigned long long* a, int n, unsigned long long key) {
  int lo = 0, hi = n;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if (a[mid] < key) lo = mid + 1; else hi = mid; }
  return lo;
}

// Merge-rank over a virtual grid of n_blocks CTAs (this one is `block`).  A CTA takes T consecutive keys of one
// shard; their positions in another shard all lie in the window between the positions of the first and the last
// of them, so two binary searches per shard and CTA (not per key) find the windows, the windows are staged in
// shared memory (s_keys, cap_keys entries; a window that does not fit is searched in place), and every thread
// finishes its key with short searches there.  Position = own position + the counts in the other shards.
__device__ __forceinline__ void merge_rank_body(const MergeParams& P, int block, int n_blocks, unsigned long long* s_keys, int cap_keys) {
  __shared__ int s_wlo[kMaxExchangeRanks], s_wn[kMaxExchangeRanks], s_woff[kMaxExchangeRanks];
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31;
  const int nfd = c.n_free * c.D, W = P.n_shards, len = P.shard_len;
  if (block == 0) {            // snapshot of the state the elites were sampled from
    for (int t = tid; t < nfd; t += T) {
      P.state0[t] = P.mean[t]; P.state0[nfd + t] = P.sigma[t]; P.state0[2 * nfd + t] = P.best[t];
    }
    if (tid == 0) reinterpret_cast<int*>(P.state0)[3 * nfd] = *P.best_valid;
  }
  for (int k = P.n_keys + block * T + tid; k < P.n_pad; k += n_blocks * T) P.keys[k] = ~0ull;
  const int chunks = (len + T - 1) / T;
  for (int item = block; item < W * chunks; item += n_blocks) {
    const int shard = item / chunks, p0 = (item - shard * chunks) * T;
    const int n_mine = min(T, len - p0);
    const unsigned long long* mine = P.gathered + (size_t)shard * len + p0;
    const unsigned long long key = (tid < n_mine) ? __ldcg(mine + tid) : ~0ull;
    // Failed keys order by candidate index alone, successful ones by cost: a chunk that holds the last successes and the
    // first failures of its shard would span most of every other shard.  Such a chunk is merged in two parts.
    const int n_ok = __syncthreads_count(tid < n_mine && key_success(key));
    const int n_parts = (n_ok > 0 && n_ok < n_mine) ? 2 : 1;
    int succ_hi = 0;
    for (int part = 0; part < n_parts; ++part) {
      const int a0 = (part == 0) ? 0 : n_ok, a1 = (n_parts == 1 || part == 1) ? n_mine : n_ok;
      __syncthreads();                                   // the previous windows have been consumed
      for (int b = tid >> 5; b < 2 * W; b += T >> 5) {   // window bounds, one warp each: 2s -> first key, 2s+1 -> last key, in shard s
        const int s2 = b >> 1;
        if (s2 == shard) continue;
        const unsigned long long bound = __ldcg(mine + ((b & 1) ? a1 - 1 : a0));
        const int pos = lower_bound_warp(P.gathered + (size_t)s2 * len, len, bound, lane);
        if (lane == 0) { if (b & 1) s_wn[s2] = pos; else s_wlo[s2] = pos; }
      }
      __syncthreads();
      if (tid == 0) {                                    // window lengths and their places in the staging buffer
        int off = 0;
        for (int s2 = 0; s2 < W; ++s2) {
          if (s2 == shard) { s_wn[s2] = 0; s_woff[s2] = -1; continue; }
          const int n = s_wn[s2] - s_wlo[s2];
          s_wn[s2] = n;
          if (off + n <= cap_keys) { s_woff[s2] = off; off += n; } else s_woff[s2] = -1;
        }
      }
      __syncthreads();
      for (int s2 = 0; s2 < W; ++s2) {
        const int off = s_woff[s2];
        if (off < 0) continue;
        const unsigned long long* w = P.gathered + (size_t)s2 * len + s_wlo[s2];
        for (int j = tid; j < s_wn[s2]; j += T) s_keys[off + j] = __ldcg(w + j);
      }
      __syncthreads();
      if (tid >= a0 && tid < a1) {
        int rank = p0 + tid;
        for (int s2 = 0; s2 < W; ++s2) {
          if (s2 == shard) continue;
          const int off = s_woff[s2];
          rank += s_wlo[s2] + (off >= 0 ? lower_bound_smem(s_keys + off, s_wn[s2], key)
                                        : lower_bound_cg(P.gathered + (size_t)s2 * len + s_wlo[s2], s_wn[s2], key));
        }
        P.keys[rank] = key;
        if (key_success(key)) succ_hi = rank + 1;
      }
    }
    succ_hi = __reduce_max_sync(0xFFFFFFFFu, succ_hi);
    if (lane == 0 && succ_hi) atomicMax(P.succ_acc, succ_hi);
  }
}

__global__ void __launch_bounds__(256)
sspp_merge_rank_kernel(const __grid_constant__ MergeParams P) {
  __shared__ unsigned long long s_keys[2048];
  merge_rank_body(P, blockIdx.x, gridDim.x, s_keys, 2048);
}

// finish over a virtual grid of n_blocks CTAs (this one is `block`): one CTA per distribution element (+1: the scalars)
__device__ __forceinline__ void merge_finish_body(const MergeParams& P, float* s_part, int block, int n_blocks, const float* vals, int vals_pitch) {
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x;
  const int nfd = c.n_free * c.D;
  const int n_succ = __ldcg(P.succ_acc);
  const int k = elite_count(n_succ, c.elite_frac);
  const int best_ok0 = c.include_best ? __ldcg(reinterpret_cast<const int*>(P.state0) + 3 * nfd) : 0;
  RegenParams gp{&c, P.seed, P.iter, P.state0, P.state0 + nfd, P.state0 + 2 * nfd, best_ok0};
  for (int t = block; t <= nfd; t += n_blocks) {
    if (t == nfd) {
      if (tid == 0) { *P.n_success = n_succ; *P.n_elite = k; }
      const unsigned long long k0 = __ldcg(P.keys);
      if (n_succ > 0) {
        if (tid == 0) { *P.best_idx = (int)(unsigned)(k0 & 0xFFFFFFFFu); *P.best_cost = key_cost(k0); *P.any_success = 1; *P.best_valid = 1; }
        for (int e = tid; e < nfd; e += T) P.best[e] = gp((unsigned)(k0 & 0xFFFFFFFFu), e);
        const float lk = log_poly((float)k + 0.5f);
        const float sumw = ces_weight_sum(k, lk, CtaSum{s_part});
        for (int j = tid; j < k; j += T) P.elite_w[j] = ces_log_weight(k, lk, j) / sumw;
      } else if (tid == 0) { *P.best_idx = -1; *P.best_cost = __int_as_float(0x7F800000); }
      continue;
    }
    if (!c.do_update) continue;
    if (n_succ == 0) { if (tid == 0) P.sigma[t] = clamp_sigma(c, P.sigma[t] * c.inc); continue; }
    const float lk = log_poly((float)k + 0.5f);
    const float sumw = ces_weight_sum(k, lk, CtaSum{s_part});
    if (vals) update_element(c, k, lk, sumw, t, P.mean, P.sigma, [&](int j) { return __ldcg(vals + (size_t)t * vals_pitch + j); }, CtaSum{s_part}, tid == 0);
    else update_element(c, k, lk, sumw, t, P.mean, P.sigma, [&](int j) { return gp((unsigned)(__ldcg(P.keys + j) & 0xFFFFFFFFu), t); }, CtaSum{s_part}, tid == 0);
  }
}

__global__ void __launch_bounds__(256)
sspp_merge_finish_kernel(const __grid_constant__ MergeParams P) {
  __shared__ float s_part[8];
  merge_finish_body(P, s_part, blockIdx.x, gridDim.x);
}

__global__ void __launch_bounds__(128)
sspp_eval_points_kernel(const __grid_constant__ ModelView M, int D, const float* __restrict__ pts, int n, float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_qs = reinterpret_cast<float*>(smem_raw);
  float* s_xf = s_qs + D * blockDim.x;
  const int tid = threadIdx.x, T = blockDim.x;
  const int i = blockIdx.x * T + tid;
  if (i >= n) return;
  for (int d = 0; d < D; ++d) s_qs[d * T + tid] = pts[(size_t)i * D + d];
  SmemXF xf{s_xf, T, tid};
  SmemQ qa{s_qs, T, tid};
  out[i] = point_collision_cost(M, qa, xf);
}

// Collision.h:51-82 check_collision_point for n configurations: per point the penetrating contact (dist < -1e-3) with the
// smallest geom pair, as FlatModel geom ids, and its distance; geom1 = -1 when the configuration is free.  Diagnostic path
// (the planner loop never calls it): scene tables read from global memory, primitives mode only.
__global__ void __launch_bounds__(128)
sspp_check_points_kernel(const __grid_constant__ ModelView M, int D, const float* __restrict__ pts, int n, const int* __restrict__ mg_src,
                         const int* __restrict__ sg_src, int* __restrict__ geom1, int* __restrict__ geom2, float* __restrict__ dist) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_qs = reinterpret_cast<float*>(smem_raw);
  float* s_xf = s_qs + D * blockDim.x;
  const int tid = threadIdx.x, T = blockDim.x;
  const int i = blockIdx.x * T + tid;
  if (i >= n) return;
  for (int d = 0; d < D; ++d) s_qs[d * T + tid] = pts[(size_t)i * D + d];
  SmemXF xf{s_xf, T, tid};
  SmemQ qa{s_qs, T, tid};
  ContactProbe cp{mg_src, sg_src, 0xFFFFFFFFu, 0.0f};
  (void)point_collision_cost(M, qa, xf, nullptr, &cp);
  const bool hit = cp.key != 0xFFFFFFFFu;
  geom1[i] = hit ? (int)(cp.key >> 16) : -1;
  geom2[i] = hit ? (int)(cp.key & 0xFFFFu) : -1;
  dist[i] = hit ? cp.dist : 0.0f;
}

// Success rows for the host (sspp.h:215-216 hands every collision-free spline back to the caller): compacts the sampled rows
// of the converged candidates of every query, in candidate order, into `rows` [Q][N][nfd] / `idx` [Q][N] / `count` [Q] -
// pointers into the handle's mapped pinned block, so the rows cross PCIe as posted stores straight from this kernel and
// only the successful ones do.  One CTA per 256 candidates: its base slot is the number of successes in front of its first
// candidate (a count over at most N status bytes), the slots inside by ballot prefix.
__global__ void __launch_bounds__(256)
sspp_success_rows_kernel(const uint8_t* __restrict__ status, const float* __restrict__ params, int n_cand, int nfd,
                         float* __restrict__ rows, int32_t* __restrict__ idx, int32_t* __restrict__ count) {
  __shared__ int s_base, s_warp[8];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int query = blockIdx.y, c0 = blockIdx.x * 256;
  const uint8_t* st = status + (size_t)query * n_cand;
  if (tid == 0) s_base = 0;
  __syncthreads();
  int before = 0;
  for (int i = tid; i < c0; i += 256) before += st[i] == ST_CONVERGED;
  before = __reduce_add_sync(0xFFFFFFFFu, before);
  if (lane == 0 && before) atomicAdd(&s_base, before);
  const int cand = c0 + tid;
  const bool ok = cand < n_cand && st[cand] == ST_CONVERGED;
  const unsigned b = __ballot_sync(0xFFFFFFFFu, ok);
  if (lane == 0) s_warp[warp] = __popc(b);
  __syncthreads();
  int slot = s_base + __popc(b & ((1u << lane) - 1u));
  for (int w = 0; w < warp; ++w) slot += s_warp[w];
  if (ok) idx[(size_t)query * n_cand + slot] = cand;
  if (blockIdx.x == gridDim.x - 1 && tid == 255) {                    // the last CTA knows the total
    int tot = s_base;
    for (int w = 0; w < 8; ++w) tot += s_warp[w];
    count[query] = tot;
  }
  // rows: a warp copies the rows of its successes one after another, lanes across the row (coalesced posted stores)
  unsigned todo = b;
  while (todo) {
    const int l = __ffs((int)todo) - 1;
    todo &= todo - 1u;
    const int src_cand = c0 + warp * 32 + l;
    const int dst_slot = __shfl_sync(0xFFFFFFFFu, slot, l);
    const float* src = params + ((size_t)query * n_cand + src_cand) * nfd;
    float* dst = rows + ((size_t)query * n_cand + dst_slot) * nfd;
    for (int t = lane; t < nfd; t += 32) dst[t] = __ldcg(src + t);
  }
}

// Batched B-spline evaluation (sspp/BSplines.py:44-51 over many curves at once): out[b][t][:] = sum_i rows[t][i] ctrl[b][i][:],
// i ascending from +0 with explicit fmaf (the numpy twin restates the same sum).  rows = curve_row() of every parameter
// (banded: at most order + 1 non-zeros, zeros skipped by value); one thread per (curve, parameter), rows staged in shared memory.
__global__ void __launch_bounds__(256)
sspp_bspline_eval_kernel(const float* __restrict__ rows, int T_, int n, const float* __restrict__ ctrl, int B_, int d, float* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_rows = reinterpret_cast<float*>(smem_raw);
  for (int e = threadIdx.x; e < T_ * n; e += blockDim.x) s_rows[e] = rows[e];
  __syncthreads();
  const long long total = (long long)B_ * T_;
  for (long long item = (long long)blockIdx.x * blockDim.x + threadIdx.x; item < total; item += (long long)gridDim.x * blockDim.x) {
    const int b = (int)(item / T_), t = (int)(item - (long long)b * T_);
    const float* c = ctrl + (size_t)b * n * d;
    const float* r = s_rows + (size_t)t * n;
    for (int e = 0; e < d; ++e) {
      float acc = 0.0f;
      for (int i = 0; i < n; ++i) { const float w = r[i]; if (w != 0.0f) acc = fmaf(w, c[(size_t)i * d + e], acc); }
      out[((size_t)b * T_ + t) * d + e] = acc;
    }
  }
}

// Diagnostics: the EXECUTED work of the last scored iteration (roofline: flop_per_point_executed).  Re-scores the stored
// candidates of query 0 .. n_query-1 (sampled rows P.params, point matrix P.base) on the collision grid with the work
// counters of sspp_math.h switched on; one thread per (candidate, collision row), totals by warp reduction + atomics.
// Never part of a timed step.
__global__ void __launch_bounds__(128)
sspp_count_kernel(const __grid_constant__ ModelView M, const __grid_constant__ IterParams P, unsigned long long* __restrict__ totals) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x, D = c.D, K = c.K, nfd = c.n_free * D;
  float* s_qs = reinterpret_cast<float*>(smem_raw);
  float* s_xf = s_qs + D * T;
  uint32_t wc[WC_COUNT];
#pragma unroll
  for (int e = 0; e < WC_COUNT; ++e) wc[e] = 0u;
  const long long total = (long long)P.n_query * c.n_cand * c.Mc;
  for (long long item = (long long)blockIdx.x * T + tid; item < total; item += (long long)gridDim.x * T) {
    const int m = (int)(item % c.Mc);
    const long long qc = item / c.Mc;                      // query * n_cand + candidate
    const int query = (int)(qc / c.n_cand);
    if (m < c.col_first || P.status[qc] == ST_ABSENT) continue;
    const float* base = P.base + (size_t)query * K * D;
    const float* par = P.params + (size_t)qc * nfd;
    const float* row = P.psi_col + (size_t)m * P.w_col;
    const int first = P.first_col[m];
    for (int d = 0; d < D; ++d) {
      float acc = 0.0f;
      for (int k = 0; k < P.w_col; ++k) {
        const int r = first + k, fr = r - c.free_begin;
        const float x = (fr >= 0 && fr < c.n_free) ? par[fr * D + d] : base[r * D + d];
        acc = fmaf(row[k], x, acc);
      }
      s_qs[d * T + tid] = acc;
    }
    SmemXF xf{s_xf, T, tid};
    SmemQ qa{s_qs, T, tid};
    (void)point_collision_cost(M, qa, xf, wc);
  }
#pragma unroll
  for (int e = 0; e < WC_COUNT; ++e) {
    const unsigned tot = __reduce_add_sync(0xFFFFFFFFu, wc[e]);
    if ((tid & 31) == 0 && tot) atomicAdd(totals + e, (unsigned long long)tot);
  }
}

// SDF refinement step: per collocation row m the configuration q_m = sum_j W[m][j] pts[j], the closest
// sample sphere and g_m = -clearance * grad / |grad| (clearance times the unit direction towards the obstacle, the
// `grad` of scripts/sdf_gradient/main.py:60-61); then G_j = (1/rows) sum_m g_m W[m][j] (main.py:66-81) by one thread
// per output in row order.  One CTA: rows are few (the collocation grid of one path).
__global__ void __launch_bounds__(128)
sspp_sdf_gradient_kernel(const __grid_constant__ ModelView M, int D, int K, int rows, const float* __restrict__ W,
                         const float* __restrict__ pts, float* __restrict__ dist, float* __restrict__ dirs, float* __restrict__ G) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  float* s_qs = reinterpret_cast<float*>(smem_raw);
  float* s_xf = s_qs + D * blockDim.x;
  const int tid = threadIdx.x, T = blockDim.x;
  for (int m0 = 0; m0 < rows; m0 += T) {
    const int m = m0 + tid;
    if (m < rows) {
      for (int d = 0; d < D; ++d) {
        float acc = 0.0f;
        for (int j = 0; j < K; ++j) acc = fmaf(W[(size_t)m * K + j], pts[(size_t)j * D + d], acc);
        s_qs[d * T + tid] = acc;
      }
      SmemXF xf{s_xf, T, tid};
      SmemQ qa{s_qs, T, tid};
      float cl; v3 g;
      point_closest_sample(M, qa, xf, &cl, &g);
      const float n = sqrtf(dot(g, g));
      const float sc = (n > 0.0f) ? -cl / n : 0.0f;
      dist[m] = cl;
      dirs[3 * m] = sc * g.x; dirs[3 * m + 1] = sc * g.y; dirs[3 * m + 2] = sc * g.z;
    }
  }
  __threadfence_block();
  __syncthreads();
  for (int o = tid; o < K * 3; o += T) {
    const int j = o / 3, a = o - 3 * j;
    float acc = 0.0f;
    for (int m = 0; m < rows; ++m) acc = fmaf(dirs[3 * m + a], W[(size_t)m * K + j], acc);
    G[o] = acc / (float)rows;
  }
}

// node value = min over the static primitives of the exact point-primitive signed distance
__global__ void __launch_bounds__(256)
sspp_sdf_bake_kernel(const __grid_constant__ ModelView M, float* __restrict__ nodes) {
  const SdfGrid& g = M.grid;
  const size_t n = (size_t)g.nx * g.ny * g.nz;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(e % g.nx), j = (int)((e / g.nx) % g.ny), k = (int)(e / ((size_t)g.nx * g.ny));
    nodes[e] = sdf_node_value(M, i, j, k);
  }
}
__global__ void __launch_bounds__(256)
sspp_sdf_pack_kernel(const __grid_constant__ ModelView M, const float* __restrict__ nodes, float* __restrict__ cells) {
  const SdfGrid& g = M.grid;
  const int cx = g.nx - 1, cy = g.ny - 1, cz = g.nz - 1;
  const size_t n = (size_t)cx * cy * cz;
  for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (size_t)gridDim.x * blockDim.x) {
    const int i = (int)(e % cx), j = (int)((e / cx) % cy), k = (int)(e / ((size_t)cx * cy));
#pragma unroll
    for (int c = 0; c < 8; ++c)
      cells[e * 8 + c] = nodes[((size_t)(k + (c >> 2)) * g.ny + (j + ((c >> 1) & 1))) * g.nx + (i + (c & 1))];
  }
}

__global__ void __launch_bounds__(256)
sspp_fma_probe_kernel(float* sink, int iters) {
  float a[8];
  const float x = 1.0f + 1e-7f * (float)threadIdx.x, y = 1e-9f * (float)blockIdx.x;
#pragma unroll
  for (int k = 0; k < 8; ++k) a[k] = (float)k;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 16; ++r)
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] = fmaf(a[k], x, y);
  }
  float s = 0.0f;
#pragma unroll
  for (int k = 0; k < 8; ++k) s += a[k];
  if (s == 123.456f) sink[0] = s;   // never true: keeps the chains alive
}

}  // namespace

cudaError_t launch_sdf_bake(const ModelView& M, float* nodes, float* cells, int n_sm, cudaStream_t stream) {
  sspp_sdf_bake_kernel<<<n_sm * 8, 256, 0, stream>>>(M, nodes);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  sspp_sdf_pack_kernel<<<n_sm * 8, 256, 0, stream>>>(M, nodes, cells);
  return cudaGetLastError();
}

cudaError_t launch_fma_probe(float* sink, int blocks, int threads, int iters, cudaStream_t stream) {
  sspp_fma_probe_kernel<<<blocks, threads, 0, stream>>>(sink, iters);
  return cudaGetLastError();
}

size_t model_smem_bytes(const ssppm::DevModelHead& M) {
  size_t b = (size_t)M.n_mbody * sizeof(DevBody);
  if (M.sdf_mode) return b + (size_t)M.n_sphere * sizeof(DevSphere);
  return b + (size_t)M.n_mgeom * sizeof(DevMGeom) + (size_t)M.n_sgeom * sizeof(DevSGeom) + (size_t)(M.n_pair + M.n_bpair) * sizeof(DevPair) +
         (size_t)M.n_mmpair * sizeof(DevMMPair);
}

size_t iterate_smem_bytes(const ssppm::DevModelHead& M, const IterParams& P, int threads) {
  const ScoreConfig& c = P.cfg;
  const int wc = row_pitch(P.w_col), wa = row_pitch(P.w_arc), dp = up4(c.D);
  size_t f = (size_t)up4(c.Mc * wc) + up4(c.Mc) + (c.arc_same ? 0 : (size_t)up4(c.Ma * wa) + up4(c.Ma));
  f += (size_t)P.cand_per_cta * c.K * dp;
  f += (size_t)P.cand_per_cta * c.Mc * 2;
  f += (size_t)P.cand_per_cta * P.mq;
  f += (size_t)c.D * threads;
  if (!M.chain) f += (size_t)M.n_mbody * 12 * threads;
  if (!P.coop) f += 2 * (size_t)P.n_pad;
  return f * sizeof(float);
}

namespace {
using IterKernel = void (*)(const ModelView, const IterParams);
struct VariantFns { IterKernel coop, solo, coop_mirror; };
template <int MAXT, int MINB>
constexpr VariantFns variant_fns() {
  return {sspp_iterate_kernel<true, MAXT, MINB>, sspp_iterate_kernel<false, MAXT, MINB>, sspp_iterate_kernel<true, MAXT, MINB, true>};
}
// same order as kIterVariants
#ifdef SSPP_FAST_BUILD   // kernel experiments: only the two default variants are compiled (the others alias them)
const VariantFns kFns[kNumIterVariants] = {variant_fns<256, 2>(), variant_fns<256, 2>(), variant_fns<320, 2>(), variant_fns<320, 2>(),
                                           variant_fns<320, 2>(), variant_fns<320, 2>(), variant_fns<256, 2>(), variant_fns<256, 2>()};
#else
const VariantFns kFns[kNumIterVariants] = {variant_fns<256, 2>(), variant_fns<256, 3>(), variant_fns<320, 2>(), variant_fns<320, 3>(),
                                           variant_fns<448, 2>(), variant_fns<384, 2>(), variant_fns<128, 7>(), variant_fns<192, 5>()};
#endif
cudaError_t set_smem(IterKernel k, size_t smem_bytes) {   // per device / context, so no caching across calls
  if (smem_bytes <= 48 * 1024) return cudaSuccess;
  return cudaFuncSetAttribute(reinterpret_cast<const void*>(k), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
}
}  // namespace

cudaError_t iterate_occupancy(int variant, int threads, size_t smem_bytes, int* ctas_per_sm) {
  if (variant < 0 || variant >= kNumIterVariants || threads > kIterVariants[variant].max_threads) return cudaErrorInvalidValue;
  cudaError_t e = set_smem(kFns[variant].coop, smem_bytes);
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, reinterpret_cast<const void*>(kFns[variant].coop), threads, smem_bytes);
}

cudaError_t launch_iterate(const ModelView& M, const IterParams& P, int variant, int threads, size_t smem_bytes, int grid,
                           cudaStream_t stream) {
  if (variant < 0 || variant >= kNumIterVariants || threads > kIterVariants[variant].max_threads) return cudaErrorInvalidValue;
  if (!P.coop) {
    cudaError_t e = set_smem(kFns[variant].solo, smem_bytes);
    if (e != cudaSuccess) return e;
    kFns[variant].solo<<<(unsigned)P.n_tiles, threads, smem_bytes, stream>>>(M, P);
    return cudaGetLastError();
  }
  const IterKernel k = P.host_out ? kFns[variant].coop_mirror : kFns[variant].coop;
  cudaError_t e = set_smem(k, smem_bytes);
  if (e != cudaSuccess) return e;
  void* args[] = {const_cast<ModelView*>(&M), const_cast<IterParams*>(&P)};
  return cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(k), dim3((unsigned)grid), dim3((unsigned)threads),
                                     args, smem_bytes, stream);
}

cudaError_t launch_eval_points(const ModelView& M, int D, const float* pts, int n, float* out, cudaStream_t stream) {
  const int T = 128;
  const size_t smem = (size_t)(D + (M.chain ? 0 : M.n_mbody * 12)) * T * sizeof(float);
  static size_t configured = 0;
  if (smem > configured && smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sspp_eval_points_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  sspp_eval_points_kernel<<<(n + T - 1) / T, T, smem, stream>>>(M, D, pts, n, out);
  return cudaGetLastError();
}

cudaError_t launch_check_points(const ModelView& M, int D, const float* pts, int n, const int* mg_src, const int* sg_src, int* geom1, int* geom2,
                                float* dist, cudaStream_t stream) {
  const int T = 128;
  const size_t smem = (size_t)(D + (M.chain ? 0 : M.n_mbody * 12)) * T * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sspp_check_points_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  sspp_check_points_kernel<<<(n + T - 1) / T, T, smem, stream>>>(M, D, pts, n, mg_src, sg_src, geom1, geom2, dist);
  return cudaGetLastError();
}

cudaError_t launch_sdf_gradient(const ModelView& M, int D, int K, int rows, const float* W, const float* pts, float* dist, float* dirs,
                                float* G, cudaStream_t stream) {
  const int T = 128;
  const size_t smem = (size_t)(D + (M.chain ? 0 : M.n_mbody * 12)) * T * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sspp_sdf_gradient_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  sspp_sdf_gradient_kernel<<<1, T, smem, stream>>>(M, D, K, rows, W, pts, dist, dirs, G);
  return cudaGetLastError();
}

cudaError_t launch_bspline_eval(const float* rows, int T, int n, const float* ctrl, int B, int d, float* out, int n_sm, cudaStream_t stream) {
  const size_t smem = (size_t)T * n * sizeof(float);
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sspp_bspline_eval_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const long long total = (long long)B * T;
  const int blocks = (int)std::max<long long>(1, std::min<long long>((total + 255) / 256, (long long)n_sm * 8));
  sspp_bspline_eval_kernel<<<blocks, 256, smem, stream>>>(rows, T, n, ctrl, B, d, out);
  return cudaGetLastError();
}

cudaError_t launch_success_rows(const IterParams& P, float* rows, int32_t* idx, int32_t* count, cudaStream_t stream) {
  const int nfd = P.cfg.n_free * P.cfg.D;
  sspp_success_rows_kernel<<<dim3((unsigned)((P.cfg.n_cand + 255) / 256), (unsigned)P.n_query), 256, 0, stream>>>(
      P.status, P.params, P.cfg.n_cand, nfd, rows, idx, count);
  return cudaGetLastError();
}

cudaError_t launch_count(const ModelView& M, const IterParams& P, unsigned long long* totals, int n_sm, cudaStream_t stream) {
  const int T = 128;
  const size_t smem = (size_t)(P.cfg.D + (M.chain ? 0 : M.n_mbody * 12)) * T * sizeof(float);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(sspp_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  sspp_count_kernel<<<n_sm * 4, T, smem, stream>>>(M, P, totals);
  return cudaGetLastError();
}

cudaError_t launch_merge_update(const MergeParams& P, cudaStream_t stream) {
  const int nfd = P.cfg.n_free * P.cfg.D;
  sspp_merge_rank_kernel<<<(std::max(P.n_keys, 1) + 255) / 256, 256, 0, stream>>>(P);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  sspp_merge_finish_kernel<<<nfd + 1, 256, 0, stream>>>(P);
  return cudaGetLastError();
}

}  // namespace sspp

// ------------------------------------------------------------------------------------------------------------------------
// Allocation-free variant of the batched B-spline evaluation (ssppcu_bspline_eval_batch_noalloc).  No counterpart in the
// reference; added for callers that cannot call cudaMalloc (CUDA graph capture,
// tight serving loops, harnesses that forbid it).  The device staging that ssppcu_bspline_eval_batch cudaMallocs per
// call (rows | control points | output) lives in this file-scope array instead: kStaticScratchFloats * 4 = 4 MiB, held by
// every GPU that loads the library until the process exits.
//   * In HOST code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is looked
//     up once per device, with that device current, by cudaGetSymbolAddress and cached in g_static_ptr[].
//   * Every call on a device shares the one array and finds the previous call's contents in it.  Nothing has to start
//     at zero: the two input regions are overwritten by the copies and the kernel stores every output element before
//     the read-back.  Calls that overlap on the same GPU would race on it; inside one process a per-device mutex
//     serialises them (the call is synchronous), across processes each has its own copy of the array.
// ------------------------------------------------------------------------------------------------------------------------
static __device__ __align__(256) float g_static_scratch[sspp::kStaticScratchFloats];

namespace sspp {
namespace {
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu[kStaticMaxDevices];
float* g_static_ptr[kStaticMaxDevices];          // device address of g_static_scratch on device i; nullptr = not looked up yet
int g_static_sms[kStaticMaxDevices];
}  // namespace

size_t static_scratch_floats_needed(int T, int n, int B, int d) {
  auto up = [](size_t x) { return (x + 63) / 64 * 64; };                    // every region starts on a 256-byte boundary
  return up((size_t)T * n) + up((size_t)B * n * d) + up((size_t)B * T * d);
}

cudaError_t bspline_eval_static(int device, const float* h_rows, int T, int n, const float* h_ctrl, int B, int d, float* h_out) {
  if (device < 0 || device >= kStaticMaxDevices) return cudaErrorInvalidDevice;
  if (static_scratch_floats_needed(T, n, B, d) > (size_t)kStaticScratchFloats) return cudaErrorInvalidValue;
  int prev = -1;
  cudaError_t e = cudaGetDevice(&prev);
  if (e != cudaSuccess) return e;
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return e;
  {
    std::lock_guard<std::mutex> lock(g_static_mu[device]);
    if (!g_static_ptr[device]) {                                             // set-up, once per device, that device current
      float* p = nullptr;
      e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch);
      if (e == cudaSuccess && !p) e = cudaErrorInvalidSymbol;
      if (e == cudaSuccess) e = cudaDeviceGetAttribute(&g_static_sms[device], cudaDevAttrMultiProcessorCount, device);
      if (e == cudaSuccess) g_static_ptr[device] = p;
    }
    if (e == cudaSuccess) {
      auto up = [](size_t x) { return (x + 63) / 64 * 64; };
      const size_t nr = (size_t)T * n, nc = (size_t)B * n * d, no = (size_t)B * T * d;
      float* d_rows = g_static_ptr[device];
      float* d_ctrl = d_rows + up(nr);
      float* d_out = d_ctrl + up(nc);
      e = cudaMemcpy(d_rows, h_rows, nr * sizeof(float), cudaMemcpyHostToDevice);
      if (e == cudaSuccess) e = cudaMemcpy(d_ctrl, h_ctrl, nc * sizeof(float), cudaMemcpyHostToDevice);
      if (e == cudaSuccess) e = launch_bspline_eval(d_rows, T, n, d_ctrl, B, d, d_out, g_static_sms[device], 0);
      if (e == cudaSuccess) e = cudaMemcpy(h_out, d_out, no * sizeof(float), cudaMemcpyDeviceToHost);
    }
  }
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  return e;
}

}  // namespace sspp
