// sspp_kernels.cu — the fused scoring iteration for sm_100a: one launch per planner iteration.
//
//   K1 perturb     Philox4x32-10 + Box-Muller noise on the sampled rows      (sspp.h:114-130, tsp_sampler.h:12-51)
//   K2 spline      q_m = Psi[m,:] . X   (X: K x D point matrix in shared memory)  (Eigen::Spline::operator(), sspp.h:137)
//   K3 kinematics  link frames from q, per thread                            (mj_kinematics via sspp.h:139-143, utility.h:149-175)
//   K4 collision   moving geoms against the scene primitives                 (mj_collision via sspp.h:144, Collision.h:84-103)
//   K5 reduce      arc length, collision and penalty sums per candidate      (sspp.h:152-169, tsp_evaluator.h:17-32)
//   K6 elites      bitonic sort of (cost,index) keys in the last CTA, CES weights (sspp.h:171-192, tsp_elites.h:13-32)
//   K7 update      weighted mean / EMA variance / sigma adaptation           (tsp_distribution.h:31-83)
//
// Work decomposition: a CTA owns `cand_per_cta` consecutive candidates of one query; its threads
// stride over the flattened (candidate, collocation point) items, so every lane carries one
// full FK + collision evaluation at a time whatever M is.  Per-point results go to shared
// memory, one warp per candidate then reduces them in a fixed order (lane-strided partial sums,
// xor butterfly) that the oracles restate.  The last CTA of a query to finish (atomic ticket)
// sorts the keys and updates the sampling distribution, so an iteration is a single launch;
// queries small enough for one CTA can run many iterations inside one launch.
#include "sspp_kernels.h"

#include <cuda_runtime.h>

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

__device__ __forceinline__ float warp_strided_sum(const float* v, int n, int lane) {
  float acc = 0.0f;
  for (int i = lane; i < n; i += 32) acc += v[i];
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, m);
  return acc;
}
__device__ __forceinline__ float warp_butterfly(float acc) {
#pragma unroll
  for (int m = 16; m > 0; m >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, m);
  return acc;
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

// Elite selection + distribution update for one query, executed by one CTA.
//   keys       sorted in place (shared or global memory), n_pad entries, unused tail = ~0
//   get_param  (global candidate index, element t) -> sampled value
template <class ParamGetter>
__device__ void elites_and_update(const ScoreConfig& c, unsigned long long* keys, int n_pad, float* elite_w,
                                  float* mean, float* sigma, float* best, int* best_valid,
                                  int* n_success_out, int* n_elite_out, int* best_idx_out, float* best_cost_out,
                                  int* any_success, ParamGetter get_param, int* s_int, float* s_flt) {
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = T >> 5;
  block_bitonic_sort(keys, n_pad);
  if (tid == 0) s_int[0] = 0;
  __syncthreads();
  int cnt = 0;
  for (int i = tid; i < n_pad; i += T) cnt += ((unsigned)(keys[i] >> 32) != 0xFFFFFFFFu) ? 1 : 0;
  cnt = __reduce_add_sync(0xFFFFFFFFu, cnt);
  if (lane == 0 && cnt) atomicAdd(&s_int[0], cnt);
  __syncthreads();
  const int n_succ = s_int[0];
  const int k = elite_count(n_succ, c.elite_frac);
  const int nfd = c.n_free * c.D;
  if (tid == 0) {
    *n_success_out = n_succ; *n_elite_out = k;
    if (n_succ > 0) {
      const unsigned long long k0 = keys[0];
      *best_idx_out = (int)(unsigned)(k0 & 0xFFFFFFFFu);
      unsigned ob = (unsigned)(k0 >> 32);
      ob = (ob & 0x80000000u) ? (ob & 0x7FFFFFFFu) : ~ob;
      *best_cost_out = __uint_as_float(ob);
      *any_success = 1;
    } else { *best_idx_out = -1; *best_cost_out = __int_as_float(0x7F800000); }
  }
  if (n_succ == 0) {
    if (c.do_update) for (int t = tid; t < nfd; t += T) sigma[t] = clamp_sigma(c, sigma[t] * c.inc);   // adapt(false)
    __syncthreads();
    return;
  }
  // CES log-weights (tsp_elites.h:24-32)
  const float lk = log_poly((float)k + 0.5f);
  for (int j = tid; j < k; j += T) elite_w[j] = lk - log_poly((float)j + 1.0f);
  __syncthreads();
  if (warp == 0) { const float s = warp_strided_sum(elite_w, k, lane); if (lane == 0) s_flt[0] = s; }
  __syncthreads();
  const float sumw = s_flt[0];
  for (int j = tid; j < k; j += T) elite_w[j] = elite_w[j] / sumw;
  __syncthreads();
  if (c.do_update) {
    for (int t = warp; t < nfd; t += nwarp) {       // one warp per distribution element
      const int d = t % c.D;
      float acc = 0.0f;
      for (int j = lane; j < k; j += 32) acc = fmaf(elite_w[j], get_param((unsigned)(keys[j] & 0xFFFFFFFFu), t), acc);
      const float em = warp_butterfly(acc);
      const float m0 = mean[t];
      float nm = fmaf(c.mean_lr, em - m0, m0);
      if (c.mode == MODE_TASK && d == 2) nm = fmaxf(nm, c.dist_zmin);
      nm = fminf(fmaxf(nm, c.lo[d]), c.hi[d]);
      const bool wrap = (c.mode == MODE_TASK && d == 3 && c.lo[3] != c.hi[3]);
      float vacc = 0.0f;
      for (int j = lane; j < k; j += 32) {
        const float v = get_param((unsigned)(keys[j] & 0xFFFFFFFFu), t);
        const float df = wrap ? wrap_angle_diff(v, nm, c.lo[3], c.hi[3]) : v - nm;
        vacc = fmaf(elite_w[j], df * df, vacc);
      }
      const float var = warp_butterfly(vacc);
      const float s0 = sigma[t];
      const float blend = fmaf(c.var_beta, var, (1.0f - c.var_beta) * (s0 * s0));
      float sg = clamp_sigma(c, sqrtf(blend));
      sg = clamp_sigma(c, sg * c.dec);               // adapt(true)
      if (lane == 0) { mean[t] = nm; sigma[t] = sg; }
    }
  }
  // forward the best candidate (last_best_)
  const unsigned bi = (unsigned)(keys[0] & 0xFFFFFFFFu);
  for (int t = tid; t < nfd; t += T) best[t] = get_param(bi, t);
  if (tid == 0) *best_valid = 1;
  __syncthreads();
}

struct StoredParams {
  const float* params; int nfd;
  __device__ __forceinline__ float operator()(unsigned idx, int t) const { return params[(size_t)idx * nfd + t]; }
};

__global__ void __launch_bounds__(256, 2)
sspp_iterate_kernel(const __grid_constant__ DevModel M, const __grid_constant__ IterParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_int[4];
  __shared__ float s_flt[4];
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x, lane = tid & 31, warp = tid >> 5, nwarp = T >> 5;
  const int query = blockIdx.y;
  const int c0 = blockIdx.x * P.cand_per_cta;
  const int nc = min(P.cand_per_cta, c.n_cand - c0);
  const int D = c.D, K = c.K, Mc = c.Mc, Ma = c.Ma, Kp = P.k_pad, Mq = P.mq, CT = P.cand_per_cta;
  const int nfd = c.n_free * D;
  const int Marc = c.arc_same ? Mc : Ma;

  // ---- shared memory carve-up (all float) ----
  float* sm = reinterpret_cast<float*>(smem_raw);
  float* s_psi_col = sm;              sm += Mc * Kp;
  float* s_psi_arc = sm;              sm += c.arc_same ? 0 : Ma * Kp;
  float* s_X = sm;                    sm += CT * K * D;
  float* s_q = sm;                    sm += CT * Mq * D;
  float* s_cc = sm;                   sm += CT * Mc;     // collision cost per point
  float* s_cw = sm;                   sm += CT * Mc;     // collision cost + penalty per point
  float* s_seg = sm;                  sm += CT * Mq;     // segment lengths
  float* s_qs = sm;                   sm += D * T;       // coordinates of the point in flight
  float* s_xf = sm;                                       // link frames (only when !chain)

  for (int e = tid; e < Mc * K; e += T) s_psi_col[(e / K) * Kp + (e % K)] = P.psi_col[e];
  if (!c.arc_same) for (int e = tid; e < Ma * K; e += T) s_psi_arc[(e / K) * Kp + (e % K)] = P.psi_arc[e];

  float* mean = P.mean + (size_t)query * nfd;
  float* sigma = P.sigma + (size_t)query * nfd;
  float* best = P.best + (size_t)query * nfd;
  const float* base = P.base + (size_t)query * K * D;
  const size_t qn = (size_t)query * c.n_cand;
  SmemXF xf{s_xf, T, tid};
  SmemQ qa{s_qs, T, tid};

  for (int it = 0; it < P.n_inner; ++it) {
    const uint32_t iter = P.iter0 + (uint32_t)it;
    const int best_ok = c.include_best ? P.best_valid[query] : 0;
    // ---- K1: candidate point matrices ----
    for (int e = tid; e < nc * K * D; e += T) s_X[e] = base[e % (K * D)];
    __syncthreads();
    for (int e = tid; e < nc * nfd; e += T) {
      const int cl = e / nfd, t = e - cl * nfd;
      const int cand = c0 + cl, gcand = P.cand_offset + cand;
      const int kind = candidate_kind(c, gcand, best_ok);
      const float v = candidate_param(c, P.seed, iter, (uint32_t)query, gcand, kind, t, mean, sigma, best);
      s_X[(cl * K + c.free_begin) * D + t] = v;
      P.params[(qn + cand) * nfd + t] = v;
    }
    __syncthreads();
    // ---- K2-K4: spline point, kinematics, collision per (candidate, point) item ----
    for (int w = tid; w < nc * Mc; w += T) {
      const int cl = w / Mc, m = w - cl * Mc;
      const float* psi = s_psi_col + m * Kp;
      const float* X = s_X + cl * K * D;
      for (int d = 0; d < D; ++d) {
        float acc = 0.0f;
        for (int k = 0; k < K; ++k) acc = fmaf(psi[k], X[k * D + d], acc);
        s_qs[d * T + tid] = acc;
        if (c.arc_same) s_q[(cl * Mq + m) * D + d] = acc;
      }
      float cc = 0.0f, cw = 0.0f;
      if (m >= c.col_first) {
        cc = point_collision_cost(M, qa, xf);
        cw = cc + point_penalty(c, qa);
      }
      s_cc[cl * Mc + m] = cc;
      s_cw[cl * Mc + m] = cw;
    }
    if (!c.arc_same) {
      for (int w = tid; w < nc * Ma; w += T) {
        const int cl = w / Ma, m = w - cl * Ma;
        const float* psi = s_psi_arc + m * Kp;
        const float* X = s_X + cl * K * D;
        for (int d = 0; d < D; ++d) {
          float acc = 0.0f;
          for (int k = 0; k < K; ++k) acc = fmaf(psi[k], X[k * D + d], acc);
          s_q[(cl * Mq + m) * D + d] = acc;
        }
      }
    }
    __syncthreads();
    // ---- K5: segment lengths, then one warp per candidate reduces ----
    for (int w = tid; w < nc * (Marc - 1); w += T) {
      const int cl = w / (Marc - 1), m = 1 + (w - cl * (Marc - 1));
      const float* a = s_q + (cl * Mq + m) * D;
      float s2 = 0.0f;
      for (int d = 0; d < D; ++d) { const float df = a[d] - a[d - D]; s2 = fmaf(df, df, s2); }
      s_seg[cl * Mq + m - 1] = sqrtf(s2);
    }
    __syncthreads();
    for (int cl = warp; cl < nc; cl += nwarp) {
      const float L = warp_strided_sum(s_seg + cl * Mq, Marc - 1, lane);
      const float Cnf = warp_strided_sum(s_cc + cl * Mc + c.col_first, Mc - c.col_first, lane);
      const float Cwf = warp_strided_sum(s_cw + cl * Mc + c.col_first, Mc - c.col_first, lane);
      if (lane == 0) {
        const int cand = c0 + cl;
        const int kind = candidate_kind(c, P.cand_offset + cand, best_ok);
        const int st = (kind == CK_ABSENT) ? ST_ABSENT : ((Cnf == 0.0f) ? ST_CONVERGED : ST_FAILED);
        const float cost = fmaf(c.w_collision, Cwf, L);
        P.L[qn + cand] = L; P.Cnf[qn + cand] = Cnf; P.Cwf[qn + cand] = Cwf; P.cost[qn + cand] = cost;
        P.status[qn + cand] = (uint8_t)st;
        P.keys[(size_t)query * P.n_pad + cand] = sort_key(cost, st, (uint32_t)(P.cand_offset + cand));
      }
    }
    // ---- K6/K7 in the last CTA of this query ----
    __threadfence();
    __syncthreads();
    if (tid == 0) s_int[1] = (atomicAdd(&P.ticket[query], 1u) == (unsigned)(P.ctas_per_query - 1)) ? 1 : 0;
    __syncthreads();
    if (s_int[1]) {
      __threadfence();
      unsigned long long* gkeys = P.keys + (size_t)query * P.n_pad;
      unsigned long long* keys = gkeys;
      if (P.sort_in_smem) {
        keys = reinterpret_cast<unsigned long long*>(smem_raw);
        for (int i = tid; i < P.n_pad; i += T) keys[i] = (i < c.n_cand) ? __ldcg(gkeys + i) : ~0ull;
      } else {
        for (int i = c.n_cand + tid; i < P.n_pad; i += T) gkeys[i] = ~0ull;
      }
      __syncthreads();
      StoredParams gp{P.params + qn * nfd, nfd};
      if (P.sort_only) block_bitonic_sort(keys, P.n_pad);
      else elites_and_update(c, keys, P.n_pad, P.elite_w + (size_t)query * P.n_pad, mean, sigma, best, P.best_valid + query,
                        P.n_success + query, P.n_elite + query, P.best_idx + query, P.best_cost + query,
                        P.any_success + query, gp, s_int, s_flt);
      if (P.sort_in_smem) for (int i = tid; i < P.n_pad; i += T) gkeys[i] = keys[i];
      if (tid == 0) P.ticket[query] = 0u;
      __syncthreads();
      if (P.sort_in_smem && P.n_inner > 1) {   // the key buffer aliases the basis matrices: reload them
        for (int e = tid; e < Mc * K; e += T) s_psi_col[(e / K) * Kp + (e % K)] = P.psi_col[e];
        if (!c.arc_same) for (int e = tid; e < Ma * K; e += T) s_psi_arc[(e / K) * Kp + (e % K)] = P.psi_arc[e];
      }
    }
    __syncthreads();
  }
}

// regenerate a candidate's sampled element from the Philox counters (population-sharded path)
struct RegenParams {
  const ScoreConfig* c; uint64_t seed; uint32_t iter; const float* mean0; const float* sigma0; const float* best0; int best_ok;
  __device__ __forceinline__ float operator()(unsigned idx, int t) const {
    const int kind = candidate_kind(*c, (int)idx, best_ok);
    return candidate_param(*c, seed, iter, 0u, (int)idx, kind, t, mean0, sigma0, best0);
  }
};

__global__ void __launch_bounds__(1024, 1)
sspp_merge_update_kernel(const __grid_constant__ MergeParams P) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ int s_int[4];
  __shared__ float s_flt[4];
  __shared__ float s_mean0[MAX_K * MAX_DOF], s_sigma0[MAX_K * MAX_DOF], s_best0[MAX_K * MAX_DOF];
  const ScoreConfig& c = P.cfg;
  const int tid = threadIdx.x, T = blockDim.x;
  const int nfd = c.n_free * c.D;
  for (int t = tid; t < nfd; t += T) { s_mean0[t] = P.mean[t]; s_sigma0[t] = P.sigma[t]; s_best0[t] = P.best[t]; }
  for (int i = P.n_keys + tid; i < P.n_pad; i += T) P.keys[i] = ~0ull;
  __syncthreads();
  RegenParams gp{&c, P.seed, P.iter, s_mean0, s_sigma0, s_best0, c.include_best ? *P.best_valid : 0};
  elites_and_update(c, P.keys, P.n_pad, P.elite_w, P.mean, P.sigma, P.best, P.best_valid, P.n_success, P.n_elite,
                    P.best_idx, P.best_cost, P.any_success, gp, s_int, s_flt);
}

__global__ void __launch_bounds__(128)
sspp_eval_points_kernel(const __grid_constant__ DevModel M, int D, const float* __restrict__ pts, int n, float* __restrict__ out) {
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

cudaError_t launch_fma_probe(float* sink, int blocks, int threads, int iters, cudaStream_t stream) {
  sspp_fma_probe_kernel<<<blocks, threads, 0, stream>>>(sink, iters);
  return cudaGetLastError();
}

size_t iterate_smem_bytes(const DevModel& M, const IterParams& P, int threads) {
  const ScoreConfig& c = P.cfg;
  size_t f = (size_t)c.Mc * P.k_pad + (c.arc_same ? 0 : (size_t)c.Ma * P.k_pad);
  f += (size_t)P.cand_per_cta * c.K * c.D;
  f += (size_t)P.cand_per_cta * P.mq * c.D;
  f += (size_t)P.cand_per_cta * c.Mc * 2;
  f += (size_t)P.cand_per_cta * P.mq;
  f += (size_t)c.D * threads;
  if (!M.chain) f += (size_t)M.n_mbody * 12 * threads;
  return f * sizeof(float);
}

cudaError_t launch_iterate(const DevModel& M, const IterParams& P, int threads, size_t smem_bytes, cudaStream_t stream) {
  static size_t configured = 0;
  if (smem_bytes > configured) {
    cudaError_t e = cudaFuncSetAttribute(sspp_iterate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
    if (e != cudaSuccess) return e;
    configured = smem_bytes;
  }
  dim3 grid((unsigned)P.ctas_per_query, (unsigned)P.n_query, 1);
  sspp_iterate_kernel<<<grid, threads, smem_bytes, stream>>>(M, P);
  return cudaGetLastError();
}

cudaError_t launch_eval_points(const DevModel& M, int D, const float* pts, int n, float* out, cudaStream_t stream) {
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

cudaError_t launch_merge_update(const MergeParams& P, cudaStream_t stream) {
  sspp_merge_update_kernel<<<1, 1024, 0, stream>>>(P);
  return cudaGetLastError();
}

}  // namespace sspp
