// spline_ab.cu - A/B evidence for one design decision (north_star: "use tensor cores only if ncu shows they beat
// FP32 FMA at this small K"): the spline evaluation q[cand][m][:] = Phi[m][:] . C[cand][:][:] of cfg2 (4096 candidates, M = 65
// collocation points, K = 10 control points, D = 7 coordinates; Phi is the clamped cubic B-spline collocation matrix, 4 non-zeros
// per row) done three ways, every result consumed in registers (one float per point goes to memory) as in the fused kernel:
//   A   FP32 FMA over the 4-wide band, one thread per point           - what sspp_iterate_kernel does (spline_point_any)
//   B   TF32 tensor cores, mma.sync.m16n8k8, dense K padded to 16, D padded to 8, one warp per candidate
//   B3  the same with the 3 x TF32 split (big + small parts of both operands, 3 MMAs per step)
// Reported per variant: time per launch (CUDA events, 200 launches), max relative error of q against an FP64 host evaluation.
// ncu on this binary gives the tensor / FMA pipe utilisation of each kernel (scripts/spline_ab.sh).
// mma.sync is the legacy tensor path on sm_100a (tcgen05 is the fast one); for a 65 x 16 x 8 product per candidate the issue
// is not tensor throughput but precision and the share of the step the contraction has at all - see DESIGN.md.
#include <cuda_runtime.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

constexpr int N = 4096, M = 65, K = 10, D = 7, P = 3, MP = 80, KP = 16, DP = 8;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); std::exit(1); } } while (0)

// Cox - de Boor on the clamped uniform knot vector of K control points, degree P
static std::vector<double> knots() {
  std::vector<double> t(K + P + 1);
  for (int i = 0; i < K + P + 1; ++i) t[i] = i <= P ? 0.0 : (i >= K ? 1.0 : (double)(i - P) / (K - P));
  return t;
}
static double basis(const std::vector<double>& t, int i, int p, double u) {
  if (p == 0) return ((t[i] <= u && u < t[i + 1]) || (u == 1.0 && t[i + 1] == 1.0 && t[i] < 1.0)) ? 1.0 : 0.0;
  double a = 0.0, b = 0.0;
  if (t[i + p] > t[i]) a = (u - t[i]) / (t[i + p] - t[i]) * basis(t, i, p - 1, u);
  if (t[i + p + 1] > t[i + 1]) b = (t[i + p + 1] - u) / (t[i + p + 1] - t[i + 1]) * basis(t, i + 1, p - 1, u);
  return a + b;
}

__constant__ float c_band[M * 4];
__constant__ int c_first[M];

// A: one thread per (candidate, point); ctrl [N][K][DP] (rows padded to 8 floats like the kernel's shared-memory rows)
__global__ void __launch_bounds__(256) eval_fma(const float* __restrict__ ctrl, float* __restrict__ out, float* __restrict__ qdump) {
  const int item = blockIdx.x * blockDim.x + threadIdx.x;
  if (item >= N * M) return;
  const int cand = item / M, m = item - cand * M;
  const int f = c_first[m];
  const float4* c4 = reinterpret_cast<const float4*>(ctrl + ((size_t)cand * K + f) * DP);
  float q[DP];
#pragma unroll
  for (int d = 0; d < DP; ++d) q[d] = 0.0f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float w = c_band[m * 4 + i];
    const float4 lo = c4[2 * i], hi = c4[2 * i + 1];
    q[0] = fmaf(w, lo.x, q[0]); q[1] = fmaf(w, lo.y, q[1]); q[2] = fmaf(w, lo.z, q[2]); q[3] = fmaf(w, lo.w, q[3]);
    q[4] = fmaf(w, hi.x, q[4]); q[5] = fmaf(w, hi.y, q[5]); q[6] = fmaf(w, hi.z, q[6]);
  }
  float s = 0.0f;
#pragma unroll
  for (int d = 0; d < D; ++d) s = fmaf(q[d], q[d], s);
  out[item] = s;
  if (qdump) for (int d = 0; d < D; ++d) qdump[(size_t)item * D + d] = q[d];
}

__device__ __forceinline__ unsigned to_tf32(float x) { unsigned r; asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x)); return r; }
__device__ __forceinline__ void mma_tf32(float (&c)[4], const unsigned (&a)[4], const unsigned (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// B / B3: one warp per candidate; phi [MP][KP] dense FP32 in shared memory (zero padded), ctrl [N][K][DP]
template <bool SPLIT>
__global__ void __launch_bounds__(256) eval_mma(const float* __restrict__ phi, const float* __restrict__ ctrl, float* __restrict__ out,
                                                float* __restrict__ qdump) {
  __shared__ float s_phi[MP * KP];
  for (int e = threadIdx.x; e < MP * KP; e += blockDim.x) s_phi[e] = phi[e];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cand = blockIdx.x * (blockDim.x >> 5) + warp;
  if (cand >= N) return;
  const int g = lane >> 2, t = lane & 3;
  // B fragments of the candidate: k-step ks covers control points 8 ks .. 8 ks + 7; b0 = C[8ks + t][g], b1 = C[8ks + t + 4][g]
  unsigned bh[2][2], bl[2][2];
#pragma unroll
  for (int ks = 0; ks < 2; ++ks)
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int k = 8 * ks + t + 4 * r;
      const float v = (k < K && g < D) ? ctrl[((size_t)cand * K + k) * DP + g] : 0.0f;
      bh[ks][r] = to_tf32(v);
      bl[ks][r] = SPLIT ? to_tf32(v - __uint_as_float(bh[ks][r])) : 0u;
    }
#pragma unroll
  for (int mt = 0; mt < MP / 16; ++mt) {
    float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int ks = 0; ks < 2; ++ks) {
      unsigned ah[4], al[4];
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        const float v = s_phi[(16 * mt + g + 8 * (r & 1)) * KP + 8 * ks + t + 4 * (r >> 1)];
        ah[r] = to_tf32(v);
        al[r] = SPLIT ? to_tf32(v - __uint_as_float(ah[r])) : 0u;
      }
      if (SPLIT) { mma_tf32(c, al, bh[ks]); mma_tf32(c, ah, bl[ks]); }
      mma_tf32(c, ah, bh[ks]);
    }
    // c0 = q[row g][col 2t], c1 = [g][2t+1], c2 = [g+8][2t], c3 = [g+8][2t+1]: the squares of a row's 7 coordinates are summed over
    // the 4 lanes that hold them
    float s0 = (2 * t < D ? c[0] * c[0] : 0.f) + (2 * t + 1 < D ? c[1] * c[1] : 0.f);
    float s1 = (2 * t < D ? c[2] * c[2] : 0.f) + (2 * t + 1 < D ? c[3] * c[3] : 0.f);
    s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, 1); s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, 2);
    s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, 1); s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, 2);
    const int m0 = 16 * mt + g, m1 = m0 + 8;
    if (t == 0) { if (m0 < M) out[(size_t)cand * M + m0] = s0; if (m1 < M) out[(size_t)cand * M + m1] = s1; }
    if (qdump) {
      if (m0 < M) { if (2 * t < D) qdump[((size_t)cand * M + m0) * D + 2 * t] = c[0]; if (2 * t + 1 < D) qdump[((size_t)cand * M + m0) * D + 2 * t + 1] = c[1]; }
      if (m1 < M) { if (2 * t < D) qdump[((size_t)cand * M + m1) * D + 2 * t] = c[2]; if (2 * t + 1 < D) qdump[((size_t)cand * M + m1) * D + 2 * t + 1] = c[3]; }
    }
  }
}

int main() {
  const std::vector<double> t = knots();
  std::vector<double> phi((size_t)M * K);
  std::vector<float> band(M * 4, 0.0f), phif((size_t)MP * KP, 0.0f);
  std::vector<int> first(M);
  for (int m = 0; m < M; ++m) {
    const double u = (double)m / (M - 1);
    int f = -1;
    for (int i = 0; i < K; ++i) { phi[(size_t)m * K + i] = basis(t, i, P, u); if (f < 0 && phi[(size_t)m * K + i] != 0.0) f = i; }
    f = std::min(f, K - 4);
    first[m] = f;
    for (int i = 0; i < 4; ++i) band[m * 4 + i] = (float)phi[(size_t)m * K + f + i];
    for (int i = 0; i < K; ++i) phif[(size_t)m * KP + i] = (float)phi[(size_t)m * K + i];
  }
  std::vector<float> ctrl((size_t)N * K * DP, 0.0f);
  unsigned long long s = 0x53535050ull;
  for (int c = 0; c < N; ++c) for (int k = 0; k < K; ++k) for (int d = 0; d < D; ++d) {
    s = s * 6364136223846793005ull + 1442695040888963407ull;
    ctrl[((size_t)c * K + k) * DP + d] = (float)(((s >> 33) & 0xFFFFFF) / 16777216.0 * 6.2831853 - 3.14159265);   // joint angles in [-pi, pi)
  }
  std::vector<double> ref((size_t)N * M * D);
  for (int c = 0; c < N; ++c) for (int m = 0; m < M; ++m) for (int d = 0; d < D; ++d) {
    double a = 0.0;
    for (int k = 0; k < K; ++k) a += (double)(float)phi[(size_t)m * K + k] * (double)ctrl[((size_t)c * K + k) * DP + d];
    ref[((size_t)c * M + m) * D + d] = a;
  }
  float *d_ctrl, *d_phi, *d_out, *d_q;
  CK(cudaMalloc(&d_ctrl, ctrl.size() * 4)); CK(cudaMalloc(&d_phi, phif.size() * 4)); CK(cudaMalloc(&d_out, (size_t)N * M * 4));
  CK(cudaMalloc(&d_q, (size_t)N * M * D * 4));
  CK(cudaMemcpy(d_ctrl, ctrl.data(), ctrl.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d_phi, phif.data(), phif.size() * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpyToSymbol(c_band, band.data(), band.size() * 4)); CK(cudaMemcpyToSymbol(c_first, first.data(), first.size() * 4));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  std::vector<float> q((size_t)N * M * D);
  const char* names[3] = {"A  fp32 fma, 4-wide band", "B  tf32 mma.sync m16n8k8", "B3 3xtf32 split mma.sync"};
  for (int v = 0; v < 3; ++v) {
    auto launch = [&](float* dump) {
      if (v == 0) eval_fma<<<(N * M + 255) / 256, 256>>>(d_ctrl, d_out, dump);
      else if (v == 1) eval_mma<false><<<N / 8, 256>>>(d_phi, d_ctrl, d_out, dump);
      else eval_mma<true><<<N / 8, 256>>>(d_phi, d_ctrl, d_out, dump);
    };
    launch(d_q); CK(cudaGetLastError()); CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(q.data(), d_q, q.size() * 4, cudaMemcpyDeviceToHost));
    double max_rel = 0.0, max_abs = 0.0;
    for (size_t i = 0; i < q.size(); ++i) {
      const double err = std::fabs((double)q[i] - ref[i]);
      max_abs = std::max(max_abs, err);
      max_rel = std::max(max_rel, err / std::max(std::fabs(ref[i]), 1e-2));
    }
    for (int i = 0; i < 20; ++i) launch(nullptr);
    CK(cudaDeviceSynchronize());
    CK(cudaEventRecord(e0));
    for (int i = 0; i < 200; ++i) launch(nullptr);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms = 0; CK(cudaEventElapsedTime(&ms, e0, e1));
    std::printf("{\"variant\": \"%s\", \"us_per_launch\": %.3f, \"points\": %d, \"max_abs_err\": %.3e, \"max_rel_err\": %.3e}\n", names[v], ms * 1e3 / 200,
                N * M, max_abs, max_rel);
  }
  return 0;
}
