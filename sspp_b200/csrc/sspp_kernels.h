// sspp_kernels.h — launch interface of the fused scoring iteration (sspp_kernels.cu).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "sspp_math.h"

namespace sspp {

// Everything one launch needs besides the scene.  All pointers are device pointers owned by the
// C-ABI handle; layouts are row-major with the query index outermost.
struct IterParams {
  ssppm::ScoreConfig cfg;
  int n_query;
  int cand_per_cta, ctas_per_query;
  int n_tiles;              // n_query * ctas_per_query
  int coop;                 // 1: persistent cooperative grid (grid barriers, grid-wide ranking); 0: one CTA per query
  int n_inner;              // iterations run inside this launch
  int cand_offset;          // global index of local candidate 0 (population sharding), else 0
  int sort_only;            // population shard: sort the local keys, leave elites/update to the merge kernel
  int n_pad;                // n_cand rounded up to a power of two (pitch of the key arrays)
  int rank_chunk;           // keys staged in shared memory per ranking pass (coop)
  int w_col, w_arc;         // band widths of the collocation matrices
  int mq;                   // pitch (rows) of the per-candidate point buffer: max(Mc, Ma)
  int smem_words;           // 4-byte words of dynamic shared memory behind the scene tables (phase scratch)
  int model_bytes;          // scene tables staged at the front of dynamic shared memory (model_smem_bytes)
  uint32_t iter0;           // iteration counter of the first inner iteration (Philox counter word)
  uint64_t seed;
  const float* psi_col;     // [Mc][w_col]  banded collocation matrix of the collision grid
  const int* first_col;     // [Mc]         first column of every row
  const float* psi_arc;     // [Ma][w_arc]  arc-length grid (unused when arc_same)
  const int* first_arc;     // [Ma]
  const float* base;        // [Q][K][D] point matrix before the sampled rows are overwritten
  float* mean;              // [Q][n_free*D]
  float* sigma;             // [Q][n_free*D]
  float* best;              // [Q][n_free*D]  forwarded best candidate (last_best_, tsp_planner.h:134)
  int* best_valid;          // [Q]
  const float* state_in;    // NULL, or a snapshot of the live block mean|sigma|best|best_valid|any_success: the launch's first
                            // iteration samples from it and re-arms the live block with it (fused restore)
  float* params;            // [Q][N][n_free*D] sampled rows of every candidate
  float* L; float* Cnf; float* Cwf; float* cost;   // [Q][N]
  uint8_t* status;          // [Q][N]
  unsigned long long* keys_raw; // [Q][n_pad] (cost bits << 32 | index) by local candidate index
  unsigned long long* keys; // [Q][n_pad] the same keys sorted ascending (tail = ~0)
  float* elite_w;           // [Q][n_pad] CES weights of the first n_elite sorted entries
  int* n_success; int* n_elite; int* best_idx; int* any_success;   // [Q]
  float* best_cost;         // [Q]
  int* succ_acc;            // [Q] scratch: success count accumulated by the ranking phase
  int* tile_counter;        // coop, more tiles than CTAs: tiles beyond the first wave are handed out from this counter (zero between iterations)
  const struct ExchangeParams* xchg;  // population shard with the fused key exchange (device-resident block), else NULL
  unsigned xchg_epoch;      // value this launch publishes in the peers' flag words (monotonic per handle, > 0)
  // result mirror (coop kernel, last inner iteration of a launch; NULL = off): host_out is the device address of a mapped pinned
  // host block laid out like the result block at res_dev, with best | best_valid | any_success behind it at mir_extra
  unsigned char* host_out; const unsigned char* res_dev;
  unsigned mir_bytes, mir_nsucc, mir_nelite, mir_bidx, mir_extra;   // byte offsets (multiples of 16): block end, n_success, n_elite, best_idx
  unsigned long long* probe; // diagnostics, NULL = off: [grid][kProbeSlots] %globaltimer stamps at the phase boundaries (coop kernel)
  // Query inputs carried in the kernel's parameter block (ssppcu_plan_step; inline_words = 0: off): the point matrices
  // [Q][K][D] followed by a state snapshot laid out like state_in.  The launch's first iteration samples from it, re-arms
  // the live state with it and stores the point matrices to `base` - no host-to-device copy precedes the launch.
  int inline_words;
  float inline_blk[256];
};
constexpr int kInlineWords = 256;
constexpr int kStaticScratchFloats = 1 << 20;   // file-scope device scratch of the allocation-free entry point: 4 MiB per GPU

// Fused key exchange of the population-sharded path (sort_only handles, cooperative kernel).  Every rank owns an
// exchange block of 16-byte lines [2 halves][world][shard_len] that its peers map (CUDA IPC across processes, plain
// pointers inside one).  After ranking its shard the kernel stores every sorted key into its line of every peer's
// block as {key low word, epoch, key high word, epoch} with ONE 16-byte store over NVLink - data and flag travel
// together, so there is no fence and no separate signal; the receiving kernel polls each line until both flag
// words carry the epoch, compacts the keys into a local array, then merges and updates in the same launch.
// Halves alternate with the epoch: a peer can be at most one iteration ahead.
constexpr int kMaxExchangeRanks = 16;
struct ExchangeParams {
  int world, rank, shard_len;
  int protocol;                                          // 1: two 8-byte relaxed.sys scalar accesses per line (default); 0: one 16-byte volatile vector access
  int n_global, n_pad;                                   // merged population, pitch of the merged key array
  uint4* peer_lines[kMaxExchangeRanks];                  // base of rank r's block
  const uint4* lines_local;                              // this rank's own block
  unsigned long long* gathered;                          // [world][shard_len] keys received this iteration (local)
  int* succ;                                             // scratch word of the merge
  int* status;                                           // mapped host word: 1 = a peer did not publish before the timeout
  unsigned long long timeout_ns;
  unsigned long long* mkeys; float* elite_w; float* state0;   // merge outputs (MergeParams::keys / elite_w / state0)
  float* vals;                                           // [n_free*D][n_pad] elites' sampled values, regenerated grid-wide before the update
};

// shared-memory bytes of the staged scene tables, and of the scoring phase for the given tile shape (the
// launch needs their sum)
size_t model_smem_bytes(const ssppm::DevModelHead& M);
size_t iterate_smem_bytes(const ssppm::DevModelHead& M, const IterParams& P, int threads);

// Launch-bound variants of the iteration kernel: __launch_bounds__(max_threads, min_ctas) sets the register
// budget (65536 / (max_threads * min_ctas), rounded down to 8) and with it the resident warps per SM.
constexpr int kProbeSlots = 12;  // phase-probe stamps per CTA (0-7: iteration phases, 8-11: fused key exchange)
struct IterVariant { int max_threads, min_ctas; };
constexpr int kNumIterVariants = 8;
constexpr IterVariant kIterVariants[kNumIterVariants] = {{256, 2}, {256, 3}, {320, 2}, {320, 3}, {448, 2}, {384, 2}, {128, 7}, {192, 5}};

// grid size: n_tiles CTAs (coop == 0) or `grid` co-resident CTAs looping over the tiles (coop == 1)
cudaError_t launch_iterate(const ssppm::ModelView& M, const IterParams& P, int variant, int threads, size_t smem_bytes, int grid,
                           cudaStream_t stream);
// co-resident CTAs per SM of the cooperative variant for this launch shape
cudaError_t iterate_occupancy(int variant, int threads, size_t smem_bytes, int* ctas_per_sm);

// Population-sharded path, after the all-gather of every shard's sorted keys: two small launches that
// every rank runs identically.  (1) merge-rank: the global position of a key = its position in its own
// shard + the number of smaller keys in every other shard (binary searches; keys are unique), which
// merges the sorted shards without a sort; (2) finish: elites, CES weights and the distribution update,
// one warp per distribution element; elites' sampled rows are regenerated from the Philox counters, so
// only 8-byte keys ever cross NVLink.
struct MergeParams {
  ssppm::ScoreConfig cfg;      // n_cand = global population
  int n_keys, n_shards, shard_len, n_pad;
  uint32_t iter; uint64_t seed;
  const unsigned long long* gathered;   // [n_shards][shard_len], each shard ascending
  unsigned long long* keys;    // [n_pad] merged ascending order (tail ~0)
  float* elite_w;              // [n_pad]
  float* state0;               // [3 * nfd + 1] snapshot of mean | sigma | best | best_valid taken by (1)
  int* succ_acc;               // zeroed by the caller before (1)
  float* mean; float* sigma; float* best; int* best_valid;
  int* n_success; int* n_elite; int* best_idx; float* best_cost; int* any_success;
};
cudaError_t launch_merge_update(const MergeParams& P, cudaStream_t stream);

// collision cost of n explicit points pts[n][D] (unit check of kinematics + narrow phase)
cudaError_t launch_eval_points(const ssppm::ModelView& M, int D, const float* pts, int n, float* out, cudaStream_t stream);

// SDF mode: bake the node grid from the handle's static primitives (M.sg[0..n_sgeom)) and pack the
// 8 corners of every cell contiguously for the kernel's two 16-byte loads per sample
cudaError_t launch_sdf_bake(const ssppm::ModelView& M, float* nodes, float* cells, int n_sm, cudaStream_t stream);

// Collision.h:51-82 for n configurations (device pointers): FlatModel geom ids of the penetrating contact with the smallest
// geom pair (-1: free) and its distance; mg_src / sg_src map the compiled geom tables back to FlatModel ids
cudaError_t launch_check_points(const ssppm::ModelView& M, int D, const float* pts, int n, const int* mg_src, const int* sg_src, int* geom1,
                                int* geom2, float* dist, cudaStream_t stream);

// SDF mode, refinement step: per collocation row the closest sample sphere's clearance and -clearance * unit field
// gradient, and their projection on the point-matrix rows G[K][3] = (1/rows) sum_m dirs[m] W[m][.]  (device pointers)
cudaError_t launch_sdf_gradient(const ssppm::ModelView& M, int D, int K, int rows, const float* W, const float* pts, float* dist,
                                float* dirs, float* G, cudaStream_t stream);

// batched B-spline evaluation out[B][T][d] = rows[T][n] . ctrl[B][n][d] (device pointers)
cudaError_t launch_bspline_eval(const float* rows, int T, int n, const float* ctrl, int B, int d, float* out, int n_sm, cudaStream_t stream);

// the same evaluation with HOST pointers and no allocation: staging in the file-scope device array of sspp_kernels.cu
// (kStaticScratchFloats); cudaErrorInvalidValue when static_scratch_floats_needed() exceeds it.  Synchronous.
size_t static_scratch_floats_needed(int T, int n, int B, int d);
cudaError_t bspline_eval_static(int device, const float* h_rows, int T, int n, const float* h_ctrl, int B, int d, float* h_out);

// sampled rows of the converged candidates, compacted in candidate order: rows [Q][N][nfd], idx [Q][N], count [Q] (device-visible
// addresses, normally mapped pinned host memory)
cudaError_t launch_success_rows(const IterParams& P, float* rows, int32_t* idx, int32_t* count, cudaStream_t stream);

// diagnostics: executed-work counters (ssppm::WorkCounter) of the stored candidates of the last iteration, added to totals[WC_COUNT]
cudaError_t launch_count(const ssppm::ModelView& M, const IterParams& P, unsigned long long* totals, int n_sm, cudaStream_t stream);

// FP32 FMA peak probe: every thread runs `iters` rounds of 8 independent FMA chains
cudaError_t launch_fma_probe(float* sink, int blocks, int threads, int iters, cudaStream_t stream);

}  // namespace sspp
