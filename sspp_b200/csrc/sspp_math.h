// sspp_math.h — single-source FP32 arithmetic of the candidate-scoring hot path.
//
// Compiled by nvcc this is the body of the sm_100a kernels (sspp_kernels.cu); compiled by g++
// (-ffp-contract=off) it is the bit-replica checker oracle/o2_replica.cpp.  Every fused
// multiply-add is an explicit fmaf(); the device build uses --fmad=false so nothing else is
// contracted, sqrt and division are IEEE on both sides, and sin/cos/log are the polynomials
// below (libm and the CUDA math library do not agree bit for bit).  No warp intrinsics in here:
// reductions live in the kernel and are restated by the oracle.
//
// Reference semantics restated (file:line under /root/reference):
//   sampleWithNoise          include/sspp.h:114-130        -> sample_joint_row()
//   Sampler::sample          include/sspp/tsp_sampler.h:12-38 -> sample_task_dim()
//   mj_set_point/yaw_to_quat include/utility.h:149-175,198-206 -> fk_body() JT_FREE4
//   checkCollision (ncon>0)  include/sspp.h:132-150        -> contact rule CR_COUNT
//   collision_point_cost     include/Collision.h:84-103    -> contact rule CR_CENTER
//   floorPenalty             include/sspp/tsp_evaluator.h:10-14 -> floor_penalty()
// MuJoCo's kinematics / narrow phase (third party, not in the reference tree) are restated
// from its documented semantics; see DESIGN.md "Geometry semantics".
#pragma once
#include <stdint.h>
#include <math.h>

#if defined(__CUDACC__)
#define SSPP_HD __host__ __device__ __forceinline__
#else
#define SSPP_HD inline
#endif

namespace ssppm {

// ----------------------------------------------------------------------------------------
// limits of the flattened model
// ----------------------------------------------------------------------------------------
enum { MAX_MBODY = 16, MAX_MGEOM = 64, MAX_GEOM_PAIRS = 32, MAX_SGEOM = 112, MAX_PAIR = 1024, MAX_MMPAIR = 64,
       MAX_DOF = 9, MAX_K = 16, MAX_SPHERE = 256 };

enum GeomType { GT_PLANE = 0, GT_SPHERE = 2, GT_CAPSULE = 3, GT_CYLINDER = 5, GT_BOX = 6 };
enum JointType { JT_NONE = 0, JT_HINGE = 1, JT_SLIDE = 2, JT_FREE4 = 3 };
enum ContactRule { CR_COUNT = 0,   // planner A: every contact with dist < margin counts 1
                   CR_CENTER = 1,  // planner B: dist < -1e-3 adds -1/(centre distance + 1e-4)   (Collision.h:93-96)
                   CR_DEPTH = 2 }; // planner B, use_center_dist = false: dist < -1e-3 adds -dist (Collision.h:97-98)

// Work counters (diagnostics: ssppcu_work_counters / sspp_count_kernel).  Every function below that takes a
// `uint32_t* wc` adds to wc[WC_*] when the pointer is non-null; the scoring kernels pass nullptr and the
// inlined code disappears.  They give the EXECUTED work of a launch (roofline: flop_per_point_executed).
enum WorkCounter { WC_POINTS = 0,      // configurations scored (kinematics + collision)
                   WC_FK_HINGE_AXIS,   // hinge about a coordinate axis of the body frame
                   WC_FK_HINGE_GEN,    // hinge about a general axis (Rodrigues)
                   WC_FK_SLIDE, WC_FK_FREE4,
                   WC_FK_R0,           // bodies under a moving parent with a non-identity rest orientation (one 3x3 product more)
                   WC_FK_ANCHOR,       // hinges with an anchor off the frame origin
                   WC_GEOM_POS,        // moving geoms placed (centre only)
                   WC_GEOM_ROT_BOX, WC_GEOM_ROT_AXIS,   // orientations placed on demand: full matrix / axis column only
                   WC_BROAD,           // broad-phase pair tests (one affine form each)
                   WC_BODY_BROAD,      // body-level broad-phase tests (the same affine form, per body and static partner)
                   WC_NP_PLANE,        // narrow phase: any moving geom against a plane
                   WC_NP_SPHERE_BOX,
                   WC_NP_BOX_CULL,     // second-level cull: the geom's bounding sphere against the partner BOX itself (not its sphere)
                   WC_NP_CAPS_BOX_EARLY, WC_NP_CAPS_BOX, WC_NP_CAPS_NEWTON,   // centre-distance exit / full segment-box / gradient evaluations inside it
                   WC_NP_BOX_BOX_FACE, WC_NP_BOX_BOX,     // 6 face axes only (separated beyond thr) / all 15 axes
                   WC_NP_CYL_BOX_FACE, WC_NP_CYL_BOX,
                   WC_NP_OTHER,        // pairings served by prim_sd
                   WC_CONTACTS,        // pairs that ended in a contact (cost != 0)
                   WC_MM_TESTS, WC_MM_NARROW,             // moving-moving pairs: bounding-sphere tests / narrow phases
                   WC_SDF_SAMPLES,     // SDF mode: sphere samples read from the grid
                   WC_COUNT };

// Every record is a whole number of 16-byte words with its hot fields grouped four by four: the kernels stage
// the tables into shared memory and read them with warp-uniform (broadcast) 128-bit loads.
struct alignas(16) DevBody {
  float pos[3];     // frame origin in the parent frame (world frame when parent < 0)
  float axis_sign;
  float R0[9];      // frame orientation in the parent frame, row-major
  float axis[3];    // joint axis in the body frame (unit)
  float anchor[3];  // joint anchor in the body frame
  float fpad_;
  int parent;       // index into the moving-body list, -1: static parent folded into pos/R0
  int jtype;        // JointType
  int qidx;         // first planned coordinate this joint reads
  int axis_kind;    // 0,1,2: the joint axis is +-e_x, +-e_y, +-e_z (axis_sign); 3: general axis
  int has_anchor;
  int r0_identity;  // R0 == I: the body frame is a pure translation of the parent frame
  int geom_begin, geom_end;  // moving geoms attached to this body
  int sph_begin, sph_end;    // SDF mode: sphere samples attached to this body
  int pad_[2];
  // body-level broad phase (bodies that carry several geoms): a sphere (centre in the body frame, radius) around all of the
  // body's geoms and one affine record (DevPair form, in the pair table at [bpair_begin, bpair_end), planes first) per static
  // partner of any of them; when the sphere is clear of every partner the body's geoms are not even placed
  float bound[4];
  int bpair_begin, bpair_mid, bpair_end, pad2_;
};
struct alignas(16) DevMGeom {
  float pos[3];     // in the body frame
  float rbound;     // bounding-sphere radius about pos
  float R[9];
  float size[3];    // sphere r | capsule,cylinder r,half-length | box half extents
  int type;
  int pair_begin, pair_mid, pair_end;   // <= 32 pairs: [pair_begin, pair_mid) plane partners, [pair_mid, pair_end) bounded ones
  int body;
  int pad_[3];
};
struct alignas(16) DevSGeom {
  float pos[3];     // world frame
  float rbound;     // < 0: unbounded (plane)
  float R[9];
  float size[3];
  int type;
  int pad_[3];
};
// One filtered (moving geom, static geom) pair with its broad-phase test precomputed by compile_model as ONE
// affine form of the moving geom's centre p:  f(p) = c[0] p.x + c[1] p.y + c[2] p.z + c[3]
//   plane partner:   c = (n, -(n . S.pos + G.rbound + thr));      culled when f(p) > 0
//   bounded partner: c = (-2 S.pos, |S.pos|^2 - (G.rbound + S.rbound + thr)^2);  culled when f(p) > -|p|^2
//                    (|p - S.pos|^2 > rr^2 expanded; |p|^2 is formed once per moving geom)
// thr = max(margin, 0) + 1e-4 (conservative: margin + slack; the expansion's rounding, ~1e-6 m, is far inside it).
struct alignas(16) DevPair { float c[4]; int s; float margin; float thr; int plane; };
// SDF mode: a moving geom is sampled by spheres (centre in the frame of moving body `body`)
struct alignas(16) DevSphere { float c[3]; float r; float margin; int body; int pad_[2]; };
// dense signed-distance grid of the static scene: node (i,j,k) sits at lo + (i,j,k) * h
struct SdfGrid {
  int nx, ny, nz;
  float lo[3], h[3], inv_h[3];
  const float* nodes;   // [nz][ny][nx] node values (host replica; device read-back)
  const float* cells;   // [(nz-1)][(ny-1)][(nx-1)][8] the 8 corners of every cell, contiguous (32 B = one sector)
};
struct alignas(16) DevMMPair { int a, b; float margin; int pad_; };

// scalars of the flattened scene (kernel-parameter resident)
struct DevModelHead {
  int n_mbody, n_mgeom, n_sgeom, n_pair, n_mmpair;
  int n_bpair;       // body-level broad-phase records, stored behind the n_pair geom pairs in the pair table
  int rule;          // ContactRule
  int chain;         // 1: parent[b] == b-1 for every b and no moving-moving pairs (no transform store)
  float static_cost; // contribution of pairs that never move (added to every point)
  int sdf_mode;      // 1: static scene = SdfGrid, moving geoms = sphere samples (pairs unused)
  int n_sphere;
  SdfGrid grid;
};
// host container of the scene (what compile_model returns and the replica oracle reads)
struct DevModel : DevModelHead {
  DevBody body[MAX_MBODY];
  DevMGeom mg[MAX_MGEOM];
  DevSGeom sg[MAX_SGEOM];
  DevPair pair[MAX_PAIR];
  DevMMPair mm[MAX_MMPAIR];
  DevSphere sph[MAX_SPHERE];   // grouped by body: body[b].sph_begin .. sph_end
};
// the same scene through pointers: device global memory (kernel parameter) or shared memory (staged copy)
struct ModelView : DevModelHead {
  const DevBody* body;
  const DevMGeom* mg;
  const DevSGeom* sg;
  const DevPair* pair;
  const DevMMPair* mm;
  const DevSphere* sph;
};

static_assert(sizeof(DevBody) == 160 && sizeof(DevMGeom) == 96 && sizeof(DevSGeom) == 80 && sizeof(DevPair) == 32 &&
              sizeof(DevSphere) == 32 && sizeof(DevMMPair) == 16, "table records are whole 16-byte words");

// ----------------------------------------------------------------------------------------
// small vector helpers (explicit fmaf everywhere)
// ----------------------------------------------------------------------------------------
struct v3 { float x, y, z; };
struct m3 { float a[9]; };  // row-major

SSPP_HD v3 mk(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
SSPP_HD v3 ld3(const float* p) { return mk(p[0], p[1], p[2]); }
SSPP_HD v3 add(v3 a, v3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
SSPP_HD v3 sub(v3 a, v3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
SSPP_HD v3 scale(v3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
SSPP_HD v3 madd(v3 a, float s, v3 b) { return mk(fmaf(a.x, s, b.x), fmaf(a.y, s, b.y), fmaf(a.z, s, b.z)); }
SSPP_HD float dot(v3 a, v3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
SSPP_HD v3 cross(v3 a, v3 b) {
  return mk(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
SSPP_HD v3 col(const m3& R, int j) { return mk(R.a[j], R.a[3 + j], R.a[6 + j]); }
SSPP_HD v3 mulv(const m3& R, v3 v) {  // R v
  return mk(fmaf(R.a[2], v.z, fmaf(R.a[1], v.y, R.a[0] * v.x)),
            fmaf(R.a[5], v.z, fmaf(R.a[4], v.y, R.a[3] * v.x)),
            fmaf(R.a[8], v.z, fmaf(R.a[7], v.y, R.a[6] * v.x)));
}
SSPP_HD v3 multv(const m3& R, v3 v) {  // R^T v
  return mk(fmaf(R.a[6], v.z, fmaf(R.a[3], v.y, R.a[0] * v.x)),
            fmaf(R.a[7], v.z, fmaf(R.a[4], v.y, R.a[1] * v.x)),
            fmaf(R.a[8], v.z, fmaf(R.a[5], v.y, R.a[2] * v.x)));
}
SSPP_HD m3 mulm(const m3& A, const float* B) {  // A B, B row-major
  m3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j)
      C.a[3 * i + j] = fmaf(A.a[3 * i + 2], B[6 + j], fmaf(A.a[3 * i + 1], B[3 + j], A.a[3 * i] * B[j]));
  return C;
}
SSPP_HD m3 ldm(const float* p) {
  m3 R;
#pragma unroll
  for (int i = 0; i < 9; ++i) R.a[i] = p[i];
  return R;
}
SSPP_HD float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// Four consecutive words of a 16-byte aligned table record in one access (the records above group their
// fields for this); the host build reads the same four words one by one.
struct alignas(16) f4 { float x, y, z, w; };
struct alignas(16) i4 { int x, y, z, w; };
SSPP_HD f4 ld4(const float* p) {
#if defined(__CUDA_ARCH__)
  return *reinterpret_cast<const f4*>(p);
#else
  f4 r; r.x = p[0]; r.y = p[1]; r.z = p[2]; r.w = p[3]; return r;
#endif
}
SSPP_HD i4 ld4i(const int* p) {
#if defined(__CUDA_ARCH__)
  return *reinterpret_cast<const i4*>(p);
#else
  i4 r; r.x = p[0]; r.y = p[1]; r.z = p[2]; r.w = p[3]; return r;
#endif
}
SSPP_HD float word_as_float(int w) {
#if defined(__CUDA_ARCH__)
  return __int_as_float(w);
#else
  union { int i; float f; } b; b.i = w; return b.f;
#endif
}
// Record access.  Tables are arrays of 16-byte aligned records; rec_f4 / rec_i4(tab, i, w) read words
// [w, w + 4) of record i (w a multiple of 4).  These overloads serve plain arrays (host container, global
// memory); the kernels add overloads for their shared-memory staged tables.  Word offsets of the groups:
enum { BW_POS = 0, BW_R0 = 4, BW_ANCHOR = 16, BW_JOINT = 20, BW_FLAGS = 24, BW_SPH = 28, BW_BOUND = 32, BW_BPAIR = 36,   // DevBody
       GW_POS = 0, GW_R = 4, GW_TYPE = 16, GW_BODY = 20,                                   // DevMGeom / DevSGeom
       PW_C = 0, PW_S = 4,                                                               // DevPair
       SW_C = 0, SW_MARGIN = 4 };                                                        // DevSphere
template <class T> SSPP_HD f4 rec_f4(const T* tab, int i, int w) { return ld4(reinterpret_cast<const float*>(tab + i) + w); }
template <class T> SSPP_HD i4 rec_i4(const T* tab, int i, int w) { return ld4i(reinterpret_cast<const int*>(tab + i) + w); }
// R[9] + 3 trailing words (size / axis) of record i, starting at word w, as three 4-word reads
struct m3t { m3 R; v3 t; };
template <class TAB>
SSPP_HD m3t rec_rot12(const TAB& tab, int i, int w) {
  const f4 a = rec_f4(tab, i, w), b = rec_f4(tab, i, w + 4), c = rec_f4(tab, i, w + 8);
  m3t o;
  o.R.a[0] = a.x; o.R.a[1] = a.y; o.R.a[2] = a.z; o.R.a[3] = a.w;
  o.R.a[4] = b.x; o.R.a[5] = b.y; o.R.a[6] = b.z; o.R.a[7] = b.w;
  o.R.a[8] = c.x; o.t.x = c.y; o.t.y = c.z; o.t.z = c.w;
  return o;
}
SSPP_HD int lowest_bit(uint32_t m) {   // index of the lowest set bit, m != 0
#if defined(__CUDA_ARCH__)
  return __ffs((int)m) - 1;
#else
  return __builtin_ctz(m);
#endif
}

// ----------------------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al., SC'11) — counter based, so samples are a pure function of
// (seed, iteration, query, candidate, stream) on the GPU and in both oracles.
// ----------------------------------------------------------------------------------------
SSPP_HD void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c[0];
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c[2];
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
}

// counter layout: {candidate, query, iteration, stream << 8 | block}
SSPP_HD void philox_block(uint64_t seed, uint32_t iter, uint32_t query, uint32_t cand,
                          uint32_t stream, uint32_t block, uint32_t out[4]) {
  out[0] = cand; out[1] = query; out[2] = iter; out[3] = (stream << 8) | (block & 0xFFu);
  philox4x32_10(out, (uint32_t)seed, (uint32_t)(seed >> 32));
}

// (0,1) uniform with 23 random bits, exactly representable
SSPP_HD float u01(uint32_t x) { return ((float)(x >> 9) + 0.5f) * 1.1920928955078125e-7f; }

// natural log for x in (0, 2), cephes logf polynomial with explicit fma
SSPP_HD float log_poly(float x) {
  union { float f; uint32_t u; } b; b.f = x;
  int e = (int)(b.u >> 23) - 126;            // x = m * 2^e, m in [0.5,1)
  b.u = (b.u & 0x007FFFFFu) | 0x3F000000u;
  float m = b.f;
  if (m < 0.70710678118654752440f) { e -= 1; m = m + m; }
  const float f = m - 1.0f;
  const float z = f * f;
  float y = 7.0376836292e-2f;
  y = fmaf(y, f, -1.1514610310e-1f);
  y = fmaf(y, f, 1.1676998740e-1f);
  y = fmaf(y, f, -1.2420140846e-1f);
  y = fmaf(y, f, 1.4249322787e-1f);
  y = fmaf(y, f, -1.6668057665e-1f);
  y = fmaf(y, f, 2.0000714765e-1f);
  y = fmaf(y, f, -2.4999993993e-1f);
  y = fmaf(y, f, 3.3333331174e-1f);
  y = y * f * z;
  const float fe = (float)e;
  y = fmaf(fe, -2.12194440e-4f, y);
  y = fmaf(-0.5f, z, y);
  return fmaf(fe, 0.693359375f, f + y);
}

// ln(1 + f) for f in (-0.2929, 0.4143): log_poly's kernel without the reduction (same coefficients, exponent 0)
SSPP_HD float log1p_poly(float f) {
  const float z = f * f;
  float y = 7.0376836292e-2f;
  y = fmaf(y, f, -1.1514610310e-1f);
  y = fmaf(y, f, 1.1676998740e-1f);
  y = fmaf(y, f, -1.2420140846e-1f);
  y = fmaf(y, f, 1.4249322787e-1f);
  y = fmaf(y, f, -1.6668057665e-1f);
  y = fmaf(y, f, 2.0000714765e-1f);
  y = fmaf(y, f, -2.4999993993e-1f);
  y = fmaf(y, f, 3.3333331174e-1f);
  y = y * f * z;
  y = fmaf(-0.5f, z, y);
  return f + y;
}
// CES log-weight of rank j among k elites: ln(k + 1/2) - ln(j + 1) (tsp_elites.h:24-32); lk = log_poly(k + 1/2).  Towards the end
// of the elite list the two logarithms nearly cancel (j = k - 1: their difference is 1 / 2k), and a difference of FP32 logarithms
// is then off by up to 1e-4 relative at k ~ 1000.  Where the ratio (k + 1/2) / (j + 1) is below sqrt(2) the weight is ln(1 + x) of
// x = (k - j - 1/2) / (j + 1), whose numerator and denominator are exact in FP32.
SSPP_HD float ces_log_weight(int k, float lk, int j) {
  const float den = (float)j + 1.0f;
  const float x = ((float)(k - j) - 0.5f) / den;
  return (x <= 0.41421354f) ? log1p_poly(x) : lk - log_poly(den);
}

// sin and cos for |x| < ~1e4: 3-term Cody-Waite reduction to [-pi/4, pi/4], cephes minimax kernels
SSPP_HD void sincos_poly(float x, float* s, float* c) {
  const float k = rintf(x * 0.63661977236758134308f);
  float r = fmaf(k, -1.5703125f, x);
  r = fmaf(k, -4.837512969970703125e-4f, r);
  r = fmaf(k, -7.54978995489188216e-8f, r);
  const float z = r * r;
  float ps = -1.9515295891e-4f;
  ps = fmaf(ps, z, 8.3321608736e-3f);
  ps = fmaf(ps, z, -1.6666654611e-1f);
  const float sr = fmaf(ps * z, r, r);
  float pc = 2.443315711809948e-5f;
  pc = fmaf(pc, z, -1.388731625493765e-3f);
  pc = fmaf(pc, z, 4.166664568298827e-2f);
  const float cr = fmaf(pc, z * z, fmaf(-0.5f, z, 1.0f));
  const int q = (int)k & 3;
  const float sv = (q & 1) ? cr : sr;
  const float cv = (q & 1) ? sr : cr;
  *s = (q & 2) ? -sv : sv;
  *c = ((q + 1) & 2) ? -cv : cv;
}

// Box-Muller: two uniforms -> two standard normals
SSPP_HD void box_muller(uint32_t a, uint32_t b, float* z0, float* z1) {
  const float r = sqrtf(-2.0f * log_poly(u01(a)));
  float s, c;
  sincos_poly(6.28318530717958647692f * u01(b), &s, &c);
  *z0 = r * c; *z1 = r * s;
}

// the j-th standard normal of (seed, iter, query, cand, stream): block j/4, lane j%4
SSPP_HD float normal_at(uint64_t seed, uint32_t iter, uint32_t query, uint32_t cand,
                        uint32_t stream, uint32_t j) {
  uint32_t w[4];
  philox_block(seed, iter, query, cand, stream, j >> 2, w);
  float z0, z1;
  if (j & 2) box_muller(w[2], w[3], &z0, &z1); else box_muller(w[0], w[1], &z0, &z1);
  return (j & 1) ? z1 : z0;
}

// four normals of one Philox block
SSPP_HD void normal4(uint64_t seed, uint32_t iter, uint32_t query, uint32_t cand,
                     uint32_t stream, uint32_t block, float z[4]) {
  uint32_t w[4];
  philox_block(seed, iter, query, cand, stream, block, w);
  box_muller(w[0], w[1], &z[0], &z[1]);
  box_muller(w[2], w[3], &z[2], &z[3]);
}

// Planner B via sampler for one coordinate (include/sspp/tsp_sampler.h:16-36).
//  dims 0..2: N(mean,sigma) rejected until inside [lo,hi]; after 100 tries U(lo,hi)
//  dim 3 (yaw): N(mean,sigma) wrapped into [lo,hi] by +-range; mean if lo == hi
// stream = free_row * 4 + dim; try t consumes normal t of that stream.
SSPP_HD float sample_task_dim(uint64_t seed, uint32_t iter, uint32_t query, uint32_t cand,
                              int row, int dim, float mean, float sigma, float lo, float hi) {
  const uint32_t stream = (uint32_t)(row * 4 + dim);
  if (dim < 3) {
    float s = 0.0f;
    int tries = 0;
    bool inside = false;
    for (int blk = 0; blk < 25 && !inside; ++blk) {
      float z[4];
      normal4(seed, iter, query, cand, stream, (uint32_t)blk, z);
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        if (!inside && tries < 99) {     // tries 1..99 may be accepted (reference breaks at try 100)
          s = fmaf(z[l], sigma, mean);
          ++tries;
          inside = !(s < lo || s > hi);
        }
      }
    }
    if (!inside) {                        // 100th draw is discarded for a uniform
      uint32_t w[4];
      philox_block(seed, iter, query, cand, stream, 25u, w);
      s = fmaf(u01(w[0]), hi - lo, lo);
    }
    return s;
  }
  if (lo == hi) return mean;
  float yaw = fmaf(normal_at(seed, iter, query, cand, stream, 0u), sigma, mean);
  const float range = hi - lo;
  for (int it = 0; it < 64 && yaw < lo; ++it) yaw += range;
  for (int it = 0; it < 64 && yaw > hi; ++it) yaw -= range;
  return yaw;
}

// ----------------------------------------------------------------------------------------
// forward kinematics of one moving body (MuJoCo mj_kinematics semantics for one joint per body)
// ----------------------------------------------------------------------------------------
SSPP_HD void rodrigues(v3 a, float th, float* Rj) {
  float s, c;
  sincos_poly(th, &s, &c);
  const float t = 1.0f - c;
  const float xy = a.x * a.y * t, xz = a.x * a.z * t, yz = a.y * a.z * t;
  Rj[0] = fmaf(a.x * a.x, t, c); Rj[1] = fmaf(-a.z, s, xy);     Rj[2] = fmaf(a.y, s, xz);
  Rj[3] = fmaf(a.z, s, xy);      Rj[4] = fmaf(a.y * a.y, t, c); Rj[5] = fmaf(-a.x, s, yz);
  Rj[6] = fmaf(-a.y, s, xz);     Rj[7] = fmaf(a.x, s, yz);      Rj[8] = fmaf(a.z * a.z, t, c);
}

// R <- R * Rot(e_K, theta): only the two other columns mix (compile-time indices keep R in registers)
template <int K>
SSPP_HD void rot_axis(m3& R, float s, float c) {
  constexpr int i1 = (K + 1) % 3, i2 = (K + 2) % 3;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const float u = R.a[3 * r + i1], v = R.a[3 * r + i2];
    R.a[3 * r + i1] = fmaf(s, v, c * u);
    R.a[3 * r + i2] = fmaf(c, v, -(s * u));
  }
}

// parent frame (Rp,pp) -> this body's frame (R,p) given planned coordinates q
template <class TAB, class QA>
SSPP_HD void fk_body(const TAB& bodies, int b, i4 jt, bool has_parent, const m3& Rp, v3 pp, QA& q, m3* R, v3* p, uint32_t* wc = nullptr) {
  const int jtype = jt.y, qidx = jt.z, axis_kind = jt.w;     // jt = (parent, jtype, qidx, axis_kind)
  if (jtype == JT_FREE4) {  // (x,y,z,yaw) -> pose; quaternion (cos y/2,0,0,sin y/2) == Rz(yaw)
    float s, c;
    sincos_poly(q.get(qidx + 3), &s, &c);
    R->a[0] = c; R->a[1] = -s; R->a[2] = 0.0f;
    R->a[3] = s; R->a[4] = c;  R->a[5] = 0.0f;
    R->a[6] = 0.0f; R->a[7] = 0.0f; R->a[8] = 1.0f;
    *p = mk(q.get(qidx), q.get(qidx + 1), q.get(qidx + 2));
    if (wc) ++wc[WC_FK_FREE4];
    return;
  }
  const f4 ps = rec_f4(bodies, b, BW_POS);             // pos, axis_sign
  const i4 fl = rec_i4(bodies, b, BW_FLAGS);           // has_anchor, r0_identity, geom_begin, geom_end
  const int has_anchor = fl.x, r0_identity = fl.y;
  m3 Rw; v3 pw;
  v3 axis = mk(0.f, 0.f, 0.f);
  if (has_parent) {
    if (r0_identity && axis_kind < 3 && jtype == JT_HINGE) Rw = Rp;
    else { const m3t ra = rec_rot12(bodies, b, BW_R0); axis = ra.t; Rw = r0_identity ? Rp : mulm(Rp, ra.R.a); }
    pw = add(pp, mulv(Rp, mk(ps.x, ps.y, ps.z)));
  } else { const m3t ra = rec_rot12(bodies, b, BW_R0); axis = ra.t; Rw = ra.R; pw = mk(ps.x, ps.y, ps.z); }
  if (wc) {
    if (has_parent && !r0_identity) ++wc[WC_FK_R0];
    if (jtype == JT_HINGE) { ++wc[axis_kind < 3 ? WC_FK_HINGE_AXIS : WC_FK_HINGE_GEN]; if (has_anchor) ++wc[WC_FK_ANCHOR]; }
    else if (jtype == JT_SLIDE) ++wc[WC_FK_SLIDE];
  }
  if (jtype == JT_HINGE) {
    m3 Rn;
    if (axis_kind < 3) {
      // rotation about a coordinate axis of the body frame: only the two other columns mix
      float s, c;
      sincos_poly(q.get(qidx), &s, &c);
      s *= ps.w;
      Rn = Rw;
      if (axis_kind == 0) rot_axis<0>(Rn, s, c); else if (axis_kind == 1) rot_axis<1>(Rn, s, c); else rot_axis<2>(Rn, s, c);
    } else {
      float Rj[9];
      rodrigues(axis, q.get(qidx), Rj);
      Rn = mulm(Rw, Rj);
    }
    if (has_anchor) { const f4 an4 = rec_f4(bodies, b, BW_ANCHOR); const v3 an = mk(an4.x, an4.y, an4.z); pw = add(pw, sub(mulv(Rw, an), mulv(Rn, an))); }
    Rw = Rn;
  } else if (jtype == JT_SLIDE) {
    pw = madd(mulv(Rw, axis), q.get(qidx), pw);
  }
  *R = Rw; *p = pw;
}

// ----------------------------------------------------------------------------------------
// primitives in the world frame and their signed distances (negative = penetration depth)
// ----------------------------------------------------------------------------------------
struct Prim { v3 p; m3 R; v3 s; int type; };

SSPP_HD float seg_point_dist2(v3 a, v3 d, float dd, v3 x) {  // segment a + t d
  const v3 w = sub(x, a);
  float t = (dd > 0.0f) ? clampf(dot(w, d) / dd, 0.0f, 1.0f) : 0.0f;
  const v3 r = sub(w, scale(d, t));
  return dot(r, r);
}

// squared distance between segments p1+s d1 and p2+t d2 (Ericson 5.1.9 clamping scheme)
SSPP_HD float seg_seg_dist2(v3 p1, v3 d1, v3 p2, v3 d2) {
  const v3 r = sub(p1, p2);
  const float a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r);
  float s, t;
  const float eps = 1e-12f;
  if (a <= eps && e <= eps) { s = 0.0f; t = 0.0f; }
  else if (a <= eps) { s = 0.0f; t = clampf(f / e, 0.0f, 1.0f); }
  else {
    const float c = dot(d1, r);
    if (e <= eps) { t = 0.0f; s = clampf(-c / a, 0.0f, 1.0f); }
    else {
      const float b = dot(d1, d2);
      const float den = fmaf(a, e, -(b * b));
      s = (den > eps) ? clampf(fmaf(b, f, -(c * e)) / den, 0.0f, 1.0f) : 0.0f;
      t = fmaf(b, s, f) / e;
      if (t < 0.0f) { t = 0.0f; s = clampf(-c / a, 0.0f, 1.0f); }
      else if (t > 1.0f) { t = 1.0f; s = clampf((b - c) / a, 0.0f, 1.0f); }
    }
  }
  const v3 w = sub(madd(d1, s, r), scale(d2, t));
  return dot(w, w);
}

// signed distance of a point (box frame) to a box of half extents h
SSPP_HD float point_box_sd(v3 x, v3 h) {
  const float qx = fabsf(x.x) - h.x, qy = fabsf(x.y) - h.y, qz = fabsf(x.z) - h.z;
  const float ox = fmaxf(qx, 0.0f), oy = fmaxf(qy, 0.0f), oz = fmaxf(qz, 0.0f);
  const float out2 = fmaf(oz, oz, fmaf(oy, oy, ox * ox));
  const float in = fmaxf(qx, fmaxf(qy, qz));
  return (out2 > 0.0f) ? sqrtf(out2) : in;
}

// distance (>= 0) from the segment p0 + t d (box frame, t in [0,1]) to a box.  The squared
// distance g(t) is convex, C1 and piecewise quadratic, so g'(t) is monotone piecewise linear with kinks only where a
// coordinate of the point crosses a face plane (t = (+-h_a - p0_a) / d_a, at most six of them).  Bracket the root of g' on
// [0,1], shrink the bracket with every kink that falls inside it, and the root of the linear piece that is left is exact.
// (The first version ran six safeguarded Newton steps; on long capsules - a 0.9 m post against a 3 cm plate - six steps
// could stop millimetres short of the minimum and miss a shallow contact that the FP64 oracle and MuJoCo's capsule-box
// routine see.)
SSPP_HD void seg_box_grad(float t, v3 p0, v3 d, v3 h, float* g, float* hs) {
  const float x = fmaf(t, d.x, p0.x), y = fmaf(t, d.y, p0.y), z = fmaf(t, d.z, p0.z);
  const float ex = x - clampf(x, -h.x, h.x), ey = y - clampf(y, -h.y, h.y), ez = z - clampf(z, -h.z, h.z);
  *g = fmaf(ez, d.z, fmaf(ey, d.y, ex * d.x));
  const float hx = (ex != 0.0f) ? d.x * d.x : 0.0f;
  const float hy = (ey != 0.0f) ? d.y * d.y : 0.0f;
  const float hz = (ez != 0.0f) ? d.z * d.z : 0.0f;
  *hs = hx + hy + hz;
}
SSPP_HD float seg_box_dist(v3 p0, v3 d, v3 h, uint32_t* wc = nullptr) {
  float g0, g1, hs;
  seg_box_grad(0.0f, p0, d, h, &g0, &hs);
  seg_box_grad(1.0f, p0, d, h, &g1, &hs);
  float t = (g0 >= 0.0f) ? 0.0f : 1.0f;
  if (g0 < 0.0f && g1 > 0.0f) {        // interior minimum; otherwise an end point is the answer
    float lo = 0.0f, hi = 1.0f, glo = g0, ghi = g1;
    const float pc[3] = {p0.x, p0.y, p0.z}, dc[3] = {d.x, d.y, d.z}, hc[3] = {h.x, h.y, h.z};
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
      for (int sgn = 0; sgn < 2; ++sgn) {
        const float tb = ((sgn ? hc[a] : -hc[a]) - pc[a]) / dc[a];    // d_a = 0: inf or nan, outside every bracket
        if (tb > lo && tb < hi) {
          float g;
          seg_box_grad(tb, p0, d, h, &g, &hs);
          if (wc) ++wc[WC_NP_CAPS_NEWTON];
          if (g > 0.0f) { hi = tb; ghi = g; } else { lo = tb; glo = g; }
        }
      }
    }
    t = lo - glo * ((hi - lo) / (ghi - glo));                           // g' is linear on [lo, hi] and glo <= 0 < ghi
    t = fminf(fmaxf(t, lo), hi);
  }
  const float x = fmaf(t, d.x, p0.x), y = fmaf(t, d.y, p0.y), z = fmaf(t, d.z, p0.z);
  const float ex = x - clampf(x, -h.x, h.x), ey = y - clampf(y, -h.y, h.y), ez = z - clampf(z, -h.z, h.z);
  return sqrtf(fmaf(ez, ez, fmaf(ey, ey, ex * ex)));
}

// running maximum of s/sqrt(l2) kept without the square root: compare s|s|/l2 by cross products
SSPP_HD void sat_max(float s, float l2, float* bs, float* bl2) {
  if (s * fabsf(s) * (*bl2) > (*bs) * fabsf(*bs) * l2) { *bs = s; *bl2 = l2; }
}

// box-box: 15-axis separating-axis test; > 0 separation along the best axis (a lower bound of
// the distance), < 0 the smallest penetration depth
// thr: once the face axes alone separate the boxes by more than thr the 9 edge axes are skipped (the
// result only has to stay above thr: callers compare it with a margin below thr)
SSPP_HD float box_box_sd(const Prim& A, const Prim& B, float thr = 3.0e38f, uint32_t* wc = nullptr) {
  float Rm[9], Ab[9];
  const v3 dlt = sub(B.p, A.p);
  const v3 t = multv(A.R, dlt);
  const float tt[3] = {t.x, t.y, t.z};
  const float ea[3] = {A.s.x, A.s.y, A.s.z}, eb[3] = {B.s.x, B.s.y, B.s.z};
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      Rm[3 * i + j] = dot(col(A.R, i), col(B.R, j));
      Ab[3 * i + j] = fabsf(Rm[3 * i + j]);
    }
  float bs = -3.0e38f, bl2 = 1.0f;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const float ra = ea[i];
    const float rb = fmaf(eb[2], Ab[3 * i + 2], fmaf(eb[1], Ab[3 * i + 1], eb[0] * Ab[3 * i]));
    const float s = fabsf(tt[i]) - (ra + rb);
    if (s > bs) bs = s;
  }
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const float ra = fmaf(ea[2], Ab[6 + j], fmaf(ea[1], Ab[3 + j], ea[0] * Ab[j]));
    const float rb = eb[j];
    const float pr = fmaf(tt[2], Rm[6 + j], fmaf(tt[1], Rm[3 + j], tt[0] * Rm[j]));
    const float s = fabsf(pr) - (ra + rb);
    if (s > bs) bs = s;
  }
  if (wc) ++wc[bs > thr ? WC_NP_BOX_BOX_FACE : WC_NP_BOX_BOX];
  if (bs > thr) return bs;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int i1 = (i + 1) % 3, i2 = (i + 2) % 3;
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int j1 = (j + 1) % 3, j2 = (j + 2) % 3;
      const float l2 = fmaf(-Rm[3 * i + j], Rm[3 * i + j], 1.0f);
      const float ra = fmaf(ea[i1], Ab[3 * i2 + j], ea[i2] * Ab[3 * i1 + j]);
      const float rb = fmaf(eb[j1], Ab[3 * i + j2], eb[j2] * Ab[3 * i + j1]);
      const float pr = fmaf(tt[i2], Rm[3 * i1 + j], -(tt[i1] * Rm[3 * i2 + j]));
      const float s = fabsf(pr) - (ra + rb);
      if (l2 > 1e-6f) sat_max(s, l2, &bs, &bl2);
    }
  }
  return bs / sqrtf(bl2);
}

// cylinder-box: support-function separating test over a fixed direction set
//   box face normals, cylinder axis, axis x face normal, direction centre -> closest box point
SSPP_HD float cyl_support(v3 a, float r, float h, v3 L, float l2) {  // support of the cylinder along L
  const float la = dot(L, a);
  return fmaf(h, fabsf(la), r * sqrtf(fmaxf(fmaf(-la, la, l2), 0.0f)));
}
SSPP_HD float cyl_box_sd(const Prim& C, const Prim& B, float thr = 3.0e38f, uint32_t* wc = nullptr) {
  const v3 a = col(C.R, 2);
  const float r = C.s.x, h = C.s.y;
  const v3 dlt = sub(C.p, B.p);
  const v3 bx[3] = {col(B.R, 0), col(B.R, 1), col(B.R, 2)};
  const float eb[3] = {B.s.x, B.s.y, B.s.z};
  float bs = -3.0e38f, bl2 = 1.0f;
#pragma unroll
  for (int i = 0; i < 3; ++i) {   // box normals (unit)
    const float s = fabsf(dot(dlt, bx[i])) - eb[i] - cyl_support(a, r, h, bx[i], 1.0f);
    if (s > bs) bs = s;
  }
  {                               // cylinder axis (unit)
    const float rb = fmaf(eb[2], fabsf(dot(a, bx[2])), fmaf(eb[1], fabsf(dot(a, bx[1])), eb[0] * fabsf(dot(a, bx[0]))));
    const float s = fabsf(dot(dlt, a)) - rb - h;
    if (s > bs) bs = s;
  }
  if (wc) ++wc[bs > thr ? WC_NP_CYL_BOX_FACE : WC_NP_CYL_BOX];
  if (bs > thr) return bs;
#pragma unroll
  for (int i = 0; i < 3; ++i) {   // axis x box normal
    const v3 L = cross(a, bx[i]);
    const float l2 = dot(L, L);
    const float rb = fmaf(eb[2], fabsf(dot(L, bx[2])), fmaf(eb[1], fabsf(dot(L, bx[1])), eb[0] * fabsf(dot(L, bx[0]))));
    const float s = fabsf(dot(dlt, L)) - rb - r * sqrtf(l2);
    if (l2 > 1e-6f) sat_max(s, l2, &bs, &bl2);
  }
  {                               // centre -> closest point of the box
    const v3 x = multv(B.R, dlt);
    const v3 cp = mk(clampf(x.x, -eb[0], eb[0]), clampf(x.y, -eb[1], eb[1]), clampf(x.z, -eb[2], eb[2]));
    const v3 Ll = sub(x, cp);     // box frame
    const float l2 = dot(Ll, Ll);
    if (l2 > 1e-12f) {
      const v3 L = mulv(B.R, Ll);
      const float rb = fmaf(eb[2], fabsf(Ll.z), fmaf(eb[1], fabsf(Ll.y), eb[0] * fabsf(Ll.x)));
      const float s = fabsf(dot(dlt, L)) - rb - cyl_support(a, r, h, L, l2);
      sat_max(s, l2, &bs, &bl2);
    }
  }
  return bs / sqrtf(bl2);
}

SSPP_HD float point_cyl_sd(v3 x, const Prim& C) {  // signed distance of a world point to a cylinder
  const v3 a = col(C.R, 2);
  const v3 w = sub(x, C.p);
  const float z = dot(w, a);
  const v3 rad = sub(w, scale(a, z));
  const float rho = sqrtf(dot(rad, rad));
  const float dz = fabsf(z) - C.s.y, dr = rho - C.s.x;
  const float oz = fmaxf(dz, 0.0f), orr = fmaxf(dr, 0.0f);
  const float out2 = fmaf(oz, oz, orr * orr);
  return (out2 > 0.0f) ? sqrtf(out2) : fmaxf(dz, dr);
}

// ----------------------------------------------------------------------------------------
// SDF mode: exact point-to-primitive signed distances (grid baking) and the trilinear read
// ----------------------------------------------------------------------------------------
SSPP_HD float point_prim_sd(v3 x, const Prim& G) {
  if (G.type == GT_PLANE) return dot(col(G.R, 2), sub(x, G.p));
  if (G.type == GT_SPHERE) { const v3 d = sub(x, G.p); return sqrtf(dot(d, d)) - G.s.x; }
  if (G.type == GT_CAPSULE) {
    const v3 ax = col(G.R, 2);
    const v3 a0 = madd(ax, -G.s.y, G.p), dd = scale(ax, 2.0f * G.s.y);
    return sqrtf(seg_point_dist2(a0, dd, dot(dd, dd), x)) - G.s.x;
  }
  if (G.type == GT_CYLINDER) return point_cyl_sd(x, G);
  return point_box_sd(multv(G.R, sub(x, G.p)), G.s);
}

// cell index and fractions of a world point; points outside the grid read the boundary cell
SSPP_HD void sdf_locate(const SdfGrid& g, v3 x, int* ix, int* iy, int* iz, float* fx, float* fy, float* fz) {
  const float gx = clampf((x.x - g.lo[0]) * g.inv_h[0], 0.0f, (float)(g.nx - 1));
  const float gy = clampf((x.y - g.lo[1]) * g.inv_h[1], 0.0f, (float)(g.ny - 1));
  const float gz = clampf((x.z - g.lo[2]) * g.inv_h[2], 0.0f, (float)(g.nz - 1));
  int i = (int)gx, j = (int)gy, k = (int)gz;
  if (i > g.nx - 2) i = g.nx - 2;
  if (j > g.ny - 2) j = g.ny - 2;
  if (k > g.nz - 2) k = g.nz - 2;
  *ix = i; *iy = j; *iz = k;
  *fx = gx - (float)i; *fy = gy - (float)j; *fz = gz - (float)k;
}
// corners ordered v000 v100 v010 v110 v001 v101 v011 v111 (x fastest)
SSPP_HD float trilerp(const float v[8], float fx, float fy, float fz) {
  const float c00 = fmaf(fx, v[1] - v[0], v[0]), c10 = fmaf(fx, v[3] - v[2], v[2]);
  const float c01 = fmaf(fx, v[5] - v[4], v[4]), c11 = fmaf(fx, v[7] - v[6], v[6]);
  const float c0 = fmaf(fy, c10 - c00, c00), c1 = fmaf(fy, c11 - c01, c01);
  return fmaf(fz, c1 - c0, c0);
}
// host / replica read from the node array
SSPP_HD float sdf_sample_nodes(const SdfGrid& g, v3 x) {
  int i, j, k; float fx, fy, fz;
  sdf_locate(g, x, &i, &j, &k, &fx, &fy, &fz);
  float v[8];
#pragma unroll
  for (int c = 0; c < 8; ++c)
    v[c] = g.nodes[((size_t)(k + (c >> 2)) * g.ny + (j + ((c >> 1) & 1))) * g.nx + (i + (c & 1))];
  return trilerp(v, fx, fy, fz);
}
// kernel read: the 8 corners of a cell are one 32-byte sector, fetched with one 256-bit read-only load
// (a warp's 32 scattered cells then cost one pass through the L1 tag stage instead of two)
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ void sdf_load_cell(const float* c, float v[8]) {
  asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(c));
}
#endif
SSPP_HD float sdf_sample(const SdfGrid& g, v3 x) {
#if defined(__CUDA_ARCH__)
  int i, j, k; float fx, fy, fz;
  sdf_locate(g, x, &i, &j, &k, &fx, &fy, &fz);
  float v[8];
  sdf_load_cell(g.cells + 8 * (((size_t)k * (g.ny - 1) + j) * (g.nx - 1) + i), v);
  return trilerp(v, fx, fy, fz);
#else
  return sdf_sample_nodes(g, x);
#endif
}
#ifndef SDF_BATCH
#define SDF_BATCH 2   // cell reads in flight per thread (costs are still added in sample order)
#endif
// contribution of one sphere sample at signed distance d
SSPP_HD float sphere_contact_cost(int rule, float d, float margin) {
  if (rule == CR_COUNT) return (d < margin) ? 1.0f : 0.0f;
  return (d < margin && d < -1e-3f) ? -d : 0.0f;
}

// signed distance between two world primitives; *ok = false for an unsupported pairing
SSPP_HD float prim_sd(const Prim& A, const Prim& B, bool* ok) {
  *ok = true;
  const bool sw = A.type > B.type;      // order by type; swap by value (keeps both in registers)
  const Prim Pv = sw ? B : A, Qv = sw ? A : B;
  const Prim* P = &Pv; const Prim* Q = &Qv;
  const int tp = P->type, tq = Q->type;
  if (tp == GT_PLANE) {
    const v3 n = col(P->R, 2);
    const float d0 = dot(n, sub(Q->p, P->p));
    if (tq == GT_SPHERE) return d0 - Q->s.x;
    if (tq == GT_CAPSULE) return d0 - Q->s.y * fabsf(dot(n, col(Q->R, 2))) - Q->s.x;
    if (tq == GT_CYLINDER) {
      const float na = dot(n, col(Q->R, 2));
      return d0 - Q->s.y * fabsf(na) - Q->s.x * sqrtf(fmaxf(fmaf(-na, na, 1.0f), 0.0f));
    }
    if (tq == GT_BOX)
      return d0 - fmaf(Q->s.z, fabsf(dot(n, col(Q->R, 2))),
                       fmaf(Q->s.y, fabsf(dot(n, col(Q->R, 1))), Q->s.x * fabsf(dot(n, col(Q->R, 0)))));
    *ok = false; return 3.0e38f;
  }
  if (tp == GT_SPHERE) {
    if (tq == GT_SPHERE) { const v3 d = sub(Q->p, P->p); return sqrtf(dot(d, d)) - P->s.x - Q->s.x; }
    if (tq == GT_CAPSULE) {
      const v3 ax = col(Q->R, 2);
      const v3 a0 = madd(ax, -Q->s.y, Q->p), dd = scale(ax, 2.0f * Q->s.y);
      return sqrtf(seg_point_dist2(a0, dd, dot(dd, dd), P->p)) - P->s.x - Q->s.x;
    }
    if (tq == GT_CYLINDER) return point_cyl_sd(P->p, *Q) - P->s.x;
    if (tq == GT_BOX) return point_box_sd(multv(Q->R, sub(P->p, Q->p)), Q->s) - P->s.x;
  }
  if (tp == GT_CAPSULE) {
    const v3 ax = col(P->R, 2);
    if (tq == GT_CAPSULE) {
      const v3 bx = col(Q->R, 2);
      const float d2 = seg_seg_dist2(madd(ax, -P->s.y, P->p), scale(ax, 2.0f * P->s.y),
                                     madd(bx, -Q->s.y, Q->p), scale(bx, 2.0f * Q->s.y));
      return sqrtf(d2) - P->s.x - Q->s.x;
    }
    if (tq == GT_BOX) {
      const v3 c = multv(Q->R, sub(P->p, Q->p));
      const v3 al = multv(Q->R, ax);
      return seg_box_dist(madd(al, -P->s.y, c), scale(al, 2.0f * P->s.y), Q->s) - P->s.x;
    }
  }
  if (tp == GT_CYLINDER && tq == GT_BOX) return cyl_box_sd(*P, *Q);
  if (tp == GT_BOX && tq == GT_BOX) return box_box_sd(*P, *Q);
  *ok = false; return 3.0e38f;
}

// Signed distance of a MOVING primitive P to a STATIC one S for the pairings the scenes use, without
// the by-value operand swap of prim_sd, and with conservative early exits: as soon as a cheap lower
// bound of the distance exceeds thr (> the pair's margin) that bound is returned instead of the exact
// distance — the contact rule maps both to "no contact".  Other pairings fall back to prim_sd.
SSPP_HD float moving_static_sd(const Prim& P, const Prim& S, float thr, uint32_t* wc = nullptr) {
  if (S.type == GT_PLANE) {
    if (wc && (P.type == GT_SPHERE || P.type == GT_CAPSULE || P.type == GT_CYLINDER || P.type == GT_BOX)) ++wc[WC_NP_PLANE];
    const v3 n = col(S.R, 2);
    const float d0 = dot(n, sub(P.p, S.p));
    if (P.type == GT_SPHERE) return d0 - P.s.x;
    if (P.type == GT_CAPSULE) return d0 - P.s.y * fabsf(dot(n, col(P.R, 2))) - P.s.x;
    if (P.type == GT_CYLINDER) {
      const float na = dot(n, col(P.R, 2));
      return d0 - P.s.y * fabsf(na) - P.s.x * sqrtf(fmaxf(fmaf(-na, na, 1.0f), 0.0f));
    }
    if (P.type == GT_BOX)
      return d0 - fmaf(P.s.z, fabsf(dot(n, col(P.R, 2))),
                       fmaf(P.s.y, fabsf(dot(n, col(P.R, 1))), P.s.x * fabsf(dot(n, col(P.R, 0)))));
  }
  if (S.type == GT_BOX) {
    if (P.type == GT_CAPSULE) {
      const v3 c = multv(S.R, sub(P.p, S.p));
      const float lb = point_box_sd(c, S.s) - P.s.y - P.s.x;     // centre distance minus half length minus radius
      if (wc) ++wc[lb > thr ? WC_NP_CAPS_BOX_EARLY : WC_NP_CAPS_BOX];
      if (lb > thr) return lb;
      const v3 al = multv(S.R, col(P.R, 2));
      return seg_box_dist(madd(al, -P.s.y, c), scale(al, 2.0f * P.s.y), S.s, wc) - P.s.x;
    }
    if (P.type == GT_BOX) return box_box_sd(P, S, thr, wc);
    if (P.type == GT_CYLINDER) return cyl_box_sd(P, S, thr, wc);
    if (P.type == GT_SPHERE) { if (wc) ++wc[WC_NP_SPHERE_BOX]; return point_box_sd(multv(S.R, sub(P.p, S.p)), S.s) - P.s.x; }
  }
  if (wc && S.type != GT_PLANE) ++wc[WC_NP_OTHER];
  bool ok;
  return prim_sd(P, S, &ok);
}

// contribution of one geom pair at signed distance d (centres ca, cb) under the model's rule
SSPP_HD float contact_cost(int rule, float d, float margin, v3 ca, v3 cb) {
  if (rule == CR_COUNT) return (d < margin) ? 1.0f : 0.0f;
  if (rule == CR_DEPTH) return (d < margin && d < -1e-3f) ? -d : 0.0f;
  if (d < margin && d < -1e-3f) {
    const v3 w = sub(cb, ca);
    return -1.0f / (sqrtf(dot(w, w)) + 1e-4f);
  }
  return 0.0f;
}

// moving geom g of table `mg` under the body frame (R, p)
template <class TAB>
SSPP_HD Prim place(const TAB& mg, int g, const m3& R, v3 p) {
  Prim P;
  const f4 gp = rec_f4(mg, g, GW_POS);
  const m3t rs = rec_rot12(mg, g, GW_R);
  P.p = add(p, mulv(R, mk(gp.x, gp.y, gp.z)));
  P.R = mulm(R, rs.R.a);
  P.s = rs.t;
  P.type = rec_i4(mg, g, GW_TYPE).x;
  return P;
}
// orientation and size of a placement, as little as the type needs: capsules and cylinders only use their
// axis (column 2), spheres nothing; element formulas are those of mulm
template <class TAB>
SSPP_HD void place_rotation(const TAB& mg, int g, int type, const m3& R, Prim* P) {
  if (type == GT_BOX) { const m3t rs = rec_rot12(mg, g, GW_R); P->s = rs.t; P->R = mulm(R, rs.R.a); return; }
  const f4 a = rec_f4(mg, g, GW_R), b = rec_f4(mg, g, GW_R + 4), c = rec_f4(mg, g, GW_R + 8);   // R0..3 | R4..7 | R8, size
  P->s = mk(c.y, c.z, c.w);
  if (type == GT_SPHERE) return;
  const float g2 = a.z, g5 = b.y, g8 = c.x;
#pragma unroll
  for (int i = 0; i < 3; ++i)
    P->R.a[3 * i + 2] = fmaf(R.a[3 * i + 2], g8, fmaf(R.a[3 * i + 1], g5, R.a[3 * i] * g2));
}
template <class TAB>
SSPP_HD Prim static_prim(const TAB& sg, int i) {
  Prim P;
  const f4 a = rec_f4(sg, i, GW_POS);
  const m3t rs = rec_rot12(sg, i, GW_R);
  P.p = mk(a.x, a.y, a.z); P.R = rs.R; P.s = rs.t; P.type = rec_i4(sg, i, GW_TYPE).x;
  return P;
}
SSPP_HD Prim static_prim(const DevSGeom& g) { return static_prim(&g, 0); }

// Collision cost of one configuration q: FK of every moving body, placement of its geoms,
// bounding-sphere cull, narrow phase, contact rule.  XF is a transform store with
// put(body, k, value) / get(body, k), k in 0..11 (R row-major then p) — shared memory on the
// GPU, a plain array in the replica; unused when model.chain is set.
// QA exposes the planned coordinates as q.get(i).
// The broad phase of a geom runs branch-free over its pair records, 32 at a time, into a bit mask (plane
// partners first, then bounded ones); the narrow phase then visits the set bits in ascending order, so the
// contact costs are summed in pair-table order.
// ssppcu_check_points (Collision.h:51-82 check_collision_point): of the contacts of one configuration that penetrate by more
// than 1 mm (dist < margin and dist < -1e-3, the reference's test on mjData::contact) the one with the smallest geom pair
// (g1 << 16 | g2, FlatModel geom ids, g1 < g2).  Only the diagnostic kernel passes a probe; the scoring paths pass none.
struct ContactProbe {
  const int* mg_src;   // DevModel::mg index -> FlatModel geom id
  const int* sg_src;   // DevModel::sg index -> FlatModel geom id
  uint32_t key;        // 0xFFFFFFFF: no such contact
  float dist;
};
SSPP_HD void probe_contact(ContactProbe* cp, int ga, int gb, float d, float margin) {
  if (!(d < margin && d < -1e-3f)) return;
  const uint32_t lo = (uint32_t)(ga < gb ? ga : gb), hi = (uint32_t)(ga < gb ? gb : ga);
  const uint32_t key = (lo << 16) | hi;
  if (key < cp->key) { cp->key = key; cp->dist = d; }
}

template <class MT, class XF, class QA>
SSPP_HD float point_collision_cost(const MT& M, QA& q, XF& xf, uint32_t* wc = nullptr, ContactProbe* cp = nullptr) {
  float cost = M.static_cost;
  if (wc) ++wc[WC_POINTS];
  m3 R; v3 p = mk(0.f, 0.f, 0.f);
#pragma unroll
  for (int k = 0; k < 9; ++k) R.a[k] = 0.0f;
#pragma unroll 1
  for (int b = 0; b < M.n_mbody; ++b) {
    const i4 jt = rec_i4(M.body, b, BW_JOINT);               // parent, jtype, qidx, axis_kind
    m3 Rp; v3 pp;
    const bool hp = jt.x >= 0;
    if (hp) {
      if (M.chain) { Rp = R; pp = p; }
      else {
#pragma unroll
        for (int k = 0; k < 9; ++k) Rp.a[k] = xf.get(jt.x, k);
        pp = mk(xf.get(jt.x, 9), xf.get(jt.x, 10), xf.get(jt.x, 11));
      }
    } else { pp = mk(0.f, 0.f, 0.f); Rp = R; }
    fk_body(M.body, b, jt, hp, Rp, pp, q, &R, &p, wc);
    if (!M.chain) {
#pragma unroll
      for (int k = 0; k < 9; ++k) xf.put(b, k, R.a[k]);
      xf.put(b, 9, p.x); xf.put(b, 10, p.y); xf.put(b, 11, p.z);
    }
    if (M.sdf_mode) {                                        // sphere samples of this body against the grid
      const i4 sp = rec_i4(M.body, b, BW_SPH);          // sph_begin, sph_end
      if (wc) wc[WC_SDF_SAMPLES] += (uint32_t)(sp.y - sp.x);
#if defined(__CUDA_ARCH__) && SDF_BATCH > 1
#pragma unroll 1
      for (int k0 = sp.x; k0 < sp.y; k0 += SDF_BATCH) {
        float v[SDF_BATCH][8];
        float fx[SDF_BATCH], fy[SDF_BATCH], fz[SDF_BATCH];
#pragma unroll
        for (int u = 0; u < SDF_BATCH; ++u) {
          const int k = (k0 + u < sp.y) ? k0 + u : sp.y - 1;
          const f4 sc = rec_f4(M.sph, k, SW_C);
          int i, j, l;
          sdf_locate(M.grid, add(p, mulv(R, mk(sc.x, sc.y, sc.z))), &i, &j, &l, &fx[u], &fy[u], &fz[u]);
          sdf_load_cell(M.grid.cells + 8 * (((size_t)l * (M.grid.ny - 1) + j) * (M.grid.nx - 1) + i), v[u]);
        }
#pragma unroll
        for (int u = 0; u < SDF_BATCH; ++u) {
          if (k0 + u < sp.y) {
            const i4 rm = rec_i4(M.sph, k0 + u, SW_C);
            const float margin = word_as_float(rec_i4(M.sph, k0 + u, SW_MARGIN).x);
            cost += sphere_contact_cost(M.rule, trilerp(v[u], fx[u], fy[u], fz[u]) - word_as_float(rm.w), margin);
          }
        }
      }
#else
#pragma unroll 1
      for (int k = sp.x; k < sp.y; ++k) {
        const f4 sc = rec_f4(M.sph, k, SW_C);                // centre, radius
        const float margin = word_as_float(rec_i4(M.sph, k, SW_MARGIN).x);
        const float d = sdf_sample(M.grid, add(p, mulv(R, mk(sc.x, sc.y, sc.z)))) - sc.w;
        cost += sphere_contact_cost(M.rule, d, margin);
      }
#endif
      continue;
    }
    const i4 fl = rec_i4(M.body, b, BW_FLAGS);               // has_anchor, r0_identity, geom_begin, geom_end
    {
      const i4 bp = rec_i4(M.body, b, BW_BPAIR);             // bpair_begin, bpair_mid, bpair_end
      if (bp.z > bp.x) {                                     // body-level broad phase: is the whole body clear of all its partners?
        const f4 bb = rec_f4(M.body, b, BW_BOUND);
        const v3 bc = add(p, mulv(R, mk(bb.x, bb.y, bb.z)));
        const float nbb = -dot(bc, bc);
        bool near = false;
#pragma unroll 1
        for (int k = bp.x; k < bp.y; ++k) {
          const f4 c4 = rec_f4(M.pair, k, PW_C);
          near = near || !(fmaf(c4.z, bc.z, fmaf(c4.y, bc.y, fmaf(c4.x, bc.x, c4.w))) > 0.0f);
        }
#pragma unroll 1
        for (int k = bp.y; k < bp.z; ++k) {
          const f4 c4 = rec_f4(M.pair, k, PW_C);
          near = near || !(fmaf(c4.z, bc.z, fmaf(c4.y, bc.y, fmaf(c4.x, bc.x, c4.w))) > nbb);
        }
        if (wc) wc[WC_BODY_BROAD] += (uint32_t)(bp.z - bp.x);
        if (!near) continue;
      }
    }
#pragma unroll 1
    for (int g = fl.z; g < fl.w; ++g) {
      const f4 gp = rec_f4(M.mg, g, GW_POS);                 // pos, rbound
      const i4 gi = rec_i4(M.mg, g, GW_TYPE);                // type, pair_begin, pair_mid, pair_end (<= 32 pairs)
      if (gi.w == gi.y) continue;                            // every partner was pruned when the scene was compiled: nothing to place
      Prim P;
      P.p = add(p, mulv(R, mk(gp.x, gp.y, gp.z)));
      P.type = gi.x;
      bool rotated = false;
      if (wc) { ++wc[WC_GEOM_POS]; wc[WC_BROAD] += (uint32_t)(gi.w - gi.y); }
      {
        const int k0 = gi.y;
        uint32_t near_mask = 0u;
        const float npp = -dot(P.p, P.p);
#pragma unroll 1
        for (int k = k0; k < gi.z; ++k) {                    // plane partners: bounding sphere beyond the half space?
          const f4 c4 = rec_f4(M.pair, k, PW_C);
          if (!(fmaf(c4.z, P.p.z, fmaf(c4.y, P.p.y, fmaf(c4.x, P.p.x, c4.w))) > 0.0f)) near_mask |= 1u << (k - k0);
        }
#pragma unroll 1
        for (int k = gi.z; k < gi.w; ++k) {                  // bounded partners: bounding spheres apart?
          const f4 c4 = rec_f4(M.pair, k, PW_C);
          if (!(fmaf(c4.z, P.p.z, fmaf(c4.y, P.p.y, fmaf(c4.x, P.p.x, c4.w))) > npp)) near_mask |= 1u << (k - k0);
        }
        while (near_mask) {
          const int j = lowest_bit(near_mask);
          near_mask &= near_mask - 1u;
          const i4 t4 = rec_i4(M.pair, k0 + j, PW_S);        // s, margin, thr, plane (floats as raw words)
          const float margin = word_as_float(t4.y), thr = word_as_float(t4.z);
          const Prim Q = static_prim(M.sg, t4.x);
          if (Q.type == GT_BOX) {
            // second-level cull before anything of the moving geom is oriented: a long thin box (a wall) has a bounding sphere
            // that most of the workspace lies inside, so its pairs pass the sphere test at every point; the geom lies inside the
            // sphere (P.p, rbound), hence its distance to the box is at least dist(P.p, box) - rbound
            if (wc) ++wc[WC_NP_BOX_CULL];
            if (point_box_sd(multv(Q.R, sub(P.p, Q.p)), Q.s) - gp.w > thr) continue;
          }
          if (!rotated) { place_rotation(M.mg, g, gi.x, R, &P); rotated = true; if (wc && gi.x != GT_SPHERE) ++wc[gi.x == GT_BOX ? WC_GEOM_ROT_BOX : WC_GEOM_ROT_AXIS]; }
          const float d = moving_static_sd(P, Q, thr, wc);
          const float cc = contact_cost(M.rule, d, margin, P.p, Q.p);
          if (wc && cc != 0.0f) ++wc[WC_CONTACTS];
          if (cp) probe_contact(cp, cp->mg_src[g], cp->sg_src[t4.x], d, margin);
          cost += cc;
        }
      }
    }
  }
#pragma unroll 1
  for (int k = 0; k < M.n_mmpair; ++k) {   // pairs of geoms that move relative to each other
    const i4 mp = rec_i4(M.mm, k, 0);      // a, b, margin
    const float margin = word_as_float(mp.z);
    const int ba = rec_i4(M.mg, mp.x, GW_BODY).x, bb = rec_i4(M.mg, mp.y, GW_BODY).x;
    const float ra = rec_f4(M.mg, mp.x, GW_POS).w, rb = rec_f4(M.mg, mp.y, GW_POS).w;
    m3 Ra, Rb; v3 pa, pb;
#pragma unroll
    for (int e = 0; e < 9; ++e) { Ra.a[e] = xf.get(ba, e); Rb.a[e] = xf.get(bb, e); }
    pa = mk(xf.get(ba, 9), xf.get(ba, 10), xf.get(ba, 11));
    pb = mk(xf.get(bb, 9), xf.get(bb, 10), xf.get(bb, 11));
    const Prim P = place(M.mg, mp.x, Ra, pa), Q = place(M.mg, mp.y, Rb, pb);
    const v3 w = sub(P.p, Q.p);
    const float rr = ra + rb + fmaxf(margin, 0.0f) + 1e-4f;
    if (wc) ++wc[WC_MM_TESTS];
    if (dot(w, w) > rr * rr) continue;
    if (wc) ++wc[WC_MM_NARROW];
    bool ok;
    const float d = prim_sd(P, Q, &ok);
    if (cp) probe_contact(cp, cp->mg_src[mp.x], cp->mg_src[mp.y], d, margin);
    cost += contact_cost(M.rule, d, margin, P.p, Q.p);
  }
  return cost;
}

// SDF mode, refinement step: the sphere sample closest to the static scene at configuration q, its clearance
// (field value - radius) and the gradient of the trilinear interpolant there (d clearance / d world position).
// Restates get_closest_points of /root/reference/scripts/sdf_gradient/main.py:37-63 on the grid instead of
// mj_geomDistance: the minimum over (moving geom x environment geom) pairs becomes the minimum over sample spheres
// of the baked field, the witness direction becomes the field gradient.
SSPP_HD v3 sdf_gradient_at(const SdfGrid& g, v3 x, float* value) {
  int i, j, k; float fx, fy, fz;
  sdf_locate(g, x, &i, &j, &k, &fx, &fy, &fz);
  float v[8];
#if defined(__CUDA_ARCH__)
  sdf_load_cell(g.cells + 8 * (((size_t)k * (g.ny - 1) + j) * (g.nx - 1) + i), v);
#else
  for (int c = 0; c < 8; ++c)
    v[c] = g.nodes[((size_t)(k + (c >> 2)) * g.ny + (j + ((c >> 1) & 1))) * g.nx + (i + (c & 1))];
#endif
  *value = trilerp(v, fx, fy, fz);
  const float a0 = v[1] - v[0], a1 = v[3] - v[2], a2 = v[5] - v[4], a3 = v[7] - v[6];
  const float ax0 = fmaf(fy, a1 - a0, a0), ax1 = fmaf(fy, a3 - a2, a2);
  const float c00 = fmaf(fx, a0, v[0]), c10 = fmaf(fx, a1, v[2]), c01 = fmaf(fx, a2, v[4]), c11 = fmaf(fx, a3, v[6]);
  const float b0 = c10 - c00, b1 = c11 - c01;
  const float c0 = fmaf(fy, b0, c00), c1 = fmaf(fy, b1, c01);
  return mk(fmaf(fz, ax1 - ax0, ax0) * g.inv_h[0], fmaf(fz, b1 - b0, b0) * g.inv_h[1], (c1 - c0) * g.inv_h[2]);
}

template <class MT, class XF, class QA>
SSPP_HD void point_closest_sample(const MT& M, QA& q, XF& xf, float* clearance, v3* grad) {
  float best = 3.0e38f;
  v3 best_g = mk(0.f, 0.f, 0.f);
  m3 R; v3 p = mk(0.f, 0.f, 0.f);
  for (int k = 0; k < 9; ++k) R.a[k] = 0.0f;
  for (int b = 0; b < M.n_mbody; ++b) {
    const i4 jt = rec_i4(M.body, b, BW_JOINT);
    m3 Rp; v3 pp;
    const bool hp = jt.x >= 0;
    if (hp) {
      if (M.chain) { Rp = R; pp = p; }
      else {
        for (int k = 0; k < 9; ++k) Rp.a[k] = xf.get(jt.x, k);
        pp = mk(xf.get(jt.x, 9), xf.get(jt.x, 10), xf.get(jt.x, 11));
      }
    } else { pp = mk(0.f, 0.f, 0.f); Rp = R; }
    fk_body(M.body, b, jt, hp, Rp, pp, q, &R, &p);
    if (!M.chain) {
      for (int k = 0; k < 9; ++k) xf.put(b, k, R.a[k]);
      xf.put(b, 9, p.x); xf.put(b, 10, p.y); xf.put(b, 11, p.z);
    }
    const i4 sp = rec_i4(M.body, b, BW_SPH);
    for (int k = sp.x; k < sp.y; ++k) {
      const f4 sc = rec_f4(M.sph, k, SW_C);
      float val;
      const v3 g = sdf_gradient_at(M.grid, add(p, mulv(R, mk(sc.x, sc.y, sc.z))), &val);
      if (val - sc.w < best) { best = val - sc.w; best_g = g; }
    }
  }
  *clearance = best; *grad = best_g;
}

// value of grid node (i,j,k): min over the static primitives (shared by the bake kernel and the replica)
template <class MT>
SSPP_HD float sdf_node_value(const MT& M, int i, int j, int k) {
  const SdfGrid& g = M.grid;
  const v3 x = mk(fmaf((float)i, g.h[0], g.lo[0]), fmaf((float)j, g.h[1], g.lo[1]), fmaf((float)k, g.h[2], g.lo[2]));
  float best = 3.0e38f;
  for (int s = 0; s < M.n_sgeom; ++s) best = fminf(best, point_prim_sd(x, static_prim(M.sg, s)));
  return best;
}

// floor penalty (include/sspp/tsp_evaluator.h:10-14) and its joint-space analogue
SSPP_HD float floor_penalty(float z, float z_min, float margin, float scl) {
  const float deficit = (z_min + margin) - z;
  return (deficit > 0.0f) ? scl * deficit * deficit : 0.0f;
}
SSPP_HD float limit_penalty(float q, float lim, float scl) {
  const float ex = fabsf(q) - lim;
  return (ex > 0.0f) ? scl * ex * ex : 0.0f;
}

// ----------------------------------------------------------------------------------------
// banded collocation rows: a degree-p B-spline row has p+1 consecutive non-zeros, so the
// kernel stores [rows][W] coefficients plus the first column of every row.  Skipping exact
// zeros leaves the fmaf chain's result unchanged bit for bit (acc starts at +0 and can never
// become -0), so the banded and the dense sums are the same number.
// ----------------------------------------------------------------------------------------
SSPP_HD float spline_coord(const float* row, int first, int W, const float* X, int D, int d) {
  float acc = 0.0f;
  for (int k = 0; k < W; ++k) acc = fmaf(row[k], X[(first + k) * D + d], acc);
  return acc;
}
// host helper: band of a dense [rows][K] matrix -> width, first column per row (first + W <= K)
inline int compute_band(const float* psi, int rows, int K, int* first) {
  int W = 1;
  for (int m = 0; m < rows; ++m) {
    int lo = K, hi = -1;
    for (int k = 0; k < K; ++k) if (psi[m * K + k] != 0.0f) { if (k < lo) lo = k; hi = k; }
    if (hi < 0) { lo = 0; hi = 0; }
    first[m] = lo;
    if (hi - lo + 1 > W) W = hi - lo + 1;
  }
  for (int m = 0; m < rows; ++m) if (first[m] + W > K) first[m] = K - W;
  return W;
}
inline void pack_band(const float* psi, int rows, int K, int W, const int* first, float* out /* [rows][W] */) {
  for (int m = 0; m < rows; ++m)
    for (int k = 0; k < W; ++k) out[m * W + k] = psi[m * K + first[m] + k];
}

// order-preserving map float -> uint32 used for the (cost, index) sort keys
SSPP_HD uint32_t float_order_bits(float f) {
  union { float f; uint32_t u; } b; b.f = f;
  return (b.u & 0x80000000u) ? ~b.u : (b.u | 0x80000000u);
}

// ----------------------------------------------------------------------------------------
// planner configuration shared by the kernel, the C-ABI and the replica oracle
// ----------------------------------------------------------------------------------------
enum { MODE_JOINT = 0,    // planner A: Gaussian noise on interior control points (sspp.h:114-130)
       MODE_TASK = 1 };   // planner B: truncated-normal via points (tsp_sampler.h:12-51)
enum { ST_CONVERGED = 0, ST_FAILED = 1, ST_ABSENT = 2 };   // tsp_types.h:16 (+ empty slot)

struct ScoreConfig {
  int mode;
  int D, K;               // coordinates per point, points per candidate (control points / interpolation points)
  int n_free, free_begin; // sampled rows [free_begin, free_begin + n_free) of the K x D point matrix
  int Mc, col_first;      // rows of the collision grid; rows < col_first are not collision checked
  int Ma, arc_same;       // rows of the arc-length grid (ignored when arc_same: the collision grid is used)
  int n_cand;             // candidates per query and iteration, special slots included
  int include_mean;       // slot 0 is the distribution mean      (tsp_planner.h:82-85)
  int include_best;       // next slot is the forwarded best path (tsp_planner.h:87)
  int do_update;          // CES update of mean / sigma           (tsp_planner.h:122-141)
  float w_collision;      // cost = L + w_collision * C_wf        (tsp_planner.h:123)
  float eval_zmin, floor_margin, floor_scale;   // Evaluator members (tsp_evaluator.h:8)
  float limit_scale;      // joint-space analogue of the floor penalty: scale*max(0,|q_d|-hi_d)^2, 0 = off
  float sample_zmin;      // cfg_.z_min applied to sampled via z   (tsp_sampler.h:47)
  float dist_zmin;        // dist_.z_min applied to the mean z     (tsp_distribution.h:64)
  float inc, dec, sigma_min, sigma_max, sigma_floor, var_beta, mean_lr;
  float lo[MAX_DOF], hi[MAX_DOF];   // sampler limits and mean clamp
  double elite_frac;
};

// what candidate slot `cand` holds
enum { CK_MEAN = 0, CK_BEST = 1, CK_SAMPLE = 2, CK_ABSENT = 3 };
SSPP_HD int candidate_kind(const ScoreConfig& c, int cand, int best_valid) {
  if (c.include_mean && cand == 0) return CK_MEAN;
  if (c.include_best && cand == c.include_mean) return best_valid ? CK_BEST : CK_ABSENT;
  return CK_SAMPLE;
}

// element t = row * D + dim of candidate `cand` (the sampled rows only)
SSPP_HD float candidate_param(const ScoreConfig& c, uint64_t seed, uint32_t iter, uint32_t query, int cand,
                              int kind, int t, const float* mean, const float* sigma, const float* best) {
  const int r = t / c.D, d = t - r * c.D;
  float v;
  if (kind == CK_MEAN) v = mean[t];
  else if (kind == CK_BEST) return best[t];
  else if (kind == CK_ABSENT) return mean[t];
  else if (c.mode == MODE_JOINT) {
    // draw order of sspp.h:121-127: coordinate-major, then control point
    v = fmaf(normal_at(seed, iter, query, (uint32_t)cand, 0u, (uint32_t)(d * c.n_free + r)), sigma[t], mean[t]);
    return v;
  } else v = sample_task_dim(seed, iter, query, (uint32_t)cand, r, d, mean[t], sigma[t], c.lo[d], c.hi[d]);
  if (c.mode == MODE_TASK && d == 2) v = fmaxf(v, c.sample_zmin);
  return v;
}

// per-point penalty added to C_wf
template <class QA>
SSPP_HD float point_penalty(const ScoreConfig& c, QA& q) {
  if (c.mode == MODE_TASK) return floor_penalty(q.get(2), c.eval_zmin, c.floor_margin, c.floor_scale);
  float pen = 0.0f;
  if (c.limit_scale > 0.0f)
    for (int d = 0; d < c.D; ++d) pen += limit_penalty(q.get(d), c.hi[d], c.limit_scale);
  return pen;
}

SSPP_HD float clamp_sigma(const ScoreConfig& c, float s) {   // tsp_distribution.h:78-81 order
  return fmaxf(fminf(fmaxf(s, c.sigma_min), c.sigma_max), c.sigma_floor);
}
SSPP_HD float wrap_angle_diff(float a, float b, float lo, float hi) {   // tsp_distribution.h:40-46
  const float range = hi - lo;
  float d = a - b;
  for (int it = 0; it < 64 && d > 0.5f * range; ++it) d -= range;
  for (int it = 0; it < 64 && d < -0.5f * range; ++it) d += range;
  return d;
}
SSPP_HD uint64_t sort_key(float cost, int status, uint32_t idx) {
  const uint32_t hi = (status == ST_CONVERGED) ? float_order_bits(cost) : 0xFFFFFFFFu;
  return ((uint64_t)hi << 32) | idx;
}
SSPP_HD int elite_count(int n_success, double frac) {   // tsp_elites.h:15
  if (n_success <= 0) return 0;
  const int k = (int)((double)n_success * frac);
  return k < 1 ? 1 : k;
}

}  // namespace ssppm
