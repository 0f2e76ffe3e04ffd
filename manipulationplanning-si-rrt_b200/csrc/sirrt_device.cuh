// sirrt_device.cuh -- device-side data layout and geometry for libsirrt_b200 (sm_100a).
//
// Everything here is FP32/FP64 CUDA-core arithmetic: the path is rigid-body geometry, not a dense
// contraction, so there is deliberately no tensor-core (tcgen05) code.  The same templates serve the
// FP32 fast path and the FP64 re-check so both evaluate the same formulas.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SIRRT_MAXJ 12  // non-fixed joints of the planned robot
#define SIRRT_MAXL 12  // collision links of the planned robot after the base strip
#define SIRRT_FULL 0xffffffffu

enum { SH_CAPSULE = 0, SH_HULL = 1, SH_BOX = 2, SH_SPHERE = 3 };
enum { RES_FREE = 0, RES_HIT = 1, RES_AMBIG = 2 };

// coal's collision predicate: distance - security_margin <= collision_distance_threshold
// (security_margin 1e-21: CollisionManager.cpp:559,611,654; threshold = Eigen dummy_precision)
#define SIRRT_SECURITY_MARGIN 1e-21
#define SIRRT_COLLISION_THRESHOLD 1e-12

#ifdef SIRRT_GJK_STATS
// diagnostics build: FP32 GJK iterations per task, by verdict (0 FREE, 1 HIT, 2 AMBIGUOUS, 3 AMBIGUOUS with a thin enclosing tetrahedron)
__device__ unsigned int g_gjk_hist[4][64];
// cycles per phase, summed over warps: [0] k_collide_bits warps, [1] FK + link table, [2] static slots, [3] dynamic tiles,
// [4] whole item, [5] angle load + sincos, [6] FK chain, [7] link table; [8] k_gjk_tasks tasks, [9] task prologue (up to the GJK call), [10] GJK + verdict, [11] kernel entry to first task,
// [12] GJK iterations, [13] hull support of A, [14] support of B + transforms, [15] simplex_closest (FP32 task GJK only)
__device__ unsigned long long g_phase[16];
#define PHASE_T(v) const long long v = clock64()
#define PHASE_ADD(i, d) do { if (lane == 0) atomicAdd(&g_phase[i], (unsigned long long)(d)); } while (0)
__device__ unsigned long long g_phase_max[16];  // largest single contribution per phase (the slowest warp decides a latency-bound launch)
#define PHASE_MAX(i, d) do { if (lane == 0) atomicMax(&g_phase_max[i], (unsigned long long)(d)); } while (0)
#else
#define PHASE_T(v)
#define PHASE_ADD(i, d)
#define PHASE_MAX(i, d)
#endif

// ---------------------------------------------------------------------------------------------
// Device tables
// ---------------------------------------------------------------------------------------------
// FK model of the planned robot (RobotObstacleFCL::fk_chain flattened; FP64, read by every
// kernel once per configuration).
struct DRobot {
  int n_joints, n_links, dof, pad;
  double base_t[3], base_R[9];
  double jt[SIRRT_MAXJ][3], jR[SIRRT_MAXJ][9];
  int link_jb[SIRRT_MAXL];      // joints applied before link l is emitted
  int link_shape[SIRRT_MAXL];   // index into the shape table
  double lo_t[SIRRT_MAXL][3], lo_R[SIRRT_MAXL][9];
};

// Geometry of one convex shape = core polytope + inflation radius.
struct DShape {
  int type;            // SH_*
  int vert_off;        // hull: first vertex in the vertex arrays
  int n_verts;
  int cell_off;        // hull: first entry of its support-cell offsets in DScene::hull_cells, -1 = none
  float radius, half_len;  // capsule / sphere
  float half[3];           // box half extents
  float bc[3], br;         // bounding sphere (inflated shape), local frame
  float amin[3], amax[3], ac[3], ar;  // coal aabb_local / aabb_center / aabb_radius
  double d_radius, d_half_len, d_half[3];
  double d_amin[3], d_amax[3], d_ac[3], d_ar;
};

// One moving thing the planned robot can hit: a box/sphere obstacle or one link of a robot obstacle.
struct DSlot {
  int shape;        // true geometry
  int shape_shift;  // QUIRK 7 functor geometry, -1 if none
  int is_static;    // identical pose at every frame
  int is_robot;     // robot-obstacle link
  int rot_identity; // every pose has R == I exactly (skip the rotation on the fast path)
  float br;         // bounding-sphere radius around bcw (covers the shifted geometry too)
  int type;         // shapes[shape].type / shapes[shape_shift].type (-1: none) -- saves the dependent load of the shape record
  int type_shift;   // where only "is a hull involved?" has to be decided
};

struct DScene {
  int T, Tp, n_tiles, n_slots;
  int n_static, n_dynamic, parity_mode, has_hull;
  const DShape* shapes;
  const DSlot* slots;
  const int* static_slots;   // slot ids
  const int* dynamic_slots;
  const float* pose_f;       // [(slot*12 + k)*Tp + f]  k<9: R row-major, k>=9: t
  const float* bcw_f;        // [(slot*3 + k)*Tp + f]   world bounding-sphere centre
  const float* tight_f;      // [(slot*6 + k)*Tp + f]   tight world AABB lo(3) hi(3) of the posed slot (covers the shifted geometry too)
  const double* pose_d;      // [(slot*Tp + f)*12 + k]
  const float4* tile_bounds; // [tile*n_dynamic + j]    swept bounding sphere of dynamic slot j over the tile
  const float* verts_f;      // xyz interleaved
  const double* verts_d;
  // support-candidate cells of the hulls (see hull_cell): per hull 6*K*K+1 offsets into hull_cand, which lists for
  // each cube-map cell of directions every vertex that can be the support vertex for a direction of that cell
  const unsigned int* hull_cells;
  const unsigned short* hull_cand;  // vertex indices local to the hull, ascending within a cell
  const float4* hull_rec;           // FP32 inline records parallel to hull_cells: SIRRT_HULL_REC x (x, y, z, index bits) per cell
};

// Cube-map cell of a (non-zero, finite) direction: dominant axis and sign pick the face, the two other components
// divided by the dominant magnitude pick one of K x K squares.  Host and device may round the ratios differently:
// the host dilates every cell by SIRRT_HULL_CELL_PAD in (u, v) when it collects the candidates.
#define SIRRT_HULL_K 16
#define SIRRT_HULL_CELLS (6 * SIRRT_HULL_K * SIRRT_HULL_K)
#define SIRRT_HULL_CELL_PAD 1e-3
#define SIRRT_HULL_REC 16   // candidate slots per inline FP32 cell record (DScene::hull_rec)
template <typename T> __host__ __device__ inline int hull_cell(const T* d) {
  const T ax = d[0] < T(0) ? -d[0] : d[0], ay = d[1] < T(0) ? -d[1] : d[1], az = d[2] < T(0) ? -d[2] : d[2];
  int a = 0;
  T m = ax;
  if (ay > m) { a = 1; m = ay; }
  if (az > m) { a = 2; m = az; }
  if (!(m > T(0)) || !(m < T(3.0e38))) return -1;  // zero / NaN / inf direction: caller scans every vertex
  // components picked by selects (a dynamically indexed d[] would live in local memory on the device), one reciprocal
  const T d1 = a == 0 ? d[1] : (a == 1 ? d[2] : d[0]), d2 = a == 0 ? d[2] : (a == 1 ? d[0] : d[1]), da = a == 0 ? d[0] : (a == 1 ? d[1] : d[2]);
  const T inv = T(1) / m;
  const T u = d1 * inv, v = d2 * inv;
  int iu = (int)((u + T(1)) * T(0.5 * SIRRT_HULL_K)), iv = (int)((v + T(1)) * T(0.5 * SIRRT_HULL_K));
  iu = iu < 0 ? 0 : (iu > SIRRT_HULL_K - 1 ? SIRRT_HULL_K - 1 : iu);
  iv = iv < 0 ? 0 : (iv > SIRRT_HULL_K - 1 ? SIRRT_HULL_K - 1 : iv);
  const int face = 2 * a + (da < T(0) ? 1 : 0);
  return (face * SIRRT_HULL_K + iu) * SIRRT_HULL_K + iv;
}

struct WorkItem {  // one ambiguous pair test for the FP64 re-check
  int config;
  int frame;       // -1: static slot, applies to every frame
  short link, shifted;
  int slot;
};

// ---------------------------------------------------------------------------------------------
// small algebra
// ---------------------------------------------------------------------------------------------
template <typename T> __host__ __device__ __forceinline__ T dot3(const T* a, const T* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
template <typename T> __host__ __device__ __forceinline__ void sub3(T* o, const T* a, const T* b) { o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2]; }
template <typename T> __host__ __device__ __forceinline__ void cross3(T* o, const T* a, const T* b) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
template <typename T> __host__ __device__ __forceinline__ void matvec(T* o, const T* R, const T* v) {
  o[0] = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
  o[1] = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
  o[2] = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
}
template <typename T> __device__ __forceinline__ void matTvec(T* o, const T* R, const T* v) {
  o[0] = R[0] * v[0] + R[3] * v[1] + R[6] * v[2];
  o[1] = R[1] * v[0] + R[4] * v[1] + R[7] * v[2];
  o[2] = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
}
template <typename T> __device__ __forceinline__ void matmul3(T* o, const T* A, const T* B) {
  T r[9];
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) r[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
#pragma unroll
  for (int i = 0; i < 9; i++) o[i] = r[i];
}
template <typename T> __device__ __forceinline__ void matTmul3(T* o, const T* A, const T* B) {  // A^T B
#pragma unroll
  for (int i = 0; i < 3; i++)
#pragma unroll
    for (int j = 0; j < 3; j++) o[3 * i + j] = A[i] * B[j] + A[3 + i] * B[3 + j] + A[6 + i] * B[6 + j];
}
__device__ __forceinline__ float tsqrt(float x) { return sqrtf(x); }
__device__ __forceinline__ double tsqrt(double x) { return sqrt(x); }
__device__ __forceinline__ float tabs(float x) { return fabsf(x); }
__device__ __forceinline__ double tabs(double x) { return fabs(x); }
template <typename T> __device__ __forceinline__ T clamp01(T t) { return t < T(0) ? T(0) : (t > T(1) ? T(1) : t); }

// ---------------------------------------------------------------------------------------------
// a1: forward kinematics, FP64 (RobotObstacleFCL.cpp:287-386).  One call per configuration, so its
// cost is amortised over the frames; doing it in FP64 removes FK error from the FP32 band budget.
// out: per link R[9], t[3].
// ---------------------------------------------------------------------------------------------
struct LinkXfD { double R[9], t[3]; };

__device__ inline void fk_links(const DRobot& rb, const double* q, LinkXfD* out) {
  double t[3] = {rb.base_t[0], rb.base_t[1], rb.base_t[2]}, R[9];
#pragma unroll
  for (int i = 0; i < 9; i++) R[i] = rb.base_R[i];
  int l = 0;
  for (int k = 0; k <= rb.n_joints; k++) {
    while (l < rb.n_links && rb.link_jb[l] == k) {
      double w[3];
      matvec(w, R, rb.lo_t[l]);  // :335
      out[l].t[0] = t[0] + w[0]; out[l].t[1] = t[1] + w[1]; out[l].t[2] = t[2] + w[2];  // :336
      matmul3(out[l].R, R, rb.lo_R[l]);  // :347
      l++;
    }
    if (k == rb.n_joints) break;
    double w[3];
    matvec(w, R, rb.jt[k]);  // :319,361
    t[0] += w[0]; t[1] += w[1]; t[2] += w[2];
    matmul3(R, R, rb.jR[k]);  // :320,362
    double sn, cs;
    sincos(q[k], &sn, &cs);
    // Eigen::AngleAxisd(a, UnitZ()).toRotationMatrix() (:321,363)
    double Rz[9] = {cs, -sn, 0, sn, cs, 0, 0, 0, (1.0 - cs) * 1.0 + cs};
    matmul3(R, R, Rz);
  }
}

// Warp form of the same chain for the latency kernels (one warp per configuration): a PARALLEL PREFIX over the joints.
// Measured on B200 (scripts/micro/gjk_hist, clock64): the per-lane sequential chain costs a warp ~7 100 cycles for six
// joints (~600 dependent FP64 operations at ~6 cycles each, every lane repeating all of them).  Here lane e builds chain
// element e on its own -- element 0 = the base transform, element k + 1 = joint k's local transform (jt_k, jR_k Rz(q_k))
// with ITS sincos -- and an inclusive scan (log2 steps, one affine composition = 36 multiply-adds per step and lane)
// leaves "base and the first e joints applied" in lane e; lane l then takes the state at link_jb[l] by shuffle and applies
// the link's own offset.  ~170 FP64 operations per lane instead of ~600.  The products associate differently from
// fk_links, so the transforms agree with it to rounding (1e-16 relative), not bit for bit: they feed the FP32 link table
// (narrowed, consumers decide outside a 2e-5 m band only) and the FP64 enclosure certificate (1e-9 m margin); every
// FP64 DECISION of the re-check recomputes its link with fk_one_link.
// my_q = angle of joint (lane - 1) (0 for lanes without one); every lane of the warp must call.
__device__ inline void fk_link_of_lane(const DRobot& rb, double my_q, int lane, LinkXfD& mine, long long* t_sincos = nullptr) {
  double sn, cs;
  sincos(my_q, &sn, &cs);
  if (t_sincos) *t_sincos = clock64() + (long long)(sn == 2.0);  // diagnostics builds: the time stamp waits for the result
  double R[9], t[3];
  if (lane >= 1 && lane <= rb.n_joints) {
    const int k = lane - 1;
    const double z = (1.0 - cs) * 1.0 + cs;  // Eigen::AngleAxisd(q, UnitZ).toRotationMatrix()(2, 2)  (:321,363)
#pragma unroll
    for (int i = 0; i < 3; i++) {
      const double a = rb.jR[k][3 * i], b = rb.jR[k][3 * i + 1], c = rb.jR[k][3 * i + 2];
      R[3 * i] = a * cs + b * sn; R[3 * i + 1] = b * cs - a * sn; R[3 * i + 2] = c * z;
      t[i] = rb.jt[k][i];
    }
  } else {
#pragma unroll
    for (int i = 0; i < 9; i++) R[i] = lane == 0 ? rb.base_R[i] : ((i & 3) == 0 ? 1.0 : 0.0);
#pragma unroll
    for (int i = 0; i < 3; i++) t[i] = lane == 0 ? rb.base_t[i] : 0.0;
  }
  for (int d = 1; d <= rb.n_joints; d <<= 1) {  // inclusive scan of X_e = X_(e-d) o X_e
    double Rp[9], tp[3];
#pragma unroll
    for (int i = 0; i < 9; i++) Rp[i] = __shfl_up_sync(SIRRT_FULL, R[i], d);
#pragma unroll
    for (int i = 0; i < 3; i++) tp[i] = __shfl_up_sync(SIRRT_FULL, t[i], d);
    if (lane >= d) {
      double w[3];
      matvec(w, Rp, t);
      t[0] = tp[0] + w[0]; t[1] = tp[1] + w[1]; t[2] = tp[2] + w[2];
      matmul3(R, Rp, R);
    }
  }
  const int l = lane < rb.n_links ? lane : 0;
  const int src = rb.link_jb[l];
  double Rs[9], ts[3];
#pragma unroll
  for (int i = 0; i < 9; i++) Rs[i] = __shfl_sync(SIRRT_FULL, R[i], src);
#pragma unroll
  for (int i = 0; i < 3; i++) ts[i] = __shfl_sync(SIRRT_FULL, t[i], src);
  double w[3];
  matvec(w, Rs, rb.lo_t[l]);  // :335
  mine.t[0] = ts[0] + w[0]; mine.t[1] = ts[1] + w[1]; mine.t[2] = ts[2] + w[2];  // :336
  matmul3(mine.R, Rs, rb.lo_R[l]);  // :347
}

#ifdef SIRRT_GJK_STATS
// diagnostics builds: the per-lane chain with the sincos spread over the lanes (what fk_link_of_lane replaced), for A/B timing
__device__ int g_fk_mode;
__device__ inline void fk_link_of_lane_serial(const DRobot& rb, double my_q, int lane, LinkXfD& mine, long long* t_sincos) {
  double my_sn, my_cs;
  sincos(my_q, &my_sn, &my_cs);
  if (t_sincos) *t_sincos = clock64() + (long long)(my_sn == 2.0);
  double t[3] = {rb.base_t[0], rb.base_t[1], rb.base_t[2]}, R[9];
#pragma unroll
  for (int i = 0; i < 9; i++) R[i] = rb.base_R[i];
  int l = 0;
  for (int k = 0; k <= rb.n_joints; k++) {
    while (l < rb.n_links && rb.link_jb[l] == k) {
      if (l == lane) {
        double w[3];
        matvec(w, R, rb.lo_t[l]);
        mine.t[0] = t[0] + w[0]; mine.t[1] = t[1] + w[1]; mine.t[2] = t[2] + w[2];
        matmul3(mine.R, R, rb.lo_R[l]);
      }
      l++;
    }
    if (k == rb.n_joints) break;
    double w[3];
    matvec(w, R, rb.jt[k]);
    t[0] += w[0]; t[1] += w[1]; t[2] += w[2];
    matmul3(R, R, rb.jR[k]);
    const double sn = __shfl_sync(SIRRT_FULL, my_sn, k), cs = __shfl_sync(SIRRT_FULL, my_cs, k);
    double Rz[9] = {cs, -sn, 0, sn, cs, 0, 0, 0, (1.0 - cs) * 1.0 + cs};
    matmul3(R, R, Rz);
  }
}
#endif

// ---------------------------------------------------------------------------------------------
// closed-form core distances (same formulas as the test-only CPU checker, written independently of
// precision)
// ---------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ T pt_seg_dist2(const T* p, const T* a, const T* b) {
  T ab[3], ap[3];
  sub3(ab, b, a); sub3(ap, p, a);
  T den = dot3(ab, ab), t = T(0);
  if (den > T(0)) t = clamp01(dot3(ap, ab) / den);
  T c[3] = {ap[0] - t * ab[0], ap[1] - t * ab[1], ap[2] - t * ab[2]};
  return dot3(c, c);
}

template <typename T> __device__ inline T seg_seg_dist2(const T* p1, const T* q1, const T* p2, const T* q2) {
  T d1[3], d2[3], r[3];
  sub3(d1, q1, p1); sub3(d2, q2, p2); sub3(r, p1, p2);
  T a = dot3(d1, d1), e = dot3(d2, d2), f = dot3(d2, r), s, t;
  const T EPS = T(1e-30);
  if (a <= EPS && e <= EPS) return dot3(r, r);
  if (a <= EPS) { s = T(0); t = clamp01(f / e); }
  else {
    T c = dot3(d1, r);
    if (e <= EPS) { t = T(0); s = clamp01(-c / a); }
    else {
      T b = dot3(d1, d2), denom = a * e - b * b;
      s = denom > T(0) ? clamp01((b * f - c * e) / denom) : T(0);
      t = (b * s + f) / e;
      if (t < T(0)) { t = T(0); s = clamp01(-c / a); }
      else if (t > T(1)) { t = T(1); s = clamp01((b - c) / a); }
    }
  }
  T c1[3] = {p1[0] + d1[0] * s, p1[1] + d1[1] * s, p1[2] + d1[2] * s};
  T c2[3] = {p2[0] + d2[0] * t, p2[1] + d2[1] * t, p2[2] + d2[2] * t};
  T d[3]; sub3(d, c1, c2);
  return dot3(d, d);
}

// parameters (s on p1q1, t on p2q2) of the closest points of two segments (same case analysis as seg_seg_dist2)
template <typename T> __device__ inline void seg_seg_params(const T* p1, const T* q1, const T* p2, const T* q2, T& s, T& t) {
  T d1[3], d2[3], r[3];
  sub3(d1, q1, p1); sub3(d2, q2, p2); sub3(r, p1, p2);
  T a = dot3(d1, d1), e = dot3(d2, d2), f = dot3(d2, r);
  const T EPS = T(1e-30);
  if (a <= EPS && e <= EPS) { s = t = T(0); return; }
  if (a <= EPS) { s = T(0); t = clamp01(f / e); return; }
  T c = dot3(d1, r);
  if (e <= EPS) { t = T(0); s = clamp01(-c / a); return; }
  T b = dot3(d1, d2), denom = a * e - b * b;
  s = denom > T(0) ? clamp01((b * f - c * e) / denom) : T(0);
  t = (b * s + f) / e;
  if (t < T(0)) { t = T(0); s = clamp01(-c / a); }
  else if (t > T(1)) { t = T(1); s = clamp01((b - c) / a); }
}

// segment a + t*d (t in [0,1]) vs box [-h,h]: g(t) = sum max(|p_i|-h_i,0)^2 is convex, g' is
// continuous piecewise linear with breakpoints at the slab crossings -> bracket and solve.
template <typename T> __device__ __forceinline__ T sb_g(const T* a, const T* d, const T* h, T t, T& gp) {
  T g = T(0), s = T(0);
#pragma unroll
  for (int i = 0; i < 3; i++) {
    T p = a[i] + t * d[i], e = tabs(p) - h[i];
    if (e > T(0)) { g += e * e; s += (p > T(0) ? e : -e) * d[i]; }
  }
  gp = T(2) * s;
  return g;
}
template <typename T> __device__ inline T seg_box_dist2(const T* a, const T* b, const T* h, T* t_out = nullptr) {
  T d[3]; sub3(d, b, a);
  T gp0, gp1, g0 = sb_g(a, d, h, T(0), gp0), g1 = sb_g(a, d, h, T(1), gp1);
  if (gp0 >= T(0)) { if (t_out) *t_out = T(0); return g0; }
  if (gp1 <= T(0)) { if (t_out) *t_out = T(1); return g1; }
  T lo = T(0), hi = T(1), glo = gp0, ghi = gp1;
#pragma unroll
  for (int i = 0; i < 3; i++) {
    if (d[i] == T(0)) continue;
#pragma unroll
    for (int sg = 0; sg < 2; sg++) {
      T t = ((sg ? h[i] : -h[i]) - a[i]) / d[i];
      if (!(t > lo && t < hi)) continue;
      T gp; sb_g(a, d, h, t, gp);
      if (gp <= T(0)) { lo = t; glo = gp; } else { hi = t; ghi = gp; }
    }
  }
  T t = lo;
  if (ghi > glo) t = lo - glo * (hi - lo) / (ghi - glo);
  t = t < lo ? lo : (t > hi ? hi : t);
  if (t_out) *t_out = t;
  T gp;
  return sb_g(a, d, h, t, gp);
}

// ---------------------------------------------------------------------------------------------
// warp-cooperative GJK between two convex cores, evaluated in the local frame of A.
//   A: hull (vertex scan striped over the 32 lanes)        [planned-robot link]
//   B: hull / box / segment / point, posed by (Rrel, trel) relative to A
// All 32 lanes hold identical simplex state; only the support scans are distributed.
// ---------------------------------------------------------------------------------------------
template <typename T> struct GShape {
  int type;        // SH_*
  const T* verts;  // hull: xyz interleaved
  const unsigned int* cells;     // hull: support-candidate cell offsets (null: scan every vertex)
  const unsigned short* cand;
  const float4* rec;             // hull, FP32 only: inline candidate records of the same cells (null: use cells / cand)
  int n;
  T half[3];       // box
  T half_len;      // capsule core
};

// FP32 hull support through the inline cell record: one load per lane (slot `lane` of the cell's record), the winner's
// coordinates come back by shuffle -- no dependent loads.  Same winner as the candidate-list scan (largest dot product,
// lowest vertex index on ties: slots hold the candidates in ascending index order).  Returns false when the cell overflows
// its record (the caller scans the candidate list instead).
__device__ __forceinline__ bool hull_rec_support(const GShape<double>&, int, const double*, double*, int, int*) { return false; }
__device__ __forceinline__ bool hull_rec_support(const GShape<float>& s, int cell, const float* d, float* out, int lane, int* id) {
  if (!s.rec) return false;
  float4 r = make_float4(0.f, 0.f, 0.f, __int_as_float(-1));
  if (lane < SIRRT_HULL_REC) r = __ldg(s.rec + (size_t)cell * SIRRT_HULL_REC + lane);
  const int i = __float_as_int(r.w);
  if (__any_sync(SIRRT_FULL, i == -2)) return false;
  float bv = -3.0e38f;
  int bl = 32;
  if (i >= 0) { bv = r.x * d[0] + r.y * d[1] + r.z * d[2]; bl = lane; }
#pragma unroll
  for (int o = SIRRT_HULL_REC / 2; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(SIRRT_FULL, bv, o);
    const int ol = __shfl_xor_sync(SIRRT_FULL, bl, o);
    if (ov > bv || (ov == bv && ol < bl)) { bv = ov; bl = ol; }
  }
  bl = __shfl_sync(SIRRT_FULL, bl, 0);  // lanes >= SIRRT_HULL_REC reduced among themselves: take the record lanes' answer
  if (bl >= 32) return false;           // empty cell (cannot happen for a built table): full scan
  out[0] = __shfl_sync(SIRRT_FULL, r.x, bl); out[1] = __shfl_sync(SIRRT_FULL, r.y, bl); out[2] = __shfl_sync(SIRRT_FULL, r.z, bl);
  *id = __shfl_sync(SIRRT_FULL, i, bl);
  return true;
}

template <typename T>
__device__ inline int warp_support(const GShape<T>& s, const T* d, T* out, int lane) {
  if (s.type == SH_HULL) {
    T bv = T(-3.0e38);
    int bi = 0x7fffffff;
    const int cell = s.cells ? hull_cell(d) : -1;
    if (cell >= 0 && hull_rec_support(s, cell, d, out, lane, &bi)) return bi;
    if (cell >= 0) {
      // every vertex that can win the full scan for a direction of this cell is listed (ascending index), so the
      // winner -- largest dot product, lowest index on ties -- is the one the full scan finds, bit for bit
      const unsigned c0 = s.cells[cell], c1 = s.cells[cell + 1];
      for (unsigned k = c0 + lane; k < c1; k += 32) {
        const int i = s.cand[k];
        T v = s.verts[3 * i] * d[0] + s.verts[3 * i + 1] * d[1] + s.verts[3 * i + 2] * d[2];
        if (v > bv) { bv = v; bi = i; }
      }
    } else {
      for (int i = lane; i < s.n; i += 32) {
        T v = s.verts[3 * i] * d[0] + s.verts[3 * i + 1] * d[1] + s.verts[3 * i + 2] * d[2];
        if (v > bv) { bv = v; bi = i; }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      T ov = __shfl_xor_sync(SIRRT_FULL, bv, o);
      int oi = __shfl_xor_sync(SIRRT_FULL, bi, o);
      if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
    }
    if (bi == 0x7fffffff) bi = 0;
    out[0] = s.verts[3 * bi]; out[1] = s.verts[3 * bi + 1]; out[2] = s.verts[3 * bi + 2];
    return bi;
  }
  if (s.type == SH_BOX) {
    int id = 0;
#pragma unroll
    for (int k = 0; k < 3; k++) {
      bool pos = d[k] >= T(0);
      out[k] = pos ? s.half[k] : -s.half[k];
      id |= (pos ? 1 : 0) << k;
    }
    return id;
  }
  if (s.type == SH_CAPSULE) {
    bool pos = d[2] >= T(0);
    out[0] = T(0); out[1] = T(0); out[2] = pos ? s.half_len : -s.half_len;
    return pos ? 1 : 0;
  }
  out[0] = out[1] = out[2] = T(0);
  return 0;
}

template <typename T> struct Simplex { T w[4][3]; int id[4]; int n; };
template <typename T> __host__ __device__ __forceinline__ void gjk_point(const GShape<T>& s, int id, T* out);

template <typename T> struct GjkEps;
template <> struct GjkEps<float> { static __device__ __forceinline__ float tri() { return 1e-10f; } static __device__ __forceinline__ float tet() { return 1e-4f; } };
template <> struct GjkEps<double> { static __device__ __forceinline__ double tri() { return 1e-24; } static __device__ __forceinline__ double tet() { return 1e-9; } };

// closest point to the origin on segment ab; returns kept-vertex mask (a=1,b=2)
template <typename T> __device__ __forceinline__ int closest_seg(const T* a, const T* b, T* out) {
  T ab[3]; sub3(ab, b, a);
  const T den = dot3(ab, ab), t = den > T(0) ? -dot3(a, ab) / den : T(0);
  if (t <= T(0)) { out[0] = a[0]; out[1] = a[1]; out[2] = a[2]; return 1; }
  if (t >= T(1)) { out[0] = b[0]; out[1] = b[1]; out[2] = b[2]; return 2; }
  out[0] = a[0] + t * ab[0]; out[1] = a[1] + t * ab[1]; out[2] = a[2] + t * ab[2];
  return 3;
}

// closest point to the origin on triangle abc; returns kept-vertex mask (a=1,b=2,c=4).  Polytopes
// hand GJK many collinear / coplanar support points: a triangle whose area is negligible against its
// edge lengths is solved as the best of its three edges.
template <typename T> __device__ inline int closest_tri(const T* a, const T* b, const T* c, T* out) {
  T ab[3], ac[3], bc[3], n[3];
  sub3(ab, b, a); sub3(ac, c, a); sub3(bc, c, b); cross3(n, ab, ac);
  const T ab2 = dot3(ab, ab), ac2 = dot3(ac, ac), bc2 = dot3(bc, bc);
  const T l2 = ab2 > ac2 ? (ab2 > bc2 ? ab2 : bc2) : (ac2 > bc2 ? ac2 : bc2);
  if (dot3(n, n) <= GjkEps<T>::tri() * l2 * l2) {
    T q[3];
    int m = closest_seg(a, b, out), keep = m;
    T best = dot3(out, out);
    m = closest_seg(a, c, q);
    if (dot3(q, q) < best) { best = dot3(q, q); out[0] = q[0]; out[1] = q[1]; out[2] = q[2]; keep = (m & 1) | ((m & 2) << 1); }
    m = closest_seg(b, c, q);
    if (dot3(q, q) < best) { out[0] = q[0]; out[1] = q[1]; out[2] = q[2]; keep = m << 1; }
    return keep;
  }
  T d1 = -dot3(ab, a), d2 = -dot3(ac, a);
  if (d1 <= T(0) && d2 <= T(0)) { out[0] = a[0]; out[1] = a[1]; out[2] = a[2]; return 1; }
  T d3 = -dot3(ab, b), d4 = -dot3(ac, b);
  if (d3 >= T(0) && d4 <= d3) { out[0] = b[0]; out[1] = b[1]; out[2] = b[2]; return 2; }
  T vc = d1 * d4 - d3 * d2;
  if (vc <= T(0) && d1 >= T(0) && d3 <= T(0)) {
    T v = d1 / (d1 - d3);
    for (int i = 0; i < 3; i++) out[i] = a[i] + v * ab[i];
    return 3;
  }
  T d5 = -dot3(ab, c), d6 = -dot3(ac, c);
  if (d6 >= T(0) && d5 <= d6) { out[0] = c[0]; out[1] = c[1]; out[2] = c[2]; return 4; }
  T vb = d5 * d2 - d1 * d6;
  if (vb <= T(0) && d2 >= T(0) && d6 <= T(0)) {
    T w = d2 / (d2 - d6);
    for (int i = 0; i < 3; i++) out[i] = a[i] + w * ac[i];
    return 5;
  }
  T va = d3 * d6 - d5 * d4;
  if (va <= T(0) && (d4 - d3) >= T(0) && (d5 - d6) >= T(0)) {
    T w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    for (int i = 0; i < 3; i++) out[i] = b[i] + w * (c[i] - b[i]);
    return 6;
  }
  T den = T(1) / (va + vb + vc), v = vb * den, w = vc * den;
  for (int i = 0; i < 3; i++) out[i] = a[i] + ab[i] * v + ac[i] * w;
  return 7;
}

// reduce simplex to the sub-simplex closest to the origin; returns 1 when the origin is enclosed.
// A tetrahedron only counts as enclosing when it has real volume (coplanar support points are the
// norm between polytopes); otherwise, and whenever the origin is outside, the answer is the best of
// the four faces.  min_face (enclosing tetrahedron only) is a lower bound of the penetration depth.
// Warp form: all 32 lanes hold the same simplex and must call together; for a tetrahedron lane (lane & 3) evaluates
// face (lane & 3) -- the side test and the closest point of ONE face instead of four per lane -- and the four answers
// are combined exactly as the sequential loop over the faces would (smallest distance, lowest face on ties).
template <typename T> __device__ __forceinline__ void simplex_pick(const Simplex<T>& s, int i, T* o) {
#pragma unroll
  for (int k = 0; k < 3; k++) o[k] = i == 0 ? s.w[0][k] : (i == 1 ? s.w[1][k] : (i == 2 ? s.w[2][k] : s.w[3][k]));
}
template <typename T> __device__ inline int simplex_closest(Simplex<T>& s, T* v, T& min_face, int lane) {
  if (s.n == 1) { v[0] = s.w[0][0]; v[1] = s.w[0][1]; v[2] = s.w[0][2]; return 0; }
  int keep = 0;
  if (s.n == 2) keep = closest_seg(s.w[0], s.w[1], v);
  else if (s.n == 3) keep = closest_tri(s.w[0], s.w[1], s.w[2], v);
  else {
    // faces F[f] = {0,1,2|3}, {0,2,3|1}, {0,3,1|2}, {1,3,2|0}: three vertices of the face, then the opposite one
    const int f = lane & 3;
    const int i0 = f == 3 ? 1 : 0, i1 = f == 0 ? 1 : (f == 1 ? 2 : 3), i2 = f == 0 ? 2 : (f == 1 ? 3 : (f == 2 ? 1 : 2)), i3 = f == 0 ? 3 : (f == 1 ? 1 : (f == 2 ? 2 : 0));
    T a[3], b[3], c[3], d[3];
    simplex_pick(s, i0, a); simplex_pick(s, i1, b); simplex_pick(s, i2, c); simplex_pick(s, i3, d);
    {
      const T *A = s.w[0], *B = s.w[1], *C = s.w[2], *D = s.w[3];
      T ab[3], ac[3], ad[3], bc[3], bd[3], cd[3], n[3];
      sub3(ab, B, A); sub3(ac, C, A); sub3(ad, D, A); sub3(bc, C, B); sub3(bd, D, B); sub3(cd, D, C);
      cross3(n, ab, ac);
      const T vol6 = dot3(n, ad);
      T l2 = dot3(ab, ab);
      T e = dot3(ac, ac); l2 = e > l2 ? e : l2;
      e = dot3(ad, ad); l2 = e > l2 ? e : l2;
      e = dot3(bc, bc); l2 = e > l2 ? e : l2;
      e = dot3(bd, bd); l2 = e > l2 ? e : l2;
      e = dot3(cd, cd); l2 = e > l2 ? e : l2;
      if (tabs(vol6) > GjkEps<T>::tet() * l2 * tsqrt(l2)) {
        T e1[3], e2[3], fn[3], ed[3];
        sub3(e1, b, a); sub3(e2, c, a); cross3(fn, e1, e2); sub3(ed, d, a);
        const T sp = -dot3(a, fn), sd = dot3(ed, fn);
        const bool inside = __all_sync(SIRRT_FULL, sp * sd > T(0));  // origin strictly on the opposite vertex's side of every face
        const T nn = dot3(fn, fn);
        T mf = nn > T(0) ? tabs(sp) / tsqrt(nn) : T(0);
#pragma unroll
        for (int o = 1; o < 4; o <<= 1) { const T of = __shfl_xor_sync(SIRRT_FULL, mf, o); mf = of < mf ? of : mf; }
        if (inside) { v[0] = v[1] = v[2] = T(0); min_face = mf; return 1; }
      }
    }
    T q[3];
    int m = closest_tri(a, b, c, q);
    const T dd = dot3(q, q);
    T bd = dd < T(3.0e38) ? dd : T(3.0e38);
    int bf = dd < T(3.0e38) ? f : 4;
#pragma unroll
    for (int o = 1; o < 4; o <<= 1) {
      const T od = __shfl_xor_sync(SIRRT_FULL, bd, o);
      const int of = __shfl_xor_sync(SIRRT_FULL, bf, o);
      if (od < bd || (od == bd && of < bf)) { bd = od; bf = of; }
    }
    if (bf < 4) {
      const int src = (lane & ~3) | bf;
      v[0] = __shfl_sync(SIRRT_FULL, q[0], src); v[1] = __shfl_sync(SIRRT_FULL, q[1], src); v[2] = __shfl_sync(SIRRT_FULL, q[2], src);
      m = __shfl_sync(SIRRT_FULL, m, src);
      const int j0 = bf == 3 ? 1 : 0, j1 = bf == 0 ? 1 : (bf == 1 ? 2 : 3), j2 = bf == 0 ? 2 : (bf == 1 ? 3 : (bf == 2 ? 1 : 2));
      keep = ((m & 1) ? 1 << j0 : 0) | ((m & 2) ? 1 << j1 : 0) | ((m & 4) ? 1 << j2 : 0);
    }
  }
  // (the simplex is indexed dynamically here and in warp_gjk and so lives in local memory; a register-resident variant
  // -- static indices, selects -- was measured on B200: same GJK cycles per task, 124 -> 166 registers and -45 % on the
  // hull-batch throughput of k_collide_bits, so it was not kept)
  int j = 0;
  for (int i = 0; i < s.n; i++)
    if (keep & (1 << i)) {
      if (i != j) { s.w[j][0] = s.w[i][0]; s.w[j][1] = s.w[i][1]; s.w[j][2] = s.w[i][2]; s.id[j] = s.id[i]; }
      j++;
    }
  s.n = j;
  return 0;
}

// FINAL = false (FP32 fast path): classify against [rsum-band, rsum+band] using the rigorous GJK
//         bounds  v.w/|v| <= d_core <= |v|  -> RES_FREE / RES_HIT / RES_AMBIG.
// FINAL = true  (FP64 re-check): converge and apply coal's predicate -> RES_FREE / RES_HIT;
//         *dist_out gets the signed distance (core distance - rsum).
template <typename T, bool FINAL>
__device__ inline int warp_gjk(const GShape<T>& A, const GShape<T>& B, const T* Rrel, const T* trel, T rsum, T band,
                               int lane, T* dist_out, int* enc_ids = nullptr, Simplex<T>* final_simplex = nullptr, T* final_v = nullptr,
                               int* n_iter = nullptr, int* last_ids = nullptr, const int* seed_ids = nullptr) {
  // enc_ids (FP32 pass only): when the answer is AMBIGUOUS because the enclosing tetrahedron is too thin to trust in
  // FP32, the four support-point ids -- the caller can certify the enclosure in FP64 from them (gjk_point / tetra_encloses)
  // last_ids (FP32 pass): the support-point ids of the simplex an AMBIGUOUS verdict ended on (-1 = unused slot); seed_ids
  // (FP64 re-check): start from that simplex instead of from scratch -- the same fixed point in two or three iterations
  // instead of a dozen (a warp-wide FP64 instruction costs a lone warp 6-16 cycles on B200: 40 us per cold re-check)
  if (enc_ids) enc_ids[0] = -1;
  Simplex<T> s;
  T v[3] = {-trel[0], -trel[1], -trel[2]};  // A is centred near its origin, B near trel
  if (dot3(v, v) == T(0)) v[0] = T(1);
  T pa[3], pb[3], dl[3], nv[3], min_face = T(0);
  auto note_last = [&]() {
    if (last_ids) { for (int i = 0; i < 4; i++) last_ids[i] = i < s.n ? s.id[i] : -1; }
  };
  const T REL = FINAL ? T(1e-15) : T(1e-6);
  auto support = [&](const T* dir, T* w) -> int {  // support of A - B in direction dir
#ifdef SIRRT_GJK_STATS
    const long long c0 = clock64() + (long long)(dir[0] == T(3e38));
#endif
    int ia = warp_support(A, dir, pa, lane);
#ifdef SIRRT_GJK_STATS
    const long long c1 = clock64() + (long long)(pa[0] == T(3e38));
#endif
    T nd[3] = {-dir[0], -dir[1], -dir[2]};
    matTvec(dl, Rrel, nd);
    T u[3];
    int ib = warp_support(B, dl, u, lane);
    matvec(pb, Rrel, u);
    w[0] = pa[0] - (pb[0] + trel[0]); w[1] = pa[1] - (pb[1] + trel[1]); w[2] = pa[2] - (pb[2] + trel[2]);
#ifdef SIRRT_GJK_STATS
    if (!FINAL) { const long long c2 = clock64() + (long long)(w[0] == T(3e38)); PHASE_ADD(13, c1 - c0); PHASE_ADD(14, c2 - c1); }
#endif
    return (ia << 12) | ib;
  };
  bool seeded = false;
  if (FINAL && seed_ids && seed_ids[0] >= 0) {
    s.n = 0;
    for (int i = 0; i < 4; i++)
      if (seed_ids[i] >= 0) {
        T xa[3], xb[3], w[3];
        gjk_point(A, seed_ids[i] >> 12, xa);
        gjk_point(B, seed_ids[i] & 4095, xb);
        matvec(w, Rrel, xb);
        s.w[s.n][0] = xa[0] - (w[0] + trel[0]); s.w[s.n][1] = xa[1] - (w[1] + trel[1]); s.w[s.n][2] = xa[2] - (w[2] + trel[2]);
        s.id[s.n] = seed_ids[i];
        s.n++;
      }
    T sv[3];
    if (simplex_closest(s, sv, min_face, lane)) {  // the seed already encloses the origin: the cores intersect
      if (dist_out) *dist_out = -rsum;
      return RES_HIT;
    }
    v[0] = sv[0]; v[1] = sv[1]; v[2] = sv[2];
    seeded = true;
  }
  if (!seeded) {
    nv[0] = -v[0]; nv[1] = -v[1]; nv[2] = -v[2];
    s.id[0] = support(nv, s.w[0]);
    s.n = 1;
    v[0] = s.w[0][0]; v[1] = s.w[0][1]; v[2] = s.w[0][2];
  }
  T vv = dot3(v, v);
  const int MAXIT = FINAL ? 256 : 48;
  for (int it = 0; it < MAXIT; it++) {
    if (n_iter) *n_iter = it + 1;  // diagnostics builds only (SIRRT_GJK_STATS)
    if (vv <= (FINAL ? T(1e-30) : T(1e-20))) {  // cores touch / intersect at a vertex
      if (dist_out) *dist_out = -rsum;
      if (FINAL) return RES_HIT;
      if (rsum > band) return RES_HIT;
      note_last();
      return RES_AMBIG;
    }
    T w[3];
    nv[0] = -v[0]; nv[1] = -v[1]; nv[2] = -v[2];
    int id = support(nv, w);
    T vw = dot3(v, w);
    if (!FINAL) {
      // separating axis: every point of A-B has v.x >= v.w, so d_core >= v.w/|v|
      if (vw > T(0)) {
        T lower = vw * rsqrtf((float)vv);
        if (lower - rsum > band) return RES_FREE;
      }
    }
    T gap = vv - vw;
    bool done = gap <= REL * vv;
    for (int i = 0; i < s.n; i++) done |= (s.id[i] == id);
    if (done) break;
    s.w[s.n][0] = w[0]; s.w[s.n][1] = w[1]; s.w[s.n][2] = w[2]; s.id[s.n] = id; s.n++;
    T nvv[3];
#ifdef SIRRT_GJK_STATS
    const long long cs0 = clock64() + (long long)(w[0] == T(3e38));
    const int enclosed = simplex_closest(s, nvv, min_face, lane);
    if (!FINAL) PHASE_ADD(15, clock64() + (long long)(nvv[0] == T(3e38)) - cs0);
    if (enclosed) {
#else
    if (simplex_closest(s, nvv, min_face, lane)) {
#endif
      if (dist_out) *dist_out = -rsum;
      if (FINAL) return RES_HIT;  // core distance 0 <= threshold + rsum
      if (rsum > band) return RES_HIT;
      if (min_face > band) return RES_HIT;
      if (enc_ids && s.n == 4) { enc_ids[0] = s.id[0]; enc_ids[1] = s.id[1]; enc_ids[2] = s.id[2]; enc_ids[3] = s.id[3]; }
      note_last();
      return RES_AMBIG;
    }
    T n2 = dot3(nvv, nvv);
    if (!(n2 < vv)) break;
    v[0] = nvv[0]; v[1] = nvv[1]; v[2] = nvv[2]; vv = n2;
  }
  T d = tsqrt(vv) - rsum;
  if (dist_out) *dist_out = d;
  if (final_simplex) { *final_simplex = s; final_v[0] = v[0]; final_v[1] = v[1]; final_v[2] = v[2]; }  // witness points (get_distances)
  if (FINAL) return ((double)d - SIRRT_SECURITY_MARGIN <= SIRRT_COLLISION_THRESHOLD) ? RES_HIT : RES_FREE;
  if (d < -band) return RES_HIT;
  note_last();
  return RES_AMBIG;  // |d| <= band, or not converged enough to certify separation
}

// Point of shape s behind a support id returned by warp_support (hull vertex index, box corner bits, capsule end).
template <typename T> __host__ __device__ __forceinline__ void gjk_point(const GShape<T>& s, int id, T* out) {
  if (s.type == SH_HULL) { out[0] = s.verts[3 * id]; out[1] = s.verts[3 * id + 1]; out[2] = s.verts[3 * id + 2]; }
  else if (s.type == SH_BOX) { for (int k = 0; k < 3; k++) out[k] = (id >> k) & 1 ? s.half[k] : -s.half[k]; }
  else if (s.type == SH_CAPSULE) { out[0] = out[1] = T(0); out[2] = id ? s.half_len : -s.half_len; }
  else { out[0] = out[1] = out[2] = T(0); }
}

// FP64 certificate for "the cores intersect": the four points of A - B named by the ids of an FP32 GJK simplex form a
// tetrahedron with real volume that has the origin strictly inside, at least `margin` from every face.  Four points of
// the Minkowski difference enclosing the origin prove the intersection whatever GJK run found them.
__host__ __device__ inline bool tetra_encloses(const GShape<double>& A, const GShape<double>& B, const double* Rrel, const double* trel,
                                      const int* ids, double margin, double* min_face) {
  double w[4][3];
  for (int i = 0; i < 4; i++) {
    double pa[3], u[3], pb[3];
    gjk_point(A, ids[i] >> 12, pa);
    gjk_point(B, ids[i] & 4095, u);
    matvec(pb, Rrel, u);
    for (int k = 0; k < 3; k++) w[i][k] = pa[k] - (pb[k] + trel[k]);
  }
  const int F[4][4] = {{0, 1, 2, 3}, {0, 2, 3, 1}, {0, 3, 1, 2}, {1, 3, 2, 0}};
  double mf = 3.0e38;
  for (int f = 0; f < 4; f++) {
    const double *a = w[F[f][0]], *b = w[F[f][1]], *c = w[F[f][2]], *d = w[F[f][3]];
    double e1[3], e2[3], fn[3], ed[3];
    sub3(e1, b, a); sub3(e2, c, a); cross3(fn, e1, e2); sub3(ed, d, a);
    const double nn = dot3(fn, fn);
    if (!(nn > 0.0)) return false;
    const double inv = 1.0 / sqrt(nn);
    const double so = -dot3(a, fn) * inv, sd = dot3(ed, fn) * inv;  // signed distances of the origin and of d from the face plane
    if (!(so * sd > 0.0)) return false;                               // origin on d's side
    if (!(fabs(sd) > 1e-7)) return false;                             // real volume: the apex is clear of the face plane
    mf = fabs(so) < mf ? fabs(so) : mf;
  }
  *min_face = mf;
  return mf > margin;
}
