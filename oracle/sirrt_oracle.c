/*
 * sirrt_oracle.c -- TEST INFRASTRUCTURE ONLY (see sirrt_oracle.h).  PARITY UNPINNED against the reference (see header);
 * cross-checked cell by cell against the independent oracle/oracle2.py (tests/test_oracle2.py).
 *
 * FP64 CPU restatement of the reference's hot path.  Each function cites the reference lines it
 * follows (paths relative to the reference root).  The collision arithmetic itself lives in the
 * un-vendored coal-library/coal @ ff01445364a7b386e8e5c46ccd530ed9d9a8e8cf (install.sh:15); its
 * published algorithm is restated here: every shape is a convex core polytope (sphere = point,
 * capsule = segment, box = 8 corners, convex mesh = hull vertices) inflated by a radius, the core
 * distance comes from GJK (or an exact closed form where one exists), and a pair collides when
 *      distance - security_margin <= collision_distance_threshold
 * with security_margin = 1e-21 (CollisionManager.cpp:559,611,654) and coal's default threshold
 * Eigen::NumTraits<double>::dummy_precision() = 1e-12.  GJK here is run to ~1e-14 (tighter than
 * coal's 1e-6 default) so that the oracle is the truth inside the 1e-6 tolerance band.
 * kNN: jlblancoc/nanoflann @ ba47cfcb127c3597d69196d87f5aa9ca8811b0a9 (install.sh:18), metric
 * L2_Simple_Adaptor restated as a brute-force scan.
 *
 * Build: see oracle/Makefile (gcc -O3 -march=native -ffp-contract=off -fopenmp).
 */
#include "sirrt_oracle.h"
#include "../include/sirrt_pack.h"
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define SECURITY_MARGIN 1e-21
#define COLLISION_THRESHOLD 1e-12
#define CULL_MARGIN 1e-5

static __thread char g_err[256];
const char* oracle_last_error(void) { return g_err; }
static void set_err(const char* m) { snprintf(g_err, sizeof g_err, "%s", m); }

/* ------------------------------------------------------------------ small linear algebra */
typedef struct { double R[9], t[3]; } Xf;

static inline double dot3(const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }
static inline void sub3(double* o, const double* a, const double* b) { o[0] = a[0] - b[0]; o[1] = a[1] - b[1]; o[2] = a[2] - b[2]; }
static inline void cross3(double* o, const double* a, const double* b) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
static inline void matvec(double* o, const double* R, const double* v) {
  o[0] = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
  o[1] = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
  o[2] = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
}
static inline void matTvec(double* o, const double* R, const double* v) {
  o[0] = R[0] * v[0] + R[3] * v[1] + R[6] * v[2];
  o[1] = R[1] * v[0] + R[4] * v[1] + R[7] * v[2];
  o[2] = R[2] * v[0] + R[5] * v[1] + R[8] * v[2];
}
static inline void matmul(double* o, const double* A, const double* B) {
  double r[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) r[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  memcpy(o, r, sizeof r);
}
static inline void xf_point(double* o, const Xf* x, const double* p) {
  matvec(o, x->R, p);
  o[0] += x->t[0]; o[1] += x->t[1]; o[2] += x->t[2];
}
/* Eigen::QuaternionBase::toRotationMatrix, no normalisation (used for every pose in the
 * reference: CollisionManager.cpp:72,132,233; RobotObstacleFCL.cpp:302) */
static void quat_to_R(double* R, double x, double y, double z, double w) {
  double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  double twx = tx * w, twy = ty * w, twz = tz * w;
  double txx = tx * x, txy = ty * x, txz = tz * x;
  double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

/* ------------------------------------------------------------------ closed-form distances */
double oracle_pt_seg_dist2(const double p[3], const double a[3], const double b[3]) {
  double ab[3], ap[3];
  sub3(ab, b, a); sub3(ap, p, a);
  double den = dot3(ab, ab), t = 0;
  if (den > 0) { t = dot3(ap, ab) / den; t = t < 0 ? 0 : (t > 1 ? 1 : t); }
  double c[3] = {ap[0] - t * ab[0], ap[1] - t * ab[1], ap[2] - t * ab[2]};
  return dot3(c, c);
}

/* closest points of two segments (Ericson, Real-Time Collision Detection 5.1.9) */
double oracle_seg_seg_dist2(const double p1[3], const double q1[3], const double p2[3], const double q2[3]) {
  double d1[3], d2[3], r[3];
  sub3(d1, q1, p1); sub3(d2, q2, p2); sub3(r, p1, p2);
  double a = dot3(d1, d1), e = dot3(d2, d2), f = dot3(d2, r), s, t;
  const double EPS = 1e-300;
  if (a <= EPS && e <= EPS) return dot3(r, r);
  if (a <= EPS) { s = 0; t = f / e; t = t < 0 ? 0 : (t > 1 ? 1 : t); }
  else {
    double c = dot3(d1, r);
    if (e <= EPS) { t = 0; s = -c / a; s = s < 0 ? 0 : (s > 1 ? 1 : s); }
    else {
      double b = dot3(d1, d2), denom = a * e - b * b;
      if (denom > 0) { s = (b * f - c * e) / denom; s = s < 0 ? 0 : (s > 1 ? 1 : s); } else s = 0;
      t = (b * s + f) / e;
      if (t < 0) { t = 0; s = -c / a; s = s < 0 ? 0 : (s > 1 ? 1 : s); }
      else if (t > 1) { t = 1; s = (b - c) / a; s = s < 0 ? 0 : (s > 1 ? 1 : s); }
    }
  }
  double c1[3] = {p1[0] + d1[0] * s, p1[1] + d1[1] * s, p1[2] + d1[2] * s};
  double c2[3] = {p2[0] + d2[0] * t, p2[1] + d2[1] * t, p2[2] + d2[2] * t};
  double d[3]; sub3(d, c1, c2);
  return dot3(d, d);
}

/* Exact squared distance between segment a-b and the axis-aligned box [-half, half].
 * g(t) = sum_i max(|a_i + t d_i| - h_i, 0)^2 is convex and piecewise quadratic in t with
 * breakpoints where a coordinate crosses a slab plane, so g' is continuous, monotone and
 * piecewise linear: bracket its root between adjacent breakpoints and solve linearly. */
static inline double sb_g(const double a[3], const double d[3], const double h[3], double t, double* gp) {
  double g = 0, s = 0;
  for (int i = 0; i < 3; i++) {
    double p = a[i] + t * d[i], e = fabs(p) - h[i];
    if (e > 0) { g += e * e; s += (p > 0 ? e : -e) * d[i]; }
  }
  *gp = 2 * s;
  return g;
}
double oracle_seg_box_dist2(const double a[3], const double b[3], const double h[3]) {
  double d[3]; sub3(d, b, a);
  double gp0, gp1, g0 = sb_g(a, d, h, 0, &gp0), g1 = sb_g(a, d, h, 1, &gp1);
  if (gp0 >= 0) return g0;
  if (gp1 <= 0) return g1;
  double lo = 0, hi = 1, glo = gp0, ghi = gp1;
  for (int i = 0; i < 3; i++) {
    if (d[i] == 0) continue;
    for (int sgn = -1; sgn <= 1; sgn += 2) {
      double t = (sgn * h[i] - a[i]) / d[i];
      if (!(t > lo && t < hi)) continue;
      double gp; sb_g(a, d, h, t, &gp);
      if (gp <= 0) { lo = t; glo = gp; } else { hi = t; ghi = gp; }
    }
  }
  double t = lo;
  if (ghi > glo) t = lo - glo * (hi - lo) / (ghi - glo);
  if (t < lo) t = lo;
  if (t > hi) t = hi;
  double gp;
  return sb_g(a, d, h, t, &gp);
}

/* ------------------------------------------------------------------ GJK (FP64, polytopes) */
typedef struct {
  const double* v; /* n x 3 local core vertices */
  int n;
  Xf xf;           /* local -> world */
} Core;

static inline int support_local(const Core* c, const double* dir_world, double* out_world) {
  double dl[3];
  matTvec(dl, c->xf.R, dir_world);
  int best = 0;
  double bv = -DBL_MAX;
  for (int i = 0; i < c->n; i++) {
    double s = c->v[3 * i] * dl[0] + c->v[3 * i + 1] * dl[1] + c->v[3 * i + 2] * dl[2];
    if (s > bv) { bv = s; best = i; }
  }
  xf_point(out_world, &c->xf, c->v + 3 * best);
  return best;
}

typedef struct { double w[4][3]; int ia[4], ib[4]; int n; } Simplex;

/* closest point to the origin on segment ab; returns the kept-vertex mask (a = 1, b = 2) */
static int closest_seg(const double* a, const double* b, double* out) {
  double ab[3]; sub3(ab, b, a);
  double den = dot3(ab, ab), t = den > 0 ? -dot3(a, ab) / den : 0;
  if (t <= 0) { memcpy(out, a, 24); return 1; }
  if (t >= 1) { memcpy(out, b, 24); return 2; }
  for (int i = 0; i < 3; i++) out[i] = a[i] + t * ab[i];
  return 3;
}

/* closest point to the origin on a triangle (Ericson 5.1.5); returns the kept-vertex mask.
 * Polytopes hand GJK many collinear / coplanar support points, so a triangle whose area is
 * negligible against its edge lengths is solved as the best of its three edges. */
static int closest_tri(const double* a, const double* b, const double* c, double* out) {
  double ab[3], ac[3], bc[3], n[3];
  sub3(ab, b, a); sub3(ac, c, a); sub3(bc, c, b); cross3(n, ab, ac);
  double l2 = fmax(dot3(ab, ab), fmax(dot3(ac, ac), dot3(bc, bc)));
  if (dot3(n, n) <= 1e-24 * l2 * l2) {
    double q[3], best;
    int m = closest_seg(a, b, out), keep = m;
    best = dot3(out, out);
    m = closest_seg(a, c, q);
    if (dot3(q, q) < best) { best = dot3(q, q); memcpy(out, q, 24); keep = (m & 1) | ((m & 2) << 1); }
    m = closest_seg(b, c, q);
    if (dot3(q, q) < best) { memcpy(out, q, 24); keep = m << 1; }
    return keep;
  }
  double d1 = -dot3(ab, a), d2 = -dot3(ac, a);
  if (d1 <= 0 && d2 <= 0) { memcpy(out, a, 24); return 1; }
  double d3 = -dot3(ab, b), d4 = -dot3(ac, b);
  if (d3 >= 0 && d4 <= d3) { memcpy(out, b, 24); return 2; }
  double vc = d1 * d4 - d3 * d2;
  if (vc <= 0 && d1 >= 0 && d3 <= 0) {
    double v = d1 / (d1 - d3);
    for (int i = 0; i < 3; i++) out[i] = a[i] + v * ab[i];
    return 3;
  }
  double d5 = -dot3(ab, c), d6 = -dot3(ac, c);
  if (d6 >= 0 && d5 <= d6) { memcpy(out, c, 24); return 4; }
  double vb = d5 * d2 - d1 * d6;
  if (vb <= 0 && d2 >= 0 && d6 <= 0) {
    double w = d2 / (d2 - d6);
    for (int i = 0; i < 3; i++) out[i] = a[i] + w * ac[i];
    return 5;
  }
  double va = d3 * d6 - d5 * d4;
  if (va <= 0 && (d4 - d3) >= 0 && (d5 - d6) >= 0) {
    double w = (d4 - d3) / ((d4 - d3) + (d5 - d6));
    for (int i = 0; i < 3; i++) out[i] = b[i] + w * (c[i] - b[i]);
    return 6;
  }
  double den = 1.0 / (va + vb + vc), v = vb * den, w = vc * den;
  for (int i = 0; i < 3; i++) out[i] = a[i] + ab[i] * v + ac[i] * w;
  return 7;
}

/* reduce the simplex to the sub-simplex supporting the point closest to the origin; returns 1 if
 * the origin is enclosed (intersection).  A tetrahedron only counts as enclosing when it has real
 * volume (coplanar support points are the norm between polytopes); otherwise, and whenever the
 * origin is outside, the answer is the best of the four faces. */
static int simplex_closest(Simplex* s, double* v) {
  if (s->n == 1) { memcpy(v, s->w[0], 24); return 0; }
  int keep;
  if (s->n == 2) keep = closest_seg(s->w[0], s->w[1], v);
  else if (s->n == 3) keep = closest_tri(s->w[0], s->w[1], s->w[2], v);
  else {
    static const int F[4][4] = {{0, 1, 2, 3}, {0, 2, 3, 1}, {0, 3, 1, 2}, {1, 3, 2, 0}};
    const double *A = s->w[0], *B = s->w[1], *C = s->w[2], *D = s->w[3];
    double ab[3], ac[3], ad[3], bc[3], bd[3], cd[3], n[3];
    sub3(ab, B, A); sub3(ac, C, A); sub3(ad, D, A); sub3(bc, C, B); sub3(bd, D, B); sub3(cd, D, C);
    cross3(n, ab, ac);
    double vol6 = dot3(n, ad);
    double l2 = fmax(fmax(dot3(ab, ab), dot3(ac, ac)), fmax(fmax(dot3(ad, ad), dot3(bc, bc)), fmax(dot3(bd, bd), dot3(cd, cd))));
    if (fabs(vol6) > 1e-9 * l2 * sqrt(l2)) {
      int inside = 1;
      for (int f = 0; f < 4 && inside; f++) {
        const double *a = s->w[F[f][0]], *b = s->w[F[f][1]], *c = s->w[F[f][2]], *d = s->w[F[f][3]];
        double e1[3], e2[3], fn[3], ed[3];
        sub3(e1, b, a); sub3(e2, c, a); cross3(fn, e1, e2); sub3(ed, d, a);
        double sp = -dot3(a, fn), sd = dot3(ed, fn);
        if (!(sp * sd > 0)) inside = 0; /* origin strictly on d's side of every face */
      }
      if (inside) { v[0] = v[1] = v[2] = 0; return 1; }
    }
    double best = DBL_MAX;
    keep = 0;
    for (int f = 0; f < 4; f++) {
      double q[3];
      int m = closest_tri(s->w[F[f][0]], s->w[F[f][1]], s->w[F[f][2]], q);
      double dd = dot3(q, q);
      if (dd < best) {
        best = dd; memcpy(v, q, 24);
        keep = 0;
        for (int k = 0; k < 3; k++) if (m & (1 << k)) keep |= 1 << F[f][k];
      }
    }
  }
  int j = 0;
  for (int i = 0; i < s->n; i++)
    if (keep & (1 << i)) {
      if (i != j) { memcpy(s->w[j], s->w[i], 24); s->ia[j] = s->ia[i]; s->ib[j] = s->ib[i]; }
      j++;
    }
  s->n = j;
  return 0;
}

/* distance between two convex core polytopes; 0 when they intersect */
/* BASELINE MODE (oracle_set_fast): the timed CPU arm answers the boolean query the reference asks of coal
 * (collide(), not distance()): GJK stops as soon as a separating direction certifies clearance or the current
 * simplex point certifies penetration of the inflated shapes, cells stop at their first colliding pair
 * (CollisionCallback::collide sets is_collided and the traversal ends, CollisionManager.cpp:605-608) and safe
 * intervals go through the AABB tree over constant-pose windows (:107-216).  Every early exit is conservative
 * by 1e-9 m, so the predicate -- and therefore every bit, interval and plan -- is identical to the exact mode
 * (tests/test_oracle.py::test_fast_mode_is_bit_identical). */
static int g_fast = 0;
void oracle_set_fast(int on) { g_fast = on; }
int oracle_get_fast(void) { return g_fast; }
#define FAST_MARGIN 1e-9

static double gjk_core_r(const Core* A, const Core* B, double rsum) {
  Simplex s; s.n = 0;
  double v[3], pa[3], pb[3];
  sub3(v, A->xf.t, B->xf.t);
  if (dot3(v, v) == 0) v[0] = 1;
  double nv[3] = {-v[0], -v[1], -v[2]};
  int ia = support_local(A, nv, pa), ib = support_local(B, v, pb);
  sub3(s.w[0], pa, pb); s.ia[0] = ia; s.ib[0] = ib; s.n = 1;
  memcpy(v, s.w[0], 24);
  double vv = dot3(v, v);
  for (int it = 0; it < 256; it++) {
    if (vv <= 1e-30) return 0;
    double w[3];
    nv[0] = -v[0]; nv[1] = -v[1]; nv[2] = -v[2];
    ia = support_local(A, nv, pa); ib = support_local(B, v, pb);
    sub3(w, pa, pb);
    if (g_fast && rsum >= 0) {
      double vw = dot3(v, w), nv1 = sqrt(vv);
      if (vw > 0 && vw / nv1 - rsum > FAST_MARGIN) return vw / nv1; /* separating axis: distance >= v.w/|v| */
      if (nv1 - rsum < -FAST_MARGIN) return nv1;                    /* v is a point of A-B: distance <= |v| */
    }
    double gap = vv - dot3(v, w);
    if (gap <= 1e-15 * vv) break; /* v is the closest point (duality gap closed) */
    int dup = 0;
    for (int i = 0; i < s.n; i++) if (s.ia[i] == ia && s.ib[i] == ib) dup = 1;
    if (dup) break;
    memcpy(s.w[s.n], w, 24); s.ia[s.n] = ia; s.ib[s.n] = ib; s.n++;
    double nvv[3];
    if (simplex_closest(&s, nvv)) return 0;
    double n2 = dot3(nvv, nvv);
    if (!(n2 < vv)) break; /* no progress: numerically converged */
    memcpy(v, nvv, 24); vv = n2;
  }
  return sqrt(vv);
}

static double gjk_core(const Core* A, const Core* B) { return gjk_core_r(A, B, -1.0); }

double oracle_gjk_distance(const double* a, int na, const double* b, int nb) {
  Core A = {a, na, {{1, 0, 0, 0, 1, 0, 0, 0, 1}, {0, 0, 0}}}, B = {b, nb, {{1, 0, 0, 0, 1, 0, 0, 0, 1}, {0, 0, 0}}};
  return gjk_core(&A, &B);
}

/* ------------------------------------------------------------------ scene model */
typedef struct {
  int type, joints_before, n_verts;
  double lo_t[3], lo_R[9], radius, half_len;
  double* verts;  /* core vertices, local frame (hull verts, or the 2 capsule end points) */
  int n_core;
  double bc[3], br; /* bounding sphere of the inflated shape, local frame */
  /* coal CollisionGeometry::computeLocalAABB: aabb_local, aabb_center, aabb_radius = |min - center| */
  double amin[3], amax[3], ac[3], ar;
} OShape;

typedef struct {
  int dof, n_joints, n_shapes, n_strip, n_links;
  double base_t[3], base_R[9];
  float* limits;
  double* jt; /* n_joints x 3 */
  double* jR; /* n_joints x 9 */
  OShape* shapes;
} ORobot;

typedef struct {
  int type, is_static, n_poses;
  double half[3], radius;
  Xf* pose;          /* n_poses */
  float* raw;        /* n_poses x 7 (for window compression) */
  double corners[24]; /* box core, local */
  double br;
} OObst;

typedef struct {
  ORobot robot;
  int is_static, n_states;
  double *times, *angles;
  Xf* link_xf; /* [T][n_links] */
  int id0;     /* first reference obstacle id of this robot's links */
} ORobObst;

typedef struct { double lo[3], hi[3]; int left, right, leaf; } BvhNode;
typedef struct { int obstacle_id, f_lo, f_hi; double lo[3], hi[3]; } BvhLeaf;

struct Oracle {
  int T, fps, dof, C, R, parity_mode;
  double start_time;
  double *start_cfg, *end_cfg, *max_vel;
  ORobot robot;
  OObst* obst;
  ORobObst* robs;
  int n_rob_links;      /* sum of stripped links over robot obstacles */
  int n_rob_shapes_all; /* sum of all shapes (QUIRK 7 row stride) */
  int max_hull_verts;
  BvhNode* bvh; int n_bvh;
  BvhLeaf* leaves; int n_leaves;
};

typedef struct { const unsigned char* p; size_t n, o; int bad; } Cur;
static const void* cur_take(Cur* c, size_t bytes) {
  if (c->o + bytes > c->n) { c->bad = 1; return NULL; }
  const void* r = c->p + c->o;
  c->o += bytes;
  return r;
}
static void cur_pad8(Cur* c) { c->o = (c->o + 7) & ~(size_t)7; }

static void shape_finish(OShape* s) {
  double lo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, hi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
  for (int i = 0; i < s->n_core; i++)
    for (int k = 0; k < 3; k++) {
      double x = s->verts[3 * i + k];
      if (x < lo[k]) lo[k] = x;
      if (x > hi[k]) hi[k] = x;
    }
  for (int k = 0; k < 3; k++) s->bc[k] = 0.5 * (lo[k] + hi[k]);
  double r2 = 0;
  for (int i = 0; i < s->n_core; i++) {
    double d[3]; sub3(d, s->verts + 3 * i, s->bc);
    double q = dot3(d, d);
    if (q > r2) r2 = q;
  }
  s->br = sqrt(r2) + s->radius;
  for (int k = 0; k < 3; k++) { s->amin[k] = lo[k] - s->radius; s->amax[k] = hi[k] + s->radius; s->ac[k] = 0.5 * (s->amin[k] + s->amax[k]); }
  double d[3]; sub3(d, s->amin, s->ac);
  s->ar = sqrt(dot3(d, d));
}

/* coal CollisionObject::computeAABB: identity rotation -> translated local AABB, otherwise the
 * cube around the transformed aabb_center with half-side aabb_radius.  Only needed to restate
 * which (link, robot-obstacle link) pairs the broadphase hands to the mis-indexed functor. */
static void coal_aabb(const OShape* s, const Xf* x, double* lo, double* hi) {
  int ident = 1;
  for (int i = 0; i < 9; i++) if (fabs(x->R[i] - ((i % 4) == 0 ? 1.0 : 0.0)) > 1e-12) ident = 0;
  if (ident) { for (int k = 0; k < 3; k++) { lo[k] = s->amin[k] + x->t[k]; hi[k] = s->amax[k] + x->t[k]; } return; }
  double c[3]; xf_point(c, x, s->ac);
  for (int k = 0; k < 3; k++) { lo[k] = c[k] - s->ar; hi[k] = c[k] + s->ar; }
}
static int coal_aabb_overlap(const OShape* sa, const Xf* xa, const OShape* sb, const Xf* xb) {
  double alo[3], ahi[3], blo[3], bhi[3];
  coal_aabb(sa, xa, alo, ahi); coal_aabb(sb, xb, blo, bhi);
  for (int k = 0; k < 3; k++) if (alo[k] > bhi[k] || ahi[k] < blo[k]) return 0;
  return 1;
}

static int read_robot(Cur* c, ORobot* r) {
  const SirrtRobotHeader* h = cur_take(c, sizeof *h);
  if (!h) return -1;
  r->dof = h->dof; r->n_joints = h->n_joints; r->n_shapes = h->n_shapes; r->n_strip = h->n_strip;
  r->n_links = r->n_shapes - r->n_strip;
  if (r->n_links < 0) r->n_links = 0;
  for (int i = 0; i < 3; i++) r->base_t[i] = (double)h->base_pos[i];
  quat_to_R(r->base_R, (double)h->base_quat[0], (double)h->base_quat[1], (double)h->base_quat[2], (double)h->base_quat[3]);
  const float* lim = cur_take(c, sizeof(float) * 2 * r->dof);
  if (!lim) return -1;
  r->limits = malloc(sizeof(float) * 2 * (r->dof > 0 ? r->dof : 1));
  memcpy(r->limits, lim, sizeof(float) * 2 * r->dof);
  cur_pad8(c);
  r->jt = malloc(sizeof(double) * 3 * (r->n_joints + 1));
  r->jR = malloc(sizeof(double) * 9 * (r->n_joints + 1));
  for (int j = 0; j < r->n_joints; j++) {
    const SirrtJointRecord* jr = cur_take(c, sizeof *jr);
    if (!jr) return -1;
    memcpy(r->jt + 3 * j, jr->t, 24);
    memcpy(r->jR + 9 * j, jr->R, 72);
  }
  r->shapes = calloc(r->n_shapes + 1, sizeof(OShape));
  for (int s = 0; s < r->n_shapes; s++) {
    const SirrtShapeRecord* sr = cur_take(c, sizeof *sr);
    if (!sr) return -1;
    OShape* o = &r->shapes[s];
    o->type = sr->type; o->joints_before = sr->joints_before; o->n_verts = sr->n_verts;
    memcpy(o->lo_t, sr->lo_t, 24); memcpy(o->lo_R, sr->lo_R, 72);
    o->radius = sr->radius; o->half_len = sr->length / 2; /* coal::Capsule(r, lz): halfLength = lz/2 */
    const float* v = cur_take(c, sizeof(float) * 3 * sr->n_verts);
    if (sr->n_verts && !v) return -1;
    cur_pad8(c);
    if (o->type == SIRRT_SHAPE_CAPSULE) {
      o->n_core = 2; o->verts = calloc(6, sizeof(double));
      o->verts[2] = -o->half_len; o->verts[5] = o->half_len;
    } else {
      o->n_core = sr->n_verts; o->radius = 0;
      o->verts = malloc(sizeof(double) * 3 * (sr->n_verts + 1));
      for (int i = 0; i < 3 * sr->n_verts; i++) o->verts[i] = (double)v[i];
    }
    shape_finish(o);
  }
  return 0;
}

static void free_robot(ORobot* r) {
  if (r->shapes) for (int s = 0; s < r->n_shapes; s++) free(r->shapes[s].verts);
  free(r->shapes); free(r->limits); free(r->jt); free(r->jR);
}

/* a1 -- RobotObstacleFCL::get_collision_object_for_robot_angles_private, RobotObstacleFCL.cpp:287-386.
 * out[link] for the base-stripped links (erase of the first two, :370). */
static void robot_fk(const ORobot* r, const double* q, Xf* out) {
  double t[3] = {r->base_t[0], r->base_t[1], r->base_t[2]}, R[9];
  memcpy(R, r->base_R, sizeof R);
  int s = 0;
  for (int k = 0; k <= r->n_joints; k++) {
    while (s < r->n_shapes && r->shapes[s].joints_before == k) {
      if (s >= r->n_strip) {
        Xf* o = &out[s - r->n_strip];
        double w[3];
        matvec(w, R, r->shapes[s].lo_t);               /* :335 origin_in_world_basis */
        o->t[0] = t[0] + w[0]; o->t[1] = t[1] + w[1]; o->t[2] = t[2] + w[2]; /* :336 */
        matmul(o->R, R, r->shapes[s].lo_R);            /* :347 */
      }
      s++;
    }
    if (k == r->n_joints) break;
    double w[3];
    matvec(w, R, r->jt + 3 * k);                       /* :319,361 */
    t[0] += w[0]; t[1] += w[1]; t[2] += w[2];
    matmul(R, R, r->jR + 9 * k);                       /* :320,362 */
    double a = q[k], sn = sin(a), cs = cos(a);
    /* Eigen::AngleAxisd(a, UnitZ()).toRotationMatrix() (:321,363) */
    double Rz[9] = {cs, -sn, 0, sn, cs, 0, 0, 0, (1.0 - cs) * 1.0 + cs};
    matmul(R, R, Rz);
  }
}

int oracle_fk(const Oracle* o, const double* q, double* out) {
  Xf x[64];
  if (o->robot.n_links > 64) return -1;
  robot_fk(&o->robot, q, x);
  for (int l = 0; l < o->robot.n_links; l++) { memcpy(out + 12 * l, x[l].R, 72); memcpy(out + 12 * l + 9, x[l].t, 24); }
  return 0;
}

/* a7 -- RobotObstacleFCL::get_collision_object_for_frame, RobotObstacleFCL.cpp:393-429 */
static void robobst_config_at_frame(const Oracle* o, const ORobObst* ro, int frame, double* q) {
  int dof = ro->robot.dof, n = ro->n_states;
  double time = (double)frame / (double)o->fps + o->start_time; /* :399 */
  if (n <= 0) { for (int i = 0; i < dof; i++) q[i] = 0; return; }
  if (ro->times[n - 1] <= time + 0.00001) { memcpy(q, ro->angles + (size_t)(n - 1) * dof, sizeof(double) * dof); return; }
  for (int i = 0; i < n - 1; i++) {
    if (ro->times[i] <= time + 0.00001 && ro->times[i + 1] > time + 0.00001) {
      const double *l = ro->angles + (size_t)i * dof, *r = ro->angles + (size_t)(i + 1) * dof;
      double lt = ro->times[i], rt = ro->times[i + 1];
      for (int a = 0; a < dof; a++) {
        double k = (r[a] - l[a]) / (rt - lt);
        double b = r[a] - (r[a] - l[a]) / (rt - lt) * rt;
        q[a] = k * time + b; /* :417-419 */
      }
      return;
    }
  }
  /* before the first state the reference hits assert(false) (:429, UB in Release): hold state 0 */
  memcpy(q, ro->angles, sizeof(double) * dof);
}

int oracle_robot_obstacle_fk(const Oracle* o, int robot, int frame, double* out) {
  if (robot < 0 || robot >= o->R || frame < 0 || frame >= o->T) return -1;
  const ORobObst* ro = &o->robs[robot];
  for (int l = 0; l < ro->robot.n_links; l++) {
    const Xf* x = &ro->link_xf[(size_t)frame * ro->robot.n_links + l];
    memcpy(out + 12 * l, x->R, 72); memcpy(out + 12 * l + 9, x->t, 24);
  }
  return 0;
}

Oracle* oracle_create(const void* pack, size_t bytes, int parity_mode) {
  Cur c = {(const unsigned char*)pack, bytes, 0, 0};
  const SirrtPackHeader* h = cur_take(&c, sizeof *h);
  if (!h || memcmp(h->magic, SIRRT_PACK_MAGIC, 8) || h->version != SIRRT_PACK_VERSION || h->total_bytes != bytes) {
    set_err("bad scene pack header");
    return NULL;
  }
  Oracle* o = calloc(1, sizeof *o);
  o->T = h->n_frames; o->fps = h->fps; o->start_time = h->start_time; o->dof = h->dof;
  o->C = h->n_obstacles; o->R = h->n_robot_obstacles; o->parity_mode = parity_mode;
  size_t vb = sizeof(double) * (o->dof > 0 ? o->dof : 1);
  o->start_cfg = malloc(vb); o->end_cfg = malloc(vb); o->max_vel = malloc(vb);
  const void* p;
  if (!(p = cur_take(&c, sizeof(double) * o->dof))) goto bad;
  memcpy(o->start_cfg, p, sizeof(double) * o->dof);
  if (!(p = cur_take(&c, sizeof(double) * o->dof))) goto bad;
  memcpy(o->end_cfg, p, sizeof(double) * o->dof);
  if (!(p = cur_take(&c, sizeof(double) * o->dof))) goto bad;
  memcpy(o->max_vel, p, sizeof(double) * o->dof);
  if (read_robot(&c, &o->robot)) goto bad;
  o->obst = calloc(o->C + 1, sizeof(OObst));
  for (int i = 0; i < o->C; i++) {
    const SirrtObstacleRecord* r = cur_take(&c, sizeof *r);
    if (!r) goto bad;
    OObst* b = &o->obst[i];
    b->type = r->type; b->is_static = r->is_static; b->n_poses = r->n_poses;
    if (b->n_poses < 1 || (!b->is_static && b->n_poses < o->T)) goto bad;
    if (b->type == SIRRT_OBS_BOX) {
      for (int k = 0; k < 3; k++) b->half[k] = (double)r->dims[k] / 2; /* coal::Box(x,y,z): halfSide = side/2 */
      int n = 0;
      for (int sx = -1; sx <= 1; sx += 2) for (int sy = -1; sy <= 1; sy += 2) for (int sz = -1; sz <= 1; sz += 2) {
        b->corners[n++] = sx * b->half[0]; b->corners[n++] = sy * b->half[1]; b->corners[n++] = sz * b->half[2];
      }
      b->br = sqrt(dot3(b->half, b->half));
    } else {
      b->radius = (double)r->dims[0]; b->br = b->radius;
      memset(b->corners, 0, sizeof b->corners);
    }
    const float* ps = cur_take(&c, sizeof(float) * 7 * b->n_poses);
    if (!ps) goto bad;
    cur_pad8(&c);
    b->raw = malloc(sizeof(float) * 7 * b->n_poses);
    memcpy(b->raw, ps, sizeof(float) * 7 * b->n_poses);
    b->pose = malloc(sizeof(Xf) * b->n_poses);
    for (int f = 0; f < b->n_poses; f++) {
      const float* q = ps + 7 * f;
      for (int k = 0; k < 3; k++) b->pose[f].t[k] = (double)q[k];
      quat_to_R(b->pose[f].R, (double)q[3], (double)q[4], (double)q[5], (double)q[6]);
    }
  }
  o->robs = calloc(o->R + 1, sizeof(ORobObst));
  o->max_hull_verts = 0;
  for (int s = 0; s < o->robot.n_shapes; s++)
    if (o->robot.shapes[s].n_core > o->max_hull_verts) o->max_hull_verts = o->robot.shapes[s].n_core;
  int id = o->C;
  for (int i = 0; i < o->R; i++) {
    const SirrtRobotObstacleHeader* rh = cur_take(&c, sizeof *rh);
    if (!rh) goto bad;
    ORobObst* ro = &o->robs[i];
    ro->is_static = rh->is_static; ro->n_states = rh->n_states;
    if (read_robot(&c, &ro->robot)) goto bad;
    const double* tm = cur_take(&c, sizeof(double) * ro->n_states);
    const double* an = cur_take(&c, sizeof(double) * (size_t)ro->n_states * ro->robot.dof);
    if (ro->n_states && (!tm || !an)) goto bad;
    ro->times = malloc(sizeof(double) * (ro->n_states + 1));
    ro->angles = malloc(sizeof(double) * ((size_t)ro->n_states * ro->robot.dof + 1));
    memcpy(ro->times, tm, sizeof(double) * ro->n_states);
    memcpy(ro->angles, an, sizeof(double) * (size_t)ro->n_states * ro->robot.dof);
    ro->id0 = id; id += ro->robot.n_links;
    o->n_rob_links += ro->robot.n_links; o->n_rob_shapes_all += ro->robot.n_shapes;
    ro->link_xf = malloc(sizeof(Xf) * ((size_t)o->T * ro->robot.n_links + 1));
    double q[64];
    for (int f = 0; f < o->T; f++) {
      robobst_config_at_frame(o, ro, f, q);
      robot_fk(&ro->robot, q, ro->link_xf + (size_t)f * ro->robot.n_links);
    }
  }
  if (c.bad || c.o != bytes) goto bad;
  return o;
bad:
  set_err("truncated or malformed scene pack");
  oracle_destroy(o);
  return NULL;
}

void oracle_destroy(Oracle* o) {
  if (!o) return;
  free_robot(&o->robot);
  if (o->obst) for (int i = 0; i < o->C; i++) { free(o->obst[i].pose); free(o->obst[i].raw); }
  free(o->obst);
  if (o->robs) for (int i = 0; i < o->R; i++) {
    free_robot(&o->robs[i].robot); free(o->robs[i].times); free(o->robs[i].angles); free(o->robs[i].link_xf);
  }
  free(o->robs); free(o->start_cfg); free(o->end_cfg); free(o->max_vel); free(o->bvh); free(o->leaves);
  free(o);
}

int oracle_info(const Oracle* o, int* dof, int* n_links, int* n_frames, int* fps, float* lim) {
  if (dof) *dof = o->dof;
  if (n_links) *n_links = o->robot.n_links;
  if (n_frames) *n_frames = o->T;
  if (fps) *fps = o->fps;
  if (lim) memcpy(lim, o->robot.limits, sizeof(float) * 2 * o->robot.dof);
  return 0;
}
int oracle_counts(const Oracle* o, int* C, int* R, int* RL) {
  if (C) *C = o->C;
  if (R) *R = o->R;
  if (RL) *RL = o->n_rob_links;
  return 0;
}

/* ------------------------------------------------------------------ pair tests */
typedef struct {
  const OShape* shape; /* planned-robot link geometry */
  Xf xf;
  double bc[3];        /* world bounding-sphere centre */
  double alo[3], ahi[3]; /* coal AABB of the posed link (baseline mode's broadphase gate) */
} WLink;

static inline const Xf* obst_pose(const OObst* b, int frame) { return &b->pose[b->is_static ? 0 : frame]; }

/* signed distance: generic "core polytope + radius" shapes */
static double shape_shape_distance(const OShape* sa, const Xf* xa, const OShape* sb, const Xf* xb) {
  double dcore;
  if (sa->n_core <= 2 && sb->n_core <= 2) {
    double a0[3], a1[3], b0[3], b1[3];
    xf_point(a0, xa, sa->verts); xf_point(a1, xa, sa->verts + 3 * (sa->n_core - 1));
    xf_point(b0, xb, sb->verts); xf_point(b1, xb, sb->verts + 3 * (sb->n_core - 1));
    dcore = sqrt(oracle_seg_seg_dist2(a0, a1, b0, b1));
  } else {
    Core A = {sa->verts, sa->n_core, *xa}, B = {sb->verts, sb->n_core, *xb};
    dcore = gjk_core_r(&A, &B, sa->radius + sb->radius);
  }
  return dcore - sa->radius - sb->radius;
}

static double link_obstacle_distance(const WLink* L, const OObst* b, const Xf* xb) {
  const OShape* s = L->shape;
  if (s->n_core <= 2) {
    double a0[3], a1[3];
    xf_point(a0, &L->xf, s->verts); xf_point(a1, &L->xf, s->verts + 3 * (s->n_core - 1));
    if (b->type == SIRRT_OBS_SPHERE) return sqrt(oracle_pt_seg_dist2(xb->t, a0, a1)) - s->radius - b->radius;
    double d0[3], d1[3], l0[3], l1[3];
    sub3(d0, a0, xb->t); sub3(d1, a1, xb->t);
    matTvec(l0, xb->R, d0); matTvec(l1, xb->R, d1);
    return sqrt(oracle_seg_box_dist2(l0, l1, b->half)) - s->radius;
  }
  Core A = {s->verts, s->n_core, L->xf};
  Core B = {b->corners, b->type == SIRRT_OBS_BOX ? 8 : 1, *xb};
  const double rs = s->radius + (b->type == SIRRT_OBS_SPHERE ? b->radius : 0);
  return gjk_core_r(&A, &B, rs) - rs;
}

/* baseline mode: the reference's broadphase hands a pair to the narrowphase only when the AABBs overlap
 * (DynamicAABBTree, CollisionManager.cpp:215-216,272-276); boxes / spheres: AABB of the posed obstacle */
static inline int aabb_culled(const WLink* L, const OObst* b, const Xf* xb) {
  if (!g_fast) return 0;
  for (int k = 0; k < 3; k++) {
    double e = b->type == SIRRT_OBS_SPHERE ? b->radius
                                           : fabs(xb->R[3 * k]) * b->half[0] + fabs(xb->R[3 * k + 1]) * b->half[1] + fabs(xb->R[3 * k + 2]) * b->half[2];
    if (xb->t[k] - e > L->ahi[k] + FAST_MARGIN || xb->t[k] + e < L->alo[k] - FAST_MARGIN) return 1;
  }
  return 0;
}

static inline int culled(const double* ca, double ra, const double* cb, double rb) {
  double d[3]; sub3(d, ca, cb);
  double s = (ra + rb) * (1 + 1e-6) + CULL_MARGIN;
  return dot3(d, d) > s * s;
}

/* QUIRK 7 (CollisionManager.cpp:97-104,206-208,263-265,356,494,628,691): the functor picked for
 * robot-obstacle link (r, j) is the one built for shape k = (id - C) of the concatenation of ALL
 * shapes of all robot obstacles. */
static const OShape* shifted_shape(const Oracle* o, int k) {
  for (int r = 0; r < o->R; r++) {
    if (k < o->robs[r].robot.n_shapes) return &o->robs[r].robot.shapes[k];
    k -= o->robs[r].robot.n_shapes;
  }
  return NULL;
}

static void make_links(const Oracle* o, const double* q, WLink* L) {
  Xf x[64];
  robot_fk(&o->robot, q, x);
  for (int l = 0; l < o->robot.n_links; l++) {
    L[l].shape = &o->robot.shapes[l + o->robot.n_strip];
    L[l].xf = x[l];
    xf_point(L[l].bc, &x[l], L[l].shape->bc);
    coal_aabb(L[l].shape, &x[l], L[l].alo, L[l].ahi);
  }
}

/* min signed distance over robot-obstacle links at a frame under the given semantics */
static double robobst_min_distance(const Oracle* o, const WLink* L, int nl, int frame, int semantics, double m) {
  int k = 0;
  for (int r = 0; r < o->R; r++) {
    const ORobObst* ro = &o->robs[r];
    for (int j = 0; j < ro->robot.n_links; j++, k++) {
      const Xf* xb = &ro->link_xf[(size_t)frame * ro->robot.n_links + j];
      const OShape* st = &ro->robot.shapes[j + ro->robot.n_strip];
      const OShape* ss = shifted_shape(o, k);
      for (int l = 0; l < nl; l++) {
        int want_true = o->parity_mode == ORACLE_FIXED || semantics == ORACLE_SEM_INTERVAL;
        int want_shift = o->parity_mode == ORACLE_REFERENCE_EXACT;
        if (want_true) {
          double cb[3]; xf_point(cb, xb, st->bc);
          if (!culled(L[l].bc, L[l].shape->br, cb, st->br)) {
            double d = shape_shape_distance(L[l].shape, &L[l].xf, st, xb);
            if (d < m) m = d;
            if (g_fast && d - SECURITY_MARGIN <= COLLISION_THRESHOLD) return m;
          }
        }
        /* the broadphase only hands the pair to the callback when the TRUE geometries' AABBs overlap */
        if (want_shift && ss && coal_aabb_overlap(L[l].shape, &L[l].xf, st, xb)) {
          double cb[3]; xf_point(cb, xb, ss->bc);
          if (!culled(L[l].bc, L[l].shape->br, cb, ss->br)) {
            double d = shape_shape_distance(L[l].shape, &L[l].xf, ss, xb);
            if (d < m) m = d;
            if (g_fast && d - SECURITY_MARGIN <= COLLISION_THRESHOLD) return m;
          }
        }
      }
    }
  }
  return m;
}

static inline int is_collision(double d) { return d - SECURITY_MARGIN <= COLLISION_THRESHOLD; }

/* deciding signed distance of one (config, frame) cell; +inf-ish when everything is culled */
static double cell_distance(const Oracle* o, const WLink* L, int frame, int semantics, const double* static_m) {
  double m = static_m ? *static_m : 1e300;
  int nl = o->robot.n_links;
  for (int c = 0; c < o->C; c++) {
    const OObst* b = &o->obst[c];
    if (b->is_static && static_m) continue;
    const Xf* xb = obst_pose(b, frame);
    for (int l = 0; l < nl; l++) {
      if (culled(L[l].bc, L[l].shape->br, xb->t, b->br) || aabb_culled(&L[l], b, xb)) continue;
      double d = link_obstacle_distance(&L[l], b, xb);
      if (d < m) m = d;
      if (g_fast && is_collision(m)) return m;
    }
  }
  if (o->R) m = robobst_min_distance(o, L, nl, frame, semantics, m);
  return m;
}

static double static_min_distance(const Oracle* o, const WLink* L) {
  double m = 1e300;
  for (int c = 0; c < o->C; c++) {
    const OObst* b = &o->obst[c];
    if (!b->is_static) continue;
    for (int l = 0; l < o->robot.n_links; l++) {
      if (culled(L[l].bc, L[l].shape->br, b->pose[0].t, b->br) || aabb_culled(&L[l], b, &b->pose[0])) continue;
      double d = link_obstacle_distance(&L[l], b, &b->pose[0]);
      if (d < m) m = d;
    }
  }
  return m;
}

int oracle_collision_bits(const Oracle* o, const double* q, int semantics, uint32_t* bits, double tol, int64_t* near) {
  WLink L[64];
  if (o->robot.n_links > 64) return -1;
  make_links(o, q, L);
  int W = (o->T + 31) / 32;
  memset(bits, 0, sizeof(uint32_t) * W);
  double sm = static_min_distance(o, L);
  int64_t nc = 0;
  for (int f = 0; f < o->T; f++) {
    double m = cell_distance(o, L, f, semantics, &sm);
    if (is_collision(m)) bits[f >> 5] |= 1u << (f & 31);
    if (fabs(m) < tol) nc++;
  }
  if (near) *near = nc;
  return 0;
}

/* a3 -- interval extraction, CollisionManager.cpp:507-546 */
static int bits_to_intervals(const uint32_t* bits, int T, int32_t* lo_hi, int cap) {
  int n = 0, last = -1, any = 0;
  for (int f = 0; f < T; f++) {
    if (!(bits[f >> 5] >> (f & 31) & 1)) continue;
    any = 1;
    if (f - last != 1) { /* :525-529 */
      if (n < cap) { lo_hi[2 * n] = last + 1; lo_hi[2 * n + 1] = f - 1; }
      n++;
    }
    last = f;
  }
  if (!any) { /* :509-517 */
    if (cap > 0) { lo_hi[0] = 0; lo_hi[1] = T - 1; }
    return 1;
  }
  if (last != T - 1) { /* :533-538 */
    if (n < cap) { lo_hi[2 * n] = last + 1; lo_hi[2 * n + 1] = T - 1; }
    n++;
  }
  return n;
}

static void bits_bvh(const Oracle* o, const double* q, uint32_t* bits);

int oracle_safe_intervals(const Oracle* o, const double* q, int32_t* lo_hi, int cap) {
  int W = (o->T + 31) / 32;
  uint32_t* bits = malloc(sizeof(uint32_t) * W);
  if (g_fast && o->bvh) bits_bvh(o, q, bits); /* baseline mode: the reference's cross-frame AABB tree */
  else if (oracle_collision_bits(o, q, ORACLE_SEM_INTERVAL, bits, 0, NULL)) { free(bits); return -1; }
  int n = bits_to_intervals(bits, o->T, lo_hi, cap);
  free(bits);
  return n;
}

/* a4 -- CollisionManager::check_collision_private, CollisionManager.cpp:348-380 */
int oracle_check_frame(const Oracle* o, const double* q, int frame) {
  if (frame < 0 || frame >= o->T) return 1; /* reference: unchecked index (UB); we answer "collision" */
  WLink L[64];
  make_links(o, q, L);
  return is_collision(cell_distance(o, L, frame, ORACLE_SEM_FRAME, NULL));
}

double oracle_pair_distance(const Oracle* o, const double* q, int link, int id, int frame) {
  WLink L[64];
  make_links(o, q, L);
  if (link < 0 || link >= o->robot.n_links || frame < 0 || frame >= o->T) return NAN;
  if (id < o->C) return link_obstacle_distance(&L[link], &o->obst[id], obst_pose(&o->obst[id], frame));
  int k = id - o->C;
  for (int r = 0; r < o->R; r++) {
    const ORobObst* ro = &o->robs[r];
    if (k < ro->robot.n_links)
      return shape_shape_distance(L[link].shape, &L[link].xf, &ro->robot.shapes[k + ro->robot.n_strip],
                                  &ro->link_xf[(size_t)frame * ro->robot.n_links + k]);
    k -= ro->robot.n_links;
  }
  return NAN;
}

/* a5 -- PlannerConnect::is_collision_motion, PlannerConnect.cpp:862-897 */
int oracle_edge_check(const Oracle* o, const double* q0, const double* q1, double t0, double t1, double vmax,
                      double* collision_coord) {
  int dof = o->dof;
  if (t0 >= t1) return -2; /* :865 */
  double dir[64], n2 = 0;
  for (int i = 0; i < dof; i++) { dir[i] = q1[i] - q0[i]; n2 += dir[i] * dir[i]; }
  double norm = sqrt(n2);
  if (vmax > 0 && (norm * ((double)o->fps) / (t1 - t0)) > (vmax + 0.001)) return -3; /* :872 */
  int steps = (int)(norm / 0.1), s2 = (int)(t1 - t0 + 0.5); /* :879 */
  if (s2 > steps) steps = s2;
  if (steps < 1) steps = 1;
  for (int s = 0; s <= steps; s++) {
    double q[64];
    for (int i = 0; i < dof; i++) q[i] = q0[i] + dir[i] * (double)s / (double)steps; /* :885 */
    int frame = (int)(t0 + (t1 - t0) * (double)s / (double)steps);                    /* :886 */
    if (oracle_check_frame(o, q, frame)) {
      if (collision_coord) memcpy(collision_coord, q, sizeof(double) * dof);
      return s;
    }
  }
  return -1;
}

/* f4 -- StateValidityCheckerFunctor::isValid, STRRT_Planner/src/OMPLValidators/StateValidityCheckerFunctor.cpp:29-47:
 * float t = state time; CollisionManager::check_collision (CollisionManager.cpp:330-336): frame = (int)(t * fps) in float32 */
int oracle_state_valid(const Oracle* o, const double* q, double time_s) {
  float t = (float)time_s;
  volatile float ff = t * (float)o->fps;
  int frame = (ff >= 2147483647.0f || ff <= -2147483648.0f || ff != ff) ? -1 : (int)ff;
  return !oracle_check_frame(o, q, frame);
}

/* f4 -- SpaceTimeMotionValidator::checkMotion, SpaceTimeMotionValidator.cpp:22-74 (variant 0: 2*delta_t*fps, :51)
 * and :76-127 (variant 1: delta_t*fps, :108; reports lastValid.second).  Times in seconds.
 * returns -1 valid, -4 s2 invalid (:25/:82), -2 delta_t <= 0 (:41/:98), -3 speed gate (:46/:103),
 * >= 0 index of the first invalid interpolated state (:54-64 / :111-123).
 * The reference takes `abs(...)` (unqualified) of the double joint differences (:33, :90); with <cmath> in scope that is
 * the double overload, and the differences are squared right after, so only an integer-truncating abs could matter --
 * the reference's logged ST-RRT* solutions are valid motions under the double reading (tests/test_oracle.py). */
int oracle_motion_check(const Oracle* o, const double* q0, const double* q1, double t0, double t1, double vmax,
                        int variant, double* last_valid_fraction) {
  int dof = o->dof;
  if (last_valid_fraction) *last_valid_fraction = 0.0; /* :79 */
  if (!oracle_state_valid(o, q1, t1)) return -4;
  double n2 = 0;
  for (int i = 0; i < dof; i++) { double d = fabs(q1[i] - q0[i]); n2 += d * d; }
  double dt = t1 - t0;
  if (dt <= 0) return -2;
  double norm = sqrt(n2);
  if (vmax > 0 && norm / dt > vmax) return -3;
  int a = (int)(norm / 0.1), b = (int)((variant == 0 ? 2.0 : 1.0) * dt * (double)o->fps + 0.5);
  int steps = (a > b ? a : b) + 2;
  for (int i = 0; i < steps; i++) {
    double f = (double)i / (double)steps, q[64];
    for (int k = 0; k < dof; k++) { volatile double m = (q1[k] - q0[k]) * f; q[k] = q0[k] + m; } /* RealVectorStateSpace::interpolate */
    volatile double mt = (t1 - t0) * f;
    double tm = t0 + mt;                                                                   /* TimeStateSpace::interpolate */
    if (!oracle_state_valid(o, q, tm)) return i;
    if (last_valid_fraction) *last_valid_fraction = f; /* :121 */
  }
  return -1;
}

/* a9 -- CollisionManager::get_distances_private + DistanceCallback, CollisionManager.cpp:387-420,549-600 */
int oracle_get_distances(const Oracle* o, const double* q, int frame, double* min_dist) {
  WLink L[64];
  make_links(o, q, L);
  if (frame < 0 || frame >= o->T) return -1;
  for (int l = 0; l < o->robot.n_links; l++) {
    double m = 1e300;
    for (int c = 0; c < o->C; c++) {
      double d = link_obstacle_distance(&L[l], &o->obst[c], obst_pose(&o->obst[c], frame));
      if (d < m) m = d;
    }
    min_dist[l] = m;
    if (is_collision(m)) return 1;
  }
  return 0;
}

/* ------------------------------------------------------------------ BVH baseline (broadphase) */
static void leaf_aabb(const Oracle* o, BvhLeaf* lf, const double* c, double r) {
  (void)o;
  for (int k = 0; k < 3; k++) { lf->lo[k] = c[k] - r; lf->hi[k] = c[k] + r; }
}

static int bvh_build_rec(Oracle* o, int* idx, int n) {
  int me = o->n_bvh++;
  BvhNode* nd = &o->bvh[me];
  for (int k = 0; k < 3; k++) { nd->lo[k] = DBL_MAX; nd->hi[k] = -DBL_MAX; }
  double clo[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, chi[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
  for (int i = 0; i < n; i++) {
    const BvhLeaf* lf = &o->leaves[idx[i]];
    for (int k = 0; k < 3; k++) {
      if (lf->lo[k] < nd->lo[k]) nd->lo[k] = lf->lo[k];
      if (lf->hi[k] > nd->hi[k]) nd->hi[k] = lf->hi[k];
      double c = 0.5 * (lf->lo[k] + lf->hi[k]);
      if (c < clo[k]) clo[k] = c;
      if (c > chi[k]) chi[k] = c;
    }
  }
  if (n == 1) { nd->leaf = idx[0]; nd->left = nd->right = -1; return me; }
  nd->leaf = -1;
  int ax = 0;
  if (chi[1] - clo[1] > chi[ax] - clo[ax]) ax = 1;
  if (chi[2] - clo[2] > chi[ax] - clo[ax]) ax = 2;
  double mid = 0.5 * (clo[ax] + chi[ax]);
  int i = 0, j = n - 1;
  while (i <= j) {
    const BvhLeaf* lf = &o->leaves[idx[i]];
    if (0.5 * (lf->lo[ax] + lf->hi[ax]) < mid) i++;
    else { int t = idx[i]; idx[i] = idx[j]; idx[j] = t; j--; }
  }
  if (i == 0 || i == n) i = n / 2;
  int l = bvh_build_rec(o, idx, i);
  int r = bvh_build_rec(o, idx + i, n - i);
  o->bvh[me].left = l; o->bvh[me].right = r;
  return me;
}

/* one leaf per (obstacle, maximal run of bit-identical float poses): CollisionManager.cpp:142-166;
 * robot-obstacle links likewise on identical transforms (:175-211) */
int oracle_bvh_build(Oracle* o) {
  free(o->bvh); free(o->leaves);
  size_t cap = (size_t)(o->C + o->n_rob_links) * o->T + 1;
  o->leaves = malloc(sizeof(BvhLeaf) * cap);
  o->n_leaves = 0;
  for (int c = 0; c < o->C; c++) {
    const OObst* b = &o->obst[c];
    if (b->is_static) {
      BvhLeaf* lf = &o->leaves[o->n_leaves++];
      lf->obstacle_id = c; lf->f_lo = 0; lf->f_hi = o->T - 1;
      leaf_aabb(o, lf, b->pose[0].t, b->br);
      continue;
    }
    for (int f = 0; f < o->T; f++) {
      int e = f;
      while (e + 1 < o->T && !memcmp(b->raw + 7 * f, b->raw + 7 * (e + 1), 7 * sizeof(float))) e++;
      BvhLeaf* lf = &o->leaves[o->n_leaves++];
      lf->obstacle_id = c; lf->f_lo = f; lf->f_hi = e;
      leaf_aabb(o, lf, b->pose[f].t, b->br);
      f = e;
    }
  }
  for (int r = 0; r < o->R; r++) {
    const ORobObst* ro = &o->robs[r];
    int nl = ro->robot.n_links;
    for (int f = 0; f < o->T; f++) {
      int e = f;
      while (e + 1 < o->T && !memcmp(ro->link_xf + (size_t)f * nl, ro->link_xf + (size_t)(e + 1) * nl, sizeof(Xf) * nl)) e++;
      for (int j = 0; j < nl; j++) {
        const OShape* st = &ro->robot.shapes[j + ro->robot.n_strip];
        double cb[3]; xf_point(cb, &ro->link_xf[(size_t)f * nl + j], st->bc);
        double br = st->br;
        if (o->parity_mode == ORACLE_REFERENCE_EXACT) { /* the shifted functor geometry may be larger */
          const OShape* ss = shifted_shape(o, ro->id0 - o->C + j);
          if (ss) { double d[3]; sub3(d, ss->bc, st->bc); double e2 = sqrt(dot3(d, d)) + ss->br; if (e2 > br) br = e2; }
        }
        BvhLeaf* lf = &o->leaves[o->n_leaves++];
        lf->obstacle_id = ro->id0 + j; lf->f_lo = f; lf->f_hi = e;
        leaf_aabb(o, lf, cb, br);
      }
      f = e;
    }
  }
  int* idx = malloc(sizeof(int) * (o->n_leaves + 1));
  for (int i = 0; i < o->n_leaves; i++) idx[i] = i;
  o->bvh = malloc(sizeof(BvhNode) * (2 * (size_t)o->n_leaves + 2));
  o->n_bvh = 0;
  if (o->n_leaves) bvh_build_rec(o, idx, o->n_leaves);
  free(idx);
  return o->n_leaves;
}

static inline void mark(uint32_t* bits, int lo, int hi) { for (int f = lo; f <= hi; f++) bits[f >> 5] |= 1u << (f & 31); }
static inline int all_marked(const uint32_t* bits, int lo, int hi) {
  for (int f = lo; f <= hi; f++) if (!(bits[f >> 5] >> (f & 31) & 1)) return 0;
  return 1;
}

/* get_safe_intervals through the tree: per link one traversal, narrowphase per overlapping leaf,
 * skip leaves whose frames are all marked already (CollisionManager.cpp:495-505, 648-740) */
static void bits_bvh(const Oracle* o, const double* q, uint32_t* bits) {
  WLink L[64];
  make_links(o, q, L);
  int W = (o->T + 31) / 32;
  memset(bits, 0, sizeof(uint32_t) * W);
  if (!o->n_leaves) return;
  int stack[128];
  for (int l = 0; l < o->robot.n_links; l++) {
    double lo[3], hi[3], r = L[l].shape->br * (1 + 1e-6) + CULL_MARGIN;
    for (int k = 0; k < 3; k++) { lo[k] = L[l].bc[k] - r; hi[k] = L[l].bc[k] + r; }
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
      const BvhNode* nd = &o->bvh[stack[--sp]];
      if (nd->lo[0] > hi[0] || nd->hi[0] < lo[0] || nd->lo[1] > hi[1] || nd->hi[1] < lo[1] || nd->lo[2] > hi[2] || nd->hi[2] < lo[2]) continue;
      if (nd->leaf < 0) { stack[sp++] = nd->left; stack[sp++] = nd->right; continue; }
      const BvhLeaf* lf = &o->leaves[nd->leaf];
      if (all_marked(bits, lf->f_lo, lf->f_hi)) continue;
      double d;
      if (lf->obstacle_id < o->C) {
        const OObst* b = &o->obst[lf->obstacle_id];
        if (aabb_culled(&L[l], b, obst_pose(b, lf->f_lo))) continue;
        d = link_obstacle_distance(&L[l], b, obst_pose(b, lf->f_lo));
      } else {
        int k = lf->obstacle_id - o->C, kk = k, r2 = 0;
        while (kk >= o->robs[r2].robot.n_links) { kk -= o->robs[r2].robot.n_links; r2++; }
        const ORobObst* ro = &o->robs[r2];
        const Xf* xb = &ro->link_xf[(size_t)lf->f_lo * ro->robot.n_links + kk];
        d = shape_shape_distance(L[l].shape, &L[l].xf, &ro->robot.shapes[kk + ro->robot.n_strip], xb);
        if (o->parity_mode == ORACLE_REFERENCE_EXACT) {
          const OShape* ss = shifted_shape(o, k);
          if (ss && coal_aabb_overlap(L[l].shape, &L[l].xf, &ro->robot.shapes[kk + ro->robot.n_strip], xb)) {
            double d2 = shape_shape_distance(L[l].shape, &L[l].xf, ss, xb);
            if (d2 < d) d = d2;
          }
        }
      }
      if (is_collision(d)) mark(bits, lf->f_lo, lf->f_hi);
    }
  }
}

int oracle_safe_intervals_batch(const Oracle* o, const float* q_soa, int64_t N, int32_t* table, int Kmax,
                                uint32_t* bits_out, int use_bvh, int n_threads) {
  int W = (o->T + 31) / 32, dof = o->dof;
  if (use_bvh && !o->bvh) { set_err("bvh not built"); return -1; }
#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
#pragma omp parallel
  {
    uint32_t* bits = malloc(sizeof(uint32_t) * W);
    int32_t* iv = malloc(sizeof(int32_t) * 2 * (Kmax > 0 ? Kmax : 1));
#pragma omp for schedule(dynamic, 16)
    for (int64_t n = 0; n < N; n++) {
      double q[64];
      for (int j = 0; j < dof; j++) q[j] = (double)q_soa[(size_t)j * N + n];
      if (use_bvh) bits_bvh(o, q, bits);
      else oracle_collision_bits(o, q, ORACLE_SEM_INTERVAL, bits, 0, NULL);
      if (bits_out) memcpy(bits_out + (size_t)n * W, bits, sizeof(uint32_t) * W);
      if (table) {
        int cnt = bits_to_intervals(bits, o->T, iv, Kmax);
        int32_t* row = table + (size_t)n * (1 + 2 * Kmax);
        row[0] = cnt;
        for (int k = 0; k < 2 * Kmax; k++) row[1 + k] = k < 2 * (cnt < Kmax ? cnt : Kmax) ? iv[k] : -1;
      }
    }
    free(bits); free(iv);
  }
  return 0;
}

int oracle_check_frames_batch(const Oracle* o, const float* q_soa, const int32_t* frames, int64_t N, uint8_t* out,
                              int n_threads) {
  int dof = o->dof;
#ifdef _OPENMP
  if (n_threads > 0) omp_set_num_threads(n_threads);
#endif
#pragma omp parallel for schedule(dynamic, 64)
  for (int64_t n = 0; n < N; n++) {
    double q[64];
    for (int j = 0; j < dof; j++) q[j] = (double)q_soa[(size_t)j * N + n];
    out[n] = (uint8_t)oracle_check_frame(o, q, frames[n]);
  }
  return 0;
}

/* ------------------------------------------------------------------ kNN (a6) */
/* nanoflann L2_Simple_Adaptor::evalMetric: result += diff*diff in dimension order */
static inline double l2_simple(const double* a, const double* b, int dim) {
  double r = 0;
  for (int i = 0; i < dim; i++) { double d = a[i] - b[i]; r += d * d; }
  return r;
}
int oracle_knn_nearest(const double* pts, int64_t n, int dim, const double* queries, int64_t nq, int64_t* idx, double* dist2) {
  for (int64_t k = 0; k < nq; k++) {
    double best = DBL_MAX; int64_t bi = -1;
    for (int64_t i = 0; i < n; i++) {
      double d = l2_simple(queries + k * dim, pts + i * dim, dim);
      if (d < best) { best = d; bi = i; }
    }
    idx[k] = bi;
    if (dist2) dist2[k] = best;
  }
  return 0;
}
/* RadiusResultSet::addPoint: dist < radius (strict), PlannerConnect.cpp:394-412 passes r^2 */
int64_t oracle_knn_radius(const double* pts, int64_t n, int dim, const double* query, double r2, int64_t* idx, int64_t cap) {
  int64_t m = 0;
  for (int64_t i = 0; i < n; i++)
    if (l2_simple(query, pts + i * dim, dim) < r2) { if (m < cap) idx[m] = i; m++; }
  return m;
}
