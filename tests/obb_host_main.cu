// Host-run soundness check (nvcc only because the header is a CUDA header) of the oriented-box cull that sits in front of
// every hull pair test (csrc/sirrt_kernels.cuh: obb_separated, SIRRT_OPT_OBB_CULL): random pairs of oriented boxes -- link-
// sized and obstacle-sized, many of them pushed to within a few micrometres of contact -- posed with FP32 transforms as the
// kernels see them.  Whenever the FP32 15-axis test says "separated by more than margin" (margin = band + 1e-5 m, what the
// kernels pass), an independent FP64 feature computation (vertex-box, edge-edge, edge-through-box) must find the boxes
// disjoint and further apart than the band: a wrong "separated" would turn a colliding pair FREE without any GJK run.
// Prints one JSON line; exit code 1 on a wrong verdict or if the sample was one-sided.
//   obb_host <n_trials>
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>

#include "sirrt_kernels.cuh"

struct Box { double R[9], t[3], h[3]; };  // world = R * local + t, local in [-h, h]

static void corner(const Box& b, int i, double* o) {
  const double l[3] = {(i & 1 ? b.h[0] : -b.h[0]), (i & 2 ? b.h[1] : -b.h[1]), (i & 4 ? b.h[2] : -b.h[2])};
  for (int k = 0; k < 3; k++) o[k] = b.R[3 * k] * l[0] + b.R[3 * k + 1] * l[1] + b.R[3 * k + 2] * l[2] + b.t[k];
}
static void to_local(const Box& b, const double* p, double* o) {
  const double d[3] = {p[0] - b.t[0], p[1] - b.t[1], p[2] - b.t[2]};
  for (int k = 0; k < 3; k++) o[k] = b.R[k] * d[0] + b.R[3 + k] * d[1] + b.R[6 + k] * d[2];
}
static double point_box(const Box& b, const double* p) {
  double l[3], s = 0;
  to_local(b, p, l);
  for (int k = 0; k < 3; k++) { const double e = std::max(fabs(l[k]) - b.h[k], 0.0); s += e * e; }
  return sqrt(s);
}
// does segment p0-p1 meet box b?  (slab clipping in the box frame)
static bool seg_hits_box(const Box& b, const double* p0, const double* p1) {
  double a[3], c[3], t0 = 0, t1 = 1;
  to_local(b, p0, a); to_local(b, p1, c);
  for (int k = 0; k < 3; k++) {
    const double d = c[k] - a[k];
    if (fabs(d) < 1e-300) { if (fabs(a[k]) > b.h[k]) return false; continue; }
    double u = (-b.h[k] - a[k]) / d, v = (b.h[k] - a[k]) / d;
    if (u > v) std::swap(u, v);
    t0 = std::max(t0, u); t1 = std::min(t1, v);
    if (t0 > t1) return false;
  }
  return true;
}
static double seg_seg(const double* p1, const double* q1, const double* p2, const double* q2) {
  double d1[3], d2[3], r[3];
  for (int k = 0; k < 3; k++) { d1[k] = q1[k] - p1[k]; d2[k] = q2[k] - p2[k]; r[k] = p1[k] - p2[k]; }
  auto dot = [](const double* a, const double* b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; };
  const double a = dot(d1, d1), e = dot(d2, d2), f = dot(d2, r), c = dot(d1, r), b = dot(d1, d2), den = a * e - b * b;
  auto cl = [](double x) { return x < 0 ? 0.0 : (x > 1 ? 1.0 : x); };
  double s = den > 0 ? cl((b * f - c * e) / den) : 0.0, t = (b * s + f) / e;
  if (t < 0) { t = 0; s = cl(-c / a); } else if (t > 1) { t = 1; s = cl((b - c) / a); }
  double dd = 0;
  for (int k = 0; k < 3; k++) { const double x = p1[k] + d1[k] * s - p2[k] - d2[k] * t; dd += x * x; }
  return sqrt(dd);
}
static const int EDGE[12][2] = {{0, 1}, {2, 3}, {4, 5}, {6, 7}, {0, 2}, {1, 3}, {4, 6}, {5, 7}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};
// exact distance of two boxes from their features; 0 when they meet
static double box_box(const Box& A, const Box& B) {
  double va[8][3], vb[8][3], d = 1e300;
  for (int i = 0; i < 8; i++) { corner(A, i, va[i]); corner(B, i, vb[i]); }
  for (int e = 0; e < 12; e++) {
    if (seg_hits_box(B, va[EDGE[e][0]], va[EDGE[e][1]]) || seg_hits_box(A, vb[EDGE[e][0]], vb[EDGE[e][1]])) return 0.0;
  }
  for (int i = 0; i < 8; i++) d = std::min(d, std::min(point_box(B, va[i]), point_box(A, vb[i])));
  for (int e = 0; e < 12; e++)
    for (int g = 0; g < 12; g++) d = std::min(d, seg_seg(va[EDGE[e][0]], va[EDGE[e][1]], vb[EDGE[g][0]], vb[EDGE[g][1]]));
  return d;
}

static void rand_rot(std::mt19937_64& rng, double* R) {
  std::uniform_real_distribution<double> U(-1.0, 1.0);
  double ax[3] = {U(rng), U(rng), U(rng)}, an = sqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]) + 1e-12, ang = 3.14159265 * U(rng);
  if ((rng() & 7) == 0) ang = 0.5 * 3.14159265 * (double)(int)(4 * U(rng));  // axis-aligned-ish poses: parallel edges, the SAT's weak spot
  for (double& x : ax) x /= an;
  if ((rng() & 7) == 1) { ax[0] = 0; ax[1] = 0; ax[2] = 1; }
  const double c = cos(ang), s = sin(ang), C = 1 - c;
  const double M[9] = {c + ax[0] * ax[0] * C, ax[0] * ax[1] * C - ax[2] * s, ax[0] * ax[2] * C + ax[1] * s,
                       ax[1] * ax[0] * C + ax[2] * s, c + ax[1] * ax[1] * C, ax[1] * ax[2] * C - ax[0] * s,
                       ax[2] * ax[0] * C - ax[1] * s, ax[2] * ax[1] * C + ax[0] * s, c + ax[2] * ax[2] * C};
  for (int i = 0; i < 9; i++) R[i] = M[i];
}

int main(int argc, char** argv) {
  const long n_trials = argc > 1 ? atol(argv[1]) : 60000;
  const float band = 2e-5f, margin = band + 1e-5f;  // what k_collide_bits / k_check_cells pass
  std::mt19937_64 rng(11);
  std::uniform_real_distribution<double> U(-1.0, 1.0), U01(0.0, 1.0);
  static WarpCtx W;  // only R, oc, oh of link 0 are read
  long n_sep = 0, wrong = 0, n_near = 0, near_sep = 0, n_far = 0, far_sep = 0, n_meet = 0;
  double worst = 1e300;
  for (long t = 0; t < n_trials; t++) {
    // link box: centre off the frame origin (the xArm6 hulls sit up to 14 cm off), long and thin; obstacle box: a cube or a slab
    float aminA[3], amaxA[3], Rf[9], tf[3], RBf[9], tBf[3];
    DShape sb{};
    double RA[9], RB[9];
    rand_rot(rng, RA); rand_rot(rng, RB);
    for (int k = 0; k < 3; k++) {
      const double c = 0.15 * U(rng), h = 0.02 + 0.17 * U01(rng) * U01(rng);
      aminA[k] = (float)(c - h); amaxA[k] = (float)(c + h);
      const double cb = (rng() & 3) ? 0.0 : 0.05 * U(rng), hb = 0.01 + 0.3 * U01(rng) * U01(rng);
      sb.amin[k] = (float)(cb - hb); sb.amax[k] = (float)(cb + hb);
      sb.ac[k] = 0.5f * (sb.amin[k] + sb.amax[k]);
      tf[k] = (float)(1.2 * U(rng));
    }
    for (int i = 0; i < 9; i++) { Rf[i] = (float)RA[i]; RBf[i] = (float)RB[i]; }
    const float acA[3] = {0.5f * (aminA[0] + amaxA[0]), 0.5f * (aminA[1] + amaxA[1]), 0.5f * (aminA[2] + amaxA[2])};
    Box A, B;
    for (int i = 0; i < 9; i++) { A.R[i] = Rf[i]; B.R[i] = RBf[i]; }
    for (int k = 0; k < 3; k++) {
      A.h[k] = 0.5 * ((double)amaxA[k] - (double)aminA[k]); B.h[k] = 0.5 * ((double)sb.amax[k] - (double)sb.amin[k]);
      A.t[k] = A.R[3 * k] * acA[0] + A.R[3 * k + 1] * acA[1] + A.R[3 * k + 2] * acA[2] + tf[k];
    }
    // B's frame origin such that its box centre lands at cA + s * u
    double u[3] = {U(rng), U(rng), U(rng)}, un = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]) + 1e-12;
    for (double& x : u) x /= un;
    auto place = [&](double s) {
      for (int k = 0; k < 3; k++) {
        B.t[k] = A.t[k] + s * u[k];  // box centre
        const double off = B.R[3 * k] * sb.ac[0] + B.R[3 * k + 1] * sb.ac[1] + B.R[3 * k + 2] * sb.ac[2];
        tBf[k] = (float)(B.t[k] - off);
        B.t[k] = (double)tBf[k] + off;  // the centre the FP32 pose really gives
      }
    };
    double s;
    if (t % 3 != 0) {  // near contact: bisect the shift for a target gap, then jitter (some end up a hair inside)
      static const double G[8] = {1e-6, 5e-6, 1.5e-5, 2.5e-5, 3.5e-5, 1e-4, 1e-3, 1e-2};
      const double g = G[rng() & 7] * (0.5 + U01(rng));
      double lo = 0, hi = 3;
      for (int it = 0; it < 44; it++) { const double m = 0.5 * (lo + hi); place(m); if (box_box(A, B) > g) hi = m; else lo = m; }
      s = hi - ((rng() & 7) == 0 ? 2e-5 * U01(rng) + g : 0.0);
    } else s = 1.2 * U01(rng);
    place(s);
    const double d = box_box(A, B);
    for (int i = 0; i < 9; i++) W.R[0][i] = Rf[i];
    for (int k = 0; k < 3; k++) {  // as setup_links: FP32 centre and half extents
      W.oc[0][k] = Rf[3 * k] * acA[0] + Rf[3 * k + 1] * acA[1] + Rf[3 * k + 2] * acA[2] + tf[k];
      W.oh[0][k] = 0.5f * (amaxA[k] - aminA[k]);
    }
    const bool sep = obb_separated(W, 0, sb, RBf, tBf, margin);
    n_sep += sep;
    if (d == 0.0) n_meet++;
    if (d < 1e-4) { n_near++; near_sep += sep; }
    if (d > 1e-3) { n_far++; far_sep += sep; }
    if (sep) { worst = std::min(worst, d); if (!(d > (double)band)) wrong++; }
  }
  printf("{\"trials\": %ld, \"separated\": %ld, \"wrong_separated\": %ld, \"closest_separated_m\": %.3g, \"near_contact\": %ld, \"near_contact_separated\": %ld, "
         "\"touching_or_inside\": %ld, \"clear_by_1mm\": %ld, \"clear_by_1mm_separated\": %ld}\n",
         n_trials, n_sep, wrong, worst, n_near, near_sep, n_meet, n_far, far_sep);
  return (wrong || n_sep < n_trials / 10 || n_near < n_trials / 10 || n_meet < 50) ? 1 : 0;
}
