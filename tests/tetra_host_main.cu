// Host-run check (nvcc only because the header is a CUDA header) of the FP64 enclosure certificate of the GJK task kernel
// (csrc/sirrt_device.cuh: tetra_encloses): random hull-vertex x box-corner quadruples around random relative poses;
// whenever the certificate says "the four Minkowski-difference points enclose the origin" an independent long-double
// barycentric test must agree and find the origin at least ~margin inside; flat / degenerate quadruples and quadruples
// with the origin outside must never be certified.  Prints one JSON line; exit code 1 on a wrong certificate or if the
// sample never produced a certified case.
//   tetra_host <n_trials>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <vector>

#include "sirrt_device.cuh"

static long double det3(const long double* a, const long double* b, const long double* c) {
  return a[0] * (b[1] * c[2] - b[2] * c[1]) - a[1] * (b[0] * c[2] - b[2] * c[0]) + a[2] * (b[0] * c[1] - b[1] * c[0]);
}

int main(int argc, char** argv) {
  const long n_trials = argc > 1 ? atol(argv[1]) : 200000;
  std::mt19937_64 rng(7);
  std::uniform_real_distribution<double> U(-1.0, 1.0);
  const int NV = 200;
  std::vector<double> verts(3 * NV);
  for (double& v : verts) v = 0.15 * U(rng);
  GShape<double> A{}, B{};
  A.type = SH_HULL; A.verts = verts.data(); A.n = NV;
  B.type = SH_BOX;
  std::vector<double> tiny(verts);
  for (double& v : tiny) v *= 1e-7;  // a speck: its tetrahedra have no real volume in metres and must never be certified
  long certified = 0, wrong = 0, exact_inside = 0, flat_certified = 0, tiny_certified = 0, tiny_inside = 0;
  for (long t = 0; t < n_trials; t++) {
    const bool speck = (t % 5) == 4;
    A.verts = speck ? tiny.data() : verts.data();
    B.half[0] = 0.02 + 0.05 * fabs(U(rng)); B.half[1] = 0.02 + 0.05 * fabs(U(rng)); B.half[2] = 0.02 + 0.05 * fabs(U(rng));
    double ax[3] = {U(rng), U(rng), U(rng)}, an = sqrt(ax[0] * ax[0] + ax[1] * ax[1] + ax[2] * ax[2]) + 1e-12, ang = 3.14159 * U(rng);
    for (double& x : ax) x /= an;
    const double c = cos(ang), s = sin(ang), C = 1 - c;
    double R[9] = {c + ax[0] * ax[0] * C, ax[0] * ax[1] * C - ax[2] * s, ax[0] * ax[2] * C + ax[1] * s,
                   ax[1] * ax[0] * C + ax[2] * s, c + ax[1] * ax[1] * C, ax[1] * ax[2] * C - ax[0] * s,
                   ax[2] * ax[0] * C - ax[1] * s, ax[2] * ax[1] * C + ax[0] * s, c + ax[2] * ax[2] * C};
    if (speck) for (int k = 0; k < 3; k++) B.half[k] *= 1e-7;
    int ids[4];
    const bool flat = (t % 7) == 0;  // degenerate quadruple: a repeated point
    for (int i = 0; i < 4; i++) ids[i] = ((int)(rng() % NV) << 12) | (int)(rng() % 8);
    if (flat) ids[3] = ids[t % 3];
    // put the origin near the centroid of the four difference points (inside about half of the time), at every depth scale
    double w0[4][3], trel0[3] = {0, 0, 0}, cen[3] = {0, 0, 0};
    for (int i = 0; i < 4; i++) {
      double pa[3], u[3], pb[3];
      gjk_point(A, ids[i] >> 12, pa); gjk_point(B, ids[i] & 4095, u); matvec(pb, R, u);
      for (int k = 0; k < 3; k++) { w0[i][k] = pa[k] - pb[k]; cen[k] += 0.25 * w0[i][k]; }
    }
    const double spread = (speck ? 1e-9 : 1.0) * pow(10.0, -10.0 + 9.5 * fabs(U(rng)));  // 1e-10 .. 0.3 m away from the centroid
    double trel[3];
    for (int k = 0; k < 3; k++) trel[k] = trel0[k] + cen[k] + spread * U(rng);
    double mf = 0;
    const bool cert = tetra_encloses(A, B, R, trel, ids, 1e-9, &mf);
    // independent test: barycentric coordinates of the origin w.r.t. the four points w0[i] - trel, in long double
    long double p[4][3];
    for (int i = 0; i < 4; i++) for (int k = 0; k < 3; k++) p[i][k] = (long double)w0[i][k] - (long double)trel[k];
    long double e1[3], e2[3], e3[3];
    for (int k = 0; k < 3; k++) { e1[k] = p[1][k] - p[0][k]; e2[k] = p[2][k] - p[0][k]; e3[k] = p[3][k] - p[0][k]; }
    const long double vol = det3(e1, e2, e3);
    bool inside = false;
    long double depth = 0;
    if (fabsl(vol) > 0) {
      inside = true;
      depth = 1e30L;
      const int F[4][4] = {{0, 1, 2, 3}, {0, 2, 3, 1}, {0, 3, 1, 2}, {1, 3, 2, 0}};
      for (int f = 0; f < 4; f++) {
        const long double *a = p[F[f][0]], *b = p[F[f][1]], *cc = p[F[f][2]], *d = p[F[f][3]];
        long double u1[3], u2[3], n[3], ed[3];
        for (int k = 0; k < 3; k++) { u1[k] = b[k] - a[k]; u2[k] = cc[k] - a[k]; ed[k] = d[k] - a[k]; }
        n[0] = u1[1] * u2[2] - u1[2] * u2[1]; n[1] = u1[2] * u2[0] - u1[0] * u2[2]; n[2] = u1[0] * u2[1] - u1[1] * u2[0];
        const long double nn = sqrtl(n[0] * n[0] + n[1] * n[1] + n[2] * n[2]);
        if (!(nn > 0)) { inside = false; break; }
        const long double so = -(a[0] * n[0] + a[1] * n[1] + a[2] * n[2]) / nn, sd = (ed[0] * n[0] + ed[1] * n[1] + ed[2] * n[2]) / nn;
        if (!(so * sd > 0)) { inside = false; break; }
        depth = fminl(depth, fabsl(so));
      }
    }
    exact_inside += inside;
    if (speck) { tiny_inside += inside; tiny_certified += cert; }
    if (cert) {
      certified++;
      if (flat) flat_certified++;
      if (!inside || depth < 0.9e-9L || fabs((double)depth - mf) > 1e-12 + 1e-9 * mf) wrong++;
    }
  }
  printf("{\"trials\": %ld, \"certified\": %ld, \"exact_inside\": %ld, \"wrong_certificates\": %ld, \"flat_certified\": %ld, \"speck_inside\": %ld, "
         "\"speck_certified\": %ld}\n", n_trials, certified, exact_inside, wrong, flat_certified, tiny_inside, tiny_certified);
  return (wrong || flat_certified || tiny_certified || tiny_inside < n_trials / 100 || certified < n_trials / 50) ? 1 : 0;
}
