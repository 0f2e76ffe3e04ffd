// Host-run check (nvcc only because the header is a CUDA header) of k_sweep's analytic frame windows
// (csrc/sirrt_sweep_window.cuh): random capsule links against SYN-like bouncing-box tiles (reflections, slow, static and
// axis-parallel motion, mass-test-like far coordinates).  Against an FP64 point-segment distance of the TRUE FP32 centres:
//   every frame with distance <= r_in  must be in the IN window   (else a collision could be missed),
//   every frame in the HIT window must have distance <  r_hit     (else a false collision would be reported).
// Prints one JSON line with the violation counts and the mean window sizes; exit code 1 on any violation.
//   sweep_window_host <n_trials>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>

#include "sirrt_sweep_window.cuh"

static double pt_seg(const double* p, const double* a, const double* b) {
  double ab[3], ap[3], den = 0, t = 0;
  for (int k = 0; k < 3; k++) { ab[k] = b[k] - a[k]; ap[k] = p[k] - a[k]; den += ab[k] * ab[k]; t += ap[k] * ab[k]; }
  t = den > 0 ? fmin(1.0, fmax(0.0, t / den)) : 0.0;
  double d2 = 0;
  for (int k = 0; k < 3; k++) { const double e = ap[k] - t * ab[k]; d2 += e * e; }
  return sqrt(d2);
}

int main(int argc, char** argv) {
  const long n_trials = argc > 1 ? atol(argv[1]) : 300000;
  std::mt19937_64 rng(11);
  std::uniform_real_distribution<double> U(-1.0, 1.0), U01(0.0, 1.0);
  long miss_in = 0, false_hit = 0, survivors = 0, n_in = 0, n_hit = 0, n_test = 0, true_in = 0, true_hit = 0, split = 0;
  double worst_in = 1e9, worst_hit = 1e9;
  for (long t = 0; t < n_trials; t++) {
    const int kind = (int)(t % 8);  // 0-3 SYN, 4 slow, 5 static, 6 parallel to the link axis, 7 far coordinates (mass test)
    const double off = kind == 7 ? 20.0 * U(rng) : 0.0;
    const double wall = 1.0;
    double p[3] = {U(rng), U(rng), U(rng)}, v[3] = {0.2 / 30 * U(rng), 0.2 / 30 * U(rng), 0.2 / 30 * U(rng)};
    if (kind == 4) for (double& x : v) x *= 1e-3;
    if (kind == 5) for (double& x : v) x = 0;
    // link: centre near the obstacle's path so that windows are exercised
    double u[3] = {U(rng), U(rng), U(rng)}, un = sqrt(u[0] * u[0] + u[1] * u[1] + u[2] * u[2]) + 1e-12;
    for (double& x : u) x /= un;
    if (kind == 6) { const double s = 0.003 + 0.004 * U01(rng); for (int k = 0; k < 3; k++) v[k] = u[k] * s; }
    const int nf = (t % 11 == 0) ? 1 + (int)(31 * U01(rng)) : 32;
    double c64[32][3];
    double q[3] = {p[0], p[1], p[2]};
    for (int f = 0; f < nf; f++) {
      for (int k = 0; k < 3; k++) c64[f][k] = (double)(float)(q[k] + (k == 0 ? off : 0.0));
      for (int k = 0; k < 3; k++) {
        q[k] += v[k];
        if (fabs(q[k]) > wall) { v[k] = -v[k]; q[k] = copysign(2.0 - fabs(q[k]), q[k]); }
      }
    }
    const double hlen = 0.04 + 0.08 * U01(rng), rl = 0.04 + 0.055 * U01(rng), h = 0.025 + 0.025 * U01(rng);
    const int fm = (int)(nf * U01(rng));
    const double dist = 0.35 * U01(rng);
    double dir[3] = {U(rng), U(rng), U(rng)}, dn = sqrt(dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2]) + 1e-12;
    float cf[3], abf[3];
    double a[3], b[3];
    for (int k = 0; k < 3; k++) {
      cf[k] = (float)(c64[fm][k] + dist * dir[k] / dn);
      abf[k] = (float)(2 * hlen * u[k]);
      a[k] = (double)cf[k] - 0.5 * (double)abf[k];
      b[k] = (double)cf[k] + 0.5 * (double)abf[k];
    }
    const float ab2 = abf[0] * abf[0] + abf[1] * abf[1] + abf[2] * abf[2];
    const float inv = ab2 > 0.f ? 1.f / ab2 : 0.f;
    const float band = 2e-5f;
    const float r_in = (float)(rl + h * sqrt(3.0)) + band, r_hit = (float)(rl + h) - band;
    float4 m[4];
    sweep_fit_model(c64, nf, m);
    if ((int)m[1].w < nf) split++;
    unsigned in_mask, hit_mask;
    sweep_windows(m[0], m[1], m[2], m[3], cf[0], cf[1], cf[2], abf[0], abf[1], abf[2], inv, r_in, r_hit, in_mask, hit_mask);
    bool any = false;
    for (int f = 0; f < nf; f++) {
      const double d = pt_seg(c64[f], a, b);
      const bool bi = (in_mask >> f) & 1u, bh = (hit_mask >> f) & 1u;
      if (d <= (double)r_in) { any = true; true_in++; if (!bi) { miss_in++; } }
      if (d < (double)r_hit) true_hit++;
      if (bh && !(d < (double)r_hit)) false_hit++;
      if (!bi) worst_in = fmin(worst_in, d - (double)r_in);
      if (bh) worst_hit = fmin(worst_hit, (double)r_hit - d);
    }
    if (nf < 32 && ((in_mask | hit_mask) >> nf)) miss_in += 1000000;  // bits beyond the tile
    if (any) { survivors++; n_in += __builtin_popcount(in_mask); n_hit += __builtin_popcount(hit_mask); n_test += __builtin_popcount(in_mask & ~hit_mask); }
  }
  printf("{\"trials\": %ld, \"survivors\": %ld, \"missed_in\": %ld, \"false_hit\": %ld, \"split_models\": %ld, \"true_in\": %ld, \"true_hit\": %ld, "
         "\"win_in\": %ld, \"win_hit\": %ld, \"win_test\": %ld, \"closest_excluded_margin\": %.3e, \"closest_hit_margin\": %.3e}\n",
         n_trials, survivors, miss_in, false_hit, split, true_in, true_hit, n_in, n_hit, n_test, worst_in, worst_hit);
  return (miss_in || false_hit || survivors == 0) ? 1 : 0;
}
