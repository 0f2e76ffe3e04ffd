// sirrt_sweep_window.cuh -- analytic frame windows of k_sweep's stage W (host + device).
//
// Within a 32-frame tile an obstacle's bounding-sphere centre follows a piecewise-linear path (the synthetic boxes and
// the mass-test cubes move at constant velocity and reflect at the walls).  sweep_fit_model() fits, on the host and in
// FP64, a TWO-piece linear model to the tile's FP32 centres and records the largest deviation `dev` of any true centre
// from it.  For one capsule link (centre c, axis ab = b - a) the squared distance from the model centre to
//   the link's supporting line   Ql(g) = |w|^2 - (w.ab)^2 / |ab|^2,   w = w0 + V g
//   the link centre              Qc(g) = |w|^2
//   the end points a, b          Qa(g) = |w + ab/2|^2,  Qb(g) = |w - ab/2|^2
// are quadratics in the frame g, so the set of frames on either side of a distance threshold is an interval:
//   IN  (superset of the frames whose centre-axis distance can be <= r_in):   Ql <= Rin^2  and  Qc <= (Rin + |ab|/2)^2
//   HIT (subset   of the frames whose centre-axis distance is certainly < r_hit):
//        Qa <= Rh^2  or  Qb <= Rh^2  or  (Ql <= Rh^2 and the foot of the perpendicular lies inside the segment)
// with Rin = r_in + dev + SWEEP_WIN_DELTA, Rh = r_hit - dev - SWEEP_WIN_DELTA.  The margins make both sets rigorous against the
// FP32 evaluation error of the coefficients (|error of Q| < 1e-5 m^2 for coordinates below 4 m, against a margin of
// 2 * R * DELTA > 3e-5 m^2 for R > 0.075 m); whatever lies between the two windows is tested frame by frame from the
// true centres (stage B2), so the model only decides how much per-frame work is left, never a collision bit near a boundary.
#pragma once
#include <cuda_runtime.h>
#include <math.h>

#define SWEEP_WIN_DELTA 2e-4f

#define SWEEP_HD __host__ __device__ __forceinline__

// integers g of [0, len-1] with  A g^2 + 2 B g + C <= 0, as a bit mask shifted by g0.
// deg (A too small for the root formula): 1 = the centre is static within the piece (it moves < 3.2e-6 m over 32 frames, far
// inside the margin): the value at g = 0 decides every frame; 2 = the quadratic is a cancelled difference (motion along the
// link axis): `superset` decides (superset: everything, subset: nothing).
SWEEP_HD unsigned sweep_quad_mask(float A, float B, float C, int deg, bool superset, int g0, int len) {
  if (len <= 0) return 0u;
  float lo = 0.f, hi = (float)(len - 1);
  if (deg == 1) {
    if (!(C <= 0.f)) return 0u;
  } else if (deg) {
    if (!superset) return 0u;
  } else {
    const float disc = B * B - A * C;
    if (!(disc >= 0.f)) return 0u;
#ifdef __CUDA_ARCH__
    // MUFU approximations (2 ulp): a root may move by 1e-5 frames, far inside what the radius margins absorb
    const float sq = disc * rsqrtf(fmaxf(disc, 1e-37f)), ia = __fdividef(1.f, A);
#else
    const float sq = sqrtf(disc), ia = 1.f / A;
#endif
    const float r1 = (-B - sq) * ia, r2 = (-B + sq) * ia;
    if (!(r1 <= r2)) return 0u;
    lo = fmaxf(lo, ceilf(r1));
    hi = fminf(hi, floorf(r2));
  }
  if (!(lo <= hi)) return 0u;
  const int l = (int)lo, h = (int)hi;
  return (0xffffffffu >> (31 - (h - l))) << (g0 + l);
}

// integers g of [0, len-1] with |wa + va g| <= h, shifted by g0
SWEEP_HD unsigned sweep_lin_mask(float wa, float va, float h, int g0, int len) {
  if (len <= 0) return 0u;
  float lo = 0.f, hi = (float)(len - 1);
  if (fabsf(va) < 1e-12f) {
    if (!(fabsf(wa) <= h)) return 0u;
  } else {
#ifdef __CUDA_ARCH__
    const float iv = __fdividef(1.f, va);
#else
    const float iv = 1.f / va;
#endif
    const float r1 = (-h - wa) * iv, r2 = (h - wa) * iv;
    lo = fmaxf(lo, ceilf(fminf(r1, r2)));
    hi = fminf(hi, floorf(fmaxf(r1, r2)));
  }
  if (!(lo <= hi)) return 0u;
  const int l = (int)lo, hh = (int)hi;
  return (0xffffffffu >> (31 - (hh - l))) << (g0 + l);
}

// one linear piece: centre(g) = P + V g for local frames g = 0 .. len-1 (tile frame g0 + g)
SWEEP_HD void sweep_piece_windows(float px, float py, float pz, float vx, float vy, float vz, int g0, int len, float cx, float cy, float cz,
                                  float abx, float aby, float abz, float inv, float ab2, float Rin, float Rh, unsigned& in_mask,
                                  unsigned& hit_mask) {
  if (len <= 0) return;
  const float wx = px - cx, wy = py - cy, wz = pz - cz;
  const float A1 = vx * vx + vy * vy + vz * vz, B1 = wx * vx + wy * vy + wz * vz, C1 = wx * wx + wy * wy + wz * wz;
  const float wa = wx * abx + wy * aby + wz * abz, va = vx * abx + vy * aby + vz * abz;
  const float A2 = A1 - va * va * inv, B2 = B1 - wa * va * inv, C2 = C1 - wa * wa * inv;
  const int deg1 = !(A1 >= 1e-14f) ? 1 : 0;                          // the centre does not move within the piece
  const int deg2 = deg1 ? 1 : (!(A2 >= 1e-4f * A1) ? 2 : 0);  // ... or moves along the link axis: A2 is a cancelled difference
  const float hl = 0.5f * sqrtf(ab2);
  const float Rc = Rin + hl;
  unsigned in = sweep_quad_mask(A1, B1, C1 - Rc * Rc, deg1, true, g0, len);
  if (in) in &= sweep_quad_mask(A2, B2, C2 - Rin * Rin, deg2, true, g0, len);
  in_mask |= in;
  if (Rh > 0.f && in) {
    const float R2 = Rh * Rh, q = 0.25f * ab2;
    unsigned h = sweep_quad_mask(A1, B1 + 0.5f * va, C1 + wa + q - R2, deg1, false, g0, len);
    h |= sweep_quad_mask(A1, B1 - 0.5f * va, C1 - wa + q - R2, deg1, false, g0, len);
    const unsigned hl_m = sweep_quad_mask(A2, B2, C2 - R2, deg2, false, g0, len);
    if (hl_m) h |= hl_m & sweep_lin_mask(wa, va, 0.499f * ab2, g0, len);
    hit_mask |= h;
  }
}

// model = 4 float4 per (tile, slot):  m0 = (Pa, dev)  m1 = (Va, split frame fs)  m2 = (Pb, frames in tile nf)  m3 = (Vb, 0)
//   piece a: tile frames 0 .. fs-1, centre = Pa + Va g;   piece b: tile frames fs .. nf-1, centre = Pb + Vb (g - fs)
// r_in  = link radius + obstacle bounding radius + band;  r_hit = link radius + obstacle inscribed radius - band
SWEEP_HD void sweep_windows(const float4& m0, const float4& m1, const float4& m2, const float4& m3, float cx, float cy, float cz, float abx,
                            float aby, float abz, float inv, float r_in, float r_hit, unsigned& in_mask, unsigned& hit_mask) {
  const float dev = m0.w;
  const float Rin = r_in + dev + SWEEP_WIN_DELTA, Rh = r_hit - dev - SWEEP_WIN_DELTA;
  const int fs = (int)m1.w, nf = (int)m2.w;
  const float ab2 = abx * abx + aby * aby + abz * abz;
  in_mask = 0u; hit_mask = 0u;
  sweep_piece_windows(m0.x, m0.y, m0.z, m1.x, m1.y, m1.z, 0, fs, cx, cy, cz, abx, aby, abz, inv, ab2, Rin, Rh, in_mask, hit_mask);
  sweep_piece_windows(m2.x, m2.y, m2.z, m3.x, m3.y, m3.z, fs, nf - fs, cx, cy, cz, abx, aby, abz, inv, ab2, Rin, Rh, in_mask, hit_mask);
  hit_mask &= in_mask;
}

// ---- host: fit the two-piece model to the nf (<= 32) centres of one tile ---------------------------------------
#include <algorithm>
static inline void sweep_fit_model(const double (*c)[3], int nf, float4 out[4]) {
  double best = 1e300;
  float bp[4][3] = {{0}};
  int bfs = nf;
  // the single-piece model (fs == nf) is tried first; a split is only adopted when it halves the deviation
  for (int it = 0; it < nf; it++) {
    const int fs = it == 0 ? nf : it;
    float Pa[3], Va[3], Pb[3], Vb[3];
    for (int k = 0; k < 3; k++) {
      Pa[k] = (float)c[0][k];
      Va[k] = fs > 1 ? (float)((c[fs - 1][k] - c[0][k]) / (fs - 1)) : 0.f;
      if (fs < nf) {
        Pb[k] = (float)c[fs][k];
        Vb[k] = nf - 1 > fs ? (float)((c[nf - 1][k] - c[fs][k]) / (nf - 1 - fs)) : 0.f;
      } else { Pb[k] = Pa[k]; Vb[k] = Va[k]; }
    }
    double dev = 0;
    for (int f = 0; f < nf; f++) {
      double d2 = 0;
      for (int k = 0; k < 3; k++) {
        const double m = f < fs ? (double)Pa[k] + (double)Va[k] * f : (double)Pb[k] + (double)Vb[k] * (f - fs);
        d2 += (m - c[f][k]) * (m - c[f][k]);
      }
      dev = std::max(dev, sqrt(d2));
    }
    if (it == 0 || dev < 0.5 * best) {
      best = dev; bfs = fs;
      for (int k = 0; k < 3; k++) { bp[0][k] = Pa[k]; bp[1][k] = Va[k]; bp[2][k] = Pb[k]; bp[3][k] = Vb[k]; }
    }
    if (best < 1e-6) break;
  }
  const float devf = (float)(best * (1.0 + 1e-6) + 1e-7);
  out[0] = make_float4(bp[0][0], bp[0][1], bp[0][2], devf);
  out[1] = make_float4(bp[1][0], bp[1][1], bp[1][2], (float)bfs);
  out[2] = make_float4(bp[2][0], bp[2][1], bp[2][2], (float)nf);
  out[3] = make_float4(bp[3][0], bp[3][1], bp[3][2], 0.f);
}
