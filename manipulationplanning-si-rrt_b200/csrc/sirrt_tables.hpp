// sirrt_tables.hpp -- HScene -> flat device tables (layout documented in sirrt_device.cuh / DESIGN.md).
#pragma once
#include <cuda_runtime.h>

#include "sirrt_host.hpp"

namespace sirrt {

struct HostTables {
  int T = 0, Tp = 0, n_tiles = 0, n_slots = 0, has_hull = 0, n_robot_slots = 0, n_static = 0, n_dynamic = 0;
  DRobot robot;
  std::vector<DShape> shapes;
  std::vector<DSlot> slots;
  std::vector<int> static_slots, dynamic_slots;
  std::vector<float> pose_f, bcw_f, aabb_f, tight_f, verts_f;
  std::vector<double> pose_d, aabb_d, verts_d;
  std::vector<float4> tile_bounds;
  std::vector<unsigned int> hull_cells;     // per hull SIRRT_HULL_CELLS + 1 offsets into hull_cand
  std::vector<unsigned short> hull_cand;
  // FP32 inline form of the same cells: SIRRT_HULL_REC float4 (x, y, z, vertex index as int bits; -1 = empty slot) per entry
  // of hull_cells, so that a support query is ONE load per lane instead of the chain offsets -> candidate list -> vertices;
  // a cell with more candidates than slots carries -2 in its last slot and is scanned through hull_cand
  std::vector<float4> hull_rec;
  double extent = 0;  // largest |coordinate| the FP32 path can meet near the arm (band scaling)
};

// Support-candidate cells of one hull.  For a cube-map cell with unit centre direction c and chord radius eps
// (|d/|d| - c| <= eps for every direction d of the dilated cell), a vertex v can only be the support vertex for such a d
// if  v.c >= h(c) - |v - w*| eps  (w* = a support vertex for c, h(c) = w*.c): from v.d >= w*.d with d = c + (d - c).
// `slack` covers the rounding of the dot products the device compares (FP32: ~2e-7 |v| per unit direction), so the
// winner of the device's full FP32 / FP64 scan is always listed.
inline void build_hull_cells(const std::vector<float>& verts, std::vector<unsigned int>& cells, std::vector<unsigned short>& cand) {
  const int n = (int)(verts.size() / 3), K = SIRRT_HULL_K;
  double rmax = 0;
  for (int i = 0; i < n; i++) rmax = std::max(rmax, std::sqrt((double)verts[3 * i] * verts[3 * i] + (double)verts[3 * i + 1] * verts[3 * i + 1] + (double)verts[3 * i + 2] * verts[3 * i + 2]));
  const double slack = 1e-5 * std::max(rmax, 1e-3);
  for (int face = 0; face < 6; face++) {
    const int a = face >> 1;
    const double sgn = (face & 1) ? -1.0 : 1.0;
    for (int iu = 0; iu < K; iu++)
      for (int iv = 0; iv < K; iv++) {
        cells.push_back((unsigned int)cand.size());
        const double u0 = -1.0 + 2.0 * iu / K - SIRRT_HULL_CELL_PAD, u1 = -1.0 + 2.0 * (iu + 1) / K + SIRRT_HULL_CELL_PAD;
        const double v0 = -1.0 + 2.0 * iv / K - SIRRT_HULL_CELL_PAD, v1 = -1.0 + 2.0 * (iv + 1) / K + SIRRT_HULL_CELL_PAD;
        auto dir = [&](double u, double v, double* o) {
          o[a] = sgn; o[(a + 1) % 3] = u; o[(a + 2) % 3] = v;
          const double nn = std::sqrt(o[0] * o[0] + o[1] * o[1] + o[2] * o[2]);
          o[0] /= nn; o[1] /= nn; o[2] /= nn;
        };
        double c[3], k4[4][3];
        dir(0.5 * (u0 + u1), 0.5 * (v0 + v1), c);
        dir(u0, v0, k4[0]); dir(u0, v1, k4[1]); dir(u1, v0, k4[2]); dir(u1, v1, k4[3]);
        double eps = 0;
        for (auto& k : k4) eps = std::max(eps, std::sqrt((k[0] - c[0]) * (k[0] - c[0]) + (k[1] - c[1]) * (k[1] - c[1]) + (k[2] - c[2]) * (k[2] - c[2])));
        eps *= 1.001;
        double h = -1e300;
        int w = 0;
        for (int i = 0; i < n; i++) {
          const double dv = verts[3 * i] * c[0] + verts[3 * i + 1] * c[1] + verts[3 * i + 2] * c[2];
          if (dv > h) { h = dv; w = i; }
        }
        for (int i = 0; i < n; i++) {
          const double dv = verts[3 * i] * c[0] + verts[3 * i + 1] * c[1] + verts[3 * i + 2] * c[2];
          const double ex = (double)verts[3 * i] - verts[3 * w], ey = (double)verts[3 * i + 1] - verts[3 * w + 1], ez = (double)verts[3 * i + 2] - verts[3 * w + 2];
          if (dv >= h - std::sqrt(ex * ex + ey * ey + ez * ez) * eps - slack) cand.push_back((unsigned short)i);
        }
      }
  }
  cells.push_back((unsigned int)cand.size());
  if (getenv("SIRRT_DEBUG_CELLS")) {
    unsigned mx = 0;
    const size_t c0 = cells.size() - 1 - SIRRT_HULL_CELLS;
    for (size_t i = c0; i + 1 < cells.size(); i++) mx = std::max(mx, cells[i + 1] - cells[i]);
    fprintf(stderr, "hull cells: %d vertices, mean %.1f candidates per cell, max %u\n", n, (double)(cells.back() - cells[c0]) / SIRRT_HULL_CELLS, mx);
  }
}

inline void to_dshape(const HShape& h, DShape& d, std::vector<float>& vf, std::vector<double>& vd, std::vector<unsigned int>& cells,
                      std::vector<unsigned short>& cand) {
  memset(&d, 0, sizeof d);
  d.type = h.type;
  d.vert_off = (int)(vf.size() / 3);
  d.n_verts = h.type == SH_HULL ? (int)(h.verts.size() / 3) : 0;
  d.cell_off = -1;
  if (h.type == SH_HULL) {
    for (float v : h.verts) { vf.push_back(v); vd.push_back((double)v); }
    if (d.n_verts > 64 && d.n_verts < 65536) {  // small hulls: the full scan is two steps anyway
      d.cell_off = (int)cells.size();
      build_hull_cells(h.verts, cells, cand);
    }
  }
  d.radius = (float)h.radius; d.half_len = (float)h.half_len;
  d.d_radius = h.radius; d.d_half_len = h.half_len;
  for (int k = 0; k < 3; k++) {
    d.half[k] = (float)h.half[k]; d.d_half[k] = h.half[k];
    d.bc[k] = (float)h.bc[k];
    d.amin[k] = (float)h.amin[k]; d.amax[k] = (float)h.amax[k]; d.ac[k] = (float)h.ac[k];
    d.d_amin[k] = h.amin[k]; d.d_amax[k] = h.amax[k]; d.d_ac[k] = h.ac[k];
  }
  d.br = (float)(h.br * (1 + 1e-6));  // round-off guard: the float radius must not be smaller
  d.ar = (float)h.ar; d.d_ar = h.ar;
}

// tight world AABB of a posed shape: centre R*ac + t, half extents sum_j |R_kj| * (amax_j - amin_j)/2
inline void tight_aabb(const HShape& s, const HXf& x, double* lo, double* hi) {
  double c[3]; h_matvec(c, x.R, s.ac);
  for (int k = 0; k < 3; k++) {
    double e = 0;
    for (int j = 0; j < 3; j++) e += fabs(x.R[3 * k + j]) * 0.5 * (s.amax[j] - s.amin[j]);
    lo[k] = c[k] + x.t[k] - e; hi[k] = c[k] + x.t[k] + e;
  }
}

// coal CollisionObject::computeAABB: identity rotation -> translated local AABB, otherwise the cube
// of half-side aabb_radius around the transformed aabb_center
inline void coal_aabb(const HShape& s, const HXf& x, double* lo, double* hi) {
  bool ident = true;
  for (int i = 0; i < 9; i++) ident &= fabs(x.R[i] - ((i % 4) == 0 ? 1.0 : 0.0)) <= 1e-12;
  if (ident) { for (int k = 0; k < 3; k++) { lo[k] = s.amin[k] + x.t[k]; hi[k] = s.amax[k] + x.t[k]; } return; }
  double c[3]; h_matvec(c, x.R, s.ac);
  for (int k = 0; k < 3; k++) { lo[k] = c[k] + x.t[k] - s.ar; hi[k] = c[k] + x.t[k] + s.ar; }
}

inline int build_tables(const HScene& sc, int parity_mode, HostTables& tb, std::string& err) {
  const HRobot& rb = sc.robot;
  if (rb.n_joints > SIRRT_MAXJ || rb.n_links > SIRRT_MAXL) { err = "planned robot exceeds SIRRT_MAXJ/SIRRT_MAXL"; return -4; }
  if (rb.n_links < 1) { err = "planned robot has no collision links after the base strip"; return -2; }
  const int T = sc.T, Tp = (T + 31) / 32 * 32, n_tiles = Tp / 32;
  tb.T = T; tb.Tp = Tp; tb.n_tiles = n_tiles;

  // ---- planned robot
  DRobot& dr = tb.robot;
  memset(&dr, 0, sizeof dr);
  dr.n_joints = rb.n_joints; dr.n_links = rb.n_links; dr.dof = rb.dof;
  memcpy(dr.base_t, rb.base_t, 24); memcpy(dr.base_R, rb.base_R, 72);
  for (int j = 0; j < rb.n_joints; j++) { memcpy(dr.jt[j], &rb.jt[3 * j], 24); memcpy(dr.jR[j], &rb.jR[9 * j], 72); }
  double reach = 0;
  for (int j = 0; j < rb.n_joints; j++) reach += sqrt(rb.jt[3 * j] * rb.jt[3 * j] + rb.jt[3 * j + 1] * rb.jt[3 * j + 1] + rb.jt[3 * j + 2] * rb.jt[3 * j + 2]);
  for (int l = 0; l < rb.n_links; l++) {
    const HShape& s = rb.shapes[l + rb.n_strip];
    dr.link_jb[l] = s.joints_before; dr.link_shape[l] = l;
    memcpy(dr.lo_t[l], s.lo_t, 24); memcpy(dr.lo_R[l], s.lo_R, 72);
    DShape d; to_dshape(s, d, tb.verts_f, tb.verts_d, tb.hull_cells, tb.hull_cand);
    tb.shapes.push_back(d);
    if (s.type == SH_HULL) tb.has_hull = 1;
  }
  double lmax = 0;
  for (int l = 0; l < rb.n_links; l++) {
    const HShape& s = rb.shapes[l + rb.n_strip];
    lmax = fmax(lmax, sqrt(s.lo_t[0] * s.lo_t[0] + s.lo_t[1] * s.lo_t[1] + s.lo_t[2] * s.lo_t[2]) + sqrt(s.bc[0] * s.bc[0] + s.bc[1] * s.bc[1] + s.bc[2] * s.bc[2]) + s.br);
  }
  tb.extent = sqrt(rb.base_t[0] * rb.base_t[0] + rb.base_t[1] * rb.base_t[1] + rb.base_t[2] * rb.base_t[2]) + reach + lmax;

  // ---- obstacle shapes, then ALL shapes of every robot obstacle (QUIRK 7 needs the stripped ones too)
  for (const HObst& o : sc.obst) { DShape d; to_dshape(o.shape, d, tb.verts_f, tb.verts_d, tb.hull_cells, tb.hull_cand); tb.shapes.push_back(d); }
  const int rob_shape_base = (int)tb.shapes.size();
  int n_rob_shapes_all = 0, n_rob_links = 0;
  std::vector<int> rob_off;
  for (const HRobObst& ro : sc.robs) {
    rob_off.push_back(n_rob_shapes_all);
    for (const HShape& s : ro.robot.shapes) {
      DShape d; to_dshape(s, d, tb.verts_f, tb.verts_d, tb.hull_cells, tb.hull_cand); tb.shapes.push_back(d);
      if (s.type == SH_HULL) tb.has_hull = 1;
    }
    n_rob_shapes_all += ro.robot.n_shapes; n_rob_links += ro.robot.n_links;
  }
  if (tb.verts_f.empty()) { tb.verts_f.assign(3, 0.f); tb.verts_d.assign(3, 0.0); }
  if (tb.hull_cells.empty()) tb.hull_cells.assign(1, 0u);
  if (tb.hull_cand.empty()) tb.hull_cand.assign(1, 0);

  const int n_slots = sc.C + n_rob_links;
  tb.n_slots = n_slots; tb.n_robot_slots = n_rob_links;
  tb.slots.assign(n_slots, DSlot{});
  tb.pose_f.assign((size_t)n_slots * 12 * Tp + 1, 0.f);
  tb.bcw_f.assign((size_t)n_slots * 3 * Tp + 1, 0.f);
  tb.pose_d.assign((size_t)n_slots * Tp * 12 + 1, 0.0);
  tb.aabb_f.assign((size_t)n_slots * 6 * Tp + 1, 0.f);
  tb.aabb_d.assign((size_t)n_slots * Tp * 6 + 1, 0.0);
  tb.tight_f.assign((size_t)n_slots * 6 * Tp + 1, 0.f);
  // tight AABB gate table: outward-rounded floats (nextafter) so that the FP32 box always contains the FP64 one
  auto put_tight = [&](int slot, int f, const double* lo, const double* hi) {
    for (int k = 0; k < 3; k++) {
      tb.tight_f[((size_t)slot * 6 + k) * Tp + f] = nextafterf((float)lo[k], -3.0e38f);
      tb.tight_f[((size_t)slot * 6 + 3 + k) * Tp + f] = nextafterf((float)hi[k], 3.0e38f);
    }
  };
  std::vector<double> bcw_d((size_t)n_slots * Tp * 3 + 1, 0.0);

  auto put_pose = [&](int slot, int f, const HXf& x, const HShape& shape) {
    for (int k = 0; k < 9; k++) { tb.pose_f[((size_t)slot * 12 + k) * Tp + f] = (float)x.R[k]; tb.pose_d[((size_t)slot * Tp + f) * 12 + k] = x.R[k]; }
    for (int k = 0; k < 3; k++) { tb.pose_f[((size_t)slot * 12 + 9 + k) * Tp + f] = (float)x.t[k]; tb.pose_d[((size_t)slot * Tp + f) * 12 + 9 + k] = x.t[k]; }
    double c[3]; h_matvec(c, x.R, shape.bc);
    for (int k = 0; k < 3; k++) { bcw_d[((size_t)slot * Tp + f) * 3 + k] = c[k] + x.t[k]; tb.bcw_f[((size_t)slot * 3 + k) * Tp + f] = (float)(c[k] + x.t[k]); }
  };

  // ---- box / sphere obstacles (CollisionManager.cpp:60-84: float32 pose -> double)
  for (int s = 0; s < sc.C; s++) {
    const HObst& o = sc.obst[s];
    DSlot& S = tb.slots[s];
    S.shape = rb.n_links + s; S.shape_shift = -1; S.is_robot = 0; S.br = tb.shapes[S.shape].br;
    bool ident = true, same = true;
    for (int f = 0; f < Tp; f++) {
      const int fs = o.is_static ? 0 : (f < T ? f : T - 1);
      const float* p = &o.raw[7 * (size_t)fs];
      HXf x;
      for (int k = 0; k < 3; k++) x.t[k] = (double)p[k];
      h_quat_to_R(x.R, (double)p[3], (double)p[4], (double)p[5], (double)p[6]);
      for (int k = 0; k < 9; k++) ident &= x.R[k] == ((k % 4) == 0 ? 1.0 : 0.0);
      same &= !memcmp(p, &o.raw[0], 7 * sizeof(float));
      put_pose(s, f, x, o.shape);
      double tlo[3], thi[3];
      tight_aabb(o.shape, x, tlo, thi);
      put_tight(s, f, tlo, thi);
    }
    S.is_static = o.is_static || same;
    S.rot_identity = ident;
  }

  // ---- robot-obstacle links (CollisionManager.cpp:47-54,171-213,255-269)
  int slot = sc.C, k_ref = 0;
  for (size_t r = 0; r < sc.robs.size(); r++) {
    const HRobObst& ro = sc.robs[r];
    const int nl = ro.robot.n_links;
    if (ro.robot.dof > 64 || nl > 64) { err = "robot obstacle too large"; return -4; }
    std::vector<HXf> xf((size_t)Tp * (nl > 0 ? nl : 1));
    double q[64];
    for (int f = 0; f < Tp; f++) {
      robobst_config_at_frame(sc, ro, f < T ? f : T - 1, q);
      host_fk(ro.robot, q, &xf[(size_t)f * nl]);
    }
    for (int j = 0; j < nl; j++, slot++, k_ref++) {
      DSlot& S = tb.slots[slot];
      const HShape& st = ro.robot.shapes[j + ro.robot.n_strip];
      S.shape = rob_shape_base + rob_off[r] + ro.robot.n_strip + j;
      S.shape_shift = (parity_mode == 0 && k_ref < n_rob_shapes_all) ? rob_shape_base + k_ref : -1;
      S.is_robot = 1; S.rot_identity = 0;
      double br = st.br;
      if (S.shape_shift >= 0) {  // the functor geometry shares the link's transform but may reach further
        const HRobObst* owner = nullptr; int kk = k_ref;
        for (const HRobObst& o2 : sc.robs) { if (kk < o2.robot.n_shapes) { owner = &o2; break; } kk -= o2.robot.n_shapes; }
        const HShape& ss = owner->robot.shapes[kk];
        double d = 0;
        for (int k = 0; k < 3; k++) d += (ss.bc[k] - st.bc[k]) * (ss.bc[k] - st.bc[k]);
        br = fmax(br, sqrt(d) + ss.br);
      }
      S.br = (float)(br * (1 + 1e-6));
      bool same = true;
      for (int f = 0; f < Tp; f++) {
        const HXf& x = xf[(size_t)f * nl + j];
        same &= !memcmp(&x, &xf[j], sizeof(HXf));
        put_pose(slot, f, x, st);
        double lo[3], hi[3];
        tight_aabb(st, x, lo, hi);
        if (S.shape_shift >= 0) {  // the gate must also admit every pair the shifted functor geometry can hit
          const HRobObst* owner = nullptr; int kk = k_ref;
          for (const HRobObst& o2 : sc.robs) { if (kk < o2.robot.n_shapes) { owner = &o2; break; } kk -= o2.robot.n_shapes; }
          double l2[3], h2[3];
          tight_aabb(owner->robot.shapes[kk], x, l2, h2);
          for (int k = 0; k < 3; k++) { lo[k] = fmin(lo[k], l2[k]); hi[k] = fmax(hi[k], h2[k]); }
        }
        put_tight(slot, f, lo, hi);
        coal_aabb(st, x, lo, hi);
        for (int k = 0; k < 3; k++) {
          tb.aabb_f[((size_t)slot * 6 + k) * Tp + f] = (float)lo[k]; tb.aabb_f[((size_t)slot * 6 + 3 + k) * Tp + f] = (float)hi[k];
          tb.aabb_d[((size_t)slot * Tp + f) * 6 + k] = lo[k]; tb.aabb_d[((size_t)slot * Tp + f) * 6 + 3 + k] = hi[k];
        }
      }
      S.is_static = same;  // the reference treats static_robot and dynamic_robot alike (CollisionManager.cpp:171-213)
    }
  }

  for (int s = 0; s < n_slots; s++) {
    DSlot& S = tb.slots[s];
    S.type = tb.shapes[S.shape].type;
    S.type_shift = S.shape_shift >= 0 ? tb.shapes[S.shape_shift].type : -1;
  }
  for (int s = 0; s < n_slots; s++) (tb.slots[s].is_static ? tb.static_slots : tb.dynamic_slots).push_back(s);
  const int nd = (int)tb.dynamic_slots.size();
  tb.n_dynamic = nd; tb.n_static = (int)tb.static_slots.size();

  // ---- swept bounding sphere of every dynamic slot over every 32-frame tile
  tb.tile_bounds.assign((size_t)n_tiles * (nd > 0 ? nd : 1), make_float4(0, 0, 0, -1));
  for (int tile = 0; tile < n_tiles; tile++)
    for (int j = 0; j < nd; j++) {
      const int s = tb.dynamic_slots[j];
      double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
      const int f1 = tile * 32 + 32 < T ? tile * 32 + 32 : T;
      for (int f = tile * 32; f < f1; f++)
        for (int k = 0; k < 3; k++) { const double c = bcw_d[((size_t)s * Tp + f) * 3 + k]; lo[k] = fmin(lo[k], c); hi[k] = fmax(hi[k], c); }
      double c[3], r2 = 0;
      for (int k = 0; k < 3; k++) c[k] = 0.5 * (lo[k] + hi[k]);
      for (int f = tile * 32; f < f1; f++) {
        double d = 0;
        for (int k = 0; k < 3; k++) { const double e = bcw_d[((size_t)s * Tp + f) * 3 + k] - c[k]; d += e * e; }
        r2 = fmax(r2, d);
      }
      const double rad = (sqrt(r2) + (double)tb.slots[s].br) * (1 + 1e-6) + 1e-5;
      tb.tile_bounds[(size_t)tile * nd + j] = make_float4((float)c[0], (float)c[1], (float)c[2], (float)rad);
    }
  // ---- inline support-candidate records (FP32)
  {
    float4 empty = make_float4(0.f, 0.f, 0.f, 0.f), over = empty;
    const int m1 = -1, m2 = -2;
    memcpy(&empty.w, &m1, 4); memcpy(&over.w, &m2, 4);
    tb.hull_rec.assign(tb.hull_cells.size() * SIRRT_HULL_REC, empty);
    for (const DShape& d : tb.shapes) {
      if (d.type != SH_HULL || d.cell_off < 0) continue;
      for (int cell = 0; cell < SIRRT_HULL_CELLS; cell++) {
        const unsigned c0 = tb.hull_cells[d.cell_off + cell], c1 = tb.hull_cells[d.cell_off + cell + 1];
        float4* rec = &tb.hull_rec[((size_t)d.cell_off + cell) * SIRRT_HULL_REC];
        if (c1 - c0 > SIRRT_HULL_REC) { rec[SIRRT_HULL_REC - 1] = over; continue; }
        for (unsigned k = c0; k < c1; k++) {
          const int i = tb.hull_cand[k];
          float4 r = make_float4(tb.verts_f[3 * ((size_t)d.vert_off + i)], tb.verts_f[3 * ((size_t)d.vert_off + i) + 1], tb.verts_f[3 * ((size_t)d.vert_off + i) + 2], 0.f);
          memcpy(&r.w, &i, 4);
          rec[k - c0] = r;
        }
      }
    }
  }
  if (tb.static_slots.empty()) tb.static_slots.push_back(0);  // keep the device arrays non-empty
  if (tb.dynamic_slots.empty()) tb.dynamic_slots.push_back(0);
  return 0;
}

}  // namespace sirrt
