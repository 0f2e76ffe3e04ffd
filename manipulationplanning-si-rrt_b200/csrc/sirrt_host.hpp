// sirrt_host.hpp -- host side of libsirrt_b200: scene-pack parsing and construction of the device
// tables.  Replaces what MDP::CollisionManager's constructor builds (CollisionManager.cpp:19-280):
// instead of per-frame and cross-frame AABB trees it produces flat, frame-major (SoA over frames)
// pose tables plus swept bounding spheres per 32-frame tile, sized to live in L2/HBM.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#include <string>
#include <vector>

#include "../../include/sirrt_pack.h"
#include "sirrt_device.cuh"

namespace sirrt {

struct HXf { double R[9], t[3]; };

struct HShape {
  int type = SH_CAPSULE;  // SH_*
  int joints_before = 0, n_verts = 0;
  double lo_t[3] = {0, 0, 0}, lo_R[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  double radius = 0, half_len = 0, half[3] = {0, 0, 0};
  std::vector<float> verts;  // hull vertices (float32 as stored in the pack), xyz interleaved
  double bc[3], br;          // bounding sphere of the inflated shape, local frame
  double amin[3], amax[3], ac[3], ar;  // coal computeLocalAABB: aabb_local / aabb_center / aabb_radius
};

struct HRobot {
  int dof = 0, n_joints = 0, n_shapes = 0, n_strip = 0, n_links = 0;
  double base_t[3], base_R[9];
  std::vector<float> limits;
  std::vector<double> jt, jR;
  std::vector<HShape> shapes;
};

struct HObst {
  int type = 0, is_static = 0, n_poses = 0;
  HShape shape;
  std::vector<float> raw;  // n_poses x 7
};

struct HRobObst {
  HRobot robot;
  int is_static = 0, n_states = 0;
  std::vector<double> times, angles;
};

struct HScene {
  int T = 0, fps = 0, dof = 0, C = 0, R = 0;
  double start_time = 0;
  std::vector<double> start_cfg, end_cfg, max_vel;
  HRobot robot;
  std::vector<HObst> obst;
  std::vector<HRobObst> robs;
};

// ---------------------------------------------------------------------------------------------
inline void h_matvec(double* o, const double* R, const double* v) {
  o[0] = R[0] * v[0] + R[1] * v[1] + R[2] * v[2];
  o[1] = R[3] * v[0] + R[4] * v[1] + R[5] * v[2];
  o[2] = R[6] * v[0] + R[7] * v[1] + R[8] * v[2];
}
inline void h_matmul(double* o, const double* A, const double* B) {
  double r[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) r[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
  memcpy(o, r, sizeof r);
}
// Eigen::QuaternionBase::toRotationMatrix without normalisation (CollisionManager.cpp:72,132,233;
// RobotObstacleFCL.cpp:302)
inline void h_quat_to_R(double* R, double x, double y, double z, double w) {
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

// local bounding sphere + coal local AABB of a shape from its core (segment / vertices / box)
inline void shape_finish(HShape& s) {
  double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
  std::vector<double> core;
  if (s.type == SH_CAPSULE) core = {0, 0, -s.half_len, 0, 0, s.half_len};
  else if (s.type == SH_SPHERE) core = {0, 0, 0};
  else if (s.type == SH_BOX) {
    for (int sx = -1; sx <= 1; sx += 2) for (int sy = -1; sy <= 1; sy += 2) for (int sz = -1; sz <= 1; sz += 2) {
      core.push_back(sx * s.half[0]); core.push_back(sy * s.half[1]); core.push_back(sz * s.half[2]);
    }
  } else for (float v : s.verts) core.push_back((double)v);
  const size_t n = core.size() / 3;
  for (size_t i = 0; i < n; i++)
    for (int k = 0; k < 3; k++) { lo[k] = fmin(lo[k], core[3 * i + k]); hi[k] = fmax(hi[k], core[3 * i + k]); }
  if (!n) for (int k = 0; k < 3; k++) lo[k] = hi[k] = 0;
  const double infl = (s.type == SH_CAPSULE || s.type == SH_SPHERE) ? s.radius : 0.0;
  for (int k = 0; k < 3; k++) s.bc[k] = 0.5 * (lo[k] + hi[k]);
  double r2 = 0;
  for (size_t i = 0; i < n; i++) {
    double d = 0;
    for (int k = 0; k < 3; k++) d += (core[3 * i + k] - s.bc[k]) * (core[3 * i + k] - s.bc[k]);
    r2 = fmax(r2, d);
  }
  s.br = sqrt(r2) + infl;
  for (int k = 0; k < 3; k++) { s.amin[k] = lo[k] - infl; s.amax[k] = hi[k] + infl; s.ac[k] = 0.5 * (s.amin[k] + s.amax[k]); }
  double d = 0;
  for (int k = 0; k < 3; k++) d += (s.amin[k] - s.ac[k]) * (s.amin[k] - s.ac[k]);
  s.ar = sqrt(d);
}

// a1 on the host (FP64): RobotObstacleFCL::get_collision_object_for_robot_angles_private,
// RobotObstacleFCL.cpp:287-386.  Used at scene-build time for robot obstacles (a7) and by
// sirrt_scene_fk.  out[] receives the base-stripped links (:370).
inline void host_fk(const HRobot& r, const double* q, HXf* out) {
  double t[3] = {r.base_t[0], r.base_t[1], r.base_t[2]}, R[9];
  memcpy(R, r.base_R, sizeof R);
  int s = 0;
  for (int k = 0; k <= r.n_joints; k++) {
    for (; s < r.n_shapes && r.shapes[s].joints_before == k; s++) {
      if (s < r.n_strip) continue;
      HXf& o = out[s - r.n_strip];
      double w[3];
      h_matvec(w, R, r.shapes[s].lo_t);  // :335
      for (int i = 0; i < 3; i++) o.t[i] = t[i] + w[i];  // :336
      h_matmul(o.R, R, r.shapes[s].lo_R);  // :347
    }
    if (k == r.n_joints) break;
    double w[3];
    h_matvec(w, R, &r.jt[3 * k]);  // :319,361
    for (int i = 0; i < 3; i++) t[i] += w[i];
    h_matmul(R, R, &r.jR[9 * k]);  // :320,362
    const double a = k < r.dof ? q[k] : 0.0, sn = sin(a), cs = cos(a);
    const double Rz[9] = {cs, -sn, 0, sn, cs, 0, 0, 0, (1.0 - cs) * 1.0 + cs};  // AngleAxisd(a, UnitZ) :321,363
    h_matmul(R, R, Rz);
  }
}

// a7: RobotObstacleFCL::get_collision_object_for_frame, RobotObstacleFCL.cpp:393-429
inline void robobst_config_at_frame(const HScene& sc, const HRobObst& ro, int frame, double* q) {
  const int dof = ro.robot.dof, n = ro.n_states;
  const double time = (double)frame / (double)sc.fps + sc.start_time;  // :399
  if (n <= 0) { for (int i = 0; i < dof; i++) q[i] = 0; return; }
  if (ro.times[n - 1] <= time + 0.00001) { memcpy(q, &ro.angles[(size_t)(n - 1) * dof], sizeof(double) * dof); return; }
  for (int i = 0; i + 1 < n; i++) {
    if (ro.times[i] <= time + 0.00001 && ro.times[i + 1] > time + 0.00001) {
      const double *l = &ro.angles[(size_t)i * dof], *r = &ro.angles[(size_t)(i + 1) * dof];
      const double lt = ro.times[i], rt = ro.times[i + 1];
      for (int a = 0; a < dof; a++) {
        const double k = (r[a] - l[a]) / (rt - lt);
        const double b = r[a] - (r[a] - l[a]) / (rt - lt) * rt;
        q[a] = k * time + b;  // :417-419
      }
      return;
    }
  }
  // before the first state the reference runs into assert(false) (:429; UB in Release): hold state 0
  memcpy(q, &ro.angles[0], sizeof(double) * dof);
}

// ---------------------------------------------------------------------------------------------
// pack reader
// ---------------------------------------------------------------------------------------------
struct Cursor {
  const unsigned char* p; size_t n, o = 0; bool bad = false;
  const void* take(size_t bytes) {
    if (bad || o + bytes > n) { bad = true; return nullptr; }
    const void* r = p + o; o += bytes; return r;
  }
  void pad8() { o = (o + 7) & ~(size_t)7; }
};

inline bool read_robot(Cursor& c, HRobot& r) {
  auto* h = (const SirrtRobotHeader*)c.take(sizeof(SirrtRobotHeader));
  if (!h) return false;
  if (h->dof < 0 || h->dof > 64 || h->n_joints < 0 || h->n_joints > 64 || h->n_shapes < 0 || h->n_shapes > 256 || h->n_strip < 0) return false;
  r.dof = h->dof; r.n_joints = h->n_joints; r.n_shapes = h->n_shapes; r.n_strip = h->n_strip;
  r.n_links = r.n_shapes > r.n_strip ? r.n_shapes - r.n_strip : 0;
  for (int i = 0; i < 3; i++) r.base_t[i] = (double)h->base_pos[i];
  h_quat_to_R(r.base_R, (double)h->base_quat[0], (double)h->base_quat[1], (double)h->base_quat[2], (double)h->base_quat[3]);
  auto* lim = (const float*)c.take(sizeof(float) * 2 * r.dof);
  if (r.dof && !lim) return false;
  r.limits.assign(lim, lim + 2 * r.dof);
  c.pad8();
  r.jt.resize(3 * (size_t)r.n_joints); r.jR.resize(9 * (size_t)r.n_joints);
  for (int j = 0; j < r.n_joints; j++) {
    auto* jr = (const SirrtJointRecord*)c.take(sizeof(SirrtJointRecord));
    if (!jr) return false;
    memcpy(&r.jt[3 * j], jr->t, 24); memcpy(&r.jR[9 * j], jr->R, 72);
  }
  r.shapes.resize(r.n_shapes);
  for (int s = 0; s < r.n_shapes; s++) {
    auto* sr = (const SirrtShapeRecord*)c.take(sizeof(SirrtShapeRecord));
    if (!sr || sr->n_verts < 0) return false;
    HShape& o = r.shapes[s];
    o.joints_before = sr->joints_before; o.n_verts = sr->n_verts;
    memcpy(o.lo_t, sr->lo_t, 24); memcpy(o.lo_R, sr->lo_R, 72);
    auto* v = (const float*)c.take(sizeof(float) * 3 * (size_t)sr->n_verts);
    if (sr->n_verts && !v) return false;
    c.pad8();
    if (sr->type == SIRRT_SHAPE_CAPSULE) {
      o.type = SH_CAPSULE; o.radius = sr->radius; o.half_len = sr->length / 2;  // coal::Capsule(r, lz): halfLength = lz/2
      o.n_verts = 0;
    } else if (sr->type == SIRRT_SHAPE_HULL) {
      if (sr->n_verts < 1 || sr->n_verts >= 4096) return false;
      o.type = SH_HULL; o.verts.assign(v, v + 3 * (size_t)sr->n_verts);
    } else return false;
    shape_finish(o);
  }
  return !c.bad;
}

inline bool parse_pack(const void* pack, size_t bytes, HScene& sc, std::string& err) {
  Cursor c{(const unsigned char*)pack, bytes};
  auto* h = (const SirrtPackHeader*)c.take(sizeof(SirrtPackHeader));
  if (!h || memcmp(h->magic, SIRRT_PACK_MAGIC, 8) || h->version != SIRRT_PACK_VERSION || h->total_bytes != bytes) {
    err = "bad scene pack header"; return false;
  }
  if (h->n_frames < 1 || h->fps < 1 || h->dof < 1 || h->dof > 64 || h->n_obstacles < 0 || h->n_robot_obstacles < 0) {
    err = "scene pack header out of range"; return false;
  }
  sc.T = h->n_frames; sc.fps = h->fps; sc.start_time = h->start_time; sc.dof = h->dof;
  sc.C = h->n_obstacles; sc.R = h->n_robot_obstacles;
  for (std::vector<double>* v : {&sc.start_cfg, &sc.end_cfg, &sc.max_vel}) {
    auto* p = (const double*)c.take(sizeof(double) * sc.dof);
    if (!p) { err = "truncated scene pack"; return false; }
    v->assign(p, p + sc.dof);
  }
  if (!read_robot(c, sc.robot)) { err = "malformed planned-robot record"; return false; }
  sc.obst.resize(sc.C);
  for (int i = 0; i < sc.C; i++) {
    auto* r = (const SirrtObstacleRecord*)c.take(sizeof(SirrtObstacleRecord));
    if (!r) { err = "truncated obstacle record"; return false; }
    HObst& b = sc.obst[i];
    b.type = r->type; b.is_static = r->is_static; b.n_poses = r->n_poses;
    if (b.n_poses < 1 || (!b.is_static && b.n_poses < sc.T)) { err = "obstacle has fewer poses than frames"; return false; }
    if (b.type == SIRRT_OBS_BOX) {
      b.shape.type = SH_BOX;
      for (int k = 0; k < 3; k++) b.shape.half[k] = (double)r->dims[k] / 2;  // coal::Box(x,y,z): halfSide = side/2
    } else if (b.type == SIRRT_OBS_SPHERE) {
      b.shape.type = SH_SPHERE; b.shape.radius = (double)r->dims[0];
    } else { err = "unknown obstacle type"; return false; }
    shape_finish(b.shape);
    auto* ps = (const float*)c.take(sizeof(float) * 7 * (size_t)b.n_poses);
    if (!ps) { err = "truncated obstacle poses"; return false; }
    c.pad8();
    b.raw.assign(ps, ps + 7 * (size_t)b.n_poses);
  }
  sc.robs.resize(sc.R);
  for (int i = 0; i < sc.R; i++) {
    auto* rh = (const SirrtRobotObstacleHeader*)c.take(sizeof(SirrtRobotObstacleHeader));
    if (!rh || rh->n_states < 0) { err = "truncated robot-obstacle record"; return false; }
    HRobObst& ro = sc.robs[i];
    ro.is_static = rh->is_static; ro.n_states = rh->n_states;
    if (!read_robot(c, ro.robot)) { err = "malformed robot-obstacle record"; return false; }
    auto* tm = (const double*)c.take(sizeof(double) * (size_t)ro.n_states);
    auto* an = (const double*)c.take(sizeof(double) * (size_t)ro.n_states * ro.robot.dof);
    if (ro.n_states && (!tm || !an)) { err = "truncated robot-obstacle trajectory"; return false; }
    if (ro.n_states) { ro.times.assign(tm, tm + ro.n_states); ro.angles.assign(an, an + (size_t)ro.n_states * ro.robot.dof); }
  }
  if (c.bad || c.o != bytes) { err = "truncated or oversized scene pack"; return false; }
  return true;
}

}  // namespace sirrt
