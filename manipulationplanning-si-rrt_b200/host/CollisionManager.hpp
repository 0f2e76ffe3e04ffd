// CollisionManager.hpp -- header-only C++ shim with the public surface of the reference's
// MDP::CollisionManager (STRRT_Planner/include/CollisionManager/CollisionManager.hpp:71-145) on top
// of the C ABI of libsirrt_b200.so (include/sirrt_b200.h).  PlannerConnect-style callers
// (MSIRRT/src/PlannerConnect.cpp:20,28,47,131,544,751,859), the OMPL validators
// (StateValidityCheckerFunctor.cpp:46,64) and check_scene.cpp:74,83 compile against it unchanged
// apart from the constructor argument: the scene arrives as a scene pack (include/sirrt_pack.h)
// instead of a parsed MDP::ConfigReader::SceneTask.
//
// Semantics kept: value type, copy deleted / move allowed (:76-79); results by value; exceptions
// (std::runtime_error) for bad scenes and device failures; not thread-safe (one instance per thread).
// There is no CPU fallback: construction throws when no CUDA device is usable.
#pragma once
#include <array>
#include <fstream>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "sirrt_b200.h"

namespace MDP {

#ifdef EIGEN_CORE_H
using Vector3d = Eigen::Vector3d;
#else
struct Vector3d { double x = 0, y = 0, z = 0; };
#endif

class CollisionManager {
 public:
  explicit CollisionManager(const std::vector<unsigned char>& scene_pack, int device = 0, int parity_mode = SIRRT_PARITY_REFERENCE_EXACT) {
    if (sirrt_scene_create_from_pack(scene_pack.data(), scene_pack.size(), device, parity_mode, &scene_) != SIRRT_OK)
      throw std::runtime_error(std::string("CollisionManager: ") + sirrt_last_error());
    sirrt_scene_set_option(scene_, SIRRT_OPT_HOST_BUFFERS, 1);  // this class and its users (B200Backend, the validators) pass host vectors only
    sirrt_scene_info(scene_, &dof_, &n_links_, &frame_count_, &fps_, nullptr);
    limits_.resize(2 * (size_t)dof_);
    sirrt_scene_info(scene_, nullptr, nullptr, nullptr, nullptr, limits_.data());
    start_.resize(dof_); end_.resize(dof_); max_vel_.resize(dof_);
    sirrt_scene_task(scene_, start_.data(), end_.data(), max_vel_.data());
  }
  static CollisionManager from_pack_file(const std::string& path, int device = 0, int parity_mode = SIRRT_PARITY_REFERENCE_EXACT) {
    std::ifstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("CollisionManager: cannot open " + path);
    std::vector<unsigned char> b((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    return CollisionManager(b, device, parity_mode);
  }
  ~CollisionManager() { if (scene_) sirrt_scene_destroy(scene_); }
  CollisionManager(const CollisionManager&) = delete;
  CollisionManager& operator=(const CollisionManager&) = delete;
  CollisionManager(CollisionManager&& o) noexcept { *this = std::move(o); }
  CollisionManager& operator=(CollisionManager&& o) noexcept {
    if (this != &o) {
      if (scene_) sirrt_scene_destroy(scene_);
      scene_ = o.scene_; o.scene_ = nullptr;
      dof_ = o.dof_; n_links_ = o.n_links_; frame_count_ = o.frame_count_; fps_ = o.fps_;
      limits_ = std::move(o.limits_); start_ = std::move(o.start_); end_ = std::move(o.end_); max_vel_ = std::move(o.max_vel_);
    }
    return *this;
  }

  struct robot_link_distance_to_obj {  // CollisionManager.hpp:91-99
    float min_distance;
    std::vector<Vector3d> neares_point_robot;
    std::vector<Vector3d> neares_point_obj;
  };

  // CollisionManager.cpp:427-443
  int get_goal_frame_low_bound() {
    for (int frame = frame_count_ - 1; frame >= 0; frame--) {
      if (check_collision_frame_no_wrapper(end_, frame)) {
        if (frame == frame_count_ - 1) throw std::runtime_error("there is collision in goal state at the last frame!");
        return frame + 1;
      }
    }
    return 0;
  }
  // CollisionManager.cpp:330-336: frame = (int)(time * fps) with a float32 product (QUIRK 13)
  bool check_collision(const std::vector<double>& robot_angles, float& time) {
    int frame = (int)(time * fps_);
    return check_collision_frame(robot_angles, frame);
  }
  // CollisionManager.cpp:338-341
  bool check_collision_frame(const std::vector<double>& robot_angles, int& frame) { return check_collision_frame_no_wrapper(robot_angles, frame); }
  bool check_collision_frame_no_wrapper(std::vector<double> robot_angles, int frame) {
    need_dof(robot_angles);
    int collided = 0;
    ck(sirrt_check_frame(scene_, robot_angles.data(), frame, &collided));
    return collided != 0;
  }
  // CollisionManager.cpp:382-420 + DistanceCallback :549-600.  Per link: min_distance and the closest point pairs
  // (neares_point_robot / neares_point_obj, coal nearest_points[0] / [1]).  Two documented deviations:
  // the reference's callback overwrites `dist` with the LAST pair its AABB-tree traversal visited and pushes one point pair per
  // visited leaf in traversal order; this backend reports the true minimum over all box / sphere obstacles of the frame
  // and returns the point pairs of ALL of them in obstacle order (the visited ones are a subset).
  std::vector<robot_link_distance_to_obj> get_distances(std::vector<double> robot_angles, float time, bool& is_collision) {
    need_dof(robot_angles);
    int n_obs = 0;
    ck(sirrt_scene_counts(scene_, &n_obs, nullptr));
    std::vector<double> d(n_links_), pd((size_t)n_links_ * n_obs), pts((size_t)n_links_ * n_obs * 6);
    int col = 0;
    if (n_obs > 0) ck(sirrt_get_distances_ex(scene_, robot_angles.data(), (int)(time * fps_), d.data(), &col, pd.data(), pts.data()));
    else ck(sirrt_get_distances(scene_, robot_angles.data(), (int)(time * fps_), d.data(), &col));
    is_collision = col != 0;
    std::vector<robot_link_distance_to_obj> out;
    for (int l = 0; l < n_links_; l++) {
      if (d[l] - 1e-21 <= 1e-12) break;  // the reference returns what it had collected when a link collides (:409-413)
      robot_link_distance_to_obj r{(float)d[l], {}, {}};
      for (int c = 0; c < n_obs; c++) {
        const double* p = &pts[((size_t)l * n_obs + c) * 6];
        r.neares_point_robot.push_back(Vector3d{p[0], p[1], p[2]});
        r.neares_point_obj.push_back(Vector3d{p[3], p[4], p[5]});
      }
      out.push_back(std::move(r));
    }
    return out;
  }
  bool check_robot_selfcollision(std::vector<double>) { return false; }     // CollisionManager.cpp:445-447
  bool check_base_joints_collisiion(std::vector<double>, int) { return false; }  // :476-479
  // CollisionManager.cpp:481-547
  std::vector<std::pair<int, int>> get_safe_intervals(const std::vector<double>& robot_angles) {
    need_dof(robot_angles);
    int cap = 64, n = 0;
    std::vector<int32_t> buf;
    for (;;) {
      buf.resize(2 * (size_t)cap);
      const int rc = sirrt_safe_intervals(scene_, robot_angles.data(), buf.data(), cap, &n);
      if (rc == SIRRT_OK) break;
      if (rc == SIRRT_E_OVERFLOW && n > cap) { cap = n; continue; }
      ck(rc);
    }
    std::vector<std::pair<int, int>> out;
    out.reserve(n);
    for (int i = 0; i < n; i++) out.emplace_back(buf[2 * i], buf[2 * i + 1]);
    return out;
  }
  // CollisionManager.cpp:422-425 (URDF limits narrowed to float, RobotObstacleFCL.cpp:82)
  std::vector<std::pair<float, float>> get_planned_robot_limits() const {
    std::vector<std::pair<float, float>> out;
    for (int j = 0; j < dof_; j++) out.emplace_back(limits_[2 * j], limits_[2 * j + 1]);
    return out;
  }

  // ---- batched extensions (not in the reference; what a GPU-aware planner calls) -----------------
  // bits: [n][words()] collision rows of n configurations given as double AoS [n][dof]
  void collision_bits(const std::vector<double>& q_aos, int semantics, std::vector<uint32_t>& bits) {
    const size_t n = q_aos.size() / dof_;
    bits.assign(n * words(), 0u);
    if (n) ck(sirrt_collision_bits_f64(scene_, q_aos.data(), (int64_t)n, semantics, bits.data()));
  }
  // PlannerConnect::is_collision_motion for many edges at once (PlannerConnect.cpp:862-897)
  void edge_check(const std::vector<double>& q0, const std::vector<double>& q1, const std::vector<double>& t0, const std::vector<double>& t1,
                  double vmax, std::vector<int32_t>& first_colliding_step, std::vector<double>* collision_coord = nullptr) {
    const size_t n = t0.size();
    first_colliding_step.assign(n, -1);
    if (collision_coord) collision_coord->assign(n * dof_, 0.0);
    if (n) ck(sirrt_edge_check_batch(scene_, q0.data(), q1.data(), t0.data(), t1.data(), (int64_t)n, vmax, first_colliding_step.data(),
                                     collision_coord ? collision_coord->data() : nullptr));
  }

  int dof() const { return dof_; }
  int link_count() const { return n_links_; }  // collision links of the planned robot (collision_robot_links_count)
  int frame_count() const { return frame_count_; }
  int fps() const { return fps_; }
  int words() const { return (frame_count_ + 31) / 32; }
  const std::vector<double>& start_configuration() const { return start_; }
  const std::vector<double>& end_configuration() const { return end_; }
  const std::vector<double>& robot_joint_max_velocity() const { return max_vel_; }
  sirrt_scene* handle() { return scene_; }

 private:
  void ck(int rc) const { if (rc != SIRRT_OK) throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error()); }
  void need_dof(const std::vector<double>& q) const { if ((int)q.size() < dof_) throw std::runtime_error("CollisionManager: configuration has fewer angles than the robot has joints"); }
  sirrt_scene* scene_ = nullptr;
  int dof_ = 0, n_links_ = 0, frame_count_ = 0, fps_ = 0;
  std::vector<float> limits_;
  std::vector<double> start_, end_, max_vel_;
};

}  // namespace MDP
