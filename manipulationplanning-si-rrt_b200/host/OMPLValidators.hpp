// OMPLValidators.hpp -- the ST-RRT* checker hooks of the reference on the B200 backend:
//   MDP::StateValidityCheckerFunctor  (STRRT_Planner/include/OMPLValidators/StateValidityCheckerFunctor.hpp:27-46,
//                                      src/OMPLValidators/StateValidityCheckerFunctor.cpp:23-80)
//   MDP::SpaceTimeMotionValidator     (include/OMPLValidators/SpaceTimeMotionValidator.hpp:26-43,
//                                      src/OMPLValidators/SpaceTimeMotionValidator.cpp:18-132)
// Same class names, method names, argument meaning and counters.  OMPL is not installed in this image, so the
// classes are written against a plain space-time state (MDP::SpaceTimeState = joint values + time in seconds,
// the two components the reference reads out of ob::CompoundState, StateValidityCheckerFunctor.cpp:35,39).
// When the OMPL headers are present (checked with __has_include) thin ob::StateValidityChecker /
// ob::MotionValidator subclasses forwarding to the same code are compiled as well (MDP::ompl_adapters).
//
// checkMotion does not loop over isValid on the host: the s2 test and all interpolated states of a motion
// (or of many motions: checkMotions) go to the GPU in ONE call, sirrt_motion_check_batch, which builds the
// (configuration, frame) cells on the device with the reference's arithmetic and returns the first invalid one.
#pragma once
#include <algorithm>
#include <cmath>
#include <utility>
#include <vector>

#include "CollisionManager.hpp"

namespace MDP {

struct SpaceTimeState {
  std::vector<double> values;  // ob::RealVectorStateSpace::StateType(0)->values
  double position = 0.0;       // ob::TimeStateSpace::StateType(1)->position, seconds
};

class StateValidityCheckerFunctor {
 public:
  // the reference builds its own CollisionManager from the SceneTask (StateValidityCheckerFunctor.cpp:23-27);
  // here the scene arrives as a pack (see CollisionManager.hpp)
  explicit StateValidityCheckerFunctor(const std::vector<unsigned char>& scene_pack, int device = 0,
                                       int parity_mode = SIRRT_PARITY_REFERENCE_EXACT)
      : collision_manager(scene_pack, device, parity_mode) {}

  // StateValidityCheckerFunctor.cpp:29-47
  bool isValid(const SpaceTimeState* state) const {
    validation_counter++;
    float t = (float)state->position;
    std::vector<double> pos(state->values.begin(), state->values.begin() + collision_manager.dof());
    return !collision_manager.check_collision(pos, t);
  }
  // StateValidityCheckerFunctor.cpp:49-70: smallest per-link distance.  The reference dereferences
  // min_element of an empty vector when the state collides (undefined); here a colliding state answers 0.
  double clearance(const SpaceTimeState* state) const {
    const double t = state->position;
    std::vector<double> pos(state->values.begin(), state->values.begin() + collision_manager.dof());
    bool is_collision = false;
    auto distances = collision_manager.get_distances(pos, (float)t, is_collision);
    if (is_collision || distances.empty()) return 0.0;
    auto min_distance = [](const auto& lhs, const auto& rhs) -> bool { return lhs.min_distance < rhs.min_distance; };
    return std::min_element(distances.begin(), distances.end(), min_distance)->min_distance;
  }
  int get_count_of_validation() const { return validation_counter; }
  std::vector<std::pair<float, float>> get_planned_robot_limits() const { return collision_manager.get_planned_robot_limits(); }

  CollisionManager& manager() const { return collision_manager; }
  void count_validations(int n) const { validation_counter += n; }

 private:
  mutable CollisionManager collision_manager;
  mutable int validation_counter = 0;
};

class SpaceTimeMotionValidator {
 public:
  // the reference takes the SpaceInformationPtr whose isValid is the functor above (SpaceTimeMotionValidator.cpp:18-20)
  explicit SpaceTimeMotionValidator(const StateValidityCheckerFunctor& si) : si(si) {}

  // SpaceTimeMotionValidator.cpp:22-74
  bool checkMotion(const SpaceTimeState* s1, const SpaceTimeState* s2) const {
    int32_t r; double f;
    run(s1, s2, 1, 0, &r, &f);
    return r == -1;
  }
  // SpaceTimeMotionValidator.cpp:76-127: lastValid.first = last valid interpolated state (s1 when none), .second its fraction
  bool checkMotion(const SpaceTimeState* s1, const SpaceTimeState* s2, std::pair<SpaceTimeState*, double>& lastValid) const {
    int32_t r; double f;
    run(s1, s2, 1, 1, &r, &f);
    lastValid.second = f;
    if (lastValid.first) {
      const int dof = si.manager().dof();
      lastValid.first->values.resize(s1->values.size());
      for (int i = 0; i < dof; i++) lastValid.first->values[i] = s1->values[i] + (s2->values[i] - s1->values[i]) * f;
      lastValid.first->position = s1->position + (s2->position - s1->position) * f;
    }
    return r == -1;
  }
  // batched form: n motions in one GPU call; result codes of sirrt_motion_check_batch
  void checkMotions(const std::vector<SpaceTimeState>& from, const std::vector<SpaceTimeState>& to, int variant,
                    std::vector<int32_t>& result, std::vector<double>* last_valid_fraction = nullptr) const {
    const size_t n = from.size();
    result.assign(n, -1);
    if (last_valid_fraction) last_valid_fraction->assign(n, 0.0);
    if (!n) return;
    const int dof = si.manager().dof();
    std::vector<double> q0(n * dof), q1(n * dof), t0(n), t1(n);
    for (size_t e = 0; e < n; e++) {
      for (int i = 0; i < dof; i++) { q0[e * dof + i] = from[e].values[i]; q1[e * dof + i] = to[e].values[i]; }
      t0[e] = from[e].position; t1[e] = to[e].position;
    }
    call(q0.data(), q1.data(), t0.data(), t1.data(), n, variant, result.data(), last_valid_fraction ? last_valid_fraction->data() : nullptr);
  }
  int get_count_of_validation() const { return validation_counter; }

 private:
  void run(const SpaceTimeState* s1, const SpaceTimeState* s2, size_t n, int variant, int32_t* r, double* f) const {
    call(s1->values.data(), s2->values.data(), &s1->position, &s2->position, n, variant, r, f);
  }
  void call(const double* q0, const double* q1, const double* t0, const double* t1, size_t n, int variant, int32_t* r, double* f) const {
    validation_counter += (int)n;
    CollisionManager& cm = si.manager();
    if (sirrt_motion_check_batch(cm.handle(), q0, q1, t0, t1, (int64_t)n, cm.robot_joint_max_velocity()[0], variant, r, f) != SIRRT_OK)
      throw std::runtime_error(std::string("libsirrt_b200: ") + sirrt_last_error());
    // the reference counts every isValid a checkMotion issues (StateValidityCheckerFunctor.cpp:32): s2, then the
    // interpolated states up to and including the first invalid one
    const int dof = cm.dof();
    long long tested = 0;
    for (size_t e = 0; e < n; e++) {
      tested += 1;
      if (r[e] >= 0) tested += r[e] + 1;
      else if (r[e] == -1) {
        double n2 = 0;
        for (int i = 0; i < dof; i++) { const double d = q1[e * dof + i] - q0[e * dof + i]; n2 += d * d; }
        const double dt = t1[e] - t0[e];
        tested += std::max((int)(std::sqrt(n2) / 0.1), (int)((variant == 0 ? 2.0 : 1.0) * dt * cm.fps() + 0.5)) + 2;
      }
    }
    si.count_validations((int)tested);
  }
  const StateValidityCheckerFunctor& si;
  mutable int validation_counter = 0;
};

}  // namespace MDP

#if defined(__has_include)
#if __has_include(<ompl/base/spaces/SpaceTimeStateSpace.h>)
// Uncompiled in this image (no OMPL): the subclasses a maintainer registers with ob::SpaceInformation
// (STRRT_Planner/src/main.cpp sets them with setStateValidityChecker / setMotionValidator).
#include <ompl/base/MotionValidator.h>
#include <ompl/base/SpaceInformation.h>
#include <ompl/base/StateValidityChecker.h>
#include <ompl/base/spaces/RealVectorStateSpace.h>
#include <ompl/base/spaces/SpaceTimeStateSpace.h>
#include <ompl/base/spaces/TimeStateSpace.h>
namespace MDP { namespace ompl_adapters {
namespace ob = ompl::base;
inline SpaceTimeState unpack(const ob::State* s, int dof) {
  SpaceTimeState o;
  const double* v = s->as<ob::CompoundState>()->as<ob::RealVectorStateSpace::StateType>(0)->values;
  o.values.assign(v, v + dof);
  o.position = s->as<ob::CompoundState>()->as<ob::TimeStateSpace::StateType>(1)->position;
  return o;
}
class StateValidityChecker : public ob::StateValidityChecker {
 public:
  StateValidityChecker(const ob::SpaceInformationPtr& si, const MDP::StateValidityCheckerFunctor& f) : ob::StateValidityChecker(si), f(f) {}
  bool isValid(const ob::State* s) const override { auto p = unpack(s, f.manager().dof()); return f.isValid(&p); }
  double clearance(const ob::State* s) const override { auto p = unpack(s, f.manager().dof()); return f.clearance(&p); }
 private:
  const MDP::StateValidityCheckerFunctor& f;
};
class MotionValidator : public ob::MotionValidator {
 public:
  MotionValidator(const ob::SpaceInformationPtr& si, const MDP::SpaceTimeMotionValidator& m, int dof) : ob::MotionValidator(si), m(m), dof(dof), sinfo(si) {}
  bool checkMotion(const ob::State* a, const ob::State* b) const override { auto pa = unpack(a, dof), pb = unpack(b, dof); return m.checkMotion(&pa, &pb); }
  bool checkMotion(const ob::State* a, const ob::State* b, std::pair<ob::State*, double>& lastValid) const override {
    auto pa = unpack(a, dof), pb = unpack(b, dof);
    std::pair<SpaceTimeState*, double> lv{nullptr, 0.0};
    const bool ok = m.checkMotion(&pa, &pb, lv);
    lastValid.second = lv.second;
    if (lastValid.first) sinfo->getStateSpace()->interpolate(a, b, lv.second, lastValid.first);
    return ok;
  }
 private:
  const MDP::SpaceTimeMotionValidator& m;
  int dof;
  ob::SpaceInformationPtr sinfo;
};
}}  // namespace MDP::ompl_adapters
#endif
#endif
