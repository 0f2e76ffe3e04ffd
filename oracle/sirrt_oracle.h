/*
 * sirrt_oracle.h -- TEST INFRASTRUCTURE ONLY.  CPU (FP64) restatement of the reference's
 * collision-checker / safe-interval / edge-validation / kNN algorithm.
 *
 * PARITY UNPINNED against the reference: it ships no unit tests or golden vectors for this path and cannot be
 * compiled in this image (coal, Eigen, nanoflann, urdfdom, orocos-kdl are absent), so no output of the shipped
 * C++ checker (CollisionManager.cpp:330-380) exists to compare with.  What stands in for it:
 *   - a second, code-independent oracle (oracle/oracle2.py: NumPy/SciPy, LP feasibility + NNLS-QP for hulls, scalar
 *     minimisation for capsule/box, matrix FK, own pack parser, no GJK) agrees with this file on >= 10 000 stratified
 *     (configuration, frame) cells per scene, >= 1 000 colliding and >= 200 within 1e-4 m of contact
 *     (tests/test_oracle2.py) -- this pins the arithmetic on BOTH sides of the decision, not the reference's semantics;
 *   - a centimetre-level sanity check: trajectories logged by an OLDER pipeline of the same project (the JSON logs
 *     under "Blender/scenes with animation" carry the keys of STRRT/STRRT_eval/evaluate_blender_scene.py:414-422, whose
 *     checker was pybullet, not the shipped ResultsWriter.cpp:69-99) stay collision-free here, passing 2-31 mm from
 *     the obstacles (tests/test_oracle.py);
 *   - the reference's call sites (every function cites them), check_scene invariants, FK sanity values.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library.  The product (libsirrt_b200.so) never links or calls it.
 */
#ifndef SIRRT_ORACLE_H
#define SIRRT_ORACLE_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct Oracle Oracle;

enum { ORACLE_REFERENCE_EXACT = 0, ORACLE_FIXED = 1 };      /* QUIRK 7 behaviour */
enum { ORACLE_SEM_INTERVAL = 0, ORACLE_SEM_FRAME = 1 };      /* which callback's semantics */

Oracle* oracle_create(const void* pack, size_t bytes, int parity_mode);
void oracle_destroy(Oracle*);
const char* oracle_last_error(void);
int oracle_info(const Oracle*, int* dof, int* n_links, int* n_frames, int* fps, float* limits_lo_hi);
int oracle_counts(const Oracle*, int* n_obstacles, int* n_robot_obstacles, int* n_robot_obstacle_links);

/* a1: link world transforms, [n_links][12] = R row-major (9) then t (3) */
int oracle_fk(const Oracle*, const double* q, double* out);
/* a7: robot-obstacle link transforms at a frame, [n_links_of_that_robot][12] */
int oracle_robot_obstacle_fk(const Oracle*, int robot, int frame, double* out);

/* collision bits of one configuration over all frames (bit f of word f/32 set = colliding).
 * near_contact (may be NULL): number of frames whose deciding signed distance is within tol. */
int oracle_collision_bits(const Oracle*, const double* q, int semantics, uint32_t* bits, double tol,
                          int64_t* near_contact);
/* a3: CollisionManager::get_safe_intervals; returns interval count (may exceed cap) */
int oracle_safe_intervals(const Oracle*, const double* q, int32_t* lo_hi, int cap);
/* a4: CollisionManager::check_collision_frame; out-of-range frame -> 1 */
int oracle_check_frame(const Oracle*, const double* q, int frame);
/* signed distance of planned-robot link l vs obstacle id (0..C-1 boxes/spheres, then
 * robot-obstacle links in reference id order) at a frame; core penetration returns -(ra+rb) */
double oracle_pair_distance(const Oracle*, const double* q, int link, int obstacle_id, int frame);
/* a5: PlannerConnect::is_collision_motion; returns first colliding step, -1 free,
 * -2 start_time>=end_time, -3 speed gate; collision_coord (may be NULL) gets q at that step */
int oracle_edge_check(const Oracle*, const double* q0, const double* q1, double t0, double t1, double vmax,
                      double* collision_coord);
/* f4: StateValidityCheckerFunctor::isValid (time in seconds, float32 frame arithmetic) and
 * SpaceTimeMotionValidator::checkMotion (variant 0 = two-argument form, 1 = lastValid form) */
int oracle_state_valid(const Oracle*, const double* q, double time_s);
int oracle_motion_check(const Oracle*, const double* q0, const double* q1, double t0, double t1, double vmax,
                        int variant, double* last_valid_fraction);
/* a9: get_distances: per planned-robot link the minimum distance over box/sphere obstacles at a
 * frame (robot obstacles are "not implemented" in the reference, CollisionManager.cpp:418).
 * returns 1 if a collision was hit (reference early-exits), else 0 */
int oracle_get_distances(const Oracle*, const double* q, int frame, double* min_dist /*[n_links]*/);

/* batched / threaded forms used as the CPU baseline (configs as float32 SoA [dof][N]) */
int oracle_bvh_build(Oracle*);  /* AABB tree over constant-pose windows (CollisionManager.cpp:107-216) */
int oracle_safe_intervals_batch(const Oracle*, const float* q_soa, int64_t N, int32_t* table, int Kmax,
                                uint32_t* bits, int use_bvh, int n_threads);
int oracle_check_frames_batch(const Oracle*, const float* q_soa, const int32_t* frames, int64_t N,
                              uint8_t* collided, int n_threads);

/* a6: exact brute-force kNN in FP64 (nanoflann L2_Simple metric) */
int oracle_knn_nearest(const double* pts, int64_t n, int dim, const double* queries, int64_t nq, int64_t* idx,
                       double* dist2);
int64_t oracle_knn_radius(const double* pts, int64_t n, int dim, const double* query, double r2, int64_t* idx,
                          int64_t cap);

/* primitives exposed for cross-checks */
double oracle_gjk_distance(const double* a, int na, const double* b, int nb);  /* 0 if intersecting */
double oracle_seg_box_dist2(const double a[3], const double b[3], const double half[3]);
double oracle_seg_seg_dist2(const double p1[3], const double q1[3], const double p2[3], const double q2[3]);
double oracle_pt_seg_dist2(const double p[3], const double a[3], const double b[3]);

/* Baseline mode for the TIMED CPU arm (process-global): boolean early exits + AABB tree, bit-identical results. */
void oracle_set_fast(int on);
int oracle_get_fast(void);
#ifdef __cplusplus
}
#endif
#endif
