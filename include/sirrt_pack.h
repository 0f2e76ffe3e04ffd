/*
 * sirrt_pack.h -- the "scene pack": a flat, little-endian, sequential binary image of one
 * planning scene.  It is the only input format of libsirrt_b200 (sirrt_scene_create_from_pack)
 * and of the test-only CPU oracle.  It carries exactly what the reference's
 * MDP::ConfigReader::SceneTask + MDP::RobotObstacleFCL::fk_chain hold after loading:
 *
 *   scene JSON      -> STRRT_Planner/src/config_read_writer/config_read.cpp:62-314
 *   obstacle poses  -> STRRT_Planner/include/config_read_writer/ObstacleJsonInfo.hpp:11-37 (float32!)
 *   URDF fk_chain   -> STRRT_Planner/src/CollisionManager/RobotObstacleFCL.cpp:54-195
 *
 * Producers: manipulationplanning-si-rrt_b200/tools/scene_pack.py (scene JSON + URDF + binary
 * STL, or the synthetic generators).  All parsing (JSON/XML/STL/Qhull) happens in the packer;
 * all arithmetic that decides parity (quaternion -> matrix, FK, interpolation of robot-obstacle
 * trajectories, collision) happens in the consumer, in FP64 on the host or FP32/FP64 on device.
 *
 * The stream is read front to back with a cursor; every record starts 8-byte aligned.
 *
 *   PackHeader
 *   f64 start_configuration[dof]     (SceneTask::start_configuration)
 *   f64 end_configuration[dof]
 *   f64 joint_max_velocity[dof]      (SceneTask::robot_joint_max_velocity; OMPL validator only)
 *   RobotRecord  planned robot
 *   ObstacleRecord x n_obstacles       (box / sphere; "C" in the survey)
 *   RobotObstacleRecord x n_robot_obstacles   ("R": dynamic_robot / static_robot)
 */
#ifndef SIRRT_PACK_H
#define SIRRT_PACK_H
#include <stdint.h>

#define SIRRT_PACK_MAGIC "SIRRTPK1"
#define SIRRT_PACK_VERSION 1u

enum { SIRRT_OBS_BOX = 0, SIRRT_OBS_SPHERE = 1 };
enum { SIRRT_SHAPE_CAPSULE = 0, SIRRT_SHAPE_HULL = 1 };

#pragma pack(push, 1)
typedef struct {
  char magic[8];      /* "SIRRTPK1" */
  uint32_t version;   /* 1 */
  uint32_t flags;     /* reserved, 0 */
  int32_t n_frames;   /* SceneTask::frame_count */
  int32_t fps;        /* SceneTask::fps */
  double start_time;  /* SceneTask::start_time (s) */
  int32_t dof;        /* planned robot: number of angles in a configuration */
  int32_t n_obstacles;        /* C */
  int32_t n_robot_obstacles;  /* R */
  int32_t reserved;
  uint64_t total_bytes;
} SirrtPackHeader;    /* 56 bytes */

/* One robot (planned robot or robot obstacle).
 * joints[]  : the NON-FIXED joints in chain order (fixed joints' origins are ignored by the
 *             reference, RobotObstacleFCL.cpp:317-325,357-367); joint k consumes angles[k].
 * shapes[]  : every link with collision geometry in chain order, INCLUDING the first
 *             n_strip (=2) base links that FK erases (RobotObstacleFCL.cpp:370) -- they still
 *             count in get_geometric_shapes() (QUIRK 7, CollisionManager.cpp:97-104). */
typedef struct {
  int32_t dof;         /* angles expected (robot_joints_order.size()) */
  int32_t n_joints;    /* non-fixed joints in the chain */
  int32_t n_shapes;    /* links with collision geometry (all) */
  int32_t n_strip;     /* 2 */
  float base_pos[3];   /* float32 narrowed (config_read.cpp:290-294) */
  float base_quat[4];  /* x,y,z,w, not normalised */
  int32_t pad;
  /* followed by:
   *   float  limits[dof][2]          (lower, upper as float32; RobotObstacleFCL.cpp:82), padded to 8
   *   SirrtJointRecord joints[n_joints]
   *   SirrtShapeRecord shapes[n_shapes], each followed by float verts[n_verts][3] padded to 8 */
} SirrtRobotHeader;    /* 48 bytes */

typedef struct {
  double t[3];   /* parent_to_joint_origin_transform.position */
  double R[9];   /* row-major, Eigen::Quaterniond(w,x,y,z).toRotationMatrix() of the urdf rpy */
} SirrtJointRecord;    /* 96 bytes */

typedef struct {
  int32_t type;            /* SIRRT_SHAPE_* ; URDF box links are stored as 8-vertex hulls */
  int32_t joints_before;   /* number of non-fixed joints applied before this link is emitted */
  int32_t n_verts;         /* hull only */
  int32_t pad;
  double lo_t[3];          /* link collision origin (cylinders only, else 0)  */
  double lo_R[9];          /* row-major (cylinders only, else identity)       */
  double radius;           /* capsule */
  double length;           /* capsule: full segment length lz (coal::Capsule(r, lz)) */
} SirrtShapeRecord;        /* 128 bytes */

typedef struct {
  int32_t type;       /* SIRRT_OBS_* */
  int32_t is_static;  /* 1: one pose for all frames */
  int32_t n_poses;    /* 1 if static else >= n_frames */
  int32_t pad;
  float dims[4];      /* box: full edge lengths x,y,z,0 ; sphere: r,0,0,0  (float32 narrowed) */
  /* followed by float poses[n_poses][7] = x,y,z,qx,qy,qz,qw (float32 narrowed), padded to 8 */
} SirrtObstacleRecord;   /* 32 bytes */

typedef struct {
  int32_t is_static;
  int32_t n_states;
  /* followed by: SirrtRobotHeader+payload, then double times[n_states],
   * then double angles[n_states][dof] */
} SirrtRobotObstacleHeader;  /* 8 bytes */
#pragma pack(pop)

#endif /* SIRRT_PACK_H */
