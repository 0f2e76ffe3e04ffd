/*
 * sirrt_b200.h -- C ABI of libsirrt_b200.so: the B200 (sm_100a) backend for SI-RRT's hot path
 * (safe intervals, per-frame collision checks, edge validation, C-space kNN).
 *
 * Every entry point names the reference interface it replaces (paths relative to the reference
 * root).  Conventions:
 *   - all functions return 0 on success and a negative SIRRT_E_* code otherwise; nothing throws
 *     across the ABI; sirrt_last_error() gives the message of the calling thread's last failure;
 *   - the caller owns every buffer; data pointers may be host (pageable or pinned) or device
 *     memory of the scene's GPU (detected with cudaPointerGetAttributes);
 *   - a sirrt_scene is bound to one CUDA device and one stream and may be used by one host
 *     thread at a time; different scenes may be used concurrently;
 *   - there is NO CPU fallback: without a usable CUDA device every compute call fails with
 *     SIRRT_E_CUDA.
 *   - configurations are "q": dof joint angles in chain order of the non-fixed joints
 *     (RobotObstacleFCL.cpp:321,363).  Frames are integer timesteps 0..n_frames-1; callers do
 *     their own time->frame conversion (SURVEY.md §3.5 QUIRK 13).
 */
#ifndef SIRRT_B200_H
#define SIRRT_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct sirrt_scene sirrt_scene;
typedef struct sirrt_knn sirrt_knn;
typedef struct sirrt_multi sirrt_multi;

enum {
  SIRRT_OK = 0,
  SIRRT_E_ARG = -1,      /* bad argument */
  SIRRT_E_PACK = -2,     /* malformed scene pack */
  SIRRT_E_CUDA = -3,     /* CUDA runtime failure / no device */
  SIRRT_E_LIMIT = -4,    /* scene exceeds a compiled-in limit (links, joints) */
  SIRRT_E_OVERFLOW = -5  /* output capacity too small */
};

/* QUIRK 7 (CollisionManager.cpp:97-104,206-208,263-265,628,691): robot-obstacle functor
 * mis-indexing.  REFERENCE_EXACT reproduces it, FIXED tests the true geometry only. */
enum { SIRRT_PARITY_REFERENCE_EXACT = 0, SIRRT_PARITY_FIXED = 1 };
/* which reference callback a bit table follows: SafeIntervalCollisionCallback
 * (CollisionManager.cpp:648-740) or CollisionCallback (:602-646); they differ only for
 * robot obstacles in REFERENCE_EXACT mode */
enum { SIRRT_SEM_INTERVAL = 0, SIRRT_SEM_FRAME = 1 };

const char* sirrt_last_error(void);
int sirrt_version(void);
int sirrt_device_count(void);
/* FP32 FMA throughput of `device` (TFLOP/s, FMA = 2) from an FFMA micro-benchmark: the roofline
 * denominator bench.py reports against (MEASURED_PEAKS.json carries no FP32 figure) */
int sirrt_measure_fp32_peak(int device, double* tflops);

/* ---- scene: replaces MDP::CollisionManager::CollisionManager(SceneTask), CollisionManager.cpp:19-280
 * (+ RobotObstacleFCL ctor, RobotObstacleFCL.cpp:21-30).  `pack` layout: include/sirrt_pack.h. */
int sirrt_scene_create_from_pack(const void* pack, size_t bytes, int device, int parity_mode, sirrt_scene** out);
void sirrt_scene_destroy(sirrt_scene*);
/* dof / links after base strip / frame_count / fps / limits[2*dof] = get_planned_robot_limits()
 * (CollisionManager.cpp:422-425); any pointer may be NULL */
int sirrt_scene_info(const sirrt_scene*, int* dof, int* n_links, int* n_frames, int* fps, float* limits_lo_hi);
/* SceneTask::start_configuration / end_configuration / robot_joint_max_velocity (config_read.hpp:31-79) */
int sirrt_scene_task(const sirrt_scene*, double* start_cfg, double* end_cfg, double* joint_max_velocity);
/* run all work of this scene on `cuda_stream` (a cudaStream_t; NULL = the scene's own stream) */
int sirrt_scene_set_stream(sirrt_scene*, void* cuda_stream);
/* half-width (m) of the FP32 uncertainty band inside which a pair is re-evaluated in FP64 */
int sirrt_scene_set_band(sirrt_scene*, double band);
/* counters of the last batched call: [0] kernels launched, [1] FP64 re-checked pair tests,
 * [2] near-contact pair tests (|d| < 1e-6 m), [3] narrowphase (exact) pair tests,
 * [4] device time of the collision sweep kernel(s), nanoseconds (CUDA events on the launch stream) */
int sirrt_scene_last_stats(const sirrt_scene*, uint64_t stats[5]);
/* tuning knobs (defaults are right for B200): see SIRRT_OPT_* */
enum {
  SIRRT_OPT_USE_SWEEP = 0,      /* 1 (default): capsule-link x box/sphere scenes use the queue-compacted sweep kernel */
  SIRRT_OPT_CHUNK_CONFIGS = 1,  /* configurations per launch of the batched path (default 262144) */
  SIRRT_OPT_WORKLIST_CAP = 2,   /* capacity of the FP64 re-check work list, entries (default 4194304) */
  SIRRT_OPT_SWEEP_MIN_CONFIGS = 3, /* batches smaller than this use the generic kernel, which splits rows over warps (default 8192) */
  SIRRT_OPT_HULL_CELLS = 5,     /* 1 (default): GJK support queries on hulls scan the direction cell's candidate vertices; 0: every vertex (same result) */
  SIRRT_OPT_PIPELINE = 6,       /* 1 (default): batches of more than 131072 configurations in HOST buffers run as a 3-stream pipeline (H2D || kernels || D2H) */
  SIRRT_OPT_PDL = 8,            /* 1: the kernels of one planner-sized call are chained by programmatic dependent launch (default 0: ~1 us faster per call on an idle GPU, 30 % slower when 48 planner threads share it) */
  SIRRT_OPT_INLINE_RECHECK = 9, /* n > 0 (default 9): single-configuration / cell calls on hull scenes do not queue the FP64 re-check kernel; the task kernel's last block re-checks up to n - 1 items itself (one launch fewer per call), more are handed back to the host; 0: always queue the kernel */
  SIRRT_OPT_SEED_RECHECK = 10,  /* 1 (default): on planner-sized hull calls the FP64 re-check GJK starts from the simplex the FP32 run ended on */
  SIRRT_OPT_HOST_BUFFERS = 11,  /* 1: the caller promises HOST memory for every buffer it passes to any entry point of this scene (the C++ shim does: std::vector storage only) -- no pointer-attribute query per pointer and call; default 0 */
  SIRRT_OPT_OBB_CULL = 12,      /* 1 (default): pair tests that involve a convex hull are preceded by a 15-axis separating-axis test of the two shapes' oriented bounding boxes (decides "certainly free" only; same results, fewer GJK runs) */
  SIRRT_OPT_SLOT_SPLIT = 7,     /* 1 (default): planner-sized batches split the dynamic obstacles of a row over several warps (same result) */
  SIRRT_OPT_DEFER_MAX_CONFIGS = 4  /* hull scenes: batches up to this many configurations hand their GJK pair tests to a machine-wide task kernel (default 4096; 0 = never) */
};
int sirrt_scene_set_option(sirrt_scene*, int key, long long value);
/* link world transforms of the planned robot for q, [n_links][12] = R row-major then t, from the
 * device FK the kernels use (RobotObstacleFCL::get_collision_object_for_robot_angles,
 * RobotObstacleFCL.cpp:197-201,287-386) */
int sirrt_scene_fk(sirrt_scene*, const double* q, double* out);
/* the same transforms from the warp-parallel form of the chain (parallel prefix over the joints) that the
 * planner-sized kernels use to build their FP32 link table; agrees with sirrt_scene_fk to rounding (the
 * products associate differently), not bit for bit -- exposed so that the tests can say how closely */
int sirrt_scene_fk_warp(sirrt_scene*, const double* q, double* out);

/* ---- single-call drop-ins (latency path) ------------------------------------------------- */
/* CollisionManager::get_safe_intervals, CollisionManager.cpp:481-547.  lo_hi receives up to `cap`
 * inclusive (lo,hi) pairs, ascending; *n_out the true count (SIRRT_E_OVERFLOW if > cap). */
int sirrt_safe_intervals(sirrt_scene*, const double* q, int32_t* lo_hi, int cap, int* n_out);
/* CollisionManager::check_collision_frame[_no_wrapper], CollisionManager.cpp:338-380.
 * frame outside [0, n_frames) answers "collided" (the reference indexes unchecked, :369). */
int sirrt_check_frame(sirrt_scene*, const double* q, int frame, int* collided);
/* CollisionManager::get_distances, CollisionManager.cpp:382-420 (+DistanceCallback :549-600):
 * per planned-robot link the minimum distance to the box/sphere obstacles at `frame`.
 * Links after the first colliding one are left at +inf (the reference returns early). */
int sirrt_get_distances(sirrt_scene*, const double* q, int frame, double* min_dist, int* is_collision);
/* the same, plus what DistanceCallback::distance collects per visited pair (CollisionManager.cpp:568-598,
 * robot_link_distance_to_obj at CollisionManager.hpp:91-99): for EVERY (link, box/sphere obstacle) pair of the frame its signed
 * distance pair_dist [n_links][C] and the closest point pair points [n_links][C][6] = point on the link surface (xyz), point on
 * the obstacle surface (xyz), world frame (coal DistanceResult::nearest_points[0], [1]).  The reference's vectors hold the
 * pairs its AABB-tree traversal happened to visit, in traversal order; here all C pairs are returned in obstacle order. */
int sirrt_get_distances_ex(sirrt_scene*, const double* q, int frame, double* min_dist, int* is_collision, double* pair_dist, double* points);
/* number of box / sphere obstacles (C) and of robot obstacles of the scene */
int sirrt_scene_counts(const sirrt_scene*, int* n_obstacles, int* n_robot_obstacles);

/* ---- batched throughput path ------------------------------------------------------------- */
/* N configurations -> collision bit rows and/or fixed-width interval table.
 *   q_soa  : float32 [dof][N]      (structure of arrays; coalesced loads)
 *   table  : int32 [N][1+2*Kmax]   row = {count, lo0, hi0, lo1, hi1, ...}, unused = -1; count is the
 *            true interval count even when it exceeds Kmax (then only Kmax pairs are stored); or NULL
 *   bits   : uint32 [N][ceil(T/32)] bit f%32 of word f/32 set = colliding at frame f; or NULL
 *   near_contact_count : pair tests within 1e-6 m of contact (after FP64 re-evaluation); or NULL */
int sirrt_safe_intervals_batch(sirrt_scene*, const float* q_soa, int64_t N, int32_t* table, int Kmax,
                               uint32_t* bits, uint64_t* near_contact_count);
/* compact interval table for host callers: int16 entries, same row layout {count, lo0, hi0, ...}, unused = -1 (halves the
 * bytes read back over the bus); host buffers only, scenes with at most 32767 frames (SIRRT_E_LIMIT otherwise) */
int sirrt_safe_intervals_batch_i16(sirrt_scene*, const float* q_soa_host, int64_t N, int16_t* table16, int Kmax, uint64_t* near_contact_count);
/* same, double-precision configurations as array of structures [N][dof] (edge tables, §7.3) with
 * explicit semantics (SIRRT_SEM_*) */
int sirrt_collision_bits_f64(sirrt_scene*, const double* q_aos, int64_t N, int semantics, uint32_t* bits);
/* windowed form of the edge table: row i is only computed for the n_tiles 32-frame tiles starting at tile first_tile[i]
 * (host array [N], clamped to the scene); bits keeps the full-row layout [N][ceil(T/32)], words outside a row's window carry
 * no information (zero, or all ones when the configuration hits a static obstacle) */
int sirrt_collision_bits_window_f64(sirrt_scene*, const double* q_aos, int64_t N, int semantics, const int32_t* first_tile, int n_tiles,
                                    uint32_t* bits);
/* N independent (configuration, frame) cells: check_collision_frame semantics */
int sirrt_check_frames_batch(sirrt_scene*, const float* q_soa, const int32_t* frames, int64_t N, uint8_t* collided);
int sirrt_check_frames_batch_f64(sirrt_scene*, const double* q_aos, const int32_t* frames, int64_t N, uint8_t* collided);
/* PlannerConnect::is_collision_motion, PlannerConnect.cpp:862-897, for n_edges edges.
 *   q0,q1 : double [n_edges][dof]; t0,t1 : double [n_edges], in frame units
 *   vmax  : speed gate (:872) in rad/s, <= 0 disables it
 *   first_colliding_step : -1 free, >= 0 index of the first colliding interpolation step,
 *                          -2 start_time >= end_time (:865), -3 speed gate (:872)
 *   collision_coord : double [n_edges][dof] configuration at that step, or NULL */
int sirrt_edge_check_batch(sirrt_scene*, const double* q0, const double* q1, const double* t0, const double* t1,
                           int64_t n_edges, double vmax, int32_t* first_colliding_step, double* collision_coord);

/* SpaceTimeMotionValidator::checkMotion for n motions (ST-RRT* hook),
 * STRRT_Planner/src/OMPLValidators/SpaceTimeMotionValidator.cpp:22-74 (variant 0) and :76-127 (variant 1, lastValid),
 * every state tested as StateValidityCheckerFunctor::isValid does (StateValidityCheckerFunctor.cpp:29-47:
 * frame = (int)((float)t * fps), CollisionManager.cpp:333).
 *   q0,q1 : double [n][dof]; t0,t1 : double [n] in SECONDS (the OMPL time component)
 *   vmax  : robot_joint_max_velocity[0] (:46), <= 0 disables the gate
 *   result: -1 valid motion, -4 s2 invalid (:25), -2 delta_t <= 0 (:41), -3 speed gate (:46),
 *           >= 0 index i of the first invalid interpolated state (fraction i/steps, :54-64)
 *   last_valid_fraction : lastValid.second of the :76-127 form ([n], or NULL) */
int sirrt_motion_check_batch(sirrt_scene*, const double* q0, const double* q1, const double* t0, const double* t1,
                             int64_t n, double vmax, int variant, int32_t* result, double* last_valid_fraction);

/* ---- kNN: replaces nanoflann KDTreeSingleIndexDynamicAdaptor<L2_Simple> as used by
 * MSIRRT/src/Tree.cpp:5,18-36 and PlannerConnect.cpp:382-412 (exact, FP64, squared L2) ------- */
/* `capacity` is the initial reservation: the store doubles when an insertion does not fit (up to 2^31 - 1 points) */
int sirrt_knn_create(int dim, int64_t capacity, int device, sirrt_knn** out);
int sirrt_knn_destroy(sirrt_knn*);
int sirrt_knn_size(const sirrt_knn*, int64_t* n);
int sirrt_knn_info(const sirrt_knn*, int* dim, int64_t* capacity, int* device); /* any pointer may be NULL */
/* on != 0: the caller promises host memory for every buffer it passes to this store (as SIRRT_OPT_HOST_BUFFERS does
 * for a scene): the single-point add / single-query calls of a planner then skip the pointer-attribute queries */
int sirrt_knn_host_buffers(sirrt_knn*, int on);
int sirrt_knn_clear(sirrt_knn*);
int sirrt_knn_add(sirrt_knn*, const double* pts /*[n][dim]*/, int64_t n);
/* nq == 1 with host buffers: single-launch latency path; nq > 1: tiled multi-query kernel (FP32 filter with a rigorous
 * bound + FP64 verification; same answers bit for bit).  queries / idx / dist2 may be host or device pointers. */
int sirrt_knn_nearest(sirrt_knn*, const double* queries /*[nq][dim]*/, int64_t nq, int64_t* idx, double* dist2);
/* strict dist2 < r2 (nanoflann RadiusResultSet); indices ascending; *n_out = true count */
int sirrt_knn_radius(sirrt_knn*, const double* query, double r2, int64_t* idx, int64_t cap, int64_t* n_out);
/* The same query in two halves, for a planner that has other GPU work to issue in between: PlannerConnect.cpp:105-150 asks
 * nearest -> steer -> get_safe_intervals -> set_parent, and the radius query that opens set_parent (:414-416, :394-412)
 * depends on the steered configuration only.  _begin launches and returns at once; _end waits and answers exactly what
 * sirrt_knn_radius answers.  One query may be pending per store; any other call on the store settles (and drops) it. */
int sirrt_knn_radius_begin(sirrt_knn*, const double* query, double r2);
int sirrt_knn_radius_end(sirrt_knn*, int64_t* idx, int64_t cap, int64_t* n_out);
/* k nearest of each query, ascending distance (ties by index); idx/dist2 [nq][k] */
int sirrt_knn_topk(sirrt_knn*, const double* queries, int64_t nq, int k, int64_t* idx, double* dist2);

/* Exact 1-NN WITHOUT ANY ALLOCATION, for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  No counterpart in the reference.
 * Same arithmetic and tie rule as sirrt_knn_nearest (FP64 squared L2 in dimension order, ties to the lowest
 * index); the kernel's scratch (block minima, tickets) is a fixed file-scope device array of 257 KB that every GPU which
 * loads the library holds until exit -- hence the limit of SIRRT_KNN_NOALLOC_MAX_QUERIES queries per call (more:
 * SIRRT_E_LIMIT; split the batch).  The call allocates nothing (no cudaMalloc / cudaFree, no store object) and is
 * ASYNCHRONOUS on cuda_stream (0 = the legacy default stream): the caller synchronises before it reads the results.
 *   pts_soa : DEVICE memory, double [dim][stride], point i = pts_soa[k * stride + i], n <= stride points used
 *   queries : DEVICE memory, double [nq][dim];  idx : DEVICE int64 [nq] (-1 when n == 0);  dist2 : DEVICE double [nq]
 * ALL CALLS SHARE THE SCRATCH ARRAY: it keeps the last call's contents (the words that must start at zero are cleared on
 * the stream in front of every launch), and two calls that overlap on one GPU -- other streams, other threads -- RACE. */
#define SIRRT_KNN_NOALLOC_MAX_QUERIES 256
int sirrt_knn_nearest_noalloc(int device, void* cuda_stream, const double* pts_soa, int64_t n, int64_t stride, int dim,
                              const double* queries, int64_t nq, int64_t* idx, double* dist2);

/* one tree grown across n_shards GPUs (one process per GPU): shard g stores the points whose global index is
 * g (mod n_shards) at local index global / n_shards.  After each rank searched its shard (sirrt_knn_topk / _nearest
 * into device buffers) and the lists were all-gathered over NCCL, this merges them on the device:
 *   cand_d2 [n_shards][nq][k] ascending per list, cand_idx [n_shards][nq][k] LOCAL indices (-1 = no candidate)
 *   out_idx / out_d2 [nq][k]: GLOBAL indices ascending by (dist2, index); asynchronous on cuda_stream */
#define SIRRT_KNN_MAX_SHARDS 64
int sirrt_knn_merge_shards(int device, void* cuda_stream, const double* cand_d2, const int64_t* cand_idx, int n_shards, int64_t nq,
                           int k, int64_t* out_idx, double* out_d2);

/* ---- multi-GPU inside one process: shards a configuration batch over several devices ------ */
int sirrt_multi_create(const void* pack, size_t bytes, const int* devices, int n_dev, int parity_mode, sirrt_multi** out);
void sirrt_multi_destroy(sirrt_multi*);
int sirrt_multi_safe_intervals_batch(sirrt_multi*, const float* q_soa_host, int64_t N, int32_t* table, int Kmax,
                                     uint32_t* bits, uint64_t* near_contact_count);

#ifdef __cplusplus
}
#endif
#endif /* SIRRT_B200_H */
