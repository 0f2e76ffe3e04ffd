"""sirrt_b200 -- Python mirror of the reference's collision-checker interface over libsirrt_b200.so.

The product is the C-ABI library (include/sirrt_b200.h; hand-written sm_100a kernels).  This module
is a thin ctypes binding used by the tests, bench.py and as an example of the reference-side
binding.  `CollisionManager` mirrors MDP::CollisionManager's public surface
(STRRT_Planner/include/CollisionManager/CollisionManager.hpp:74,99-117): same method names,
argument meaning and error behaviour.  There is NO CPU fallback: importing fails if the library
is missing, every compute call raises if no CUDA device is usable.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_DIR, "libsirrt_b200.so")

PARITY_REFERENCE_EXACT, PARITY_FIXED = 0, 1
SEM_INTERVAL, SEM_FRAME = 0, 1
OPT_USE_SWEEP, OPT_CHUNK_CONFIGS, OPT_WORKLIST_CAP, OPT_SWEEP_MIN_CONFIGS, OPT_DEFER_MAX_CONFIGS, OPT_HULL_CELLS, OPT_PIPELINE, OPT_SLOT_SPLIT, OPT_PDL, OPT_INLINE_RECHECK, OPT_SEED_RECHECK, OPT_HOST_BUFFERS, OPT_OBB_CULL = 0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(nvcc, sm_100a).  sirrt_b200 has no CPU or PyTorch fallback."
    )

_lib = C.CDLL(LIB_PATH)
_vp, _i32, _i64, _f64, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_double, C.c_size_t

EXPORTS = {
    "sirrt_last_error": (C.c_char_p, []),
    "sirrt_version": (_i32, []),
    "sirrt_device_count": (_i32, []),
    "sirrt_measure_fp32_peak": (_i32, [_i32, _vp]),
    "sirrt_scene_create_from_pack": (_i32, [_vp, _sz, _i32, _i32, _vp]),
    "sirrt_scene_destroy": (None, [_vp]),
    "sirrt_scene_info": (_i32, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "sirrt_scene_task": (_i32, [_vp, _vp, _vp, _vp]),
    "sirrt_scene_set_stream": (_i32, [_vp, _vp]),
    "sirrt_scene_set_band": (_i32, [_vp, _f64]),
    "sirrt_scene_last_stats": (_i32, [_vp, _vp]),
    "sirrt_scene_set_option": (_i32, [_vp, _i32, C.c_longlong]),
    "sirrt_scene_fk": (_i32, [_vp, _vp, _vp]),
    "sirrt_scene_fk_warp": (_i32, [_vp, _vp, _vp]),
    "sirrt_knn_host_buffers": (_i32, [_vp, _i32]),
    "sirrt_knn_radius_begin": (_i32, [_vp, _vp, _f64]),
    "sirrt_knn_radius_end": (_i32, [_vp, _vp, _i64, _vp]),
    "sirrt_knn_nearest_noalloc": (_i32, [_i32, _vp, _vp, _i64, _i64, _i32, _vp, _i64, _vp, _vp]),
    "sirrt_safe_intervals": (_i32, [_vp, _vp, _vp, _i32, _vp]),
    "sirrt_check_frame": (_i32, [_vp, _vp, _i32, _vp]),
    "sirrt_get_distances": (_i32, [_vp, _vp, _i32, _vp, _vp]),
    "sirrt_get_distances_ex": (_i32, [_vp, _vp, _i32, _vp, _vp, _vp, _vp]),
    "sirrt_scene_counts": (_i32, [_vp, _vp, _vp]),
    "sirrt_safe_intervals_batch": (_i32, [_vp, _vp, _i64, _vp, _i32, _vp, _vp]),
    "sirrt_safe_intervals_batch_i16": (_i32, [_vp, _vp, _i64, _vp, _i32, _vp]),
    "sirrt_collision_bits_f64": (_i32, [_vp, _vp, _i64, _i32, _vp]),
    "sirrt_collision_bits_window_f64": (_i32, [_vp, _vp, _i64, _i32, _vp, _i32, _vp]),
    "sirrt_check_frames_batch": (_i32, [_vp, _vp, _vp, _i64, _vp]),
    "sirrt_check_frames_batch_f64": (_i32, [_vp, _vp, _vp, _i64, _vp]),
    "sirrt_edge_check_batch": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _f64, _vp, _vp]),
    "sirrt_knn_merge_shards": (_i32, [_i32, _vp, _vp, _vp, _i32, _i64, _i32, _vp, _vp]),
    "sirrt_motion_check_batch": (_i32, [_vp, _vp, _vp, _vp, _vp, _i64, _f64, _i32, _vp, _vp]),
    "sirrt_knn_create": (_i32, [_i32, _i64, _i32, _vp]),
    "sirrt_knn_destroy": (_i32, [_vp]),
    "sirrt_knn_size": (_i32, [_vp, _vp]),
    "sirrt_knn_info": (_i32, [_vp, _vp, _vp, _vp]),
    "sirrt_knn_clear": (_i32, [_vp]),
    "sirrt_knn_add": (_i32, [_vp, _vp, _i64]),
    "sirrt_knn_nearest": (_i32, [_vp, _vp, _i64, _vp, _vp]),
    "sirrt_knn_radius": (_i32, [_vp, _vp, _f64, _vp, _i64, _vp]),
    "sirrt_knn_topk": (_i32, [_vp, _vp, _i64, _i32, _vp, _vp]),
    "sirrt_multi_create": (_i32, [_vp, _sz, _vp, _i32, _i32, _vp]),
    "sirrt_multi_destroy": (None, [_vp]),
    "sirrt_multi_safe_intervals_batch": (_i32, [_vp, _vp, _i64, _vp, _i32, _vp, _vp]),
}
for _name, (_res, _args) in EXPORTS.items():
    _f = getattr(_lib, _name)  # AttributeError here = the library does not export what the header declares
    _f.restype = _res
    _f.argtypes = _args


class SirrtError(RuntimeError):
    def __init__(self, code: int):
        self.code = code
        super().__init__(f"libsirrt_b200 error {code}: {_lib.sirrt_last_error().decode(errors='replace')}")


def _ck(rc: int):
    if rc != 0:
        raise SirrtError(rc)


def _ptr(a):
    """numpy array / torch tensor (host or cuda) / None -> raw address."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data
    if hasattr(a, "data_ptr"):
        return a.data_ptr()
    raise TypeError(type(a))


def device_count() -> int:
    return _lib.sirrt_device_count()


def measure_fp32_peak(device: int = 0) -> float:
    v = C.c_double()
    _ck(_lib.sirrt_measure_fp32_peak(device, C.byref(v)))
    return v.value


class Scene:
    """Owns the device tables of one planning scene on one GPU (sirrt_scene)."""

    def __init__(self, pack: bytes, device: int = 0, parity_mode: int = PARITY_REFERENCE_EXACT):
        self._h = _vp()
        buf = bytes(pack)
        _ck(_lib.sirrt_scene_create_from_pack(buf, len(buf), device, parity_mode, C.byref(self._h)))
        v = [C.c_int() for _ in range(4)]
        _ck(_lib.sirrt_scene_info(self._h, *[C.byref(x) for x in v], None))
        self.dof, self.n_links, self.n_frames, self.fps = [x.value for x in v]
        self.limits = np.zeros((self.dof, 2), np.float32)
        _ck(_lib.sirrt_scene_info(self._h, None, None, None, None, self.limits.ctypes.data))
        self.words = (self.n_frames + 31) // 32
        self.device = device
        self.start_configuration = np.zeros(self.dof)
        self.end_configuration = np.zeros(self.dof)
        self.joint_max_velocity = np.zeros(self.dof)
        _ck(_lib.sirrt_scene_task(self._h, self.start_configuration.ctypes.data, self.end_configuration.ctypes.data,
                                  self.joint_max_velocity.ctypes.data))

    def close(self):
        if getattr(self, "_h", None):
            _lib.sirrt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- knobs
    def set_stream(self, cuda_stream: int | None):
        _ck(_lib.sirrt_scene_set_stream(self._h, cuda_stream))

    def set_band(self, band: float):
        _ck(_lib.sirrt_scene_set_band(self._h, band))

    def set_option(self, key: int, value: int):
        _ck(_lib.sirrt_scene_set_option(self._h, key, value))

    def last_stats(self):
        st = np.zeros(5, np.uint64)
        _ck(_lib.sirrt_scene_last_stats(self._h, st.ctypes.data))
        return {"launches": int(st[0]), "rechecked": int(st[1]), "near_contact": int(st[2]), "exact_tests": int(st[3]),
                "sweep_kernel_ms": float(st[4]) * 1e-6}

    # -- single-call drop-ins
    def fk(self, q):
        q = np.ascontiguousarray(q, np.float64)
        out = np.zeros((self.n_links, 12))
        _ck(_lib.sirrt_scene_fk(self._h, q.ctypes.data, out.ctypes.data))
        return out[:, :9].reshape(-1, 3, 3), out[:, 9:]

    def fk_warp(self, q):
        """The warp-parallel FK of the planner-sized kernels (sirrt_scene_fk_warp)."""
        q = np.ascontiguousarray(q, np.float64)
        out = np.zeros((self.n_links, 12))
        _ck(_lib.sirrt_scene_fk_warp(self._h, q.ctypes.data, out.ctypes.data))
        return out[:, :9].reshape(-1, 3, 3), out[:, 9:]

    def safe_intervals(self, q, cap: int = 1024):
        q = np.ascontiguousarray(q, np.float64)
        assert q.shape == (self.dof,)
        buf = np.zeros(2 * cap, np.int32)
        n = C.c_int()
        _ck(_lib.sirrt_safe_intervals(self._h, q.ctypes.data, buf.ctypes.data, cap, C.byref(n)))
        return [(int(buf[2 * i]), int(buf[2 * i + 1])) for i in range(n.value)]

    def check_frame(self, q, frame: int) -> bool:
        q = np.ascontiguousarray(q, np.float64)
        out = C.c_int()
        _ck(_lib.sirrt_check_frame(self._h, q.ctypes.data, int(frame), C.byref(out)))
        return bool(out.value)

    def get_distances(self, q, frame: int):
        q = np.ascontiguousarray(q, np.float64)
        d = np.zeros(self.n_links)
        col = C.c_int()
        _ck(_lib.sirrt_get_distances(self._h, q.ctypes.data, int(frame), d.ctypes.data, C.byref(col)))
        return bool(col.value), d

    def counts(self):
        c, r = C.c_int(), C.c_int()
        _ck(_lib.sirrt_scene_counts(self._h, C.byref(c), C.byref(r)))
        return c.value, r.value

    def get_distances_ex(self, q, frame: int):
        """(is_collision, min_dist [L], pair_dist [L][C], link points [L][C][3], obstacle points [L][C][3])"""
        q = np.ascontiguousarray(q, np.float64)
        n_obs = self.counts()[0]
        d = np.zeros(self.n_links)
        pd = np.zeros((self.n_links, n_obs))
        pts = np.zeros((self.n_links, n_obs, 6))
        col = C.c_int()
        _ck(_lib.sirrt_get_distances_ex(self._h, q.ctypes.data, int(frame), d.ctypes.data, C.byref(col), pd.ctypes.data, pts.ctypes.data))
        return bool(col.value), d, pd, pts[:, :, :3].copy(), pts[:, :, 3:].copy()

    # -- batched path (numpy host arrays or torch cuda tensors)
    def safe_intervals_batch(self, q_soa, kmax: int = 8, want_bits: bool = True, want_table: bool = True):
        q = np.ascontiguousarray(q_soa, np.float32)
        assert q.ndim == 2 and q.shape[0] == self.dof
        N = q.shape[1]
        table = np.zeros((N, 1 + 2 * kmax), np.int32) if want_table else None
        bits = np.zeros((N, self.words), np.uint32) if want_bits else None
        near = C.c_uint64()
        _ck(_lib.sirrt_safe_intervals_batch(self._h, q.ctypes.data, N, _ptr(table), kmax, _ptr(bits), C.byref(near)))
        return table, bits, near.value

    def safe_intervals_batch_i16(self, q_soa, kmax: int = 8):
        """compact (int16) interval table, host buffers"""
        q = np.ascontiguousarray(q_soa, np.float32)
        assert q.ndim == 2 and q.shape[0] == self.dof
        table = np.zeros((q.shape[1], 1 + 2 * kmax), np.int16)
        near = C.c_uint64()
        _ck(_lib.sirrt_safe_intervals_batch_i16(self._h, q.ctypes.data, q.shape[1], table.ctypes.data, kmax, C.byref(near)))
        return table, near.value

    def safe_intervals_batch_i16_raw(self, q_ptr, N: int, table_ptr, kmax: int):
        near = C.c_uint64()
        _ck(_lib.sirrt_safe_intervals_batch_i16(self._h, q_ptr, N, table_ptr, kmax, C.byref(near)))
        return near.value

    def safe_intervals_batch_raw(self, q_ptr, N: int, table_ptr, kmax: int, bits_ptr):
        """Pointer-level call (device or pinned host memory): no allocation, no conversion."""
        near = C.c_uint64()
        _ck(_lib.sirrt_safe_intervals_batch(self._h, q_ptr, N, table_ptr, kmax, bits_ptr, C.byref(near)))
        return near.value

    def collision_bits_f64(self, q_aos, semantics: int = SEM_INTERVAL):
        q = np.ascontiguousarray(q_aos, np.float64).reshape(-1, self.dof)
        bits = np.zeros((q.shape[0], self.words), np.uint32)
        _ck(_lib.sirrt_collision_bits_f64(self._h, q.ctypes.data, q.shape[0], semantics, bits.ctypes.data))
        return bits

    def collision_bits_window_f64(self, q_aos, first_tile, n_tiles: int, semantics: int = SEM_FRAME):
        """Edge-table rows restricted to n_tiles 32-frame tiles from first_tile[i]; words outside the window carry no information."""
        q = np.ascontiguousarray(q_aos, np.float64).reshape(-1, self.dof)
        ft = np.ascontiguousarray(first_tile, np.int32).reshape(-1)
        assert ft.size == q.shape[0]
        bits = np.zeros((q.shape[0], self.words), np.uint32)
        _ck(_lib.sirrt_collision_bits_window_f64(self._h, q.ctypes.data, q.shape[0], semantics, ft.ctypes.data, int(n_tiles), bits.ctypes.data))
        return bits

    def check_frames_batch(self, q_soa, frames):
        q = np.ascontiguousarray(q_soa, np.float32)
        fr = np.ascontiguousarray(frames, np.int32)
        assert q.shape == (self.dof, fr.size)
        out = np.zeros(fr.size, np.uint8)
        _ck(_lib.sirrt_check_frames_batch(self._h, q.ctypes.data, fr.ctypes.data, fr.size, out.ctypes.data))
        return out

    def check_frames_batch_f64(self, q_aos, frames):
        q = np.ascontiguousarray(q_aos, np.float64).reshape(-1, self.dof)
        fr = np.ascontiguousarray(frames, np.int32)
        assert q.shape[0] == fr.size
        out = np.zeros(fr.size, np.uint8)
        _ck(_lib.sirrt_check_frames_batch_f64(self._h, q.ctypes.data, fr.ctypes.data, fr.size, out.ctypes.data))
        return out

    def edge_check_batch(self, q0, q1, t0, t1, vmax: float = 3.1415, want_coord: bool = True):
        q0 = np.ascontiguousarray(q0, np.float64).reshape(-1, self.dof)
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.dof)
        t0 = np.ascontiguousarray(t0, np.float64).reshape(-1)
        t1 = np.ascontiguousarray(t1, np.float64).reshape(-1)
        n = q0.shape[0]
        assert q1.shape[0] == n and t0.size == n and t1.size == n
        first = np.zeros(n, np.int32)
        coord = np.zeros((n, self.dof)) if want_coord else None
        _ck(_lib.sirrt_edge_check_batch(self._h, q0.ctypes.data, q1.ctypes.data, t0.ctypes.data, t1.ctypes.data, n, float(vmax),
                                        first.ctypes.data, _ptr(coord)))
        return first, coord

    def motion_check_batch(self, q0, q1, t0, t1, vmax: float, variant: int = 0):
        """SpaceTimeMotionValidator::checkMotion (SpaceTimeMotionValidator.cpp:22-74 / :76-127), times in seconds."""
        q0 = np.ascontiguousarray(q0, np.float64).reshape(-1, self.dof)
        q1 = np.ascontiguousarray(q1, np.float64).reshape(-1, self.dof)
        t0 = np.ascontiguousarray(t0, np.float64).reshape(-1)
        t1 = np.ascontiguousarray(t1, np.float64).reshape(-1)
        n = q0.shape[0]
        assert q1.shape[0] == n and t0.size == n and t1.size == n
        res = np.zeros(n, np.int32)
        frac = np.zeros(n)
        _ck(_lib.sirrt_motion_check_batch(self._h, q0.ctypes.data, q1.ctypes.data, t0.ctypes.data, t1.ctypes.data, n, float(vmax),
                                          int(variant), res.ctypes.data, frac.ctypes.data))
        return res, frac


class KnnIndex:
    """Exact FP64 nearest-neighbour store (replaces the nanoflann kd-tree of MSIRRT/src/Tree.cpp)."""

    def __init__(self, dim: int, capacity: int, device: int = 0):
        self._h = _vp()
        self.dim = dim
        _ck(_lib.sirrt_knn_create(dim, capacity, device, C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None):
            _lib.sirrt_knn_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __len__(self):
        n = C.c_int64()
        _ck(_lib.sirrt_knn_size(self._h, C.byref(n)))
        return n.value

    def clear(self):
        _ck(_lib.sirrt_knn_clear(self._h))

    @property
    def capacity(self) -> int:
        cap = C.c_int64()
        _ck(_lib.sirrt_knn_info(self._h, None, C.byref(cap), None))
        return cap.value

    def add(self, pts):
        pts = np.ascontiguousarray(pts, np.float64).reshape(-1, self.dim)
        _ck(_lib.sirrt_knn_add(self._h, pts.ctypes.data, pts.shape[0]))

    def nearest(self, queries):
        q = np.ascontiguousarray(queries, np.float64).reshape(-1, self.dim)
        idx = np.zeros(q.shape[0], np.int64)
        d2 = np.zeros(q.shape[0])
        _ck(_lib.sirrt_knn_nearest(self._h, q.ctypes.data, q.shape[0], idx.ctypes.data, d2.ctypes.data))
        return idx, d2

    def radius(self, query, r2: float, cap: int | None = None):
        q = np.ascontiguousarray(query, np.float64).reshape(self.dim)
        cap = len(self) if cap is None else cap
        idx = np.zeros(max(cap, 1), np.int64)
        n = C.c_int64()
        _ck(_lib.sirrt_knn_radius(self._h, q.ctypes.data, float(r2), idx.ctypes.data, cap, C.byref(n)))
        return idx[: min(n.value, cap)], n.value

    def radius_begin(self, query, r2: float):
        q = np.ascontiguousarray(query, np.float64).reshape(self.dim)
        _ck(_lib.sirrt_knn_radius_begin(self._h, q.ctypes.data, float(r2)))

    def radius_end(self, cap: int | None = None):
        cap = len(self) if cap is None else cap
        idx = np.zeros(max(cap, 1), np.int64)
        n = C.c_int64()
        _ck(_lib.sirrt_knn_radius_end(self._h, idx.ctypes.data, cap, C.byref(n)))
        return idx[: min(n.value, cap)], n.value

    def topk(self, queries, k: int):
        q = np.ascontiguousarray(queries, np.float64).reshape(-1, self.dim)
        idx = np.zeros((q.shape[0], k), np.int64)
        d2 = np.zeros((q.shape[0], k))
        _ck(_lib.sirrt_knn_topk(self._h, q.ctypes.data, q.shape[0], k, idx.ctypes.data, d2.ctypes.data))
        return idx, d2


class MultiScene:
    """One scene replica per device inside one process; batches are split into contiguous shards."""

    def __init__(self, pack: bytes, devices, parity_mode: int = PARITY_REFERENCE_EXACT):
        self._h = _vp()
        buf = bytes(pack)
        dev = (C.c_int * len(devices))(*devices)
        _ck(_lib.sirrt_multi_create(buf, len(buf), dev, len(devices), parity_mode, C.byref(self._h)))
        self.n_dev = len(devices)

    def close(self):
        if getattr(self, "_h", None):
            _lib.sirrt_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def safe_intervals_batch(self, q_soa, words: int, kmax: int = 8, want_bits: bool = True):
        q = np.ascontiguousarray(q_soa, np.float32)
        N = q.shape[1]
        table = np.zeros((N, 1 + 2 * kmax), np.int32)
        bits = np.zeros((N, words), np.uint32) if want_bits else None
        near = C.c_uint64()
        _ck(_lib.sirrt_multi_safe_intervals_batch(self._h, q.ctypes.data, N, table.ctypes.data, kmax, _ptr(bits), C.byref(near)))
        return table, bits, near.value


class CollisionManager:
    """Mirror of MDP::CollisionManager (CollisionManager.hpp:71-145) over the B200 backend.

    Method names, argument meaning and return shapes follow the reference; `scene_pack` is the
    packed SceneTask (tools/scene_pack.py).  Errors that the reference raises as exceptions are
    raised here as SirrtError / RuntimeError.
    """

    def __init__(self, scene_pack: bytes, device: int = 0, parity_mode: int = PARITY_REFERENCE_EXACT):
        self.scene = Scene(scene_pack, device, parity_mode)

    def get_planned_robot_limits(self):
        """CollisionManager.cpp:422-425: [(lower, upper)] as float32."""
        return [(float(lo), float(hi)) for lo, hi in self.scene.limits]

    def get_safe_intervals(self, robot_angles):
        """CollisionManager.cpp:481-547: inclusive, ascending (lo, hi) frame pairs."""
        return self.scene.safe_intervals(robot_angles)

    def check_collision_frame(self, robot_angles, frame: int) -> bool:
        """CollisionManager.cpp:338-341."""
        return self.scene.check_frame(robot_angles, frame)

    check_collision_frame_no_wrapper = check_collision_frame

    def check_collision(self, robot_angles, time: float) -> bool:
        """CollisionManager.cpp:330-336: frame = (int)(time * fps) evaluated in float32 (QUIRK 13)."""
        frame = int(np.float32(time) * np.float32(self.scene.fps))
        return self.scene.check_frame(robot_angles, frame)

    def get_distances(self, robot_angles, time: float):
        """CollisionManager.cpp:382-420: (is_collision, per-link min distance); frame = (int)(time*fps)."""
        frame = int(np.float32(time) * np.float32(self.scene.fps))
        return self.scene.get_distances(robot_angles, frame)

    def check_robot_selfcollision(self, robot_angles) -> bool:
        return False  # CollisionManager.cpp:445-447: disabled in the reference

    def check_base_joints_collisiion(self, robot_angles, frame: int) -> bool:
        return False  # CollisionManager.cpp:476-479

    def get_goal_frame_low_bound(self) -> int:
        """CollisionManager.cpp:427-443: scan check_collision_frame(end_configuration, f) from the last
        frame down; first collision at f -> f + 1; raises if the last frame collides; 0 if none."""
        T = self.scene.n_frames
        bits = self.scene.collision_bits_f64(self.scene.end_configuration[None, :], SEM_FRAME)[0]
        col = np.unpackbits(bits.view(np.uint8), bitorder="little")[:T]
        hit = np.nonzero(col)[0]
        if hit.size == 0:
            return 0
        if hit[-1] == T - 1:
            raise RuntimeError("there is collision in goal state at the last frame!")
        return int(hit[-1]) + 1
