"""ctypes binding of the CPU oracle (TEST INFRASTRUCTURE ONLY -- see oracle/sirrt_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_DIR = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_DIR, "libsirrt_oracle.so")

REFERENCE_EXACT, FIXED = 0, 1
SEM_INTERVAL, SEM_FRAME = 0, 1


def build(force: bool = False) -> str:
    src = os.path.join(_DIR, "sirrt_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _DIR, "-s"])
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB):
            build()
        L = C.CDLL(_LIB)
        vp, i32, i64, f64 = C.c_void_p, C.c_int, C.c_int64, C.c_double
        L.oracle_create.restype = vp
        L.oracle_create.argtypes = [vp, C.c_size_t, i32]
        L.oracle_destroy.argtypes = [vp]
        L.oracle_last_error.restype = C.c_char_p
        L.oracle_info.argtypes = [vp] + [vp] * 5
        L.oracle_counts.argtypes = [vp] + [vp] * 3
        L.oracle_fk.argtypes = [vp, vp, vp]
        L.oracle_robot_obstacle_fk.argtypes = [vp, i32, i32, vp]
        L.oracle_collision_bits.argtypes = [vp, vp, i32, vp, f64, vp]
        L.oracle_safe_intervals.argtypes = [vp, vp, vp, i32]
        L.oracle_check_frame.argtypes = [vp, vp, i32]
        L.oracle_pair_distance.restype = f64
        L.oracle_pair_distance.argtypes = [vp, vp, i32, i32, i32]
        L.oracle_edge_check.argtypes = [vp, vp, vp, f64, f64, f64, vp]
        L.oracle_state_valid.argtypes = [vp, vp, f64]
        L.oracle_state_valid.restype = i32
        L.oracle_motion_check.argtypes = [vp, vp, vp, f64, f64, f64, i32, vp]
        L.oracle_motion_check.restype = i32
        L.oracle_get_distances.argtypes = [vp, vp, i32, vp]
        L.oracle_bvh_build.argtypes = [vp]
        L.oracle_set_fast.argtypes = [i32]
        L.oracle_get_fast.restype = i32
        L.oracle_safe_intervals_batch.argtypes = [vp, vp, i64, vp, i32, vp, i32, i32]
        L.oracle_check_frames_batch.argtypes = [vp, vp, vp, i64, vp, i32]
        L.oracle_knn_nearest.argtypes = [vp, i64, i32, vp, i64, vp, vp]
        L.oracle_knn_radius.restype = i64
        L.oracle_knn_radius.argtypes = [vp, i64, i32, vp, f64, vp, i64]
        L.oracle_gjk_distance.restype = f64
        L.oracle_gjk_distance.argtypes = [vp, i32, vp, i32]
        L.oracle_seg_box_dist2.restype = f64
        L.oracle_seg_box_dist2.argtypes = [vp, vp, vp]
        L.oracle_seg_seg_dist2.restype = f64
        L.oracle_seg_seg_dist2.argtypes = [vp, vp, vp, vp]
        L.oracle_pt_seg_dist2.restype = f64
        L.oracle_pt_seg_dist2.argtypes = [vp, vp, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def set_fast(on: bool):
    """Baseline mode of the TIMED CPU arm (boolean early exits + AABB tree); results are bit-identical."""
    lib().oracle_set_fast(int(bool(on)))


class Oracle:
    def __init__(self, pack: bytes, parity_mode: int = REFERENCE_EXACT):
        self._buf = bytes(pack)
        self.h = lib().oracle_create(self._buf, len(self._buf), parity_mode)
        if not self.h:
            raise RuntimeError(lib().oracle_last_error().decode())
        v = [C.c_int() for _ in range(4)]
        lib().oracle_info(self.h, *[C.byref(x) for x in v], None)
        self.dof, self.n_links, self.n_frames, self.fps = [x.value for x in v]
        self.limits = np.zeros((self.dof, 2), np.float32)
        lib().oracle_info(self.h, None, None, None, None, _p(self.limits))
        c = [C.c_int() for _ in range(3)]
        lib().oracle_counts(self.h, *[C.byref(x) for x in c])
        self.n_obstacles, self.n_robot_obstacles, self.n_robot_obstacle_links = [x.value for x in c]
        self.words = (self.n_frames + 31) // 32

    def close(self):
        if self.h:
            lib().oracle_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def fk(self, q):
        q = _d(q)
        out = np.zeros((self.n_links, 12))
        lib().oracle_fk(self.h, _p(q), _p(out))
        return out[:, :9].reshape(-1, 3, 3), out[:, 9:]

    def robot_obstacle_fk(self, robot, frame):
        out = np.zeros((64, 12))
        if lib().oracle_robot_obstacle_fk(self.h, robot, frame, _p(out)):
            raise ValueError("bad robot/frame")
        return out

    def collision_bits(self, q, semantics=SEM_INTERVAL, tol=1e-6):
        q = _d(q)
        bits = np.zeros(self.words, np.uint32)
        near = C.c_int64()
        lib().oracle_collision_bits(self.h, _p(q), semantics, _p(bits), tol, C.byref(near))
        return bits, near.value

    def safe_intervals(self, q, cap=4096):
        q = _d(q)
        buf = np.zeros(2 * cap, np.int32)
        n = lib().oracle_safe_intervals(self.h, _p(q), _p(buf), cap)
        return [(int(buf[2 * i]), int(buf[2 * i + 1])) for i in range(min(n, cap))]

    def check_frame(self, q, frame) -> bool:
        q = _d(q)
        return bool(lib().oracle_check_frame(self.h, _p(q), int(frame)))

    def pair_distance(self, q, link, obstacle_id, frame) -> float:
        q = _d(q)
        return lib().oracle_pair_distance(self.h, _p(q), link, obstacle_id, frame)

    def edge_check(self, q0, q1, t0, t1, vmax=3.1415):
        q0, q1 = _d(q0), _d(q1)
        cc = np.zeros(self.dof)
        s = lib().oracle_edge_check(self.h, _p(q0), _p(q1), float(t0), float(t1), float(vmax), _p(cc))
        return s, cc

    def state_valid(self, q, time_s) -> bool:
        q = _d(q)
        return bool(lib().oracle_state_valid(self.h, _p(q), float(time_s)))

    def motion_check(self, q0, q1, t0, t1, vmax, variant=0):
        q0, q1 = _d(q0), _d(q1)
        frac = np.zeros(1)
        r = lib().oracle_motion_check(self.h, _p(q0), _p(q1), float(t0), float(t1), float(vmax), int(variant), _p(frac))
        return r, float(frac[0])

    def get_distances(self, q, frame):
        q = _d(q)
        out = np.zeros(self.n_links)
        rc = lib().oracle_get_distances(self.h, _p(q), int(frame), _p(out))
        return rc, out

    def bvh_build(self) -> int:
        return lib().oracle_bvh_build(self.h)

    def safe_intervals_batch(self, q_soa, kmax=8, want_bits=True, use_bvh=False, n_threads=0):
        q = np.ascontiguousarray(q_soa, np.float32)
        assert q.shape[0] == self.dof
        N = q.shape[1]
        table = np.zeros((N, 1 + 2 * kmax), np.int32)
        bits = np.zeros((N, self.words), np.uint32) if want_bits else None
        rc = lib().oracle_safe_intervals_batch(self.h, _p(q), N, _p(table), kmax, _p(bits) if want_bits else None,
                                               int(use_bvh), n_threads)
        if rc:
            raise RuntimeError(lib().oracle_last_error().decode())
        return table, bits

    def check_frames_batch(self, q_soa, frames, n_threads=0):
        q = np.ascontiguousarray(q_soa, np.float32)
        fr = np.ascontiguousarray(frames, np.int32)
        out = np.zeros(q.shape[1], np.uint8)
        lib().oracle_check_frames_batch(self.h, _p(q), _p(fr), q.shape[1], _p(out), n_threads)
        return out


def knn_nearest(pts, queries):
    pts, queries = _d(pts), _d(queries)
    nq = queries.shape[0]
    idx = np.zeros(nq, np.int64)
    d2 = np.zeros(nq)
    lib().oracle_knn_nearest(_p(pts), pts.shape[0], pts.shape[1], _p(queries), nq, _p(idx), _p(d2))
    return idx, d2


def knn_radius(pts, query, r2):
    pts, query = _d(pts), _d(query)
    idx = np.zeros(pts.shape[0], np.int64)
    n = lib().oracle_knn_radius(_p(pts), pts.shape[0], pts.shape[1], _p(query), float(r2), _p(idx), idx.size)
    return idx[:n]


def gjk_distance(a, b) -> float:
    a, b = _d(a), _d(b)
    return lib().oracle_gjk_distance(_p(a), a.shape[0], _p(b), b.shape[0])


def seg_box_dist2(a, b, half) -> float:
    a, b, half = _d(a), _d(b), _d(half)
    return lib().oracle_seg_box_dist2(_p(a), _p(b), _p(half))


def seg_seg_dist2(p1, q1, p2, q2) -> float:
    p1, q1, p2, q2 = _d(p1), _d(q1), _d(p2), _d(q2)
    return lib().oracle_seg_seg_dist2(_p(p1), _p(q1), _p(p2), _p(q2))


def pt_seg_dist2(p, a, b) -> float:
    p, a, b = _d(p), _d(a), _d(b)
    return lib().oracle_pt_seg_dist2(_p(p), _p(a), _p(b))
