"""oracle2.py -- TEST INFRASTRUCTURE ONLY: a second, code-independent CPU oracle (NumPy / SciPy).

Purpose: every bit-exact claim of the CUDA path is "GPU == oracle/sirrt_oracle.c",
and that C file restates un-vendored coal arithmetic (GJK + closed forms) that no reference output pins on the COLLIDING
side.  This module answers the same questions -- link transforms, signed pair distances, collision booleans per
(configuration, frame) cell, safe intervals -- by DIFFERENT numerical routes and shares no code with sirrt_oracle.c
(it does not load the C library, has its own parser of the pack format of include/sirrt_pack.h, and uses no GJK):

  forward kinematics     4x4 homogeneous products (RobotObstacleFCL.cpp:287-386: base pose, non-fixed joints only,
                         rotation about local +Z, link collision origin, first two collision links dropped :370);
                         fk_urdf() additionally walks a URDF file the textbook way (KDL-like: <axis>, fixed-joint
                         origins honoured) -- the reference's own cross-check idea (RobotObstacleFCL.cpp:372-378)
  capsule vs sphere      point-segment distance by projection
  capsule vs box         golden-section minimisation over the segment parameter of the point-to-box distance (convex)
  capsule vs capsule     golden-section over one segment parameter of the point-to-segment distance (convex)
  hull vs box / hull     BOOLEAN by LP feasibility (scipy.optimize.linprog: lambda, mu >= 0, sum = 1, A lambda = B mu);
                         DISTANCE by a small QP solved as NNLS with the two simplex constraints as heavily weighted rows
                         (scipy.optimize.nnls, Lawson-Hanson active set); penetration depth bounded from above by the
                         smallest overlap along the face normals of both bodies (scipy.spatial.ConvexHull)
  hull vs sphere/capsule the same QP (a capsule core is a 2-vertex hull), radii subtracted
  predicate              collision <=> distance - 1e-21 <= 1e-12 (CollisionManager.cpp:611,654 security_margin; coal's
                         collision_distance_threshold) -- identical outside any band wider than the solvers' 1e-9
  robot obstacles        configuration at a frame by the reference's interpolation (RobotObstacleFCL.cpp:393-429);
                         QUIRK 7 (CollisionManager.cpp:97-104,206-212,263-268,628,691,721): REFERENCE_EXACT tests the
                         mis-indexed functor geometry behind coal's AABB broadphase gate, FIXED the true geometry
  intervals              complement runs of the colliding frames (CollisionManager.cpp:507-546)

Only tests/ may import this module.
"""
from __future__ import annotations

import math
import struct
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import numpy as np
from scipy.optimize import linprog, nnls
from scipy.spatial import ConvexHull

REFERENCE_EXACT, FIXED = 0, 1
SEM_INTERVAL, SEM_FRAME = 0, 1
THRESHOLD = 1e-12 + 1e-21  # coal collision_distance_threshold + security_margin
CAPSULE, HULL, BOX, SPHERE = "capsule", "hull", "box", "sphere"


# ------------------------------------------------------------------------------------------------ pack parser
@dataclass
class Geom:
    kind: str
    radius: float = 0.0       # capsule / sphere
    half_len: float = 0.0     # capsule core: local (0, 0, +-half_len)
    half: np.ndarray = field(default_factory=lambda: np.zeros(3))   # box half extents
    verts: np.ndarray = field(default_factory=lambda: np.zeros((0, 3)))  # hull vertices (local)
    _faces: Optional[np.ndarray] = None
    _aabb: Optional[Tuple[np.ndarray, np.ndarray]] = None
    _ball: Optional[Tuple[np.ndarray, float]] = None

    def core_points(self) -> np.ndarray:
        """vertices of the convex core (local frame): hull vertices, box corners, capsule end points, sphere centre"""
        if self.kind == HULL:
            return self.verts
        if self.kind == BOX:
            s = np.array([[a, b, c] for a in (-1, 1) for b in (-1, 1) for c in (-1, 1)], float)
            return s * self.half
        if self.kind == CAPSULE:
            return np.array([[0, 0, -self.half_len], [0, 0, self.half_len]], float)
        return np.zeros((1, 3))

    def core_radius(self) -> float:
        return self.radius if self.kind in (CAPSULE, SPHERE) else 0.0

    def face_normals(self) -> np.ndarray:
        """outward unit face normals of the core polytope (local frame), for the penetration-depth bound"""
        if self._faces is None:
            if self.kind == BOX:
                self._faces = np.vstack([np.eye(3), -np.eye(3)])
            elif self.kind == HULL and len(self.verts) >= 4:
                try:
                    eq = ConvexHull(self.verts).equations[:, :3]
                    self._faces = np.unique(np.round(eq, 9), axis=0)
                except Exception:
                    self._faces = np.zeros((0, 3))
            else:
                self._faces = np.zeros((0, 3))
        return self._faces

    def local_aabb(self) -> Tuple[np.ndarray, np.ndarray]:
        """coal computeLocalAABB of the shape (inflated by its radius)"""
        if self._aabb is None:
            p = self.core_points()
            r = self.core_radius()
            self._aabb = (p.min(0) - r, p.max(0) + r)
        return self._aabb

    def local_ball(self) -> Tuple[np.ndarray, float]:
        """a bounding ball of the inflated shape in the local frame (vertex mean, largest vertex distance + radius)"""
        if self._ball is None:
            p = self.core_points()
            c = p.mean(0)
            self._ball = (c, float(np.linalg.norm(p - c, axis=1).max()) + self.core_radius())
        return self._ball


@dataclass
class RobotModel:
    dof: int
    n_strip: int
    base_pos: np.ndarray
    base_quat: np.ndarray
    limits: np.ndarray
    joint_t: np.ndarray     # [nj, 3]
    joint_R: np.ndarray     # [nj, 3, 3]
    shapes: List[Geom] = field(default_factory=list)
    shape_joints_before: List[int] = field(default_factory=list)
    shape_lo: List[np.ndarray] = field(default_factory=list)   # 4x4 link collision origin


class _Reader:
    def __init__(self, b: bytes):
        self.b, self.o = b, 0

    def take(self, fmt: str):
        v = struct.unpack_from("<" + fmt, self.b, self.o)
        self.o += struct.calcsize("<" + fmt)
        return v

    def array(self, dtype, n: int) -> np.ndarray:
        a = np.frombuffer(self.b, dtype=dtype, count=n, offset=self.o).copy()
        self.o += a.nbytes
        return a

    def align8(self):
        self.o = (self.o + 7) & ~7


def quat_to_matrix(x: float, y: float, z: float, w: float) -> np.ndarray:
    """Eigen::Quaternion::toRotationMatrix: no normalisation (the reader does not normalise either)"""
    tx, ty, tz = 2 * x, 2 * y, 2 * z
    return np.array([[1 - (ty * y + tz * z), tx * y - tz * w, tx * z + ty * w],
                     [tx * y + tz * w, 1 - (tx * x + tz * z), ty * z - tx * w],
                     [tx * z - ty * w, ty * z + tx * w, 1 - (tx * x + ty * y)]])


def _read_robot(r: _Reader) -> RobotModel:
    dof, nj, ns, strip = r.take("4i")
    base_pos = np.array(r.take("3f"), np.float32).astype(float)
    base_quat = np.array(r.take("4f"), np.float32).astype(float)
    r.take("i")
    limits = r.array(np.float32, 2 * dof).reshape(dof, 2)
    r.align8()
    jt, jR = np.zeros((nj, 3)), np.zeros((nj, 3, 3))
    for k in range(nj):
        jt[k] = r.array(np.float64, 3)
        jR[k] = r.array(np.float64, 9).reshape(3, 3)
    m = RobotModel(dof, strip, base_pos, base_quat, limits, jt, jR)
    for _ in range(ns):
        typ, jb, nv, _pad = r.take("4i")
        lo_t = r.array(np.float64, 3)
        lo_R = r.array(np.float64, 9).reshape(3, 3)
        radius, length = r.take("2d")
        verts = r.array(np.float32, 3 * nv).reshape(nv, 3).astype(float)
        r.align8()
        lo = np.eye(4)
        lo[:3, :3], lo[:3, 3] = lo_R, lo_t
        g = Geom(CAPSULE, radius=radius, half_len=length / 2) if typ == 0 else Geom(HULL, verts=verts)
        m.shapes.append(g)
        m.shape_joints_before.append(jb)
        m.shape_lo.append(lo)
    return m


@dataclass
class Obstacle:
    geom: Geom
    is_static: bool
    poses: np.ndarray  # float32 -> float64 [n, 7]


@dataclass
class RobotObstacle:
    model: RobotModel
    is_static: bool
    times: np.ndarray
    angles: np.ndarray


class Pack:
    def __init__(self, blob: bytes):
        r = _Reader(bytes(blob))
        magic = r.take("8s")[0]
        if magic != b"SIRRTPK1":
            raise ValueError("not a scene pack")
        _ver, _flags, self.n_frames, self.fps = r.take("2I2i")
        self.start_time = r.take("d")[0]
        self.dof, n_obs, n_rob, _res = r.take("4i")
        r.take("Q")
        self.start_cfg = r.array(np.float64, self.dof)
        self.end_cfg = r.array(np.float64, self.dof)
        self.max_vel = r.array(np.float64, self.dof)
        self.robot = _read_robot(r)
        self.obstacles: List[Obstacle] = []
        for _ in range(n_obs):
            typ, is_static, n_poses, _pad = r.take("4i")
            dims = np.array(r.take("4f"), np.float32).astype(float)
            poses = r.array(np.float32, 7 * n_poses).reshape(n_poses, 7).astype(float)
            r.align8()
            g = Geom(BOX, half=dims[:3] / 2) if typ == 0 else Geom(SPHERE, radius=dims[0])
            self.obstacles.append(Obstacle(g, bool(is_static), poses))
        self.robot_obstacles: List[RobotObstacle] = []
        for _ in range(n_rob):
            is_static, n_states = r.take("2i")
            model = _read_robot(r)
            times = r.array(np.float64, n_states)
            angles = r.array(np.float64, n_states * model.dof).reshape(n_states, model.dof)
            self.robot_obstacles.append(RobotObstacle(model, bool(is_static), times, angles))


# ------------------------------------------------------------------------------------------------ kinematics
def rot_z(a: float) -> np.ndarray:
    c, s = math.cos(a), math.sin(a)
    m = np.eye(4)
    m[0, 0], m[0, 1], m[1, 0], m[1, 1] = c, -s, s, c
    return m


def fk_chain(m: RobotModel, q) -> List[np.ndarray]:
    """world transforms (4x4) of EVERY collision shape of the robot, chain order (RobotObstacleFCL.cpp:287-368)"""
    T = np.eye(4)
    T[:3, :3] = quat_to_matrix(*m.base_quat)
    T[:3, 3] = m.base_pos
    out, done = [], 0
    order = sorted(range(len(m.shapes)), key=lambda i: (m.shape_joints_before[i], i))
    assert order == list(range(len(m.shapes))), "shapes are stored in chain order"
    for i in range(len(m.shapes)):
        while done < m.shape_joints_before[i]:
            J = np.eye(4)
            J[:3, :3], J[:3, 3] = m.joint_R[done], m.joint_t[done]
            T = T @ J @ rot_z(float(q[done]) if done < len(q) else 0.0)
            done += 1
        out.append(T @ m.shape_lo[i])
    return out


def _rpy(r: float, p: float, y: float) -> np.ndarray:
    cr, sr, cp, sp_, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return np.array([[cy * cp, cy * sp_ * sr - sy * cr, cy * sp_ * cr + sy * sr],
                     [sy * cp, sy * sp_ * sr + cy * cr, sy * sp_ * cr - cy * sr],
                     [-sp_, cp * sr, cp * cr]])


def fk_urdf(urdf_path: str, q, base_pos=(0, 0, 0), base_quat=(0, 0, 0, 1)):
    """Textbook URDF forward kinematics straight from the XML numbers (what KDL does; the reference compares its own FK
    with KDL's at RobotObstacleFCL.cpp:372-378): every joint origin is applied, revolute joints rotate about their <axis>.
    Returns {link name: 4x4 world transform of the link FRAME} following the first child chain from the root link."""
    root = ET.parse(urdf_path).getroot()
    joints = {}
    children = set()
    for j in root.findall("joint"):
        parent, child = j.find("parent").get("link"), j.find("child").get("link")
        o = j.find("origin")
        xyz = [float(v) for v in (o.get("xyz", "0 0 0") if o is not None else "0 0 0").split()]
        rpy = [float(v) for v in (o.get("rpy", "0 0 0") if o is not None else "0 0 0").split()]
        ax = j.find("axis")
        axis = [float(v) for v in (ax.get("xyz") if ax is not None else "1 0 0").split()]
        joints.setdefault(parent, []).append((child, j.get("type"), xyz, rpy, axis))
        children.add(child)
    links = [l.get("name") for l in root.findall("link")]
    cur = next(l for l in links if l not in children)
    T = np.eye(4)
    T[:3, :3] = quat_to_matrix(*base_quat)
    T[:3, 3] = base_pos
    out, k = {cur: T.copy()}, 0
    while cur in joints:
        child, typ, xyz, rpy, axis = joints[cur][0]
        J = np.eye(4)
        J[:3, :3], J[:3, 3] = _rpy(*rpy), xyz
        T = T @ J
        if typ in ("revolute", "continuous"):
            a = np.array(axis, float)
            a /= np.linalg.norm(a)
            ang = float(q[k])
            k += 1
            K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
            Rm = np.eye(4)
            Rm[:3, :3] = np.eye(3) + math.sin(ang) * K + (1 - math.cos(ang)) * (K @ K)
            T = T @ Rm
        out[child] = T.copy()
        cur = child
    return out


# ------------------------------------------------------------------------------------------------ distances
_GR = (math.sqrt(5.0) - 1.0) / 2.0


def _golden(fun, n_iter: int = 90):
    """minimum of a convex function on [0, 1] by golden-section search; returns (t, f(t)) including the end points"""
    lo, hi = 0.0, 1.0
    x1, x2 = hi - _GR * (hi - lo), lo + _GR * (hi - lo)
    f1, f2 = fun(x1), fun(x2)
    for _ in range(n_iter):
        if f1 < f2:
            hi, x2, f2 = x2, x1, f1
            x1 = hi - _GR * (hi - lo)
            f1 = fun(x1)
        else:
            lo, x1, f1 = x1, x2, f2
            x2 = lo + _GR * (hi - lo)
            f2 = fun(x2)
    cands = [(fun(0.0), 0.0), (fun(1.0), 1.0), (f1, x1), (f2, x2)]
    f, t = min(cands)
    return t, f


def point_box_dist(p: np.ndarray, half: np.ndarray) -> float:
    e = np.maximum(np.abs(p) - half, 0.0)
    return float(math.sqrt(float(e @ e)))


def segment_box_distance(a: np.ndarray, b: np.ndarray, half: np.ndarray) -> float:
    """a, b in the box frame.  phi(t) = dist(a + t (b - a), box) is convex -> scalar minimisation"""
    d = b - a
    _, f = _golden(lambda t: point_box_dist(a + t * d, half))
    return f


def point_segment_distance(p: np.ndarray, a: np.ndarray, b: np.ndarray) -> Tuple[float, np.ndarray]:
    d = b - a
    den = float(d @ d)
    t = 0.0 if den <= 0 else min(1.0, max(0.0, float((p - a) @ d) / den))
    c = a + t * d
    return float(np.linalg.norm(p - c)), c


def segment_segment_distance(a1, b1, a2, b2) -> float:
    d = b1 - a1
    _, f = _golden(lambda t: point_segment_distance(a1 + t * d, a2, b2)[0])
    return f


def hull_intersect_lp(A: np.ndarray, B: np.ndarray) -> bool:
    """conv(A) and conv(B) share a point  <=>  the LP {lambda, mu >= 0, sum lambda = sum mu = 1, A^T lambda = B^T mu} is
    feasible (world-frame vertex arrays [n,3]).  Coordinates are centred on the mean of B and scaled to unit size.  A
    grazing contact can leave the simplex runs undecided ("unknown" status); then the phase-1 form of the same LP
    (always feasible: minimise the L1 norm of the slack s in A^T lambda - B^T mu = s) decides: intersecting iff its
    optimum is zero to the solver's 1e-9."""
    c = B.mean(0)
    sc = max(float(np.abs(A - c).max()), float(np.abs(B - c).max()), 1e-9)
    A, B = (A - c) / sc, (B - c) / sc
    na, nb = len(A), len(B)
    Aeq = np.zeros((5, na + nb))
    Aeq[:3, :na], Aeq[:3, na:] = A.T, -B.T
    Aeq[3, :na] = 1.0
    Aeq[4, na:] = 1.0
    beq = np.array([0, 0, 0, 1.0, 1.0])
    for method, opts in (("highs", {"primal_feasibility_tolerance": 1e-10, "dual_feasibility_tolerance": 1e-10, "presolve": False}), ("highs-ds", {})):
        res = linprog(np.zeros(na + nb), A_eq=Aeq, b_eq=beq, bounds=(0, None), method=method, options=opts)
        if res.status == 0:
            return True
        if res.status == 2:
            return False
    A1 = np.hstack([Aeq, np.vstack([np.eye(3), np.zeros((2, 3))]), np.vstack([-np.eye(3), np.zeros((2, 3))])])
    cost = np.concatenate([np.zeros(na + nb), np.ones(6)])
    res = linprog(cost, A_eq=A1, b_eq=beq, bounds=(0, None), method="highs")
    if res.status != 0:
        raise RuntimeError(f"linprog (phase 1): {res.message}")
    return bool(res.fun <= 1e-9)


def hull_distance_qp(A: np.ndarray, B: np.ndarray, w: float = 1e4) -> Tuple[float, np.ndarray, np.ndarray]:
    """min |A^T lambda - B^T mu| over the two simplices, solved as NNLS with the simplex constraints as rows of weight w
    (coordinates are scaled to unit size first, so the constraint violation is O(1/w^2)); returns (distance, pa, pb)"""
    c = B.mean(0)
    s = max(float(np.abs(A - c).max()), float(np.abs(B - c).max()), 1e-9)
    As, Bs = (A - c) / s, (B - c) / s
    na, nb = len(A), len(B)
    M = np.zeros((5, na + nb))
    M[:3, :na], M[:3, na:] = As.T, -Bs.T
    M[3, :na] = w
    M[4, na:] = w
    x, _ = nnls(M, np.array([0, 0, 0, w, w]), maxiter=20 * (na + nb))
    la, mu = x[:na] / x[:na].sum(), x[na:] / x[na:].sum()
    pa, pb = la @ A, mu @ B
    return float(np.linalg.norm(pa - pb)), pa, pb


def penetration_bound(A: np.ndarray, nA: np.ndarray, B: np.ndarray, nB: np.ndarray) -> float:
    """upper bound of the penetration depth of two intersecting polytopes: the smallest translation that separates them
    along a face normal of either body (world-frame vertices and outward normals)"""
    best = math.inf
    if len(nA):
        best = min(best, float(((A @ nA.T).max(0) - (B @ nA.T).min(0)).min()))
    if len(nB):
        best = min(best, float(((B @ nB.T).max(0) - (A @ nB.T).min(0)).min()))
    return best


# ------------------------------------------------------------------------------------------------ the oracle
@dataclass
class Posed:
    geom: Geom
    T: np.ndarray  # 4x4 world

    def world_core(self) -> np.ndarray:
        return self.geom.core_points() @ self.T[:3, :3].T + self.T[:3, 3]

    def world_normals(self) -> np.ndarray:
        return self.geom.face_normals() @ self.T[:3, :3].T

    def bound(self) -> Tuple[np.ndarray, float]:
        c, r = self.geom.local_ball()
        return self.T[:3, :3] @ c + self.T[:3, 3], r

    def coal_aabb(self) -> Tuple[np.ndarray, np.ndarray]:
        """coal CollisionObject::computeAABB: identity rotation -> translated local AABB, else a cube of the local AABB's
        half diagonal around the rotated local centre"""
        lo, hi = self.geom.local_aabb()
        R, t = self.T[:3, :3], self.T[:3, 3]
        if np.array_equal(R, np.eye(3)):
            return lo + t, hi + t
        c = (lo + hi) / 2
        rad = float(np.linalg.norm(lo - c))
        cw = R @ c + t
        return cw - rad, cw + rad


def signed_distance(a: Posed, b: Posed) -> float:
    """signed distance of two posed shapes: > 0 separated, <= 0 touching / penetrating.  For penetrating polytope cores
    the value is minus an UPPER bound of the depth (enough to say how far inside the collision a pair is)."""
    ka, kb = a.geom.kind, b.geom.kind
    if ka != CAPSULE and kb == CAPSULE:
        return signed_distance(b, a)
    if ka == CAPSULE and kb in (SPHERE, BOX, CAPSULE):
        p = a.world_core()
        if kb == SPHERE:
            return point_segment_distance(b.T[:3, 3], p[0], p[1])[0] - a.geom.radius - b.geom.radius
        if kb == BOX:
            Rb, tb = b.T[:3, :3], b.T[:3, 3]
            return segment_box_distance(Rb.T @ (p[0] - tb), Rb.T @ (p[1] - tb), b.geom.half) - a.geom.radius
        pb = b.world_core()
        return segment_segment_distance(p[0], p[1], pb[0], pb[1]) - a.geom.radius - b.geom.radius
    A, B = a.world_core(), b.world_core()
    rsum = a.geom.core_radius() + b.geom.core_radius()
    # cheap certificate of separation: a plane normal to the line between the vertex means (no solver needed far apart)
    n = B.mean(0) - A.mean(0)
    nn = float(np.linalg.norm(n))
    if nn > 0:
        gap = float((B @ n).min() - (A @ n).max()) / nn
        if gap - rsum > 0.05:
            return gap - rsum  # a LOWER bound of the distance, only ever returned when it is centimetres from contact
    d, _, _ = hull_distance_qp(A, B)
    if d > 1e-7:
        return d - rsum
    # cores touch or intersect: LP decides, the face-normal bound says how deep
    if len(A) >= 4 or len(B) >= 4:
        if not hull_intersect_lp(A, B):
            return d - rsum
        na, nb = a.world_normals(), b.world_normals()
        if len(na) + len(nb) == 0:
            return -rsum
        return -penetration_bound(A, na, B, nb) - rsum
    return d - rsum


def closest_points(a: Posed, b: Posed) -> Tuple[float, np.ndarray, np.ndarray]:
    """(signed distance, closest point on a's surface, closest point on b's surface) for SEPARATED shapes -- coal's
    DistanceResult::nearest_points, collected per pair by the reference's DistanceCallback (CollisionManager.cpp:568-598).
    Core closest points: projection / scalar minimisation for capsule pairs, the NNLS-QP for polytopes; radii are then
    stepped off along the line between them."""
    ka, kb = a.geom.kind, b.geom.kind
    ra, rb = a.geom.core_radius(), b.geom.core_radius()
    if ka == CAPSULE and kb == SPHERE:
        p = a.world_core()
        _, ca = point_segment_distance(b.T[:3, 3], p[0], p[1])
        cb = b.T[:3, 3].copy()
    elif ka == CAPSULE and kb == BOX:
        p = a.world_core()
        Rb, tb = b.T[:3, :3], b.T[:3, 3]
        la, lb = Rb.T @ (p[0] - tb), Rb.T @ (p[1] - tb)
        t, _ = _golden(lambda u: point_box_dist(la + u * (lb - la), b.geom.half), 120)
        ps = la + t * (lb - la)
        ca = p[0] + t * (p[1] - p[0])
        cb = Rb @ np.clip(ps, -b.geom.half, b.geom.half) + tb
    elif ka == CAPSULE and kb == CAPSULE:
        p, q = a.world_core(), b.world_core()
        s, _ = _golden(lambda u: point_segment_distance(p[0] + u * (p[1] - p[0]), q[0], q[1])[0], 120)
        ca = p[0] + s * (p[1] - p[0])
        _, cb = point_segment_distance(ca, q[0], q[1])
    else:
        _, ca, cb = hull_distance_qp(a.world_core(), b.world_core())
    n = cb - ca
    ln = float(np.linalg.norm(n))
    n = n / ln if ln > 0 else n
    return ln - ra - rb, ca + ra * n, cb - rb * n


def collides(a: Posed, b: Posed) -> Tuple[bool, float]:
    """(collision boolean, signed distance).  Polytope-polytope pairs without radii are decided by LP feasibility alone."""
    d = signed_distance(a, b)
    return d <= THRESHOLD, d


class Oracle2:
    def __init__(self, pack: bytes, parity_mode: int = REFERENCE_EXACT):
        self.p = Pack(pack)
        self.parity_mode = parity_mode
        self.dof, self.n_frames, self.fps = self.p.dof, self.p.n_frames, self.p.fps
        self.n_links = len(self.p.robot.shapes) - self.p.robot.n_strip
        self.C = len(self.p.obstacles)
        # QUIRK 7 bookkeeping: all shapes of all robot obstacles in one list (collision_pairs columns), and the
        # base-stripped links (obstacle ids C, C+1, ...)
        self.rob_all_shapes: List[Geom] = []
        self.rob_links: List[Tuple[int, int]] = []  # (robot, shape index within robot) per stripped link id
        for r, ro in enumerate(self.p.robot_obstacles):
            self.rob_all_shapes += ro.model.shapes
            for s in range(ro.model.n_strip, len(ro.model.shapes)):
                self.rob_links.append((r, s))
        self._rob_cache = {}

    # a1
    def fk(self, q):
        Ts = fk_chain(self.p.robot, q)[self.p.robot.n_strip:]
        return np.array([T[:3, :3] for T in Ts]), np.array([T[:3, 3] for T in Ts])

    def links(self, q) -> List[Posed]:
        m = self.p.robot
        return [Posed(g, T) for g, T in list(zip(m.shapes, fk_chain(m, q)))[m.n_strip:]]

    def obstacle(self, c: int, f: int) -> Posed:
        o = self.p.obstacles[c]
        pose = o.poses[0 if o.is_static else f]
        T = np.eye(4)
        T[:3, :3] = quat_to_matrix(*pose[3:7])
        T[:3, 3] = pose[:3]
        return Posed(o.geom, T)

    # a7
    def robot_obstacle_angles(self, r: int, f: int) -> np.ndarray:
        ro = self.p.robot_obstacles[r]
        time = f / self.fps + self.p.start_time
        tt, aa = ro.times, ro.angles
        if tt[-1] <= time + 0.00001:
            return aa[-1]
        for i in range(len(tt) - 1):
            if tt[i] <= time + 0.00001 and tt[i + 1] > time + 0.00001:
                k = (aa[i + 1] - aa[i]) / (tt[i + 1] - tt[i])
                b = aa[i + 1] - (aa[i + 1] - aa[i]) / (tt[i + 1] - tt[i]) * tt[i + 1]
                return k * time + b
        raise ValueError("frame before the first trajectory state (undefined in the reference, RobotObstacleFCL.cpp:429)")

    def robot_obstacle_links(self, r: int, f: int) -> List[np.ndarray]:
        """world transforms of ALL shapes of robot obstacle r at frame f"""
        key = (r, 0 if self.p.robot_obstacles[r].is_static else f)
        if key not in self._rob_cache:
            if len(self._rob_cache) > 4096:
                self._rob_cache.clear()
            self._rob_cache[key] = fk_chain(self.p.robot_obstacles[r].model, self.robot_obstacle_angles(r, key[1]))
        return self._rob_cache[key]

    # a3 / a4 core: every pair test of one cell
    def cell_pairs(self, q, f: int, semantics: int = SEM_INTERVAL):
        """yields (link, obstacle id, kind, hit, signed distance) for every pair test the reference's callback would make
        whose bounding spheres are within reach (the rest is far apart and reported as not hit, distance None)"""
        links = self.links(q)
        lb = [l.bound() for l in links]
        out = []
        for c in range(self.C):
            ob = self.obstacle(c, f)
            oc, orad = ob.bound()
            for li, l in enumerate(links):
                if np.linalg.norm(lb[li][0] - oc) > lb[li][1] + orad + 1e-3:
                    continue
                hit, d = collides(l, ob)
                out.append((li, c, "true", hit, d))
        for k, (r, s) in enumerate(self.rob_links):
            Ts = self.robot_obstacle_links(r, f)
            true_ob = Posed(self.p.robot_obstacles[r].model.shapes[s], Ts[s])
            want_true = self.parity_mode == FIXED or semantics == SEM_INTERVAL
            want_shift = self.parity_mode == REFERENCE_EXACT
            shift_geom = self.rob_all_shapes[k] if k < len(self.rob_all_shapes) else None
            tc, trad = true_ob.bound()
            for li, l in enumerate(links):
                if want_true and np.linalg.norm(lb[li][0] - tc) <= lb[li][1] + trad + 1e-3:
                    hit, d = collides(l, true_ob)
                    out.append((li, self.C + k, "true", hit, d))
                if want_shift and shift_geom is not None:
                    # broadphase gate: coal AABBs of the link and of the TRUE obstacle geometry must overlap
                    alo, ahi = l.coal_aabb()
                    blo, bhi = true_ob.coal_aabb()
                    if np.any(alo > bhi) or np.any(ahi < blo):
                        continue
                    sh = Posed(shift_geom, Ts[s])
                    sc, srad = sh.bound()
                    if np.linalg.norm(lb[li][0] - sc) > lb[li][1] + srad + 1e-3:
                        continue
                    hit, d = collides(l, sh)
                    out.append((li, self.C + k, "shift", hit, d))
        return out

    def check_cell(self, q, f: int, semantics: int = SEM_FRAME) -> Tuple[bool, float]:
        """(collision boolean of the cell, smallest |signed distance| over its pair tests = distance to the decision
        boundary; inf when nothing is within reach)"""
        if f < 0 or f >= self.n_frames:
            return True, math.inf
        hit, margin = False, math.inf
        for _, _, _, h, d in self.cell_pairs(q, f, semantics):
            hit |= h
            margin = min(margin, abs(d))
        return hit, margin

    def collision_row(self, q) -> np.ndarray:
        return np.array([self.check_cell(q, f, SEM_INTERVAL)[0] for f in range(self.n_frames)], bool)

    @staticmethod
    def intervals_from_row(row: np.ndarray) -> List[Tuple[int, int]]:
        """complement runs of the colliding frames, inclusive (CollisionManager.cpp:507-546)"""
        safe = ~np.asarray(row, bool)
        pad = np.concatenate([[False], safe, [False]]).astype(np.int8)
        d = np.diff(pad)
        return [(int(a), int(b) - 1) for a, b in zip(np.nonzero(d == 1)[0], np.nonzero(d == -1)[0])]
