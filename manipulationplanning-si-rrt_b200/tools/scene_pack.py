"""Scene packer: scene JSON + URDF + binary STL (or a synthetic generator) -> flat "scene pack".

The pack (layout: include/sirrt_pack.h) is the only input of libsirrt_b200 and of the test-only
oracle.  This module does the *parsing* the reference does with rapidjson / urdfdom / assimp /
qhull and nothing that decides parity numerically beyond what those parsers do:

* scene JSON           -- config_read.cpp:62-314 (all six obstacle types; JSON doubles narrowed
                          to float32 for obstacle poses/dimensions and the robot base, QUIRK 5/12)
* URDF -> fk_chain     -- RobotObstacleFCL.cpp:54-195 (first child only; fixed-joint origins
                          ignored; collision <origin> honoured only for cylinders; cylinder ->
                          capsule; mesh -> Qhull "Qt" convex hull; QUIRKS 1-4 of SURVEY.md §3.5)
* rpy -> quaternion    -- urdfdom Rotation::setFromRPY (+normalize), then
                          Eigen::Quaternion::toRotationMatrix (restated in quat_to_matrix)
* legacy trajectories  -- STRRT/scene_task_multiple_robots.json stores bare angle vectors; they
                          are converted to {time: k/fps} states (SURVEY.md §8d input caveats)

Synthetic generators follow Blender/scripts/generate_mass_test.py:30-60 (bouncing cubes).
"""
from __future__ import annotations

import json
import math
import os
import struct
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np

MAGIC = b"SIRRTPK1"
VERSION = 1
OBS_BOX, OBS_SPHERE = 0, 1
SHAPE_CAPSULE, SHAPE_HULL = 0, 1
N_STRIP = 2  # RobotObstacleFCL.cpp:370

ASSETS = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "assets")


# --------------------------------------------------------------------------- math helpers
def rpy_to_quat(roll: float, pitch: float, yaw: float):
    """urdfdom Rotation::setFromRPY followed by normalize(); returns (x, y, z, w)."""
    phi, the, psi = roll / 2.0, pitch / 2.0, yaw / 2.0
    x = math.sin(phi) * math.cos(the) * math.cos(psi) - math.cos(phi) * math.sin(the) * math.sin(psi)
    y = math.cos(phi) * math.sin(the) * math.cos(psi) + math.sin(phi) * math.cos(the) * math.sin(psi)
    z = math.cos(phi) * math.cos(the) * math.sin(psi) - math.sin(phi) * math.sin(the) * math.cos(psi)
    w = math.cos(phi) * math.cos(the) * math.cos(psi) + math.sin(phi) * math.sin(the) * math.sin(psi)
    s = math.sqrt(x * x + y * y + z * z + w * w)
    if s == 0.0:
        return 0.0, 0.0, 0.0, 1.0
    return x / s, y / s, z / s, w / s


def quat_to_matrix(x: float, y: float, z: float, w: float) -> np.ndarray:
    """Eigen::QuaternionBase::toRotationMatrix (no normalisation), row-major 3x3 float64."""
    tx, ty, tz = 2.0 * x, 2.0 * y, 2.0 * z
    twx, twy, twz = tx * w, ty * w, tz * w
    txx, txy, txz = tx * x, ty * x, tz * x
    tyy, tyz, tzz = ty * y, tz * y, tz * z
    return np.array(
        [
            [1.0 - (tyy + tzz), txy - twz, txz + twy],
            [txy + twz, 1.0 - (txx + tzz), tyz - twx],
            [txz - twy, tyz + twx, 1.0 - (txx + tyy)],
        ],
        dtype=np.float64,
    )


def _floats(s: Optional[str], n: int = 3):
    if s is None:
        return [0.0] * n
    v = [float(t) for t in s.split()]
    assert len(v) == n, s
    return v


# --------------------------------------------------------------------------- data model
@dataclass
class Shape:
    type: int
    joints_before: int
    lo_t: np.ndarray = field(default_factory=lambda: np.zeros(3))
    lo_R: np.ndarray = field(default_factory=lambda: np.eye(3))
    radius: float = 0.0
    length: float = 0.0
    verts: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float32))
    name: str = ""


@dataclass
class Robot:
    dof: int
    base_pos: np.ndarray  # float32[3]
    base_quat: np.ndarray  # float32[4] x,y,z,w
    limits: np.ndarray  # float32[dof,2]
    joints_t: np.ndarray  # float64[nj,3]
    joints_R: np.ndarray  # float64[nj,3,3]
    shapes: List[Shape]
    joint_names: List[str] = field(default_factory=list)


@dataclass
class Obstacle:
    type: int
    is_static: bool
    dims: np.ndarray  # float32[4]
    poses: np.ndarray  # float32[n,7]
    name: str = ""


@dataclass
class RobotObstacle:
    robot: Robot
    is_static: bool
    times: np.ndarray  # float64[n]
    angles: np.ndarray  # float64[n,dof]
    name: str = ""


@dataclass
class Scene:
    n_frames: int
    fps: int
    start_time: float
    start_configuration: np.ndarray
    end_configuration: np.ndarray
    joint_max_velocity: np.ndarray
    robot: Robot
    obstacles: List[Obstacle] = field(default_factory=list)
    robot_obstacles: List[RobotObstacle] = field(default_factory=list)


# --------------------------------------------------------------------------- STL / hull
def read_binary_stl_vertices(path: str) -> np.ndarray:
    """All triangle corner coordinates of a binary STL as float32 [3*ntri, 3]."""
    with open(path, "rb") as f:
        data = f.read()
    ntri = struct.unpack_from("<I", data, 80)[0]
    assert 84 + 50 * ntri == len(data), f"{path}: not a binary STL"
    rec = np.frombuffer(data, dtype=np.dtype([("n", "<f4", 3), ("v", "<f4", (3, 3)), ("a", "<u2")]), count=ntri, offset=84)
    return np.ascontiguousarray(rec["v"].reshape(-1, 3))


def convex_hull_vertices(points_f32: np.ndarray) -> np.ndarray:
    """Hull vertex set with Qhull option "Qt" (RobotObstacleFCL.cpp:117) -> float32 [V,3]."""
    from scipy.spatial import ConvexHull

    uniq = np.unique(points_f32, axis=0)
    hull = ConvexHull(uniq.astype(np.float64), qhull_options="Qt")
    return np.ascontiguousarray(uniq[np.sort(hull.vertices)], dtype=np.float32)


def box_hull(dx: float, dy: float, dz: float) -> np.ndarray:
    h = np.array([dx, dy, dz], np.float64) / 2.0
    s = np.array([[sx, sy, sz] for sx in (-1, 1) for sy in (-1, 1) for sz in (-1, 1)], np.float64)
    return (s * h).astype(np.float32)


# --------------------------------------------------------------------------- URDF
def parse_urdf(urdf_path: str, joints_order: Sequence[str], base_pos, base_quat) -> Robot:
    """Restates RobotObstacleFCL::load_joint_collision_models (RobotObstacleFCL.cpp:54-195)."""
    root = ET.parse(urdf_path).getroot()
    urdf_dir = os.path.dirname(os.path.abspath(urdf_path))
    links = {l.get("name"): l for l in root.findall("link")}
    children = {}  # parent link -> list of (joint name, joint element)
    has_parent = set()
    for j in root.findall("joint"):
        p = j.find("parent").get("link")
        c = j.find("child").get("link")
        children.setdefault(p, []).append((j.get("name"), j))
        has_parent.add(c)
    roots = [n for n in links if n not in has_parent]
    if len(roots) != 1:
        raise RuntimeError("Failed to parse urdf file")  # RobotObstacleFCL.cpp:50
    for v in children.values():
        v.sort(key=lambda t: t[0])  # urdfdom keeps joints in a std::map keyed by name

    dof = len(joints_order)
    limits = np.zeros((dof, 2), np.float32)
    joints_t, joints_R, joint_names, shapes = [], [], [], []
    cur = roots[0]
    while cur is not None:
        link = links[cur]
        col = link.find("collision")
        geom = col.find("geometry") if col is not None else None
        shape = None
        if geom is not None and len(geom):
            g = geom[0]
            if g.tag == "mesh":
                v = read_binary_stl_vertices(os.path.join(urdf_dir, g.get("filename")))
                shape = Shape(SHAPE_HULL, len(joints_t), verts=convex_hull_vertices(v), name=cur)
            elif g.tag == "box":
                d = _floats(g.get("size"))
                shape = Shape(SHAPE_HULL, len(joints_t), verts=box_hull(*d), name=cur)
            elif g.tag == "cylinder":
                o = col.find("origin")
                xyz = _floats(o.get("xyz") if o is not None else None)
                rpy = _floats(o.get("rpy") if o is not None else None)
                shape = Shape(
                    SHAPE_CAPSULE,
                    len(joints_t),
                    lo_t=np.array(xyz, np.float64),
                    lo_R=quat_to_matrix(*rpy_to_quat(*rpy)),
                    radius=float(g.get("radius")),
                    length=float(g.get("length")),
                    name=cur,
                )
        if shape is not None:
            shapes.append(shape)
        ch = children.get(cur, [])
        if not ch:
            cur = None
            continue
        jname, j = ch[0]  # child_joints.at(0)
        lim = j.find("limit")
        if jname in joints_order and lim is not None:
            i = list(joints_order).index(jname)
            limits[i, 0] = np.float32(float(lim.get("lower", "0")))
            limits[i, 1] = np.float32(float(lim.get("upper", "0")))
        if j.get("type") != "fixed":
            o = j.find("origin")
            xyz = _floats(o.get("xyz") if o is not None else None)
            rpy = _floats(o.get("rpy") if o is not None else None)
            joints_t.append(xyz)
            joints_R.append(quat_to_matrix(*rpy_to_quat(*rpy)))
            joint_names.append(jname)
        cur = j.find("child").get("link")

    return Robot(
        dof=dof,
        base_pos=np.asarray(base_pos, np.float32),
        base_quat=np.asarray(base_quat, np.float32),
        limits=limits,
        joints_t=np.asarray(joints_t, np.float64).reshape(-1, 3),
        joints_R=np.asarray(joints_R, np.float64).reshape(-1, 3, 3),
        shapes=shapes,
        joint_names=joint_names,
    )


# --------------------------------------------------------------------------- scene JSON
def _resolve(path: str, search_dirs: Sequence[str]) -> str:
    if os.path.isabs(path) and os.path.exists(path):
        return path
    for d in search_dirs:
        p = os.path.normpath(os.path.join(d, path))
        if os.path.exists(p):
            return p
    raise FileNotFoundError(path)


def load_scene_json(scene_path: str, repo_root: Optional[str] = None) -> Scene:
    """config_read.cpp:62-314.  URDF paths are relative to the CWD the planner is run from
    (the reference repo root); pass it as repo_root."""
    with open(scene_path) as f:
        d = json.load(f)
    for k in ("start_configuration", "end_configuration", "frame_count", "fps", "obstacles", "robot_urdf",
              "robot_base_coords", "robot_base_quat_rot", "robot_joints_order"):
        if k not in d:
            raise KeyError(f"scene json lacks required key {k!r}")  # check_json_key
    search = [os.getcwd(), os.path.dirname(os.path.abspath(scene_path))]
    if repo_root:
        search.insert(0, repo_root)
    T, fps = int(d["frame_count"]), int(d["fps"])
    order = list(d["robot_joints_order"])
    robot = parse_urdf(_resolve(d["robot_urdf"], search), order, d["robot_base_coords"], d["robot_base_quat_rot"])
    scene = Scene(
        n_frames=T,
        fps=fps,
        start_time=float(d.get("start_time", 0.0)),
        start_configuration=np.asarray(d["start_configuration"], np.float64),
        end_configuration=np.asarray(d["end_configuration"], np.float64),
        joint_max_velocity=np.asarray(d.get("robot_joint_max_velocity", [math.pi] * len(order)), np.float64),
        robot=robot,
    )
    for o in d["obstacles"]:
        t = o["type"]
        if t in ("static_box", "dynamic_box", "static_sphere", "dynamic_sphere"):
            static = t.startswith("static")
            poses = np.asarray(o["positions"][:1] if static else o["positions"], np.float64).astype(np.float32)
            assert poses.shape[1] == 7
            if not static and poses.shape[0] < T:
                raise ValueError(f"obstacle {o.get('name')}: {poses.shape[0]} poses < frame_count {T}")
            if t.endswith("box"):
                dims = np.array(list(o["dimensions"]) + [0.0], np.float64).astype(np.float32)
                scene.obstacles.append(Obstacle(OBS_BOX, static, dims, poses, o.get("name", "")))
            else:
                dims = np.array([o["radius"], 0, 0, 0], np.float64).astype(np.float32)
                scene.obstacles.append(Obstacle(OBS_SPHERE, static, dims, poses, o.get("name", "")))
        elif t in ("dynamic_robot", "static_robot"):
            r = parse_urdf(_resolve(o["urdf_file_path"], search), list(o["robot_joints_order"]),
                           o["robot_base_coords"], o["robot_base_quat_rot"])
            traj = o["trajectory"]
            if traj and not isinstance(traj[0], dict):  # legacy: bare angle vectors, one per frame
                times = np.arange(len(traj), dtype=np.float64) / float(fps)
                angles = np.asarray(traj, np.float64)
            else:
                times = np.asarray([s["time"] for s in traj], np.float64)
                angles = np.asarray([s["robot_angles"] for s in traj], np.float64)
            scene.robot_obstacles.append(RobotObstacle(r, t == "static_robot", times, angles.reshape(len(times), -1), o.get("name", "")))
        # unknown types are silently skipped (config_read.cpp:78-250)
    return scene


# --------------------------------------------------------------------------- pack writer / reader
def _pad8(b: bytearray):
    while len(b) % 8:
        b.append(0)


def _robot_bytes(r: Robot) -> bytearray:
    b = bytearray()
    b += struct.pack("<4i", r.dof, len(r.joints_t), len(r.shapes), N_STRIP)
    b += np.asarray(r.base_pos, "<f4").tobytes() + np.asarray(r.base_quat, "<f4").tobytes()
    b += struct.pack("<i", 0)
    b += np.asarray(r.limits, "<f4").reshape(-1).tobytes()
    _pad8(b)
    for t, R in zip(r.joints_t, r.joints_R):
        b += np.asarray(t, "<f8").tobytes() + np.asarray(R, "<f8").reshape(-1).tobytes()
    for s in r.shapes:
        b += struct.pack("<4i", s.type, s.joints_before, len(s.verts), 0)
        b += np.asarray(s.lo_t, "<f8").tobytes() + np.asarray(s.lo_R, "<f8").reshape(-1).tobytes()
        b += struct.pack("<2d", s.radius, s.length)
        b += np.asarray(s.verts, "<f4").reshape(-1).tobytes()
        _pad8(b)
    return b


def write_pack(scene: Scene) -> bytes:
    dof = scene.robot.dof
    body = bytearray()
    for v in (scene.start_configuration, scene.end_configuration, scene.joint_max_velocity):
        v = np.asarray(v, "<f8")
        assert v.shape == (dof,), (v.shape, dof)
        body += v.tobytes()
    body += _robot_bytes(scene.robot)
    for o in scene.obstacles:
        poses = np.asarray(o.poses, "<f4").reshape(-1, 7)
        n = 1 if o.is_static else poses.shape[0]
        body += struct.pack("<4i", o.type, int(o.is_static), n, 0)
        body += np.asarray(o.dims, "<f4").tobytes()
        body += poses[:n].tobytes()
        _pad8(body)
    for ro in scene.robot_obstacles:
        body += struct.pack("<2i", int(ro.is_static), len(ro.times))
        body += _robot_bytes(ro.robot)
        body += np.asarray(ro.times, "<f8").tobytes()
        body += np.asarray(ro.angles, "<f8").reshape(-1).tobytes()
    total = 56 + len(body)
    hdr = struct.pack("<8sIIiidiiiiQ", MAGIC, VERSION, 0, scene.n_frames, scene.fps, scene.start_time, dof,
                      len(scene.obstacles), len(scene.robot_obstacles), 0, total)
    assert len(hdr) == 56
    return bytes(hdr) + bytes(body)


class _Cur:
    def __init__(self, b: bytes):
        self.b, self.o = b, 0

    def take(self, fmt):
        v = struct.unpack_from(fmt, self.b, self.o)
        self.o += struct.calcsize(fmt)
        return v

    def arr(self, dt, n):
        a = np.frombuffer(self.b, dtype=dt, count=n, offset=self.o).copy()
        self.o += a.nbytes
        return a

    def pad8(self):
        self.o = (self.o + 7) & ~7


def _read_robot(c: _Cur) -> Robot:
    dof, nj, ns, nstrip = c.take("<4i")
    base_pos, base_quat = c.arr("<f4", 3), c.arr("<f4", 4)
    c.take("<i")
    limits = c.arr("<f4", 2 * dof).reshape(dof, 2)
    c.pad8()
    jt, jR = [], []
    for _ in range(nj):
        jt.append(c.arr("<f8", 3))
        jR.append(c.arr("<f8", 9).reshape(3, 3))
    shapes = []
    for _ in range(ns):
        ty, jb, nv, _p = c.take("<4i")
        lo_t, lo_R = c.arr("<f8", 3), c.arr("<f8", 9).reshape(3, 3)
        radius, length = c.take("<2d")
        verts = c.arr("<f4", 3 * nv).reshape(nv, 3)
        c.pad8()
        shapes.append(Shape(ty, jb, lo_t, lo_R, radius, length, verts))
    return Robot(dof, base_pos, base_quat, limits, np.array(jt).reshape(-1, 3), np.array(jR).reshape(-1, 3, 3), shapes)


def read_pack(b: bytes) -> Scene:
    magic, ver, _fl, T, fps, st, dof, C, R, _r, total = struct.unpack_from("<8sIIiidiiiiQ", b, 0)
    if magic != MAGIC or ver != VERSION or total != len(b):
        raise ValueError("bad scene pack")
    c = _Cur(b)
    c.o = 56
    sc, ec, mv = c.arr("<f8", dof), c.arr("<f8", dof), c.arr("<f8", dof)
    scene = Scene(T, fps, st, sc, ec, mv, _read_robot(c))
    for _ in range(C):
        ty, st_, n, _p = c.take("<4i")
        dims = c.arr("<f4", 4)
        poses = c.arr("<f4", 7 * n).reshape(n, 7)
        c.pad8()
        scene.obstacles.append(Obstacle(ty, bool(st_), dims, poses))
    for _ in range(R):
        st_, n = c.take("<2i")
        r = _read_robot(c)
        times = c.arr("<f8", n)
        angles = c.arr("<f8", n * r.dof).reshape(n, r.dof)
        scene.robot_obstacles.append(RobotObstacle(r, bool(st_), times, angles))
    assert c.o == len(b), (c.o, len(b))
    return scene


def pack_scene_json(scene_path: str, repo_root: Optional[str] = None) -> bytes:
    return write_pack(load_scene_json(scene_path, repo_root))


# --------------------------------------------------------------------------- synthetic scenes
def arm7_robot() -> Robot:
    order = [f"joint{i}" for i in range(1, 8)]
    return parse_urdf(os.path.join(ASSETS, "arm7cyl.urdf"), order, [0, 0, 0], [0, 0, 0, 1])


def bouncing_boxes(n_obs: int, n_frames: int, fps: int, rng: np.random.Generator, variant: str = "A"):
    """SYN obstacle set: cube edge ~U(0.05,0.1), start ~U(-1,1)^3, velocity
    ~U(-0.2,0.2)^3 m/s, reflect at |x|>1 (generate_mass_test.py:30-60).  Variant A: identity
    rotation; B: constant random angular velocity; C: spheres r~U(0.025,0.05)."""
    obstacles = []
    dt = 1.0 / fps
    for _ in range(n_obs):
        edge = rng.uniform(0.05, 0.1)
        p = rng.uniform(-1.0, 1.0, 3)
        v = rng.uniform(-0.2, 0.2, 3)
        axis = rng.normal(size=3)
        axis /= np.linalg.norm(axis)
        omega = rng.uniform(0.2, 2.0)
        poses = np.zeros((n_frames, 7), np.float64)
        for f in range(n_frames):
            poses[f, :3] = p
            if variant == "B":
                a = omega * f * dt
                poses[f, 3:6] = axis * math.sin(a / 2)
                poses[f, 6] = math.cos(a / 2)
            else:
                poses[f, 6] = 1.0
            p = p + v * dt
            for k in range(3):
                if abs(p[k]) > 1.0:
                    v[k] = -v[k]
                    p[k] = math.copysign(2.0 - abs(p[k]), p[k])
        if variant == "C":
            dims = np.array([rng.uniform(0.025, 0.05), 0, 0, 0], np.float32)
            obstacles.append(Obstacle(OBS_SPHERE, False, dims, poses.astype(np.float32)))
        else:
            dims = np.array([edge, edge, edge, 0], np.float32)
            obstacles.append(Obstacle(OBS_BOX, False, dims, poses.astype(np.float32)))
    return obstacles


def synthetic_scene(n_obs: int = 64, n_frames: int = 1000, fps: int = 30, seed: int = 42, variant: str = "A",
                    robot: Optional[Robot] = None) -> Scene:
    rng = np.random.default_rng(seed)
    robot = robot or arm7_robot()
    z = np.zeros(robot.dof)
    return Scene(n_frames, fps, 0.0, z, z.copy(), np.full(robot.dof, math.pi), robot,
                 bouncing_boxes(n_obs, n_frames, fps, rng, variant))


def random_configs(robot: Robot, n: int, seed: int = 42) -> np.ndarray:
    """q ~ U(lo, hi) per joint as float32 SoA [dof][n]."""
    rng = np.random.default_rng(seed + 1)
    lo = robot.limits[:, 0].astype(np.float64)[:, None]
    hi = robot.limits[:, 1].astype(np.float64)[:, None]
    return np.ascontiguousarray((lo + rng.random((robot.dof, n)) * (hi - lo)).astype(np.float32))


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser(description="scene JSON -> scene pack")
    ap.add_argument("scene_json")
    ap.add_argument("out_pack")
    ap.add_argument("--root", default=None, help="directory URDF paths are relative to")
    a = ap.parse_args()
    blob = pack_scene_json(a.scene_json, a.root)
    with open(a.out_pack, "wb") as f:
        f.write(blob)
    print(f"{a.out_pack}: {len(blob)} bytes")
