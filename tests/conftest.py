import gzip
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def sp():
    return g.load_tools()


@pytest.fixture(scope="session")
def orc():
    m = g.load_oracle()
    m.build()
    return m


@pytest.fixture(scope="session")
def pkg():
    """The product library through its ctypes binding.  GPU tests only: never falls back."""
    g.load_build().build_lib()
    return g.load_package()


def golden_pack(name: str) -> bytes:
    with gzip.open(os.path.join(GOLDEN, name + ".pack.gz"), "rb") as f:
        return f.read()


@pytest.fixture(scope="session")
def golden_vectors():
    with open(os.path.join(GOLDEN, "golden_vectors.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def logged_cells(sp):
    """(q [n][dof] float64, frames [n] int32) cells of trajectories logged by an OLDER pipeline of the reference project
    (its checker was pybullet, not the shipped CollisionManager -- tests/golden/make_logged_paths.py): a centimetre-level
    sanity check, not a pin.  ST-RRT* solutions are expanded
    edge by edge exactly as SpaceTimeMotionValidator::checkMotion interpolates them
    (SpaceTimeMotionValidator.cpp:48-55: steps = max(int(|dq|/0.1), int(2*dt*fps+0.5)) + 2, states i/steps, plus
    the end state checked first at :25); DRGBT logs are executed states.  Frame = (int)(float32 time * fps)
    (CollisionManager.cpp:333 with `float t` from StateValidityCheckerFunctor.cpp:34)."""
    with gzip.open(os.path.join(GOLDEN, "reference_logged_paths.json.gz"), "rb") as f:
        logs = json.load(f)

    def get(key: str):
        ent = logs[key]
        sc = sp.read_pack(golden_pack(ent["scene"]))
        t, q = np.asarray(ent["t"], np.float64), np.asarray(ent["q"], np.float64)
        if ent["planner"].startswith("STRRT"):
            ts, qs = [], []
            for a in range(len(t) - 1):
                dq, dt = float(np.sqrt(((q[a + 1] - q[a]) ** 2).sum())), float(t[a + 1] - t[a])
                steps = max(int(dq / 0.1), int(2 * dt * sc.fps + 0.5)) + 2
                for i in list(range(steps)) + [steps]:
                    ts.append(t[a] + dt * (i / steps))
                    qs.append(q[a] + (q[a + 1] - q[a]) * (i / steps))
            t, q = np.asarray(ts), np.asarray(qs)
        frames = (t.astype(np.float32) * np.float32(sc.fps)).astype(np.int32)
        keep = frames < sc.n_frames
        return ent["scene"], np.ascontiguousarray(q[keep]), np.ascontiguousarray(frames[keep])

    get.keys = sorted(logs)
    return get


@pytest.fixture(scope="session")
def packs(sp):
    return make_packs(sp)


def make_packs(sp):
    """Lazily built scene packs shared by the tests."""
    cache = {}

    def get(name: str) -> bytes:
        if name in cache:
            return cache[name]
        if name in ("mass_test", "flying_boxes", "multiple_robots", "scene_task"):
            pk = golden_pack(name)
        elif name == "mass_test_cyl":  # MT obstacles, capsule xArm6 (the multi-agent generators' model)
            sc = sp.read_pack(golden_pack("mass_test"))
            sc.robot = sp.read_pack(golden_pack("xarm6cyl_robot")).robot
            pk = sp.write_pack(sc)
        elif name == "multiple_robots_cyl":  # capsule robots everywhere: robot-vs-robot capsule pairs
            sc = sp.read_pack(golden_pack("multiple_robots"))
            cyl = sp.read_pack(golden_pack("xarm6cyl_robot")).robot
            cyl.base_pos, cyl.base_quat = sc.robot.base_pos, sc.robot.base_quat
            sc.robot = cyl
            for ro in sc.robot_obstacles:
                r2 = sp.read_pack(golden_pack("xarm6cyl_robot")).robot
                r2.base_pos, r2.base_quat = ro.robot.base_pos, ro.robot.base_quat
                ro.robot = r2
            pk = sp.write_pack(sc)
        elif name.startswith("syn"):  # syn<A|B|C>_<n_obs>_<n_frames>
            var, n_obs, n_frames = name[3], *[int(x) for x in name[5:].split("_")]
            pk = sp.write_pack(sp.synthetic_scene(n_obs, n_frames, 30, 42, var))
        else:
            raise KeyError(name)
        cache[name] = pk
        return pk

    return get


def random_q(sp, pack: bytes, n: int, seed: int) -> np.ndarray:
    return sp.random_configs(sp.read_pack(pack).robot, n, seed)
