"""SI-RRT planner and result log (f3).

CPU part: host/SIRRTPlanner.hpp instantiated on the CPU oracle (tests/planner_oracle_main.cpp, test
infrastructure) -- plans exist, pass the reference's own check_path (PlannerConnect.cpp:934-954), are
collision-free when re-validated state by state, are deterministic per seed, and the log carries the reference's
JSON keys (ResultsWriter.cpp:52-124, MSIRRT/src/main.cpp:70-116).
GPU part: the product executable (host/sirrt_plan, libsirrt_b200 backend with and without the §7.3 edge table)
produces IDENTICAL plans to the oracle-backed planner for the same seeds."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, golden_pack

BUILD = os.path.join(ROOT, "tests", "_build")
PKG = os.path.join(ROOT, "manipulationplanning-si-rrt_b200")

REFERENCE_LOG_KEYS = [  # ResultsWriter.cpp:69-99
    "planner_type", "path_to_scene_json", "path_to_config_json", "full_execution_time_ns", "planner_config_time_ns",
    "algorithm_solving_time_ns", "collision_check_time_ns", "distance_check_time_ns", "forward_kinematics_in_collision_check_time_ns",
    "forward_kinematics_in_distance_check_time_ns", "total_forward_kinematics_time_ns", "get_safe_interval_time_ns",
    "collision_check_init_time_ns", "safe_intervals_init_time_ns", "number_of_collision_checks", "number_of_get_safe_interval",
    "number_of_broadphase_collision_checks", "number_of_narrowphase_collision_checks", "number_of_broadphase_get_safe_interval",
    "number_of_narrowphase_get_safe_interval", "number_of_distances_checks", "number_of_forward_kinematics", "number_of_detected_collisions",
    "number_of_safe_intervals", "number_of_sum_of_safe_interval_frames", "final_planner_data", "path_results"]


@pytest.fixture(scope="session")
def planner_oracle(orc):
    os.makedirs(BUILD, exist_ok=True)
    exe = os.path.join(BUILD, "planner_oracle")
    src = os.path.join(ROOT, "tests", "planner_oracle_main.cpp")
    deps = [src, os.path.join(PKG, "host", "SIRRTPlanner.hpp"), os.path.join(PKG, "host", "ResultsLog.hpp")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        odir = os.path.join(ROOT, "oracle")
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fopenmp", "-o", exe, src, "-L", odir, "-lsirrt_oracle", f"-Wl,-rpath,{odir}"])
    return exe


@pytest.fixture(scope="session")
def scenes(sp, orc, tmp_path_factory):
    """name -> pack path.  Reference scenes + two synthetic dynamic-obstacle tasks (capsule xArm6 among bouncing
    boxes, the config-5 recipe: start free at frame 0, goal free at the last frame, |dq|_1 > pi)."""
    d = tmp_path_factory.mktemp("packs")
    out = {}
    for name in ("scene_task", "mass_test"):
        p = d / f"{name}.pack"
        p.write_bytes(golden_pack(name))
        out[name] = str(p)
    cyl = sp.read_pack(golden_pack("xarm6cyl_robot")).robot
    for k, n_obs in enumerate([40, 90]):
        for seed in range(5, 60):  # first seed whose scene passes the check_scene rule for some start / goal pair
            sc = sp.synthetic_scene(n_obs, 600, 30, seed, "A", robot=cyl)
            o = orc.Oracle(sp.write_pack(sc))
            q = sp.random_configs(cyl, 200, seed).T.astype(np.float64)
            free0 = [i for i in range(len(q)) if not o.check_frame(q[i], 0)]
            freeT = [i for i in range(len(q)) if not o.check_frame(q[i], 599)]
            pair = next(((a, b) for a in free0 for b in freeT if np.abs(q[a] - q[b]).sum() > np.pi), None)
            if pair:
                break
        assert pair, "no synthetic scene with a valid start / goal pair"
        sc.start_configuration, sc.end_configuration = q[pair[0]].copy(), q[pair[1]].copy()
        p = d / f"dyn{k}.pack"
        p.write_bytes(sp.write_pack(sc))
        out[f"dyn{k}"] = str(p)
    return out


def run_json(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    return json.loads(r.stdout.strip().splitlines()[-1])


def revalidate(orc, pack_path, path, fps):
    """Every motion of the path, interpolated like is_collision_motion (PlannerConnect.cpp:879-886), is free."""
    with open(pack_path, "rb") as f:
        o = orc.Oracle(f.read())
    for a, b in zip(path[:-1], path[1:]):
        q0, q1 = np.array(a["robot_angles"]), np.array(b["robot_angles"])
        t0, t1 = a["time"] * fps, b["time"] * fps
        assert t1 >= t0 - 1e-9
        steps = max(int(np.linalg.norm(q1 - q0) / 0.1), int(t1 - t0 + 0.5), 1)
        for s in range(steps + 1):
            q = q0 + (q1 - q0) * s / steps
            fr = int(t0 + (t1 - t0) * s / steps + 1e-9)
            assert not o.check_frame(q, fr), f"path state collides at frame {fr}"


@pytest.mark.parametrize("name", ["scene_task", "mass_test", "dyn0", "dyn1"])
def test_oracle_backed_planner_finds_valid_plans(planner_oracle, scenes, orc, name):
    for seed in (1, 2, 3):
        d = run_json([planner_oracle, scenes[name], str(seed), "--max-time", "20"])
        fp = d["final_planner_data"]
        assert fp["has_result"], f"{name} seed {seed}: no plan"
        assert d["b200"]["check_path"] == 1.0
        path = fp["final_path"]
        with open(scenes[name], "rb") as f:
            o = orc.Oracle(f.read())
        assert np.allclose(path[0]["time"], 0.0) and len(path) >= 2
        assert fp["path_cost"] == path[-1]["time"]
        assert all(b["time"] >= a["time"] for a, b in zip(path[:-1], path[1:]))
        revalidate(orc, scenes[name], path, o.fps)


@pytest.mark.parametrize("name", ["mass_test", "dyn0", "dyn1"])
def test_radius_prefetch_hook_announces_every_query_exactly(planner_oracle, scenes, name):
    """The planner template tells a backend that offers knn_radius_prefetch which radius query set_parent will ask
    (PlannerConnect.cpp:414-416) BEFORE its safe-interval call -- B200Backend launches it there, so that the two GPU calls run
    side by side.  Checked on the CPU with a recording backend (planner_oracle --check-prefetch): every hint is followed by
    exactly that query (same tree, same coordinates bit for bit, same radius, no vertex added in between), and the plan does
    not depend on the hook."""
    for seed in (1, 2):
        plain = run_json([planner_oracle, scenes[name], str(seed), "--max-time", "20", "--baseline"])
        hooked = run_json([planner_oracle, scenes[name], str(seed), "--max-time", "20", "--baseline", "--check-prefetch"])
        assert plain["final_planner_data"] == hooked["final_planner_data"]
        b = hooked["b200"]
        assert b["radius_hints"] >= 1 and b["radius_hint_mismatches"] == 0 and b["radius_hints_overwritten"] == 0, b
        assert b["radius_queries_announced_exactly"] == b["radius_hints"] <= b["radius_queries"], b


def test_planner_is_deterministic_per_seed_and_log_schema(planner_oracle, scenes):
    a = run_json([planner_oracle, scenes["dyn0"], "7", "--max-time", "20"])
    b = run_json([planner_oracle, scenes["dyn0"], "7", "--max-time", "20"])
    c = run_json([planner_oracle, scenes["dyn0"], "8", "--max-time", "20"])
    assert a["final_planner_data"] == b["final_planner_data"]
    assert a["final_planner_data"] != c["final_planner_data"]
    for k in REFERENCE_LOG_KEYS:
        assert k in a, f"log key {k} of the reference schema is missing"
    assert a["planner_type"] == "MSIRRT_jump"
    for st in a["final_planner_data"]["final_path"]:
        assert set(st) == {"time", "robot_angles"}


EXACT_COUNTERS = ["number_of_collision_checks", "number_of_get_safe_interval", "number_of_detected_collisions", "number_of_safe_intervals",
                  "number_of_sum_of_safe_interval_frames", "number_of_forward_kinematics", "number_of_broadphase_get_safe_interval",
                  "number_of_broadphase_collision_checks"]


def test_log_counters_follow_the_reference_definitions(planner_oracle, scenes, orc):
    """ResultsWriter counters the planner can reproduce (CollisionManager.cpp:350-377, :492-545): one FK per state
    check and per interval query, `links` broad-phase queries per interval query, one detected collision per colliding
    motion check, and interval counts / frame sums consistent with the scene's frame count."""
    with open(scenes["dyn1"], "rb") as f:
        o = orc.Oracle(f.read())
    d = run_json([planner_oracle, scenes["dyn1"], "3", "--max-time", "20"])
    n_si, n_cc = d["number_of_get_safe_interval"], d["number_of_collision_checks"]
    assert n_si > 0 and n_cc > 0
    assert d["number_of_forward_kinematics"] == n_si + n_cc
    assert d["number_of_broadphase_get_safe_interval"] == o.n_links * n_si
    assert d["number_of_broadphase_collision_checks"] == o.n_links * n_cc
    assert 0 < d["number_of_detected_collisions"] <= d["b200"]["motion_checks"]
    assert 0 < d["number_of_safe_intervals"] and 0 < d["number_of_sum_of_safe_interval_frames"] <= o.n_frames * n_si
    # every interval holds at least one frame and two intervals of one query are separated by a colliding frame
    assert d["number_of_sum_of_safe_interval_frames"] >= d["number_of_safe_intervals"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["scene_task", "mass_test", "dyn0", "dyn1"])
def test_b200_planner_gives_identical_plans(planner_oracle, scenes, pkg, name):
    g = __import__("__graft_entry__")
    g.load_build().build_host()
    exe = os.path.join(PKG, "host", "sirrt_plan")
    for seed in (1, 2, 3, 4):
        ref = run_json([planner_oracle, scenes[name], str(seed), "--max-time", "20"])
        for extra in ([], ["--no-edge-table"]):
            got = run_json([exe, scenes[name], "/tmp", str(seed), "--max-time", "20", "--stdout"] + extra)
            assert got["final_planner_data"] == ref["final_planner_data"], f"{name} seed {seed} {extra}: plans differ"
            for k in REFERENCE_LOG_KEYS:
                assert k in got
            for k in EXACT_COUNTERS:  # same tree growth, same queries, same answers -> the same counters
                assert got[k] == ref[k], f"{name} seed {seed} {extra}: {k} {got[k]} != {ref[k]}"
