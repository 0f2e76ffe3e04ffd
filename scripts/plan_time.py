#!/usr/bin/env python
"""SI-RRT p50 plan time (BASELINE.json metric): host/sirrt_plan on the B200 backend (with and
without the §7.3 edge table) next to the SAME planner on the CPU oracle (tests/_build/planner_oracle, the
reference's single-threaded execution model), same seeds, with plan equality.  Run on the GPU box:
    python scripts/plan_time.py --seeds 10 > gpurun_out/plan_time.json
Times are `algorithm_solving_time_ns` (the key the reference logs, ResultsWriter.cpp:74)."""
import argparse, gzip, json, os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g


def build_oracle_planner():
    orc = g.load_oracle(); orc.build()
    exe = os.path.join(ROOT, "tests", "_build", "planner_oracle")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    odir = os.path.join(ROOT, "oracle")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-fopenmp", "-o", exe, os.path.join(ROOT, "tests", "planner_oracle_main.cpp"), "-L", odir, "-lsirrt_oracle", f"-Wl,-rpath,{odir}"])
    return exe, orc


def golden(name):
    with gzip.open(os.path.join(ROOT, "tests", "golden", name + ".pack.gz"), "rb") as f:
        return f.read()


def synthetic_task(sp, orc, robot, n_obs, frames, seed0):
    for seed in range(seed0, seed0 + 200):
        sc = sp.synthetic_scene(n_obs, frames, 30, seed, "A", robot=robot)
        o = orc.Oracle(sp.write_pack(sc))
        q = sp.random_configs(robot, 200, seed).T.astype(np.float64)
        free0 = [i for i in range(len(q)) if not o.check_frame(q[i], 0)]
        freeT = [i for i in range(len(q)) if not o.check_frame(q[i], frames - 1)]
        pair = next(((a, b) for a in free0 for b in freeT if np.abs(q[a] - q[b]).sum() > np.pi), None)
        if pair:
            sc.start_configuration, sc.end_configuration = q[pair[0]].copy(), q[pair[1]].copy()
            return sp.write_pack(sc)
    raise RuntimeError("no valid synthetic task")


def run(cmd):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    if r.returncode != 0:
        return None
    return json.loads(r.stdout.strip().splitlines()[-1])


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=10)
    ap.add_argument("--max-time", type=float, default=20.0)
    ap.add_argument("--scenes", default="scene_task,mass_test,flying_boxes,multiple_robots,dyn40,dyn120,dyn200")
    ap.add_argument("--exact-cpu", action="store_true", help="also time the CPU planner on the exact (unaccelerated) oracle mode")
    a = ap.parse_args()
    g.load_build().build_lib(); g.load_build().build_host()
    sp = g.load_tools()
    ora_exe, orc = build_oracle_planner()
    gpu_exe = os.path.join(ROOT, "manipulationplanning-si-rrt_b200", "host", "sirrt_plan")
    cyl = sp.read_pack(golden("xarm6cyl_robot")).robot
    tmp = tempfile.mkdtemp()
    out = {"metric": "SI-RRT p50 plan time (algorithm_solving_time_ns)", "seeds": a.seeds,
           "cpu_arm": "host/SIRRTPlanner.hpp on oracle/ in baseline mode: AABB tree over constant-pose windows for safe intervals, AABB gate + boolean "
                      "GJK with first-hit exit per state (the reference's coal query pattern), single thread (the reference's execution model)",
           "scenes": {}}
    for name in a.scenes.split(","):
        if name.startswith("dyn"):
            pk = synthetic_task(sp, orc, cyl, int(name[3:]), 600, 5)
        else:
            pk = golden(name)
        path = os.path.join(tmp, name + ".pack")
        with open(path, "wb") as f:
            f.write(pk)
        rows = {"cpu_oracle": [], "b200_edge_table": [], "b200_per_edge_batches": []}
        mode = []  # REFERENCE_EXACT everywhere (the robot-obstacle scene's start / goal are free in both parity modes)
        same, solved, counters = 0, 0, {}
        for seed in range(1, a.seeds + 1):
            ref = run([ora_exe, path, str(seed), "--max-time", str(a.max_time), "--baseline"] + mode)
            if a.exact_cpu:
                slow = run([ora_exe, path, str(seed), "--max-time", str(a.max_time)] + mode)
                if slow:
                    rows.setdefault("cpu_oracle_exact_mode", []).append(slow["algorithm_solving_time_ns"] / 1e6)
            t1 = run([gpu_exe, path, tmp, str(seed), "--max-time", str(a.max_time), "--stdout"] + mode)
            t2 = run([gpu_exe, path, tmp, str(seed), "--max-time", str(a.max_time), "--stdout", "--no-edge-table"] + mode)
            if not (ref and t1 and t2):
                continue
            rows["cpu_oracle"].append(ref["algorithm_solving_time_ns"] / 1e6)
            rows["b200_edge_table"].append(t1["algorithm_solving_time_ns"] / 1e6)
            rows["b200_per_edge_batches"].append(t2["algorithm_solving_time_ns"] / 1e6)
            solved += bool(ref["final_planner_data"]["has_result"])
            same += ref["final_planner_data"] == t1["final_planner_data"] == t2["final_planner_data"]
            counters = {"safe_interval_calls": ref["number_of_get_safe_interval"], "state_checks": ref["number_of_collision_checks"], **t1.get("b200", {})}
        out["scenes"][name] = {"runs": len(rows["cpu_oracle"]), "solved": solved, "identical_plans": same,
                               "p50_ms": {k: float(np.median(v)) if v else None for k, v in rows.items()},
                               "mean_ms": {k: float(np.mean(v)) if v else None for k, v in rows.items()}, "last_run_counters": counters}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
