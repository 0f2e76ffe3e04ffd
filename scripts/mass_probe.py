#!/usr/bin/env python
"""Where the mass-test time goes: per-call latencies (host/sirrt_latency) and planner counters (host/sirrt_plan) on
mass-test recipe scenes (xArm6 hull robot, k bouncing cubes, 600 frames).  Run on the GPU box."""
import gzip, json, os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g

g.load_build().build_lib(); g.load_build().build_host()
sp, pkg = g.load_tools(), g.load_package()
base = sp.read_pack(gzip.open(os.path.join(ROOT, "tests", "golden", "mass_test.pack.gz")).read())
robot = base.robot
tmp = os.environ.get("MASS_PROBE_DIR") or tempfile.mkdtemp()
rng = np.random.default_rng(42)
lo, hi = robot.limits[:, 0].astype(np.float64), robot.limits[:, 1].astype(np.float64)
host = os.path.join(ROOT, "manipulationplanning-si-rrt_b200", "host")
for k in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "20,100,300").split(",")]:
    sc = sp.synthetic_scene(k, 600, 30, 1000 + k, "A", robot=robot)
    sc.joint_max_velocity = base.joint_max_velocity.copy()
    dev = pkg.Scene(sp.write_pack(sc))
    while True:
        a = lo + rng.random(robot.dof) * (hi - lo); b = lo + rng.random(robot.dof) * (hi - lo)
        if np.abs(a - b).sum() > np.pi and not dev.check_frame(a, 0) and not dev.check_frame(b, 599):
            break
    dev.close()
    sc.start_configuration, sc.end_configuration = a, b
    path = os.path.join(tmp, f"k{k}.pack")
    open(path, "wb").write(sp.write_pack(sc))
    lat = subprocess.run([os.path.join(host, "sirrt_latency"), path, "200"], capture_output=True, text=True)
    print(json.dumps({"k": k, "latency": json.loads(lat.stdout) if lat.returncode == 0 else lat.stderr[-300:]}), flush=True)
    for seed in (1, 2):
        r = subprocess.run([os.path.join(host, "sirrt_plan"), path, tmp, str(seed), "--max-time", "5", "--stdout"], capture_output=True, text=True)
        try:
            d = json.loads(r.stdout.strip().splitlines()[-1])
            print(json.dumps({"k": k, "seed": seed, "solve_ms": d["algorithm_solving_time_ns"] / 1e6, "solved": d["final_planner_data"]["has_result"],
                              **{kk: d[kk] for kk in d if kk.startswith("number_of") or kk.endswith("_ns")}, "b200": d.get("b200")}), flush=True)
        except Exception as e:
            print("plan failed", e, r.stderr[-300:], flush=True)
