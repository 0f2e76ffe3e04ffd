#!/usr/bin/env python
"""Synthetic mass test (BASELINE.json config 5): independent random dynamic-obstacle planning tasks sharded across the
GPUs of one box, next to the reference's CPU model (mass_test/do_mass_test.py:47-60: one planner PROCESS per
(scene, seed) in a ProcessPoolExecutor of NUM_CPUS workers).

Task recipe (Blender/scripts/generate_mass_test.py:17-60,66-78; do_mass_test.py:47-48): xArm6 with its visual-mesh
hulls (the robot of STRRT/mass_test.json), 600 frames @ 30 fps, k in {1, 20, 40, ..., 300} bouncing cubes of edge
U(0.05, 0.1) m inside [-1, 1]^3 at up to 0.2 m/s per axis, start / goal pairs with |dq|_1 > pi that pass the
check_scene rule (start free at frame 0, goal free at the last frame; check_scene.cpp:51-103 -- evaluated here with the
GPU checker), several planner seeds per pair.  Everything is seeded with 42.

    python scripts/mass_test.py --gpus 1 --tasks 1024 --threads 16 --cpu-sample 128 > gpurun_out/mass_test.json

GPU arm: host/sirrt_mass, ONE process per GPU, `--threads` planner threads per process, tasks dealt round-robin by scene
to GPUs -- no inter-GPU traffic.  CPU arm: tests/_build/planner_oracle --baseline, one process per task, nproc workers,
on a bounded sample of the same tasks.  Reports tasks/s (wall clock, both arms), p50 solve time and plan equality on the
sampled tasks (path_cost and waypoint count; the planner tests check waypoint-by-waypoint equality)."""
import argparse, gzip, json, os, subprocess, sys, tempfile, time
from concurrent.futures import ThreadPoolExecutor
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g


def make_tasks(sp, pkg, n_tasks, pairs_per_scene, seeds_per_pair, tmp, frames=600):
    base = sp.read_pack(gzip.open(os.path.join(ROOT, "tests", "golden", "mass_test.pack.gz")).read())
    robot = base.robot
    counts = [1] + list(range(20, 301, 20))
    rng = np.random.default_rng(42)
    lo, hi = robot.limits[:, 0].astype(np.float64), robot.limits[:, 1].astype(np.float64)
    tasks, packs = [], []
    while len(tasks) < n_tasks:
        for k in counts:
            sc = sp.synthetic_scene(k, frames, 30, int(rng.integers(1 << 30)), "A", robot=robot)
            sc.joint_max_velocity = base.joint_max_velocity.copy()
            dev = pkg.Scene(sp.write_pack(sc))
            got = 0
            for _ in range(200):
                a = lo + rng.random(robot.dof) * (hi - lo)
                b = lo + rng.random(robot.dof) * (hi - lo)
                if np.abs(a - b).sum() <= np.pi:
                    continue
                if dev.check_frame(a, 0) or dev.check_frame(b, frames - 1):
                    continue
                sc.start_configuration, sc.end_configuration = a.copy(), b.copy()
                path = os.path.join(tmp, f"k{k}_p{len(packs)}.pack")
                with open(path, "wb") as f:
                    f.write(sp.write_pack(sc))
                packs.append(path)
                for _s in range(seeds_per_pair):
                    tasks.append((path, int(rng.integers(1, 1 << 30)), k))
                got += 1
                if got == pairs_per_scene:
                    break
            dev.close()
            if len(tasks) >= n_tasks:
                break
    return tasks[:n_tasks]


def run_gpu_arm(tasks, gpus, threads, max_time, tmp, sync="spin"):
    exe = os.path.join(ROOT, "manipulationplanning-si-rrt_b200", "host", "sirrt_mass")
    # ONE task list for every per-GPU process, scenes ordered heaviest (most cubes) first; the processes drain it through a
    # shared counter (--queue), a scene at a time, so no GPU idles while another still has scenes queued
    order = sorted(range(len(tasks)), key=lambda i: (-tasks[i][2], tasks[i][0], i))
    lst, queue = os.path.join(tmp, "tasks_all.txt"), os.path.join(tmp, "queue.bin")
    with open(lst, "w") as f:
        for i in order:
            f.write(f"{tasks[i][0]} {tasks[i][1]}\n")
    with open(queue, "wb") as f:
        f.write(b"\0" * 8)
    t0 = time.perf_counter()
    logs = [(open(os.path.join(tmp, f"gpu{d}.out"), "w+"), open(os.path.join(tmp, f"gpu{d}.err"), "w+")) for d in range(gpus)]
    env = dict(os.environ, SIRRT_SYNC=sync)
    procs = [subprocess.Popen([exe, lst, str(d), str(threads), "--max-time", str(max_time), "--queue", queue], stdout=logs[d][0], stderr=logs[d][1], env=env)
             for d in range(gpus)]
    for p in procs:
        p.wait()
    wall = time.perf_counter() - t0
    outs = []
    for o, e in logs:
        o.seek(0); e.seek(0)
        outs.append((o.read(), e.read()))
    rows, summaries = {}, []
    for d, (so, se) in enumerate(outs):
        if procs[d].returncode != 0:
            sys.stderr.write(se[-2000:])
        for line in so.splitlines():
            r = json.loads(line)
            if r.get("summary"):
                summaries.append(r)
            else:
                r["device"] = d
                rows[order[r["task"]]] = r
    return wall, rows, summaries


def run_cpu_arm(tasks, sample, workers, max_time):
    exe = os.path.join(ROOT, "tests", "_build", "planner_oracle")

    def one(i):
        p, seed, _ = tasks[i]
        r = subprocess.run([exe, p, str(seed), "--max-time", str(max_time), "--baseline"], capture_output=True, text=True)
        return i, (json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 and r.stdout.strip() else None)

    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=workers) as ex:  # threads only wait on the child processes (one process per task, like os.system in do_mass_test.py)
        res = dict(ex.map(one, sample))
    return time.perf_counter() - t0, res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--tasks", type=int, default=1024)
    ap.add_argument("--pairs-per-scene", type=int, default=8)
    ap.add_argument("--seeds-per-pair", type=int, default=8)
    ap.add_argument("--threads", type=int, default=0, help="planner threads per GPU process (0 = host cores / GPUs, clamped to [4, 32])")
    ap.add_argument("--cpu-sample", type=int, default=128)
    ap.add_argument("--max-time", type=float, default=5.0)
    ap.add_argument("--sync", default="spin", help="spin | yield (SIRRT_SYNC of the planner processes; yield when threads exceed cores)")
    ap.add_argument("--also", default="", help="extra GPU arms to time on the same tasks, e.g. '8:yield,16:yield' (threads per GPU : sync)")
    a = ap.parse_args()
    if a.threads <= 0:
        a.threads = max(4, min(32, (os.cpu_count() or 16) // a.gpus))
    g.load_build().build_lib(); g.load_build().build_host()
    sp, pkg = g.load_tools(), g.load_package()
    import importlib.util
    spec = importlib.util.spec_from_file_location("plan_time", os.path.join(ROOT, "scripts", "plan_time.py"))
    pt = importlib.util.module_from_spec(spec); spec.loader.exec_module(pt)
    pt.build_oracle_planner()
    tmp = tempfile.mkdtemp()
    t0 = time.perf_counter()
    tasks = make_tasks(sp, pkg, a.tasks, a.pairs_per_scene, a.seeds_per_pair, tmp)
    gen_s = time.perf_counter() - t0
    wall, rows, summaries = run_gpu_arm(tasks, a.gpus, a.threads, a.max_time, tmp, a.sync)
    extra = []
    for spec in [x for x in a.also.split(",") if x]:
        th, sy = spec.split(":")
        w2, r2, _ = run_gpu_arm(tasks, a.gpus, int(th), a.max_time, tmp, sy)
        ms = [r["solve_ms"] for r in r2.values()]
        same_out = sum(1 for i in r2 if i in rows and r2[i]["solved"] and rows[i]["solved"] and r2[i]["path_hash"] == rows[i]["path_hash"])
        extra.append({"threads_per_gpu": int(th), "sync": sy, "wall_s": w2, "tasks_per_s": len(r2) / w2, "solved": sum(1 for r in r2.values() if r["solved"]),
                      "p50_solve_ms": float(np.median(ms)), "solved_by_both_with_identical_path": same_out,
                      "solved_by_both": sum(1 for i in r2 if i in rows and r2[i]["solved"] and rows[i]["solved"]),
                      "mismatches": [{"task": i, "pack": os.path.basename(tasks[i][0]), "seed": tasks[i][1], "iters": [rows[i]["iterations"], r2[i]["iterations"]],
                                      "cost": [rows[i]["cost"], r2[i]["cost"]], "device": [rows[i]["device"], r2[i]["device"]]}
                                     for i in r2 if i in rows and r2[i]["solved"] and rows[i]["solved"] and r2[i]["path_hash"] != rows[i]["path_hash"]]})
    ncpu = os.cpu_count() or 1
    step = max(1, len(tasks) // max(1, a.cpu_sample))
    sample = list(range(0, len(tasks), step))[: a.cpu_sample]
    cpu_wall, cpu = run_cpu_arm(tasks, sample, ncpu, a.max_time) if a.cpu_sample > 0 else (0.0, {})
    # plans are a function of (scene, seed) alone; the wall-clock limit is the only thing that can make the arms differ
    # (one arm solves a task the other runs out of time on), so equality is judged where both arms solved the task
    same = both = both_solved = 0
    for i, r in cpu.items():
        if r is None or i not in rows:
            continue
        both += 1
        fp = r["final_planner_data"]
        gr = rows[i]
        if fp["has_result"] and gr["solved"]:
            both_solved += 1
            same += abs(fp["path_cost"] - gr["cost"]) < 1e-12 and len(fp["final_path"]) == gr["waypoints"]
    gms = [r["solve_ms"] for r in rows.values()]
    cms = [r["algorithm_solving_time_ns"] / 1e6 for r in cpu.values() if r]
    gms_s = [rows[i]["solve_ms"] for i in cpu if i in rows and cpu[i]]
    by_k = {}
    for i, r in rows.items():
        by_k.setdefault(tasks[i][2], []).append(r["solve_ms"])
    out = {
        "metric": "mass test: independent SI-RRT planning tasks per second", "n_gpus": a.gpus, "tasks": len(tasks), "task_generation_s": gen_s,
        "recipe": "xArm6 (visual-mesh hulls), 600 frames @ 30 fps, k in {1,20,...,300} bouncing cubes, |dq|_1 > pi, check_scene rule, seed 42",
        "gpu_arm_variants": extra,
        "gpu_arm": {"processes": a.gpus, "threads_per_gpu": a.threads, "sync": a.sync, "wall_s": wall, "tasks_per_s": len(rows) / wall if wall > 0 else None,
                    "solved": sum(1 for r in rows.values() if r["solved"]), "finished": len(rows), "p50_solve_ms": float(np.median(gms)) if gms else None,
                    "mean_solve_ms": float(np.mean(gms)) if gms else None, "per_process": summaries},
        "cpu_arm": {"kind": "port (oracle baseline mode), one process per task", "workers": ncpu, "sample": len(sample), "wall_s": cpu_wall,
                    "tasks_per_s": len(cms) / cpu_wall if cpu_wall > 0 else None, "solved": sum(1 for r in cpu.values() if r and r["final_planner_data"]["has_result"]),
                    "p50_solve_ms": float(np.median(cms)) if cms else None, "mean_solve_ms": float(np.mean(cms)) if cms else None,
                    "gpu_p50_solve_ms_same_tasks": float(np.median(gms_s)) if gms_s else None},
        "plan_equality_on_sample": {"compared": both, "solved_by_both_arms": both_solved, "identical_plans_among_those": int(same)},
        "p50_solve_ms_by_cube_count": {str(k): float(np.median(v)) for k, v in sorted(by_k.items())},
        "max_time_s": a.max_time,
    }
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
