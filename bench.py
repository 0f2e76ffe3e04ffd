#!/usr/bin/env python
"""bench.py -- config x timestep collision checks/s of the safe-interval hot path on B200, and SI-RRT p50 plan time.

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): synthetic batched safe-interval sweep,
7-DOF capsule arm, ONE batch of 2^20 configurations x 1000 frames x 64 moving boxes.  One "step" = one full pass over
the batch: collision bit rows + FP64 re-check of the near-contact band + fixed-width interval table (Kmax=8).

N = 1   the whole batch on one GPU.
N > 1   the SAME batch split into N contiguous blocks ("split N configs into G contiguous blocks"), one
        rank per GPU, scene tables replicated; every step ends with an NCCL all-gather of the interval-table shards so
        that every rank holds all 2^20 rows -> "scaling": "strong".  Rank 0 then recomputes 512 random rows of the
        gathered table on its own GPU and asserts equality (multi-GPU correctness inside the driver-run record).  The
        weak-scaling figure (every rank sweeps a full 2^20 batch, no collective) is kept under results.weak_scaling.

  value     whole-job checks/s, inputs and outputs resident in HBM (CUDA events on the launch stream, max over ranks)
  e2e       the same pass through the C ABI with pinned HOST buffers (H2D of q, D2H of the table inside the timed region)
  roofline  dominant kernel (k_sweep): executed FP32 FLOP and warp instructions per check come from the committed ncu
            capture profiles/ksweep_constants.json (scripts/ncu_summary.py --emit-constants), which records the SHA-256 of
            the kernel sources it was taken from; if the working tree's sources differ the constants are NOT used
  cpu_baseline  the oracle's AABB-tree restatement of the reference algorithm on the host cores (bounded sample)
  results.plan_time  SI-RRT p50 plan time (algorithm_solving_time_ns, ResultsWriter.cpp:74): host/sirrt_plan on the B200
            backend next to the same planner on the CPU oracle, 3 seeds x 4 reference scenes, with plan identity

`--impl reference` times the CPU restatement alone (the reference itself cannot be built in this image: coal, Eigen,
nanoflann, urdfdom, orocos-kdl are absent and not vendored).
"""
from __future__ import annotations

import argparse
import gzip
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402

METRIC = "config x timestep collision checks/sec"
UNIT = "checks/s"
# SURVEY.md §8d accounting (frozen before measuring; DESIGN.md "Roofline accounting")
F_CULL, F_FK, F_EXACT_CAPSULE_BOX, F_EXACT_CAPSULE_SPHERE = 12.0, 1100.0, 150.0, 30.0
N_SM, SCHEDULERS_PER_SM = 148, 4
CONSTANTS = os.path.join(ROOT, "profiles", "ksweep_constants.json")
KERNEL_SOURCES = ["sirrt_sweep.cuh", "sirrt_sweep_window.cuh", "sirrt_device.cuh", "sirrt_kernels.cuh"]


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced shard [n0, n1) of n units for `rank` of `world` (no overlap, no gap)."""
    return n * rank // world, n * (rank + 1) // world


def source_hashes():
    """SHA-256 of the sources k_sweep is compiled from (the roofline constants are only valid for these)"""
    out = {}
    for f in KERNEL_SOURCES:
        with open(os.path.join(g.PKG_DIR, "csrc", f), "rb") as fh:
            out[f] = hashlib.sha256(fh.read()).hexdigest()
    return out


def load_constants():
    """executed work of k_sweep per check from the committed ncu capture; None (with the reason) when it is stale"""
    if not os.path.exists(CONSTANTS):
        return None, "profiles/ksweep_constants.json missing"
    with open(CONSTANTS) as f:
        c = json.load(f)
    if c.get("source_sha256") != source_hashes():
        return None, "stale: kernel sources changed since the ncu capture behind profiles/ksweep_constants.json (re-run scripts/ncu_summary.py --emit-constants)"
    return c, None


def gather_table_shards(table_shard, n_total: int, rank: int, world: int, dist, group=None):
    """all ranks contribute the interval table of their contiguous block [shard_range) and receive all n_total rows.
    Equal blocks go through one all_gather_into_tensor; blocks that differ by a row are padded to the largest."""
    import torch

    sizes = [shard_range(n_total, r, world)[1] - shard_range(n_total, r, world)[0] for r in range(world)]
    n_max = max(sizes)
    cols = table_shard.shape[1]
    if min(sizes) == n_max:
        out = torch.empty((n_total, cols), dtype=table_shard.dtype, device=table_shard.device)
        dist.all_gather_into_tensor(out, table_shard.contiguous(), group=group)
        return out
    pad = torch.zeros((n_max, cols), dtype=table_shard.dtype, device=table_shard.device)
    pad[: table_shard.shape[0]] = table_shard
    buf = torch.empty((world * n_max, cols), dtype=table_shard.dtype, device=table_shard.device)
    dist.all_gather_into_tensor(buf, pad, group=group)
    return torch.cat([buf[r * n_max: r * n_max + sizes[r]] for r in range(world)])


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """SM clocks / throttle reasons of one GPU sampled DURING the timed region: NVML in-process every 2 ms (a timed region of a
    few tens of milliseconds at N = 8 still gets samples; `nvidia-smi -lms` needs ~100 ms to deliver its first line)."""

    def __init__(self, index: int):
        self.index, self.rows, self.t, self.h, self.nv = index, [], None, None, None
        self.stop_flag = threading.Event()

    def start(self):
        try:
            import pynvml as nv

            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = self.index
            if vis:
                ent = vis.split(",")[self.index].strip()
                phys = int(ent) if ent.isdigit() else None
                self.h = nv.nvmlDeviceGetHandleByUUID(ent) if phys is None else nv.nvmlDeviceGetHandleByIndex(phys)
            else:
                self.h = nv.nvmlDeviceGetHandleByIndex(phys)
            self.nv = nv
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
            self._sample()  # one sample at the start of the region whatever its length
            self.t = threading.Thread(target=self._loop, daemon=True)
            self.t.start()
        except Exception:
            self.nv = None

    def _sample(self):
        nv = self.nv
        sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
        try:
            rs = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.h))
        except Exception:
            rs = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
        self.rows.append((sm, rs))

    def _loop(self):
        while not self.stop_flag.wait(0.002):
            try:
                self._sample()
            except Exception:
                return

    def stop(self):
        if not self.nv:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["NVML unavailable"], "samples": 0}
        try:
            self._sample()  # and one at its end
        except Exception:
            pass
        self.stop_flag.set()
        if self.t:
            self.t.join(timeout=1)
        nv = self.nv
        bits = {"hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8), "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
                "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20), "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4)}
        reasons = sorted(n for n, b in bits.items() if any(r & b for _, r in self.rows))
        sm = [r[0] for r in self.rows]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.mx, "reasons": reasons, "samples": len(sm), "source": "NVML, 2 ms period"}


def build_workload(a):
    sp = g.load_tools()
    scene = sp.synthetic_scene(a.obstacles, a.frames, 30, 42, a.variant)
    pack = sp.write_pack(scene)
    q = sp.random_configs(scene.robot, a.configs, 42)
    return sp, scene, pack, q


def cpu_baseline(pack, q, budget_s: float, threads: int):
    """The oracle's restatement of the reference ALGORITHM (FK + AABB tree over constant-pose windows + exact
    narrowphase, CollisionManager.cpp:107-216,485-547), all host threads, on a bounded sample of the same workload."""
    orc = g.load_oracle()
    orc.build()
    o = orc.Oracle(pack)
    o.bvh_build()
    probe = np.ascontiguousarray(q[:, : 8 * threads])
    t = time.perf_counter()
    o.safe_intervals_batch(probe, 8, want_bits=False, use_bvh=True, n_threads=threads)
    dt = max(time.perf_counter() - t, 1e-6)
    n = int(min(q.shape[1], max(8 * threads, probe.shape[1] * budget_s / dt)))
    sample = np.ascontiguousarray(q[:, :n])
    t = time.perf_counter()
    o.safe_intervals_batch(sample, 8, want_bits=False, use_bvh=True, n_threads=threads)
    dt = time.perf_counter() - t
    return n * o.n_frames / dt, n, dt


def run_reference(a, rank, world):
    if rank != 0:
        return
    sp, scene, pack, q = build_workload(a)
    threads = os.cpu_count() or 1
    per_step = a.ref_budget / max(1, a.steps + a.warmup)
    rates, n_used = [], 0
    for i in range(a.warmup + a.steps):
        r, n_used, dt = cpu_baseline(pack, q, per_step, threads)
        if i >= a.warmup:
            rates.append((n_used * a.frames, dt))
    tot_checks = sum(c for c, _ in rates)
    tot_t = sum(t for _, t in rates)
    v = tot_checks / tot_t
    sample = f"{n_used} of {a.configs} configurations x {a.frames} frames per step"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": 1e3 * tot_t / max(1, a.steps), "higher_is_better": True, "scaling": "strong" if a.gpus > 1 else "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(a, max(1, a.gpus)),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                         "note": "oracle/sirrt_oracle.c AABB-tree restatement of the reference algorithm, OpenMP over configurations; the reference binary is unbuildable here"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(a, world):
    return {"workload": f"synthetic batched safe-interval sweep: 7-DOF capsule arm (arm7cyl.urdf), ONE batch of {a.configs} configs x {a.frames} frames x {a.obstacles} moving "
                        f"{'spheres' if a.variant == 'C' else 'boxes'} (variant {a.variant}), Kmax={a.kmax}",
            "configs_total": a.configs, "configs_per_gpu": (a.configs + world - 1) // world, "frames": a.frames, "obstacles": a.obstacles, "links": 7, "kmax": a.kmax,
            "sharding": ("whole batch on one GPU" if world == 1 else
                         f"the batch split into {world} contiguous blocks, scene replicated, NCCL all-gather of the interval-table shards every step"),
            "l2": "flushed (256 MiB write) between timed steps"}


# ---------------------------------------------------------------------------------------------- plan time
def plan_time_leg(seeds: int, max_time: float, device: int):
    """SI-RRT p50 plan time (the other half of BASELINE.json's metric): host/sirrt_plan (B200 backend) and the same planner
    on the CPU oracle (tests/_build/planner_oracle, single thread = the reference's execution model), same seeds; a plan
    counts as identical when final_planner_data (has_result + every waypoint) is equal."""
    gpu_exe = os.path.join(g.PKG_DIR, "host", "sirrt_plan")
    cpu_exe = os.path.join(ROOT, "tests", "_build", "planner_oracle")
    if not (os.path.exists(gpu_exe) and os.path.exists(cpu_exe)):
        return {"unavailable": "host/sirrt_plan or tests/_build/planner_oracle not built (run __graft_entry__.build())"}
    out = {"metric": "SI-RRT p50 plan time, ms (algorithm_solving_time_ns, ResultsWriter.cpp:74)", "seeds": seeds, "max_time_s": max_time, "scenes": {}}
    tmp = tempfile.mkdtemp()

    def run(cmd):
        try:
            r = subprocess.run(cmd, capture_output=True, text=True, timeout=max_time * 4 + 30)
            return json.loads(r.stdout.strip().splitlines()[-1]) if r.returncode == 0 else None
        except Exception:
            return None

    for name in ("scene_task", "mass_test", "flying_boxes", "multiple_robots"):
        with gzip.open(os.path.join(ROOT, "tests", "golden", name + ".pack.gz"), "rb") as f:
            path = os.path.join(tmp, name + ".pack")
            with open(path, "wb") as o:
                o.write(f.read())
        gpu_ms, cpu_ms, same, solved = [], [], 0, 0
        for seed in range(1, seeds + 1):
            c = run([cpu_exe, path, str(seed), "--max-time", str(max_time), "--baseline"])
            d = run([gpu_exe, path, tmp, str(seed), "--max-time", str(max_time), "--stdout", "--device", str(device)])
            if not (c and d):
                continue
            cpu_ms.append(c["algorithm_solving_time_ns"] / 1e6)
            gpu_ms.append(d["algorithm_solving_time_ns"] / 1e6)
            solved += bool(d["final_planner_data"]["has_result"])
            same += c["final_planner_data"] == d["final_planner_data"]
        out["scenes"][name] = {"runs": len(gpu_ms), "solved": solved, "identical_plans": f"{same}/{len(gpu_ms)}",
                               "p50_ms_b200": float(np.median(gpu_ms)) if gpu_ms else None, "p50_ms_cpu": float(np.median(cpu_ms)) if cpu_ms else None}
    return out


def latency_leg(calls: int = 300):
    """Per-call latency of the planner-in-the-loop entry points over the C ABI (host/sirrt_latency, a C++ program: no Python in
    the loop), microseconds, on the three shipped dynamic scenes -- what the SI-RRT plan time is made of."""
    exe = os.path.join(g.PKG_DIR, "host", "sirrt_latency")
    if not os.path.exists(exe):
        return {"unavailable": "host/sirrt_latency not built (run __graft_entry__.build())"}
    out = {"unit": "us per call (mean of %d calls after warm-up)" % calls, "scenes": {}}
    tmp = tempfile.mkdtemp()
    for name in ("mass_test", "flying_boxes", "multiple_robots"):
        path = os.path.join(tmp, name + ".pack")
        with gzip.open(os.path.join(ROOT, "tests", "golden", name + ".pack.gz"), "rb") as f, open(path, "wb") as o:
            o.write(f.read())
        try:
            r = subprocess.run([exe, path, str(calls)], capture_output=True, text=True, timeout=120)
            out["scenes"][name] = json.loads(r.stdout.strip().splitlines()[-1])["us"] if r.returncode == 0 else {"failed": r.stderr[-200:]}
        except Exception as e:  # a leg of the report, never the bench itself
            out["scenes"][name] = {"failed": str(e)[:200]}
    return out


def mass_test_leg(tasks: int, device_count: int):
    """BASELINE.json config 5 in small: scripts/mass_test.py (host/sirrt_mass: one process per GPU, 16 planner threads each, a
    shared task queue) on `tasks` synthetic planning tasks (one start / goal pair x 6 planner seeds for each of the 16 cube counts) with a 5 s budget, next to the reference's execution model (one CPU
    planner process per task, one worker per core) on a 16-task sample.  Condensed; the full record is the script's own JSON."""
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", "mass_test.py"), "--gpus", str(device_count), "--tasks", str(tasks), "--pairs-per-scene", "1", "--seeds-per-pair", "6", "--threads", "16",
                            "--sync", "yield", "--cpu-sample", "16", "--max-time", "5"], capture_output=True, text=True, timeout=600)
        d = json.loads(r.stdout)
    except Exception as e:  # the leg is a side figure: never fail the bench line over it
        return {"unavailable": f"scripts/mass_test.py failed: {e}"}
    ga, ca = d["gpu_arm"], d["cpu_arm"]
    return {"metric": d["metric"], "tasks": d["tasks"], "max_time_s": d["max_time_s"], "recipe": d["recipe"],
            "b200": {"gpus": d["n_gpus"], "threads_per_gpu": ga["threads_per_gpu"], "tasks_per_s": ga["tasks_per_s"], "solved": ga["solved"], "p50_solve_ms": ga["p50_solve_ms"]},
            "cpu": {"workers": ca["workers"], "sample": ca["sample"], "tasks_per_s": ca["tasks_per_s"], "solved": ca["solved"], "p50_solve_ms": ca["p50_solve_ms"]},
            "plan_equality_on_sample": d["plan_equality_on_sample"]}


def knn_leg(pkg, device: int, hbm_gbs: float, fp32_tflops: float):
    """C-space nearest-neighbour throughput (MSIRRT/src/Tree.cpp:18-36, PlannerConnect.cpp:382-412): queries/s through
    the C ABI with device-resident queries and results (1-NN: tiled multi-query kernel; top-8: single-pass per-warp lists) and
    with host buffers one query at a time (radius: what the planner calls once per iteration), D = 6.  Every figure is the wall
    time of the synchronous API call, averaged over repetitions; 64 answers per size are checked against a NumPy brute force."""
    import time

    import torch
    from scipy.spatial import cKDTree

    oracle = g.load_oracle()
    D = 6
    rng = np.random.default_rng(3)
    out = {"dim": D, "unit": "queries/s", "sizes": {},
           "roofline_note": "1-NN: each 256-query tile streams the FP32 row copy once (32 B per point) and executes 18 FLOP per (query, point) pair; "
                            "hbm_frac = tiles * 32 N / time / hbm peak, fp32_frac = 18 * pairs / time / measured FFMA peak. The kernel is bound by FP32 issue slots, "
                            "not by HBM: the store (32 MB at N = 1e6) stays in the 126 MB L2."}
    for n in (1000, 10000, 100000, 1000000):
        pts = rng.uniform(-3.2, 3.2, (n, D))
        idx = pkg.KnnIndex(D, n, device)
        idx.add(pts)
        nq = 65536 if n <= 100000 else 16384
        qh = rng.uniform(-3.2, 3.2, (nq, D))
        q = torch.from_numpy(qh).cuda()
        oi = torch.empty(nq, dtype=torch.int64, device="cuda")
        od = torch.empty(nq, dtype=torch.float64, device="cuda")

        def timed(fn, reps):
            fn()
            t0 = time.perf_counter()
            for _ in range(reps):
                fn()
            return (time.perf_counter() - t0) / reps

        t_nn = timed(lambda: pkg._ck(pkg._lib.sirrt_knn_nearest(idx._h, q.data_ptr(), nq, oi.data_ptr(), od.data_ptr())), 5)
        ref = np.empty(64, np.int64)
        for j in range(64):   # NumPy brute force, dimension-order accumulation (no oracle on this path)
            acc = np.zeros(n)
            for c in range(D):
                e = qh[j, c] - pts[:, c]
                acc = acc + e * e
            ref[j] = int(np.argmin(acc))
        ok = bool(np.array_equal(oi[:64].cpu().numpy(), ref))
        k = 8
        nqk = 4096 if n <= 100000 else 1024
        ti = torch.empty((nqk, k), dtype=torch.int64, device="cuda")
        td = torch.empty((nqk, k), dtype=torch.float64, device="cuda")
        t_k = timed(lambda: pkg._ck(pkg._lib.sirrt_knn_topk(idx._h, q.data_ptr(), nqk, k, ti.data_ptr(), td.data_ptr())), 3)
        ok = ok and bool(np.array_equal(ti[:64, 0].cpu().numpy(), oi[:64].cpu().numpy()))
        hits = []
        t0 = time.perf_counter()
        for j in range(200):
            hits.append(idx.radius(qh[j], 1.0, 4096)[1])
        t_r = (time.perf_counter() - t0) / 200
        t0 = time.perf_counter()
        for j in range(200):
            idx.nearest(qh[j : j + 1])
        t_1 = (time.perf_counter() - t0) / 200
        idx.close()
        # cpu baseline of this leg: the oracle's brute force (one thread) on a bounded sample, and a kd-tree (SciPy cKDTree, one thread: the data structure nanoflann implements)
        nq_cpu = int(max(8, min(nq, 2e8 // n)))
        t0 = time.perf_counter()
        oracle.knn_nearest(pts, qh[:nq_cpu])
        t_cpu = (time.perf_counter() - t0) / nq_cpu
        tree = cKDTree(pts)
        t0 = time.perf_counter()
        tree.query(qh[:4096], k=1, workers=1)
        t_kd = (time.perf_counter() - t0) / 4096
        pairs = float(nq) * n
        out["sizes"][str(n)] = {
            "nearest_batch_qps": nq / t_nn, "nearest_batch_queries": nq, "pairs_per_s": pairs / t_nn,
            "hbm_frac": ((nq + 255) // 256) * 32.0 * n / t_nn / 1e9 / hbm_gbs, "fp32_frac": 18.0 * pairs / t_nn / 1e12 / fp32_tflops if fp32_tflops else None,
            "top8_batch_qps": nqk / t_k, "top8_batch_queries": nqk,
            "radius_single_qps": 1.0 / t_r, "radius_mean_hits": float(np.mean(hits)), "nearest_single_qps": 1.0 / t_1,
            "single_query_hbm_frac": 48.0 * n / t_1 / 1e9 / hbm_gbs,
            "cpu_bruteforce_qps_1thread": 1.0 / t_cpu, "cpu_kdtree_qps_1thread": 1.0 / t_kd, "self_check_64_answers": ok}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--configs", type=int, default=1 << 20)
    ap.add_argument("--frames", type=int, default=1000)
    ap.add_argument("--obstacles", type=int, default=64)
    ap.add_argument("--variant", default="A", choices=["A", "B", "C"])
    ap.add_argument("--kmax", type=int, default=8)
    ap.add_argument("--cpu-budget", type=float, default=15.0, help="seconds of CPU work for the cpu_baseline leg")
    ap.add_argument("--ref-budget", type=float, default=60.0, help="seconds of CPU work for --impl reference")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-plan-time", action="store_true")
    ap.add_argument("--no-knn", action="store_true", help="skip the nearest-neighbour throughput leg (results.knn)")
    ap.add_argument("--mass-tasks", type=int, default=96, help="planning tasks of the mass-test leg (results.mass_test); 0 = skip")
    ap.add_argument("--no-weak", action="store_true", help="N > 1: skip the weak-scaling side figure")
    ap.add_argument("--plan-seeds", type=int, default=3)
    ap.add_argument("--generic", action="store_true", help="force the generic kernel instead of the sweep kernel")
    ap.add_argument("--print-source-hash", action="store_true", help="print the SHA-256 of the k_sweep sources and exit")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 0)
    if a.print_source_hash:
        print(json.dumps(source_hashes()))
        return

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if a.impl == "reference":
        run_reference(a, rank, world)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 backend has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    pkg = g.load_package()
    sp, scene, pack, q = build_workload(a)   # the same batch on every rank; rank r sweeps block r of it
    N, T, W, K = a.configs, a.frames, (a.frames + 31) // 32, a.kmax
    n0, n1 = shard_range(N, rank, world)
    n_loc = n1 - n0
    dev = pkg.Scene(pack, local)
    if a.generic:
        dev.set_option(pkg.OPT_USE_SWEEP, 0)
    # one explicit stream for everything of the step -- the sweep kernels (sirrt_scene_set_stream), the L2 flush, NCCL's
    # dependency edges and the timing events.  torch's DEFAULT stream is handle 0, which sirrt_scene_set_stream reads as "use
    # the scene's own stream": events on the default stream would then only see the kernels through the host's blocking
    # wait, and under N > 1 (the all-gather keeps the stream busy past the host's return) they under-measured the step.
    stream = torch.cuda.Stream(device=local)
    torch.cuda.set_stream(stream)
    dev.set_stream(stream.cuda_stream)
    assert stream.cuda_stream != 0
    q_loc = np.ascontiguousarray(q[:, n0:n1])
    q_dev = torch.from_numpy(q_loc).cuda()
    table_loc = torch.empty((n_loc, 1 + 2 * K), dtype=torch.int32, device="cuda")
    bits_dev = torch.empty((n_loc, W), dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    fp32_peak = pkg.measure_fp32_peak(local)
    gathered = [None]

    def step():
        near = dev.safe_intervals_batch_raw(q_dev.data_ptr(), n_loc, table_loc.data_ptr(), K, bits_dev.data_ptr())
        if world > 1:
            gathered[0] = gather_table_shards(table_loc, N, rank, world, dist)
        return near

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn, steps):
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        acc = {"kern_ms": 0.0, "launches": 0, "near": 0, "rechecked": 0, "exact": 0}
        sync_all()
        t_wall = time.perf_counter()
        for s0, s1 in ev:
            flush.fill_(1)  # evict L2 (126 MB) between timed steps; outside the event pair
            s0.record(stream)
            acc["near"] = fn()
            s1.record(stream)
            st = dev.last_stats()
            acc["kern_ms"] += st["sweep_kernel_ms"]; acc["launches"] += st["launches"]; acc["rechecked"] = st["rechecked"]; acc["exact"] = st["exact_tests"]
        sync_all()
        # host wall clock around the same loop (includes the L2 flushes between steps): cross-check of the event times
        acc["wall_ms_per_step"] = 1e3 * (time.perf_counter() - t_wall) / max(1, steps)
        acc["step_ms"] = [s0.elapsed_time(s1) for s0, s1 in ev]
        total_ms = sum(acc["step_ms"])
        if world > 1:
            t = torch.tensor([total_ms, acc["wall_ms_per_step"]], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            total_ms, acc["wall_ms_per_step"] = float(t[0].item()), float(t[1].item())
        return total_ms / max(1, steps), acc

    for _ in range(a.warmup):
        step()
    sampler = ClockSampler(local)
    sync_all()
    sampler.start()
    ms_per_step, acc = timed(step, a.steps)
    clocks = sampler.stop()
    value = N * T / (ms_per_step * 1e-3)
    launches = acc["launches"] + (a.steps if world > 1 else 0)  # + NCCL's all-gather kernel per step
    if world > 1:
        for k in ("near", "rechecked", "exact"):
            t = torch.tensor([acc[k]], dtype=torch.int64, device="cuda")
            dist.all_reduce(t)
            acc[k] = int(t.item())

    # ---- N > 1: rank 0 re-computes random rows of the GATHERED table on its own GPU
    multi_check = None
    if world > 1:
        full = gathered[0]
        assert full.shape == (N, 1 + 2 * K)
        if rank == 0:
            rows = np.sort(np.random.default_rng(7).choice(N, min(512, N), replace=False))
            rt, _, _ = dev.safe_intervals_batch(np.ascontiguousarray(q[:, rows]), K, want_bits=False)
            got = full[torch.from_numpy(rows).cuda()].cpu().numpy()
            assert np.array_equal(got, rt), "multi-GPU: gathered interval table differs from rank 0's own result on sampled rows"
            per_rank = np.bincount(np.minimum(rows * world // N, world - 1), minlength=world).tolist()
            multi_check = {"rows": int(len(rows)), "rows_per_rank_block": per_rank, "equal": True}

    # ---- N > 1: the weak-scaling side figure (every rank sweeps a full batch, no collective)
    weak = None
    if world > 1 and not a.no_weak:
        qf = torch.from_numpy(q).cuda()
        tf = torch.empty((N, 1 + 2 * K), dtype=torch.int32, device="cuda")
        bf = torch.empty((N, W), dtype=torch.int32, device="cuda")
        wfn = lambda: dev.safe_intervals_batch_raw(qf.data_ptr(), N, tf.data_ptr(), K, bf.data_ptr())  # noqa: E731
        wfn()
        sync_all()
        wms, _ = timed(wfn, max(1, min(a.steps, 3)))
        weak = {"value": world * N * T / (wms * 1e-3), "unit": UNIT, "ms_per_step": wms, "configs_per_gpu": N}
        del qf, tf, bf

    # ---- end to end through the C ABI with pinned host buffers (this rank's block in, its table block out)
    q_pin = torch.from_numpy(q_loc).pin_memory()
    table_pin = torch.empty((n_loc, 1 + 2 * K), dtype=torch.int32).pin_memory()
    e2e_steps = max(1, min(a.steps, 5))
    dev.safe_intervals_batch_raw(q_pin.data_ptr(), n_loc, table_pin.data_ptr(), K, None)
    sync_all()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        dev.safe_intervals_batch_raw(q_pin.data_ptr(), n_loc, table_pin.data_ptr(), K, None)
    torch.cuda.synchronize()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    assert np.array_equal(table_pin.numpy(), table_loc.cpu().numpy()), "e2e result differs from the device-resident result"
    e2e_value = N * T / (e2e_ms * 1e-3)
    # the same with the compact (int16) interval table: half the D2H bytes (sirrt_safe_intervals_batch_i16)
    table16_pin = torch.empty((n_loc, 1 + 2 * K), dtype=torch.int16).pin_memory()
    dev.safe_intervals_batch_i16_raw(q_pin.data_ptr(), n_loc, table16_pin.data_ptr(), K)
    sync_all()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        dev.safe_intervals_batch_i16_raw(q_pin.data_ptr(), n_loc, table16_pin.data_ptr(), K)
    torch.cuda.synchronize()
    e2e16_ms = 1e3 * (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t = torch.tensor([e2e16_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e16_ms = float(t.item())
    assert np.array_equal(table16_pin.numpy().astype(np.int32), table_pin.numpy()), "compact table differs from the int32 table"

    if rank == 0:
        hbm_peak, peak_src, _ = measured_peaks()
        bits = bits_dev.cpu().numpy().view(np.uint32)
        density = float(np.unpackbits(bits.view(np.uint8)).sum()) / (n_loc * T)
        k_ms = acc["kern_ms"] / max(1, a.steps)  # rank 0's k_sweep launches of one step
        f_exact = F_EXACT_CAPSULE_SPHERE if a.variant == "C" else F_EXACT_CAPSULE_BOX
        f_check = 7 * a.obstacles * F_CULL + F_FK / T + (acc["exact"] / (N * T)) * f_exact
        brute_tflops = n_loc * T * f_check / (k_ms * 1e-3) / 1e12 if k_ms > 0 else None
        consts, stale = load_constants()
        exec_ok = consts is not None and (not a.generic) and a.variant == "A" and a.obstacles == 64 and a.frames == 1000
        f_exec = i_exec = traffic = None
        if exec_ok:
            checks = consts["configs_per_launch"] * consts["frames"]
            f_exec, i_exec = consts["flop"] / checks, consts["warp_inst"] / checks
            traffic = consts["dram_read_bytes"] + consts["dram_write_bytes"]
        exec_tflops = n_loc * T * f_exec / (k_ms * 1e-3) / 1e12 if (k_ms > 0 and exec_ok) else None
        sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
        issue_peak = N_SM * SCHEDULERS_PER_SM * sm_hz  # warp instructions / s at the clock seen under load
        issue_rate = n_loc * T * i_exec / (k_ms * 1e-3) if (k_ms > 0 and exec_ok) else None
        alg_bytes = n_loc * 7 * 4 + a.obstacles * 12 * 4 * ((T + 31) // 32 * 32) + n_loc * W * 4
        roofline = {
            "bound": "fp32", "kernel": "k_sweep" if not a.generic else "k_collide_bits", "achieved": exec_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
            "frac": (exec_tflops / fp32_peak) if exec_tflops and fp32_peak else None,
            "traffic": traffic, "traffic_note": (f"bytes per k_sweep launch of {consts['configs_per_launch']} configurations (ncu dram__bytes_read.sum + dram__bytes_write.sum, "
                                                 f"{consts.get('summary')})" if exec_ok else stale),
            "peak_source": "FFMA micro-benchmark measured in this run (sirrt_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 figure",
            "kernel_ms": k_ms, "flop_per_check_executed": f_exec, "executed_source": (consts.get("summary") if exec_ok else stale),
            "issue": {"achieved": issue_rate, "peak": issue_peak, "unit": "warp-inst/s", "frac": (issue_rate / issue_peak) if issue_rate else None,
                      "warp_inst_per_check": i_exec,
                      "note": "the cull pipeline is compare/ballot/queue work: warp-instruction issue slots are the binding resource, not the FMA pipe or HBM; "
                              "a smarter cull LOWERS executed FLOP/check, so checks/s and warp-inst/check are the figures to follow, not the FP32 fraction"},
            "bruteforce": {"flop_per_check": f_check, "equivalent_tflops": brute_tflops,
                           "note": "SURVEY §8d brute-force accounting (L*C*12 per check): what a kernel without culling would have to execute for the same answer"},
            "hbm": {"achieved": alg_bytes / (k_ms * 1e-3) / 1e9 if k_ms > 0 else None, "peak": hbm_peak, "unit": "GB/s", "peak_source": peak_src,
                    "frac": (alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak) if k_ms > 0 else None, "algorithmic_bytes": alg_bytes},
            "scope": "rank 0's block" if world > 1 else "whole batch",
        }
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong" if world > 1 else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(a, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(q_loc.nbytes) * world, "d2h_bytes_per_step": int(table_pin.numel() * 4) * world,
                    "ms_per_step": e2e_ms, "note": "whole-job bytes; every rank moves its own block" if world > 1 else None},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
            "timing_check": {"wall_ms_per_step_incl_l2_flush": acc["wall_ms_per_step"], "event_ms_per_step_rank0": acc["step_ms"],
                             "note": "host wall clock around the timed loop (max over ranks), next to the CUDA-event times the value is computed from"},
            "results": {"colliding_cell_fraction": density, "near_contact_pair_tests_1e-6": int(acc["near"]), "fp64_rechecked_pair_tests": int(acc["rechecked"]),
                        "exact_pair_tests": int(acc["exact"])},
        }
        out["results"]["e2e_int16_table"] = {"value": N * T / (e2e16_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e16_ms,
                                             "d2h_bytes_per_step": int(table16_pin.numel() * 2) * world}
        if multi_check:
            out["results"]["multi_gpu_check"] = multi_check
        if weak:
            out["results"]["weak_scaling"] = weak
        if not a.no_plan_time and world == 1:
            out["results"]["plan_time"] = plan_time_leg(a.plan_seeds, 5.0, local)
        if not a.no_plan_time and world == 1 and local == 0:
            out["results"]["latency"] = latency_leg()
        if not a.no_knn and world == 1:
            out["results"]["knn"] = knn_leg(pkg, local, hbm_peak, roofline["peak"])
        if a.mass_tasks > 0 and world == 1:
            out["results"]["mass_test"] = mass_test_leg(a.mass_tasks, 1)
        if not a.no_cpu_baseline:
            threads = os.cpu_count() or 1
            v, n, dt = cpu_baseline(pack, q, a.cpu_budget, threads)
            out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": threads, "kind": "port",
                                   "sample": f"{n} of {N} configurations x {T} frames ({dt:.1f} s)"}
        print(json.dumps(out))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
