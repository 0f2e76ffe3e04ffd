#!/usr/bin/env python
"""bench.py -- config x timestep collision checks/s of the safe-interval hot path on B200.

Workload (BASELINE.json configs[2], the configuration the metric is quoted on): synthetic batched
safe-interval sweep, 7-DOF capsule arm, 2^20 configurations x 1000 frames x 64 moving boxes PER GPU
(weak scaling: every rank sweeps its own 2^20 configurations against a replica of the scene; there
is no data-path collective).  One "step" = one full pass: collision bit rows + FP64 re-check of the
near-contact band + fixed-width interval table (Kmax=8) for the whole batch.

  value     whole-job checks/s, inputs and outputs resident in HBM (CUDA events on the launch stream)
  e2e       same pass through the C ABI with pinned HOST buffers (H2D of q, D2H of the table inside)
  roofline  dominant kernel (the collision sweep) against the FP32 FMA peak measured in this run
            and, for context, against the measured HBM copy bandwidth (MEASURED_PEAKS.json)
  cpu_baseline  the oracle's AABB-tree restatement of the reference algorithm on the host cores

`--impl reference` times that CPU restatement alone (the reference itself cannot be built in this
image: coal, Eigen, nanoflann, urdfdom, orocos-kdl are absent and not vendored).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
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
# Executed work of k_sweep per check on SYN variant A, frozen from the ncu capture of the current kernel
# version (profiles/r04_ksweep_summary.txt: 131072 configs x 1000 frames):
#   F_EXEC  FP32 FLOP/check  = (fadd + fmul + 2*ffma thread instructions, predicated on) / checks
#   I_EXEC  warp instructions/check = smsp__inst_executed.sum / checks
F_EXEC, I_EXEC, EXEC_SOURCE = 135.7, 16.23, "profiles/r04_ksweep_summary.txt"
# DRAM traffic of one k_sweep launch (131072 configurations x 1000 frames), same ncu --set full capture:
# dram__bytes_read.sum 4.264 MB + dram__bytes_write.sum 0.165 MB (collision rows stay in the 126 MB L2)
TRAFFIC_PER_LAUNCH_BYTES, TRAFFIC_LAUNCH_CONFIGS = 4264192 + 165376, 131072
N_SM, SCHEDULERS_PER_SM = 148, 4


def shard_range(n: int, rank: int, world: int):
    """Contiguous, balanced shard [n0, n1) of n units for `rank` of `world` (no overlap, no gap)."""
    return n * rank // world, n * (rank + 1) // world


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        return float(d.get("hbm_gbs", 6650.0)), "measured (MEASURED_PEAKS.json)", float(d.get("sm_max_mhz", 1965.0))
    return 6650.0, "fallback (B200_PROFILING.md)", 1965.0


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""

    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons), "samples": len(sm)}


def build_workload(a, seed_off=0):
    sp = g.load_tools()
    scene = sp.synthetic_scene(a.obstacles, a.frames, 30, 42, a.variant)
    pack = sp.write_pack(scene)
    q = sp.random_configs(scene.robot, a.configs, 42 + seed_off)
    return sp, scene, pack, q


def cpu_baseline(pack, q, budget_s: float, threads: int):
    """The oracle's restatement of the reference ALGORITHM (FK + AABB tree over constant-pose
    windows + exact narrowphase, CollisionManager.cpp:107-216,485-547), all host threads, on a
    bounded sample of the same workload."""
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
        "ms_per_step": 1e3 * tot_t / max(1, a.steps), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": workload_config(a, 1),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample,
                         "note": "oracle/sirrt_oracle.c AABB-tree restatement of the reference algorithm, OpenMP over configurations; the reference binary is unbuildable here"},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def workload_config(a, world):
    return {"workload": f"synthetic batched safe-interval sweep: 7-DOF capsule arm (arm7cyl.urdf), {a.configs} configs x {a.frames} frames x {a.obstacles} moving "
                        f"{'spheres' if a.variant == 'C' else 'boxes'} (variant {a.variant}) per GPU, Kmax={a.kmax}",
            "configs_per_gpu": a.configs, "frames": a.frames, "obstacles": a.obstacles, "links": 7, "kmax": a.kmax,
            "sharding": f"independent config shards x{world}, scene replicated, no collective", "l2": "flushed (256 MiB write) between timed steps"}


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
    ap.add_argument("--generic", action="store_true", help="force the generic kernel instead of the sweep kernel")
    a = ap.parse_args()
    a.warmup = max(a.warmup, 0)

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
    sp, scene, pack, q = build_workload(a, rank)
    N, T, W, K = a.configs, a.frames, (a.frames + 31) // 32, a.kmax
    dev = pkg.Scene(pack, local)
    if a.generic:
        dev.set_option(pkg.OPT_USE_SWEEP, 0)
    stream = torch.cuda.current_stream()
    dev.set_stream(stream.cuda_stream)
    q_dev = torch.from_numpy(q).cuda()
    table_dev = torch.empty((N, 1 + 2 * K), dtype=torch.int32, device="cuda")
    bits_dev = torch.empty((N, W), dtype=torch.int32, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    fp32_peak = pkg.measure_fp32_peak(local)

    def step():
        return dev.safe_intervals_batch_raw(q_dev.data_ptr(), N, table_dev.data_ptr(), K, bits_dev.data_ptr())

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(a.warmup):
        step()
    sampler = ClockSampler(local)
    sync_all()
    sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(a.steps)]
    kern_ms, launches, near, rechecked, exact = 0.0, 0, 0, 0, 0
    for s0, s1 in ev:
        flush.fill_(1)  # evict L2 (126 MB) between timed steps; outside the event pair
        s0.record(stream)
        near = step()
        s1.record(stream)
        st = dev.last_stats()
        kern_ms += st["sweep_kernel_ms"]; launches += st["launches"]; rechecked = st["rechecked"]; exact = st["exact_tests"]
    sync_all()
    clocks = sampler.stop()
    total_ms = sum(s0.elapsed_time(s1) for s0, s1 in ev)
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_per_step = total_ms / max(1, a.steps)
    value = world * N * T / (ms_per_step * 1e-3)

    # ---- end to end through the C ABI with pinned host buffers
    q_pin = torch.from_numpy(q).pin_memory()
    table_pin = torch.empty((N, 1 + 2 * K), dtype=torch.int32).pin_memory()
    e2e_steps = max(1, min(a.steps, 5))
    dev.safe_intervals_batch_raw(q_pin.data_ptr(), N, table_pin.data_ptr(), K, None)
    sync_all()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        dev.safe_intervals_batch_raw(q_pin.data_ptr(), N, table_pin.data_ptr(), K, None)
    torch.cuda.synchronize()
    e2e_ms = 1e3 * (time.perf_counter() - t0) / e2e_steps
    if world > 1:
        t = torch.tensor([e2e_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_ms = float(t.item())
    assert np.array_equal(table_pin.numpy(), table_dev.cpu().numpy()), "e2e result differs from the device-resident result"
    e2e_value = world * N * T / (e2e_ms * 1e-3)

    if rank == 0:
        hbm_peak, peak_src, _ = measured_peaks()
        bits = bits_dev.cpu().numpy().view(np.uint32)
        density = float(np.unpackbits(bits.view(np.uint8)).sum()) / (N * T)
        k_ms = kern_ms / max(1, a.steps)
        f_exact = F_EXACT_CAPSULE_SPHERE if a.variant == "C" else F_EXACT_CAPSULE_BOX
        f_check = 7 * a.obstacles * F_CULL + F_FK / T + (exact / (N * T)) * f_exact
        brute_tflops = N * T * f_check / (k_ms * 1e-3) / 1e12 if k_ms > 0 else None
        exec_ok = (not a.generic) and a.variant == "A" and a.obstacles == 64 and a.frames == 1000
        exec_tflops = N * T * F_EXEC / (k_ms * 1e-3) / 1e12 if (k_ms > 0 and exec_ok) else None
        sm_hz = (clocks.get("sm_mhz") or 1965.0) * 1e6
        issue_peak = N_SM * SCHEDULERS_PER_SM * sm_hz  # warp instructions / s at the clock seen under load
        issue_rate = N * T * I_EXEC / (k_ms * 1e-3) if (k_ms > 0 and exec_ok) else None
        alg_bytes = N * 7 * 4 + a.obstacles * 12 * 4 * ((T + 31) // 32 * 32) + N * W * 4
        roofline = {
            "bound": "fp32", "kernel": "k_sweep" if not a.generic else "k_collide_bits", "achieved": exec_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
            "frac": (exec_tflops / fp32_peak) if exec_tflops and fp32_peak else None,
            "traffic": TRAFFIC_PER_LAUNCH_BYTES if exec_ok else None, "traffic_note": f"bytes per k_sweep launch of {TRAFFIC_LAUNCH_CONFIGS} configurations (ncu dram__bytes_read.sum + dram__bytes_write.sum, {EXEC_SOURCE})",
            "peak_source": "FFMA micro-benchmark measured in this run (sirrt_measure_fp32_peak); MEASURED_PEAKS.json has no FP32 figure",
            "kernel_ms": k_ms, "flop_per_check_executed": F_EXEC if exec_ok else None, "executed_source": EXEC_SOURCE,
            "issue": {"achieved": issue_rate, "peak": issue_peak, "unit": "warp-inst/s", "frac": (issue_rate / issue_peak) if issue_rate else None,
                      "warp_inst_per_check": I_EXEC if exec_ok else None,
                      "note": "the cull pipeline is compare/ballot/queue work: warp-instruction issue slots are the binding resource, not the FMA pipe or HBM"},
            "bruteforce": {"flop_per_check": f_check, "equivalent_tflops": brute_tflops,
                           "note": "SURVEY §8d brute-force accounting (L*C*12 per check): what a kernel without culling would have to execute for the same answer"},
            "hbm": {"achieved": alg_bytes / (k_ms * 1e-3) / 1e9 if k_ms > 0 else None, "peak": hbm_peak, "unit": "GB/s", "peak_source": peak_src,
                    "frac": (alg_bytes / (k_ms * 1e-3) / 1e9 / hbm_peak) if k_ms > 0 else None, "algorithmic_bytes": alg_bytes},
        }
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(a, world),
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(q.nbytes), "d2h_bytes_per_step": int(table_pin.numel() * 4), "ms_per_step": e2e_ms},
            "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
            "results": {"colliding_cell_fraction": density, "near_contact_pair_tests_1e-6": int(near), "fp64_rechecked_pair_tests": int(rechecked),
                        "exact_pair_tests": int(exact)},
        }
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
