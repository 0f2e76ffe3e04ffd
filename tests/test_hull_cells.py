"""CPU-run check of the GJK support-candidate cells (csrc/sirrt_tables.hpp: build_hull_cells, csrc/sirrt_device.cuh:
hull_cell): tests/hull_cells_host_main.cu is compiled with nvcc and run on the HOST -- for every hull of the shipped
xArm6 scene and ~1.4 million directions (random over seven decades of magnitude, cell corners / edges with a hair of
jitter, cube edges, hull vertices, edge and cross-product directions) the candidate scan must pick the vertex the full
scan picks, in FP32 and in FP64."""
import json
import os
import shutil
import subprocess

import pytest

from conftest import ROOT


def test_support_cells_pick_the_full_scan_winner(packs, tmp_path):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = os.path.join(ROOT, "tests", "_build", "hull_cells_host")
    src = os.path.join(ROOT, "tests", "hull_cells_host_main.cu")
    csrc = os.path.join(ROOT, "manipulationplanning-si-rrt_b200", "csrc")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    deps = [src] + [os.path.join(csrc, f) for f in ("sirrt_device.cuh", "sirrt_tables.hpp", "sirrt_host.hpp")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-I", csrc, "-I", os.path.join(ROOT, "include"),
                               "-o", exe, src])
    (tmp_path / "s.pack").write_bytes(packs("flying_boxes"))
    r = subprocess.run([exe, str(tmp_path / "s.pack"), "100000"], capture_output=True, text=True, timeout=600)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert r.returncode == 0 and out["hulls"] >= 5 and out["directions"] > 700000, (out, r.stderr)
    assert out["mismatch_fp32"] == 0 and out["mismatch_fp64"] == 0, out
    assert out["mean_candidates_per_cell"] < 32, out   # the point of the cells: one lane-step per support query on average


def test_enclosure_certificate_is_sound(tmp_path):
    """tests/tetra_host_main.cu, run on the HOST: the FP64 enclosure certificate of the GJK task kernel
    (csrc/sirrt_device.cuh: tetra_encloses) against an independent long-double barycentric test on 200 000 random
    hull-vertex x box-corner quadruples: never certifies a quadruple whose tetrahedron does not contain the origin, a
    degenerate (repeated-point) quadruple or a speck without real volume; the reported depth equals the exact one."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = os.path.join(ROOT, "tests", "_build", "tetra_host")
    src = os.path.join(ROOT, "tests", "tetra_host_main.cu")
    csrc = os.path.join(ROOT, "manipulationplanning-si-rrt_b200", "csrc")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    deps = [src, os.path.join(csrc, "sirrt_device.cuh")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-I", csrc, "-I", os.path.join(ROOT, "include"),
                               "-o", exe, src])
    r = subprocess.run([exe, "200000"], capture_output=True, text=True, timeout=600)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert r.returncode == 0, (out, r.stderr)
    assert out["wrong_certificates"] == 0 and out["flat_certified"] == 0 and out["speck_certified"] == 0
    assert out["certified"] > 50000 and out["speck_inside"] > 10000   # the sample exercises both sides


def test_sweep_frame_windows_are_rigorous(tmp_path):
    """tests/sweep_window_host_main.cu, run on the HOST: the analytic frame windows of k_sweep's stage W
    (csrc/sirrt_sweep_window.cuh) on random capsule links against bouncing-box tiles (reflections, slow, static,
    axis-parallel and far-coordinate motion), checked against an FP64 point-segment distance of the true FP32 centres:
    no frame that can touch is missing from the IN window, no frame of the HIT window is farther than r_hit."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = os.path.join(ROOT, "tests", "_build", "sweep_window_host")
    src = os.path.join(ROOT, "tests", "sweep_window_host_main.cu")
    csrc = os.path.join(ROOT, "manipulationplanning-si-rrt_b200", "csrc")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    deps = [src, os.path.join(csrc, "sirrt_sweep_window.cuh")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-I", csrc, "-o", exe, src])
    r = subprocess.run([exe, "400000"], capture_output=True, text=True, timeout=600)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert r.returncode == 0, (out, r.stderr)
    assert out["missed_in"] == 0 and out["false_hit"] == 0, out
    assert out["true_in"] > 100000 and out["true_hit"] > 10000, out   # both windows are exercised


def test_oriented_box_cull_is_sound(tmp_path):
    """tests/obb_host_main.cu, run on the HOST: the FP32 15-axis oriented-box test in front of every hull pair test
    (csrc/sirrt_kernels.cuh: obb_separated, SIRRT_OPT_OBB_CULL) against an independent FP64 feature-based box-box distance on
    200 000 random pairs, two thirds of them pushed to within 1e-6 ... 1e-2 m of contact (some a hair inside): "separated by
    more than band + 1e-5 m" is never said of a pair that meets or lies within the band -- a wrong verdict there would turn a
    colliding pair FREE without a GJK run -- and every pair clear by a millimetre is culled."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    exe = os.path.join(ROOT, "tests", "_build", "obb_host")
    src = os.path.join(ROOT, "tests", "obb_host_main.cu")
    csrc = os.path.join(ROOT, "manipulationplanning-si-rrt_b200", "csrc")
    os.makedirs(os.path.dirname(exe), exist_ok=True)
    deps = [src, os.path.join(csrc, "sirrt_kernels.cuh"), os.path.join(csrc, "sirrt_device.cuh")]
    if not os.path.exists(exe) or os.path.getmtime(exe) < max(os.path.getmtime(d) for d in deps):
        subprocess.check_call([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-I", csrc, "-I", os.path.join(ROOT, "include"),
                               "-o", exe, src], stderr=subprocess.DEVNULL)
    r = subprocess.run([exe, "200000"], capture_output=True, text=True, timeout=600)
    out = json.loads(r.stdout.strip().splitlines()[-1])
    assert r.returncode == 0, (out, r.stderr)
    assert out["wrong_separated"] == 0 and out["closest_separated_m"] > 2e-5, out
    assert out["near_contact"] > 80000 and out["touching_or_inside"] > 10000, out        # the sample sits where it matters
    assert out["clear_by_1mm_separated"] == out["clear_by_1mm"] > 50000, out               # and the cull does its job
