#!/usr/bin/env python
"""Per-call latency of the planner-in-the-loop entry points (single-config drop-ins, small batches, kNN) on the
B200 backend, next to the CPU oracle on the same inputs.  Run on the GPU box: python scripts/latency.py"""
import gzip, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g


def bench(fn, n=400, warm=20):
    for _ in range(warm):
        fn()
    t = time.perf_counter()
    for _ in range(n):
        fn()
    return (time.perf_counter() - t) / n * 1e6


def main():
    sp, orc, pkg = g.load_tools(), g.load_oracle(), g.load_package()
    orc.build()
    out = {}
    cyl = sp.read_pack(gzip.open(os.path.join(ROOT, "tests/golden/xarm6cyl_robot.pack.gz")).read()).robot
    scenes = {"mass_test_hull": gzip.open(os.path.join(ROOT, "tests/golden/mass_test.pack.gz")).read(),
              "dyn120_capsule": sp.write_pack(sp.synthetic_scene(120, 600, 30, 5, "A", robot=cyl))}
    for name, pk in scenes.items():
        dev, ref = pkg.Scene(pk), orc.Oracle(pk)
        rob = sp.read_pack(pk).robot
        q = sp.random_configs(rob, 64, 3).T.astype(np.float64).copy()
        i = [0]
        def nxt():
            i[0] = (i[0] + 1) % 64
            return q[i[0]]
        r = {}
        r["safe_intervals_us"] = {"b200": bench(lambda: dev.safe_intervals(nxt())), "cpu_oracle": bench(lambda: ref.safe_intervals(nxt()), 100, 5)}
        r["check_frame_us"] = {"b200": bench(lambda: dev.check_frame(nxt(), 100)), "cpu_oracle": bench(lambda: ref.check_frame(nxt(), 100), 200, 5)}
        qa = np.ascontiguousarray(q[:32]); fr = np.arange(32, dtype=np.int32) * 7
        r["check_frames_batch32_us"] = {"b200": bench(lambda: dev.check_frames_batch_f64(qa, fr))}
        r["collision_bits_f64_32rows_us"] = {"b200": bench(lambda: dev.collision_bits_f64(qa, 1), 200)}
        out[name] = r
    k = pkg.KnnIndex(6, 1 << 20)
    pts = np.random.default_rng(1).uniform(-3, 3, (5000, 6))
    k.add(pts)
    qq = pts[:64] + 0.01
    j = [0]
    def nq():
        j[0] = (j[0] + 1) % 64
        return qq[j[0]]
    out["knn_5000pts"] = {"nearest_us": bench(lambda: k.nearest(nq()[None, :])), "radius_us": bench(lambda: k.radius(nq(), 4.0)),
                          "add_us": bench(lambda: k.add(nq()[None, :]))}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
