"""GPU-box diagnostic: list (row, frame) cells where the CUDA bits differ from the oracle and the
oracle's deciding pair distances there.  usage: python scripts/diag_parity.py <pack-name> <n> <mode>"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import __graft_entry__ as g
import conftest

name, n, mode = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
seed = int(sys.argv[4]) if len(sys.argv) > 4 else 19
sp, orc, pkg = g.load_tools(), g.load_oracle(), g.load_package()
packs = conftest.make_packs(sp)
pk = packs(name)
q = sp.random_configs(sp.read_pack(pk).robot, n, seed)
dev, ref = pkg.Scene(pk, 0, mode), orc.Oracle(pk, mode)
dev.set_option(pkg.OPT_SWEEP_MIN_CONFIGS, 1)
for sweep in (1, 0):
    dev.set_option(pkg.OPT_USE_SWEEP, sweep)
    t, b, near = dev.safe_intervals_batch(q, 8)
    rt, rb = ref.safe_intervals_batch(q, 8)
    T = ref.n_frames
    diff = b ^ rb
    rows = np.nonzero(diff.any(1))[0]
    print(f"sweep={sweep}: {len(rows)} rows differ; stats {dev.last_stats()}")
    n_ids = ref.n_obstacles + ref.n_robot_obstacle_links
    for r in rows[:6]:
        d = np.unpackbits(diff[r].view(np.uint8), bitorder="little")[:T]
        gbits = np.unpackbits(b[r].view(np.uint8), bitorder="little")[:T]
        fr = np.nonzero(d)[0]
        print(f" row {r}: {len(fr)} frames differ, first {fr[:8]}, gpu says {gbits[fr[:8]]}")
        qq = q[:, r].astype(np.float64)
        for f in fr[:2]:
            ds = [(ref.pair_distance(qq, l, i, int(f)), l, i) for l in range(ref.n_links) for i in range(n_ids)]
            ds.sort()
            print("   frame", f, "closest true-geometry pairs (d, link, id):", [(round(x, 7), l, i) for x, l, i in ds[:4]])
