import sys, os, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g
import conftest
import importlib.util
spec = importlib.util.spec_from_file_location("oracle2", os.path.join(ROOT, "oracle", "oracle2.py")); o2m = importlib.util.module_from_spec(spec); sys.modules["oracle2"] = o2m; spec.loader.exec_module(o2m)
sp, orc = g.load_tools(), g.load_oracle()
packs = conftest.make_packs(sp)
name = sys.argv[1]; n = int(sys.argv[2]); mode = int(sys.argv[3]) if len(sys.argv) > 3 else 0
pk = packs(name)
O1 = orc.Oracle(pk, mode); O2 = o2m.Oracle2(pk, mode)
q = conftest.random_q(sp, pk, n, 5).astype(np.float64)
rng = np.random.default_rng(1)
t = time.perf_counter(); nd = 0; nh = 0
for i in range(n):
    R1, t1 = O1.fk(q[:, i]); R2, t2 = O2.fk(q[:, i])
    assert np.abs(R1 - R2).max() < 1e-12 and np.abs(t1 - t2).max() < 1e-12, (np.abs(R1-R2).max(), np.abs(t1-t2).max())
    f = int(rng.integers(0, O1.n_frames))
    h1 = O1.check_frame(q[:, i], f)
    h2, m = O2.check_cell(q[:, i], f, o2m.SEM_FRAME)
    nh += h1
    if h1 != h2: nd += 1; print("DIFF", i, f, h1, h2, m)
print(name, "cells", n, "hits", nh, "diff", nd, "time", time.perf_counter() - t)
