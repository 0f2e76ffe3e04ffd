import sys, os, numpy as np
sys.path.insert(0, "/root/repo")
import __graft_entry__ as g
sp, pkg = g.load_tools(), g.load_package()
pk = open("/tmp/mp/k100.pack","rb").read()
dev = pkg.Scene(pk)
sc = sp.read_pack(pk)
q = sp.random_configs(sc.robot, 60, 7).T.astype(np.float64)
for i in range(30):
    iv = dev.safe_intervals(q[i])
    print(len(iv), dev.last_stats())
