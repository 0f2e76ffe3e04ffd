import os, sys, json, subprocess, tempfile, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import __graft_entry__ as g
g.load_build().build_lib(); g.load_build().build_host()
sp, pkg = g.load_tools(), g.load_package()
import importlib.util
spec = importlib.util.spec_from_file_location("mass_test", os.path.join(ROOT, "scripts", "mass_test.py")); mt = importlib.util.module_from_spec(spec); spec.loader.exec_module(mt)
tmp = tempfile.mkdtemp()
tasks = mt.make_tasks(sp, pkg, 1024, 8, 8, tmp)
idx = int(sys.argv[1]) if len(sys.argv) > 1 else 476
pack, seed, k = tasks[idx]
print("task", idx, os.path.basename(pack), seed, k, flush=True)
import shutil; shutil.copy(pack, os.path.join(ROOT, "gpurun_out", "flaky.pack"))
exe = os.path.join(ROOT, "manipulationplanning-si-rrt_b200", "host", "sirrt_mass")
def run(n_lines, threads, sync, extra_env=None):
    lst = os.path.join(tmp, "one.txt")
    open(lst, "w").write("".join(f"{pack} {seed}\n" for _ in range(n_lines)))
    env = dict(os.environ, SIRRT_SYNC=sync, **(extra_env or {}))
    r = subprocess.run([exe, lst, "0", str(threads), "--max-time", "5"], capture_output=True, text=True, env=env)
    rows = [json.loads(l) for l in r.stdout.splitlines() if '"task"' in l]
    return collections.Counter((x["iterations"], x["path_hash"], x["solved"]) for x in rows)

import random
group = [t for t in tasks if t[2] in (120, 140, 160)]
lines = group * 3
random.Random(1).shuffle(lines)
lines.sort(key=lambda t: t[0])   # seeds of a pack stay together, as in the mass test list
def run_list(threads, sync):
    lst = os.path.join(tmp, "grp.txt")
    open(lst, "w").write("".join(f"{p} {sd}\n" for p, sd, _ in lines))
    env = dict(os.environ, SIRRT_SYNC=sync)
    r = subprocess.run([exe, lst, "0", str(threads), "--max-time", "5"], capture_output=True, text=True, env=env)
    rows = [json.loads(l) for l in r.stdout.splitlines() if '"task"' in l]
    by = collections.defaultdict(set)
    for x in rows:
        if x["solved"]:
            by[(os.path.basename(x["pack"]), x["seed"])].add((x["iterations"], x["path_hash"], x["safe_interval_calls"], x["state_checks"]))
    bad = {k: v for k, v in by.items() if len(v) > 1}
    return len(rows), len(by), bad
for th, sy in ((16, "yield"), (16, "spin"), (4, "spin")):
    print(th, sy, run_list(th, sy), flush=True)
