#!/bin/bash
# 1-GPU mass test (first TASKS tasks) at 48 / 16 threads (yield) and 48 threads (nap) + C-ABI latencies
TAG=${1:-m1}; TASKS=${2:-384}
OUT=gpurun_out
mkdir -p $OUT /tmp/packs
python - <<PY
import gzip
for n in ("mass_test", "flying_boxes", "multiple_robots"):
    open(f"/tmp/packs/{n}.pack", "wb").write(gzip.open(f"tests/golden/{n}.pack.gz").read())
PY
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
: > $OUT/${TAG}_latency.json
for n in mass_test flying_boxes multiple_robots; do
  timeout 120 manipulationplanning-si-rrt_b200/host/sirrt_latency /tmp/packs/$n.pack 500 >> $OUT/${TAG}_latency.json 2>> $OUT/${TAG}_latency.err
done
cat $OUT/${TAG}_latency.json
timeout 600 python scripts/mass_test.py --gpus 1 --tasks $TASKS --threads 48 --sync yield --cpu-sample 0 --also 16:yield,48:nap,32:yield > $OUT/${TAG}_mass_test.json 2> $OUT/${TAG}_err.log
python - <<PY
import json
d = json.load(open("$OUT/${TAG}_mass_test.json"))
print({k: d["gpu_arm"][k] for k in ("threads_per_gpu", "sync", "tasks_per_s", "solved", "p50_solve_ms")})
for v in d["gpu_arm_variants"]: print({k: v[k] for k in ("threads_per_gpu", "sync", "tasks_per_s", "solved", "p50_solve_ms", "solved_by_both_with_identical_path", "solved_by_both")})
PY
