#!/bin/bash
# One box visit for the latency path: gpurun --timeout 1200 -- 'bash scripts/gpu_latency.sh TAG [seeds]'
#   host/sirrt_latency on the three shipped dynamic scenes (C ABI call latencies), the planner tests, plan time
TAG=${1:-lat}; SEEDS=${2:-10}
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
timeout 600 python -m pytest tests/test_planner.py tests/test_gpu_parity.py -m gpu -q -k "planner or knn or latency or edge_cases or collision_manager" > $OUT/${TAG}_tests.log 2>&1; tail -3 $OUT/${TAG}_tests.log
timeout 900 python scripts/plan_time.py --seeds $SEEDS --max-time 20 > $OUT/${TAG}_plan_time.json 2> $OUT/${TAG}_plan_time.err; echo "plan_time rc=$?"
python - <<PY
import json
try:
    d = json.load(open("$OUT/${TAG}_plan_time.json"))
    for k, v in d.get("scenes", d).items():
        print(k, {a: b for a, b in v.items() if "p50" in a or "identical" in a or "solved" in a} if isinstance(v, dict) else v)
except Exception as e:
    print("plan_time parse failed", e)
PY
