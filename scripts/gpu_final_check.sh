#!/bin/bash
# short check visit: latency-path GPU tests (incl. the warp FK test), C-ABI latencies, plan time on the shipped scenes
TAG=${1:-fc}; SEEDS=${2:-5}
OUT=gpurun_out
mkdir -p $OUT /tmp/packs
python - <<PY
import gzip
for n in ("mass_test", "flying_boxes", "multiple_robots"):
    open(f"/tmp/packs/{n}.pack", "wb").write(gzip.open(f"tests/golden/{n}.pack.gz").read())
PY
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_planner.py -m gpu -q -x -k "warp_fk or slot_split or deferred or single_call or check_frames or reference_scenes or knn or collision_manager or edge_cases or planner or concurrent or validator" > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 $OUT/${TAG}_tests.log
: > $OUT/${TAG}_latency.json
for n in mass_test flying_boxes multiple_robots; do
  timeout 120 manipulationplanning-si-rrt_b200/host/sirrt_latency /tmp/packs/$n.pack 500 >> $OUT/${TAG}_latency.json 2>> $OUT/${TAG}_latency.err
done
cut -c1-330 $OUT/${TAG}_latency.json
timeout 600 python scripts/plan_time.py --seeds $SEEDS --max-time 20 --scenes scene_task,mass_test,flying_boxes,multiple_robots,dyn40,dyn120 > $OUT/${TAG}_plan_time.json 2> $OUT/${TAG}_plan_time.err; echo "plan_time rc=$?"
python - <<PY
import json
try:
    d = json.load(open("$OUT/${TAG}_plan_time.json"))
    for k, v in d.get("scenes", d).items():
        print(k, {a: b for a, b in v.items() if "p50" in a or "identical" in a or "solved" in a} if isinstance(v, dict) else v)
except Exception as e:
    print("plan_time parse failed", e)
PY
