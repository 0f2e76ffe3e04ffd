#!/bin/bash
# One box visit for the latency kernels: gpurun --timeout 900 -- 'bash scripts/gpu_lat2.sh TAG'
#   latency-path GPU tests, host/sirrt_latency on the shipped dynamic scenes, the GJK iteration histogram (diagnostics build,
#   scripts/micro/build_gjk_hist.sh) and an ncu launch list of the latency program on the flying-boxes scene
TAG=${1:-lat2}
OUT=gpurun_out
mkdir -p $OUT /tmp/packs
python - <<PY
import gzip
for n in ("mass_test", "flying_boxes", "multiple_robots"):
    open(f"/tmp/packs/{n}.pack", "wb").write(gzip.open(f"tests/golden/{n}.pack.gz").read())
PY
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_planner.py -m gpu -q -x -k "slot_split or deferred or windowed or single_call or check_frames or edge_check or planner or collision_manager or edge_cases or concurrent" > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -4 $OUT/${TAG}_tests.log
: > $OUT/${TAG}_latency.json
for n in mass_test flying_boxes multiple_robots; do
  timeout 120 manipulationplanning-si-rrt_b200/host/sirrt_latency /tmp/packs/$n.pack 500 >> $OUT/${TAG}_latency.json 2>> $OUT/${TAG}_latency.err
done
cat $OUT/${TAG}_latency.json
if [ -x scripts/micro/_build/gjk_hist ]; then
  : > $OUT/${TAG}_gjk_hist.json
  for n in mass_test flying_boxes multiple_robots; do
    timeout 120 scripts/micro/_build/gjk_hist /tmp/packs/$n.pack 64 >> $OUT/${TAG}_gjk_hist.json 2>> $OUT/${TAG}_latency.err
  done
  python - <<PY
import json
for line in open("$OUT/${TAG}_gjk_hist.json"):
    d = json.loads(line)
    print(d["call"], d["calls"], {k: (v["tasks"], v["mean_iter"], v["max_iter"]) for k, v in d.items() if isinstance(v, dict) and "mean_iter" in v},
          d.get("collide_bits_cycles_per_warp"), d.get("gjk_task_cycles"))
PY
fi
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file $OUT/${TAG}_launches_fb.csv manipulationplanning-si-rrt_b200/host/sirrt_latency /tmp/packs/flying_boxes.pack 40 > $OUT/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("$OUT/${TAG}_launches_fb.csv")) if len(r) > 10 and r[0].isdigit()]
agg = collections.defaultdict(list)
for r in rows: agg[(r[4].split("(")[0], r[8])].append(float(r[-1]))
for k, v in agg.items(): print(k, len(v), "mean ns %.0f" % (sum(v) / len(v)), "median %.0f" % sorted(v)[len(v) // 2])
PY
