#!/bin/bash
# A/B of latency-path knobs under load: the 1-GPU mass test (first TASKS tasks, 48 and 16 threads, yield), each configuration twice
#   gpurun --timeout 1500 -- 'bash scripts/gpu_mass_ab.sh TAG TASKS "ENV1" "ENV2" ...'   (each ENV a string of VAR=value pairs)
TAG=${1:-mab}; TASKS=${2:-384}; shift 2
OUT=gpurun_out
mkdir -p $OUT
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_planner.py -m gpu -q -x -k "slot_split or deferred or single_call or check_frames or reference_scenes or hull_support or edge_check or windowed or concurrent or collision_manager or edge_cases or planner" > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 $OUT/${TAG}_tests.log
mkdir -p /tmp/packs
python - <<PY
import gzip
for n in ("mass_test", "flying_boxes", "multiple_robots"):
    open(f"/tmp/packs/{n}.pack", "wb").write(gzip.open(f"tests/golden/{n}.pack.gz").read())
PY
: > $OUT/${TAG}_latency.json
for n in mass_test flying_boxes multiple_robots; do
  timeout 120 manipulationplanning-si-rrt_b200/host/sirrt_latency /tmp/packs/$n.pack 500 >> $OUT/${TAG}_latency.json 2>> $OUT/${TAG}_err.log
done
cut -c1-260 $OUT/${TAG}_latency.json
for rep in 1 2; do
  i=0
  for cfg in "$@"; do
    i=$((i+1))
    env $cfg timeout 600 python scripts/mass_test.py --gpus 1 --tasks $TASKS --threads 48 --sync yield --cpu-sample 0 --also 16:yield > $OUT/${TAG}_cfg${i}_rep${rep}.json 2> $OUT/${TAG}_err.log
    python - <<PY
import json
d = json.load(open("$OUT/${TAG}_cfg${i}_rep${rep}.json"))
print("rep $rep [$cfg]", {k: round(d["gpu_arm"][k], 1) for k in ("threads_per_gpu", "tasks_per_s", "solved", "p50_solve_ms")}, [{k: round(v[k], 1) for k in ("threads_per_gpu", "tasks_per_s", "solved", "p50_solve_ms", "solved_by_both_with_identical_path", "solved_by_both")} for v in d["gpu_arm_variants"]])
PY
  done
done
