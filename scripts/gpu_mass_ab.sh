#!/bin/bash
# A/B of the latency-path knobs under load: the 1-GPU mass test (first TASKS tasks, 48 threads, yield) with PDL / slot split on and off
TAG=${1:-mab}; TASKS=${2:-384}
OUT=gpurun_out
mkdir -p $OUT
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
for cfg in "1 1" "0 1" "1 0" "0 0"; do
  set -- $cfg
  SIRRT_OPT_PDL=$1 SIRRT_OPT_SLOT_SPLIT=$2 timeout 600 python scripts/mass_test.py --gpus 1 --tasks $TASKS --threads 48 --sync yield --cpu-sample 0 --also 16:yield > $OUT/${TAG}_pdl$1_split$2.json 2> $OUT/${TAG}_err.log
  python - <<PY
import json
d = json.load(open("$OUT/${TAG}_pdl$1_split$2.json"))
print("pdl=$1 split=$2", {k: d["gpu_arm"][k] for k in ("threads_per_gpu", "tasks_per_s", "solved", "p50_solve_ms")}, [{k: v[k] for k in ("threads_per_gpu", "tasks_per_s", "solved", "p50_solve_ms")} for v in d["gpu_arm_variants"]])
PY
done
