#!/bin/bash
# One GPU-box visit: GPU tests, a bench line, the ncu launch list and one full capture of k_sweep.
#   gpurun --timeout 900 -- 'bash scripts/gpu_round.sh TAG [pytest-args]'
TAG=${1:-x}; shift
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q "$@" > $OUT/${TAG}_gpu_tests.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_gpu_tests.log
tail -5 $OUT/${TAG}_gpu_tests.log
timeout 300 python bench.py --steps 5 --warmup 3 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_bench.json").read().strip().splitlines()[-1])
    print({k: d[k] for k in ("value", "ms_per_step")}, d["e2e"], d["roofline"].get("kernel_ms"), d["results"])
except Exception as e:
    print("bench parse failed", e); print(open("$OUT/${TAG}_bench.err").read()[-2000:])
PY
python bench.py --print-source-hash > $OUT/${TAG}_src_hashes.json
SMALL="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-plan-time --no-knn --mass-tasks 0 --configs 131072"
$SMALL > $OUT/${TAG}_small_plain.log 2>&1; echo "small plain rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/${TAG}_launches.csv $SMALL > $OUT/${TAG}_ncu_l.log 2>&1; echo "ncu launches rc=$?"
timeout 600 ncu --set full --metrics smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_dfma_pred_on.sum --clock-control none --import-source on -k regex:k_sweep -c 1 -o $OUT/${TAG}_ksweep -f $SMALL > $OUT/${TAG}_ncu_f.log 2>&1; echo "ncu full rc=$?"
ls -la $OUT | grep ${TAG}
