#!/bin/bash
# Lean 8-GPU box visit (the box time is charged N-fold): multi-GPU tests, the strong-scaling bench at N only, the mass test
#   gpurun --gpus 8 --timeout 900 -- 'bash scripts/gpu_multi_lean.sh 8 TAG'
N=${1:-8}; TAG=${2:-m8}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi -L > $OUT/${TAG}_gpus.txt; nproc >> $OUT/${TAG}_gpus.txt
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -q > $OUT/${TAG}_multi_tests.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_multi_tests.log
tail -3 $OUT/${TAG}_multi_tests.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --steps 5 --warmup 3 --cpu-budget 4 > $OUT/${TAG}_bench_${N}gpu.json 2> $OUT/${TAG}_bench_${N}gpu.err
echo "bench $N rc=$?"
python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_bench_${N}gpu.json").read().strip().splitlines()[-1])
    print($N, {k: d[k] for k in ("value", "ms_per_step", "scaling")}, d["e2e"]["value"], {k: v for k, v in d["results"].items() if k in ("multi_gpu_check", "weak_scaling")})
except Exception as e:
    print("bench parse failed", e); print(open("$OUT/${TAG}_bench_${N}gpu.err").read()[-1500:])
PY
timeout 600 python scripts/mass_test.py --gpus $N --tasks 1024 --threads 32 --sync yield --cpu-sample 32 --also 16:yield,48:yield > $OUT/${TAG}_mass_test.json 2> $OUT/${TAG}_mass_test.err; echo "mass rc=$?"
python - <<PY
import json
try:
    d = json.load(open("$OUT/${TAG}_mass_test.json"))
    print(d["gpu_arm"]["threads_per_gpu"], d["gpu_arm"]["tasks_per_s"], d["gpu_arm"]["solved"], d["gpu_arm"]["p50_solve_ms"], [(v["threads_per_gpu"], v["tasks_per_s"], v["p50_solve_ms"], v["solved_by_both_with_identical_path"], v["solved_by_both"]) for v in d["gpu_arm_variants"]], d["cpu_arm"]["tasks_per_s"], d["plan_equality_on_sample"])
except Exception as e:
    print("mass parse failed", e); print(open("$OUT/${TAG}_mass_test.err").read()[-1500:])
PY
