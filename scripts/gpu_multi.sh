#!/bin/bash
# One multi-GPU box visit: gpurun --gpus N --timeout 1500 -- 'bash scripts/gpu_multi.sh N TAG [mass]'
#   multi-GPU tests, the strong-scaling bench under torchrun (N and, for N > 2, also 2 and 4), optionally the mass test
N=${1:-2}; TAG=${2:-m}; MASS=${3:-}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi -L > $OUT/${TAG}_gpus.txt; nproc >> $OUT/${TAG}_gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -q > $OUT/${TAG}_multi_tests.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_multi_tests.log
tail -3 $OUT/${TAG}_multi_tests.log
for n in 1 2 4 8; do
  if [ $n -gt $N ]; then break; fi
  if [ $n -eq 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 --cpu-budget 4 --no-plan-time --no-knn --mass-tasks 0 > $OUT/${TAG}_bench_1gpu.json 2> $OUT/${TAG}_bench_1gpu.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n --steps 5 --warmup 3 --cpu-budget 4 > $OUT/${TAG}_bench_${n}gpu.json 2> $OUT/${TAG}_bench_${n}gpu.err
  fi
  echo "bench $n rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open("$OUT/${TAG}_bench_${n}gpu.json").read().strip().splitlines()[-1])
    print($n, {k: d[k] for k in ("value", "ms_per_step", "scaling")}, d["e2e"]["value"], {k: v for k, v in d["results"].items() if k in ("multi_gpu_check", "weak_scaling")})
except Exception as e:
    print("bench parse failed", e); print(open("$OUT/${TAG}_bench_${n}gpu.err").read()[-1500:])
PY
done
if [ -n "$MASS" ]; then
  timeout 1200 python scripts/mass_test.py --gpus $N --tasks 1024 --threads 16 --sync yield --cpu-sample 64 --also 32:yield,48:yield,32:block > $OUT/${TAG}_mass_test.json 2> $OUT/${TAG}_mass_test.err; echo "mass rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("$OUT/${TAG}_mass_test.json"))
    print(d["gpu_arm"]["tasks_per_s"], d["gpu_arm"]["solved"], [(v["threads_per_gpu"], v["tasks_per_s"], v["solved_by_both_with_identical_path"], v["solved_by_both"]) for v in d["gpu_arm_variants"]], d["cpu_arm"]["tasks_per_s"], d["plan_equality_on_sample"])
except Exception as e:
    print("mass parse failed", e); print(open("$OUT/${TAG}_mass_test.err").read()[-1500:])
PY
fi
ls -la $OUT | grep ${TAG}
