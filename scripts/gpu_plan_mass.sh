#!/bin/bash
# One box visit: SI-RRT plan time (scripts/plan_time.py) and the 1-GPU mass test on the first TASKS tasks
#   gpurun --timeout 1500 -- 'bash scripts/gpu_plan_mass.sh TAG [seeds] [tasks]'
TAG=${1:-pm}; SEEDS=${2:-5}; TASKS=${3:-384}
OUT=gpurun_out
mkdir -p $OUT
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
timeout 900 python scripts/plan_time.py --seeds $SEEDS --max-time 20 --scenes scene_task,mass_test,flying_boxes,multiple_robots,dyn40,dyn120 > $OUT/${TAG}_plan_time.json 2> $OUT/${TAG}_plan_time.err; echo "plan_time rc=$?"
python - <<PY
import json
try:
    d = json.load(open("$OUT/${TAG}_plan_time.json"))
    for k, v in d.get("scenes", d).items():
        print(k, {a: b for a, b in v.items() if "p50" in a or "identical" in a or "solved" in a} if isinstance(v, dict) else v)
except Exception as e:
    print("plan_time parse failed", e)
PY
timeout 900 python scripts/mass_test.py --gpus 1 --tasks $TASKS --threads 16 --sync yield --cpu-sample 0 --also 32:yield,48:yield > $OUT/${TAG}_mass_test.json 2> $OUT/${TAG}_mass_test.err; echo "mass rc=$?"
python - <<PY
import json
try:
    d = json.load(open("$OUT/${TAG}_mass_test.json"))
    print(json.dumps(d.get("gpu_arm"))[:500]); print(json.dumps(d.get("gpu_arm_variants"))[:1200])
except Exception as e:
    print("mass parse failed", e); print(open("$OUT/${TAG}_mass_test.err").read()[-1500:])
PY
