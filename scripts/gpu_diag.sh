#!/bin/bash
# diagnostics only: scripts/micro/_build/gjk_hist on the three shipped dynamic scenes -> gpurun_out/TAG_gjk_hist.json
TAG=${1:-diag}
OUT=gpurun_out
mkdir -p $OUT /tmp/packs
python - <<PY
import gzip
for n in ("mass_test", "flying_boxes", "multiple_robots"):
    open(f"/tmp/packs/{n}.pack", "wb").write(gzip.open(f"tests/golden/{n}.pack.gz").read())
PY
: > $OUT/${TAG}_gjk_hist.json
for n in mass_test flying_boxes; do
  timeout 120 scripts/micro/_build/gjk_hist /tmp/packs/$n.pack 64 >> $OUT/${TAG}_gjk_hist.json
done
python - <<PY
import json
for line in open("$OUT/${TAG}_gjk_hist.json"):
    d = json.loads(line)
    print(d["call"], {k: (v["tasks"], v["mean_iter"], v["max_iter"], v["hist"]) for k, v in d.items() if isinstance(v, dict) and "mean_iter" in v}, d.get("gjk_task_cycles"), d.get("max_cycles_over_all_calls"))
PY
