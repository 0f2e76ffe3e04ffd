#!/bin/bash
# SM clock while the single-stream latency program runs (is a mostly idle GPU held at its maximum clock?)
TAG=${1:-clk}
OUT=gpurun_out
mkdir -p $OUT /tmp/packs
python - <<PY
import gzip
for n in ("flying_boxes",):
    open(f"/tmp/packs/{n}.pack", "wb").write(gzip.open(f"tests/golden/{n}.pack.gz").read())
PY
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks.mem,power.draw,pstate,clocks_throttle_reasons.active --format=csv -lms 100 > $OUT/${TAG}_clocks.csv &
SMI=$!
sleep 1
manipulationplanning-si-rrt_b200/host/sirrt_latency /tmp/packs/flying_boxes.pack 20000 > $OUT/${TAG}_latency.json
sleep 0.5
kill $SMI
cat $OUT/${TAG}_latency.json
python - <<PY
rows = [l.strip().split(", ") for l in open("$OUT/${TAG}_clocks.csv").read().strip().splitlines()[1:]]
clk = [int(r[0].split()[0]) for r in rows]
print("samples", len(clk), "sm MHz min/median/max", min(clk), sorted(clk)[len(clk)//2], max(clk), "first", clk[:5], "pstate", set(r[4] for r in rows), "power", rows[len(rows)//2][3])
PY
