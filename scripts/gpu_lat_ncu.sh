#!/bin/bash
# One box visit: latency numbers + ncu --set full captures of the three kernels of one get_safe_intervals call (flying boxes)
#   gpurun --timeout 900 -- 'bash scripts/gpu_lat_ncu.sh TAG [scene]'
TAG=${1:-latn}; SC=${2:-flying_boxes}
OUT=gpurun_out
mkdir -p $OUT /tmp/packs
python - <<PY
import gzip
for n in ("mass_test", "flying_boxes", "multiple_robots"):
    open(f"/tmp/packs/{n}.pack", "wb").write(gzip.open(f"tests/golden/{n}.pack.gz").read())
PY
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
LAT=manipulationplanning-si-rrt_b200/host/sirrt_latency
: > $OUT/${TAG}_latency.json
for n in mass_test flying_boxes multiple_robots; do
  timeout 120 $LAT /tmp/packs/$n.pack 500 >> $OUT/${TAG}_latency.json 2>> $OUT/${TAG}_latency.err
done
cat $OUT/${TAG}_latency.json
echo "tests skipped (full suite ran in the same round)"
for k in k_collide_bits k_gjk_tasks k_recheck; do
  timeout 300 ncu --set full --import-source on --clock-control none --cache-control none -k regex:$k --launch-skip 30 --launch-count 1 -f -o $OUT/${TAG}_$k $LAT /tmp/packs/$SC.pack 60 > $OUT/${TAG}_ncu_$k.log 2>&1; echo "ncu $k rc=$?"
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $LAT /tmp/packs/$SC.pack 40 > $OUT/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
python - <<PY
import csv, collections
rows = [r for r in csv.reader(open("$OUT/${TAG}_launches.csv")) if len(r) > 10 and r[0].isdigit()]
agg = collections.defaultdict(list)
for r in rows: agg[(r[4].split("(")[0], r[8])].append(float(r[-1]))
for k, v in agg.items(): print(k, len(v), "mean ns %.0f" % (sum(v) / len(v)), "median %.0f" % sorted(v)[len(v) // 2])
PY
