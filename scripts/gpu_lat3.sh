#!/bin/bash
# Short latency visit: gpurun --timeout 600 -- 'bash scripts/gpu_lat3.sh TAG'   (key GPU parity tests, sirrt_latency, phase diagnostics)
TAG=${1:-lat3}
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
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "slot_split or deferred or single_call or check_frames or reference_scenes or robot_obstacles or get_distances or motion_check or hull_support or edge_check or windowed or concurrent or collision_manager or edge_cases" > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -2 $OUT/${TAG}_tests.log
bash scripts/gpu_diag.sh $TAG
