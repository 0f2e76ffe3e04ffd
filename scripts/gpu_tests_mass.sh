#!/bin/bash
TAG=${1:-tm}; TASKS=${2:-384}
OUT=gpurun_out
mkdir -p $OUT
python -c "import __graft_entry__ as g; g.load_build().build_host()" > $OUT/${TAG}_build.log 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_planner.py -m gpu -q -x -k "slot_split or deferred or single_call or check_frames or reference_scenes or robot_obstacles or motion_check or hull_support or edge_check or windowed or concurrent or collision_manager or edge_cases or planner" > $OUT/${TAG}_tests.log 2>&1; echo "tests rc=$?"; tail -3 $OUT/${TAG}_tests.log
bash scripts/gpu_mass1.sh $TAG $TASKS
