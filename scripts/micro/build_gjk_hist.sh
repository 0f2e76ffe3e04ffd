#!/bin/bash
# diagnostics build: the library with -DSIRRT_GJK_STATS and scripts/micro/gjk_hist against it (outputs under scripts/micro/_build/)
set -e
cd "$(dirname "$0")/../.."
P=manipulationplanning-si-rrt_b200
mkdir -p scripts/micro/_build
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC,-pthread -shared --cudart=static -DSIRRT_GJK_STATS \
  -o scripts/micro/_build/libsirrt_b200.so $P/csrc/sirrt_api.cu 2>/dev/null
g++ -O2 -std=c++17 -pthread -I include -I $P/host -o scripts/micro/_build/gjk_hist scripts/micro/gjk_hist.cpp -L scripts/micro/_build -lsirrt_b200 '-Wl,-rpath,$ORIGIN'
