#!/bin/bash
# Instrumented build of the library (in-kernel phase timers, -DNSB_TC_PROFILE) into the git-ignored build_prof/:
#   bash scripts/build_prof.sh && NSB_LIB=build_prof/libnsb200_prof.so NSB_TC_PROFILE_PRINT=1 python scripts/prof_step.py
set -e
cd "$(dirname "$0")/.."
mkdir -p build_prof
C=nice_slam_annotation_b200/csrc
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr --extended-lambda -DNSB_TC_PROFILE"
pids=()
for f in $C/*.cu; do
  nvcc $FLAGS -c $f -o build_prof/$(basename ${f%.cu}).o &
  pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
nvcc -shared -o build_prof/libnsb200_prof.so build_prof/*.o -lcudart
ls -la build_prof/libnsb200_prof.so
