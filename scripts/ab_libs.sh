#!/bin/bash
# A/B of library builds: scripts/ab_libs.sh <f32|f64|both> <reps> lib1.so lib2.so ...   ("default" = the in-tree build)
# Prints the device time per update of the raw plan.step loop (scripts/probe_kernel.py) for each build, alternating.
cd "$(dirname "$0")/.."
dt=$1; reps=$2; shift 2
for rep in $(seq 1 $reps); do
  for lib in "$@"; do
    if [ "$lib" = default ]; then p=""; else p=$PWD/$lib; fi
    echo -n "rep $rep ${lib}: "
    PCS_LIB_PATH=$p timeout 300 python scripts/probe_kernel.py $dt 2>&1 | grep "us/iter" | tr '\n' ' '
    echo
  done
done
