#!/bin/bash
# A/B of library builds on C2 (1024^2 TV denoising, 200 updates per solve): scripts/ab_c2.sh <reps> lib1.so lib2.so ...
cd "$(dirname "$0")/.."
reps=$1; shift
for rep in $(seq 1 $reps); do
  for lib in "$@"; do
    if [ "$lib" = default ]; then p=""; else p=$PWD/$lib; fi
    echo "rep $rep ${lib}:"
    PCS_LIB_PATH=$p timeout 300 python -c "
import bench
r = bench.c2_measure()
for t in ('f64', 'f32'):
    print('   ', t, round(r[t]['us_per_update'], 2), 'us/update', r[t]['kernel']['kernel'])
" 2>&1 | grep us/update
  done
done
