#!/bin/bash
# sweep of the long/short segment ratio of the uneven one-wave split: scripts/sweep_split.sh <f32|f64> r1 r2 ...
cd "$(dirname "$0")/.."
dt=$1; shift
for r in "$@"; do
  echo -n "ratio $r: "
  PCS_PDS_SPLIT_RATIO=$r timeout 300 python scripts/probe_kernel.py $dt 2>&1 | grep "us/iter" | tr '\n' ' '
  echo
done
