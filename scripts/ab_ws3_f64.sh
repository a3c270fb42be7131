#!/bin/bash
# fp64: three-role kernel, one block per SM (PCS_PDS_WS3_F64=1) vs the single-role kernel, same library
cd "$(dirname "$0")/.."
for rep in $(seq 1 ${1:-2}); do
  for m in 0 1; do
    echo -n "rep $rep PCS_PDS_WS3_F64=$m: "
    PCS_PDS_WS3_F64=$m timeout 300 python scripts/probe_kernel.py f64 2>&1 | grep -E "us/iter|rror" | tr '\n' ' '
    echo
  done
done
