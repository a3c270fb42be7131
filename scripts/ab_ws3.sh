#!/bin/bash
# A/B of the three-role fp32 kernel (PCS_PDS_WS3=1) against the two-role one, same library: scripts/ab_ws3.sh <reps>
cd "$(dirname "$0")/.."
for rep in $(seq 1 ${1:-3}); do
  for m in 0 1; do
    echo -n "rep $rep PCS_PDS_WS3=$m: "
    PCS_PDS_WS3=$m timeout 300 python scripts/probe_kernel.py f32 2>&1 | grep "us/iter" | tr '\n' ' '
    echo
  done
done
