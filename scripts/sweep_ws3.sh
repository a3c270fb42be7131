#!/bin/bash
# long/short ratio sweep for the three-role kernel
cd "$(dirname "$0")/.."
for r in "$@"; do
  echo -n "ws3 ratio $r: "
  PCS_PDS_WS3=1 PCS_PDS_SPLIT_RATIO=$r timeout 300 python scripts/probe_kernel.py f32 2>&1 | grep "us/iter" | tr '\n' ' '
  echo
done
