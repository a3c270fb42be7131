#!/bin/bash
# A/B of the extra-column warp (32 strips, no overlap) against the overlapping 33-strip geometry, 4096^2
cd "$(dirname "$0")/.."
for lib in "" variants/lib_noxcol.so; do
  echo "=== lib: ${lib:-default (xcol)} ==="
  for rep in 1 2; do
    PCS_LIB_PATH=${lib:+$PWD/$lib} python scripts/probe_gap.py 2>&1 | grep -v Warning
    PCS_LIB_PATH=${lib:+$PWD/$lib} python scripts/probe_denoise.py 2>&1 | grep -v Warning
  done
  for r in 1.1 1.2 1.3; do
    echo "--- uneven split on the single-role kernels, ratio $r"
    PCS_PDS_SPLIT_SR=1 PCS_PDS_SPLIT_RATIO=$r PCS_LIB_PATH=${lib:+$PWD/$lib} python scripts/probe_gap.py 2>&1 | grep f64
  done
done
