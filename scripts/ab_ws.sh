#!/bin/bash
# A/B of two builds of the library on the 4096^2 deconvolution probe (fp32 = warp-specialised kernel)
cd "$(dirname "$0")/.."
for rep in 1 2; do
  for lib in "" variants/lib_noxcol.so; do
    echo "=== lib: ${lib:-default} ==="
    PCS_LIB_PATH=${lib:+$PWD/$lib} python scripts/probe_gap.py 2>&1 | grep "raw plan"
  done
done
