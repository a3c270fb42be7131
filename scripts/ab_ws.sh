#!/bin/bash
# A/B of two builds of the library on the 4096^2 deconvolution probe (fp32 = warp-specialised kernel)
cd "$(dirname "$0")/.."
VARIANT=${1:-variants/lib_onepoller.so}
for rep in 1 2 3; do
  for lib in "" $VARIANT; do
    echo "=== lib: ${lib:-default} ==="
    PCS_LIB_PATH=${lib:+$PWD/$lib} python scripts/probe_gap.py 2>&1 | grep "f32: raw plan"
  done
done
