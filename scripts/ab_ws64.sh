#!/bin/bash
# A/B of fp64 warp-specialised builds (variants/lib_ws64_*.so) against the default single-role fp64 kernel
cd "$(dirname "$0")/.."
for lib in variants/lib_ws64_*.so; do
  echo "=== parity: $lib ==="
  PCS_LIB_PATH=$PWD/$lib timeout 300 python -m pytest tests/test_gpu_fused.py -m gpu -q -x 2>&1 | tail -3
done
for rep in 1 2 3; do
  for lib in "" variants/lib_ws64_*.so; do
    echo "=== lib: ${lib:-default} ==="
    PCS_LIB_PATH=${lib:+$PWD/$lib} timeout 200 python scripts/probe_gap.py 2>&1 | grep "f64: raw plan"
  done
done
