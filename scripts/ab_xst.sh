#!/bin/bash
cd "$(dirname "$0")/.."
for rep in 1 2; do
  for lib in default variants/lib_xst3.so; do
    if [ "$lib" = default ]; then p=""; else p=$PWD/$lib; fi
    echo -n "rep $rep $lib f32: "; PCS_LIB_PATH=$p timeout 300 python scripts/probe_kernel.py f32 2>&1 | grep "us/iter" | tr '\n' ' '; echo
    for m in 1 2; do echo -n "rep $rep $lib f64 WS3_F64=$m: "; PCS_PDS_WS3_F64=$m PCS_LIB_PATH=$p timeout 300 python scripts/probe_kernel.py f64 2>&1 | grep "us/iter" | tr '\n' ' '; echo; done
  done
done
