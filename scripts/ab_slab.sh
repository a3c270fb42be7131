#!/bin/bash
# Multi-GPU A/B of library builds on the weak-scaling slab bench: scripts/ab_slab.sh <ngpus> <dtype> <reps> lib...
cd "$(dirname "$0")/.."
n=$1; dt=$2; reps=$3; shift 3
for rep in $(seq 1 $reps); do
  for lib in "$@"; do
    if [ "$lib" = default ]; then p=""; else p=$PWD/$lib; fi
    echo -n "rep $rep n=$n $dt ${lib}: "
    PCS_LIB_PATH=$p timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
      --master-port $((29500 + RANDOM % 500)) bench.py --gpus $n --steps 5 --warmup 2 --dtype $dt --no-secondary 2>/dev/null \
      | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); print('ms/iter', round(d['config']['ms_per_iteration']*1e3,2), 'us  value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'parity', d.get('parity_vs_single_gpu',{}).get('f64'), d.get('parity_vs_single_gpu',{}).get('f32'))"
  done
done
