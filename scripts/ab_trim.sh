#!/bin/bash
# Slab hand-over slack sweep: scripts/ab_trim.sh <ngpus> <dtype> trim1 trim2 ...   (PCS_SLAB_TRIM rows)
cd "$(dirname "$0")/.."
n=$1; dt=$2; shift 2
for trim in "$@"; do
  echo -n "n=$n $dt trim=$trim: "
  PCS_SLAB_TRIM=$trim timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 \
    --master-port $((29500 + RANDOM % 500)) bench.py --gpus $n --steps 5 --warmup 2 --dtype $dt --no-secondary 2>/dev/null \
    | python -c "import sys,json; d=json.loads(sys.stdin.readlines()[-1]); p=d.get('parity_vs_single_gpu',{}); print('us/iter', round(d['config']['ms_per_iteration']*1e3,2), ' value', round(d['value'],1), ' e2e', round(d['e2e']['value'],1), ' identical f64/f32', p.get('f64',{}).get('identical'), p.get('f32',{}).get('identical'), p.get('f32',{}).get('x_maxabs'))"
done
