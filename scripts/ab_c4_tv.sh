#!/bin/bash
# C4: row-marching TV kernel (default) vs the tile kernel (PCS_CPS_TV_TILE=1)
cd "$(dirname "$0")/.."
for rep in 1 2; do for tp in 0 1; do
  echo -n "rep $rep PCS_CPS_TV_TILE=$tp: "
  PCS_CPS_TV_TILE=$tp python bench.py --workload c4 --c4-problems 2 --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'it/s', round(d['config']['ms_per_problem_iteration']*1e3,1), 'us per problem-iteration, checksum', d['config']['checksum_abs_x'])"
done; done
