#!/bin/bash
cd "$(dirname "$0")/.."
for rep in 1 2; do for lib in "$@"; do
  if [ "$lib" = default ]; then p=""; else p=$PWD/$lib; fi
  echo -n "rep $rep $lib: "
  PCS_LIB_PATH=$p python bench.py --workload c4 --c4-problems 2 --steps 3 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'it/s', round(d['config']['ms_per_problem_iteration']*1e3,1), 'us per problem-iteration, checksum', d['config']['checksum_abs_x'])"
done; done
