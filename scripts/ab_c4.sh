#!/bin/bash
cd "$(dirname "$0")/.."
for tp in 0 1; do
  echo "=== PCS_CPS_BALL_TWO_PASS=$tp"
  PCS_CPS_BALL_TWO_PASS=$tp python bench.py --workload c4 --c4-problems 2 --steps 3 --warmup 2 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], 'it/s', d['config']['ms_per_problem_iteration']*1e3, 'us per problem-iteration', d['config']['checksum_abs_x'])"
done
