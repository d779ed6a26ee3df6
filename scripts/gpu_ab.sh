#!/usr/bin/env bash
# same-box A/B of the configs[2] step: the tree before this part of the round (ab_old/) against the current tree
set -u
mkdir -p gpurun_out
for rep in 1 2; do
for which in old new; do
  if [ $which = old ]; then d=ab_old; else d=.; fi
  (cd $d && LIFTER_BENCH_NO_CPU=1 timeout 200 python bench.py --steps 40 --warmup 10 2>/dev/null | python -c "
import sys, json
d = json.loads([l for l in sys.stdin if l.startswith('{')][-1])
print('$which rep $rep ms', round(d['ms_per_step'], 4), 'sustained', round(d['sustained']['ms_per_step'], 4), 'gemm us', round(d['roofline']['us_per_launch'], 2))")
done
done
