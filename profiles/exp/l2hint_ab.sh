#!/bin/bash
# L2 eviction priority of the staged source boxes (FIMEXCU_TMA_L2HINT: 0 none, 1 evict_last = library default, 2 evict_first), S->T1
for rep in 1 2; do
  for v in 0 1 2; do
    FIMEXCU_TMA_L2HINT=$v python bench.py --steps 40 --warmup 5 --no-e2e --no-job --no-cpu-baseline --no-secondary 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('l2hint=$v ms_per_step', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4))"
  done
done
