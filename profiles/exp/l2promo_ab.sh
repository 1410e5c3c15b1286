#!/bin/bash
# L2 promotion of the source tensor map of the staged kernels: none / 64 B / 128 B (shipped) / 256 B
# (FIMEXCU_TMA_L2PROMO = 0..3, read once by make_source_map(); the library default is 3)
for rep in 1 2; do
  for v in 3 2 0 1; do
    FIMEXCU_TMA_L2PROMO=$v python bench.py --steps 40 --warmup 5 --no-e2e --no-job --no-cpu-baseline --no-secondary 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('promo=$v ms_per_step', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4))"
  done
done
