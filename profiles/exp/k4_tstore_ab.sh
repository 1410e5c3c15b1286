#!/bin/bash
# A/B of the headline kernel's store path: st.global.cs per level (shipped) against TMA tensor stores (FIMEXCU_K4_TSTORE=1)
# (the variant is not in the library: git apply profiles/exp/k4_tstore.patch and rebuild first)
for rep in 1 2; do
  for v in 0 1; do
    if [ $v = 1 ]; then export FIMEXCU_K4_TSTORE=1; else unset FIMEXCU_K4_TSTORE; fi
    python bench.py --steps 40 --warmup 5 --no-e2e --no-job --no-cpu-baseline --no-secondary 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); print('tstore=$v ms_per_step', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4))"
  done
done
FIMEXCU_K4_TSTORE=1 python -m pytest tests/test_gpu_horizontal.py -q -x -k "staged_kernel_padded or staged_and_direct or full_slab_properties or oracle_levels" 2>&1 | tail -3
