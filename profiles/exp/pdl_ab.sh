#!/bin/bash
# A/B: programmatic dependent launch of the staged bilinear kernel (default) against plain stream order (FIMEXCU_NO_PDL=1)
for rep in 1 2; do
  for v in 0 1; do
    if [ $v = 1 ]; then export FIMEXCU_NO_PDL=1; else unset FIMEXCU_NO_PDL; fi
    python bench.py --steps 40 --warmup 5 --no-e2e --no-cpu-baseline --no-secondary 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); j=d['job']; print('no_pdl=$v ms_per_step', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4), 'job ms', round(j['device_resident']['ms'],3), 'two streams', round(j['device_resident_two_streams']['ms'],3))"
  done
done
python -m pytest tests/test_gpu_horizontal.py tests/test_gpu_fullsize.py -q -x 2>&1 | tail -3
