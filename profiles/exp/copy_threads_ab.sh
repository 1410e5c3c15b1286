#!/bin/bash
# plug-in sequence against the number of host threads that copy the pageable input into the pinned bounce chunks (same box, alternating)
for rep in 1 2; do
  for t in 4 8 12; do
    FIMEXCU_E2E_COPY_THREADS=$t FIMEXCU_E2E_TRACE=1 python bench.py --steps 6 --warmup 3 --no-secondary --no-job --no-cpu-baseline 2> gpurun_out/ct_$t.err | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); e=d['e2e']; print('copy threads $t: plug-in ms', round(e['ms_per_step'],2), 'pinned ms', round(e['pinned_preallocated']['ms_per_step'],2), 'i16 ms', round(e['packed_i16']['ms_per_step'],2))"
    grep "pageable in" gpurun_out/ct_$t.err | sed -n 4p | cut -c1-200
  done
done
