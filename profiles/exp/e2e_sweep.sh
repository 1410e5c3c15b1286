#!/bin/bash
# host-pointer path (plug-in call sequence, pageable input) against the pipelining knobs; one process per setting
for t in 1 2 4 8 16; do
  FIMEXCU_E2E_COPY_THREADS=$t python bench.py --steps 6 --warmup 3 --no-secondary --no-job --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); e=d['e2e']
print('copy threads $t  plugin %.2f ms  pinned %.2f ms  i16 %.2f ms' % (e['ms_per_step'], e['pinned_preallocated']['ms_per_step'], e['packed_i16']['ms_per_step']))"
done
