#!/bin/bash
# A/B on one box: pageable -> pinned bounce copies with memcpy (default) or with streaming stores (FIMEXCU_E2E_NT_COPY=1)
# (the variant is not in the library: host_copy() in fxcu_apply.cu used hostfill::copy_lines + _mm_sfence per range when the variable was set)
for rep in 1 2 3; do
  for v in 0 1; do
    if [ $v = 1 ]; then export FIMEXCU_E2E_NT_COPY=1; else unset FIMEXCU_E2E_NT_COPY; fi
    python bench.py --steps 8 --warmup 3 --no-secondary --no-job --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.readlines()[-1]); e=d['e2e']; print('nt=$v plug-in ms', round(e['ms_per_step'],2), 'pinned ms', round(e['pinned_preallocated']['ms_per_step'],2), 'i16 ms', round(e['packed_i16']['ms_per_step'],2))"
  done
done
