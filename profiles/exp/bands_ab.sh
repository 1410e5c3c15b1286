# e2e of the host entry point with and without the row-band copy-back, against the pipelining knobs
run() {
  env "$@" python bench.py --steps 6 --warmup 3 --no-secondary --no-job --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); e=d['e2e']
print('$*  plugin %.2f ms (%.2f Gp/s)  pinned %.2f ms  i16 %.2f ms  d2h %d' % (e['ms_per_step'], e['value'], e['pinned_preallocated']['ms_per_step'], e['packed_i16']['ms_per_step'], e['d2h_bytes_per_step']))"
}
run FIMEXCU_E2E_BANDS=0
run FIMEXCU_E2E_BANDS=1
for c in 4 6 12 16; do run FIMEXCU_E2E_CHUNKS=$c; done
run FIMEXCU_E2E_STREAMS=2
run FIMEXCU_E2E_STREAMS=4 FIMEXCU_E2E_CHUNKS=12
