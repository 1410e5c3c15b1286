# e2e of the host entry point against its knobs, one process per setting, same box
run() {
  env "$@" python bench.py --steps 6 --warmup 3 --no-secondary --no-job --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); e=d['e2e']
print('$*  plugin %.2f ms (%.2f Gp/s)  pinned %.2f ms  i16 %.2f ms  d2h %d' % (e['ms_per_step'], e['value'], e['pinned_preallocated']['ms_per_step'], e['packed_i16']['ms_per_step'], e['d2h_bytes_per_step']))"
}
for r in 0 1 2 0 1 2; do run FIMEXCU_E2E_RAMP=$r; done
run FIMEXCU_E2E_BANDS=0 FIMEXCU_E2E_RAMP=0
