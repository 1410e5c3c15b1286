#!/usr/bin/env python
"""Print the key metrics of an .ncu-rep (run here, no GPU needed):  python profiles/ncu_summary.py file.ncu-rep"""
import csv
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__t_requests_pipe_lsu_mem_global_op_st.sum', 'l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum',
        'l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct', 'l1tex__m_l1tex2xbar_write_bytes.sum', 'l1tex__m_xbar2l1tex_read_bytes.sum',
        'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum', 'lts__t_sector_hit_rate.pct', 'sm__cycles_elapsed.max',
        'smsp__cycles_active.avg']


def main(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print('--- kernel', r[hdr.index('Kernel Name')][:90], 'grid', r[hdr.index('Grid Size')], 'block', r[hdr.index('Block Size')])
        for k in KEYS:
            if k in hdr:
                print(f'{k:75s} {r[hdr.index(k)]:>20s} {units[hdr.index(k)]}')
        stalls = []
        for i, k in enumerate(hdr):
            if k.startswith('smsp__pcsamp_warps_issue_stalled') and 'not_issued' not in k:
                try:
                    stalls.append((float(r[i]), k.replace('smsp__pcsamp_warps_issue_stalled_', '')))
                except ValueError:
                    pass
        tot = sum(v for v, _ in stalls) or 1
        print('stalls:', ', '.join(f'{n} {100 * v / tot:.0f}%' for v, n in sorted(stalls, reverse=True)[:6]))


if __name__ == '__main__':
    main(sys.argv[1])
