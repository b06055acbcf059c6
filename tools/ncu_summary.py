#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed):  python tools/ncu_summary.py gpurun_out/prof.ncu-rep [out.md]"""
import csv
import subprocess
import sys

WANT = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sectors.sum', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'smsp__inst_executed.sum',
        'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_lg.sum', 'l1tex__data_pipe_lsu_wavefronts.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__cycles_elapsed.max',
        'lts__t_sectors_srcunit_tex_op_read.sum', 'lts__t_sectors_srcunit_tex_lookup_miss.sum',
        'l1tex__t_sector_hit_rate.pct']


def main():
    rep = sys.argv[1]
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    h, units = rows[0], rows[1]
    out = []
    ki = h.index('Kernel Name')
    for r in rows[2:]:
        out.append('### %s' % r[ki].split('(')[0])
        for w in WANT:
            if w in h:
                i = h.index(w)
                out.append('- %s = %s %s' % (w, r[i], units[i]))
        out.append('')
    text = '\n'.join(out)
    if len(sys.argv) > 2:
        with open(sys.argv[2], 'w') as f:
            f.write('# ncu --set full summary of %s\n\n' % rep + text + '\n')
    print(text)


if __name__ == '__main__':
    main()
