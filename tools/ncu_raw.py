#!/usr/bin/env python
"""Key per-kernel metrics of an ncu --set full report: python tools/ncu_raw.py report.ncu-rep"""
import csv
import subprocess
import sys

out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
H = rows[0]
want = [('gpu__time_duration.sum', 'dur'), ('dram__bytes_read.sum', 'rd'), ('dram__bytes_write.sum', 'wr'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'),
        ('sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'fma%'),
        ('sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'alu%'),
        ('sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'lsu%'),
        ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue%'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%'),
        ('launch__registers_per_thread', 'regs'), ('smsp__inst_executed.sum', 'winst'),
        ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'bankconf'),
        ('lts__t_sector_hit_rate.pct', 'l2hit%'), ('launch__waves_per_multiprocessor', 'waves'),
        ('sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'fp64%'),
        ('sm__inst_executed_pipe_tensor_subpipe_imma.avg.pct_of_peak_sustained_active', 'imma%'),
        ('lts__t_sectors_srcunit_tex_op_read.sum', 'l2rd_sectors'),
        # warps stalled per issued instruction, by reason (sum + 1 ~ resident warps per scheduler / issue rate)
        ('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'st_long'),
        ('smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'st_math'),
        ('smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'st_short'),
        ('smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'st_bar'),
        ('smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'st_wait'),
        ('smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio', 'st_notsel'),
        ('smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio', 'st_lg'),
        ('smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio', 'st_mio'),
        ('smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio', 'st_disp'),
        ('smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio', 'st_noinst')]
ki = H.index('Kernel Name')
for r in rows[2:]:
    parts = []
    for m, short in want:
        if m in H:
            v = r[H.index(m)]
            try:
                v = "%.4g" % float(v.replace(',', ''))
            except ValueError:
                pass
            parts.append("%s=%s%s" % (short, v, rows[1][H.index(m)] if short in ('dur', 'rd', 'wr') else ''))
    print(r[ki].split('(')[0][:34].ljust(34), ' '.join(parts))
