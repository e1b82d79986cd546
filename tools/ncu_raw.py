#!/usr/bin/env python
"""Print the handful of ncu raw-page metrics we track.  usage: ncu_raw.py prof_raw.csv"""
import csv, re, sys
rows = list(csv.reader(open(sys.argv[1])))
H, U = rows[0], rows[1]
want = re.compile(r'^(Kernel Name|gpu__time_duration.sum|dram__bytes_(read|write).sum|lts__t_sector_hit_rate.pct|l1tex__t_sector_hit_rate.pct|sm__warps_active.avg.pct_of_peak_sustained_active|launch__registers_per_thread|launch__grid_size|smsp__thread_inst_executed_per_inst_executed.ratio|sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active|smsp__issue_active.avg.pct_of_peak_sustained_active|smsp__average_warps_issue_stalled_(long_scoreboard|wait|short_scoreboard|not_selected|branch_resolving|no_instruction|math_pipe_throttle|lg_throttle|selected)_per_issue_active.ratio|lts__t_bytes.sum|lts__throughput.avg.pct_of_peak_sustained_elapsed|l1tex__throughput.avg.pct_of_peak_sustained_elapsed|sm__throughput.avg.pct_of_peak_sustained_elapsed|smsp__inst_executed.sum|sass__inst_executed_local_(loads|stores)|l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct|l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate.pct|lts__t_sectors_srcunit_tex_op_read.sum|l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum|l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum|l1tex__data_pipe_lsu_wavefronts.sum|dram__throughput.avg.pct_of_peak_sustained_elapsed)$')
for i, h in enumerate(H):
    if want.search(h):
        print("%-86s %-9s %s" % (h, U[i][:9], " | ".join(r[i][:38] for r in rows[2:])))
