#!/usr/bin/env python
"""Condenses `ncu -i X.ncu-rep --page raw --csv` exports into the per-kernel summary CSVs kept in profiles/.
usage: python profiles/summarize.py out.csv raw1.csv [raw2.csv ...]"""
import csv
import sys

KEYS = """gpu__time_duration.sum launch__grid_size launch__block_size launch__registers_per_thread launch__occupancy_limit_registers
sm__warps_active.avg.pct_of_peak_sustained_active sm__inst_issued.avg.pct_of_peak_sustained_active
smsp__thread_inst_executed_per_inst_executed.ratio smsp__inst_executed.sum sm__throughput.avg.pct_of_peak_sustained_elapsed
dram__bytes_read.sum dram__bytes_write.sum dram__throughput.avg.pct_of_peak_sustained_elapsed lts__t_bytes.sum lts__t_sector_hit_rate.pct
l1tex__t_sector_hit_rate.pct l1tex__throughput.avg.pct_of_peak_sustained_active l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed
sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active
smsp__warps_eligible.avg.per_cycle_active
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_wait_per_issue_active.ratio
smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio
smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio
smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio
smsp__sass_inst_executed_op_local_ld.sum smsp__sass_inst_executed_op_local_st.sum""".split()

with open(sys.argv[1], "w", newline="") as out:
    w = csv.writer(out)
    for path in sys.argv[2:]:
        rows = list(csv.reader(open(path)))
        hdr, units, data = rows[0], rows[1], rows[2:]
        name = data[0][hdr.index("Kernel Name")]
        w.writerow(["# kernel", name, f"one column per captured launch ({path.split('/')[-1]})"])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                w.writerow([k, units[i]] + [r[i] for r in data])
