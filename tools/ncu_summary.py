#!/usr/bin/env python
"""Summarise an `ncu --page raw --csv` export: one block per profiled launch with the metrics DESIGN.md argues from
(duration, DRAM bytes, L1/L2 hit rates, issue utilisation, FP32 pipe, warp execution efficiency, occupancy, stalls).

  python tools/ncu_summary.py gpurun_out/full_torus1m_raw.csv > profiles/r01_wf_ncu_torus1m.txt
"""
import csv
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "dram read"),
    ("dram__bytes_write.sum", "dram write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1 throughput %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("sm__inst_executed_pipe_fp32.avg.pct_of_peak_sustained_active", "fp32 pipe %"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma pipe %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu pipe %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu (sfu) pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu pipe %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active threads / warp instr"),
    ("smsp__thread_inst_executed_per_inst_executed.pct", "warp execution efficiency %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__occupancy_limit_registers", "occupancy limit (regs, blocks)"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "stall long scoreboard %"),
    ("smsp__warp_issue_stalled_short_scoreboard_per_warp_active.pct", "stall short scoreboard %"),
    ("smsp__warp_issue_stalled_wait_per_warp_active.pct", "stall wait %"),
    ("smsp__warp_issue_stalled_math_pipe_throttle_per_warp_active.pct", "stall math pipe %"),
    ("smsp__warp_issue_stalled_not_selected_per_warp_active.pct", "stall not selected %"),
    ("smsp__warp_issue_stalled_barrier_per_warp_active.pct", "stall barrier %"),
    ("smsp__warp_issue_stalled_branch_resolving_per_warp_active.pct", "stall branch %"),
    ("smsp__warp_issue_stalled_lg_throttle_per_warp_active.pct", "stall lg throttle %"),
    ("smsp__warp_issue_stalled_mio_throttle_per_warp_active.pct", "stall mio throttle %"),
    ("smsp__warp_issue_stalled_dispatch_stall_per_warp_active.pct", "stall dispatch %"),
    ("smsp__warp_issue_stalled_no_instruction_per_warp_active.pct", "stall no instruction %"),
]


def main(path):
    rows = list(csv.reader(open(path, newline="")))
    hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    names, units = rows[hdr], rows[hdr + 1]
    col = {n: i for i, n in enumerate(names)}
    print(f"# {path}: ncu --set full --clock-control none, one block per profiled launch")
    for r in rows[hdr + 2:]:
        if len(r) < len(names):
            continue
        print(f"\n== {r[col['Kernel Name']]}  (id {r[col['ID']]})")
        for key, label in KEYS:
            if key in col:
                print(f"  {label:34s} {r[col[key]]:>16s} {units[col[key]]}")


if __name__ == "__main__":
    main(sys.argv[1])
