#!/usr/bin/env python
"""Condense an ncu report (--page raw --csv) into the handful of metrics the
roofline discussion in DESIGN.md uses.  usage: ncu_summary.py report.ncu-rep [kernel-substr]"""
import csv
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
    ("lts__t_sector_hit_rate.pct", "l2hit%"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
    ("launch__registers_per_thread", "regs"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
    ("sm__inst_executed.sum", "inst"),
    ("smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "st_long_sb"),
    ("smsp__average_warp_latency_issue_stalled_lg_throttle.ratio", "st_lg_thr"),
    ("smsp__average_warp_latency_issue_stalled_no_instruction.ratio", "st_noinst"),
    ("smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio", "st_math"),
    ("smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio", "st_short_sb"),
    ("smsp__average_warp_latency_issue_stalled_barrier.ratio", "st_barrier"),
    ("smsp__average_warp_latency_issue_stalled_wait.ratio", "st_wait"),
    ("smsp__average_warp_latency_issue_stalled_not_selected.ratio", "st_notsel"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "fp64%"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "fp64cyc%"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"),
    ("smsp__average_warp_latency_issue_stalled_mio_throttle.ratio", "st_mio"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem_wavefronts"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conflicts"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex%"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts%"),
    ("sm__cycles_elapsed.max", "cycles"),
    ("l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1_ld_sect"),
    ("l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1_ld_req"),
    ("l1tex__t_sector_hit_rate.pct", "l1hit%"),
    ("sm__icc_hit_rate.pct", "icc%"),   # may be absent
]


def main():
    rep = sys.argv[1]
    sub = sys.argv[2] if len(sys.argv) > 2 else ""
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    ki = h.index("Kernel Name")
    idx = [(h.index(k), n) for k, n in KEYS if k in h]
    for r in rows[2:]:
        if sub and sub not in r[ki]:
            continue
        name = r[ki].split("(")[0][-48:]
        print(name, "grid", r[h.index("Grid Size")] if "Grid Size" in h else "")
        print("   " + "  ".join(f"{n}={r[i]}{units[i] if n in ('time','dram_rd','dram_wr') else ''}" for i, n in idx))


main()
