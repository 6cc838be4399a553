#!/usr/bin/env python
"""Text summary of an `ncu --set full` capture (raw page as CSV): per kernel the duration, DRAM bytes, and the
utilisation figures DESIGN.md quotes.   ncu -i X.ncu-rep --page raw --csv > raw.csv; python tools/ncu_summary.py raw.csv"""
import csv
import sys

METRICS = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
           ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
           ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe active %"),
           ("sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_elapsed", "tc pipe active %"),
           ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed", "XU (MUFU) %"),
           ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
           ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
           ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 throughput %"),
           ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2 -> SM bytes"),
           ("launch__registers_per_thread", "registers / thread"),
           ("launch__grid_size", "grid"), ("launch__block_size", "block"),
           ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall long scoreboard / issue"),
           ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall wait / issue")]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    seen = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0].split("::")[-1]
        seen.setdefault(name, []).append(r)
    for name, rs in seen.items():
        r = rs[-1]                                  # the last captured launch of each kernel
        print(f"== {name}  ({len(rs)} launches captured; last one shown)")
        for m, label in METRICS:
            if m in col:
                print(f"   {label:34s} {r[col[m]]:>16s} {units[col[m]]}")
        print()


if __name__ == "__main__":
    main()
