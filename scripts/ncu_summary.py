#!/usr/bin/env python
"""Summarise an ncu report: per kernel duration, registers, issue-slot use, pipe use, stall reasons.
    python scripts/ncu_summary.py gpurun_out/x.ncu-rep [kernel-substring]"""
import csv
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "duration"), ("launch__registers_per_thread", "registers"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots used %"),
        ("smsp__inst_executed.sum", "warp instructions"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma pipe %"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu pipe %"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu pipe %"),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex %"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "shared wavefronts"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared bank conflicts"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("dram__bytes_read.sum", "dram read"),
        ("dram__bytes_write.sum", "dram write"), ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram %")]
STALLS = ["long_scoreboard", "short_scoreboard", "wait", "barrier", "mio_throttle", "lg_throttle", "math_pipe_throttle",
          "not_selected", "branch_resolving", "dispatch_stall", "no_instruction", "drain", "membar", "sleeping", "tex_throttle"]


def main():
    rep, sub = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        if sub not in name:
            continue
        print("== " + name)
        for k, label in KEYS:
            if k in hdr:
                print("   %-24s %s %s" % (label, r[hdr.index(k)], units[hdr.index(k)]))
        st = []
        for s in STALLS:
            k = "smsp__average_warps_issue_stalled_%s_per_issue_active.ratio" % s
            if k in hdr:
                st.append((float(r[hdr.index(k)] or 0), s))
        print("   stalls (warps per issue): " + ", ".join("%s %.2f" % (s, v) for v, s in sorted(st, reverse=True)[:7]))


if __name__ == "__main__":
    main()
