#!/usr/bin/env python
"""Summarise an ncu --set full report of one step (scripts/gpu_profile_all.sh) for profiles/:

    extract_profile.py <report.ncu-rep> <out.csv> [workload]

writes the per-kernel metrics the DESIGN/README tables quote to <out.csv> and updates
profiles/kernel_traffic.json[workload] (DRAM read+write bytes per launch, read by bench.py)."""
import csv
import io
import json
import pathlib
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__warps_active.avg.per_cycle_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
]
UNIT = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def main():
    rep, out = sys.argv[1], sys.argv[2]
    workload = sys.argv[3] if len(sys.argv) > 3 else "cfg2_ht"
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units = rows[0], rows[1]
    unit_of = dict(zip(head, units))
    stall = [h for h in head if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
    traffic = {}
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel"] + [f"{k} [{unit_of.get(k, '')}]" for k in KEYS] + ["top stalls (warps per issue)"])
        for r in rows[2:]:
            d = dict(zip(head, r))
            name = d["Kernel Name"].split("(")[0].replace("void ", "").split("<")[0]
            st = sorted(((float(d[h] or 0), h.split("stalled_")[1].replace("_per_issue_active.ratio", "")) for h in stall),
                        reverse=True)[:4]
            w.writerow([name] + [d.get(k, "") for k in KEYS] + [" ".join(f"{n}:{v:.2f}" for v, n in st)])
            b = sum(float(d[k].replace(",", "")) * UNIT.get(unit_of[k], 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
            traffic[name] = int(b)
    tp = pathlib.Path(__file__).resolve().parent.parent / "profiles" / "kernel_traffic.json"
    allt = json.loads(tp.read_text()) if tp.exists() else {}
    allt[workload] = traffic
    tp.write_text(json.dumps(allt, indent=1) + "\n")
    print(json.dumps(traffic, indent=1))


if __name__ == "__main__":
    main()
