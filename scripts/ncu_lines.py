#!/usr/bin/env python
"""Per-source-line hot spots of one kernel from an ncu report taken with --import-source on.

usage: ncu_lines.py <report.ncu-rep> <kernel-regex> [top-n]

The SASS page of the report (instructions executed, stall samples per instruction) is joined with
the line table of the library that was profiled (nvdisasm --print-line-info on the cubin inside
bear_b200/libbear_cuda.so, which must be the same build).
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def line_table(kernel_re):
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(REPO, "bear_b200/libbear_cuda.so")], cwd=d,
                       check=True, stdout=subprocess.DEVNULL)
        cubin = [f for f in os.listdir(d) if f.startswith("kernels.") and f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(d, cubin)], check=True,
                             capture_output=True, text=True).stdout
    func, cur, table = None, None, {}
    for ln in txt.split("\n"):
        s = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if s:
            func, cur = s.group(1), None
            table[func] = []
            continue
        f = re.match(r'\s*//## File "([^"]+)", line (\d+)', ln)
        if f:
            cur = int(f.group(2))
            continue
        a = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if a and func:
            table[func].append((cur, a.group(2)))
    def base(mangled):  # _ZN8bearcuda10mix_kernelENS_6LayoutEi -> mix_kernel
        i, last = mangled.find("_ZN") + 3, mangled
        while i >= 3 and i < len(mangled) and mangled[i].isdigit():
            j = i
            while mangled[j].isdigit():
                j += 1
            n = int(mangled[i:j])
            last, i = mangled[j:j + n], j + n
        return last
    names = [k for k in table if re.fullmatch(kernel_re, base(k))]
    if not names:
        raise SystemExit(f"kernel regex matches nothing in {sorted(table)}")
    return [table[n] for n in names]   # one per template instantiation


def sass_page(report, kernel_re):
    out = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv", "--kernel-name", f"regex:^{kernel_re}(<.*>)?$"],
                         check=True, capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    heads = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    h = rows[heads[0]]
    end = heads[1] if len(heads) > 1 else len(rows)
    ie, smp = h.index("Instructions Executed"), h.index("# Samples")
    stall_cols = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    body = [r for r in rows[heads[0] + 1:end] if len(r) > ie and r[ie].isdigit()]
    return body, ie, smp, stall_cols


def main():
    report, kre = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    body, ie, smp, stall_cols = sass_page(report, kre)
    tables = [t for t in line_table(kre) if len(t) == len(body)]
    if not tables:
        raise SystemExit(f"no instantiation in the library has the report's {len(body)} instructions: not the same build")
    table = tables[0]
    samples, insts, stalls = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
    for (line, _), r in zip(table, body):
        samples[line] += int(r[smp])
        insts[line] += int(r[ie])
        for i, c in stall_cols:
            if r[i].isdigit() and int(r[i]):
                stalls[line][c[6:]] += int(r[i])
    ts, ti = sum(samples.values()), sum(insts.values())
    src = open(os.path.join(REPO, "bear_b200/csrc/kernels.cu")).read().split("\n")
    print(f"{kre}: {ti} warp instructions, {ts} samples")
    for line, s in samples.most_common(top):
        why = ",".join(f"{k}:{v}" for k, v in stalls[line].most_common(3))
        text = src[line - 1].strip()[:90] if line else ""
        print(f"{100 * s / ts:5.1f}% smp {100 * insts[line] / ti:5.1f}% inst  {line}: {text}   [{why}]")


if __name__ == "__main__":
    main()
