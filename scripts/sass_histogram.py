#!/usr/bin/env python
"""Per-kernel SASS opcode histogram of libbear_cuda.so (cuobjdump -sass), for profiles/.
    python scripts/sass_histogram.py > profiles/r2_sass_opcodes.txt"""
import collections
import re
import subprocess
import sys
from pathlib import Path

LIB = Path(__file__).resolve().parent.parent / "bear_b200" / "libbear_cuda.so"
KEY = ["UBLKCP", "UTMALDG", "SYNCS", "LDGSTS", "REDG", "FFMA2", "FMUL2", "FADD2", "FFMA", "FADD", "LDS", "STS", "LDG", "STG", "BAR",
       "BRA", "BRX", "SHFL", "ATOMS", "DFMA", "UTCHMMA", "LDTM", "HMMA"]


def main():
    out = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True).stdout
    kernels, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
            cur = kernels.setdefault(re.sub(r"\(.*", "", name).replace("bearcuda::", ""), collections.Counter())
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m and cur is not None:
            cur[m.group(1)] += 1
    print("SASS opcode counts per kernel of %s (static instruction counts; `cuobjdump -sass`)" % LIB.name)
    print("UBLKCP = bulk copy (TMA 1-D), SYNCS = mbarrier, LDGSTS = cp.async, REDG = red.global, FFMA2/FMUL2/FADD2 = packed f32x2;")
    print("no UTMALDG (tensor-map TMA), no tensor-core opcodes (UTCHMMA/HMMA/LDTM): by design, see DESIGN.md section 3\n")
    print("%-34s %6s " % ("kernel", "total") + " ".join("%7s" % k for k in KEY))
    for name, c in kernels.items():
        print("%-34s %6d " % (name[:34], sum(c.values())) + " ".join("%7d" % c[k] for k in KEY))


if __name__ == "__main__":
    sys.exit(main())
