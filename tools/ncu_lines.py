#!/usr/bin/env python
"""Per-source-line instruction and stall-sample totals of one kernel in an .ncu-rep (needs -lineinfo + --import-source on).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep kernel_regex [N]
"""
import csv
import io
import subprocess
import sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
SORT = 1 if len(sys.argv) > 4 and sys.argv[4] == "samples" else 0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name",
                      f"regex:{kern}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = next(r for r in rows if r and r[0] == "Line No")
ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
lines = {}
for r in rows:
    if len(r) == len(hdr) and r[0].isdigit() and r[2] == "-":  # the per-line summary rows
        key = (int(r[0]), r[1].strip())
        a = lines.setdefault(key, [0, 0])
        a[0] += int(r[ie] or 0)
        a[1] += int(r[ss] or 0)
ti, ts = sum(v[0] for v in lines.values()), sum(v[1] for v in lines.values())
print(f"instructions {ti}, samples {ts}")
for (ln, src), (i, s) in sorted(lines.items(), key=lambda kv: -kv[1][SORT])[:top]:
    print(f"{100.0 * i / max(ti, 1):5.1f}% instr {100.0 * s / max(ts, 1):5.1f}% samples  L{ln:4d} {src[:105]}")
