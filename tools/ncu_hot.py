#!/usr/bin/env python
"""Per-instruction stall breakdown of one kernel launch in an .ncu-rep (top N instructions by samples).

    python tools/ncu_hot.py gpurun_out/prof.ncu-rep [N] [launch-index]
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
launch = sys.argv[3] if len(sys.argv) > 3 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", launch, "--launch-count", "1"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
body = [r for r in rows[2:] if len(r) == len(hdr) and r[0].startswith("0x")]
total = sum(int(r[col["# Samples"]]) for r in body)
tot_by = {s: sum(int(r[col[s]]) for r in body) for s in stalls}
print(rows[0][1])
print("total samples", total, " instructions", len(body), " executed",
      sum(int(r[col["Instructions Executed"]]) for r in body))
print("by reason:", ", ".join(f"{s[6:]} {100.0 * v / total:.1f}%" for s, v in sorted(tot_by.items(), key=lambda kv: -kv[1]) if v))
order = sorted(range(len(body)), key=lambda i: -int(body[i][col["# Samples"]]))[:top]
for i in order:
    r = body[i]
    n = int(r[col["# Samples"]])
    why = sorted(((int(r[col[s]]), s[6:]) for s in stalls), reverse=True)[:3]
    print(f"{100.0 * n / total:5.1f}%  #{i:4d} {r[col['Source']].strip():60s} " + " ".join(f"{w}:{100.0 * c / max(n, 1):.0f}%" for c, w in why if c))
    if i > 0:
        print(f"            prev: {body[i - 1][col['Source']].strip()}")
