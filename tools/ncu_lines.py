#!/usr/bin/env python
"""tools/ncu_lines.py <report.ncu-rep> [top] — warp-stall samples of an ncu --set full --import-source on capture,
aggregated per CUDA source line (needs -lineinfo)."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
lines = []
for r in rows:
    if len(r) < 10 or not r[0].strip().isdigit():
        continue
    try:
        lines.append((int(r[6] or 0), int(r[7] or 0), int(r[0]), r[1][:120]))
    except ValueError:
        pass
tot = sum(l[0] for l in lines) or 1
print("total samples", tot)
for s, e, ln, src in sorted(lines, reverse=True)[:top]:
    print(f"{100 * s / tot:6.2f}%  exec {e:>11}  L{ln}: {src}")
