#!/usr/bin/env python
"""Aggregates `ncu --page source --csv --print-source cuda,sass` by CUDA source line: instructions executed, average
active lanes and stall samples per line (top N by instructions)."""
import csv
import sys
path, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, hdr, lines = None, None, {}
for r in csv.reader(open(path)):
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if r[0] in ("Function Name", "Kernel Name") or hdr is None or cur is None:
        continue
    if r[2] != "-":        # SASS row under a source line; the source-line row itself carries the totals
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    ie = float(r[hdr.index("Instructions Executed")] or 0)
    te = float(r[hdr.index("Thread Instructions Executed")] or 0)
    sm = float(r[hdr.index("# Samples")] or 0)
    if ie:
        lines[(cur, ln)] = (ie, te, sm, r[1].strip()[:110])
ti = sum(v[0] for v in lines.values()); ts = sum(v[2] for v in lines.values()) or 1
tt = sum(v[1] for v in lines.values())
print(f"total warp-instr {ti:.3g}  thread-instr {tt:.3g}  avg lanes {tt / ti:.2f}")
for (f, ln), (ie, te, sm, src) in sorted(lines.items(), key=lambda x: -x[1][2 if len(sys.argv) > 3 and sys.argv[3] == "samples" else 0])[:top]:
    print(f"{f[:16]:16s}:{ln:4d} inst {100 * ie / ti:5.1f}%  lanes {te / ie:5.1f}  samples {100 * sm / ts:5.1f}%  {src}")
