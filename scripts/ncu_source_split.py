#!/usr/bin/env python
"""Splits an `ncu --page source --csv --print-source cuda,sass` dump of several launches into one per-source-line summary
per launch (what scripts/ncu_lines.py prints): <prefix>_<k>_<kernel>.txt.  usage: ncu_source_split.py dump.csv prefix [top]"""
import csv
import gzip
import re
import sys

path, prefix, top = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 30
launches, cur, pending_file = [], None, None
opener = gzip.open if path.endswith(".gz") else open
for r in csv.reader(opener(path, "rt")):
    if not r:
        continue
    # layout of the dump: per launch, one section per source file — "File Path" row, "Function Name" row, "Line No" header, then
    # source rows (address column "-") with their SASS rows underneath.  A file that repeats (or a new function) starts the next launch.
    if r[0] == "File Path":
        pending_file = r[1].split("/")[-1]; continue
    if r[0] in ("Kernel Name", "Function Name") and len(r) > 1:
        if cur is None or cur["name"] != r[1] or pending_file in cur["files"]:
            cur = {"name": r[1], "files": set(), "file": None, "hdr": None, "lines": {}}
            launches.append(cur)
        cur["file"] = pending_file; cur["files"].add(pending_file); cur["hdr"] = None
        continue
    if cur is None:
        continue
    if r[0] == "Line No":
        cur["hdr"] = r; continue
    if cur["hdr"] is None or cur["file"] is None or len(r) < 3 or r[2] != "-":
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    h = cur["hdr"]
    ie = float(r[h.index("Instructions Executed")] or 0)
    te = float(r[h.index("Thread Instructions Executed")] or 0)
    sm = float(r[h.index("# Samples")] or 0)
    if ie:
        cur["lines"][(cur["file"], ln)] = (ie, te, sm, r[1].strip()[:110])
for k, L in enumerate(launches):
    lines = L["lines"]
    if not lines:
        continue
    short = re.sub(r"[^A-Za-z0-9<>]", "", re.sub(r"\(.*", "", L["name"]))[:40]
    ti = sum(v[0] for v in lines.values()); ts = sum(v[2] for v in lines.values()) or 1; tt = sum(v[1] for v in lines.values())
    with open(f"{prefix}_{k}_{short}.txt", "w") as f:
        f.write(f"{L['name']}\ntotal warp-instr {ti:.4g}  thread-instr {tt:.4g}  avg lanes {tt / ti:.2f}\n-- by instructions\n")
        for (fn, ln), (ie, te, sm, src) in sorted(lines.items(), key=lambda x: -x[1][0])[:top]:
            f.write(f"{fn[:16]:16s}:{ln:4d} inst {100 * ie / ti:5.1f}%  lanes {te / ie:5.1f}  samples {100 * sm / ts:5.1f}%  {src}\n")
        f.write("-- by stall samples\n")
        for (fn, ln), (ie, te, sm, src) in sorted(lines.items(), key=lambda x: -x[1][2])[:top]:
            f.write(f"{fn[:16]:16s}:{ln:4d} inst {100 * ie / ti:5.1f}%  lanes {te / ie:5.1f}  samples {100 * sm / ts:5.1f}%  {src}\n")
