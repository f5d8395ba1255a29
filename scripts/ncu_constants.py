#!/usr/bin/env python
"""Writes profiles/<round>/ncu_constants.json from the round's `ncu --set full` raw-page CSV dumps: per-launch DRAM bytes, warp
instructions and L1 data-pipe wavefronts of the dominant kernels, tagged with the commit they were captured at.  bench.py reads
the file for the `roofline.traffic` / `roofline.issue` / `roofline.l1_data_pipe` fields (they are NOT measured live: ncu cannot
run inside a timed bench).   usage: ncu_constants.py <dir with *_raw.csv> <tag> <out.json> <commit>"""
import csv
import json
import os
import sys

src, tag, out, commit = sys.argv[1], sys.argv[2], sys.argv[3], sys.argv[4]


def num(s):
    return float(s.replace(",", "") or 0)


def launches(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    res = []
    for r in rows[2:]:
        g = lambda k: num(r[hdr.index(k)]) if k in hdr else None
        u = lambda k: units[hdr.index(k)] if k in hdr else ""
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        dr = g("dram__bytes_read.sum") * scale.get(u("dram__bytes_read.sum"), 1.0)
        dw = g("dram__bytes_write.sum") * scale.get(u("dram__bytes_write.sum"), 1.0)
        t = g("gpu__time_duration.sum") * {"us": 1e-3, "ms": 1.0, "ns": 1e-6, "s": 1e3}.get(u("gpu__time_duration.sum"), 1.0)
        name = r[hdr.index("Kernel Name")].split("(")[0].replace("void ", "").replace("<unnamed>::", "")
        res.append(dict(name=name, ms=t, dram_bytes=dr + dw, warp_inst=g("smsp__inst_executed.sum"), lanes=g("smsp__thread_inst_executed_per_inst_executed.ratio"),
                        l1_wavefronts=g("l1tex__data_pipe_lsu_wavefronts.sum"), l1_wavefront_pct=g("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"),
                        l1_hit=g("l1tex__t_sector_hit_rate.pct"), l2_hit=g("lts__t_sector_hit_rate.pct"), regs=g("launch__registers_per_thread"),
                        sm_pct=g("sm__throughput.avg.pct_of_peak_sustained_elapsed"), warps_active_pct=g("sm__warps_active.avg.pct_of_peak_sustained_active")))
    return res


C = {"_commit": commit, "_note": "per-launch numbers from ncu --set full --clock-control none (cold caches, serialised launches)"}
# c2: rays per launch are the bench's frame (16 588 800 / 9 521 192 packet-rule, 2 073 600 / 1 190 056 single-rule rays)
RAYS = {"c2_packet": {"intersect": 16588800, "occluded": 9521192}, "c2_single": {"intersect": 2073600, "occluded": 1190056}}
for key in ("c2_packet", "c2_single"):
    p = os.path.join(src, f"{tag}_{key}_raw.csv")
    if not os.path.exists(p):
        continue
    for L in launches(p):
        kind = "intersect" if "intersect" in L["name"] else "occluded"
        C[f"{key}_{kind}"] = dict(L, rays=RAYS[key][kind], source=f"profiles/r2/{tag}_{key}_raw.csv @ {commit}")
for wl in ("c3", "c4", "c5"):
    p = os.path.join(src, f"{tag}_{wl}_wf_raw.csv")
    if not os.path.exists(p):
        continue
    Ls = launches(p)
    tr = [L for L in Ls if L["name"].startswith("wfTrace")]
    if tr:
        big = max(tr, key=lambda L: L["ms"])
        C[f"{wl}_wftrace"] = dict(kernel=big["name"], ms=big["ms"], dram_bytes_per_launch=big["dram_bytes"], dram_gbs=big["dram_bytes"] / (big["ms"] * 1e-3) / 1e9,
                                  l1_data_pipe={"wavefronts_per_launch": big["l1_wavefronts"], "pct_of_peak": big["l1_wavefront_pct"], "l1_hit_pct": big["l1_hit"], "l2_hit_pct": big["l2_hit"]},
                                  lanes=big["lanes"], warp_inst=big["warp_inst"], regs=big["regs"], sm_pct=big["sm_pct"],
                                  source=f"profiles/r2/{tag}_{wl}_wf_raw.csv @ {commit} (largest wfTrace launch of the capture)")
        C[f"{wl}_launches"] = Ls
json.dump(C, open(out, "w"), indent=1)
print("wrote", out, [k for k in C if not k.startswith("_")])
