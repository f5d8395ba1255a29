"""Traversal counters per ray class for the render workloads (nodes popped / triangles tested per ray, from the oracle's
counters — the same definitions as the reference's PRT_ENABLE_STATS, bvh.cpp:464-466,555-557), measured on a reduced-
resolution render of each config.  Feeds RENDER_CONSTANTS in bench.py (roofline bytes per ray)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import refprt
from prt_b200 import scenes
cfgs = {"c1": (lambda: scenes.cornell_teapot(128, 128, teapot_refl=scenes.DIFFUSE, light="directional"), 16),
        "c3": (lambda: scenes.cornell_teapot(192, 108, teapot_refl=scenes.SPECULAR, light="infinite"), 32),
        "c4": (lambda: scenes.atrium(192, 108), 16),
        "c5": (lambda: scenes.city(192, 108), 8)}
out = {}
for name in sys.argv[1:] or list(cfgs):
    mk, spp = cfgs[name]
    o = refprt.OracleScene(mk())
    refprt.oracle_stats(enable=True, reset=True)
    o.render(spp, 1, nthreads=os.cpu_count())
    s = refprt.oracle_stats()
    k = (s["single_nodes"] / max(1, s["single_rays"]), s["single_prims"] / max(1, s["single_rays"]),
         s["occl_nodes"] / max(1, s["occl_rays"]), s["occl_prims"] / max(1, s["occl_rays"]),
         s["packet_nodes"] / max(1, s["packet_rays"]), s["packet_prims"] / max(1, s["packet_rays"]))
    out[name] = tuple(round(x, 3) for x in k)
    print(name, out[name], s, flush=True)
print(json.dumps(out))
