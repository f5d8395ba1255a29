"""Pop-time culling study on the oracle (oracle/prt_oracle.cpp, bvhIntersectSingleCull): node pops / leaf triangles per
single-rule closest-hit ray with (0) the reference's push-time-only culling, (1) entry-distance culling at pop, (2) the
provably result-preserving swizzle-axis bound; and whether the rendered image stays bit-identical."""
import os, sys, ctypes as C
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import refprt
from prt_b200 import scenes
cfgs = {"c1": (lambda: scenes.cornell_teapot(128, 128, teapot_refl=scenes.DIFFUSE, light="directional"), 16),
        "c3": (lambda: scenes.cornell_teapot(192, 108, teapot_refl=scenes.SPECULAR, light="infinite"), 32),
        "c4": (lambda: scenes.atrium(192, 108), 16),
        "c5": (lambda: scenes.city(192, 108), 8)}
lib = refprt._load_oracle()
for name in sys.argv[1:] or list(cfgs):
    mk, spp = cfgs[name]
    o = refprt.OracleScene(mk())
    base = None
    for mode in (0, 1, 2):
        lib.orc_set_popcull(mode)
        refprt.oracle_stats(enable=True, reset=True)
        img = o.render(spp, 1, nthreads=1)
        s = refprt.oracle_stats()
        if base is None: base = img.copy()
        same = np.array_equal(img.view(np.uint32), base.view(np.uint32))
        print(name, "mode", mode, "nodes/ray %.2f prims/ray %.2f" % (s["single_nodes"] / max(1, s["single_rays"]), s["single_prims"] / max(1, s["single_rays"])),
              "rays", s["single_rays"], "bit-identical" if same else "DIFFERS (%d px)" % int((img != base).any(axis=-1).sum()), flush=True)
    lib.orc_set_popcull(0)
