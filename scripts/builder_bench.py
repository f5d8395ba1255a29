"""Host BVH build time: ours (prt_cuda_scene_add_mesh on a host-only context: reference-identical binned SAH, cached
centroids, parallel subtrees, + Node64 flatten) vs the unmodified reference's Bvh::build, same meshes, this machine."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import prt_b200
from prt_b200 import scenes
from oracle import refprt
for name, mk in (("cornell+teapot (15.7 k)", lambda: scenes.cornell_teapot(64, 64)), ("bunny stand-in (82 k)", lambda: scenes.bunny_standin(64, 64)),
                 ("atrium (262 k)", lambda: scenes.atrium(64, 64, tex_size=16)), ("city (4.9 M)", lambda: scenes.city(64, 64))):
    d = mk()
    t0 = time.perf_counter(); ctx = prt_b200.load_scene(d, device=-1); t1 = time.perf_counter(); ctx.close()
    with refprt.quiet_stdout():
        t2 = time.perf_counter(); r = refprt.RefScene(d); t3 = time.perf_counter()
    print(f"{name:26s} ours {t1 - t0:7.2f} s   reference {t3 - t2:7.2f} s   ({os.cpu_count()} host threads available)", flush=True)
