"""Times prt_cuda_intersect / prt_cuda_occluded with pinned host buffers for several pipeline chunk sizes."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import prt_b200
from prt_b200 import scenes
import bench
W, H = 1920, 1080
desc = scenes.bunny_standin(W, H)
for chunk in ("0", "4000000", "524288", "131072", "65536"):
    if chunk == "0": os.environ.pop("PRT_CHUNK_RAYS", None)
    else: os.environ["PRT_CHUNK_RAYS"] = chunk
    ctx = prt_b200.load_scene(desc, 0)
    d_rays, d_hits = bench.c2_scene_and_rays(ctx, torch, 1)
    d_sh = bench.shadow_rays_from(torch, d_rays, d_hits, desc["light"]["dir"], ctx.radius())
    n1, n2 = d_rays.shape[0], d_sh.shape[0]
    h_rays = torch.empty((n1, 8), dtype=torch.float32).pin_memory(); h_rays.copy_(d_rays.cpu())
    h_sh = torch.empty((n2, 8), dtype=torch.float32).pin_memory(); h_sh.copy_(d_sh.cpu())
    h_hits = torch.empty((n1, 6), dtype=torch.float32).pin_memory(); h_occ = torch.empty(n2, dtype=torch.uint8).pin_memory()
    L = ctx.L
    def f1(): ctx.check(L.prt_cuda_intersect(ctx.h, n1, h_rays.data_ptr(), None, 0, h_hits.data_ptr()))
    def f2(): ctx.check(L.prt_cuda_occluded(ctx.h, n2, h_sh.data_ptr(), 0, h_occ.data_ptr()))
    res = []
    for f in (f1, f2):
        for _ in range(3): f()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(20): f()
        res.append((time.perf_counter() - t0) / 20 * 1e3)
    print(f"chunk {chunk:>8}: intersect {res[0]:.3f} ms  occluded {res[1]:.3f} ms  -> {(n1 + n2) / (res[0] + res[1]) / 1e3:.0f} Mrays/s", flush=True)
    ctx.close()
