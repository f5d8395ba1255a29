"""Host-buffer calls with zero-copy input (PRT_ZC=0/1/2, read once per process): time per call for several chunk sizes."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import prt_b200
from prt_b200 import scenes
import bench
W, H = 1920, 1080
desc = scenes.bunny_standin(W, H)
ctx = prt_b200.load_scene(desc, 0)
d_rays, d_hits = bench.c2_scene_and_rays(ctx, torch, 1)
d_sh = bench.shadow_rays_from(torch, d_rays, d_hits, desc["light"]["dir"], ctx.radius())
n1, n2 = d_rays.shape[0], d_sh.shape[0]
h_rays = torch.empty((n1, 8), dtype=torch.float32).pin_memory(); h_rays.copy_(d_rays.cpu())
h_sh = torch.empty((n2, 8), dtype=torch.float32).pin_memory(); h_sh.copy_(d_sh.cpu())
h_hits = torch.empty((n1, 6), dtype=torch.float32).pin_memory(); h_occ = torch.empty(n2, dtype=torch.uint8).pin_memory()
L = ctx.L
def timeit(f, reps=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
def f1(): ctx.check(L.prt_cuda_intersect(ctx.h, n1, h_rays.data_ptr(), None, 0, h_hits.data_ptr()))
def f2(): ctx.check(L.prt_cuda_occluded(ctx.h, n2, h_sh.data_ptr(), 0, h_occ.data_ptr()))
zc = os.environ.get("PRT_ZC", "default"); dg = os.environ.get("PRT_D2H_GROUP", "default")
h_hits.zero_(); h_occ.zero_()
t1 = timeit(f1); t2 = timeit(f2)
ok1 = bool((h_hits.view(torch.int32) == d_hits.cpu().view(torch.int32)).all())
d_occ = torch.empty(n2, dtype=torch.uint8, device="cuda")
ctx.check(L.prt_cuda_occluded_dev(ctx.h, n2, d_sh.data_ptr(), 0, d_occ.data_ptr())); torch.cuda.synchronize()
ok2 = bool((h_occ == d_occ.cpu()).all())
print(f"PRT_ZC={zc} D2H_GROUP={dg}: intersect {t1:.3f} ms, occluded {t2:.3f} ms, e2e {(n1 + n2) / (t1 + t2) / 1e3:.0f} Mrays/s, identical {ok1} {ok2}", flush=True)
for c in ("65536", "131072", "524288", "1048576", "4194304"):
    os.environ["PRT_CHUNK_RAYS"] = c
    print(f"  chunk {c}: intersect {timeit(f1):.3f} ms occluded {timeit(f2):.3f} ms", flush=True)
