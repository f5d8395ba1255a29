"""Where a host-buffer call spends its time: pure copies of the same sizes (one direction, both at once) beside the calls."""
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
d_tmp = torch.empty((n1, 8), dtype=torch.float32, device="cuda"); d_h = torch.empty((n1, 6), dtype=torch.float32, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
L = ctx.L
def timeit(f, reps=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
def h2d():
    with torch.cuda.stream(s1): d_tmp.copy_(h_rays, non_blocking=True)
    s1.synchronize()
def d2h():
    with torch.cuda.stream(s2): h_hits.copy_(d_h, non_blocking=True)
    s2.synchronize()
def both():
    with torch.cuda.stream(s1): d_tmp.copy_(h_rays, non_blocking=True)
    with torch.cuda.stream(s2): h_hits.copy_(d_h, non_blocking=True)
    s1.synchronize(); s2.synchronize()
def both_chunked(k=8):
    a = n1 // k
    for i in range(k):
        with torch.cuda.stream(s1): d_tmp[i * a:(i + 1) * a].copy_(h_rays[i * a:(i + 1) * a], non_blocking=True)
        with torch.cuda.stream(s2): h_hits[i * a:(i + 1) * a].copy_(d_h[i * a:(i + 1) * a], non_blocking=True)
    s1.synchronize(); s2.synchronize()
def f1(): ctx.check(L.prt_cuda_intersect(ctx.h, n1, h_rays.data_ptr(), None, 0, h_hits.data_ptr()))
def f2(): ctx.check(L.prt_cuda_occluded(ctx.h, n2, h_sh.data_ptr(), 0, h_occ.data_ptr()))
mb_in, mb_out = n1 * 32 / 1e6, n1 * 24 / 1e6
t = timeit(h2d); print(f"H2D {mb_in:.0f} MB alone: {t:.3f} ms = {mb_in / t:.1f} GB/s")
t = timeit(d2h); print(f"D2H {mb_out:.0f} MB alone: {t:.3f} ms = {mb_out / t:.1f} GB/s")
t = timeit(both); print(f"both at once: {t:.3f} ms = {(mb_in + mb_out) / t:.1f} GB/s total")
t = timeit(both_chunked); print(f"both, 8 chunks each: {t:.3f} ms")
t1 = timeit(f1); print(f"prt_cuda_intersect host buffers ({n1} rays): {t1:.3f} ms")
t2 = timeit(f2); print(f"prt_cuda_occluded host buffers ({n2} rays, {n2 * 32 / 1e6:.0f} MB in): {t2:.3f} ms")
print(f"e2e {(n1 + n2) / (t1 + t2) / 1e3:.0f} Mrays/s")
for c in ("65536", "131072", "524288", "1048576"):
    os.environ["PRT_CHUNK_RAYS"] = c
    print(f"chunk {c}: intersect {timeit(f1):.3f} ms occluded {timeit(f2):.3f} ms")
