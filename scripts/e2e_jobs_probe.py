"""prt_cuda_trace_jobs on the c2 frame (closest-hit batch + occlusion batch in one call) beside the two-call form and the
pure copies of the same bytes; PRT_CHUNK_RAYS swept in-process (PRT_D2H_GROUP is read once: run the script per value)."""
import os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import prt_b200
from prt_b200 import scenes, capi
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
d_in = torch.empty((n1 + n2, 8), dtype=torch.float32, device="cuda"); d_out = torch.empty((n1, 6), dtype=torch.float32, device="cuda")
h_in = torch.empty((n1 + n2, 8), dtype=torch.float32).pin_memory()
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
L = ctx.L
jobs = (capi.prt_trace_job * 2)()
jobs[0].kind, jobs[0].n, jobs[0].mode, jobs[0].rays, jobs[0].out = 0, n1, 0, h_rays.data_ptr(), h_hits.data_ptr()
jobs[1].kind, jobs[1].n, jobs[1].mode, jobs[1].rays, jobs[1].out = 1, n2, 0, h_sh.data_ptr(), h_occ.data_ptr()
def timeit(f, reps=20):
    for _ in range(3): f()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): f()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
def copies():
    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
    with torch.cuda.stream(s2): h_hits.copy_(d_out, non_blocking=True)
    s1.synchronize(); s2.synchronize()
def up():
    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
    s1.synchronize()
def two(): 
    ctx.check(L.prt_cuda_intersect(ctx.h, n1, h_rays.data_ptr(), None, 0, h_hits.data_ptr()))
    ctx.check(L.prt_cuda_occluded(ctx.h, n2, h_sh.data_ptr(), 0, h_occ.data_ptr()))
def one(): ctx.check(L.prt_cuda_trace_jobs(ctx.h, 2, jobs))
nr = n1 + n2
print(f"PRT_D2H_GROUP={os.environ.get('PRT_D2H_GROUP', '2 (default)')}")
t = timeit(up); print(f"upload of all rays alone ({nr * 32 / 1e6:.0f} MB): {t:.3f} ms = {nr * 32 / 1e6 / t:.1f} GB/s -> bound {nr / t / 1e3:.0f} Mrays/s")
t = timeit(copies); print(f"upload + hit download at once: {t:.3f} ms -> bound {nr / t / 1e3:.0f} Mrays/s")
t = timeit(two); print(f"two calls: {t:.3f} ms = {nr / t / 1e3:.0f} Mrays/s")
t = timeit(one); print(f"trace_jobs: {t:.3f} ms = {nr / t / 1e3:.0f} Mrays/s")
for c in ("393216", "524288", "699056", "786432", "1048576"):
    os.environ["PRT_CHUNK_RAYS"] = c
    t = timeit(one); print(f"trace_jobs, chunk {c}: {t:.3f} ms = {nr / t / 1e3:.0f} Mrays/s")
