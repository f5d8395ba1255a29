"""Schedule of one prt_cuda_trace_jobs call on the c2 frame (PRT_PIPE_TRACE=1 makes the library print when each chunk's
upload, traversal and the last download finished)."""
import os, sys
os.environ["PRT_PIPE_TRACE"] = "1"
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import prt_b200
from prt_b200 import scenes, capi
import bench
desc = scenes.bunny_standin(1920, 1080)
ctx = prt_b200.load_scene(desc, 0)
d_rays, d_hits = bench.c2_scene_and_rays(ctx, torch, 1)
d_sh = bench.shadow_rays_from(torch, d_rays, d_hits, desc["light"]["dir"], ctx.radius())
n1, n2 = d_rays.shape[0], d_sh.shape[0]
h_rays = torch.empty((n1, 8), dtype=torch.float32).pin_memory(); h_rays.copy_(d_rays.cpu())
h_sh = torch.empty((n2, 8), dtype=torch.float32).pin_memory(); h_sh.copy_(d_sh.cpu())
h_hits = torch.empty((n1, 6), dtype=torch.float32).pin_memory(); h_occ = torch.empty(n2, dtype=torch.uint8).pin_memory()
jobs = (capi.prt_trace_job * 2)()
jobs[0].kind, jobs[0].n, jobs[0].mode, jobs[0].rays, jobs[0].out = 0, n1, 0, h_rays.data_ptr(), h_hits.data_ptr()
jobs[1].kind, jobs[1].n, jobs[1].mode, jobs[1].rays, jobs[1].out = 1, n2, 0, h_sh.data_ptr(), h_occ.data_ptr()
for _ in range(6):
    ctx.check(ctx.L.prt_cuda_trace_jobs(ctx.h, 2, jobs))
