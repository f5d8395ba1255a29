# A/B of the host-buffer pipeline's upload chunk size and download grouping (scripts/e2e_phase_probe.py)
for cfg in "262144 2" "131072 4" "131072 3" "196608 2" "262144 3" "98304 4" "65536 8" "262144 2"; do
  set -- $cfg
  PRT_CHUNK_RAYS=$1 PRT_D2H_GROUP=$2 python - <<PY
import os, sys, time, torch
sys.path.insert(0, ".")
import prt_b200, bench
from prt_b200 import scenes
desc = scenes.bunny_standin(1920, 1080)
ctx = prt_b200.load_scene(desc, 0)
d_rays, d_hits = bench.c2_scene_and_rays(ctx, torch, 1)
n1 = d_rays.shape[0]
h_rays = torch.empty((n1, 8), dtype=torch.float32).pin_memory(); h_rays.copy_(d_rays.cpu())
h_hits = torch.empty((n1, 6), dtype=torch.float32).pin_memory()
f = lambda: ctx.check(ctx.L.prt_cuda_intersect(ctx.h, n1, h_rays.data_ptr(), None, 0, h_hits.data_ptr()))
for _ in range(5): f()
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(40): f()
print("chunk $1 group $2: intersect %.3f ms" % ((time.perf_counter() - t0) / 40 * 1e3))
PY
done
