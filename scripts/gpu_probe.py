"""First GPU probe: traversal parity vs oracle/reference + raw kernel timings.  Run under gpurun."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import prt_b200
from prt_b200 import scenes
from oracle import refprt

def cmp_hits(a, b, name):
    same = np.array_equal(a.view(np.uint32), b.view(np.uint32))
    ta, pa, ma = refprt.hits_ids(a); tb, pb, mb = refprt.hits_ids(b)
    ha = ta[:, 0] != -1; hb = tb[:, 0] != -1
    idm = (ha != hb) | (ha & ((pa != pb) | (ma != mb)))
    print(f"{name}: {'BIT-IDENTICAL' if same else 'DIFF'} id-mismatch {idm.sum()}/{len(ta)} hitfrac {ha.mean():.4f}", flush=True)
    return same

W, H = 480, 270
desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
ctx = prt_b200.load_scene(desc, 0)
O = refprt.OracleScene(desc)
R = refprt.RefScene(desc) if refprt.available() else None
rays, avg, ohits = O.primary_packets((0, 0, W, H), 1, 5, nthreads=8)
g = ctx.intersect(rays, 1, avg)
cmp_hits(g, ohits, "packet primary vs oracle")
if R is not None:
    cmp_hits(g, R.intersect_packet(rays, avg, nthreads=8), "packet primary vs REFERENCE")
# device-generated primary rays vs oracle
n = W * H * 8
d_rays = torch.empty(n * 8, dtype=torch.float32, device="cuda"); d_avg = torch.empty(n // 8 * 3, dtype=torch.float32, device="cuda"); d_hits = torch.empty(n * 6, dtype=torch.float32, device="cuda")
ctx.primary_dev((0, 0, W, H), 1, 5, 1, d_rays.data_ptr(), d_avg.data_ptr(), d_hits.data_ptr()); ctx.synchronize()
print("device-generated rays identical:", np.array_equal(d_rays.cpu().numpy().reshape(-1, 8), rays), "avg:", np.array_equal(d_avg.cpu().numpy().reshape(-1, 3), avg))
cmp_hits(d_hits.cpu().numpy().reshape(-1, 6), ohits, "primary_dev packet vs oracle")
srays, shits = O.primary_single((0, 0, W, H), 2, 9, nthreads=8)
cmp_hits(ctx.intersect(srays, 0), shits, "single primary vs oracle")
# secondary-like rays
t, _, _ = refprt.hits_ids(shits); hit = t[:, 0] != -1
rng = np.random.RandomState(1); m = int(hit.sum()) // 8 * 8
org = (srays[hit, 0:3] + srays[hit, 3:6] * t[hit, 0:1])[:m]
dirs = rng.randn(m, 3).astype(np.float32); dirs /= np.linalg.norm(dirs, axis=1, keepdims=True)
sec = np.zeros((m, 8), np.float32); sec[:, 0:3] = org; sec[:, 3:6] = dirs; sec[:, 6] = 2 * O.radius()
cmp_hits(ctx.intersect(sec, 0), O.intersect_single(sec, nthreads=8), "secondary single vs oracle")
savg = sec[:, 3:6].reshape(-1, 8, 3).mean(1)
cmp_hits(ctx.intersect(sec, 1, savg), O.intersect_packet(sec, savg, nthreads=8), "secondary packet vs oracle")
a = ctx.occluded(sec, 0); b = O.occluded_single(sec, nthreads=8); print("occluded single mismatches", int((a != b).sum()), "of", m, "occl frac", a.mean())
a = ctx.occluded(sec, 1); b = O.occluded_packet(sec, np.full(m // 8, 255, np.uint8), nthreads=8); print("occluded packet mismatches", int((a != b).sum()))
# axis-aligned / degenerate rays (zero direction components -> NaN slab lanes)
ax = sec[:4096].copy()
ax[:, 3:6] = np.eye(3, dtype=np.float32)[rng.randint(0, 3, 4096)] * rng.choice([-1.0, 1.0], (4096, 1)).astype(np.float32)
ax[:, 0:3] = np.round(ax[:, 0:3] * 4) / 4
cmp_hits(ctx.intersect(ax, 0), O.intersect_single(ax), "axis-aligned single vs oracle")
axavg = ax[:, 3:6].reshape(-1, 8, 3).mean(1)
cmp_hits(ctx.intersect(ax, 1, axavg), O.intersect_packet(ax, axavg), "axis-aligned packet vs oracle")
print("axis occl single", int((ctx.occluded(ax, 0) != O.occluded_single(ax)).sum()), "packet", int((ctx.occluded(ax, 1) != O.occluded_packet(ax, np.full(512, 255, np.uint8))).sum()))

# ---- timings, 1920x1080 ----
W, H = 1920, 1080
ctx.set_camera(desc["camera"]["pos"], desc["camera"]["dir"], W, H)
for mode, ppp, name in ((1, 2, "packet"), (0, 16, "single")):
    n = W * H * ppp * (8 if mode == 1 else 1)
    d_hits = torch.empty(n * 6, dtype=torch.float32, device="cuda")
    d_rays = torch.empty(n * 8, dtype=torch.float32, device="cuda"); d_avg = torch.empty(max(n // 8, 1) * 3, dtype=torch.float32, device="cuda")
    for it in range(4):
        ctx.primary_dev((0, 0, W, H), ppp, 5 + it, mode, d_rays.data_ptr(), d_avg.data_ptr(), d_hits.data_ptr()); ctx.synchronize()
    ms = ctx.counters()["lastKernelMs"]
    # lastKernelMs only set by host paths; time with wall clock around sync instead
    t0 = time.time()
    for it in range(5):
        ctx.primary_dev((0, 0, W, H), ppp, 50 + it, mode, None, None, d_hits.data_ptr())
    ctx.synchronize(); dt = (time.time() - t0) / 5
    print(f"primary gen+trace {name}: {n/dt/1e6:.1f} Mrays/s ({dt*1e3:.2f} ms for {n} rays)")
    t0 = time.time()
    for it in range(5):
        ctx.intersect_dev(n, d_rays.data_ptr(), d_avg.data_ptr(), mode, d_hits.data_ptr())
    ctx.synchronize(); dt = (time.time() - t0) / 5
    print(f"intersect_dev {name}: {n/dt/1e6:.1f} Mrays/s ({dt*1e3:.2f} ms)")
    d_occ = torch.empty(n, dtype=torch.uint8, device="cuda")
    t0 = time.time()
    for it in range(5):
        ctx.occluded_dev(n, d_rays.data_ptr(), mode, d_occ.data_ptr())
    ctx.synchronize(); dt = (time.time() - t0) / 5
    print(f"occluded_dev(primary rays, maxT=1e5) {name}: {n/dt/1e6:.1f} Mrays/s")
if R is not None:
    t0 = time.time(); R.primary_packets((0, 0, W, H), 1, 5, nthreads=os.cpu_count(), want_rays=False); dt = time.time() - t0
    print(f"REFERENCE packet primary, {os.cpu_count()} threads: {W*H*8/dt/1e6:.1f} Mrays/s")
