"""Generates tests/golden/surface.npz and tests/golden/render_c1_full.npz from the UNMODIFIED reference
(oracle/_ref/libprt_ref.so).  Run once where /root/reference exists:  python tests/golden/make_golden_r2.py

surface.npz          Scene::getSurfaceProperties (scene.cpp:30-45) on primary + secondary hit records of C1's scene.
render_c1_full.npz   BASELINE configs[0] at its full size: Cornell box + teapot, 512 x 512, 16 spp, diffuse + emissive +
                     directional light (PathTracer::TraceBlock, per-packet seeding).  The frame is stored as one CRC-32 per
                     image row (bit-identity of all 512 rows) plus 24 rows in full for diagnosis."""
import os, sys, zlib
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from prt_b200 import scenes
from oracle import refprt

W, H = 160, 90
desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
sc = refprt.RefScene(desc)
rays, hits = sc.primary_single((0, 0, W, H), 1, 3, nthreads=8)
t = hits[:, 0]; hit = t != -1
rng = np.random.RandomState(5)
d = rng.randn(int(hit.sum()), 3).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
sec = np.zeros((int(hit.sum()), 8), np.float32)
sec[:, 0:3] = rays[hit, 0:3] + rays[hit, 3:6] * t[hit, None]; sec[:, 3:6] = d; sec[:, 6] = 1e5
allhits = np.concatenate([hits, sc.intersect_single(sec, nthreads=8)])
surf = sc.surface_properties(allhits)
np.savez_compressed(os.path.join(HERE, "surface.npz"), W=W, H=H, hits=allhits, surface=surf)
print("surface.npz:", allhits.shape, surf.shape, "misses", int((allhits[:, 0] == -1).sum()))

W = H = 512
SPP, SEED = 16, 1
c1 = refprt.RefScene(scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional"))
img = c1.render(SPP, SEED, nthreads=8)
crc = np.array([zlib.crc32(img[y].tobytes()) for y in range(H)], np.uint32)
rows = np.arange(10, H, 21, dtype=np.int32)[:24]
np.savez_compressed(os.path.join(HERE, "render_c1_full.npz"), W=W, H=H, spp=SPP, seed=SEED, row_crc=crc, rows=rows, row_pixels=img[rows])
print("render_c1_full.npz: mean", float(img.mean()), "rows", rows.tolist())
