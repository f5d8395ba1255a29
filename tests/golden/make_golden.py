"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libprt_ref.so, built by `make -C oracle ref`
in the container that has /root/reference).  Run once:  python tests/golden/make_golden.py
The vectors pin oracle/prt_oracle.cpp (and through it the CUDA path) to outputs of the reference itself:
  * kat.npz       — SoA intersectTriangle on the reference's own known-answer case (tests/tests.cpp:113-126) + random rays
  * bvh.npz       — Bvh::build output (nodes, primRemapping) for the Cornell box, the teapot and a 9-triangle mesh
  * hits_c1.npz   — camera rays (packet + single), Scene::intersect / Scene::occluded results on Cornell+teapot
  * render_*.npz  — PathTracer::TracePacket images (per-packet seeding) for the C1 / shipped / C3-style setups
  * light.npz     — InfiniteAreaLight tables + sample() outputs
"""
import os, sys
import numpy as np
HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from prt_b200 import scenes
from oracle import refprt

rng = np.random.RandomState(2024)

# ---- KAT + random triangle tests ----
kat_rays = np.tile(np.array([0.4, 0.4, -1, 0, 0, 1, 1e30, 0], np.float32), (8, 1))
kat_tri = np.array([0, 0, 0, 1, 0, 0, 0, 1, 0], np.float32)
kat_out = refprt.intersect_triangle_soa(kat_rays, kat_tri)
assert kat_out[0, 0] == 1.0
rr, rt, ro = [], [], []
for _ in range(64):
    rays = np.zeros((8, 8), np.float32)
    rays[:, 0:3] = rng.uniform(-2, 2, (8, 3)); d = rng.randn(8, 3); rays[:, 3:6] = d / np.linalg.norm(d, axis=1, keepdims=True); rays[:, 6] = 1e30
    tri = rng.uniform(-1, 1, 9).astype(np.float32)
    rr.append(rays); rt.append(tri); ro.append(refprt.intersect_triangle_soa(rays, tri))
np.savez_compressed(os.path.join(HERE, "kat.npz"), kat_rays=kat_rays, kat_tri=kat_tri, kat_out=kat_out, rays=np.array(rr), tris=np.array(rt), out=np.array(ro))

# ---- scenes ----
W, H = 96, 54
c1 = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
R = refprt.RefScene(c1)
nine = dict(indices=np.arange(27, dtype=np.uint32).reshape(9, 3), positions=rng.uniform(-1, 1, (27, 3)).astype(np.float32), normals=None,
            texcoords=None, prim_material=np.zeros(9, np.uint32), materials=[dict(diffuse=(1, 1, 1))])
R9 = refprt.RefScene(dict(meshes=[nine]))
bv = {}
for name, sc, mi, npr in (("cornell", R, 0, 36), ("teapot", R, 1, 15704), ("nine", R9, 0, 9)):
    nodes, dec, remap = sc.bvh_dump(mi, npr)
    bv[name + "_nodes"] = nodes[:, :7]; bv[name + "_primcount"] = dec[:, 1]; bv[name + "_axis"] = dec[:, 2]; bv[name + "_remap"] = remap
    leaf = dec[:, 1] != 15
    bv[name + "_trivec"] = np.where(leaf, dec[:, 0], 0)
np.savez_compressed(os.path.join(HERE, "bvh.npz"), nine_positions=nine["positions"], **bv)

rays, avg, hits = R.primary_packets((0, 0, W, H), 1, 5)
srays, shits = R.primary_single((0, 0, W, H), 1, 9)
t = shits[:, 0]; hit = t != -1
m = int(hit.sum()) // 8 * 8
org = (srays[hit, 0:3] + srays[hit, 3:6] * t[hit, None])[:m]
d = rng.randn(m, 3).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
sec = np.zeros((m, 8), np.float32); sec[:, 0:3] = org; sec[:, 3:6] = d; sec[:, 6] = 2 * R.radius()
savg = sec[:, 3:6].reshape(-1, 8, 3).mean(1).astype(np.float32)
mb = rng.randint(1, 256, m // 8).astype(np.uint8)
np.savez_compressed(os.path.join(HERE, "hits_c1.npz"), width=W, height=H, radius=np.float32(R.radius()), camera=R.camera_basis(),
                    packet_rays=rays, packet_avg=avg, packet_hits=hits, single_rays=srays, single_hits=shits,
                    sec_rays=sec, sec_avg=savg, sec_hits_single=R.intersect_single(sec), sec_hits_packet=R.intersect_packet(sec, savg),
                    sec_occ_single=R.occluded_single(sec), sec_mask=mb, sec_occ_packet=R.occluded_packet(sec, mb))

for name, kw, spp in (("c1", dict(teapot_refl=scenes.DIFFUSE, light="directional"), 16),
                      ("shipped", dict(teapot_refl=scenes.SPECULAR, light=None), 16),
                      ("c3", dict(teapot_refl=scenes.SPECULAR, light="infinite"), 32)):
    sc = refprt.RefScene(scenes.cornell_teapot(64, 48, **kw))
    np.savez_compressed(os.path.join(HERE, f"render_{name}.npz"), image=sc.render(spp, 7, nthreads=8), spp=spp, seed=7)

c3 = refprt.RefScene(scenes.cornell_teapot(16, 16, light="infinite"))
v, hz, img = c3.infinite_light_tables()
u = rng.rand(4096, 2).astype(np.float32)
dd, cc = c3.light_sample(u)
np.savez_compressed(os.path.join(HERE, "light.npz"), vertP=v, horiP=hz, u=u, dir=dd, color=cc)
# textured config (C4 stand-in, small): alpha test inside traversal, bump, de-gamma
c4 = scenes.atrium(64, 48, detail=0.25, tex_size=64)
R4 = refprt.RefScene(c4)
r4, a4, h4 = R4.primary_packets((0, 0, 64, 48), 1, 3)
np.savez_compressed(os.path.join(HERE, "render_c4.npz"), image=R4.render(16, 7, nthreads=8), spp=16, seed=7, packet_hits=h4,
                    alpha_flags=np.array(R4.alpha_flags[0], np.uint8))
print("golden vectors written:", sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))
