"""Scene descriptions for the BASELINE.json configs, as plain dicts of numpy arrays.

A description is what the reference's setup functions (main.cpp:22-105) produce *after* loading:
indexed triangle meshes (Mesh buffers, mesh.h:83-98), a material table, optional textures, one light,
one camera.  The same dict feeds the CUDA path (prt_b200.Scene) and, in tests, the reference harness.

    desc = {"meshes": [ {indices (P,3) u32, positions (V,3) f32, normals (V,3) f32|None,
                         texcoords (V,2) f32|None, prim_material (P,) u32,
                         materials: [ {diffuse, emissive, refl, diffuse_tex, bump_tex} ]} ],
            "textures": [ {width, height, comp, texels (u8 or f32)} ],
            "light": None | {"type": "directional", dir, intensity} | {"type": "infinite", image (H,W,4) f32},
            "camera": {pos, dir, width, height}}

Bundled data: teapot.npz / cornell.npz (see data/make_assets.py).  Everything else is procedural and
seeded; the Stanford bunny, Sponza, San-Miguel and skylight-day.exr are not available offline
(.MISSING_LARGE_BLOBS, main.cpp:60,78,67), so stand-ins of the same scale are generated.
"""
import os
import numpy as np

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")

DIFFUSE, SPECULAR, REFRACTION = 0, 1, 2


def _mat(diffuse, emissive=(0, 0, 0), refl=DIFFUSE, diffuse_tex=-1, bump_tex=-1):
    return dict(diffuse=tuple(float(x) for x in diffuse), emissive=tuple(float(x) for x in emissive), refl=int(refl),
                diffuse_tex=int(diffuse_tex), bump_tex=int(bump_tex))


def cornell_mesh():
    d = np.load(os.path.join(_DATA, "cornell.npz"))
    mats = [_mat(d["diffuse"][i], d["emissive"][i], int(d["refl"][i])) for i in range(d["diffuse"].shape[0])]
    return dict(indices=d["indices"].astype(np.uint32), positions=d["positions"].astype(np.float32), normals=None,
                texcoords=None, prim_material=d["prim_material"].astype(np.uint32), materials=mats)


def teapot_mesh(refl=SPECULAR):
    """Teapot as placed by setupCornellBox (main.cpp:28-48): material diffuse 0.9, specular by default."""
    d = np.load(os.path.join(_DATA, "teapot.npz"))
    idx = d["indices"].astype(np.uint32)
    # loadObj sets m_hasTexcoord = true (mesh.cpp:206) with effectively undefined uv contents for this path
    # (mesh.cpp:198 copies vertexCount floats); the material has no texture so uv never reaches a result.
    # We define them as zeros on both sides.
    uv = np.zeros((d["positions"].shape[0], 2), np.float32)
    return dict(indices=idx, positions=d["positions"].astype(np.float32), normals=d["normals"].astype(np.float32),
                texcoords=uv, prim_material=np.zeros(idx.shape[0], np.uint32), materials=[_mat((0.9, 0.9, 0.9), refl=refl)])


def sky_image(w=256, h=128, seed=7):
    """Seeded lat-long RGBA32F sky: vertical gradient + sun lobe + mild noise (stand-in for skylight-day.exr)."""
    rng = np.random.RandomState(seed)
    v = (np.arange(h, dtype=np.float32) + 0.5) / h
    u = (np.arange(w, dtype=np.float32) + 0.5) / w
    uu, vv = np.meshgrid(u, v)
    horizon = np.clip(1.0 - np.abs(vv - 0.5) * 2.0, 0, 1)
    base = np.stack([0.35 + 0.4 * horizon, 0.5 + 0.35 * horizon, 0.9 + 0.1 * horizon], -1).astype(np.float32)
    base[vv > 0.5] *= 0.25
    su, sv = 0.62, 0.22
    d2 = ((uu - su) * 2.0) ** 2 + (vv - sv) ** 2
    sun = np.exp(-d2 / 0.0009).astype(np.float32)[..., None] * np.array([60.0, 54.0, 40.0], np.float32)
    img = base + sun + 0.02 * rng.rand(h, w, 3).astype(np.float32)
    out = np.ones((h, w, 4), np.float32)
    out[..., :3] = img
    return np.ascontiguousarray(out)


CORNELL_CAMERA = dict(pos=(0.0, 0.965, 2.6), dir=(0.0, 0.0, -1.0))  # main.cpp:53


def _normalize(v):
    v = np.asarray(v, np.float32)
    # same operation order as vecmath.h:1200-1204 (1/len then multiply), in fp32
    inv = np.float32(1.0) / np.sqrt(np.float32(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]), dtype=np.float32)
    return (inv * v).astype(np.float32)


def cornell_teapot(width, height, teapot_refl=SPECULAR, light=None, sky_seed=7):
    """Configs C1 (teapot_refl=DIFFUSE, light='directional') and C3 (SPECULAR, light='infinite')."""
    desc = dict(meshes=[cornell_mesh(), teapot_mesh(teapot_refl)], textures=[], light=None,
                camera=dict(CORNELL_CAMERA, width=int(width), height=int(height)))
    if light == "directional":
        desc["light"] = dict(type="directional", dir=_normalize((0.2, 0.4, 1.0)), intensity=np.array((16.7, 15.6, 11.7), np.float32))
    elif light == "infinite":
        desc["light"] = dict(type="infinite", image=sky_image(seed=sky_seed))
    return desc


def icosphere(subdiv):
    t = (1.0 + 5 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
                  [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                  [8, 6, 7], [9, 8, 1]], np.int64)
    for _ in range(subdiv):
        e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]], 0)
        es = np.sort(e, 1)
        key = es[:, 0] * (v.shape[0] + 1) + es[:, 1]
        uk, inv = np.unique(key, return_inverse=True)
        a = uk // (v.shape[0] + 1); b = uk % (v.shape[0] + 1)
        mid = v[a] + v[b]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = v.shape[0]
        v = np.concatenate([v, mid], 0)
        n = f.shape[0]
        m01 = base + inv[:n]; m12 = base + inv[n:2 * n]; m20 = base + inv[2 * n:]
        f = np.concatenate([np.stack([f[:, 0], m01, m20], 1), np.stack([f[:, 1], m12, m01], 1),
                            np.stack([f[:, 2], m20, m12], 1), np.stack([m01, m12, m20], 1)], 0)
    return v, f


def vertex_normals_area(pos, idx):
    """Area-weighted smooth normals (for procedural meshes only; both sides receive the resulting array)."""
    p = pos.astype(np.float64)
    fn = np.cross(p[idx[:, 1]] - p[idx[:, 0]], p[idx[:, 2]] - p[idx[:, 0]])
    n = np.zeros_like(p)
    for k in range(3):
        np.add.at(n, idx[:, k], fn)
    ln = np.linalg.norm(n, axis=1, keepdims=True)
    ln[ln == 0] = 1.0
    return (n / ln).astype(np.float32)


def bunny_standin(width=1920, height=1080, subdiv=6, seed=3):
    """Config C2 stand-in: noise-displaced icosphere, 20*4^6 = 81 920 triangles (Stanford bunny ~69k)."""
    v, f = icosphere(subdiv)
    rng = np.random.RandomState(seed)
    disp = np.zeros(v.shape[0])
    for k in range(1, 6):
        w = rng.randn(3, 8) * (2.0 ** k) * 0.6
        ph = rng.rand(8) * 6.283
        disp += (0.5 ** k) * np.sin(v @ w + ph).sum(1) * 0.12
    pos = (v * (1.0 + disp)[:, None] * 0.8).astype(np.float32)
    pos[:, 1] += 1.0
    idx = f.astype(np.uint32)
    # a perfectly flat mesh has a zero-thickness box, which the reference's strict single-ray box test rejects
    # (vecmath.h:1449-1466); tilt the ground slightly so it is visible to both ray classes
    ground = np.array([[-6, 0, -6], [6, -0.01, -6], [6, -0.03, 6], [-6, -0.02, 6]], np.float32)
    gidx = np.array([[0, 2, 1], [0, 3, 2]], np.uint32)
    nrm = vertex_normals_area(pos, idx)
    m0 = dict(indices=idx, positions=pos, normals=nrm, texcoords=None, prim_material=np.zeros(idx.shape[0], np.uint32),
              materials=[_mat((0.8, 0.7, 0.6))])
    m1 = dict(indices=gidx, positions=ground, normals=None, texcoords=None, prim_material=np.zeros(2, np.uint32),
              materials=[_mat((0.5, 0.5, 0.5))])
    return dict(meshes=[m0, m1], textures=[],
                light=dict(type="directional", dir=_normalize((0.2, 1.0, 0.2)), intensity=np.array((16.7, 15.6, 11.7), np.float32)),
                camera=dict(pos=(0.0, 1.25, 2.3), dir=(0.0, -0.12, -1.0), width=int(width), height=int(height)))
