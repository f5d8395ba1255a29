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


# ------------------------------------------------------------------------------------------------------------------
# Procedural stand-ins for Sponza (C4) and San-Miguel (C5): those downloads are unavailable offline (main.cpp:60,78).
# ------------------------------------------------------------------------------------------------------------------
def _grid(origin, du, dv, nu, nv, uv_scale=(1.0, 1.0), disp=None):
    """(nu x nv) quad patch -> positions, indices, uvs.  origin + s*du + t*dv, s,t in [0,1]."""
    s, t = np.meshgrid(np.linspace(0, 1, nu + 1), np.linspace(0, 1, nv + 1), indexing="xy")
    p = np.asarray(origin, np.float64)[None, None, :] + s[..., None] * np.asarray(du, np.float64) + t[..., None] * np.asarray(dv, np.float64)
    if disp is not None:
        n = np.cross(du, dv); n = n / np.linalg.norm(n)
        p = p + disp(s, t)[..., None] * n
    uv = np.stack([s * uv_scale[0], t * uv_scale[1]], -1)
    i = (np.arange(nv)[:, None] * (nu + 1) + np.arange(nu)[None, :]).reshape(-1)
    tri = np.concatenate([np.stack([i, i + 1, i + nu + 2], 1), np.stack([i, i + nu + 2, i + nu + 1], 1)], 0)
    return p.reshape(-1, 3), tri, uv.reshape(-1, 2)


def _tube(center_fn, radius_fn, nseg, nring, uv_scale=(1.0, 1.0)):
    """Swept circle: center_fn(t) -> (3,), frame from finite differences; returns positions, indices, uvs."""
    t = np.linspace(0, 1, nring + 1)
    a = np.linspace(0, 2 * np.pi, nseg + 1)
    c = np.array([center_fn(x) for x in t])
    tang = np.gradient(c, axis=0); tang /= np.linalg.norm(tang, axis=1, keepdims=True)
    ref = np.where(np.abs(tang[:, 1:2]) > 0.9, np.array([[1.0, 0, 0]]), np.array([[0, 1.0, 0]]))
    b1 = np.cross(tang, ref); b1 /= np.linalg.norm(b1, axis=1, keepdims=True)
    b2 = np.cross(tang, b1)
    r = np.array([radius_fn(x) for x in t])
    p = c[:, None, :] + r[:, None, None] * (np.cos(a)[None, :, None] * b1[:, None, :] + np.sin(a)[None, :, None] * b2[:, None, :])
    uv = np.stack(np.broadcast_arrays(a[None, :] / (2 * np.pi) * uv_scale[0], t[:, None] * uv_scale[1]), -1)
    i = (np.arange(nring)[:, None] * (nseg + 1) + np.arange(nseg)[None, :]).reshape(-1)
    tri = np.concatenate([np.stack([i, i + 1, i + nseg + 2], 1), np.stack([i, i + nseg + 2, i + nseg + 1], 1)], 0)
    return p.reshape(-1, 3), tri, uv.reshape(-1, 2)


def _noise_tex(rng, size, base, contrast, alpha_holes=False):
    """Seeded RGBA8 diffuse map: multi-octave value noise tinted by `base`; optional circular alpha holes."""
    img = np.zeros((size, size), np.float32)
    for o in range(2, 8):
        g = rng.rand(2 ** o + 1, 2 ** o + 1).astype(np.float32)
        x = np.linspace(0, 2 ** o, size, endpoint=False)
        xi = x.astype(int); xf = (x - xi).astype(np.float32)
        row = g[:, xi] * (1 - xf) + g[:, xi + 1] * xf
        img += (row[xi, :] * (1 - xf)[:, None] + row[xi + 1, :] * xf[:, None]) * (0.5 ** (o - 1))
    img = (img - img.min()) / (img.max() - img.min())
    rgb = np.clip(np.asarray(base, np.float32)[None, None, :] * (1 - contrast + contrast * img[..., None]), 0, 1)
    out = np.full((size, size, 4), 255, np.uint8)
    out[..., :3] = (rgb * 255).astype(np.uint8)
    if alpha_holes:
        yy, xx = np.mgrid[0:size, 0:size].astype(np.float32) / size
        cell = 8
        fx, fy = (xx * cell) % 1 - 0.5, (yy * cell) % 1 - 0.5
        d = np.sqrt(fx * fx + fy * fy)
        a = np.clip((d - 0.25) / 0.1, 0, 1)           # soft edge so bilinear alpha crosses the 127 threshold inside texels
        out[..., 3] = (a * 255).astype(np.uint8)
    return out


def _normal_map(rng, size, strength):
    """Seeded RGBA8 tangent-space normal map (what a Sponza 'map_bump' is); converted to a height map like
    Texture::load(bumpTexture=true) does (texture.cpp:185-200, 239-247)."""
    h = _noise_tex(rng, size, (1, 1, 1), 1.0)[..., 0].astype(np.float32) / 255.0
    gx = np.gradient(h, axis=1) * strength * size / 64.0
    gy = np.gradient(h, axis=0) * strength * size / 64.0
    n = np.stack([-gx, -gy, np.ones_like(h)], -1)
    n /= np.linalg.norm(n, axis=-1, keepdims=True)
    out = np.full((size, size, 4), 255, np.uint8)
    out[..., :3] = np.clip((n * 0.5 + 0.5) * 255, 0, 255).astype(np.uint8)
    return out


def convert_normal_to_bump(normal_rgba):
    """Texture::convertNormalToBump (texture.cpp:185-200) in fp32 numpy: bump = u8(255 * clamp(nz^4, 0, 1))."""
    k = np.float32(1.0 / 255.0)
    two, half = np.float32(2.0), np.float32(0.5)
    nx = two * (normal_rgba[..., 0].astype(np.float32) * k - half)
    ny = two * (normal_rgba[..., 1].astype(np.float32) * k - half)
    nz = two * (normal_rgba[..., 2].astype(np.float32) * k - half)
    inv = np.float32(1.0) / np.sqrt((nx * nx + ny * ny) + nz * nz, dtype=np.float32)
    z = inv * nz
    z4 = ((z * z) * z) * z
    return (np.float32(255.0) * np.clip(z4, np.float32(0), np.float32(1))).astype(np.uint8)


def atrium(width=1920, height=1080, seed=11, tex_size=1024, detail=1.0):
    """Config C4: Sponza-scale atrium, ~262k triangles at detail=1, one mesh with 12 textured materials
    (RGBA8 diffuse maps, two with alpha holes -> alphaTest; single-channel bump maps converted from normal maps),
    smooth vertex normals, directional light exactly as setupSponza (main.cpp:66), camera inside."""
    rng = np.random.RandomState(seed)
    textures, materials = [], []
    palette = [(0.75, 0.7, 0.62), (0.62, 0.33, 0.26), (0.4, 0.5, 0.35), (0.7, 0.65, 0.5), (0.35, 0.38, 0.55), (0.8, 0.78, 0.74),
               (0.55, 0.45, 0.3), (0.5, 0.2, 0.2), (0.3, 0.45, 0.5), (0.85, 0.8, 0.6), (0.9, 0.3, 0.2), (0.2, 0.6, 0.3)]
    for i, base in enumerate(palette):
        alpha = i >= 10
        d = _noise_tex(rng, tex_size, base, 0.6, alpha_holes=alpha)
        textures.append(dict(width=tex_size, height=tex_size, comp=4, texels=d))
        bump_id = -1
        if i % 2 == 0 and not alpha:
            b = convert_normal_to_bump(_normal_map(rng, tex_size // 2, 1.5))
            textures.append(dict(width=tex_size // 2, height=tex_size // 2, comp=1, texels=np.ascontiguousarray(b)))
            bump_id = len(textures) - 1
        materials.append(_mat((0.9, 0.9, 0.9), diffuse_tex=len(textures) - (2 if bump_id >= 0 else 1), bump_tex=bump_id))
    P, T, UV, M = [], [], [], []
    base = 0

    def add(p, t, uv, mat):
        nonlocal base
        P.append(p); T.append(t + base); UV.append(uv); M.append(np.full(t.shape[0], mat, np.uint32)); base += p.shape[0]

    k = lambda n: max(2, int(round(n * detail)))
    X, Y, Z = 20.0, 10.0, 8.0     # half-length, height, half-width
    bumpy = lambda amp, f: (lambda s, t: amp * np.sin(s * f) * np.cos(t * f * 0.7))
    add(*_grid((-X, 0, -Z), (2 * X, 0, 0), (0, 0, 2 * Z), k(160), k(104), (10, 4), bumpy(0.03, 40)), 0)                 # floor
    add(*_grid((-X, 0, -Z), (0, Y, 0), (2 * X, 0, 0), k(48), k(96), (3, 6), bumpy(0.05, 25)), 1)                        # wall -z
    add(*_grid((-X, 0, Z), (2 * X, 0, 0), (0, Y, 0), k(96), k(48), (6, 3), bumpy(0.05, 25)), 1)                         # wall +z
    add(*_grid((-X, 0, -Z), (0, 0, 2 * Z), (0, Y, 0), k(48), k(48), (3, 3)), 2)                                         # wall -x
    add(*_grid((X, 0, -Z), (0, Y, 0), (0, 0, 2 * Z), k(48), k(48), (3, 3)), 2)                                          # wall +x
    # ceiling with an open skylight strip in the middle (two side slabs)
    add(*_grid((-X, Y, -Z), (0, 0, Z * 0.6), (2 * X, 0, 0), k(24), k(96), (2, 8)), 3)
    add(*_grid((-X, Y, Z * 0.4), (0, 0, Z * 0.6), (2 * X, 0, 0), k(24), k(96), (2, 8)), 3)
    # colonnades: two rows of 12 columns
    for row, z in enumerate((-Z * 0.55, Z * 0.55)):
        for c in range(12):
            x = -X + (c + 0.5) * (2 * X / 12)
            add(*_tube(lambda t, x=x, z=z: np.array([x, t * Y * 0.8, z]), lambda t: 0.45 + 0.08 * np.cos(t * 30) * (t > 0.9) + 0.1 * (t < 0.05), k(48), k(40), (2, 6)),
                4 + (c + row) % 3)
    # arches between neighbouring columns
    for row, z in enumerate((-Z * 0.55, Z * 0.55)):
        for c in range(11):
            x0 = -X + (c + 0.5) * (2 * X / 12); x1 = x0 + 2 * X / 12
            add(*_tube(lambda t, x0=x0, x1=x1, z=z: np.array([x0 + (x1 - x0) * t, Y * 0.8 + 1.2 * np.sin(np.pi * t), z]), lambda t: 0.3, k(32), k(36), (1, 3)), 7 + c % 3)
    # hanging banners with alpha holes
    for b in range(20):
        x = -X + (b + 0.5) * (2 * X / 20); z = (-1) ** b * Z * 0.25
        add(*_grid((x - 0.8, Y * 0.35, z), (1.6, 0, 0.3), (0, Y * 0.4, 0), k(16), k(16), (1, 1), bumpy(0.08, 9)), 10 + b % 2)
    # decorations
    for s, (cx, cz) in enumerate(((-8.0, 0.0), (0.0, 0.0), (8.0, 0.5))):
        v, f = icosphere(5 if detail >= 1 else 3)
        r = 1.0 + 0.15 * np.sin(v[:, 0] * 7 + s) * np.cos(v[:, 1] * 5)
        p = v * r[:, None] * 1.2 + np.array([cx, 1.4, cz])
        uv = np.stack([np.arctan2(v[:, 2], v[:, 0]) / (2 * np.pi) + 0.5, np.arccos(np.clip(v[:, 1], -1, 1)) / np.pi], -1) * 3
        add(p, f, uv, 5 + s)
    pos = np.concatenate(P).astype(np.float32); idx = np.concatenate(T).astype(np.uint32)
    uv = np.concatenate(UV).astype(np.float32); pm = np.concatenate(M)
    mesh = dict(indices=idx, positions=pos, normals=vertex_normals_area(pos, idx), texcoords=uv, prim_material=pm, materials=materials)
    return dict(meshes=[mesh], textures=textures,
                light=dict(type="directional", dir=_normalize((0.05, 1.0, 0.1)), intensity=np.array((16.7, 15.6, 11.7), np.float32)),
                camera=dict(pos=(-15.0, 3.2, 0.4), dir=(1.0, -0.05, -0.05), width=int(width), height=int(height)))


def city(width=3840, height=2160, seed=5, scale=1.0):
    """Config C5: San-Miguel-scale scene, ~5M triangles at scale=1 (displaced terrain grid + 400 displaced blobs),
    untextured diffuse, directional light of setupSanMiguelLowPoly (main.cpp:84)."""
    rng = np.random.RandomState(seed)
    n = max(8, int(1200 * scale))
    terr = lambda s, t: 1.5 * np.sin(s * 9) * np.cos(t * 7) + 0.4 * np.sin(s * 41 + 1) * np.sin(t * 37) + 0.08 * np.sin(s * 190) * np.cos(t * 170)
    p, t, _ = _grid((-60, 0, -60), (120, 0, 0), (0, 0, 120), n, n, disp=lambda s, tt: -terr(s, tt))
    P, T = [p], [t]
    base = p.shape[0]
    nblob = max(1, int(400 * scale * scale))
    v, f = icosphere(4)
    for b in range(nblob):
        c = np.array([rng.uniform(-55, 55), 0, rng.uniform(-55, 55)])
        c[1] = terr((c[0] + 60) / 120, (c[2] + 60) / 120) + rng.uniform(0.5, 3.0)
        r = (1.0 + 0.25 * np.sin(v[:, 0] * rng.uniform(3, 9) + b) * np.cos(v[:, 2] * rng.uniform(3, 9))) * rng.uniform(0.6, 2.2)
        stretch = np.array([1.0, rng.uniform(0.7, 3.0), 1.0])
        P.append(v * r[:, None] * stretch + c); T.append(f + base); base += v.shape[0]
    pos = np.concatenate(P).astype(np.float32); idx = np.concatenate(T).astype(np.uint32)
    nmat = 8
    mats = [_mat(tuple(0.35 + 0.55 * rng.rand(3))) for _ in range(nmat)]
    pm = (np.arange(idx.shape[0]) // 4096 % nmat).astype(np.uint32)
    mesh = dict(indices=idx, positions=pos, normals=vertex_normals_area(pos, idx), texcoords=None, prim_material=pm, materials=mats)
    return dict(meshes=[mesh], textures=[],
                light=dict(type="directional", dir=_normalize((0.2, 1.0, 0.2)), intensity=np.array((16.7, 15.6, 11.7), np.float32)),
                camera=dict(pos=(-40.0, 9.0, -38.0), dir=(0.7, -0.18, 0.68), width=int(width), height=int(height)))
