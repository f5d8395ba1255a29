"""GPU parity of the wavefront path tracer (prt_cuda_render through the C ABI) against the oracle — the scalar
restatement of PathTracer::TraceBlock/ComputeRadiance (path_tracer.cpp:17-308), itself pinned bit-exactly to the
reference (tests/test_oracle_vs_ref.py, tests/golden/).

Tolerance (FP32 images, north_star "per-pixel within a stated FP32 tolerance at equal spp and seed"): ZERO.
Every fp32 operation on the path is reproduced in the reference's order without FMA, and sinf / cosf / powf are
restated bit-exactly from the glibc the reference links (prt_shade.cuh, prt_powf.h), so every pixel must be
bit-identical; _check still reports the error magnitude when that fails.
"""
import os

import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

import prt_b200
from prt_b200 import scenes

pytestmark = pytest.mark.gpu


def _check(img, ref, min_identical=1.0):
    assert img.shape == ref.shape
    assert not np.isnan(img).any()
    same = (img.view(np.uint32) == ref.view(np.uint32)).all(axis=2)
    frac = float(same.mean())
    err = np.abs(img - ref).max(axis=2)
    tol = 2e-2 * np.maximum(1.0, np.abs(ref).max(axis=2))
    assert frac >= min_identical, f"only {frac:.5f} of pixels bit-identical; max err {err.max()}"
    assert (err <= tol).all(), f"{int((err > tol).sum())} pixels outside tolerance, max err {err.max()}"
    return frac


@pytest.mark.parametrize("name,kw,spp", [
    ("c1_diffuse_directional", dict(teapot_refl=scenes.DIFFUSE, light="directional"), 16),
    ("shipped_no_light_specular", dict(teapot_refl=scenes.SPECULAR, light=None), 16),
    ("c3_specular_infinite", dict(teapot_refl=scenes.SPECULAR, light="infinite"), 32),
    ("diffuse_infinite", dict(teapot_refl=scenes.DIFFUSE, light="infinite"), 16),
])
def test_render_matches_oracle(oracle_mod, name, kw, spp):
    W, H = 160, 120
    desc = scenes.cornell_teapot(W, H, **kw)
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    img = ctx.render(spp, 7)
    ref = orc.render(spp, 7, nthreads=8)
    frac = _check(img, ref)
    print(name, "bit-identical pixel fraction", frac)
    ctx.close()


@pytest.mark.parametrize("park", ["0", "1"])
def test_render_parked_trace_launches(oracle_mod, park):
    """Multi-mesh scenes: the two-launch trace (mesh 0, then the parked rays that go on to the teapot; prt_trace.cuh PHASE,
    on by default for batches >= 1 Mi packets) and the single launch both reproduce the oracle bit for bit — a third mesh
    (a copy of the Cornell box shifted behind the camera) makes parked rays cross more than one later mesh."""
    W, H = 160, 120
    desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.SPECULAR, light="infinite")
    extra = dict(desc["meshes"][0])
    extra["positions"] = np.asarray(extra["positions"], np.float32) + np.array([0.3, 0.0, 3.5], np.float32)
    desc["meshes"] = list(desc["meshes"]) + [extra]
    os.environ["PRT_WF_PARK"] = park
    try:
        ctx = prt_b200.load_scene(desc, 0)
        img = ctx.render(32, 11)
        launches = ctx.counters()["kernelLaunches"]
        ctx.close()
    finally:
        del os.environ["PRT_WF_PARK"]
    assert launches == (45 + 28 if park == "1" else 45)
    _check(img, oracle_mod.OracleScene(desc).render(32, 11, nthreads=8))


def test_render_region_and_spp_below_8(oracle_mod):
    desc = scenes.cornell_teapot(96, 64, teapot_refl=scenes.DIFFUSE, light="directional")
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    # samples < 8 renders black (samples/8 == 0, path_tracer.cpp:65)
    img = ctx.render(4, 1)
    assert (img == 0).all()
    # a sub-region leaves the rest of the caller's image untouched and equals the same pixels of a full render
    full = ctx.render(8, 3)
    part = np.full((64, 96, 3), -5.0, np.float32)
    ctx.render(8, 3, region=(10, 20, 50, 40), out=part)
    assert np.array_equal(part[20:40, 10:50], full[20:40, 10:50])
    assert (part[:20] == -5.0).all() and (part[:, 50:] == -5.0).all()
    ref = orc.render(8, 3, nthreads=8)
    _check(full, ref)
    ctx.close()


def test_render_deterministic_and_partition_invariant():
    """Per-packet seeding makes the image independent of how rows are partitioned (the multi-GPU contract, §8e)."""
    desc = scenes.cornell_teapot(64, 48, teapot_refl=scenes.SPECULAR, light="infinite")
    ctx = prt_b200.load_scene(desc, 0)
    a = ctx.render(16, 5)
    b = ctx.render(16, 5)
    assert np.array_equal(a, b)
    c = np.zeros_like(a)
    for y0 in range(0, 48, 12):
        ctx.render(16, 5, region=(0, y0, 64, y0 + 12), out=c)
    assert np.array_equal(a, c)
    ctx.close()


def test_render_bands_equal_full_frame():
    """Band-interleaved rendering (the multi-GPU partition) emulated as 3 ranks on one GPU: de-tiled result must be
    bit-identical to the single full-frame render (odd height exercises the zero padding rows)."""
    torch = pytest.importorskip("torch")
    from prt_b200 import multigpu
    W, H, band, world = 72, 45, 4, 3
    desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.SPECULAR, light="infinite")
    ctx = prt_b200.load_scene(desc, 0)
    full = ctx.render(16, 9)
    nb, nbpr = multigpu.band_layout(H, band, world)
    locs = []
    for r in range(world):
        buf = torch.full((nbpr * band, W, 3), -1.0, dtype=torch.float32, device="cuda")
        ctx.render_bands_dev(band, r, world, 16, 9, 1.0, buf.data_ptr())
        ctx.synchronize()
        locs.append(buf)
    img = multigpu.detile(torch.cat([b.reshape(-1) for b in locs]), world, band, W, H).cpu().numpy()
    assert np.array_equal(img, full)
    # padding rows are zero
    rows = multigpu.local_rows(H, band, world - 1, world)
    pad = [i for i, y in enumerate(rows) if y < 0]
    assert pad and (locs[world - 1][pad] == 0).all()
    ctx.close()


def test_textured_scene(oracle_mod):
    """C4 stand-in: alpha-tested traversal, bump taps and the de-gamma powf (material.cpp:24-28, restated bit-exactly
    as glibc evaluates it, prt_powf.h): hits, occlusion and the rendered image are all bit-identical to the oracle and to
    the committed reference image."""
    W, H = 96, 54
    desc = scenes.atrium(W, H, detail=0.25, tex_size=64)
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    assert ctx.alpha_flags == orc.alpha_flags
    rays, avg, ohits = orc.primary_packets((0, 0, W, H), 1, 3)
    g = ctx.intersect(rays, 1, avg)
    assert np.array_equal(g.view(np.uint32), ohits.view(np.uint32))
    srays, shits = orc.primary_single((0, 0, W, H), 2, 4)
    assert np.array_equal(ctx.intersect(srays, 0).view(np.uint32), shits.view(np.uint32))
    L = np.asarray(desc["light"]["dir"], np.float32)
    kfar = np.float32(2.0) * np.float32(orc.radius())
    hit = shits[:, 0] != -1
    m = int(hit.sum()) // 8 * 8
    sh = np.zeros((m, 8), np.float32)
    sh[:, 0:3] = (srays[hit, 0:3] + srays[hit, 3:6] * shits[hit, 0:1])[:m] + kfar * L
    sh[:, 3:6] = -L; sh[:, 6] = kfar - np.float32(0.0008)
    assert np.array_equal(ctx.occluded(sh, 0), orc.occluded_single(sh))
    assert np.array_equal(ctx.occluded(sh, 1), orc.occluded_packet(sh, np.full(m // 8, 255, np.uint8)))
    img = ctx.render(16, 2)
    ref = orc.render(16, 2, nthreads=8)
    assert np.array_equal(img.view(np.uint32), ref.view(np.uint32)), f"{int((img != ref).any(axis=2).sum())} pixels differ"
    g4 = np.load(os.path.join(G, "render_c4.npz"))
    ctx2 = prt_b200.load_scene(scenes.atrium(64, 48, detail=0.25, tex_size=64), 0)
    assert np.array_equal(ctx2.render(int(g4["spp"]), int(g4["seed"])).view(np.uint32), g4["image"].view(np.uint32))
    ctx2.close()
    ctx.close()
