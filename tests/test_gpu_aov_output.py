"""GPU parity of the G-buffer pass and the output stage (prt_cuda_gbuffer / prt_cuda_image_to_ppm /
prt_cuda_image_to_half_bgr through the C ABI) against the oracle and the committed reference outputs."""
import os
import numpy as np
import pytest

import prt_b200
from prt_b200 import scenes

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "aov.npz")


def _eq(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


def test_gbuffer_untextured_bit_identical(oracle_mod):
    gold = np.load(G)
    desc = scenes.cornell_teapot(64, 48, teapot_refl=scenes.DIFFUSE, light="directional")
    ctx = prt_b200.load_scene(desc, 0)
    for t in (0, 1, 2):
        img = ctx.gbuffer(t, 7, exposure=1.5 if t == 0 else 1.0)
        assert _eq(img, gold[f"gbuffer_c1_{t}"]), t
    # a sub-region leaves the rest of the caller's image untouched
    base = np.full((48, 64, 3), 9.0, np.float32)
    part = ctx.gbuffer(2, 7, region=(8, 4, 40, 20), out=base.copy())
    assert _eq(part[4:20, 8:40], gold["gbuffer_c1_2"][4:20, 8:40])
    mask = np.ones((48, 64), bool); mask[4:20, 8:40] = False
    assert (part[mask] == 9.0).all()
    ctx.close()


def test_gbuffer_full_size_matches_oracle(oracle_mod):
    desc = scenes.cornell_teapot(640, 360, light="infinite")
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    for t in (0, 2):
        assert _eq(ctx.gbuffer(t, 11), orc.gbuffer(t, 11, nthreads=8)), t
    ctx.close()


def test_gbuffer_textured(oracle_mod):
    """Normals (bump map taps) and the diffuse AOV (powf de-gamma, material.cpp:24-28) are bit-identical."""
    gold = np.load(G)
    ctx = prt_b200.load_scene(scenes.atrium(64, 48, detail=0.25, tex_size=64), 0)
    assert _eq(ctx.gbuffer(2, 7), gold["gbuffer_c4_2"])
    assert _eq(ctx.gbuffer(0, 7, exposure=1.5), gold["gbuffer_c4_0"])
    ctx.close()


def test_output_stage(oracle_mod):
    gold = np.load(G)
    ctx = prt_b200.Context(0)
    img = gold["image"]
    for tm, key in ((True, "ppm_tonemap"), (False, "ppm_linear")):
        assert np.array_equal(ctx.image_to_ppm(img, tm), gold[key])          # byte-identical to the reference's PPM payload
    planes = ctx.image_to_half_bgr(img)
    want = oracle_mod.half_bits(img)
    assert np.array_equal(planes[2], want[..., 0]) and np.array_equal(planes[1], want[..., 1]) and np.array_equal(planes[0], want[..., 2])
    finite = np.isfinite(gold["exr_rgba"][..., 0])
    assert np.array_equal(planes[2].view(np.float16).astype(np.float32)[finite], gold["exr_rgba"][..., 0][finite])
    # full-size image, random values over the whole half range
    rng = np.random.RandomState(9)
    big = (np.exp(rng.uniform(-20, 12, (1080, 1920, 3))) * rng.choice([-1, 1], (1080, 1920, 3))).astype(np.float32)
    assert np.array_equal(ctx.image_to_half_bgr(big)[2], oracle_mod.half_bits(big[..., 0]))
    for tm in (True, False):
        assert np.array_equal(ctx.image_to_ppm(big, tm), oracle_mod.image_ppm(big, tm, kind="oracle"))
    ctx.close()
