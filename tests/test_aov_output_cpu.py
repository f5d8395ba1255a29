"""CPU checks of the G-buffer / output-stage oracle: the restatement in oracle/prt_oracle.cpp
against the committed golden vectors (outputs of the unmodified reference, tests/golden/make_golden_aov.py) and, where
oracle/_ref exists, live against the reference.  Bit-exact."""
import os
import numpy as np
import pytest

from prt_b200 import scenes

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "aov.npz")


def _eq(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


@pytest.fixture(scope="module")
def gold():
    return np.load(G)


def test_oracle_gbuffer_matches_golden(oracle_mod, gold):
    for name, desc in (("c1", scenes.cornell_teapot(64, 48, teapot_refl=scenes.DIFFUSE, light="directional")),
                       ("c4", scenes.atrium(64, 48, detail=0.25, tex_size=64))):
        orc = oracle_mod.OracleScene(desc)
        for t in (0, 1, 2):
            img = orc.gbuffer(t, 7, exposure=1.5 if t == 0 else 1.0, nthreads=4)
            assert _eq(img, gold[f"gbuffer_{name}_{t}"]), (name, t)
    assert _eq(gold["gbuffer_c4_1"], gold["gbuffer_c4_2"])        # kMeshNormal and kNormal are one expression in the reference


def test_oracle_output_stage_matches_golden(oracle_mod, gold):
    img = gold["image"]
    assert np.array_equal(oracle_mod.image_ppm(img, True, kind="oracle"), gold["ppm_tonemap"])
    assert np.array_equal(oracle_mod.image_ppm(img, False, kind="oracle"), gold["ppm_linear"])
    assert _eq(oracle_mod.image_exr(img, kind="oracle"), gold["exr_rgba"])
    # the half values tinyexr wrote are what the restated conversion gives, including where it differs from IEEE
    # ties-to-even (tinyexr rounds half up: 1.00048828125 -> 0x3c01, numpy float16 gives 0x3c00)
    hb = oracle_mod.half_bits(img)
    finite = np.isfinite(gold["exr_rgba"][..., :3])
    assert np.array_equal(hb.view(np.float16).astype(np.float32)[finite], gold["exr_rgba"][..., :3][finite])
    assert oracle_mod.half_bits(np.array([1.00048828125], np.float32))[0] == 0x3c01


def test_oracle_output_stage_matches_reference_live(oracle_mod):
    if not oracle_mod.available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    rng = np.random.RandomState(5)
    img = (rng.rand(33, 47, 3).astype(np.float32) ** 4) * np.float32(20.0)
    img[1, :5, 0] = [0.0, -3.0, np.inf, 1e-7, 65519.9]
    for tm in (True, False):
        assert np.array_equal(oracle_mod.image_ppm(img, tm, kind="oracle"), oracle_mod.image_ppm(img, tm, kind="ref"))
    assert _eq(oracle_mod.image_exr(img, kind="oracle"), oracle_mod.image_exr(img, kind="ref"))
    desc = scenes.cornell_teapot(40, 24, light="infinite")
    ref, orc = oracle_mod.RefScene(desc), oracle_mod.OracleScene(desc)
    for t in (0, 2):
        assert _eq(ref.gbuffer(t, 3, nthreads=2), orc.gbuffer(t, 3, nthreads=2))
