"""CPU-only checks of the product's host side: the C-ABI library loads and exports every symbol include/prt_cuda.h
declares, the host BVH builder / light tables / camera agree with the oracle, and compute entry points fail loudly
without a GPU (no CPU fallback).  No compute calls are made."""
import os
import re
import numpy as np
import pytest

import prt_b200
from prt_b200 import capi, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _eq(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "prt_cuda.h")).read()
    declared = set(re.findall(r"\b(prt_cuda_\w+|prt_packet_seed)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = capi.lib()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/prt_cuda.h but not exported"
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)


def test_packet_seed_matches_definition():
    def fmix(h):
        h &= 0xffffffff; h ^= h >> 16; h = (h * 0x85ebca6b) & 0xffffffff; h ^= h >> 13; h = (h * 0xc2b2ae35) & 0xffffffff; h ^= h >> 16
        return h
    lib = capi.lib()
    for seed, pix, ppp, p in ((0, 0, 1, 0), (7, 123456, 32, 31), (0xffffffff, 2073599, 8, 3)):
        want = fmix((seed * 0x9E3779B9 + pix * ppp + p) & 0xffffffff) | 1
        assert lib.prt_packet_seed(seed, pix, ppp, p) == want


@pytest.mark.parametrize("make", [lambda: scenes.cornell_teapot(64, 36, light="infinite"), lambda: scenes.bunny_standin(64, 36)])
def test_host_builder_light_camera_match_oracle(oracle_mod, make):
    desc = make()
    ctx = prt_b200.load_scene(desc, device=-1)            # host-only context
    orc = oracle_mod.OracleScene(desc)
    assert ctx.radius() == orc.radius()
    assert _eq(ctx.camera_basis(), orc.camera_basis())
    for mi, m in enumerate(desc["meshes"]):
        npr = np.asarray(m["indices"]).shape[0]
        nodes, remap = ctx.bvh(mi, npr)
        on, od, orr = orc.bvh_dump(mi, npr)
        leaf = od[:, 1] != 15
        assert _eq(nodes[:, :7], on[:, :7])
        assert np.array_equal(nodes[leaf, 7], on[leaf, 7]) and np.array_equal(nodes[~leaf, 7] >> 26, on[~leaf, 7] >> 26)
        assert np.array_equal(remap, orr)
    if desc["light"]["type"] == "infinite":
        v, h = ctx.light_tables()
        ov, oh, _ = orc.infinite_light_tables()
        assert _eq(v, ov) and _eq(h, oh)
    ctx.close()


def test_builder_edge_cases_match_oracle(oracle_mod):
    rng = np.random.RandomState(11)
    for n in (1, 8, 9, 33, 1000):
        pos = rng.uniform(-1, 1, (3 * n, 3)).astype(np.float32)
        if n == 33:
            pos[:] = pos[:3].repeat(n, 0).reshape(3, n, 3).transpose(1, 0, 2).reshape(-1, 3)      # coincident centroids
        mesh = dict(indices=np.arange(3 * n, dtype=np.uint32).reshape(n, 3), positions=pos, normals=None, texcoords=None,
                    prim_material=np.zeros(n, np.uint32), materials=[dict(diffuse=(1, 1, 1))])
        ctx = prt_b200.load_scene(dict(meshes=[mesh]), device=-1)
        orc = oracle_mod.OracleScene(dict(meshes=[mesh]))
        nodes, remap = ctx.bvh(0, n)
        on, od, orr = orc.bvh_dump(0, n)
        assert _eq(nodes[:, :7], on[:, :7]) and np.array_equal(remap, orr), n
        ctx.close()


def test_no_cpu_fallback_and_argument_errors():
    desc = scenes.cornell_teapot(32, 32)
    ctx = prt_b200.load_scene(desc, device=-1)
    rays = np.zeros((8, 8), np.float32)
    for call in (lambda: ctx.intersect(rays, 0), lambda: ctx.occluded(rays, 0), lambda: ctx.render(8, 1),
                 lambda: ctx.trace_jobs([(prt_b200.PRT_JOB_INTERSECT, rays, 0, None), (prt_b200.PRT_JOB_OCCLUDED, rays, 0, None)])):
        with pytest.raises(prt_b200.PrtError, match="no CPU fallback"):
            call()
    with pytest.raises(prt_b200.PrtError):
        ctx.add_mesh(dict(indices=np.zeros((0, 3), np.uint32), positions=np.zeros((3, 3), np.float32), prim_material=np.zeros(0, np.uint32),
                          materials=[dict(diffuse=(1, 1, 1))]))
    with pytest.raises(prt_b200.PrtError):
        ctx.add_mesh(dict(indices=np.array([[0, 1, 2]], np.uint32), positions=np.eye(3, dtype=np.float32), prim_material=np.zeros(1, np.uint32),
                          materials=[dict(diffuse=(1, 1, 1), diffuse_tex=5)]))
    ctx.close()
    if capi.lib().prt_cuda_device_count() == 0:
        with pytest.raises(prt_b200.PrtError, match="no CPU fallback"):
            prt_b200.Context(0)
