"""GPU parity of the traversal kernels (prt_cuda_intersect / prt_cuda_occluded / prt_cuda_primary_dev through the
C ABI) against the oracle.  Bar: hit/miss, primId, meshId AND t,i,j,k bit-identical (north_star asks ≥ 99.99 %
id agreement and 1e-5 relative t; we assert exact equality and report the count)."""
import numpy as np
import pytest

import prt_b200
from prt_b200 import scenes

pytestmark = pytest.mark.gpu


def _assert_hits_equal(g, o, what):
    gu, ou = g.view(np.uint32), o.view(np.uint32)
    bad = (gu != ou).any(axis=1)
    assert not bad.any(), f"{what}: {int(bad.sum())}/{len(bad)} hit records differ; first {np.where(bad)[0][:5]}"


@pytest.fixture(scope="module")
def cornell(oracle_mod):
    W, H = 320, 180
    desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    yield desc, ctx, orc, W, H
    ctx.close()


def _secondary(orc, srays, shits, seed, maxT):
    t = shits[:, 0]
    hit = t != -1
    rng = np.random.RandomState(seed)
    m = int(hit.sum()) // 8 * 8
    org = (srays[hit, 0:3] + srays[hit, 3:6] * t[hit, None])[:m]
    d = rng.randn(m, 3).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((m, 8), np.float32)
    rays[:, 0:3] = org; rays[:, 3:6] = d; rays[:, 6] = maxT
    return rays


def test_known_answer_triangle(oracle_mod):
    """The reference's only hot-path KAT (tests/tests.cpp:113-126): tri (0,0,0),(1,0,0),(0,1,0), ray o=(0.4,0.4,-1)
    d=(0,0,1) -> t == 1.0f exactly — here through the real kernels.  A second, far-away triangle gives the mesh
    box thickness: with the KAT triangle alone the box is flat in z and the reference's strict single-ray box test
    (vecmath.h:1449-1466, entry == exit) rejects the ray at the root — that case is checked against the oracle."""
    def mesh(two):
        pos = [[0, 0, 0], [1, 0, 0], [0, 1, 0]] + ([[5, 5, 5], [6, 5, 5], [5, 6, 6]] if two else [])
        idx = [[0, 1, 2]] + ([[3, 4, 5]] if two else [])
        return dict(indices=np.array(idx, np.uint32), positions=np.array(pos, np.float32), normals=None, texcoords=None,
                    prim_material=np.zeros(len(idx), np.uint32), materials=[dict(diffuse=(1, 1, 1))])
    rays = np.tile(np.array([0.4, 0.4, -1, 0, 0, 1, 100.0, 0], np.float32), (8, 1))
    avg = np.array([[0, 0, 1]], np.float32)
    ctx = prt_b200.load_scene(dict(meshes=[mesh(True)]), 0)
    for mode in (0, 1):
        h = ctx.intersect(rays, mode, avg)
        assert (h[:, 0] == 1.0).all()
        assert np.allclose(h[:, 1:4], [0.2, 0.4, 0.4], atol=1e-6)
        assert (h.view(np.uint32)[:, 4:] == 0).all()
    assert ctx.occluded(rays, 0).all() and ctx.occluded(rays, 1).all()
    short = rays.copy(); short[:, 6] = 0.5
    assert not ctx.occluded(short, 0).any() and not ctx.occluded(short, 1).any()
    assert (ctx.intersect(short, 0)[:, 0] == -1).all()
    ctx.close()
    # flat-box case: GPU must reproduce whatever the reference rules give (single: miss, packet: hit)
    d1 = dict(meshes=[mesh(False)])
    ctx = prt_b200.load_scene(d1, 0)
    orc = oracle_mod.OracleScene(d1)
    _assert_hits_equal(ctx.intersect(rays, 0), orc.intersect_single(rays), "flat box single")
    _assert_hits_equal(ctx.intersect(rays, 1, avg), orc.intersect_packet(rays, avg), "flat box packet")
    assert (ctx.intersect(rays, 0)[:, 0] == -1).all() and (ctx.intersect(rays, 1, avg)[:, 0] == 1.0).all()
    assert np.array_equal(ctx.occluded(rays, 0), orc.occluded_single(rays))
    ctx.close()


def test_primary_packets(cornell):
    desc, ctx, orc, W, H = cornell
    rays, avg, ohits = orc.primary_packets((0, 0, W, H), 2, 5, nthreads=8)
    _assert_hits_equal(ctx.intersect(rays, 1, avg), ohits, "packet primary")


def test_primary_dev_generation(cornell):
    torch = pytest.importorskip("torch")
    desc, ctx, orc, W, H = cornell
    for mode in (1, 0):
        if mode == 1:
            rays, avg, ohits = orc.primary_packets((3, 5, 200, 100), 2, 11, nthreads=8)
        else:
            rays, ohits = orc.primary_single((3, 5, 200, 100), 3, 11, nthreads=8); avg = None
        n = rays.shape[0]
        d_r = torch.empty(n * 8, dtype=torch.float32, device="cuda")
        d_a = torch.empty(max(n // 8, 1) * 3, dtype=torch.float32, device="cuda")
        d_h = torch.empty(n * 6, dtype=torch.float32, device="cuda")
        ctx.primary_dev((3, 5, 200, 100), 2 if mode == 1 else 3, 11, mode, d_r.data_ptr(), d_a.data_ptr(), d_h.data_ptr())
        ctx.synchronize()
        assert np.array_equal(d_r.cpu().numpy().reshape(-1, 8).view(np.uint32), rays.view(np.uint32))
        if avg is not None:
            assert np.array_equal(d_a.cpu().numpy().reshape(-1, 3).view(np.uint32), avg.view(np.uint32))
        _assert_hits_equal(d_h.cpu().numpy().reshape(-1, 6), ohits, f"primary_dev mode {mode}")


def test_secondary_and_occlusion(cornell):
    desc, ctx, orc, W, H = cornell
    srays, shits = orc.primary_single((0, 0, W, H), 2, 9, nthreads=8)
    _assert_hits_equal(ctx.intersect(srays, 0), shits, "single primary")
    sec = _secondary(orc, srays, shits, 1, 2 * orc.radius())
    _assert_hits_equal(ctx.intersect(sec, 0), orc.intersect_single(sec, nthreads=8), "secondary single")
    avg = sec[:, 3:6].reshape(-1, 8, 3).mean(1)
    _assert_hits_equal(ctx.intersect(sec, 1, avg), orc.intersect_packet(sec, avg, nthreads=8), "secondary packet")
    assert np.array_equal(ctx.occluded(sec, 0), orc.occluded_single(sec, nthreads=8))
    full = np.full(sec.shape[0] // 8, 255, np.uint8)
    assert np.array_equal(ctx.occluded(sec, 1), orc.occluded_packet(sec, full, nthreads=8))
    # shadow rays built as path_tracer.cpp:209-212 toward the scene's light
    L = np.asarray(desc["light"]["dir"], np.float32)
    kfar = np.float32(2.0) * np.float32(orc.radius())
    sh = sec.copy()
    sh[:, 0:3] = sec[:, 0:3] + kfar * L
    sh[:, 3:6] = -L
    sh[:, 6] = kfar - np.float32(0.0008)
    assert np.array_equal(ctx.occluded(sh, 0), orc.occluded_single(sh, nthreads=8))
    assert np.array_equal(ctx.occluded(sh, 1), orc.occluded_packet(sh, full, nthreads=8))


def test_edge_cases(cornell):
    """Ties / near-edge enumeration: zero direction components (NaN slab lanes), origins on box planes,
    rays along triangle edges and through vertices, zero-length and empty batches, maxT shorter than the hit."""
    desc, ctx, orc, W, H = cornell
    rng = np.random.RandomState(3)
    n = 4096
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = np.round(rng.uniform(-1, 2, (n, 3)) * 4) / 4
    rays[:, 3:6] = np.eye(3, dtype=np.float32)[rng.randint(0, 3, n)] * rng.choice([-1.0, 1.0], (n, 1)).astype(np.float32)
    rays[:, 6] = 10.0
    # aim a block of rays exactly at Cornell vertices / edge midpoints
    pos = np.asarray(desc["meshes"][0]["positions"], np.float32)
    idx = np.asarray(desc["meshes"][0]["indices"])
    tgt = np.concatenate([pos[idx[:, 0]], 0.5 * (pos[idx[:, 0]] + pos[idx[:, 1]]), (pos[idx[:, 0]] + pos[idx[:, 1]] + pos[idx[:, 2]]) / 3])
    o = np.array([0.0, 0.965, 2.6], np.float32)
    d = tgt - o; d /= np.linalg.norm(d, axis=1, keepdims=True)
    k = d.shape[0]
    rays[:k, 0:3] = o; rays[:k, 3:6] = d
    rays[k:2 * k, 6] = 0.5
    avg = rays[:, 3:6].reshape(-1, 8, 3).mean(1)
    _assert_hits_equal(ctx.intersect(rays, 0), orc.intersect_single(rays), "edge single")
    _assert_hits_equal(ctx.intersect(rays, 1, avg), orc.intersect_packet(rays, avg), "edge packet")
    assert np.array_equal(ctx.occluded(rays, 0), orc.occluded_single(rays))
    assert np.array_equal(ctx.occluded(rays, 1), orc.occluded_packet(rays, np.full(n // 8, 255, np.uint8)))
    # empty and ragged batches
    assert ctx.intersect(np.zeros((0, 8), np.float32), 0).shape == (0, 6)
    assert ctx.occluded(np.zeros((0, 8), np.float32), 0).shape == (0,)
    _assert_hits_equal(ctx.intersect(rays[:37], 0), orc.intersect_single(rays[:37]), "ragged single")
    with pytest.raises(prt_b200.PrtError):
        ctx.intersect(rays[:37], 1, avg[:5])      # packet mode needs n % 8 == 0


def test_trace_jobs_batched_call(cornell):
    """prt_cuda_trace_jobs: several closest-hit / any-hit batches through one copy pipeline give exactly the per-call results
    (and the oracle's), including empty jobs, a ragged tail, packet-mode jobs and multi-chunk jobs; bad jobs are refused."""
    desc, ctx, orc, W, H = cornell
    srays, shits = orc.primary_single((0, 0, W, H), 6, 21, nthreads=8)      # 345 600 rays: two pipeline chunks
    sec = _secondary(orc, srays, shits, 5, 2 * orc.radius())
    avg = sec[:, 3:6].reshape(-1, 8, 3).mean(1)
    I, O = prt_b200.PRT_JOB_INTERSECT, prt_b200.PRT_JOB_OCCLUDED
    jobs = [(I, srays, 0, None), (O, sec, 0, None), (I, np.zeros((0, 8), np.float32), 0, None), (I, sec, 1, avg), (O, sec[:1037], 1, None),
            (O, np.zeros((0, 8), np.float32), 0, None), (I, srays[:37], 0, None)]
    out = ctx.trace_jobs(jobs)
    _assert_hits_equal(out[0], shits, "jobs[0] single primary")
    assert np.array_equal(out[1], orc.occluded_single(sec, nthreads=8))
    assert out[2].shape == (0, 6) and out[5].shape == (0,)
    _assert_hits_equal(out[3], orc.intersect_packet(sec, avg, nthreads=8), "jobs[3] packet secondary")
    assert np.array_equal(out[4], ctx.occluded(sec[:1037], 1))
    _assert_hits_equal(out[6], shits[:37], "jobs[6] ragged")
    import os
    os.environ["PRT_CHUNK_RAYS"] = "40000"                   # profiling override: many chunks per job, download groups, ragged last chunks
    try:
        out2 = ctx.trace_jobs(jobs)
    finally:
        del os.environ["PRT_CHUNK_RAYS"]
    for a, b in zip(out, out2):
        assert np.array_equal(a.view(np.uint8), b.view(np.uint8))
    assert ctx.trace_jobs([]) == []
    with pytest.raises(prt_b200.PrtError):
        ctx.trace_jobs([(I, sec[:37], 1, avg[:5])])          # packet mode needs n % 8 == 0
    with pytest.raises(prt_b200.PrtError):
        ctx.trace_jobs([(7, sec[:8], 0, None)])              # unknown kind


def test_bunny_scale_triangle_id_harness(oracle_mod):
    """Config C2 stand-in (81 920-triangle displaced icosphere): primary + occlusion rays, ids must match."""
    W, H = 480, 270
    desc = scenes.bunny_standin(W, H)
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    rays, avg, ohits = orc.primary_packets((0, 0, W, H), 1, 3, nthreads=8)
    g = ctx.intersect(rays, 1, avg)
    _assert_hits_equal(g, ohits, "C2 packet primary")
    srays, shits = orc.primary_single((0, 0, W, H), 1, 3, nthreads=8)
    _assert_hits_equal(ctx.intersect(srays, 0), shits, "C2 single primary")
    L = np.asarray(desc["light"]["dir"], np.float32)
    kfar = np.float32(2.0) * np.float32(orc.radius())
    hit = shits[:, 0] != -1
    m = int(hit.sum()) // 8 * 8
    p = (srays[hit, 0:3] + srays[hit, 3:6] * shits[hit, 0:1])[:m]
    sh = np.zeros((m, 8), np.float32)
    sh[:, 0:3] = p + kfar * L; sh[:, 3:6] = -L; sh[:, 6] = kfar - np.float32(0.0008)
    assert np.array_equal(ctx.occluded(sh, 0), orc.occluded_single(sh, nthreads=8))
    assert np.array_equal(ctx.occluded(sh, 1), orc.occluded_packet(sh, np.full(m // 8, 255, np.uint8), nthreads=8))
    ctx.close()


def test_probe_peaks_are_plausible():
    """The roofline denominators bench.py reports beside the HBM peak: L2 read bandwidth above HBM's, FP32 issue rate
    within the SM count x 4 schedulers x clock envelope."""
    ctx = prt_b200.Context(0)
    l2, issue = ctx.probe_peaks()
    ctx.close()
    assert 4000.0 < l2 < 60000.0, l2
    assert 300.0 < issue < 148 * 4 * 2.7, issue
