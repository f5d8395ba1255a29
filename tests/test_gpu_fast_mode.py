"""The opt-in fast traversal mode (prt_cuda_set_option, prt_b200/csrc/prt_wide.cuh: 8-wide nodes with quantised boxes over a
tree with smaller leaves) against the exact mode, which is bit-identical to the reference (every other GPU test).

What the fast mode promises: the same closest t, found with the same triangle arithmetic — so wherever both modes report the
same triangle the whole record is bit-identical — and never a miss where the exact mode hits.  What it does not: which of two
triangles at exactly the same t is reported (a tie), and hits the reference's own box tests reject at grazing incidence.
Every differing ray is enumerated and put in one of those classes; anything else fails the test."""
import numpy as np
import pytest

import prt_b200
from prt_b200 import capi, scenes

pytestmark = pytest.mark.gpu
SINGLE = prt_b200.PRT_MODE_SINGLE


def _rays_for(ctx, orc_desc, W, H, seed):
    """single-rule primary rays (device-generated) + incoherent secondary rays from their hits"""
    import torch
    n = W * H
    d_r = torch.zeros((n, 8), dtype=torch.float32, device="cuda"); d_h = torch.zeros((n, 6), dtype=torch.float32, device="cuda")
    ctx.primary_dev((0, 0, W, H), 1, seed, SINGLE, d_r.data_ptr(), None, d_h.data_ptr()); ctx.synchronize()
    rays, hits = d_r.cpu().numpy(), d_h.cpu().numpy()
    t = hits[:, 0]; hit = t != -1
    rng = np.random.RandomState(seed)
    d = rng.randn(int(hit.sum()), 3).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
    sec = np.zeros((int(hit.sum()), 8), np.float32)
    sec[:, 0:3] = rays[hit, 0:3] + rays[hit, 3:6] * t[hit, None]; sec[:, 3:6] = d; sec[:, 6] = 1e5
    return np.concatenate([rays, sec])


def _moller_trumbore_f64(desc, rays, recs):
    """t of each ray against the triangle its record names, in fp64 (nan where the ray misses that triangle)."""
    w = recs.view(np.uint32)
    t = np.full(len(recs), np.nan)
    for m, mesh in enumerate(desc["meshes"]):
        sel = np.where(w[:, 5] == m)[0]
        if not len(sel):
            continue
        idx = np.asarray(mesh["indices"]).reshape(-1, 3)[w[sel, 4]]
        P = np.asarray(mesh["positions"], np.float64)
        v0, e1, e2 = P[idx[:, 0]], P[idx[:, 1]] - P[idx[:, 0]], P[idx[:, 2]] - P[idx[:, 0]]
        o, d = rays[sel, 0:3].astype(np.float64), rays[sel, 3:6].astype(np.float64)
        pv = np.cross(d, e2); det = (e1 * pv).sum(1)
        with np.errstate(divide="ignore", invalid="ignore"):
            tv = o - v0; u = (tv * pv).sum(1) / det
            qv = np.cross(tv, e1); v = (d * qv).sum(1) / det
            tt = (e2 * qv).sum(1) / det
        ok = (det != 0) & (u >= 0) & (v >= 0) & (u + v <= 1) & (tt > 0)
        t[sel] = np.where(ok, tt, np.nan)
    return t


def classify(exact, fast, rays, desc):
    """-> dict of counts; raises on anything the fast mode must never do.

    ties      : same t bits, another triangle (coplanar overlapping triangles, shared edges)
    grazing   : the fast mode reports a closer hit (or a hit at all) that the reference's strict box tests reject
    artefacts : the exact mode's record is farther / present where the fast one is not, AND fp64 Moller-Trumbore says the
                exact record is not a real intersection of that ray with that triangle at that t (the reference's sheared
                test is ill-conditioned for a few rays: dominant axis with a tiny component, origin a rounding below a wall).
                A ray whose exact record IS a real hit and which the fast mode misses is "lost": never allowed."""
    e, f = exact.view(np.uint32), fast.view(np.uint32)
    diff = (e != f).any(axis=1)
    et, ft = exact[:, 0], fast[:, 0]
    emiss, fmiss = et == -1, ft == -1
    tie = diff & ~emiss & ~fmiss & (e[:, 0] == f[:, 0])
    closer = diff & ~fmiss & (emiss | (ft < et))
    farther = diff & ~emiss & (fmiss | (ft > et))
    artefact = np.zeros_like(diff)
    if farther.any():
        k = np.where(farther)[0]
        t64 = _moller_trumbore_f64(desc, rays[k], exact[k])
        real = np.isfinite(t64) & (np.abs(t64 - et[k].astype(np.float64)) <= 1e-5 * np.maximum(np.abs(t64), 1e-3))
        artefact[k[~real]] = True
        lost = k[real]
        assert not len(lost), f"fast mode lost {len(lost)} hits the exact mode finds; first {lost[:5]}"
    assert (diff == (tie | closer | artefact)).all()
    n = len(diff)
    return dict(rays=n, differing=int(diff.sum()), ties=int(tie.sum()), grazing=int(closer.sum()), reference_artefacts=int(artefact.sum()),
                agreement=1.0 - diff.mean(), agreement_without_ties=1.0 - (diff & ~tie).mean(), artefact_rays=np.where(artefact)[0])


@pytest.mark.parametrize("name,make,W,H", [
    ("c2 bunny stand-in", lambda W, H: scenes.bunny_standin(W, H), 640, 360),
    ("c1/c3 cornell + teapot", lambda W, H: scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional"), 480, 270),
    ("c4 atrium (textures, alpha)", lambda W, H: scenes.atrium(W, H, tex_size=128, detail=0.5), 480, 270),
    ("c5 city at a third of the detail", lambda W, H: scenes.city(W, H, scale=0.3), 640, 360),
])
def test_fast_mode_closest_and_any_hit_against_exact(name, make, W, H):
    desc = make(W, H)
    ctx = prt_b200.load_scene(desc, 0, fast=True)
    rays = _rays_for(ctx, desc, W, H, 3)
    exact = ctx.intersect(rays, SINGLE)
    srays = rays.copy(); srays[:, 6] = np.where(exact[:, 0] != -1, exact[:, 0] * np.float32(1.5), np.float32(50.0))
    occ_exact = ctx.occluded(srays, SINGLE)
    ctx.set_traversal(True)
    fast = ctx.intersect(rays, SINGLE)
    occ_fast = ctx.occluded(srays, SINGLE)
    ctx.set_traversal(False)
    assert np.array_equal(ctx.intersect(rays, SINGLE).view(np.uint32), exact.view(np.uint32)), "switching back does not restore the exact mode"
    c = classify(exact, fast, rays, desc)
    print(f"\\n{name}: { {k: v for k, v in c.items() if k != 'artefact_rays'} }")
    # north_star: >= 99.99 % with the differing rays enumerated.  Ties are counted apart: the same t found on another of several
    # triangles lying in one plane (the Cornell box's floor under the short box's base: 67 of 205 627 rays) is the same surface
    # point; everything else (grazing, reference artefacts) must stay under 1e-4, and the total under 1e-3.
    assert c["agreement_without_ties"] >= 0.9999 and c["agreement"] >= 0.999, c
    # any hit: a boolean over a superset of boxes — the fast mode may find an occluder at grazing incidence, never lose one
    lost = np.where((occ_exact == 1) & (occ_fast == 0))[0]      # allowed only where the exact occluder is itself a reference artefact
    assert set(lost.tolist()) <= set(c["artefact_rays"].tolist()), f"{len(lost)} occluded rays reported free in fast mode"
    extra = int(((occ_exact == 0) & (occ_fast == 1)).sum())
    print(f"{name}: any-hit rays {len(occ_exact)}, occluded only in fast mode (grazing) {extra}")
    assert extra <= 1e-4 * len(occ_exact)
    ctx.close()


@pytest.mark.parametrize("name,make,W,H", [
    ("c2 bunny stand-in", lambda W, H: scenes.bunny_standin(W, H), 320, 180),
    ("c1/c3 cornell + teapot", lambda W, H: scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional"), 240, 136),
])
def test_fast_mode_packet_rule_batches(name, make, W, H):
    """Packet-rule rays (8 per pixel, abs dominant axis, SoA root test; any hit sees the triangles under flat boxes) through
    the wide kernels: same classes of differences as the single-rule rays, nothing else."""
    import torch
    PACKET = prt_b200.PRT_MODE_PACKET
    desc = make(W, H)
    ctx = prt_b200.load_scene(desc, 0, fast=True)
    n = W * H * 8
    d_r = torch.zeros((n, 8), dtype=torch.float32, device="cuda"); d_a = torch.zeros((n // 8, 3), dtype=torch.float32, device="cuda")
    d_h = torch.zeros((n, 6), dtype=torch.float32, device="cuda")
    ctx.primary_dev((0, 0, W, H), 1, 5, PACKET, d_r.data_ptr(), d_a.data_ptr(), d_h.data_ptr()); ctx.synchronize()
    rays, avg, exact = d_r.cpu().numpy(), d_a.cpu().numpy(), d_h.cpu().numpy()
    assert np.array_equal(ctx.intersect(rays, PACKET, avg).view(np.uint32), exact.view(np.uint32))
    srays = rays.copy(); srays[:, 6] = np.where(exact[:, 0] != -1, exact[:, 0] * np.float32(1.5), np.float32(50.0))
    occ_exact = ctx.occluded(srays, PACKET)
    ctx.set_traversal(True)
    fast = ctx.intersect(rays, PACKET, avg)
    occ_fast = ctx.occluded(srays, PACKET)
    c = classify(exact, fast, rays, desc)
    print(f"\\n{name} (packet rules): {c}")
    assert c["agreement_without_ties"] >= 0.9999 and c["agreement"] >= 0.999, c
    lost = np.where((occ_exact == 1) & (occ_fast == 0))[0]
    assert set(lost.tolist()) <= set(c["artefact_rays"].tolist()), f"{len(lost)} occluded rays reported free in fast mode"
    assert int(((occ_exact == 0) & (occ_fast == 1)).sum()) <= 1e-4 * len(occ_exact)
    assert (occ_exact == 1).sum() > 1000
    ctx.close()


def test_fast_mode_on_the_full_size_city():
    """C5 at its full 4.9 M triangles (the scene the mode exists for): every differing ray in a known class, none lost."""
    W, H = 960, 540
    desc = scenes.city(W, H)
    ctx = prt_b200.load_scene(desc, 0, fast=True)
    rays = _rays_for(ctx, desc, W, H, 11)
    exact = ctx.intersect(rays, SINGLE)
    srays = rays.copy(); srays[:, 6] = np.where(exact[:, 0] != -1, exact[:, 0] * np.float32(1.5), np.float32(500.0))
    occ_exact = ctx.occluded(srays, SINGLE)
    ctx.set_traversal(True)
    fast = ctx.intersect(rays, SINGLE)
    occ_fast = ctx.occluded(srays, SINGLE)
    c = classify(exact, fast, rays, desc)
    print(f"\\nc5 full size: {c}")
    assert c["agreement_without_ties"] >= 0.9999 and c["agreement"] >= 0.999, c
    # an any-hit ray may turn "free" only where the exact mode's occluder is itself a reference artefact (the same bogus triangle
    # record its closest-hit pass reports for that ray)
    lost = np.where((occ_exact == 1) & (occ_fast == 0))[0]
    print(f"any hit: occluded only in exact mode {len(lost)} (rays {lost[:8]}), reference-artefact rays {c['artefact_rays'][:8]}")
    assert set(lost.tolist()) <= set(c["artefact_rays"].tolist())
    assert int(((occ_exact == 0) & (occ_fast == 1)).sum()) <= 1e-4 * len(occ_exact)
    assert (exact[:, 0] != -1).mean() > 0.5
    ctx.close()


def test_fast_mode_render_is_close_to_exact():
    """The images differ only where a path met a tie or a grazing case: a handful of pixels."""
    W, H = 320, 180
    for desc, spp in ((scenes.cornell_teapot(W, H, teapot_refl=scenes.SPECULAR, light="infinite"), 16), (scenes.city(W, H, scale=0.2), 16)):
        ctx = prt_b200.load_scene(desc, 0, fast=True)
        exact = ctx.render(spp, 4)
        ctx.set_traversal(True)
        fast = ctx.render(spp, 4)
        differing = (exact != fast).any(axis=2).mean()
        rel = np.abs(fast - exact).sum() / max(np.abs(exact).sum(), 1e-9)
        print(f"\\npixels differing {differing:.5f}, relative L1 difference {rel:.2e}")
        assert differing < 0.02 and rel < 1e-2
        assert exact.max() > 0
        ctx.close()


def test_fast_mode_option_rules():
    desc = scenes.cornell_teapot(32, 32)
    ctx = prt_b200.load_scene(desc, 0)
    with pytest.raises(prt_b200.PrtError, match="BUILD_FAST"):
        ctx.set_traversal(True)                                   # the structure was not built
    with pytest.raises(prt_b200.PrtError, match="before the first mesh"):
        ctx.set_option(capi.PRT_OPTION_BUILD_FAST_TRAVERSAL, 1)
    assert ctx.get_option(capi.PRT_OPTION_TRAVERSAL) == capi.PRT_TRAVERSAL_EXACT
    ctx.close()


@pytest.mark.skipif(capi.lib().prt_cuda_device_count() < 2, reason="needs 2 GPUs")
def test_fast_mode_on_a_multi_gpu_context():
    """The option reaches every GPU of a prt_cuda_create_multi context: the gathered fast-mode frame is the 1-GPU fast-mode frame."""
    desc = scenes.cornell_teapot(200, 117, teapot_refl=scenes.SPECULAR, light="infinite")
    one = prt_b200.load_scene(desc, 0, fast=True)
    many = prt_b200.load_scene(desc, [0, 1], fast=True)
    exact = one.render(16, 7)
    one.set_traversal(True); many.set_traversal(True)
    a, b = one.render(16, 7), many.render(16, 7)
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    assert (a != exact).any(axis=2).mean() < 0.02
    rays = _rays_for(one, desc, 200, 117, 3)
    assert np.array_equal(one.intersect(rays, SINGLE).view(np.uint32), many.intersect(rays, SINGLE).view(np.uint32))
    one.close(); many.close()
