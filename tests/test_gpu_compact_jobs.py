"""Compact ray layouts of prt_cuda_trace_jobs (include/prt_cuda.h, prt_ray_layout): rays sent as 12 bytes (shared origin /
shared direction) or not at all (camera rays generated on the device, shadow rays rebuilt from an earlier job's hits) must
give results bit-identical to the same rays sent as 32-byte records — and to the oracle.  Also the page-locked allocation
helper and the pageable-buffer path of the host-buffer calls."""
import ctypes as C
import numpy as np
import pytest

import prt_b200
from prt_b200 import capi, scenes

pytestmark = pytest.mark.gpu

INT, OCC = prt_b200.PRT_JOB_INTERSECT, prt_b200.PRT_JOB_OCCLUDED
SINGLE, PACKET = prt_b200.PRT_MODE_SINGLE, prt_b200.PRT_MODE_PACKET


def _eq(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


@pytest.fixture(scope="module")
def bunny(oracle_mod):
    W, H = 480, 270
    desc = scenes.bunny_standin(W, H)
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    yield desc, ctx, orc, W, H
    ctx.close()


def _shadow_rays(rays, hits, L, radius):
    """path_tracer.cpp:236-238 in numpy fp32 (unfused): org = (t*dir + org) + kFar*L, dir = -L, maxT = kFar - 0.0008."""
    L = np.asarray(L, np.float32)
    kfar = np.float32(2.0) * np.float32(radius)
    hit = hits[:, 0] != -1
    sh = np.zeros((int(hit.sum()), 8), np.float32)
    sh[:, 0:3] = (hits[hit, 0:1] * rays[hit, 3:6] + rays[hit, 0:3]) + kfar * L
    sh[:, 3:6] = -L
    sh[:, 6] = kfar - np.float32(0.0008)
    return sh, hit


@pytest.mark.parametrize("mode", [SINGLE, PACKET])
def test_primary_job_matches_oracle_and_full_layout(bunny, mode):
    desc, ctx, orc, W, H = bunny
    region = (3, 5, W - 7, H - 2)
    if mode == SINGLE:
        rays, ohits = orc.primary_single(region, 2, 9)
        avg = None
    else:
        rays, avg, ohits = orc.primary_packets(region, 1, 9)
    full, = ctx.trace_jobs([dict(kind=INT, mode=mode, rays=rays, avg_dir=avg)])
    prim, = ctx.trace_jobs([dict(kind=INT, mode=mode, primary=(region, 2 if mode == SINGLE else 1, 9))])
    assert _eq(full, ohits), "32-byte path differs from the oracle"
    assert _eq(prim, full), "device-generated primary rays give different hit records"


def test_shared_origin_and_shared_direction_layouts(bunny):
    desc, ctx, orc, W, H = bunny
    region = (0, 0, W, H)
    rays, ohits = orc.primary_single(region, 1, 4)
    L = desc["light"]["dir"]
    sh, hitmask = _shadow_rays(rays, ohits, L, ctx.radius())
    full_h, full_o = ctx.trace_jobs([dict(kind=INT, mode=SINGLE, rays=rays), dict(kind=OCC, mode=SINGLE, rays=sh)])
    cam = desc["camera"]["pos"]
    comp_h, comp_o = ctx.trace_jobs([
        dict(kind=INT, mode=SINGLE, dirs=rays[:, 3:6], origin=cam, max_t=rays[0, 6]),
        dict(kind=OCC, mode=SINGLE, origins=sh[:, 0:3], direction=sh[0, 3:6], max_t=sh[0, 6])])
    assert _eq(full_h, ohits) and np.array_equal(full_o, orc.occluded_single(sh))
    assert _eq(comp_h, full_h), "shared-origin layout changes hit records"
    assert np.array_equal(comp_o, full_o), "shared-direction layout changes occlusion bits"
    # packet rules through the shared-origin layout
    prays, pavg, phits = orc.primary_packets((0, 0, W, 64), 1, 4)
    got, = ctx.trace_jobs([dict(kind=INT, mode=PACKET, dirs=prays[:, 3:6], origin=cam, max_t=prays[0, 6], avg_dir=pavg)])
    assert _eq(got, phits)


@pytest.mark.parametrize("mode", [SINGLE, PACKET])
def test_shadow_rays_from_hits_job(bunny, mode):
    """job 1 builds its rays on the device from job 0's hits; result must equal the explicit shadow-ray batch."""
    desc, ctx, orc, W, H = bunny
    region = (0, 0, W, H) if mode == SINGLE else (0, 0, W, 96)
    L = np.asarray(desc["light"]["dir"], np.float32)
    if mode == SINGLE:
        rays, ohits = orc.primary_single(region, 1, 21)
    else:
        rays, _, ohits = orc.primary_packets(region, 1, 21)
    hits, occ = ctx.trace_jobs([dict(kind=INT, mode=mode, primary=(region, 1, 21)), dict(kind=OCC, mode=SINGLE, from_job=0, light_dir=L)])
    assert _eq(hits, ohits)
    sh, hitmask = _shadow_rays(rays, ohits, L, ctx.radius())
    want = np.full(rays.shape[0], 2, np.uint8)
    want[hitmask] = orc.occluded_single(sh)
    assert np.array_equal(occ, want), f"{int((occ != want).sum())} occlusion results differ"
    assert (occ[~hitmask] == 2).all() and hitmask.sum() > 1000


def test_argument_errors_of_compact_jobs(bunny):
    desc, ctx, orc, W, H = bunny
    with pytest.raises(prt_b200.PrtError, match="fromJob"):
        ctx.trace_jobs([dict(kind=OCC, mode=SINGLE, from_job=0, light_dir=(0, 1, 0))])
    with pytest.raises(prt_b200.PrtError, match="closest-hit"):
        ctx.trace_jobs([dict(kind=OCC, mode=SINGLE, primary=((0, 0, 8, 8), 1, 1))])
    with pytest.raises(prt_b200.PrtError, match="bad region"):
        ctx.trace_jobs([dict(kind=INT, mode=SINGLE, primary=((0, 0, W + 1, 8), 1, 1))])


def test_pinned_registered_and_pageable_buffers_agree(bunny):
    """prt_cuda_host_alloc / prt_cuda_host_register buffers and plain pageable numpy memory give the same records."""
    desc, ctx, orc, W, H = bunny
    rays, ohits = orc.primary_single((0, 0, W, H), 1, 2)
    n = rays.shape[0]
    L = ctx.L
    pr, ph = capi.PinnedBuffer((n, 8), np.float32), capi.PinnedBuffer((n, 6), np.float32)
    pr.array[:] = rays
    ctx.check(L.prt_cuda_intersect(ctx.h, n, pr.ptr, None, SINGLE, ph.ptr))
    assert _eq(ph.array, ohits)
    pr.free(); ph.free()
    reg_r, reg_h = rays.copy(), np.zeros((n, 6), np.float32)
    assert L.prt_cuda_host_register(reg_r.ctypes.data, reg_r.nbytes) == 0 and L.prt_cuda_host_register(reg_h.ctypes.data, reg_h.nbytes) == 0
    ctx.check(L.prt_cuda_intersect(ctx.h, n, reg_r.ctypes.data, None, SINGLE, reg_h.ctypes.data))
    assert L.prt_cuda_host_unregister(reg_r.ctypes.data) == 0 and L.prt_cuda_host_unregister(reg_h.ctypes.data) == 0
    assert _eq(reg_h, ohits)
    assert _eq(ctx.intersect(rays, SINGLE), ohits)          # pageable


def test_primary_dev_outputs_match_oracle(bunny):
    """prt_cuda_primary_dev now runs through the dynamic-fetch kernel: rays, avgDir and hits, both rule sets."""
    import torch
    desc, ctx, orc, W, H = bunny
    region = (16, 8, 272, 200)
    n = (region[2] - region[0]) * (region[3] - region[1])
    rays, avg, hits = orc.primary_packets(region, 1, 5)
    d_r = torch.zeros((n * 8, 8), dtype=torch.float32, device="cuda"); d_a = torch.zeros((n, 3), dtype=torch.float32, device="cuda")
    d_h = torch.zeros((n * 8, 6), dtype=torch.float32, device="cuda")
    ctx.primary_dev(region, 1, 5, PACKET, d_r.data_ptr(), d_a.data_ptr(), d_h.data_ptr()); ctx.synchronize()
    assert _eq(d_r.cpu().numpy()[:, :7], rays[:, :7]) and _eq(d_a.cpu().numpy(), avg) and _eq(d_h.cpu().numpy(), hits)
    srays, shits = orc.primary_single(region, 3, 5)
    d_r = torch.zeros((n * 3, 8), dtype=torch.float32, device="cuda"); d_h = torch.zeros((n * 3, 6), dtype=torch.float32, device="cuda")
    ctx.primary_dev(region, 3, 5, SINGLE, d_r.data_ptr(), None, d_h.data_ptr()); ctx.synchronize()
    assert _eq(d_r.cpu().numpy()[:, :7], srays[:, :7]) and _eq(d_h.cpu().numpy(), shits)
    # rays only (no trace)
    d_r.zero_()
    ctx.primary_dev(region, 3, 5, SINGLE, d_r.data_ptr(), None, None); ctx.synchronize()
    assert _eq(d_r.cpu().numpy()[:, :7], srays[:, :7])
