"""prt_cuda_primary_shadow_noalloc (include/prt_cuda.h): the entry point that allocates nothing — its scratch (work counters,
intermediate hit records) is a static __device__ array of the library — against the existing entry point that does the same
work with the context's own buffers: prt_cuda_trace_jobs with a PRT_RAYS_PRIMARY closest-hit job and a PRT_RAYS_FROM_HITS
any-hit job.  Byte work: the outputs must be identical (the tolerance of every traversal test here is zero).

Collected last on purpose; runs alone in a fresh process as well as after the other tests."""
import numpy as np
import pytest

import prt_b200
from prt_b200 import capi, scenes

pytestmark = pytest.mark.gpu
INT, OCC, SINGLE = prt_b200.PRT_JOB_INTERSECT, prt_b200.PRT_JOB_OCCLUDED, prt_b200.PRT_MODE_SINGLE


def _device_bytes_allocated(torch):
    free, total = torch.cuda.mem_get_info()
    return total - free


def test_noalloc_entry_point_matches_trace_jobs_twice_in_a_row():
    import torch
    W, H = 480, 270
    desc = scenes.bunny_standin(W, H)
    L = np.asarray(desc["light"]["dir"], np.float32)
    ctx = prt_b200.load_scene(desc, 0)
    calls = [((0, 0, W, H), 1, 7), ((16, 8, 272, 200), 3, 11)]                  # (region, samples per pixel, seed): two calls in a row
    want = []
    for region, pp, seed in calls:                                             # existing entry point (also commits the scene)
        hits, occ = ctx.trace_jobs([dict(kind=INT, mode=SINGLE, primary=(region, pp, seed)), dict(kind=OCC, mode=SINGLE, from_job=0, light_dir=L)])
        want.append(occ)
        assert (occ == 1).sum() > 1000 and (occ == 0).sum() > 1000 and (occ == 2).sum() > 100
    outs = [torch.full((len(w),), 255, dtype=torch.uint8, device="cuda") for w in want]
    warm = torch.zeros(64, dtype=torch.uint8, device="cuda")
    ctx.primary_shadow_noalloc((0, 0, 8, 8), 1, 1, L, warm.data_ptr()); ctx.synchronize()   # first use: the driver loads what it loads lazily
    torch.cuda.synchronize()
    before = _device_bytes_allocated(torch)
    for (region, pp, seed), d_out in zip(calls, outs):                         # back to back on the context's stream, no sync in between
        ctx.primary_shadow_noalloc(region, pp, seed, L, d_out.data_ptr())
    ctx.synchronize()
    assert _device_bytes_allocated(torch) == before, "device memory in use changed across the no-alloc calls"
    for k, (d_out, w) in enumerate(zip(outs, want)):
        got = d_out.cpu().numpy()
        assert np.array_equal(got, w), f"call {k}: {int((got != w).sum())} of {len(w)} bytes differ from prt_cuda_trace_jobs"
    # the first call again: the shared scratch keeps the previous call's contents, the result must not depend on them
    again = torch.full((len(want[0]),), 255, dtype=torch.uint8, device="cuda")
    ctx.primary_shadow_noalloc(calls[0][0], calls[0][1], calls[0][2], L, again.data_ptr()); ctx.synchronize()
    assert np.array_equal(again.cpu().numpy(), want[0])
    ctx.close()


def test_noalloc_entry_point_rejects_what_it_cannot_do():
    import torch
    W, H = 1024, 512
    desc = scenes.cornell_teapot(W, H, teapot_refl=scenes.DIFFUSE, light="directional")
    ctx = prt_b200.load_scene(desc, 0)
    d_out = torch.zeros(capi.PRT_NOALLOC_MAX_RAYS, dtype=torch.uint8, device="cuda")
    L = (0.2, 0.4, 1.0)
    with pytest.raises(prt_b200.PrtError, match="uncommitted"):               # committing allocates: refused
        ctx.primary_shadow_noalloc((0, 0, 64, 64), 1, 1, L, d_out.data_ptr())
    ctx.intersect(np.zeros((0, 8), np.float32), SINGLE)                        # commits
    with pytest.raises(prt_b200.PrtError, match="PRT_NOALLOC_MAX_RAYS"):
        ctx.primary_shadow_noalloc((0, 0, W, 257), 1, 1, L, d_out.data_ptr())  # 263 168 rays > 262 144
    with pytest.raises(prt_b200.PrtError, match="bad region"):
        ctx.primary_shadow_noalloc((0, 0, W + 1, 8), 1, 1, L, d_out.data_ptr())
    with pytest.raises(prt_b200.PrtError, match="null"):
        ctx.primary_shadow_noalloc((0, 0, 8, 8), 1, 1, L, None)
    ctx.primary_shadow_noalloc((0, 0, W, 256), 1, 1, L, d_out.data_ptr()); ctx.synchronize()   # exactly the maximum
    got = d_out.cpu().numpy()
    hits, occ = ctx.trace_jobs([dict(kind=INT, mode=SINGLE, primary=((0, 0, W, 256), 1, 1)), dict(kind=OCC, mode=SINGLE, from_job=0, light_dir=L)])
    assert np.array_equal(got, occ)
    ctx.close()
