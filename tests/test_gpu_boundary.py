"""The rest of the drop-in boundary on the GPU: Scene::getSurfaceProperties over a batch, contexts bound to
their device (two GPUs interleaved, calls from another thread), one context driving several GPUs (prt_cuda_create_multi:
NCCL-gathered render == single-GPU render, ray batches split across GPUs), the one-process-per-GPU communicator, and the
traversal counters of the counting build against the oracle's."""
import os
import subprocess
import sys
import threading
import numpy as np
import pytest

import prt_b200
from prt_b200 import capi, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SINGLE, PACKET = prt_b200.PRT_MODE_SINGLE, prt_b200.PRT_MODE_PACKET


def _eq(a, b):
    return np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))


def _ndev():
    return capi.lib().prt_cuda_device_count()


def _surface_as_words(s):
    return np.ascontiguousarray(s).view(np.uint32).reshape(-1, 16)


@pytest.mark.parametrize("make", [lambda: scenes.cornell_teapot(256, 144, teapot_refl=scenes.DIFFUSE, light="directional"),
                                  lambda: scenes.atrium(256, 144, tex_size=128, detail=0.25)])
def test_surface_properties_match_oracle(oracle_mod, make):
    """prt_cuda_surface_properties vs Mesh::getSurfaceProperties (mesh.cpp:311-372) restated by the oracle, on C1's scene
    (geometric normals on the box, vertex normals on the teapot, no uv) and on the textured scene (uv, bump frame)."""
    desc = make()
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    W, H = desc["camera"]["width"], desc["camera"]["height"]
    rays, hits = orc.primary_single((0, 0, W, H), 1, 3)
    # secondary hits too: they reach triangles primary rays never see
    t = hits[:, 0]; hit = t != -1
    rng = np.random.RandomState(5)
    d = rng.randn(int(hit.sum()), 3).astype(np.float32); d /= np.linalg.norm(d, axis=1, keepdims=True)
    sec = np.zeros((int(hit.sum()), 8), np.float32)
    sec[:, 0:3] = rays[hit, 0:3] + rays[hit, 3:6] * t[hit, None]; sec[:, 3:6] = d; sec[:, 6] = 1e5
    hits2 = orc.intersect_single(sec)
    allhits = np.concatenate([hits, hits2])
    got = ctx.surface_properties(allhits)
    want = orc.surface_properties(allhits)
    gw, ww = _surface_as_words(got), want.view(np.uint32)
    bad = (gw != ww).any(axis=1)
    assert not bad.any(), f"{int(bad.sum())}/{len(bad)} surface records differ; first {np.where(bad)[0][:5]}"
    miss = allhits[:, 0] == -1
    assert miss.any() and (gw[miss] == 0).all()
    nmesh = len(desc["meshes"])
    assert len(np.unique(allhits.view(np.uint32)[~miss, 5])) == nmesh
    with pytest.raises(prt_b200.PrtError, match="does not exist"):
        bogus = allhits[~miss][:1].copy(); bogus.view(np.uint32)[0, 4] = 0x7fffffff
        ctx.surface_properties(bogus)
    ctx.close()


def test_surface_properties_match_reference_golden():
    """Same call against records the unmodified reference produced (tests/golden/surface.npz, make_golden_surface.py)."""
    g = np.load(os.path.join(ROOT, "tests", "golden", "surface.npz"))
    desc = scenes.cornell_teapot(int(g["W"]), int(g["H"]), teapot_refl=scenes.DIFFUSE, light="directional")
    ctx = prt_b200.load_scene(desc, 0)
    got = ctx.surface_properties(g["hits"])
    assert np.array_equal(_surface_as_words(got), g["surface"].view(np.uint32))
    ctx.close()


def test_context_is_bound_to_its_device_across_threads(oracle_mod):
    """calls from a thread other than the creating one (the reference's ThreadPool pattern, main.cpp:132-158)."""
    desc = scenes.cornell_teapot(128, 72, teapot_refl=scenes.DIFFUSE, light="directional")
    ctx = prt_b200.load_scene(desc, 0)
    orc = oracle_mod.OracleScene(desc)
    rays, hits = orc.primary_single((0, 0, 128, 72), 1, 1)
    want_img = orc.render(8, 2, nthreads=4)
    out = {}

    def work(k):
        out[k] = (ctx.intersect(rays, SINGLE), ctx.render(8, 2)) if k == 0 else None

    th = threading.Thread(target=work, args=(0,)); th.start(); th.join()
    assert _eq(out[0][0], hits) and _eq(out[0][1], want_img)
    ctx.close()


@pytest.mark.skipif(_ndev() < 2, reason="needs 2 GPUs")
def test_two_contexts_on_two_gpus_interleaved(oracle_mod):
    """ctx0 used after ctx1 was created used to allocate on GPU 1 and launch on a GPU-0 stream."""
    desc = scenes.cornell_teapot(128, 72, teapot_refl=scenes.DIFFUSE, light="directional")
    orc = oracle_mod.OracleScene(desc)
    rays, hits = orc.primary_single((0, 0, 128, 72), 1, 1)
    want_img = orc.render(8, 2, nthreads=4)
    c0 = prt_b200.load_scene(desc, 0)
    c1 = prt_b200.load_scene(desc, 1)
    for _ in range(2):
        assert _eq(c0.intersect(rays, SINGLE), hits)
        assert _eq(c1.intersect(rays, SINGLE), hits)
        assert _eq(c1.render(8, 2), want_img)
        assert _eq(c0.render(8, 2), want_img)
    c0.close(); c1.close()


@pytest.mark.skipif(_ndev() < 2, reason="needs 2 GPUs")
def test_multi_gpu_context_render_and_batches_match_single_gpu(oracle_mod):
    """prt_cuda_create_multi: bands per GPU + ncclAllGather + de-tile == the single-GPU frame, bit for bit; sub-regions;
    host-buffer batches split across the GPUs."""
    n = min(_ndev(), 4)
    desc = scenes.cornell_teapot(200, 117, teapot_refl=scenes.SPECULAR, light="infinite")
    one = prt_b200.load_scene(desc, 0)
    many = prt_b200.load_scene(desc, list(range(n)))
    assert many.comm_info() == (0, n, n)
    ref = one.render(16, 7)
    got = many.render(16, 7)
    assert _eq(got, ref), f"{n}-GPU gathered frame differs from the 1-GPU frame in {int((got != ref).any(axis=2).sum())} pixels"
    sub = (13, 9, 150, 101)
    a = np.full((117, 200, 3), -1.0, np.float32); b = a.copy()
    one.render(8, 3, region=sub, out=a); many.render(8, 3, region=sub, out=b)
    assert _eq(a, b) and (a[0] == -1).all() and (a[9:101, 13:150] != -1).all()
    orc = oracle_mod.OracleScene(desc)
    rays, avg, hits = orc.primary_packets((0, 0, 200, 117), 1, 5)
    assert _eq(many.intersect(rays, PACKET, avg), hits)
    srays = rays.copy(); srays[:, 6] = 3.0
    assert np.array_equal(many.occluded(srays, SINGLE), one.occluded(srays, SINGLE))
    h2, o2 = many.trace_jobs([dict(kind=0, mode=SINGLE, primary=((0, 0, 200, 117), 1, 5)), dict(kind=1, mode=SINGLE, from_job=0, light_dir=(0.2, 1.0, 0.2))])
    h1, o1 = one.trace_jobs([dict(kind=0, mode=SINGLE, primary=((0, 0, 200, 117), 1, 5)), dict(kind=1, mode=SINGLE, from_job=0, light_dir=(0.2, 1.0, 0.2))])
    assert _eq(h1, h2) and np.array_equal(o1, o2)
    c = many.counters()
    assert c["lastGatherMs"] > 0.0
    one.close(); many.close()


_RANK_SCRIPT = r"""
import os, sys, zlib
import numpy as np
sys.path.insert(0, {root!r})
import torch, torch.distributed as dist
import prt_b200
from prt_b200 import capi, scenes
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
ident = torch.zeros(capi.PRT_COMM_ID_BYTES, dtype=torch.uint8)
if rank == 0:
    ident = torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8).clone()
ident = ident.cuda(); dist.broadcast(ident, 0)
desc = scenes.cornell_teapot(200, 117, teapot_refl=scenes.SPECULAR, light="infinite")
ctx = prt_b200.load_scene(desc, rank)
ctx.comm_init_rank(bytes(ident.cpu().numpy().tobytes()), rank, world)
img = ctx.render(16, 7)
np.save(os.path.join({out!r}, f"rank{{rank}}.npy"), img)
dist.barrier(); dist.destroy_process_group()
"""


@pytest.mark.skipif(_ndev() < 2, reason="needs 2 GPUs")
def test_one_process_per_gpu_communicator_gives_the_single_gpu_frame(tmp_path):
    """prt_cuda_comm_init_rank under torchrun (the launch bench.py gets): every rank ends up with the full frame, and it is
    the 1-GPU frame."""
    n = min(_ndev(), 4)
    script = tmp_path / "rank.py"
    script.write_text(_RANK_SCRIPT.format(root=ROOT, out=str(tmp_path)))
    subprocess.check_call([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
                           "--master-port", "29713", str(script)], timeout=600)
    desc = scenes.cornell_teapot(200, 117, teapot_refl=scenes.SPECULAR, light="infinite")
    one = prt_b200.load_scene(desc, 0)
    ref = one.render(16, 7)
    for r in range(n):
        assert _eq(np.load(tmp_path / f"rank{r}.npy"), ref), f"rank {r}'s gathered frame differs from the 1-GPU frame"
    one.close()


def test_counting_build_agrees_with_the_oracle_counters(oracle_mod):
    """libprt_cuda_stats.so (-DPRT_STATS=1): node records fetched + leaves entered per single-rule closest-hit ray must equal
    the nodes the reference algorithm traverses, triangle tests its primsTraversed (stats.h:10-33, bvh.cpp:464-466,555-557);
    results stay bit-identical to the product library."""
    if not os.path.exists(capi.STATS_LIB_PATH):
        pytest.skip("counting build not present")
    W, H = 320, 180
    desc = scenes.bunny_standin(W, H)
    ctx = prt_b200.load_scene(desc, 0, stats=True)
    orc = oracle_mod.OracleScene(desc)
    rays, hits = orc.primary_single((0, 0, W, H), 1, 1)
    ok = np.isfinite(1.0 / rays[:, 3:6]).all(axis=1)             # rays with a zero direction component take the uncounted exact path
    rays, hits = rays[ok], hits[ok]
    oracle_mod.oracle_stats(enable=True, reset=True)
    orc.intersect_single(rays)
    o = oracle_mod.oracle_stats(enable=False)
    ctx.reset_counters()
    got = ctx.intersect(rays, SINGLE)
    assert _eq(got, hits)
    s = ctx.counters()["stats"]["closest_single"]
    assert s["rays"] == rays.shape[0] == o["single_rays"]
    # the reference pops a node per box entered: interior nodes (one 64-byte record here) + leaves; single-mesh scene, so the
    # root is an interior node and every pop is one or the other
    assert s["nodeSteps"] + s["leafVisits"] == o["single_nodes"], (s, o)
    assert s["triTests"] == o["single_prims"], (s, o)
    assert s["uvLoads"] == 0
    ctx.close()
