"""Host-side multi-rank logic on CPU: band layout, de-tiling, and the framebuffer all-gather over gloo (world_size 2)."""
import os
import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from prt_b200 import multigpu


@pytest.mark.parametrize("H,band,world", [(45, 4, 3), (64, 8, 2), (1080, 8, 8), (7, 8, 4), (16, 1, 5)])
def test_band_layout_covers_every_row_once(H, band, world):
    seen = np.zeros(H, int)
    nb, nbpr = multigpu.band_layout(H, band, world)
    for r in range(world):
        rows = multigpu.local_rows(H, band, r, world)
        assert len(rows) == nbpr * band
        for y in rows:
            if y >= 0:
                seen[y] += 1
    assert (seen == 1).all()


def _fake_local(H, W, band, rank, world):
    """What prt_cuda_render_bands_dev would write if pixel (y,x) had value (y, x, y*W+x)."""
    rows = multigpu.local_rows(H, band, rank, world)
    buf = torch.zeros((len(rows), W, 3), dtype=torch.float32)
    for i, y in enumerate(rows):
        if y >= 0:
            buf[i, :, 0] = y
            buf[i, :, 1] = torch.arange(W, dtype=torch.float32)
            buf[i, :, 2] = y * W + torch.arange(W, dtype=torch.float32)
    return buf


def test_detile_single_process():
    H, W, band, world = 45, 10, 4, 3
    g = torch.cat([_fake_local(H, W, band, r, world).reshape(-1) for r in range(world)])
    img = multigpu.detile(g, world, band, W, H)
    assert img.shape == (H, W, 3)
    assert torch.equal(img[:, :, 0], torch.arange(H, dtype=torch.float32)[:, None].expand(H, W))
    assert torch.equal(img[:, :, 2], (torch.arange(H)[:, None] * W + torch.arange(W)[None, :]).float())


def _worker(rank, world, port, H, W, band, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        img = multigpu.gather_framebuffer(_fake_local(H, W, band, rank, world), band, W, H)
        ok = torch.equal(img[:, :, 2], (torch.arange(H)[:, None] * W + torch.arange(W)[None, :]).float()) and img.shape == (H, W, 3)
        out[rank] = 1 if ok else 0
    finally:
        dist.destroy_process_group()


def test_gather_framebuffer_gloo_world2():
    world, H, W, band = 2, 37, 12, 8
    out = mp.get_context("spawn").Array("i", [0] * world)
    port = 29500 + (os.getpid() % 2000)
    procs = [mp.get_context("spawn").Process(target=_worker, args=(r, world, port, H, W, band, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert list(out) == [1] * world
