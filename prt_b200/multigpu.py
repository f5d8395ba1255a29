"""Multi-GPU plumbing of the render path: one process per GPU, scene replicated, image bands dealt
round-robin to ranks, ONE collective — the framebuffer all-gather over NCCL/NVLink (gloo on CPU in tests).

The rank-local buffer written by prt_cuda_render_bands_dev is already the all-gather send buffer (tile-major, equal
size on every rank, padded with zero rows), so nothing is repacked before the collective; `detile` is the one
permute that turns the gathered [world, bands_per_rank, band, W, 3] block back into the row-major Image layout.
"""
import torch


def band_layout(height, band_rows, world):
    nb = (height + band_rows - 1) // band_rows
    nb_per_rank = (nb + world - 1) // world
    return nb, nb_per_rank


def local_rows(height, band_rows, rank, world):
    """Global row index of every row of rank `rank`'s local buffer (-1 for padding rows)."""
    nb, nbpr = band_layout(height, band_rows, world)
    rows = []
    for j in range(nbpr):
        b = j * world + rank
        for r in range(band_rows):
            y = b * band_rows + r
            rows.append(y if y < height else -1)
    return rows


def detile(gathered, world, band_rows, width, height):
    """gathered: [world * nb_per_rank * band_rows, W, 3] (rank-major, as all_gather_into_tensor produces) -> [H, W, 3]."""
    nb, nbpr = band_layout(height, band_rows, world)
    g = gathered.view(world, nbpr, band_rows, width, 3).permute(1, 0, 2, 3, 4).reshape(nbpr * world * band_rows, width, 3)
    return g[:height]


def gather_framebuffer(local_fb, band_rows, width, height, group=None):
    """All-gathers the rank-local band buffers and returns the full row-major image on every rank."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local_fb.view(-1, width, 3)[:height]
    out = torch.empty((world * local_fb.numel(),), dtype=local_fb.dtype, device=local_fb.device)
    dist.all_gather_into_tensor(out, local_fb.reshape(-1), group=group)
    return detile(out, world, band_rows, width, height)
