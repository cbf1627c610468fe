"""Multi-GPU partition of the path: one process per GPU, scene replicated, work split by sample
index or by image tile, float4 images combined with one reduce.

The reference has no multi-GPU code at all (no cudaSetDevice / NCCL anywhere); pixels and samples
are independent, so the only exchange step is the final image combine:

  * sample split: rank r renders samples [begin, end) of every pixel in SUM mode (xyz = sum of
    radiance, w = sample count, ``atn_set_gather_mode(ATN_GATHER_SUM)``); ``reduce(SUM)`` then one divide by w;
  * tile split: rank r renders an interleaved set of tiles; the other pixels stay (0,0,0,0), so the
    same ``reduce(SUM)`` assembles the frame (bit-identical to a single-GPU render, since no pixel
    receives more than one non-zero term).

The functions take any ``torch.distributed`` process group: NCCL over NVLink on the GPU box, gloo in
the CPU tests.
"""
from __future__ import annotations

from typing import List, Tuple


def partition_samples(total_spp: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous sample-index range of ``rank``; ranges differ by at most one sample."""
    if not (0 <= rank < world) or total_spp < 0:
        raise ValueError("bad partition arguments")
    base, extra = divmod(total_spp, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def partition_tiles(width: int, height: int, world: int, rank: int, tile: int = 64) -> List[Tuple[int, int, int, int]]:
    """Interleaved (round-robin in scanline order) tile rectangles (x0, y0, x1, y1) of ``rank``."""
    if not (0 <= rank < world) or tile <= 0:
        raise ValueError("bad partition arguments")
    rects = []
    k = 0
    for y0 in range(0, height, tile):
        for x0 in range(0, width, tile):
            if k % world == rank:
                rects.append((x0, y0, min(x0 + tile, width), min(y0 + tile, height)))
            k += 1
    return rects


def combine_images(image, group=None, dst: int = 0):
    """``reduce(SUM)`` of the float4 sum images onto ``dst`` (in place).  On CUDA the reduce is asynchronous on torch's
    current stream while the renderer clears and writes its image on its own stream: synchronize torch's stream before
    the next ``reset()`` / render on that handle."""
    import torch.distributed as dist
    dist.reduce(image, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return image


def resolve(image):
    """Sum image (xyz = radiance sum, w = count) -> mean radiance, w = count; pixels with w = 0 stay 0."""
    import torch
    w = image[..., 3:4]
    out = image.clone()
    out[..., :3] = torch.where(w > 0, image[..., :3] / torch.clamp(w, min=1.0), torch.zeros_like(image[..., :3]))
    return out


def render_partitioned(renderer, width: int, height: int, total_spp: int, max_bounce: int, rr_bounce: int,
                       world: int, rank: int, mode: str = "samples", tile: int = 64):
    """Render this rank's share with an ``aten_b200.PathTracing`` (already in SUM mode, image reset)."""
    if mode == "samples":
        b, e = partition_samples(total_spp, world, rank)
        if e > b:
            renderer.render_region(width, height, 0, 0, width, height, b, e, max_bounce, rr_bounce)
    elif mode == "tiles":
        # the same tiles partition_tiles() lists, rendered as one wavefront (atn_render_tiles)
        renderer.render_tiles(width, height, tile, world, rank, 0, total_spp, max_bounce, rr_bounce)
    else:
        raise ValueError(f"unknown partition mode {mode!r}")
