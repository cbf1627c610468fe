"""Host logic of the multi-GPU path, world_size 2 over gloo on CPU: partitions cover the work exactly
once and the reduce assembles the same image as a single process."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from aten_b200 import multigpu as mg


def test_partitions_cover_exactly_once():
    for world in (1, 2, 3, 8):
        cover = np.zeros(16, int)
        for r in range(world):
            b, e = mg.partition_samples(16, world, r)
            cover[b:e] += 1
        assert (cover == 1).all()
        img = np.zeros((1080, 1920), int)
        for r in range(world):
            for (x0, y0, x1, y1) in mg.partition_tiles(1920, 1080, world, r):
                img[y0:y1, x0:x1] += 1
        assert (img == 1).all()
    assert mg.partition_samples(3, 8, 7) == (3, 3)            # ragged: more ranks than samples
    with pytest.raises(ValueError):
        mg.partition_samples(4, 2, 2)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _synthetic_sample(W, H, s):
    """Deterministic stand-in for "radiance of sample s" (the real renderer needs a GPU)."""
    rng = np.random.default_rng(100 + s)
    return rng.random((H, W, 3)).astype(np.float32)


def _worker(rank, world, port, mode, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    W, H, SPP = 96, 40, 6
    img = torch.zeros((H, W, 4))

    class FakeRenderer:   # same call the real PathTracing.render_region receives
        def render_region(self, w, h, x0, y0, x1, y1, sb, se, mb, rr):
            for s in range(sb, se):
                smp = torch.from_numpy(_synthetic_sample(W, H, s))
                img[y0:y1, x0:x1, :3] += smp[y0:y1, x0:x1]
                img[y0:y1, x0:x1, 3] += 1.0

        def render_tiles(self, w, h, tile, world_, rank_, sb, se, mb, rr):      # PathTracing.render_tiles: one wavefront over
            for (x0, y0, x1, y1) in mg.partition_tiles(w, h, world_, rank_, tile):   # exactly the tiles partition_tiles lists
                self.render_region(w, h, x0, y0, x1, y1, sb, se, mb, rr)

    mg.render_partitioned(FakeRenderer(), W, H, SPP, 5, 3, world, rank, mode=mode, tile=16)
    mg.combine_images(img, dst=0)
    if rank == 0:
        q.put(mg.resolve(img).numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["samples", "tiles"])
def test_two_rank_combine_equals_single_process(mode):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    W, H, SPP = 96, 40, 6
    want = np.zeros((H, W, 3), np.float32)
    for s in range(SPP):
        want += _synthetic_sample(W, H, s)
    want /= SPP
    assert (got[..., 3] == SPP).all()
    if mode == "tiles":
        # every pixel has exactly one non-zero contribution: bit-identical to the single-process sum
        assert (got[..., :3].view(np.uint32) == want.view(np.uint32)).all()
    else:
        assert np.allclose(got[..., :3], want, rtol=1e-6, atol=1e-7)
