#!/usr/bin/env python3
"""BASELINE.json configs[3] on N GPUs of one box: the instanced dragon + teapot + suzanne scene (stand-in: bunny.obj
is not in the snapshot) with Disney and refraction materials, 3840x2160, 64 samples per pixel in total, partitioned
across the ranks by sample index (rank r renders samples [64 r / N, 64 (r + 1) / N) of every pixel in sum mode), the
float4 sum images combined with one NCCL reduce over NVLink onto rank 0 -- strong scaling: the frame is fixed.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_c4_multigpu.py [--steps K] [--warmup W] [--spp 64]

Rank 0 prints one JSON line; with N = 1 the script runs as a plain process.  The combined image is checked against the
same frame rendered by rank 0 alone (relative RMSE <= 1e-5: only the order of the float additions differs).  The reference's Disney
material yields a non-finite sample about once per 8 M (reproduced bit for bit); those pixels are
counted and left out of the means."""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aten_b200 import PathTracing, multigpu  # noqa: E402
from tests import oracle_lib as ol  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--spp", type=int, default=64)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    W, H, BOUNCE, RR = 3840, 2160, 5, 3
    scene = ol.ref_scene("instanced_dragon", W, H)
    r = PathTracing(local)
    r.UpdateSceneData(scene)
    r.set_gather_mode(2)                      # sum mode: xyz = radiance sum, w = sample count
    ptr, _ = r.device_image()

    class _Img:
        __cuda_array_interface__ = {"shape": (H, W, 4), "typestr": "<f4", "data": (ptr, False), "version": 3}
    image = torch.as_tensor(_Img(), device=f"cuda:{local}")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        r.reset()
        multigpu.render_partitioned(r, W, H, a.spp, BOUNCE, RR, world, rank, mode="samples")
        st = r.stats()
        if world > 1:
            multigpu.combine_images(image)
            torch.cuda.current_stream().synchronize()   # the next step clears the image on the renderer's own stream
        return st

    for _ in range(a.warmup):
        step()
    barrier()
    t0 = time.perf_counter()
    rays = samples = 0
    for _ in range(a.steps):
        st = step()
        rays += st["closest_rays"] + st["shadow_rays"]; samples += st["camera_samples"]
    barrier()
    wall = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([wall], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); wall = float(t)
        w = torch.tensor([rays, samples], dtype=torch.float64, device="cuda"); dist.all_reduce(w); rays, samples = (float(x) for x in w.tolist())
    if rank == 0:
        full = multigpu.resolve(image).cpu().numpy()
        counts_ok = bool((full[..., 3] == a.spp).all())
        # the same frame on rank 0 alone: the partitioned result differs from it by the order of the float additions only
        r.reset(); r.render_region(W, H, 0, 0, W, H, 0, a.spp, BOUNCE, RR)
        r.synchronize()            # render calls are asynchronous; the resolve below runs on torch's stream
        one = multigpu.resolve(image).cpu().numpy()
        bad = ~np.isfinite(full[..., :3]).all(axis=2)
        same_bad = bool((bad == ~np.isfinite(one[..., :3]).all(axis=2)).all())
        ok = ~bad
        d = full[ok][:, :3].astype(np.float64) - one[ok][:, :3]
        rel = float(np.sqrt(np.mean(d * d)) / np.sqrt(np.mean(one[ok][:, :3].astype(np.float64) ** 2)))
        print(json.dumps({"config": "C4 stand-in: instanced dragon + teapots + suzannes, Disney / refraction, 3840x2160, "
                                    f"{a.spp} spp in total, sample-index partition, NCCL reduce of the float4 sum images",
                          "n_gpus": world, "steps": a.steps, "ms_per_frame": round(1e3 * wall / a.steps, 2),
                          "mrays_per_s": round(rays / wall / 1e6, 1), "msamples_per_s": round(samples / wall / 1e6, 1),
                          "scaling": "strong", "timing": "barrier + synchronize on both sides, max over ranks, reduce inside",
                          "every_pixel_has_all_samples": counts_ok, "relative_rmse_vs_one_gpu": rel,
                          "non_finite_pixels_as_in_the_reference": int(bad.sum()), "same_non_finite_pixels_as_one_gpu": same_bad,
                          "mean_radiance": float(full[ok][:, :3].mean())}), flush=True)
        assert counts_ok and same_bad and rel <= 1e-5
    if world > 1:
        dist.destroy_process_group()
    r.close()


if __name__ == "__main__":
    main()
