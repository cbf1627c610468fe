#!/usr/bin/env python3
"""Large random-ray parity campaign: closest hits of the CUDA traversal against the CPU restatement (reference visit
order) on scenes built on the spot by the reference's own loaders / SBVH builder.  Prints one JSON line per scene with
the number of rays, hit/miss disagreements, differing primitive ids, the largest relative distance gap among them
(max_rel_dt; all three must be 0 with the exact walk of near ties on, the default) and the share of rays that walk
decided (exact_rays).  ATEN_B200_EXACT_TIES=0 shows what the front-to-back walk alone reports.
    python scripts/parity_campaign.py [rays_per_scene] > profiles/rNN_parity_campaign.jsonl"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aten_b200 import PathTracing, scene_io as sio  # noqa: E402
from tests import oracle_lib as ol  # noqa: E402


def rays_for(scene, n, seed):
    rng = np.random.default_rng(seed)
    cfg = scene.config[0]
    lo, hi = np.array(cfg["scene_bbox_min"], np.float64), np.array(cfg["scene_bbox_max"], np.float64)
    pad = 0.05 * (hi - lo)
    rays = np.zeros(n, sio.RAY_DTYPE)
    rays["org"] = (lo - pad + (hi - lo + 2 * pad) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    # a tenth of the rays axis aligned or nearly so (the ill-conditioned slab cases)
    k = n // 10
    axis = rng.integers(0, 3, k)
    d[:k] = 0.0
    d[np.arange(k), axis] = rng.choice([-1.0, 1.0], k)
    d[k // 2:k] += 1e-4 * rng.normal(size=(k - k // 2, 3))
    d[:k] /= np.linalg.norm(d[:k], axis=1, keepdims=True)
    rays["dir"] = d.astype(np.float32)
    return rays


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 4_000_000
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import flat_scene as fs
    form = {"0": "default", "1": "streaming"}.get(os.environ.get("ATEN_B200_STREAM", ""), "picked at upload")
    for name in ("cornell_mtrl", "sponza_atrium", "instanced", "instanced_dragon", "dragon_grid4", "dragon_grid4 flattened"):
        if name.endswith(" flattened"):     # 16 distinct dragon meshes, 1.6 M triangles, 0.2 GB of nodes + triangles: beyond the L2
            scene = fs.flatten_instances(ol.ref_scene(name.split()[0], 8, 8))
        else:
            scene = ol.ref_scene(name, 8, 8)
        r = PathTracing(0)
        r.UpdateSceneData(scene)
        bad_miss = bad_id = exact = 0
        worst = 0.0
        total = 0
        for chunk in range(0, n, 1_000_000):
            m = min(1_000_000, n - chunk)
            rays = rays_for(scene, m, seed=chunk + 17)
            got = r.trace_closest(rays)
            exact += r.stats()["exact_rays"]
            want = ol.trace_closest(scene, rays)
            hit_g, hit_w = got["objid"] >= 0, want["objid"] >= 0
            bad_miss += int((hit_g != hit_w).sum())
            both = hit_g & hit_w
            diff = both & ((got["primid"] != want["primid"]) | (got["objid"] != want["objid"]) | (got["t"].view(np.uint32) != want["t"].view(np.uint32)))
            bad_id += int(diff.sum())
            if diff.any():
                rel = np.abs(got["t"][diff].astype(np.float64) - want["t"][diff]) / want["t"][diff]
                worst = max(worst, float(rel.max()))
            same = both & ~diff
            assert (got["t"][same].view(np.uint32) == want["t"][same].view(np.uint32)).all()
            assert (got["a"][same].view(np.uint32) == want["a"][same].view(np.uint32)).all()
            total += m
        r.close()
        print(json.dumps({"scene": name, "kernel_form": form, "triangles": int(len(scene.triangles)), "rays": total, "hit_miss_disagreements": bad_miss,
                          "differing_primitives": bad_id, "rate": bad_id / total, "max_rel_dt": worst,
                          "exact_ties": os.environ.get("ATEN_B200_EXACT_TIES", "1") != "0", "exact_rays": int(exact), "exact_share": exact / total}), flush=True)


if __name__ == "__main__":
    main()
