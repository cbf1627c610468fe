#!/usr/bin/env python3
"""CPU experiment behind the near-tie rule of the traversal kernel (no GPU): runs the kernel's per-ray logic with the
near-tie bookkeeping (tests/emu_traversal.cpp) and the reference-order walk (oracle) over random and path-like rays and
reports, per band setting, the share of flagged rays and the rays that are NOT flagged yet differ from the reference
(must be 0 for the rule to be sound).
    python scripts/exp_near_ties.py [rays] [scene ...]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scripts"))
from aten_b200 import scene_io as sio  # noqa: E402
from tests import emu_lib as emu, oracle_lib as ol  # noqa: E402
from parity_campaign import rays_for  # noqa: E402


def surface_rays(scene, hits, rays, seed):
    """Secondary rays: from the hit points of `rays`, uniformly distributed directions."""
    rng = np.random.default_rng(seed)
    ok = hits["objid"] >= 0
    o = rays["org"][ok].astype(np.float64) + rays["dir"][ok].astype(np.float64) * hits["t"][ok][:, None].astype(np.float64)
    d = rng.normal(size=o.shape); d /= np.linalg.norm(d, axis=1, keepdims=True)
    out = np.zeros(len(o), sio.RAY_DTYPE)
    out["org"] = (o + 1e-3 * d).astype(np.float32)
    out["dir"] = d.astype(np.float32)
    return out


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    names = sys.argv[2:] or ["sponza_atrium", "cornell_mtrl", "instanced"]
    bands = [(2e-5, 2.5, 1e-6, 1e-5), (5e-5, 3.0, 2e-6, 5e-5), (1e-4, 4.0, 2e-6, 2e-4)]
    for name in names:
        scene = ol.ref_scene(name, 8, 8)
        r0 = rays_for(scene, n, seed=5)
        want0 = ol.trace_closest(scene, r0)
        r1 = surface_rays(scene, want0, r0, seed=6)
        for label, rays in (("random", r0), ("surface", r1)):
            want = want0 if label == "random" else ol.trace_closest(scene, rays)
            for band in bands:
                got, flags, st = emu.trace_closest_flagged(scene, rays, band)
                differ = (got["primid"] != want["primid"]) | (got["objid"] != want["objid"]) | (got["t"].view(np.uint32) != want["t"].view(np.uint32))
                missed = differ & (flags == 0)
                rel = 0.0
                if missed.any():
                    both = missed & (got["objid"] >= 0) & (want["objid"] >= 0)
                    if both.any():
                        rel = float((np.abs(got["t"][both].astype(np.float64) - want["t"][both]) / want["t"][both]).max())
                print(json.dumps({"scene": name, "rays": label, "n": int(len(rays)), "band": band, "flagged": int(flags.sum()),
                                  "flagged_share": float(flags.mean()), "differ": int(differ.sum()), "differ_not_flagged": int(missed.sum()),
                                  "worst_rel_gap_not_flagged": rel, "node_steps_per_ray": st["nodes"] / len(rays)}), flush=True)


if __name__ == "__main__":
    main()
