#!/usr/bin/env python3
"""Per-bounce comparison of the CUDA path (bounce tap) with the CPU restatement on a scene built by the reference:
for pixels whose final radiance differs, the first bounce and field that differ.
    python scripts/debug_records.py <scene> W H"""
import collections
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aten_b200 import PathTracing  # noqa: E402
from tests import oracle_lib as ol  # noqa: E402


def main():
    name, W, H = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
    scene = ol.ref_scene(name, W, H)
    with PathTracing(0) as r:
        r.UpdateSceneData(scene)
        tap = r.set_bounce_tap(5)
        r.render(W, H, 1, 5, 3)
        got = r.readback().reshape(-1, 4)
        tap = tap.copy()
    want, recs, _ = ol.render(scene, frames=1, records=True)
    want = want.reshape(-1, 4)
    d = np.abs(got[:, :3] - want[:, :3]).max(axis=1)
    bad = np.nonzero(d > 1e-4 * np.maximum(np.abs(want[:, :3]).max(axis=1), 1e-3))[0]
    print(json.dumps({"pixels": W * H, "differing": int(len(bad))}))
    why = collections.Counter()
    shown = 0
    for p in bad:
        for b in range(5):
            g, w = tap[p, b], recs[p, b]
            if (g["flags"] & 1) != (w["flags"] & 1):
                why[(b, "path alive on one side only")] += 1; break
            if not (w["flags"] & 1):
                continue
            fields = []
            same_hit = g["hit"]["primid"] == w["hit"]["primid"] and g["hit"]["objid"] == w["hit"]["objid"]
            if not same_hit: fields.append("hit")
            ray_d = np.abs(np.asarray(g["ray"]["dir"], np.float64) - np.asarray(w["ray"]["dir"], np.float64)).max()
            if ray_d > 1e-5: fields.append("ray direction differs by more than 1e-5")
            for f, tol in (("light_contrib", 1e-3), ("throughput", 1e-3)):
                a_, b_ = np.asarray(g[f], np.float64), np.asarray(w[f], np.float64)
                if (np.abs(a_ - b_) > tol * np.maximum(np.abs(b_), 1e-3)).any(): fields.append(f)
            if g["shadow_active"] != w["shadow_active"]: fields.append("shadow_active")
            elif g["shadow_visible"] != w["shadow_visible"]: fields.append("shadow_visible")
            if fields:
                key = (b, ",".join(fields), "material %d" % int(w["hit"]["mtrlid"]))
                why[key] += 1
                break
        else:
            why[("-", "no bounce differs by more than 1e-3", "")] += 1
    for k, v in why.most_common(15):
        print(v, k)


if __name__ == "__main__":
    main()
