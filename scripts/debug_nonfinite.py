#!/usr/bin/env python3
"""Find non-finite pixels of single samples (frame indices) of a scene on the GPU and show what the CPU restatement
(bit-exact against the reference) computes for the same pixels and frame.
    python scripts/debug_nonfinite.py <scene> W H first_sample last_sample"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aten_b200 import PathTracing  # noqa: E402
from tests import oracle_lib as ol  # noqa: E402


def main():
    name, W, H, s0, s1 = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    scene = ol.ref_scene(name, W, H)
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    r.set_gather_mode(2)
    found = []
    for s in range(s0, s1):
        r.reset()
        r.render_region(W, H, 0, 0, W, H, s, s + 1, 5, 3)
        img = r.readback()
        bad = ~np.isfinite(img[..., :3]).all(axis=2)
        if bad.any():
            ys, xs = np.nonzero(bad)
            found.append((s, list(zip(xs.tolist(), ys.tolist()))[:8], int(bad.sum())))
            print(json.dumps({"sample": s, "bad_pixels": int(bad.sum()), "first": found[-1][1], "values": img[ys[0], xs[0]].tolist()}), flush=True)
            if len(found) >= 3:
                break
    for s, px, _ in found[:2]:
        want, recs, _ = ol.render(scene, frames=1, first_frame=s, records=True)
        for (x, y) in px[:3]:
            rec = recs[y * W + x]
            print(json.dumps({"sample": s, "pixel": [x, y], "restatement": want[y, x].tolist(),
                              "bounces": [{"prim": int(b["hit"]["primid"]), "mtrl": int(b["hit"]["mtrlid"]), "t": float(b["hit"]["t"]),
                                           "thr": [float(v) for v in b["throughput"]], "pdfb": float(b["pdfb"]),
                                           "contrib": [float(v) for v in b["contrib"]], "lc": [float(v) for v in b["light_contrib"]]}
                                          for b in rec]}), flush=True)
    r.close()


if __name__ == "__main__":
    main()
