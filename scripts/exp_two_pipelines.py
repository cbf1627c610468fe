"""Experiment: do the traversal kernels (issue / L1 bound) and the shade kernels (latency bound) of two independent
wavefronts fill each other's gaps when they run side by side on one GPU?  Two renderers (two streams, two sets of path
buffers) each render half of the bench frame's samples at the same time, with the kernels' grids cut to a share of
the SM (ATEN_B200_TRACE_BPS / ATEN_B200_SHADE_BPS blocks per SM); compared with one renderer doing all 16 samples.
    python scripts/exp_two_pipelines.py [trace_bps shade_bps]"""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if len(sys.argv) > 2:
    os.environ["ATEN_B200_TRACE_BPS"] = sys.argv[1]; os.environ["ATEN_B200_SHADE_BPS"] = sys.argv[2]
import numpy as np
from aten_b200 import PathTracing, scene_io as sio
W, H, SPP = 1920, 1080, int(os.environ.get("EXP_SPP", "16"))
scene = sio.load_scene(os.path.join(ROOT, "tests/golden/sponza_atrium.atnscene.gz")).with_size(W, H)
def mk():
    r = PathTracing(0); r.UpdateSceneData(scene); r.set_gather_mode(2); return r
def run(rs, spp_each, reps=4):
    best = 1e9; rays = 0
    for it in range(reps):
        for r in rs: r.reset()
        for r in rs: r.synchronize()
        t0 = time.perf_counter()
        for k, r in enumerate(rs): r.render_region(W, H, 0, 0, W, H, k * spp_each, (k + 1) * spp_each, 5, 3)
        for r in rs: r.synchronize()
        dt = time.perf_counter() - t0
        if it: best = min(best, dt)
        rays = sum(r.stats()["closest_rays"] + r.stats()["shadow_rays"] for r in rs)
    return rays / best / 1e6, best * 1e3
out = {"trace_bps": os.environ.get("ATEN_B200_TRACE_BPS"), "shade_bps": os.environ.get("ATEN_B200_SHADE_BPS")}
a = mk()
out["one_renderer_all_samples"] = run([a], SPP)
b = mk()
out["spp_total"] = SPP
out["two_renderers_half_each"] = run([a, b], SPP // 2)
print(json.dumps(out), flush=True)
