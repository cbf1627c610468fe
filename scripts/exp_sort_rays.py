"""Experiment: how much does ordering the secondary rays help the traversal kernel?"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from aten_b200 import PathTracing, scene_io as sio
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = 1280, 720
scene = sio.load_scene(os.path.join(ROOT, "tests/golden/sponza_atrium.atnscene.gz")).with_size(W, H)
r = PathTracing(0); r.UpdateSceneData(scene)
tap = r.set_bounce_tap(5)
r.render(W, H, 1, 5, 3)
r.set_bounce_tap(None)
def timeit(rays, label):
    n = len(rays)
    t_rays = torch.from_numpy(np.ascontiguousarray(rays).view(np.float32).reshape(-1, 6)).cuda()
    t_hits = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    best = 1e9
    for it in range(5):
        r.synchronize(); t0 = time.perf_counter()
        r.trace_closest_device(t_rays.data_ptr(), None, n, t_hits.data_ptr()); r.synchronize()
        if it: best = min(best, time.perf_counter() - t0)
    print(f"{label:40s} n={n:8d}  {n/best/1e6:8.1f} Mrays/s")
def morton(p, lo, hi, bits=10):
    q = np.clip(((p - lo) / (hi - lo) * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)
for b in (0, 1, 2):
    rec = tap[:, b]
    v = (rec["flags"] & 1) != 0
    rays = np.ascontiguousarray(rec["ray"][v])
    print("bounce", b)
    timeit(rays, "queue (pixel) order")
    d = rays["dir"]; octant = (d[:, 0] > 0).astype(np.int64) | ((d[:, 1] > 0).astype(np.int64) << 1) | ((d[:, 2] > 0).astype(np.int64) << 2)
    timeit(rays[np.argsort(octant, kind="stable")], "sorted by direction octant (stable)")
    lo = rays["org"].min(0); hi = rays["org"].max(0) + 1e-3
    m = morton(rays["org"], lo, hi)
    timeit(rays[np.argsort((octant << 30) | m, kind="stable")], "octant, then Morton(origin)")
    timeit(rays[np.argsort(m, kind="stable")], "Morton(origin)")
    # 6-face direction bucket (dominant axis + sign) then morton
    ax = np.abs(d).argmax(1); sg = (d[np.arange(len(d)), ax] > 0).astype(np.int64)
    face = ax * 2 + sg
    timeit(rays[np.argsort((face << 30) | m, kind="stable")], "cube face, then Morton(origin)")
    rng = np.random.default_rng(0)
    timeit(rays[rng.permutation(len(rays))], "random shuffle")
