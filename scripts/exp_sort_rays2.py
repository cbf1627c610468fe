"""Experiment (round 2): what would ordering the secondary rays of a LARGE wavefront buy the traversal kernel?
Same idea as exp_sort_rays.py, at 4K x 1 spp (8.3 M paths: the ray density of four 1080p samples per wavefront), with
coarse keys that a counting sort inside the pipeline could afford.  Times atn_trace_closest on device-resident rays."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from aten_b200 import PathTracing, scene_io as sio
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3840, 2160)
scene = sio.load_scene(os.path.join(ROOT, "tests/golden/sponza_atrium.atnscene.gz")).with_size(W, H)
r = PathTracing(0); r.UpdateSceneData(scene)
tap = r.set_bounce_tap(3)
r.render(W, H, 1, 3, 3)
r.set_bounce_tap(None)
def timeit(rays, label, out):
    n = len(rays)
    t_rays = torch.from_numpy(np.ascontiguousarray(rays).view(np.float32).reshape(-1, 6)).cuda()
    t_hits = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    best = 1e9
    for it in range(4):
        r.synchronize(); t0 = time.perf_counter()
        r.trace_closest_device(t_rays.data_ptr(), None, n, t_hits.data_ptr()); r.synchronize()
        if it: best = min(best, time.perf_counter() - t0)
    out[label] = round(n / best / 1e6, 1)
    print(f"{label:44s} n={n:8d}  {n/best/1e6:8.1f} Mrays/s", flush=True)
def cell(p, lo, hi, bits):
    q = np.clip(((p - lo) / (hi - lo) * (1 << bits)).astype(np.int64), 0, (1 << bits) - 1)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)
for b in (1, 2):
    rec = tap[:, b]
    v = (rec["flags"] & 1) != 0
    rays = np.ascontiguousarray(rec["ray"][v])
    out = {"bounce": b, "rays": int(len(rays))}
    timeit(rays, "queue (pixel) order", out)
    d = rays["dir"]; octant = (d[:, 0] > 0).astype(np.int64) | ((d[:, 1] > 0).astype(np.int64) << 1) | ((d[:, 2] > 0).astype(np.int64) << 2)
    timeit(rays[np.argsort(octant, kind="stable")], "octant (stable: pixel order inside)", out)
    lo = rays["org"].min(0); hi = rays["org"].max(0) + 1e-3
    for bits in (3, 4, 5, 7):
        m = cell(rays["org"], lo, hi, bits)
        timeit(rays[np.argsort((octant << 30) | m, kind="stable")], f"octant, then {bits}-bit/axis cell (stable)", out)
        timeit(rays[np.argsort((m << 3) | octant, kind="stable")], f"{bits}-bit/axis cell, then octant (stable)", out)
    # windowed: octant buckets inside windows of 64 K consecutive queue entries (keeps path state nearly in order)
    for win in (256, 1024, 4096, 16384, 65536):
        key = (np.arange(len(rays)) // win) * 8 + octant
        timeit(rays[np.argsort(key, kind="stable")], f"octant inside windows of {win}", out)
    print(json.dumps(out), flush=True)
