"""DEVELOPMENT TOOL: node steps / triangle tests / stack depth of the QuadNode traversal on a golden scene,
from the CPU emulation in tests/emu_traversal.cpp, with its hits checked against the oracle.
usage: emu_wide.py [scene] [n_random_rays | 0 path]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aten_b200 import scene_io as sio  # noqa: E402
from tests import emu_lib, oracle_lib as ol  # noqa: E402


def random_rays(scene, n, seed=1):
    rng = np.random.default_rng(seed)
    cfg = scene.config[0]
    lo, hi = np.array(cfg["scene_bbox_min"], np.float64), np.array(cfg["scene_bbox_max"], np.float64)
    rays = np.zeros(n, sio.RAY_DTYPE)
    rays["org"] = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["dir"] = d.astype(np.float32)
    return rays


if __name__ == "__main__":
    name = sys.argv[1] if len(sys.argv) > 1 else "sponza_atrium"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 400000
    scene = sio.load_scene(os.path.join(ROOT, "tests", "golden", f"{name}.atnscene.gz"))
    if len(sys.argv) > 3 and sys.argv[3] == "path":       # rays of real paths (all bounces), from the oracle's records
        _, recs, _ = ol.render(scene.with_size(320, 180), frames=2, records=True)
        r = recs.reshape(-1)
        rays = np.ascontiguousarray(r["ray"][np.abs(r["ray"]["dir"]).sum(axis=1) > 0])
        n = len(rays)
    else:
        rays = random_rays(scene, n)
    ref = ol.trace_closest(scene, rays)
    for label, kw in (("entry-distance order", {}), ("leaves first", {"leaves_first": True}), ("no instance filter", {"entry_filter": False})):
        hits, st = emu_lib.trace_closest(scene, rays, **kw)
        bad = int(((hits["primid"] != ref["primid"]) & (ref["objid"] >= 0)).sum() + (hits["objid"] != ref["objid"]).sum())
        print(f"{label:22s} rays {n} nodes/ray {st['nodes']/n:6.2f} tris/ray {st['tris']/n:5.2f} entries/ray {st['entries']/n:5.2f} "
              f"culled pops/ray {st['pops_culled']/n:5.2f} max stack {st['max_sp']} (bound {st['stack_bound']}) "
              f"QuadNodes {st['node_count']} slots {st['slots_used']} not nested {st['not_nested']}  id differences vs oracle {bad}")
