"""DEVELOPMENT TOOL: runs scripts/emu_wide.cpp (CPU emulation of the traversal over the PairNode and
QuadNode encodings) on golden scenes and compares its hits with the oracle.  Not used by the product."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from aten_b200 import scene_io as sio  # noqa: E402
import oracle_lib as ol  # noqa: E402

SO = "/tmp/libemu_wide.so"


def build():
    subprocess.run(["g++", "-std=c++17", "-O2", "-fopenmp", "-fPIC", "-ffp-contract=off", "-shared", "-o", SO,
                    os.path.join(ROOT, "scripts", "emu_wide.cpp")], check=True)


def run(fn, scene, rays, *extra):
    lib = C.CDLL(SO)
    rays = np.ascontiguousarray(rays, dtype=sio.RAY_DTYPE)
    hits = np.zeros(len(rays), dtype=sio.HIT_DTYPE)
    stats = np.zeros(16, np.uint64)
    blob = scene.blob()
    rc = getattr(lib, fn)(C.byref(blob), C.c_void_p(rays.ctypes.data), None, C.c_size_t(len(rays)),
                          C.c_void_p(hits.ctypes.data), C.c_void_p(stats.ctypes.data), *[C.c_int(e) for e in extra])
    assert rc == 0
    return hits, stats


def random_rays(scene, n, seed=1):
    rng = np.random.default_rng(seed)
    cfg = scene.config[0]
    lo, hi = np.array(cfg['scene_bbox_min'], np.float64), np.array(cfg['scene_bbox_max'], np.float64)
    rays = np.zeros(n, sio.RAY_DTYPE)
    rays["org"] = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["dir"] = d.astype(np.float32)
    return rays


def compare(a, b):
    bad = (a["primid"] != b["primid"]) | (a["objid"] != b["objid"])
    hit = a["objid"] >= 0
    bad_f = hit & ((a["t"] != b["t"]) | (a["a"] != b["a"]) | (a["b"] != b["b"]))
    return int(bad.sum()), int((bad_f & ~bad).sum())


if __name__ == "__main__":
    build()
    name = sys.argv[1] if len(sys.argv) > 1 else "sponza_atrium"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 400000
    scene = sio.load_scene(os.path.join(ROOT, "tests", "golden", f"{name}.atnscene.gz"))
    if len(sys.argv) > 3 and sys.argv[3] == "path":       # rays of real paths (all bounces), from the oracle's records
        sc = scene.with_size(320, 180)
        _, recs, _ = ol.render(sc, frames=2, records=True)
        r = recs.reshape(-1)
        keep = np.abs(r["ray"]["dir"]).sum(axis=1) > 0
        rays = np.ascontiguousarray(r["ray"][keep])
        n = len(rays)
        print("path rays:", n)
    else:
        rays = random_rays(scene, n)
    ref = ol.trace_closest(scene, rays)
    prev = None
    for label, fn, extra in (("pair", "emu_trace_pair", ()), ("wide t0-sorted", "emu_trace_wide", (0, 1)),
                             ("wide leaves-first", "emu_trace_wide", (1, 1))):
        hits, st = run(fn, scene, rays, *extra)
        bad, badf = compare(hits, ref)
        if prev is not None:
            print("   vs pair emulation: id/float mismatches", compare(hits, prev))
        else:
            prev = hits
        print(f"{label:18s} nodes/ray {st[0]/n:6.2f} tris/ray {st[1]/n:5.2f} entries/ray {st[2]/n:5.2f} filtered {st[3]/n:5.2f} "
              f"culled pops/ray {st[4]/n:5.2f} max_sp {st[5]} nodes {st[6]} slots {st[7]} replaced {st[8]} not_nested {st[9]} "
              f"stack bound {st[10]}  id mismatches {bad} float mismatches {badf}")
