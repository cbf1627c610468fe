#!/usr/bin/env python3
"""Generates the golden vectors under tests/golden/ from the REFERENCE ITSELF (oracle/_ref, i.e.
nackdai/aten's own CPU path tracer compiled in place by oracle/build_ref.py).

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py

Files (all little endian, gzip):
  cmj.f32.gz                      aten::CMJ known-answer stream (ref_harness cmj)
  <scene>.atnscene.gz             flat scene blob at the golden resolution (ref_harness scene)
  <scene>.trace.gz                AtnBounceRecord[H*W][5] of frame FRAME, replayed from the reference's
                                  public functions and self-checked against stock render()
  <scene>.image.f32.gz            float4 image of frames 0..FRAMES-1 through aten::PathTracing::render
                                  with FilmProgressive
  <scene>.rays.gz / .hits.gz      random rays {org,dir,tmax} and the reference's Traverse<Closest> hits
  <scene>.ao.f32.gz               float4 film of ONE frame (index FRAME) of the stock aten::AORenderer (ref_harness ao)
"""
import gzip
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from tests import oracle_lib as ol  # noqa: E402
from aten_b200 import scene_io as sio  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
SCENES = {  # name: (W, H)
    "cornell": (64, 64),
    "cornell_mtrl": (64, 48),
    "sponza_atrium": (64, 36),
    "instanced": (64, 36),
    "cornell_sphere": (64, 64),     # a sphere area light + a sphere object in the top layer (sphere.cpp:30-150, arealight.h:88-130)
    "cornell_alpha": (64, 64),      # alpha blending on + stencil materials: shadow rays re-traced through ignored hits (pathtracing_impl.h:266-351)
}
FRAME = 2
FRAMES = 3
N_RAYS = 4096


def gz_write(path, data: bytes):
    with gzip.GzipFile(path, "wb", mtime=0) as fp:
        fp.write(data)


def random_rays(scene: sio.Scene, n: int, seed: int) -> tuple:
    """Incoherent rays: origins inside the scene box, uniform directions; t_max = inf for half of
    them, a finite bound for the rest (exercises the t_max culling path of aabb::hit)."""
    rng = np.random.default_rng(seed)
    lo = scene.config["scene_bbox_min"][0].astype(np.float64)
    hi = scene.config["scene_bbox_max"][0].astype(np.float64)
    org = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros(n, sio.RAY_DTYPE)
    rays["org"] = org
    rays["dir"] = d.astype(np.float32)
    diag = float(np.linalg.norm(hi - lo))
    tmax = np.full(n, np.finfo(np.float32).max, np.float32)
    tmax[n // 2:] = (rng.random(n - n // 2) * diag).astype(np.float32)
    return rays, tmax


SVGF_SCENES = ("cornell_mtrl", "instanced", "sponza_atrium")
SVGF_FRAMES = 3
SVGF_MOTION = (0.02, -0.03)     # constant motion vector fed to frames > 0 (screen fractions)


def make_svgf():
    """Film images of frames 0..2 and the AOV set of frame 2 from the stock aten::SVGFRenderer (one thread)."""
    for name in SVGF_SCENES:
        W, H = SCENES[name]
        imgs, aov = ol.ref_render_svgf(name, W, H, frames=SVGF_FRAMES, motion=SVGF_MOTION)
        gz_write(os.path.join(HERE, f"{name}.svgf.f32.gz"), np.stack(imgs).tobytes() + aov.tobytes())
        print(name, "svgf", [float(i[..., :3].mean()) for i in imgs])


RESTIR_SCENES = ("cornell_mtrl", "instanced", "sponza_atrium")
RESTIR_FRAMES = 4


def make_restir():
    """Contribution images of frames 0..3 and the final reservoirs from the reference's own ReSTIR steps run in the
    device renderer's order (ref_harness restir), with the same constant motion vector as the SVGF fixtures."""
    for name in RESTIR_SCENES:
        W, H = SCENES[name]
        imgs, res = ol.ref_render_restir(name, W, H, frames=RESTIR_FRAMES, motion=SVGF_MOTION)
        gz_write(os.path.join(HERE, f"{name}.restir.f32.gz"), np.stack(imgs).tobytes() + res.tobytes())
        print(name, "restir", [float(i[..., :3].mean()) for i in imgs])


def main():
    if not ol.have_ref():
        sys.exit("oracle/_ref is not built: run python oracle/build_ref.py")
    if len(sys.argv) > 1 and sys.argv[1] == "svgf":     # only the SVGF fixtures (the others are unchanged)
        make_svgf()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "restir":   # only the ReSTIR fixtures
        make_restir()
        return
    meta = {}
    only = sys.argv[2:] if len(sys.argv) > 2 and sys.argv[1] == "only" else None     # `only <scene> ...`: add scenes, keep the rest
    if only:
        meta = json.load(open(os.path.join(HERE, "meta.json")))
    else:
        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "cmj.bin")
            ol.ref_run("cmj", p)
            gz_write(os.path.join(HERE, "cmj.f32.gz"), open(p, "rb").read())
    for name, (W, H) in SCENES.items():
        if only and name not in only:
            continue
        with tempfile.TemporaryDirectory() as d:
            ps = os.path.join(d, "s.atnscene")
            info = ol.ref_run("scene", name, W, H, ps)
            gz_write(os.path.join(HERE, f"{name}.atnscene.gz"), open(ps, "rb").read())
            scene = sio.load_scene(ps)
        recs, img1, tinfo = ol.ref_trace(name, W, H, frame=FRAME)
        gz_write(os.path.join(HERE, f"{name}.trace.gz"), recs.tobytes())
        img, rinfo = ol.ref_render(name, W, H, frames=FRAMES)
        gz_write(os.path.join(HERE, f"{name}.image.f32.gz"), img.tobytes())
        rays, tmax = random_rays(scene, N_RAYS, seed=hash(name) % 1000 if False else len(name) * 7 + 1)
        hits = ol.ref_rays(name, rays, tmax)
        buf = np.zeros((N_RAYS, 7), np.float32)
        buf[:, :3] = rays["org"]; buf[:, 3:6] = rays["dir"]; buf[:, 6] = tmax
        gz_write(os.path.join(HERE, f"{name}.rays.gz"), buf.tobytes())
        gz_write(os.path.join(HERE, f"{name}.hits.gz"), hits.tobytes())
        ao, _ = ol.ref_render_ao(name, W, H, frame=FRAME)
        gz_write(os.path.join(HERE, f"{name}.ao.f32.gz"), ao.tobytes())
        meta[name] = {"w": W, "h": H, "frame": FRAME, "frames": FRAMES, "depth": 5, "rr": 3, "scene": info,
                      "closest_rays": tinfo["closest_rays"], "shadow_rays": tinfo["shadow_rays"],
                      "hit_fraction": float((hits["objid"] >= 0).mean())}
        print(name, meta[name])
    json.dump(meta, open(os.path.join(HERE, "meta.json"), "w"), indent=1, sort_keys=True)
    if only:
        return
    make_svgf()
    make_restir()


if __name__ == "__main__":
    main()
