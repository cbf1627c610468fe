#!/usr/bin/env python3
"""Throughput of the hot path on the other BASELINE.json configs (stand-ins where the snapshot lacks
the assets).  Scenes are built on the spot by the reference's own loaders / SBVH builder
through oracle/_ref/ref_harness, so this runs only where that binary travelled (the GPU box).

    python scripts/bench_configs.py [--quick] > profiles/rNN_configs.jsonl
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from aten_b200 import PathTracing, scene_io as sio  # noqa: E402
from tests import oracle_lib as ol  # noqa: E402


def render_bench(name, W, H, spp, steps=3, warmup=2, label=""):
    scene = ol.ref_scene(name, W, H)
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    r.set_gather_mode(2)
    for _ in range(warmup):
        r.reset(); r.render(W, H, spp, 5, 3)
    ms = rays = samples = nodes = tris = 0
    for _ in range(steps):
        r.reset(); r.render(W, H, spp, 5, 3)
        st = r.stats()
        ms += st["device_ms"]; rays += st["closest_rays"] + st["shadow_rays"]; samples += st["camera_samples"]
    r.set_profiling(2)             # node / triangle counts come from the counting form of the kernels, outside the timing
    r.reset(); r.render(W, H, spp, 5, 3)
    st = r.stats(); nodes = st["nodes_visited"] * steps; tris = st["tris_tested"] * steps
    r.set_profiling(False)
    r.reset(); r.render(W, H, spp, 5, 3)
    img = r.readback()
    r.close()
    out = {"config": label, "scene": name, "w": W, "h": H, "spp": spp, "bounces": 5, "triangles": int(len(scene.triangles)),
           "nodes": scene.node_count, "instances": int((scene.objects["type"] == 1).sum()),
           "ms_per_step": round(ms / steps, 3), "mrays_per_s": round(rays / ms / 1e3, 1),
           "msamples_per_s": round(samples / ms / 1e3, 1), "rays_per_sample": round(rays / samples, 2),
           "quad_nodes_per_ray": round(nodes / rays, 2), "tri_tests_per_ray": round(tris / rays, 2),
           "mean_radiance": float(img[..., :3].mean() / spp), "finite": bool(np.isfinite(img).all())}
    return out


def ao_bench(name, W, H, spp, steps=3, warmup=2, label=""):
    """Row N1: the ambient-occlusion renderer (1 AO ray of radius 1 per sample, as the reference fixes it)."""
    scene = ol.ref_scene(name, W, H)
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    r.set_gather_mode(2)
    for _ in range(warmup):
        r.reset(); r.render_ao(W, H, spp, 1, 1.0)
    ms = rays = samples = 0
    for _ in range(steps):
        r.reset(); r.render_ao(W, H, spp, 1, 1.0)
        st = r.stats()
        ms += st["device_ms"]; rays += st["closest_rays"] + st["shadow_rays"]; samples += st["camera_samples"]
    img = r.readback()
    r.close()
    t0 = time.perf_counter()
    ref, info = ol.ref_render_ao(name, 480, 270, frame=0, threads=os.cpu_count() or 1)
    print(json.dumps({"config": label, "renderer": "ao", "scene": name, "w": W, "h": H, "spp": spp,
                      "ms_per_step": round(ms / steps, 3), "mrays_per_s": round(rays / ms / 1e3, 1),
                      "msamples_per_s": round(samples / ms / 1e3, 1), "mean_ao": float(img[..., 0].mean() / spp),
                      "reference_cpu": {"msamples_per_s": round(480 * 270 / info["render_ms"] / 1e3, 4),
                                        "sample": f"stock aten::AORenderer, {name} 480x270, 1 frame, {info['threads']} threads"}}), flush=True)


def svgf_bench(name, W, H, frames=8, warmup=1, label=""):
    """Row N1: the SVGF renderer, a sequence of `frames` frames (1 sample per pixel each) from frame 0.  The
    image-space half (prepare, temporal reprojection, variance estimation, 5 a-trous passes, copy) is timed as the
    difference to the path-traced sample alone; its algorithmic traffic is 690 B per pixel and frame
    (DESIGN.md "SVGF")."""
    scene = ol.ref_scene(name, W, H)
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    r.set_gather_mode(2)
    for _ in range(warmup):
        r.set_frame(0)
        for f in range(frames):
            r.render_svgf(W, H, 5, 3)
    r.set_frame(0)
    ms = rays = 0
    for f in range(frames):
        r.render_svgf(W, H, 5, 3)
        st = r.stats()
        if f > 0:       # frame 0 has no temporal pass
            ms += st["device_ms"]; rays += st["closest_rays"] + st["shadow_rays"]
    img = r.readback()
    pt_ms = 0
    for f in range(1, frames):
        r.set_frame(f); r.render(W, H, 1, 5, 3)
        pt_ms += r.stats()["device_ms"]
    r.close()
    n = frames - 1
    filt_ms = (ms - pt_ms) / n
    info = {}
    ol.ref_render_svgf(name, 480, 270, frames=3, threads=os.cpu_count() or 1, info=info)
    print(json.dumps({"config": label, "renderer": "svgf", "scene": name, "w": W, "h": H, "frames": frames,
                      "ms_per_frame": round(ms / n, 3), "path_tracing_ms": round(pt_ms / n, 3), "filter_ms": round(filt_ms, 3),
                      "frames_per_s": round(1e3 * n / ms, 1), "msamples_per_s": round(W * H * n / ms / 1e3, 1),
                      "mrays_per_s": round(rays / ms / 1e3, 1),
                      "filter_algorithmic_gb_per_s": round(690.0 * W * H / (filt_ms * 1e-3) / 1e9, 1),
                      "finite": bool(np.isfinite(img).all()),
                      "reference_cpu": {"msamples_per_s": round(480 * 270 * 3 / info["render_ms"] / 1e3, 4),
                                        "sample": f"stock aten::SVGFRenderer, {name} 480x270, 3 frames, {info['threads']} threads"}}), flush=True)


def restir_bench(name, W, H, frames=8, warmup=1, label=""):
    """Row N1: the ReSTIR renderer, a sequence of `frames` frames (1 sample per pixel each) from frame 0."""
    scene = ol.ref_scene(name, W, H)
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    r.set_gather_mode(2)
    for _ in range(warmup):
        r.set_frame(0)
        for f in range(frames):
            r.render_restir(W, H, 5, 3)
    r.set_frame(0)
    ms = rays = 0
    for f in range(frames):
        r.render_restir(W, H, 5, 3)
        st = r.stats()
        if f > 1:       # frames 0 and 1 have no temporal pass
            ms += st["device_ms"]; rays += st["closest_rays"] + st["shadow_rays"]
    img = r.readback()
    pt_ms = 0
    for f in range(2, frames):
        r.set_frame(f); r.render(W, H, 1, 5, 3)
        pt_ms += r.stats()["device_ms"]
    r.close()
    n = frames - 2
    import time
    t0 = time.perf_counter()
    ol.ref_render_restir(name, 480, 270, frames=3)
    ref_s = time.perf_counter() - t0
    print(json.dumps({"config": label, "renderer": "restir", "scene": name, "w": W, "h": H, "frames": frames,
                      "ms_per_frame": round(ms / n, 3), "path_tracer_same_frame_ms": round(pt_ms / n, 3),
                      "frames_per_s": round(1e3 * n / ms, 1), "msamples_per_s": round(W * H * n / ms / 1e3, 1),
                      "mrays_per_s": round(rays / ms / 1e3, 1), "finite": bool(np.isfinite(img).all()),
                      "reference_cpu": {"msamples_per_s": round(480 * 270 * 3 / ref_s / 1e6, 4),
                                        "sample": f"the reference's ReSTIR steps (ref_harness restir), {name} 480x270, 3 frames, 1 thread, incl. scene build"}}), flush=True)


def ray_stress(name, n_rays, label=""):
    """Incoherent diffuse secondary rays: origins = surface points found by a first random pass,
    directions cosine-distributed about nothing in particular (uniform sphere), closest-hit and any-hit."""
    import torch
    scene = ol.ref_scene(name, 8, 8)
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    rng = np.random.default_rng(1)
    lo = scene.config["scene_bbox_min"][0]; hi = scene.config["scene_bbox_max"][0]
    seed_n = min(n_rays, 2_000_000)
    rays = np.zeros(seed_n, sio.RAY_DTYPE)
    rays["org"] = (lo + (hi - lo) * rng.random((seed_n, 3))).astype(np.float32)
    d = rng.normal(size=(seed_n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["dir"] = d.astype(np.float32)
    hits = r.trace_closest(rays)
    ok = hits["objid"] >= 0
    # secondary rays start on the surfaces that were hit, pushed back a little along the incoming ray
    p = rays["org"][ok] + rays["dir"][ok] * (hits["t"][ok] * np.float32(0.999))[:, None]
    reps = int(np.ceil(n_rays / max(1, len(p))))
    org = np.tile(p, (reps, 1))[:n_rays]
    d = rng.normal(size=(n_rays, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    sec = np.zeros(n_rays, sio.RAY_DTYPE)
    sec["org"] = org.astype(np.float32); sec["dir"] = d.astype(np.float32)
    diag = float(np.linalg.norm(hi - lo))
    dist = (rng.random(n_rays) * diag).astype(np.float32)
    t_rays = torch.from_numpy(sec.view(np.float32).reshape(-1, 6)).cuda()
    t_dist = torch.from_numpy(dist).cuda()
    t_hits = torch.empty((n_rays, 8), dtype=torch.float32, device="cuda")
    t_vis = torch.empty(n_rays, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    res = {}
    for kind in ("closest", "any"):
        best = 1e30
        for it in range(4):          # first pass is the warm-up; wall clock around launch + sync (tens of ms per pass)
            r.synchronize()
            t0 = time.perf_counter()
            if kind == "closest":
                r.trace_closest_device(t_rays.data_ptr(), None, n_rays, t_hits.data_ptr())
            else:
                r.trace_any_device(t_rays.data_ptr(), t_dist.data_ptr(), n_rays, t_vis.data_ptr())
            r.synchronize()
            if it > 0:
                best = min(best, (time.perf_counter() - t0) * 1e3)
        res[kind] = best
    hit_frac = float((t_hits[:, 1].view(torch.int32) >= 0).float().mean())
    r.close()
    print(json.dumps({"config": label, "scene": name, "rays": n_rays, "triangles": int(len(scene.triangles)),
                      "instances": int((scene.objects["type"] == 1).sum()), "nodes": scene.node_count,
                      "closest_mrays_per_s": round(n_rays / res["closest"] / 1e3, 1),
                      "any_mrays_per_s": round(n_rays / res["any"] / 1e3, 1), "hit_fraction": round(hit_frac, 3),
                      "visible_fraction": round(float(t_vis.float().mean()), 3)}), flush=True)


def ray_stress_flat(name, n_rays, label="", parity_rays=200_000, copies=None, render=True):
    """configs[4] proper: 100 DISTINCT dragon meshes (scripts/flat_scene.py: own vertices, triangles and bottom-layer
    node list per copy -- ~1.4 GB of QuadNodes + TriLeafs, not L2-resident), rays that start on the surfaces
    (area-uniform origins, cosine-distributed directions).  Reports closest-hit and any-hit Mrays/s, the bytes the
    traversal fetched on its own layout (128 B per QuadNode step, 64 B per triangle test, 24 B ray in, 32 B hit out)
    against the measured HBM copy peak, and a parity check of a sample of the same rays against the CPU restatement."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "scripts"))
    import flat_scene as fs
    t0 = time.perf_counter()
    src = ol.ref_scene(name, 8, 8)
    t1 = time.perf_counter()
    scene = fs.flatten_instances(src, max_copies=copies)
    del src
    t2 = time.perf_counter()
    r = PathTracing(0)
    r.UpdateSceneData(scene)
    t3 = time.perf_counter()
    # Renderer::updateBVH on this scene: list 0 + objects + matrices handed over again (host wall clock, best of 3)
    upd = 1e30
    for _ in range(3):
        tu = time.perf_counter()
        r.updateBVH(scene.node_lists[0], scene.objects, scene.matrices)
        r.synchronize()
        upd = min(upd, time.perf_counter() - tu)
    rays, dist = fs.surface_rays(scene, n_rays, seed=1)
    hits = torch.empty((n_rays, 8), dtype=torch.float32, device="cuda")
    vis = torch.empty(n_rays, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize()
    res = {}
    for kind in ("closest", "any"):
        best = None
        for it in range(4):          # first pass is the warm-up; CUDA events on the handle's stream
            if kind == "closest":
                r.trace_closest_device(rays.data_ptr(), None, n_rays, hits.data_ptr())
            else:
                r.trace_any_device(rays.data_ptr(), dist.data_ptr(), n_rays, vis.data_ptr())
            st = r.stats()
            if it > 0 and (best is None or st["device_ms"] < best["device_ms"]):
                best = st
        # node / triangle counts: one more pass with the counting form of the kernel (atn_set_profiling), not timed
        r.set_profiling(2)
        if kind == "closest":
            r.trace_closest_device(rays.data_ptr(), None, n_rays, hits.data_ptr())
        else:
            r.trace_any_device(rays.data_ptr(), dist.data_ptr(), n_rays, vis.data_ptr())
        cst = r.stats()
        r.set_profiling(False)
        best = dict(best); best["nodes_visited"] = cst["nodes_visited"]; best["tris_tested"] = cst["tris_tested"]
        res[kind] = best
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 0) or 0)

    def line(st, out_bytes):
        b = 128 * st["nodes_visited"] + 64 * st["tris_tested"] + (24 + out_bytes) * n_rays
        gbs = b / st["device_ms"] / 1e6
        return {"mrays_per_s": round(n_rays / st["device_ms"] / 1e3, 1), "ms": round(st["device_ms"], 2),
                "quad_nodes_per_ray": round(st["nodes_visited"] / n_rays, 2), "tri_tests_per_ray": round(st["tris_tested"] / n_rays, 2),
                "fetched_gb_per_s": round(gbs, 1), "frac_of_hbm_copy_peak": round(gbs / hbm, 3) if hbm else None}
    # parity of a sample of exactly these rays
    k = min(parity_rays, n_rays)
    sel = torch.arange(k, device="cuda", dtype=torch.int64) * (n_rays // k)
    h_rays = np.zeros(k, sio.RAY_DTYPE)
    rr = rays[sel].cpu().numpy()
    h_rays["org"] = rr[:, 0:3]; h_rays["dir"] = rr[:, 3:6]
    got = np.frombuffer(hits[sel].cpu().numpy().tobytes(), sio.HIT_DTYPE)
    want = ol.trace_closest(scene, h_rays)
    hit = want["objid"] >= 0
    differ = (got["objid"] != want["objid"]) | (hit & (got["primid"] != want["primid"])) | (got["t"].view(np.uint32) != want["t"].view(np.uint32))
    hitmiss = int(((got["objid"] >= 0) != hit).sum())
    rel = float(np.max(np.abs(got["t"][differ & hit].astype(np.float64) - want["t"][differ & hit]) / want["t"][differ & hit])) if (differ & hit).any() else 0.0
    out = {"config": label, "scene": name + " flattened", "rays": n_rays, "triangles": int(len(scene.triangles)),
           "meshes": int(len(scene.node_lists) - 1), "reference_nodes": scene.node_count,
           "closest": line(res["closest"], 32), "any": line(res["any"], 1),
           "hit_fraction": round(float((hits[:, 1].view(torch.int32) >= 0).float().mean()), 3),
           "visible_fraction": round(float(vis.float().mean()), 3),
           "hbm_copy_peak_gbs": hbm or None,
           "parity_vs_restatement": {"rays": k, "differing": int(differ.sum()), "hit_miss_disagreements": hitmiss, "max_rel_dt": rel},
           "seconds": {"reference_build": round(t1 - t0, 1), "flatten": round(t2 - t1, 1), "upload_encode": round(t3 - t2, 1)},
           "top_layer_update_ms": round(1e3 * upd, 3)}
    # the same scene through the whole path tracer (point light + sky IBL as the instanced stand-in has them)
    del rays, dist, hits, vis
    torch.cuda.empty_cache()
    W, H, spp = 1920, 1080, 4
    r.close()
    if not render:
        return out
    r = PathTracing(0)
    r.UpdateSceneData(scene.with_size(W, H))
    r.set_gather_mode(2)
    ms = nr = ns = 0
    for it in range(3):
        r.reset(); r.render(W, H, spp, 5, 3)
        st = r.stats()
        if it > 0:
            ms += st["device_ms"]; nr += st["closest_rays"] + st["shadow_rays"]; ns += st["camera_samples"]
    img = r.readback()
    out["render_1080p_4spp"] = {"mrays_per_s": round(nr / ms / 1e3, 1), "msamples_per_s": round(ns / ms / 1e3, 1),
                                "rays_per_sample": round(nr / ns, 2), "finite": bool(np.isfinite(img).all()),
                                "mean_radiance": float(img[..., :3].mean() / spp)}
    r.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--rays", type=int, default=0, help="--only flat: number of rays (default 100 M)")
    ap.add_argument("--copies", type=int, default=0, help="--only flat: number of distinct meshes (default all 100)")
    a = ap.parse_args()
    if not ol.have_ref():
        sys.exit("oracle/_ref not present")
    q = a.quick
    if a.only == "restir":
        restir_bench("sponza_atrium", 1920, 1080, frames=8 if not q else 4, label="N1 ReSTIR, sponza_lod 1080p, 8-frame sequence")
        return
    if a.only == "flat":
        print(json.dumps(ray_stress_flat("dragon_grid10", a.rays or (100_000_000 if not q else 5_000_000), copies=a.copies or (None if not q else 9),
                        label="C5: incoherent secondary rays into 100 distinct dragon meshes (10 M triangles, ~1.4 GB of nodes + triangles)"
                        if not (a.copies or q) else f"incoherent secondary rays into {a.copies or 9} distinct dragon meshes")), flush=True)
        return
    if a.only in ("ao", "small"):
        ao_bench("sponza_atrium", 1920, 1080, 16 if not q else 2, label="N1 ambient occlusion, sponza_lod 1080p 16spp")
        if a.only == "ao":
            return
    if a.only == "small":       # the configs whose wavefronts are one sample of a frame
        print(json.dumps(render_bench("cornell", 512, 512, 1, steps=5, warmup=3, label="C1 Cornell 512x512 1spp")), flush=True)
        svgf_bench("sponza_atrium", 1920, 1080, frames=8, label="N1 SVGF, sponza_lod 1080p, 8-frame sequence")
        restir_bench("sponza_atrium", 1920, 1080, frames=8, label="N1 ReSTIR, sponza_lod 1080p, 8-frame sequence")
        return
    if a.only == "svgf":
        svgf_bench("sponza_atrium", 1920, 1080, frames=8 if not q else 3, label="N1 SVGF, sponza_lod 1080p, 8-frame sequence")
        return
    print(json.dumps(render_bench("cornell", 512, 512, 1, steps=5, warmup=3, label="C1 Cornell 512x512 1spp")), flush=True)
    print(json.dumps(render_bench("sponza_atrium", 1920, 1080, 16 if not q else 2, label="C2 sponza_lod 1080p 16spp (bench.py workload)")), flush=True)
    print(json.dumps(render_bench("sponza_tex", 1920, 1080, 16 if not q else 2, label="C3 stand-in: sponza_lod + albedo/normal maps + synthetic-sky IBL, 1080p 16spp")), flush=True)
    print(json.dumps(render_bench("instanced_dragon", 3840, 2160, 8 if not q else 1, label="C4 stand-in: instanced dragon+teapot+suzanne, Disney/refraction, 4K, 8 spp (one GPU's share of 64 spp on 8)")), flush=True)
    print(json.dumps(render_bench("dragon_grid10", 1920, 1080, 4 if not q else 1, label="C5 stand-in (rendered): 100 instanced dragons = 10 M triangles, 1080p 4spp")), flush=True)
    ao_bench("sponza_atrium", 1920, 1080, 16 if not q else 2, label="N1 ambient occlusion, sponza_lod 1080p 16spp")
    svgf_bench("sponza_atrium", 1920, 1080, frames=8 if not q else 3, label="N1 SVGF, sponza_lod 1080p, 8-frame sequence")
    restir_bench("sponza_atrium", 1920, 1080, frames=8 if not q else 4, label="N1 ReSTIR, sponza_lod 1080p, 8-frame sequence")
    ray_stress("dragon_grid10", 100_000_000 if not q else 5_000_000,
               label="C5 stand-in: 100 M incoherent secondary rays into 100 instanced dragons (10 M triangles), closest-hit and any-hit")
    print(json.dumps(ray_stress_flat("dragon_grid10", 100_000_000 if not q else 5_000_000, copies=None if not q else 9,
                    label="C5: 100 M incoherent secondary rays into 100 distinct dragon meshes (10 M triangles, ~1.4 GB of nodes + triangles)")), flush=True)


if __name__ == "__main__":
    main()
