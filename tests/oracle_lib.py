"""TEST INFRASTRUCTURE: ctypes access to the checkers under oracle/ -- the CPU restatement
(oracle/_build/libpt_oracle.so) and the reference itself compiled in place (oracle/_ref/ref_harness).
Only tests/, __graft_entry__.smoke() and bench.py's reference arm use this module."""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess
import tempfile

import numpy as np

from aten_b200 import scene_io as sio

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_BIN = os.path.join(ORACLE_DIR, "_ref", "ref_harness")
REF_ASSET = os.path.join(ORACLE_DIR, "_ref", "asset")
_lib = None


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


def lib():
    global _lib
    if _lib is None:
        path = os.path.join(ORACLE_DIR, "_build", "libpt_oracle.so")
        if not os.path.exists(path):
            build_oracle()
        _lib = C.CDLL(path)
    return _lib


def trace_closest(scene: sio.Scene, rays: np.ndarray, t_max=None, counters=False):
    rays = np.ascontiguousarray(rays, dtype=sio.RAY_DTYPE)
    hits = np.zeros(len(rays), dtype=sio.HIT_DTYPE)
    blob = scene.blob()
    cnt = np.zeros(4, np.uint64)
    tm = None if t_max is None else np.ascontiguousarray(t_max, dtype=np.float32)
    lib().oracle_trace_closest(C.byref(blob), C.c_void_p(rays.ctypes.data),
                               C.c_void_p(tm.ctypes.data) if tm is not None else None,
                               C.c_size_t(len(rays)), C.c_void_p(hits.ctypes.data),
                               C.c_void_p(cnt.ctypes.data) if counters else None)
    return (hits, cnt) if counters else hits


def trace_any(scene: sio.Scene, rays: np.ndarray, dist: np.ndarray):
    rays = np.ascontiguousarray(rays, dtype=sio.RAY_DTYPE)
    dist = np.ascontiguousarray(dist, dtype=np.float32)
    vis = np.zeros(len(rays), dtype=np.uint8)
    blob = scene.blob()
    lib().oracle_trace_any(C.byref(blob), C.c_void_p(rays.ctypes.data), C.c_void_p(dist.ctypes.data),
                           C.c_size_t(len(rays)), C.c_void_p(vis.ctypes.data))
    return vis


def render(scene: sio.Scene, frames=1, max_depth=5, rr_depth=3, first_frame=0, records=False, image=None, threads=None, aov=None):
    """`aov`: optional dict that receives 'normal_depth' and 'albedo_meshid' float32 [H, W, 4] (first-hit AOVs)."""
    W, H = scene.width, scene.height
    if threads:
        lib().oracle_set_threads(int(threads))
    img = np.zeros((H * W, 4), np.float32) if image is None else np.ascontiguousarray(image, np.float32).reshape(H * W, 4)
    recs = np.zeros((H * W, max_depth), sio.BOUNCE_DTYPE) if records else None
    seeds = scene.seeds_for(W * H)
    stats = np.zeros(6, np.uint64)
    nd = np.zeros((H * W, 4), np.float32)
    am = np.zeros((H * W, 4), np.float32)
    blob = scene.blob()
    cam = np.ascontiguousarray(scene.camera)
    lib().oracle_render(C.byref(blob), C.c_void_p(cam.ctypes.data), C.c_void_p(seeds.ctypes.data), C.c_uint32(len(seeds)),
                        C.c_int32(W), C.c_int32(H), C.c_uint32(first_frame), C.c_int32(frames), C.c_int32(max_depth),
                        C.c_int32(rr_depth), C.c_void_p(img.ctypes.data),
                        C.c_void_p(recs.ctypes.data) if records else None, C.c_void_p(stats.ctypes.data),
                        C.c_void_p(nd.ctypes.data) if aov is not None else None,
                        C.c_void_p(am.ctypes.data) if aov is not None else None)
    if aov is not None:
        aov["normal_depth"] = nd.reshape(H, W, 4)
        aov["albedo_meshid"] = am.reshape(H, W, 4)
    return img.reshape(H, W, 4), recs, stats


def render_ao(scene: sio.Scene, frames=1, first_frame=0, num_rays=1, radius=1.0, image=None, host_row_break=False):
    W, H = scene.width, scene.height
    img = np.zeros((H * W, 4), np.float32) if image is None else np.ascontiguousarray(image, np.float32).reshape(H * W, 4)
    seeds = scene.seeds_for(W * H)
    blob = scene.blob()
    cam = np.ascontiguousarray(scene.camera)
    lib().oracle_render_ao(C.byref(blob), C.c_void_p(cam.ctypes.data), C.c_void_p(seeds.ctypes.data), C.c_uint32(len(seeds)),
                           C.c_int32(W), C.c_int32(H), C.c_uint32(first_frame), C.c_int32(frames), C.c_int32(num_rays),
                           C.c_float(radius), C.c_int32(1 if host_row_break else 0), C.c_void_p(img.ctypes.data))
    return img.reshape(H, W, 4)


def ref_render_ao(name, W, H, frame=0, threads=8):
    """One frame of the stock aten::AORenderer at frame index `frame` (film starts at (0,0,0,1))."""
    with tempfile.TemporaryDirectory() as d:
        pi = os.path.join(d, "i.f32")
        info = ref_run("ao", name, W, H, frame, threads, pi)
        img = np.fromfile(pi, np.float32).reshape(H, W, 4)
    return img, info


def cmj(idx, scramble, n):
    out = np.zeros(n, np.float32)
    lib().oracle_cmj(C.c_uint32(idx), C.c_uint32(scramble), C.c_int32(n), C.c_void_p(out.ctypes.data))
    return out


def camera_rays(scene: sio.Scene, frame=0):
    W, H = scene.width, scene.height
    rays = np.zeros(W * H, sio.RAY_DTYPE)
    seeds = scene.seeds_for(W * H)
    cam = np.ascontiguousarray(scene.camera)
    lib().oracle_camera_rays(C.c_void_p(cam.ctypes.data), C.c_void_p(seeds.ctypes.data), C.c_uint32(len(seeds)),
                             C.c_int32(W), C.c_int32(H), C.c_uint32(frame), C.c_void_p(rays.ctypes.data))
    return rays


# ---- the reference itself (oracle/_ref) ------------------------------------------------------
def have_ref() -> bool:
    return os.path.exists(REF_BIN) and os.path.isdir(REF_ASSET)


def ref_run(*args, timeout=3600):
    """Run oracle/_ref/ref_harness; returns the JSON line it prints."""
    cmd = [REF_BIN, REF_ASSET] + [str(a) for a in args]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=timeout)
    if r.returncode != 0:
        raise RuntimeError(f"ref_harness failed ({r.returncode}): {' '.join(cmd)}\n{r.stderr[-2000:]}")
    for line in reversed(r.stdout.strip().splitlines()):
        if line.startswith("{"):
            return json.loads(line)
    return {}


def ref_scene(name, W, H) -> sio.Scene:
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "s.atnscene")
        ref_run("scene", name, W, H, p)
        return sio.load_scene(p)


def ref_trace(name, W, H, frame=0, depth=5, rr=3):
    with tempfile.TemporaryDirectory() as d:
        pr, pi = os.path.join(d, "r.bin"), os.path.join(d, "i.f32")
        info = ref_run("trace", name, W, H, frame, depth, rr, pr, pi)
        recs = np.fromfile(pr, sio.BOUNCE_DTYPE).reshape(H * W, depth)
        img = np.fromfile(pi, np.float32).reshape(H, W, 4)
    return recs, img, info


def ref_render(name, W, H, frames=1, depth=5, rr=3, threads=8):
    with tempfile.TemporaryDirectory() as d:
        pi = os.path.join(d, "i.f32")
        info = ref_run("render", name, W, H, frames, depth, rr, threads, pi)
        img = np.fromfile(pi, np.float32).reshape(H, W, 4)
    return img, info


def ref_rays(name, rays: np.ndarray, t_max: np.ndarray):
    with tempfile.TemporaryDirectory() as d:
        pin, pout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        buf = np.zeros((len(rays), 7), np.float32)
        buf[:, 0:3] = rays["org"]; buf[:, 3:6] = rays["dir"]; buf[:, 6] = t_max
        buf.tofile(pin)
        ref_run("rays", name, pin, pout)
        return np.fromfile(pout, sio.HIT_DTYPE)


def fill_basic_aovs(normal, p, w2c, albedo, meshid):
    f = lambda a, n: (C.c_float * n)(*[float(x) for x in a])
    nd = (C.c_float * 4)(); am = (C.c_float * 4)()
    lib().oracle_fill_basic_aovs(f(normal, 3), f(p, 3), f(np.asarray(w2c).reshape(-1), 16), f(albedo, 4), C.c_int32(meshid), nd, am)
    return list(nd), list(am)


def fill_basic_aovs_miss(bg):
    nd = (C.c_float * 4)(); am = (C.c_float * 4)()
    lib().oracle_fill_basic_aovs_miss((C.c_float * 4)(*[float(x) for x in bg]), nd, am)
    return list(nd), list(am)


def pixel_width_at_distance(vfov, width, height, distance):
    fn = lib().oracle_pixel_width_at_distance
    fn.restype = C.c_float
    return float(fn(C.c_float(vfov), C.c_int32(width), C.c_int32(height), C.c_float(distance)))


# ---- N1: SVGF -------------------------------------------------------------------------------
SVGF_STATE_PLANES = 11    # aov[0] x4, aov[1] x4, a-trous ping / pong, temporary colour


def render_svgf(scene: sio.Scene, frames=1, max_depth=5, rr_depth=3, motion=None, host_order=False, state=None, first_frame=0, iters=5):
    """CPU restatement of aten::SVGFRenderer, frames first_frame .. first_frame+frames-1 (1 sample each).
    Returns (film image of the last frame [H,W,4], state [11,H*W,4]); pass `state` back in to continue."""
    W, H = scene.width, scene.height
    n = W * H
    st = np.zeros((SVGF_STATE_PLANES, n, 4), np.float32) if state is None else np.ascontiguousarray(state, np.float32)
    img = np.zeros((n, 4), np.float32)
    seeds = scene.seeds_for(n)
    md = None if motion is None else np.ascontiguousarray(motion, np.float32).reshape(n, 4)
    blob = scene.blob()
    cam = np.ascontiguousarray(scene.camera)
    rc = lib().oracle_render_svgf(C.byref(blob), C.c_void_p(cam.ctypes.data), C.c_void_p(seeds.ctypes.data), C.c_uint32(len(seeds)),
                                  C.c_int32(W), C.c_int32(H), C.c_uint32(first_frame), C.c_int32(frames), C.c_int32(max_depth),
                                  C.c_int32(rr_depth), C.c_void_p(md.ctypes.data) if md is not None else None,
                                  C.c_int32(1 if host_order else 0), C.c_int32(iters), C.c_void_p(st.ctypes.data), C.c_void_p(img.ctypes.data))
    assert rc == 0
    return img.reshape(H, W, 4), st


def svgf_last_aov(state, frames_rendered):
    """The AOV set [4, n, 4] (normal_depth, albedo_meshid, color_variance, moment_tw) of the last frame rendered."""
    k = (frames_rendered - 1) & 1
    return state[k * 4:(k + 1) * 4]


def ref_render_svgf(name, W, H, frames=1, depth=5, rr=3, motion=(0.0, 0.0), iters=5, threads=1, info=None):
    """Stock aten::SVGFRenderer on one thread: ([frames] film images [H,W,4], AOV set of the last frame [4,H*W,4])."""
    with tempfile.TemporaryDirectory() as d:
        prefix = os.path.join(d, "svgf")
        res = ref_run("svgf", name, W, H, frames, depth, rr, prefix, motion[0], motion[1], iters, threads)
        if info is not None:
            info.update(res)
        imgs = [np.fromfile(f"{prefix}.f{k}.f32", np.float32).reshape(H, W, 4) for k in range(frames)]
        aov = np.fromfile(f"{prefix}.aov.f32", np.float32).reshape(4, H * W, 4)
    return imgs, aov


# ---- N1: ReSTIR -----------------------------------------------------------------------------
def render_restir(scene: sio.Scene, frames=1, max_depth=5, rr_depth=3, motion=None):
    """CPU restatement of the ReSTIR sequence, frames 0 .. frames-1 (1 sample each).
    Returns (contribution image of the last frame [H,W,4], reservoirs [H*W,5] = w_sum, M, y, W, target pdf)."""
    W, H = scene.width, scene.height
    n = W * H
    img = np.zeros((n, 4), np.float32)
    res = np.zeros((n, 5), np.float32)
    seeds = scene.seeds_for(n)
    md = None if motion is None else np.ascontiguousarray(motion, np.float32).reshape(n, 4)
    blob = scene.blob()
    cam = np.ascontiguousarray(scene.camera)
    rc = lib().oracle_render_restir(C.byref(blob), C.c_void_p(cam.ctypes.data), C.c_void_p(seeds.ctypes.data), C.c_uint32(len(seeds)),
                                    C.c_int32(W), C.c_int32(H), C.c_int32(frames), C.c_int32(max_depth), C.c_int32(rr_depth),
                                    C.c_void_p(md.ctypes.data) if md is not None else None, C.c_void_p(img.ctypes.data),
                                    C.c_void_p(res.ctypes.data))
    assert rc == 0
    return img.reshape(H, W, 4), res


def ref_render_restir(name, W, H, frames=1, depth=5, rr=3, motion=(0.0, 0.0)):
    """The reference's own ReSTIR steps in the device renderer's order (ref_harness restir), one thread:
    ([frames] contribution images [H,W,4], reservoirs of the last frame [H*W,5])."""
    with tempfile.TemporaryDirectory() as d:
        prefix = os.path.join(d, "restir")
        ref_run("restir", name, W, H, frames, depth, rr, prefix, motion[0], motion[1])
        imgs = [np.fromfile(f"{prefix}.f{k}.f32", np.float32).reshape(H, W, 4) for k in range(frames)]
        res = np.fromfile(f"{prefix}.reservoirs.f32", np.float32).reshape(H * W, 5)
    return imgs, res


# ---- texture fetch modes -------------------------------------------------------------
def set_texture_fetch(device_semantics: bool):
    """Switch the restatement between the reference's host texel rule (default) and its GPU renderer's fetch semantics."""
    lib().oracle_set_texture_fetch(C.c_int32(1 if device_semantics else 0))


def tex_fetch(texels: np.ndarray, filter_mode: int, address_mode: int, uv: np.ndarray) -> np.ndarray:
    """tex_fetch_device (aten_b200/csrc/atn_math.cuh) on one float32 [h, w, 4] texture for uv [n, 2]."""
    t = np.ascontiguousarray(texels, np.float32)
    uv = np.ascontiguousarray(uv, np.float32)
    out = np.zeros((len(uv), 4), np.float32)
    lib().oracle_tex_fetch(C.c_void_p(t.ctypes.data), C.c_int32(t.shape[1]), C.c_int32(t.shape[0]), C.c_int32(filter_mode),
                           C.c_int32(address_mode), C.c_void_p(uv.ctypes.data), C.c_int32(len(uv)), C.c_void_p(out.ctypes.data))
    return out
