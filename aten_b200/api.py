"""Python host side of the C ABI (include/aten_b200.h) -- a thin ctypes mirror of the reference's
renderer object ``idaten::PathTracing`` (reference src/libidaten/kernel/pathtracing.h:107-152,
src/libidaten/kernel/renderer.h:61-116): same method names, argument meaning and call order as the
calls ``device_renderer`` makes (src/device_renderer/main.cpp:144-148, 199-202).

There is no CPU path: without the CUDA library or without a GPU every call raises ``AtnError``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import scene_io as sio

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libaten_b200.so")

EXPORTS = [
    "atn_create", "atn_destroy", "atn_last_error", "atn_version", "atn_upload_scene", "atn_update_camera",
    "atn_update_top_layer", "atn_update_materials", "atn_update_lights", "atn_update_config", "atn_set_seeds",
    "atn_set_frame", "atn_get_frame", "atn_reset", "atn_set_progressive", "atn_set_gather_mode", "atn_set_exact_ties", "atn_set_wavefront_paths", "atn_set_hit_distance_limit",
    "atn_render", "atn_render_region", "atn_render_tiles", "atn_synchronize", "atn_readback", "atn_device_image", "atn_get_stats",
    "atn_set_profiling", "atn_get_stream", "atn_trace_closest", "atn_trace_any", "atn_trace_closest_device",
    "atn_trace_any_device", "atn_set_bounce_tap", "atn_trace_closest_static", "atn_readback_aov", "atn_render_ao",
    "atn_render_svgf", "atn_set_motion_depth", "atn_readback_svgf", "atn_update_texture",
    "atn_render_restir", "atn_readback_restir", "atn_probe_read_bandwidth",
    "atn_upload_scene_cached", "atn_set_alpha_blending", "atn_set_texture_fetch", "atn_sbvh_read", "atn_sbvh_write", "atn_free",
]


GATHER_OVERWRITE, GATHER_PROGRESSIVE, GATHER_SUM = 0, 1, 2      # AtnGatherMode (include/aten_b200.h)


class AtnError(RuntimeError):
    pass


class AtnSbvhInfo(C.Structure):
    _fields_ = [("version", C.c_uint8 * 4), ("max_depth", C.c_uint32), ("voxel_material_count", C.c_uint32),
                ("boxmin", C.c_float * 3), ("boxmax", C.c_float * 3)]


def sbvh_read(path: str, offset_tri_idx: int = 0, material_names=None):
    """``sbvh::importTree`` through the C ABI (``atn_sbvh_read``): (nodes NODE_DTYPE, info dict).  No GPU needed."""
    lib = load_library()
    nodes = C.c_void_p()
    count = C.c_uint32()
    info = AtnSbvhInfo()
    names = None
    n_names = 0
    if material_names is not None:
        n_names = len(material_names)
        names = (C.c_char_p * max(1, n_names))(*[m.encode() for m in material_names])
    _check(lib, lib.atn_sbvh_read(C.c_char_p(os.fsencode(path)), C.c_int32(offset_tri_idx), names, C.c_uint32(n_names),
                                  C.byref(nodes), C.byref(count), C.byref(info)))
    try:
        buf = C.string_at(nodes.value, count.value * sio.NODE_DTYPE.itemsize)
        arr = np.frombuffer(buf, sio.NODE_DTYPE).copy()
    finally:
        lib.atn_free(nodes)
    return arr, {"version": tuple(info.version), "max_depth": int(info.max_depth), "voxel_materials": int(info.voxel_material_count),
                 "boxmin": np.array(info.boxmin[:], np.float32), "boxmax": np.array(info.boxmax[:], np.float32)}


def sbvh_write(path: str, nodes: np.ndarray, max_depth: int, boxmin, boxmax, voxel_materials=None):
    """``sbvh::exportTree``'s file layout through the C ABI (``atn_sbvh_write``)."""
    lib = load_library()
    nodes = np.ascontiguousarray(nodes, sio.NODE_DTYPE)
    vm = sorted((voxel_materials or {}).items())
    ids = (C.c_int32 * max(1, len(vm)))(*[k for k, _ in vm])
    names = (C.c_char_p * max(1, len(vm)))(*[v.encode() for _, v in vm])
    bmin = (C.c_float * 3)(*[float(x) for x in boxmin]); bmax = (C.c_float * 3)(*[float(x) for x in boxmax])
    _check(lib, lib.atn_sbvh_write(C.c_char_p(os.fsencode(path)), C.c_void_p(nodes.ctypes.data), C.c_uint32(len(nodes)),
                                   C.c_uint32(max_depth), bmin, bmax, ids, names, C.c_uint32(len(vm))))


def probe_read_bandwidth(device: int, nbytes: int, passes: int):
    """(GB/s, ms) of one persistent launch reading a resident buffer of `nbytes` `passes` times (measurement aid)."""
    lib = load_library()
    g, ms = C.c_float(), C.c_float()
    _check(lib, lib.atn_probe_read_bandwidth(C.c_int(device), C.c_size_t(nbytes), C.c_int(passes), C.byref(g), C.byref(ms)))
    return g.value, ms.value


class AtnStats(C.Structure):
    _fields_ = [("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64), ("camera_samples", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("device_ms", C.c_float), ("trace_ms", C.c_float),
                ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64), ("exact_rays", C.c_uint64)]


_lib = None


def load_library(path: str | None = None):
    """dlopen libaten_b200.so; fails loudly if it was not built (python -m aten_b200.build)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("ATEN_B200_LIB") or LIB_PATH     # env override: kernel-variant experiments only
    if not os.path.exists(p):
        raise AtnError(f"{p} is missing: build it with `python -m aten_b200.build` (needs nvcc); "
                       "aten_b200 has no CPU fallback")
    lib = C.CDLL(p)
    lib.atn_last_error.restype = C.c_char_p
    lib.atn_version.restype = C.c_char_p
    for name in EXPORTS:
        fn = getattr(lib, name)          # AttributeError if the symbol is not exported
        if name not in ("atn_last_error", "atn_version"):
            fn.restype = C.c_int
    if path is None:
        _lib = lib
    return lib


def _check(lib, rc):
    if rc != 0:
        raise AtnError(f"[{rc}] {lib.atn_last_error().decode(errors='replace')}")


class PathTracing:
    """``idaten::PathTracing`` over the C ABI.

    >>> r = PathTracing(device=0)
    >>> r.UpdateSceneData(scene)            # Renderer::UpdateSceneData
    >>> r.render(w, h, max_samples, max_bounce)
    >>> img = r.readback()                  # float32 [h, w, 4]; replaces the GL surface
    """

    def __init__(self, device: int = 0):
        self._lib = load_library()
        self._h = C.c_void_p()
        _check(self._lib, self._lib.atn_create(C.c_int(device), C.byref(self._h)))
        self.device = device
        self.width = self.height = 0
        self._scene = None
        self._tap = None

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.atn_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- Renderer::UpdateSceneData (reference renderer.cpp:12-131) ------------------------------
    def UpdateSceneData(self, scene: sio.Scene, width: int | None = None, height: int | None = None, seeds=None,
                        cache_path: str | None = None):
        """``cache_path``: file of the encoded bottom trees (``atn_upload_scene_cached``); ``self.cache_hit`` says
        afterwards whether it was used (True) or (re)written (False)."""
        width = width or scene.width
        height = height or scene.height
        blob = scene.blob()
        cam = np.ascontiguousarray(scene.camera)
        self.cache_hit = None
        if cache_path:
            hit = C.c_int(0)
            _check(self._lib, self._lib.atn_upload_scene_cached(self._h, C.byref(blob), C.c_void_p(cam.ctypes.data),
                                                                C.c_int32(width), C.c_int32(height),
                                                                C.c_char_p(os.fsencode(cache_path)), C.byref(hit)))
            self.cache_hit = bool(hit.value)
        else:
            _check(self._lib, self._lib.atn_upload_scene(self._h, C.byref(blob), C.c_void_p(cam.ctypes.data),
                                                         C.c_int32(width), C.c_int32(height)))
        self.width, self.height = width, height
        self._scene = scene
        s = scene.seeds_for(width * height) if seeds is None else np.ascontiguousarray(seeds, np.uint32)
        self.set_seeds(s)

    def set_seeds(self, seeds: np.ndarray):
        seeds = np.ascontiguousarray(seeds, np.uint32)
        _check(self._lib, self._lib.atn_set_seeds(self._h, C.c_void_p(seeds.ctypes.data), C.c_uint32(len(seeds))))

    def updateCamera(self, camera: np.ndarray):
        cam = np.ascontiguousarray(camera, sio.CAMERA_DTYPE)
        _check(self._lib, self._lib.atn_update_camera(self._h, C.c_void_p(cam.ctypes.data)))

    def updateBVH(self, top_nodes: np.ndarray, objects: np.ndarray, matrices: np.ndarray):
        top = np.ascontiguousarray(top_nodes, sio.NODE_DTYPE)
        obj = np.ascontiguousarray(objects, sio.OBJ_DTYPE)
        mtx = np.ascontiguousarray(matrices, np.float32).reshape(-1, 16)
        _check(self._lib, self._lib.atn_update_top_layer(
            self._h, C.c_void_p(top.ctypes.data), C.c_uint32(len(top)), C.c_void_p(obj.ctypes.data),
            C.c_uint32(len(obj)), C.c_void_p(mtx.ctypes.data), C.c_uint32(len(mtx))))

    def updateMaterial(self, materials: np.ndarray):
        m = np.ascontiguousarray(materials, sio.MTRL_DTYPE)
        _check(self._lib, self._lib.atn_update_materials(self._h, C.c_void_p(m.ctypes.data), C.c_uint32(len(m))))

    def updateLight(self, lights: np.ndarray):
        l = np.ascontiguousarray(lights, sio.LIGHT_DTYPE)
        _check(self._lib, self._lib.atn_update_lights(self._h, C.c_void_p(l.ctypes.data), C.c_uint32(len(l))))

    def UpdateTexture(self, idx: int, texels: np.ndarray):
        """idaten::Renderer::UpdateTexture (reference src/libidaten/kernel/renderer.cpp:236-247): float32 [h, w, 4]
        texels for texture `idx`, same size as at upload."""
        t = np.ascontiguousarray(texels, np.float32)
        if t.ndim != 3 or t.shape[2] != 4:
            raise ValueError("texels must be float32 [h, w, 4]")
        d = sio.AtnTextureDesc(C.c_void_p(t.ctypes.data), t.shape[1], t.shape[0], 0, 0)
        _check(self._lib, self._lib.atn_update_texture(self._h, C.c_uint32(idx), C.byref(d)))

    def UpdateSceneRenderingConfig(self, config: np.ndarray):
        c = np.ascontiguousarray(config, sio.CONFIG_DTYPE)
        _check(self._lib, self._lib.atn_update_config(self._h, C.c_void_p(c.ctypes.data)))

    # -- frame control ---------------------------------------------------------------------------
    def reset(self):
        _check(self._lib, self._lib.atn_reset(self._h))

    def GetFrameCount(self) -> int:
        f = C.c_uint32()
        _check(self._lib, self._lib.atn_get_frame(self._h, C.byref(f)))
        return f.value

    def set_frame(self, frame: int):
        _check(self._lib, self._lib.atn_set_frame(self._h, C.c_uint32(frame)))

    def SetEnableProgressive(self, b: bool):
        """PathTracing::SetEnableProgressive (reference pathtracing.h:127-135): True = running mean across calls
        (GATHER_PROGRESSIVE), False = every call overwrites the image with the mean of its own samples (GATHER_OVERWRITE)."""
        _check(self._lib, self._lib.atn_set_progressive(self._h, C.c_int(1 if b else 0)))

    def set_gather_mode(self, mode: int):
        """AtnGatherMode: GATHER_OVERWRITE / GATHER_PROGRESSIVE / GATHER_SUM (sum of samples + count in .w, for the
        multi-GPU reduce)."""
        _check(self._lib, self._lib.atn_set_gather_mode(self._h, C.c_int(mode)))

    def set_texture_fetch(self, device_semantics: bool):
        """False (default): the reference's host texel rule (parity runs); True: its GPU renderer's tex2DLod semantics with
        each texture's filter / address mode, in software (parity unpinned, see include/aten_b200.h)."""
        _check(self._lib, self._lib.atn_set_texture_fetch(self._h, C.c_int(1 if device_semantics else 0)))

    def set_alpha_blending(self, b: bool):
        """SceneRenderingConfig::enable_alpha_blending (also taken from the scene at upload): shadow rays restart behind
        surfaces that are translucent by alpha, up to 10 lookups (HitTestToTargetLight)."""
        _check(self._lib, self._lib.atn_set_alpha_blending(self._h, C.c_int(1 if b else 0)))

    def set_exact_ties(self, b: bool):
        """True (default): rays whose two nearest primitives tie within the error of the reference's biased box test are
        decided by the reference's own depth-first walk (bit-identical primitive ids); False: they keep the nearest."""
        _check(self._lib, self._lib.atn_set_exact_ties(self._h, C.c_int(1 if b else 0)))

    def set_wavefront_paths(self, max_paths: int):
        """Upper bound on the paths of one wavefront (memory: ~265 B per path; default 32 Mi).  Before UpdateSceneData."""
        _check(self._lib, self._lib.atn_set_wavefront_paths(self._h, C.c_uint64(int(max_paths))))

    def setHitDistanceLimit(self, d: float):
        _check(self._lib, self._lib.atn_set_hit_distance_limit(self._h, C.c_float(d)))

    def set_profiling(self, b):
        """False / True (= 1): CUDA events around the traversal kernels (stats trace_ms); 2: count node fetches and triangle
        tests (stats nodes_visited, tris_tested; a slightly slower form of the kernels); 3: both."""
        _check(self._lib, self._lib.atn_set_profiling(self._h, C.c_int(int(b))))

    # -- render ----------------------------------------------------------------------------------
    def render(self, width: int, height: int, maxSamples: int, maxBounce: int, rrBounce: int = 3):
        """PathTracingImplBase::render (reference pathtracing.cpp:49-89); rrBounce = 3 is the
        reference's constant (pathtracing.cpp:97)."""
        _check(self._lib, self._lib.atn_render(self._h, C.c_int32(width), C.c_int32(height), C.c_int32(maxSamples),
                                               C.c_int32(maxBounce), C.c_int32(rrBounce)))

    def render_ao(self, width: int, height: int, maxSamples: int, num_rays: int = 1, radius: float = 1.0):
        """idaten::AORenderer::render (reference src/libidaten/ao/ao.cpp:9-51); the reference fixes
        num_rays = 1 and radius = 1."""
        _check(self._lib, self._lib.atn_render_ao(self._h, C.c_int32(width), C.c_int32(height), C.c_int32(maxSamples),
                                                  C.c_int32(num_rays), C.c_float(radius)))

    def render_svgf(self, width: int, height: int, maxBounce: int, rrBounce: int = 3, atrous_iters: int = 5):
        """One frame of idaten::SVGFPathTracing::render (reference src/libidaten/svgf/svgf.cpp:28-160): 1 sample per
        pixel + temporal reprojection + variance estimation + a-trous filter.  Frame 0 starts a sequence."""
        _check(self._lib, self._lib.atn_render_svgf(self._h, C.c_int32(width), C.c_int32(height), C.c_int32(maxBounce),
                                                    C.c_int32(rrBounce), C.c_int32(atrous_iters)))

    def render_restir(self, width: int, height: int, maxBounce: int, rrBounce: int = 3):
        """One frame of idaten::ReSTIRPathTracing::OnRender (reference src/libidaten/restir/restir.cpp:47-123)."""
        _check(self._lib, self._lib.atn_render_restir(self._h, C.c_int32(width), C.c_int32(height), C.c_int32(maxBounce), C.c_int32(rrBounce)))

    def readback_restir(self) -> np.ndarray:
        """float32 [h*w, 5]: w_sum, M, y, W, target pdf of the reservoirs behind the last frame's pixel colours."""
        n = self.width * self.height
        out = np.empty((n, 5), np.float32)
        _check(self._lib, self._lib.atn_readback_restir(self._h, C.c_void_p(out.ctypes.data), C.c_size_t(n)))
        return out

    def set_motion_depth(self, motion: np.ndarray | None):
        """SetMotionDepthBuffer without the GL frame buffer: float32 [h*w, 4] motion vectors in .xy, or None."""
        if motion is None:
            _check(self._lib, self._lib.atn_set_motion_depth(self._h, None, C.c_size_t(0)))
            return
        m = np.ascontiguousarray(motion, np.float32).reshape(-1, 4)
        _check(self._lib, self._lib.atn_set_motion_depth(self._h, C.c_void_p(m.ctypes.data), C.c_size_t(len(m))))

    def readback_svgf(self):
        """float32 [4, h*w, 4]: normal_depth, albedo_meshid, color_variance, moment_temporalweight of the last frame."""
        n = self.width * self.height
        out = np.empty((4, n, 4), np.float32)
        _check(self._lib, self._lib.atn_readback_svgf(self._h, *[C.c_void_p(out[k].ctypes.data) for k in range(4)], C.c_size_t(n)))
        return out

    def render_region(self, width, height, x0, y0, x1, y1, sample_begin, sample_end, maxBounce, rrBounce=3):
        _check(self._lib, self._lib.atn_render_region(
            self._h, C.c_int32(width), C.c_int32(height), C.c_int32(x0), C.c_int32(y0), C.c_int32(x1), C.c_int32(y1),
            C.c_int32(sample_begin), C.c_int32(sample_end), C.c_int32(maxBounce), C.c_int32(rrBounce)))

    def render_tiles(self, width, height, tile, world, rank, sample_begin, sample_end, maxBounce, rrBounce=3):
        """The interleaved tile set of `rank` (tiles g with g % world == rank, scanline order) as one wavefront."""
        _check(self._lib, self._lib.atn_render_tiles(
            self._h, C.c_int32(width), C.c_int32(height), C.c_int32(tile), C.c_int32(world), C.c_int32(rank),
            C.c_int32(sample_begin), C.c_int32(sample_end), C.c_int32(maxBounce), C.c_int32(rrBounce)))

    def synchronize(self):
        _check(self._lib, self._lib.atn_synchronize(self._h))

    def readback(self, out: np.ndarray | None = None) -> np.ndarray:
        n = self.width * self.height
        if out is None:
            out = np.empty((self.height, self.width, 4), np.float32)
        _check(self._lib, self._lib.atn_readback(self._h, C.c_void_p(out.ctypes.data), C.c_size_t(n)))
        return out

    def readback_aov(self):
        """(normal_depth, albedo_meshid), float32 [h, w, 4] each: the reference's two basic AOV buffers."""
        n = self.width * self.height
        nd = np.empty((self.height, self.width, 4), np.float32)
        am = np.empty((self.height, self.width, 4), np.float32)
        _check(self._lib, self._lib.atn_readback_aov(self._h, C.c_void_p(nd.ctypes.data), C.c_void_p(am.ctypes.data), C.c_size_t(n)))
        return nd, am

    def readback_into(self, host_ptr: int):
        """Read back into caller-owned (e.g. pinned) host memory given as an address."""
        _check(self._lib, self._lib.atn_readback(self._h, C.c_void_p(host_ptr), C.c_size_t(self.width * self.height)))

    def device_image(self):
        p = C.c_void_p()
        n = C.c_size_t()
        _check(self._lib, self._lib.atn_device_image(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def stream(self) -> int:
        p = C.c_void_p()
        _check(self._lib, self._lib.atn_get_stream(self._h, C.byref(p)))
        return p.value or 0

    def stats(self) -> dict:
        s = AtnStats()
        _check(self._lib, self._lib.atn_get_stats(self._h, C.byref(s)))
        return {k: getattr(s, k) for k, _ in AtnStats._fields_}

    def set_bounce_tap(self, max_bounce: int | None):
        """Install (or remove with None) the per-bounce parity tap; returns the record array."""
        if max_bounce is None:
            self._tap = None
            _check(self._lib, self._lib.atn_set_bounce_tap(self._h, None, C.c_int32(0)))
            return None
        self._tap = np.zeros((self.width * self.height, max_bounce), sio.BOUNCE_DTYPE)
        _check(self._lib, self._lib.atn_set_bounce_tap(self._h, C.c_void_p(self._tap.ctypes.data), C.c_int32(max_bounce)))
        return self._tap

    # -- ray queries -----------------------------------------------------------------------------
    def trace_closest(self, rays: np.ndarray, t_max: np.ndarray | None = None) -> np.ndarray:
        rays = np.ascontiguousarray(rays, sio.RAY_DTYPE)
        hits = np.zeros(len(rays), sio.HIT_DTYPE)
        tm = None if t_max is None else np.ascontiguousarray(t_max, np.float32)
        _check(self._lib, self._lib.atn_trace_closest(
            self._h, C.c_void_p(rays.ctypes.data), C.c_void_p(tm.ctypes.data) if tm is not None else None,
            C.c_size_t(len(rays)), C.c_void_p(hits.ctypes.data)))
        return hits

    STATIC_MAX_RAYS = 65536

    def trace_closest_static(self, rays: np.ndarray, t_max: np.ndarray | None = None) -> np.ndarray:
        """``trace_closest`` through ``atn_trace_closest_static``: for hosts that may not allocate device
        memory (fixed, already spent memory budget; frame loops that cannot stall on cudaMalloc).  Rays
        and hits are staged in a fixed array inside the library, so a call takes at most
        ``STATIC_MAX_RAYS`` rays, and calls on the same GPU must not overlap."""
        rays = np.ascontiguousarray(rays, sio.RAY_DTYPE)
        hits = np.zeros(len(rays), sio.HIT_DTYPE)
        tm = None if t_max is None else np.ascontiguousarray(t_max, np.float32)
        _check(self._lib, self._lib.atn_trace_closest_static(
            self._h, C.c_void_p(rays.ctypes.data), C.c_void_p(tm.ctypes.data) if tm is not None else None,
            C.c_size_t(len(rays)), C.c_void_p(hits.ctypes.data)))
        return hits

    def trace_any(self, rays: np.ndarray, dist: np.ndarray) -> np.ndarray:
        rays = np.ascontiguousarray(rays, sio.RAY_DTYPE)
        dist = np.ascontiguousarray(dist, np.float32)
        vis = np.zeros(len(rays), np.uint8)
        _check(self._lib, self._lib.atn_trace_any(self._h, C.c_void_p(rays.ctypes.data), C.c_void_p(dist.ctypes.data),
                                                  C.c_size_t(len(rays)), C.c_void_p(vis.ctypes.data)))
        return vis

    def trace_closest_device(self, rays_ptr: int, t_max_ptr: int | None, n: int, hits_ptr: int):
        _check(self._lib, self._lib.atn_trace_closest_device(self._h, C.c_void_p(rays_ptr),
                                                             C.c_void_p(t_max_ptr) if t_max_ptr else None,
                                                             C.c_size_t(n), C.c_void_p(hits_ptr)))

    def trace_any_device(self, rays_ptr: int, dist_ptr: int, n: int, vis_ptr: int):
        _check(self._lib, self._lib.atn_trace_any_device(self._h, C.c_void_p(rays_ptr), C.c_void_p(dist_ptr),
                                                         C.c_size_t(n), C.c_void_p(vis_ptr)))
