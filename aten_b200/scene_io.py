"""Flat scene blobs (`.atnscene`) and the ctypes view of ``AtnSceneBlob``.

The file format is what ``idaten_b200::FlatScene::Save`` (include/idaten_b200_shim.h) writes: the
arrays ``Renderer::UpdateSceneData`` reads from ``aten::context`` (reference
src/libidaten/kernel/renderer.cpp:12-131) in the reference's own struct layouts, little endian::

    char[8]  "ATNSCN01"
    u32[10]  objects, matrices, materials, lights, triangles, vertices, node_lists, textures, seeds, flags (bit 0: alpha blending)
    AtnSceneConfig (56 B)   AtnCameraParameter (124 B)   i32 width, height
    objects[64 B]  matrices[64 B]  materials[248 B]  lights[80 B]  triangles[32 B]
    positions[16 B]  normals[16 B]
    u32 node_list_sizes[node_lists]   nodes[48 B] (lists concatenated)
    per texture: i32 w, h, filter, address; float4 texels[w*h]
    u32 seeds[seeds]         (optional; otherwise std::mt19937(0), see mt19937_seeds)

``.atnscene.gz`` is the same stream gzip-compressed (tests/golden).
"""
from __future__ import annotations

import ctypes as C
import gzip
import os
from dataclasses import dataclass, field

import numpy as np

# ---------------------------------------------------------------------------------------------
# numpy dtypes of the POD structs in include/aten_b200.h
# ---------------------------------------------------------------------------------------------
RAY_DTYPE = np.dtype([("org", "<f4", 3), ("dir", "<f4", 3)])
HIT_DTYPE = np.dtype([("t", "<f4"), ("objid", "<i4"), ("mtrlid", "<i4"), ("meshid", "<i4"),
                      ("primid", "<i4"), ("a", "<f4"), ("b", "<f4"), ("isVoxel", "<i4")])
NODE_DTYPE = np.dtype([("boxmin", "<f4", 3), ("hit", "<f4"), ("boxmax", "<f4", 3), ("miss", "<f4"),
                       ("object_id", "<f4"), ("primid", "<f4"), ("exid", "<f4"), ("meshid", "<f4")])
TRI_DTYPE = np.dtype([("idx", "<i4", 3), ("padding", "<f4"), ("area", "<f4"), ("needNormal", "<i4"),
                      ("mtrlid", "<i4"), ("mesh_id", "<i4")])
OBJ_DTYPE = np.dtype([("type", "<i4"), ("area", "<f4"), ("object_id", "<i4"), ("mtx_id", "<i4"),
                      ("triangle_id", "<i4"), ("triangle_num", "<i4"), ("light_id", "<i4"), ("pad0", "<i4"),
                      ("sphere_center", "<f4", 3), ("sphere_radius", "<f4"), ("sphere_mtrl_id", "<i4"),
                      ("pad1", "<i4", 3)])
MTRL_DTYPE = np.dtype([("baseColor", "<f4", 4), ("type", "<i4"), ("attrib", "<u4"), ("id", "<u2"),
                       ("isIdealRefraction", "u1"), ("is_medium", "u1"), ("albedoMap", "<i4"),
                       ("normalMap", "<i4"), ("roughnessMap", "<i4"), ("stencil_type", "<i4"),
                       ("standard", "<f4", 12), ("carpaint_tail", "u1", 16), ("medium", "u1", 32),
                       ("toon", "u1", 100), ("feature_line", "u1", 8)])
LIGHT_DTYPE = np.dtype([("pos", "<f4", 4), ("dir", "<f4", 4), ("type", "<i4"), ("light_color", "<f4", 3),
                        ("innerAngle", "<f4"), ("outerAngle", "<f4"), ("attrib", "<u4"), ("scale", "<f4"),
                        ("intensity", "<f4"), ("arealight_objid", "<i4"), ("envmapidx", "<i4"), ("padding", "<i4")])
CAMERA_DTYPE = np.dtype([("origin", "<f4", 3), ("lookat", "<f4", 3), ("aspect", "<f4"), ("center", "<f4", 3),
                         ("u", "<f4", 3), ("v", "<f4", 3), ("dir", "<f4", 3), ("right", "<f4", 3), ("up", "<f4", 3),
                         ("dist", "<f4"), ("vfov", "<f4"), ("width", "<i4"), ("height", "<i4"),
                         ("znear", "<f4"), ("zfar", "<f4")])
CONFIG_DTYPE = np.dtype([("bvh_hit_min", "<f4"), ("bg_color", "<f4", 3), ("envmap_tex_idx", "<i4"),
                         ("avgIllum", "<f4"), ("multiplyer", "<f4"), ("enable_env_map", "<i4"),
                         ("scene_bbox_min", "<f4", 3), ("scene_bbox_max", "<f4", 3)])
BOUNCE_DTYPE = np.dtype([("ray", RAY_DTYPE), ("hit", HIT_DTYPE),
                         ("shadow_org", "<f4", 3), ("shadow_dist", "<f4"),
                         ("shadow_dir", "<f4", 3), ("shadow_active", "<i4"),
                         ("light_contrib", "<f4", 3), ("shadow_visible", "<i4"),
                         ("throughput", "<f4", 3), ("pdfb", "<f4"),
                         ("contrib", "<f4", 3), ("flags", "<i4")])
assert (RAY_DTYPE.itemsize, HIT_DTYPE.itemsize, NODE_DTYPE.itemsize, TRI_DTYPE.itemsize, OBJ_DTYPE.itemsize,
        MTRL_DTYPE.itemsize, LIGHT_DTYPE.itemsize, CAMERA_DTYPE.itemsize, CONFIG_DTYPE.itemsize,
        BOUNCE_DTYPE.itemsize) == (24, 32, 48, 32, 64, 248, 80, 124, 56, 136)


def mt19937_seeds(count: int, seed: int = 0) -> np.ndarray:
    """``aten::initSampler`` (reference src/libaten/sampler/sampler.cpp:8-18): ``count`` raw 32-bit
    outputs of ``std::mt19937(seed)``."""
    rs = np.random.RandomState(seed)          # init_genrand(seed), same as std::mt19937
    return np.frombuffer(rs.bytes(4 * count), dtype="<u4").copy()


# ---------------------------------------------------------------------------------------------
# ctypes mirror of AtnSceneBlob
# ---------------------------------------------------------------------------------------------
class AtnTextureDesc(C.Structure):
    _fields_ = [("texels", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32),
                ("filter_mode", C.c_int32), ("address_mode", C.c_int32)]


class AtnSceneConfig(C.Structure):
    _fields_ = [("bvh_hit_min", C.c_float), ("bg_color", C.c_float * 3), ("envmap_tex_idx", C.c_int32),
                ("avgIllum", C.c_float), ("multiplyer", C.c_float), ("enable_env_map", C.c_int32),
                ("scene_bbox_min", C.c_float * 3), ("scene_bbox_max", C.c_float * 3)]


class AtnSceneBlob(C.Structure):
    _fields_ = [("objects", C.c_void_p), ("object_count", C.c_uint32),
                ("matrices", C.c_void_p), ("matrix_count", C.c_uint32),
                ("materials", C.c_void_p), ("material_count", C.c_uint32),
                ("lights", C.c_void_p), ("light_count", C.c_uint32),
                ("triangles", C.c_void_p), ("triangle_count", C.c_uint32),
                ("positions", C.c_void_p), ("normals", C.c_void_p), ("vertex_count", C.c_uint32),
                ("node_lists", C.POINTER(C.c_void_p)), ("node_list_sizes", C.c_void_p),
                ("node_list_count", C.c_uint32),
                ("textures", C.POINTER(AtnTextureDesc)), ("texture_count", C.c_uint32),
                ("config", AtnSceneConfig), ("enable_alpha_blending", C.c_int32)]


assert C.sizeof(AtnSceneConfig) == 56


@dataclass
class Scene:
    """A scene as flat numpy arrays (see the module docstring)."""
    objects: np.ndarray
    matrices: np.ndarray
    materials: np.ndarray
    lights: np.ndarray
    triangles: np.ndarray
    positions: np.ndarray
    normals: np.ndarray
    node_lists: list
    textures: list          # list of (texels float32[h,w,4], filter, address)
    config: np.ndarray      # CONFIG_DTYPE scalar array (shape (1,))
    camera: np.ndarray      # CAMERA_DTYPE (shape (1,))
    width: int
    height: int
    seeds: np.ndarray
    alpha_blending: bool = False      # SceneRenderingConfig::enable_alpha_blending
    _keep: list = field(default_factory=list, repr=False)

    def seeds_for(self, count: int) -> np.ndarray:
        if self.seeds is not None and len(self.seeds) >= count:
            return np.ascontiguousarray(self.seeds[:count])
        return mt19937_seeds(count)

    def with_size(self, width: int, height: int) -> "Scene":
        """Same scene and view, other resolution: PinholeCamera::CreateCameraParam
        (reference src/libaten/camera/pinhole.cpp:33-76) re-derives aspect, u and dist."""
        import copy
        s = copy.copy(self)
        cam = self.camera.copy()
        c = cam[0]
        theta = np.float32(np.float32(np.pi) * c["vfov"] / np.float32(180.0))
        aspect = np.float32(width) / np.float32(height)
        half_h = np.float32(np.tan(np.float32(theta / np.float32(2))))
        half_w = np.float32(aspect * half_h)
        c["aspect"] = aspect
        c["u"] = (half_w * c["right"]).astype(np.float32)
        c["v"] = (half_h * c["up"]).astype(np.float32)
        c["dist"] = np.float32(height) / (np.float32(2.0) * half_h)
        c["width"] = width
        c["height"] = height
        s.camera = cam
        s.width, s.height = width, height
        s.seeds = None
        s._keep = []
        return s

    def blob(self) -> AtnSceneBlob:
        """ctypes view; arrays are kept alive by ``self``."""
        b = AtnSceneBlob()
        keep = []

        def ptr(a):
            a = np.ascontiguousarray(a)
            keep.append(a)
            return a.ctypes.data if a.size else None

        b.objects, b.object_count = ptr(self.objects), len(self.objects)
        b.matrices, b.matrix_count = ptr(self.matrices), len(self.matrices)
        b.materials, b.material_count = ptr(self.materials), len(self.materials)
        b.lights, b.light_count = ptr(self.lights), len(self.lights)
        b.triangles, b.triangle_count = ptr(self.triangles), len(self.triangles)
        b.positions, b.normals, b.vertex_count = ptr(self.positions), ptr(self.normals), len(self.positions)
        n = len(self.node_lists)
        arr = (C.c_void_p * max(n, 1))()
        sizes = np.zeros(max(n, 1), dtype=np.uint32)
        for i, l in enumerate(self.node_lists):
            arr[i] = ptr(l)
            sizes[i] = len(l)
        keep.append(arr)
        b.node_lists = C.cast(arr, C.POINTER(C.c_void_p))
        b.node_list_sizes = ptr(sizes)
        b.node_list_count = n
        nt = len(self.textures)
        tarr = (AtnTextureDesc * max(nt, 1))()
        for i, (tex, flt, adr) in enumerate(self.textures):
            tarr[i].texels = ptr(tex)
            tarr[i].height, tarr[i].width = tex.shape[0], tex.shape[1]
            tarr[i].filter_mode, tarr[i].address_mode = flt, adr
        keep.append(tarr)
        b.textures = C.cast(tarr, C.POINTER(AtnTextureDesc))
        b.texture_count = nt
        C.memmove(C.byref(b.config), ptr(self.config), 56)
        b.enable_alpha_blending = 1 if self.alpha_blending else 0
        self._keep = keep
        return b

    @property
    def node_count(self) -> int:
        return int(sum(len(l) for l in self.node_lists))


def load_scene(path: str) -> Scene:
    opener = gzip.open if str(path).endswith(".gz") else open
    with opener(path, "rb") as fp:
        buf = fp.read()
    if buf[:8] != b"ATNSCN01":
        raise ValueError(f"{path}: not an .atnscene file")
    off = 8
    counts = np.frombuffer(buf, "<u4", 10, off); off += 40
    n_obj, n_mtx, n_mtrl, n_light, n_tri, n_vtx, n_lists, n_tex, n_seed, flags = (int(c) for c in counts)

    def take(dtype, n):
        nonlocal off
        a = np.frombuffer(buf, dtype, n, off).copy()
        off += a.nbytes
        return a

    config = take(CONFIG_DTYPE, 1)
    camera = take(CAMERA_DTYPE, 1)
    wh = take("<i4", 2)
    objects = take(OBJ_DTYPE, n_obj)
    matrices = take("<f4", n_mtx * 16).reshape(n_mtx, 16)
    materials = take(MTRL_DTYPE, n_mtrl)
    lights = take(LIGHT_DTYPE, n_light)
    triangles = take(TRI_DTYPE, n_tri)
    positions = take("<f4", n_vtx * 4).reshape(n_vtx, 4)
    normals = take("<f4", n_vtx * 4).reshape(n_vtx, 4)
    sizes = take("<u4", n_lists)
    node_lists = [take(NODE_DTYPE, int(s)) for s in sizes]
    textures = []
    for _ in range(n_tex):
        hdr = take("<i4", 4)
        w, h = int(hdr[0]), int(hdr[1])
        textures.append((take("<f4", w * h * 4).reshape(h, w, 4), int(hdr[2]), int(hdr[3])))
    seeds = take("<u4", n_seed) if n_seed else None
    return Scene(objects, matrices, materials, lights, triangles, positions, normals, node_lists, textures,
                 config, camera, int(wh[0]), int(wh[1]), seeds, bool(flags & 1))


def save_scene(scene: Scene, path: str, with_seeds: bool = False) -> None:
    opener = gzip.open if str(path).endswith(".gz") else open
    seeds = scene.seeds if (with_seeds and scene.seeds is not None) else np.zeros(0, np.uint32)
    with opener(path, "wb") as fp:
        fp.write(b"ATNSCN01")
        fp.write(np.array([len(scene.objects), len(scene.matrices), len(scene.materials), len(scene.lights),
                           len(scene.triangles), len(scene.positions), len(scene.node_lists),
                           len(scene.textures), len(seeds), 1 if scene.alpha_blending else 0], "<u4").tobytes())
        fp.write(scene.config.tobytes()); fp.write(scene.camera.tobytes())
        fp.write(np.array([scene.width, scene.height], "<i4").tobytes())
        for a in (scene.objects, scene.matrices, scene.materials, scene.lights, scene.triangles,
                  scene.positions, scene.normals):
            fp.write(np.ascontiguousarray(a).tobytes())
        fp.write(np.array([len(l) for l in scene.node_lists], "<u4").tobytes())
        for l in scene.node_lists:
            fp.write(np.ascontiguousarray(l).tobytes())
        for tex, flt, adr in scene.textures:
            fp.write(np.array([tex.shape[1], tex.shape[0], flt, adr], "<i4").tobytes())
            fp.write(np.ascontiguousarray(tex, dtype="<f4").tobytes())
        fp.write(np.ascontiguousarray(seeds, dtype="<u4").tobytes())
