"""TEST INFRASTRUCTURE: hand-made scenes in the reference's own array layouts (threaded node lists, triangle and object
parameters) for cases none of the reference's assets reaches: a maximally unbalanced bottom tree whose traversal stack
need exceeds the shared-memory stack plus the small local frame of the traversal kernel."""
from __future__ import annotations

import numpy as np

from aten_b200 import scene_io as sio


def _thread(tree, tri_lo, tri_hi):
    """Nested (left, right) tuples with triangle numbers at the leaves -> threaded node list in depth-first preorder
    (hit = first child, miss = next subtree; a bottom leaf carries isleaf = 1 and its triangle number, sbvh.h:19-66)."""
    nodes = []

    def build(t):
        me = len(nodes)
        n = np.zeros((), sio.NODE_DTYPE)
        n["object_id"] = -1.0; n["primid"] = -1.0; n["exid"] = -1.0; n["meshid"] = -1.0
        nodes.append(n)
        if isinstance(t, tuple):
            a, (alo, ahi) = build(t[0])
            b, (blo, bhi) = build(t[1])
            lo, hi = np.minimum(alo, blo), np.maximum(ahi, bhi)
        else:
            lo, hi = tri_lo[t], tri_hi[t]
            n["object_id"], n["primid"] = 1.0, float(t)
        n["boxmin"], n["boxmax"] = lo, hi
        return me, (lo, hi)

    import sys
    sys.setrecursionlimit(10000)
    build(tree)
    out = np.array(nodes, sio.NODE_DTYPE)
    cnt = len(out)
    size = np.ones(cnt, np.int64)
    leaf = out["primid"] >= 0
    for i in range(cnt - 1, -1, -1):
        if not leaf[i]:
            a = i + 1
            b = a + size[a]
            size[i] = 1 + size[a] + size[b]
    nxt = np.arange(cnt) + size
    nxt = np.where(nxt >= cnt, -1, nxt)
    out["miss"] = nxt.astype(np.float32)
    out["hit"] = np.where(leaf, nxt, np.arange(cnt) + 1).astype(np.float32)
    return out


def chain_scene(base: sio.Scene, depth: int, seed: int = 3) -> sio.Scene:
    """One mesh of 2 * (depth + 1) triangles in pairs stacked along +z under a maximally unbalanced tree: spine node i
    holds the inner node of pair i (the FAR pairs first) and spine node i + 1, so a ray travelling along +z follows the
    spine down to the nearest pair and leaves one subtree per level on its stack -- about `depth` entries, far beyond
    the 16 of the traversal kernel's shared-memory stack.  Threaded layout as the reference's builders emit it;
    materials, camera and config come from `base`."""
    rng = np.random.default_rng(seed)
    pairs = depth + 1
    n = 2 * pairs
    pos = np.zeros((3 * n, 4), np.float32)
    nml = np.zeros((3 * n, 4), np.float32)
    tri = np.zeros(n, sio.TRI_DTYPE)
    for i in range(n):
        z = np.float32(pairs - i // 2) + np.float32(0.25 * (i & 1))       # pair 0 is the farthest
        c = rng.uniform(-0.3, 0.3, 2).astype(np.float32)
        pos[3 * i + 0, :3] = (c[0] - 1.0, c[1] - 1.0, z)
        pos[3 * i + 1, :3] = (c[0] + 1.5, c[1] - 1.0, z)
        pos[3 * i + 2, :3] = (c[0] - 1.0, c[1] + 1.5, z)
        nml[3 * i:3 * i + 3, :3] = (0.0, 0.0, -1.0)
        tri[i]["idx"] = (3 * i, 3 * i + 1, 3 * i + 2)
        tri[i]["area"] = 0.5 * 2.5 * 2.5
        tri[i]["mtrlid"] = 0
        tri[i]["mesh_id"] = 0
    lo = np.array([pos[3 * i:3 * i + 3, :3].min(axis=0) for i in range(n)], np.float32)
    hi = np.array([pos[3 * i:3 * i + 3, :3].max(axis=0) for i in range(n)], np.float32)
    tree = (2 * depth, 2 * depth + 1)
    for i in range(depth - 1, -1, -1):
        tree = ((2 * i, 2 * i + 1), tree)
    nodes = _thread(tree, lo, hi)

    objects = np.zeros(2, sio.OBJ_DTYPE)
    objects[0]["type"] = 0; objects[0]["area"] = float(tri["area"].sum()); objects[0]["object_id"] = -1; objects[0]["mtx_id"] = -1
    objects[0]["triangle_id"] = 0; objects[0]["triangle_num"] = n; objects[0]["light_id"] = -1; objects[0]["sphere_mtrl_id"] = -1
    objects[1]["type"] = 1; objects[1]["object_id"] = 0; objects[1]["mtx_id"] = -1; objects[1]["triangle_id"] = -1
    objects[1]["light_id"] = -1; objects[1]["sphere_mtrl_id"] = -1
    top = np.zeros(1, sio.NODE_DTYPE)
    top[0]["boxmin"], top[0]["boxmax"] = nodes[0]["boxmin"], nodes[0]["boxmax"]
    top[0]["hit"] = top[0]["miss"] = -1.0
    top[0]["object_id"] = 1.0; top[0]["primid"] = -1.0; top[0]["meshid"] = -1.0
    top[0]["exid"] = np.array([1], "<u4").view("<f4")[0]               # mainExid = list 1 (threaded_bvh.h:29-37)
    cfg = base.config.copy()
    cfg["scene_bbox_min"][0] = nodes[0]["boxmin"]; cfg["scene_bbox_max"][0] = nodes[0]["boxmax"]
    lights = base.lights[:0].copy()
    return sio.Scene(objects, np.zeros((0, 16), np.float32), base.materials.copy(), lights, tri, pos, nml, [top, nodes], [],
                     cfg, base.camera.copy(), base.width, base.height, None)


def chain_rays(scene: sio.Scene, count: int, seed: int = 11) -> np.ndarray:
    """Rays that run up the stack of triangles (mostly along +z, some oblique, some from the far side)."""
    rng = np.random.default_rng(seed)
    rays = np.zeros(count, sio.RAY_DTYPE)
    o = np.zeros((count, 3)); d = np.zeros((count, 3))
    o[:, :2] = rng.uniform(-1.5, 1.5, (count, 2)); o[:, 2] = rng.uniform(-2.0, 0.5, count)
    d[:, :2] = rng.normal(scale=0.05, size=(count, 2)); d[:, 2] = 1.0
    k = count // 4
    o[:k, 2] = scene.config["scene_bbox_max"][0][2] + rng.uniform(0.5, 2.0, k); d[:k, 2] = -1.0
    d[k:2 * k, :2] = rng.normal(scale=0.5, size=(k, 2))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["org"] = o.astype(np.float32); rays["dir"] = d.astype(np.float32)
    return rays
