#!/usr/bin/env python3
"""BENCH / TEST INFRASTRUCTURE: turn a scene of INSTANCES into a scene of DISTINCT meshes (config C5).

BASELINE.json's configs[4] asks for rays into a 10 M-triangle SBVH.  Two facts shape what can be built:
  * the reference's node format stores every link and primitive number as a float
    (src/libaten/accelerator/threaded_bvh.h:13-56), so one node list cannot address more than 2^24 nodes --
    a single 10 M-triangle mesh (one triangle per leaf, ~2.2 nodes per triangle) does not fit in it;
  * its SBVH builder needs hours for a mesh of that size.
So the 10 M triangles are 100 DISTINCT dragon meshes of 100 k triangles each: every copy has its own vertices,
triangles and its own bottom-layer node list in device memory (about 1.4 GB of QuadNodes + TriLeafs, far beyond
the 126 MB L2), unlike the instanced stand-in whose 100 instances share one 14 MB tree.

A copy is the source mesh moved by a translation (and optionally a positive uniform scale).  The source's SBVH
stays valid for the copy: p -> p * s + T in float32 is monotone per coordinate, so every "box contains point"
relation of the source tree holds for the moved boxes and points as well.  Rotations of the source instances are
dropped (a rotated box is not the box of the rotated content).  The top layer over the copies is a plain median
split written in the reference's threaded layout (hit = first child, miss = next subtree, leaves carry the
object index and the list number in the exid bit field, threaded_bvh.h:29-37).
"""
from __future__ import annotations

import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
from aten_b200 import scene_io as sio  # noqa: E402


def _top_layer(boxes_min: np.ndarray, boxes_max: np.ndarray, object_ids, list_ids) -> np.ndarray:
    """Threaded top-layer list over leaf boxes (median split on the widest axis of the centroids)."""
    nodes = []

    def build(idx):
        """appends the subtree over leaves `idx` in depth-first preorder; returns its node index"""
        me = len(nodes)
        n = np.zeros((), sio.NODE_DTYPE)
        n["boxmin"] = boxes_min[idx].min(axis=0)
        n["boxmax"] = boxes_max[idx].max(axis=0)
        n["object_id"] = -1.0; n["primid"] = -1.0; n["exid"] = -1.0; n["meshid"] = -1.0
        nodes.append(n)
        if len(idx) == 1:
            k = int(idx[0])
            n["object_id"] = np.float32(object_ids[k])
            n["exid"] = np.array([list_ids[k] & 0x7fff], "<u4").view("<f4")[0]      # mainExid:15, no LOD, has external tree
            return me
        c = 0.5 * (boxes_min[idx] + boxes_max[idx])
        axis = int(np.argmax(c.max(axis=0) - c.min(axis=0)))
        order = idx[np.argsort(c[:, axis], kind="stable")]
        half = len(order) // 2
        build(order[:half])
        build(order[half:])
        return me

    build(np.arange(len(boxes_min)))
    out = np.array(nodes, sio.NODE_DTYPE)
    # links: hit = first child for inner nodes, next subtree for leaves; miss = next subtree
    n = len(out)
    size = np.ones(n, np.int64)
    leaf = out["object_id"] >= 0
    # subtree sizes by a reverse sweep: in preorder a node's children are me+1 and me+1+size(me+1)
    for i in range(n - 1, -1, -1):
        if not leaf[i]:
            a = i + 1
            b = a + size[a]
            size[i] = 1 + size[a] + size[b]
    nxt = np.arange(n) + size
    # the "next subtree" of the last nodes is the end of the walk (-1)
    nxt = np.where(nxt >= n, -1, nxt)
    out["miss"] = nxt.astype(np.float32)
    out["hit"] = np.where(leaf, nxt, np.arange(n) + 1).astype(np.float32)
    return out


def flatten_instances(src: sio.Scene, scale: float = 1.0, max_copies: int | None = None, min_triangles: int = 1000) -> sio.Scene:
    """Every instance of `src` whose mesh has at least `min_triangles` triangles becomes a mesh of its own at the
    instance's translation; smaller ones (floor planes) are dropped.  Materials, lights, textures, camera and
    rendering config are kept."""
    objs = src.objects
    top = src.node_lists[0]
    leaves = top[top["object_id"] >= 0]
    s = np.float32(scale)
    new_pos, new_nml, new_tri, new_lists, poly, inst = [], [], [], [], [], []
    lo, hi = [], []
    v_off = t_off = 0
    for lf in leaves:
        io = int(lf["object_id"])
        o = objs[io]
        if o["type"] != 1 or not (lf["exid"] >= 0):
            continue
        real = objs[int(o["object_id"])]
        if int(real["triangle_num"]) < min_triangles:
            continue
        if max_copies is not None and len(poly) >= max_copies:
            break
        k = int(np.array([lf["exid"]], "<f4").view("<u4")[0] & 0x7fff)
        T = (src.matrices[int(o["mtx_id"])].reshape(4, 4)[:3, 3] if o["mtx_id"] >= 0 else np.zeros(3)).astype(np.float32)
        t0, tn = int(real["triangle_id"]), int(real["triangle_num"])
        tri = src.triangles[t0:t0 + tn].copy()
        v0, v1 = int(tri["idx"].min()), int(tri["idx"].max()) + 1
        pos = src.positions[v0:v1].copy()
        pos[:, :3] = pos[:, :3] * s + T                      # float32: one multiply, one add per coordinate
        tri["idx"] += v_off - v0
        tri["area"] *= s * s
        lst = src.node_lists[k].copy()
        lst["boxmin"] = lst["boxmin"] * s + T
        lst["boxmax"] = lst["boxmax"] * s + T
        is_leaf = lst["primid"] >= 0
        lst["primid"] = np.where(is_leaf, lst["primid"] + np.float32(t_off - t0), lst["primid"]).astype(np.float32)
        assert t_off + tn < (1 << 24), "primitive numbers are floats in the reference's node format"
        p = np.zeros((), sio.OBJ_DTYPE)
        p[()] = real
        p["triangle_id"] = t_off
        p["area"] = real["area"] * s * s
        poly.append(p)
        q = np.zeros((), sio.OBJ_DTYPE)
        q[()] = o
        inst.append(q)
        new_pos.append(pos); new_nml.append(src.normals[v0:v1]); new_tri.append(tri); new_lists.append(lst)
        lo.append(lst["boxmin"][0]); hi.append(lst["boxmax"][0])       # node 0 is the root of a list
        v_off += v1 - v0
        t_off += tn
    n = len(poly)
    if n == 0:
        raise ValueError("no instance to flatten")
    objects = np.zeros(2 * n, sio.OBJ_DTYPE)
    for j in range(n):
        objects[j] = poly[j]
        objects[n + j] = inst[j]
        objects[n + j]["object_id"] = j
        objects[n + j]["mtx_id"] = 2 * j                    # identity L2W / W2L pair, as the host emits them
    matrices = np.tile(np.eye(4, dtype=np.float32).reshape(1, 16), (2 * n, 1))
    lo, hi = np.array(lo, np.float32), np.array(hi, np.float32)
    top_new = _top_layer(lo, hi, object_ids=[n + j for j in range(n)], list_ids=[1 + j for j in range(n)])
    cfg = src.config.copy()
    cfg["scene_bbox_min"][0] = lo.min(axis=0)
    cfg["scene_bbox_max"][0] = hi.max(axis=0)
    return sio.Scene(objects, matrices, src.materials.copy(), src.lights.copy(), np.concatenate(new_tri),
                     np.concatenate(new_pos), np.concatenate(new_nml), [top_new] + new_lists, list(src.textures),
                     cfg, src.camera.copy(), src.width, src.height, None)


def surface_rays(scene: sio.Scene, n_rays: int, seed: int = 1, device="cuda"):
    """Incoherent diffuse secondary rays as configs[4] words them: origins uniformly distributed over the surface
    area (triangle picked in proportion to its area, uniform barycentrics), pushed 1e-3 of the triangle's size
    along the face normal; directions cosine-distributed about that normal.  Works on a flattened scene, whose
    positions are in world space.  Returns torch tensors on `device`: rays [n, 6] float32 and the hit-distance
    limits for the any-hit pass, U(0, scene diagonal)."""
    import torch
    g = torch.Generator(device=device); g.manual_seed(seed)
    idx = torch.from_numpy(scene.triangles["idx"].astype(np.int64)).to(device)
    pos = torch.from_numpy(np.ascontiguousarray(scene.positions[:, :3])).to(device)
    a, b, c = pos[idx[:, 0]], pos[idx[:, 1]], pos[idx[:, 2]]
    nrm = torch.linalg.cross(b - a, c - a)
    area2 = nrm.norm(dim=1)
    cdf = torch.cumsum(area2.double(), 0)
    rays = torch.empty((n_rays, 6), dtype=torch.float32, device=device)
    chunk = 8_000_000
    for s0 in range(0, n_rays, chunk):
        m = min(chunk, n_rays - s0)
        u = torch.rand((m, 5), generator=g, device=device, dtype=torch.float64)
        t = torch.searchsorted(cdf, u[:, 0] * cdf[-1]).clamp_(max=len(cdf) - 1)
        su = torch.sqrt(u[:, 1]).float()
        w0, w1 = 1.0 - su, su * (1.0 - u[:, 2].float())
        w2 = 1.0 - w0 - w1
        p = a[t] * w0[:, None] + b[t] * w1[:, None] + c[t] * w2[:, None]
        nn = nrm[t] / area2[t].clamp_min(1e-30)[:, None]
        # orthonormal frame about nn, cosine-weighted direction
        h = torch.where(nn[:, 0:1].abs() < 0.9, torch.tensor([1.0, 0.0, 0.0], device=device), torch.tensor([0.0, 1.0, 0.0], device=device))
        tx = torch.nn.functional.normalize(torch.linalg.cross(h.expand_as(nn), nn), dim=1)
        ty = torch.linalg.cross(nn, tx)
        r = torch.sqrt(u[:, 3]).float(); phi = (2.0 * np.pi * u[:, 4]).float()
        d = tx * (r * torch.cos(phi))[:, None] + ty * (r * torch.sin(phi))[:, None] + nn * torch.sqrt((1.0 - u[:, 3]).clamp_min(0)).float()[:, None]
        size = torch.sqrt(area2[t])
        rays[s0:s0 + m, 0:3] = p + nn * (1e-3 * size)[:, None]
        rays[s0:s0 + m, 3:6] = torch.nn.functional.normalize(d, dim=1)
    lo = torch.tensor(scene.config["scene_bbox_min"][0]); hi = torch.tensor(scene.config["scene_bbox_max"][0])
    diag = float((hi - lo).norm())
    dist = torch.rand(n_rays, generator=g, device=device, dtype=torch.float32) * diag
    return rays, dist
