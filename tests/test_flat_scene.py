"""Config C5's scene shape -- many DISTINCT meshes, each with its own bottom-layer node list (scripts/flat_scene.py) --
at a size the checkers finish in seconds: the instanced golden scene with every instance turned into a mesh of
its own.  The flattened scene is a new input in the reference's layouts, so it goes through the same three
checks as the others: CPU restatement vs brute force, the encoder emulation vs the restatement, and (gpu) the
CUDA traversal vs the restatement, bit-exact."""
import os
import sys

import numpy as np
import pytest

from aten_b200 import scene_io as sio
from tests import emu_lib
from tests import oracle_lib as ol
from tests.test_gpu_parity import assert_hits_equal

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
import flat_scene as fs  # noqa: E402


@pytest.fixture(scope="module")
def flat(golden):
    return fs.flatten_instances(golden.scene("instanced"), min_triangles=100)


def random_rays(scene, n, seed):
    rng = np.random.default_rng(seed)
    lo = scene.config["scene_bbox_min"][0].astype(np.float64); hi = scene.config["scene_bbox_max"][0].astype(np.float64)
    rays = np.zeros(n, sio.RAY_DTYPE)
    rays["org"] = (lo - 0.2 * (hi - lo) + 1.4 * (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["dir"] = d.astype(np.float32)
    return rays


def test_flat_scene_layout(flat, golden):
    src = golden.scene("instanced")
    n = (len(flat.objects)) // 2
    assert n == 5 and len(flat.node_lists) == 1 + n
    assert len(flat.triangles) == 3 * 6400 + 2 * 15488           # three teapots, two suzannes, the floor is dropped
    assert all(flat.triangles["idx"].max() < len(flat.positions) for _ in [0])
    top = flat.node_lists[0]
    leaves = top[top["object_id"] >= 0]
    assert sorted(leaves["object_id"].astype(int)) == list(range(n, 2 * n))
    assert sorted(int(b) & 0x7fff for b in leaves["exid"].view("<u4")) == list(range(1, n + 1))
    # every triangle of a copy lies in the root box of the copy's own list
    for j in range(n):
        o = flat.objects[j]
        tri = flat.triangles[o["triangle_id"]:o["triangle_id"] + o["triangle_num"]]
        p = flat.positions[np.unique(tri["idx"]), :3]
        root = flat.node_lists[1 + j][0]
        assert (p >= root["boxmin"]).all() and (p <= root["boxmax"]).all()
        prim = flat.node_lists[1 + j]["primid"]
        prim = prim[prim >= 0].astype(int)
        assert prim.min() >= o["triangle_id"] and prim.max() < o["triangle_id"] + o["triangle_num"]
    assert len(src.node_lists[1]) == len(flat.node_lists[1])


def test_flat_scene_oracle_vs_brute_force(flat):
    rays = random_rays(flat, 200, 3)
    h = ol.trace_closest(flat, rays)
    P = flat.positions[:, :3].astype(np.float64); I = flat.triangles["idx"]
    a, b, c = P[I[:, 0]], P[I[:, 1]], P[I[:, 2]]
    e1, e2 = b - a, c - a
    for k in range(len(rays)):
        og = rays["org"][k].astype(np.float64); dr = rays["dir"][k].astype(np.float64)
        u = np.cross(dr, e2); det = (u * e1).sum(1)
        with np.errstate(all="ignore"):
            inv = 1.0 / det; r = og - a
            be = (u * r).sum(1) * inv; v = np.cross(r, e1); ga = (v * dr).sum(1) * inv; t = (v * e2).sum(1) * inv
        ok = (be >= 0) & (ga >= 0) & (be + ga <= 1) & (t > 1e-9)
        tb = t[ok].min() if ok.any() else np.inf
        th = float(h["t"][k]) if h["objid"][k] >= 0 else np.inf
        assert (np.isinf(tb) and np.isinf(th)) or abs(tb - th) <= 1e-4 * tb, (k, tb, th)


def test_flat_scene_encoder_emulation(flat):
    rays = random_rays(flat, 50_000, 4)
    want = ol.trace_closest(flat, rays).copy()
    miss = want["objid"] < 0
    want["primid"][miss] = -1; want["mtrlid"][miss] = -1; want["meshid"][miss] = -1
    got, st = emu_lib.trace_closest_resolved(flat, rays)
    assert (want["objid"] >= 0).mean() > 0.1
    assert_hits_equal(got, want)
    assert st["flagged"] <= 0.01 * len(rays)
    assert st["max_sp"] <= st["stack_bound"]


@pytest.mark.gpu
@pytest.mark.parametrize("form", ["0", "1"])
def test_flat_scene_gpu_vs_oracle(flat, monkeypatch, form):
    """Both forms of the traversal kernel (the library picks one at upload from the size of the encoded set; the
    override is read there): "0" the default form, "1" the load-streaming form that scenes far beyond the L2 get."""
    from aten_b200 import PathTracing
    monkeypatch.setenv("ATEN_B200_STREAM", form)
    rays = random_rays(flat, 200_000, 5)
    rng = np.random.default_rng(6)
    diag = float(np.linalg.norm(flat.config["scene_bbox_max"][0] - flat.config["scene_bbox_min"][0]))
    dist = (rng.random(len(rays)) * diag).astype(np.float32)
    with PathTracing(0) as r:
        r.UpdateSceneData(flat)
        got = r.trace_closest(rays)
        vis = r.trace_any(rays, dist)
    want = ol.trace_closest(flat, rays)
    assert_hits_equal(got, want)
    assert (vis == ol.trace_any(flat, rays, dist)).all()


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["sponza_atrium", "instanced"])
def test_streaming_form_renders_the_same_image(golden, monkeypatch, name):
    """The path tracer through the load-streaming form of the traversal kernel (closest-hit and shadow policies):
    both forms are exact traversals with the same tie rule, so the images are bit-identical."""
    from aten_b200 import PathTracing
    from tests.conftest import bits
    m = golden.meta[name]
    imgs = []
    for form in ("0", "1"):
        monkeypatch.setenv("ATEN_B200_STREAM", form)
        with PathTracing(0) as r:
            r.UpdateSceneData(golden.scene(name))
            r.set_gather_mode(2)
            r.render(m["w"], m["h"], 2, 5, 3)
            imgs.append(r.readback().copy())
    assert (bits(imgs[0]) == bits(imgs[1])).all()
    assert np.isfinite(imgs[0]).all() and imgs[0][..., :3].max() > 0


def many_instances(flat, copies):
    """`copies` translated instances per mesh of the flattened scene: new instance objects, L2W / W2L matrix pairs and a new
    top layer over the same bottom lists (what the host hands to Renderer::updateBVH after adding instances)."""
    import copy
    n = len(flat.objects) // 2
    s = copy.copy(flat)
    objs = [flat.objects[j] for j in range(n)]
    mats, lo, hi, oid, lid = [], [], [], [], []
    ext = (flat.config["scene_bbox_max"][0] - flat.config["scene_bbox_min"][0]).astype(np.float32)
    for c in range(copies):
        for j in range(n):
            t = np.array([(c % 4) * 1.1 * ext[0], 0.0, (c // 4) * 1.1 * ext[2]], np.float32)
            o = flat.objects[n + j].copy()
            o["object_id"] = j; o["mtx_id"] = len(mats)
            L = np.eye(4, dtype=np.float32); L[:3, 3] = t
            W = np.eye(4, dtype=np.float32); W[:3, 3] = -t
            mats += [L.reshape(16), W.reshape(16)]
            root = flat.node_lists[1 + j][0]
            lo.append(root["boxmin"] + t); hi.append(root["boxmax"] + t)
            oid.append(n + len(oid)); lid.append(1 + j)
            objs.append(o)
    s.objects = np.array(objs, sio.OBJ_DTYPE)
    s.matrices = np.array(mats, np.float32)
    lo, hi = np.array(lo, np.float32), np.array(hi, np.float32)
    s.node_lists = [fs._top_layer(lo, hi, oid, lid)] + list(flat.node_lists[1:])
    cfg = flat.config.copy()
    cfg["scene_bbox_min"][0] = lo.min(axis=0); cfg["scene_bbox_max"][0] = hi.max(axis=0)
    s.config = cfg
    s._keep = []
    return s


@pytest.mark.gpu
def test_top_layer_update_with_more_instances_than_uploaded(flat):
    """Renderer::updateBVH with a top layer eight times the uploaded one (the device node array has to grow): only
    list 0, objects and matrices are handed over; the result equals a fresh upload of that scene and the restatement."""
    from aten_b200 import PathTracing
    big = many_instances(flat, 8)
    assert len(big.node_lists[0]) > 6 * len(flat.node_lists[0])
    rays = random_rays(big, 100_000, 11)
    want = ol.trace_closest(big, rays)
    assert (want["objid"] >= 0).mean() > 0.1 and len(np.unique(want["objid"][want["objid"] >= 0])) > 20
    with PathTracing(0) as r, PathTracing(0) as fresh:
        r.UpdateSceneData(flat)
        r.updateBVH(big.node_lists[0], big.objects, big.matrices)
        got = r.trace_closest(rays)
        assert_hits_equal(got, want)
        fresh.UpdateSceneData(big)
        assert fresh.trace_closest(rays).tobytes() == got.tobytes()
        # and back to the small top layer
        r.updateBVH(flat.node_lists[0], flat.objects, flat.matrices)
        small_rays = random_rays(flat, 50_000, 12)
        assert_hits_equal(r.trace_closest(small_rays), ol.trace_closest(flat, small_rays))


def test_top_layer_update_encoding_equals_full_encoding(flat):
    """CPU: the top layer re-encoded alone against the kept bottom trees (atn_update_top_layer's path) is byte-identical
    to what a full encoding of the changed scene holds -- QuadNodes, instance leaves with their world boxes, root, stack
    bound -- for a top layer eight times the original and for the unchanged one."""
    assert emu_lib.top_update_equals_fresh(flat, many_instances(flat, 8)) == 0
    assert emu_lib.top_update_equals_fresh(flat, flat) == 0
