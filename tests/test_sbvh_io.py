"""`.sbvh` tree files (reference src/libaten/accelerator/sbvh.cpp:1220-1436): reader/writer round trip,
error paths, and -- where the reference's asset travelled with oracle/_ref -- the shipped
sponza_lod.sbvh against the node list the reference itself imported (golden scene blob)."""
import os

import numpy as np
import pytest

from aten_b200 import sbvh_io
from aten_b200.scene_io import NODE_DTYPE
from tests import oracle_lib as ol

ASSET = os.path.join(ol.REF_ASSET, "sponza", "sponza_lod.sbvh")


def _nodes(n, seed=1):
    rng = np.random.default_rng(seed)
    a = np.zeros(n, NODE_DTYPE)
    a["boxmin"] = rng.random((n, 3)); a["boxmax"] = a["boxmin"] + rng.random((n, 3))
    a["hit"] = np.arange(1, n + 1); a["hit"][-1] = -1
    a["miss"] = -1
    a["object_id"] = np.where(np.arange(n) % 2, 1, -1)
    a["primid"] = np.where(np.arange(n) % 2, np.arange(n) // 2, -1)
    a["exid"] = -1; a["meshid"] = -1
    return a


def test_round_trip_with_material_table(tmp_path):
    nodes = _nodes(33)
    nodes["meshid"][4] = 7                      # a voxel-LOD node carrying a material id
    p = str(tmp_path / "t.sbvh")
    sbvh_io.write_sbvh(p, nodes, 9, [0, 0, 0], [1, 2, 3], {7: "floor", 2: "wall_x"})
    f = sbvh_io.read_sbvh(p)
    assert f.max_depth == 9 and f.version == (0, 0, 0, 1) and f.voxel_materials == {7: "floor", 2: "wall_x"}
    assert f.nodes.tobytes() == nodes.tobytes() and f.boxmax.tolist() == [1, 2, 3]
    # importTree semantics: triangle offset on leaves only, material ids through the name table
    g = sbvh_io.read_sbvh(p, offset_tri_idx=100, material_index_by_name={"floor": 3, "wall_x": 5}.get)
    leaf = nodes["primid"] >= 0
    assert (g.nodes["primid"][leaf] == nodes["primid"][leaf] + 100).all() and (g.nodes["primid"][~leaf] == -1).all()
    assert g.nodes["meshid"][4] == 3


def test_rejects_bad_files(tmp_path):
    p = tmp_path / "bad.sbvh"
    p.write_bytes(b"NOPE" + b"\0" * 60)
    with pytest.raises(ValueError, match="magic"):
        sbvh_io.read_sbvh(str(p))
    nodes = _nodes(8)
    q = str(tmp_path / "t.sbvh")
    sbvh_io.write_sbvh(q, nodes, 3, [0, 0, 0], [1, 1, 1])
    raw = open(q, "rb").read()
    (tmp_path / "short.sbvh").write_bytes(raw[:-10])
    with pytest.raises(ValueError, match="truncated"):
        sbvh_io.read_sbvh(str(tmp_path / "short.sbvh"))
    (tmp_path / "empty.sbvh").write_bytes(b"")
    with pytest.raises(ValueError):
        sbvh_io.read_sbvh(str(tmp_path / "empty.sbvh"))


@pytest.mark.skipif(not os.path.exists(ASSET), reason="reference asset sponza_lod.sbvh not present")
def test_shipped_sponza_tree_equals_what_the_reference_imported(golden):
    f = sbvh_io.read_sbvh(ASSET)
    assert f.version == (0, 0, 0, 1) and f.max_depth == 23 and len(f.nodes) == 37999
    scene = golden.scene("sponza_atrium")
    lists = [l for l in scene.node_lists if len(l) == len(f.nodes)]
    assert len(lists) == 1
    ref = lists[0]
    for k in ("boxmin", "boxmax", "hit", "miss", "object_id", "primid", "exid"):
        assert (ref[k].view(np.uint32) == f.nodes[k].view(np.uint32)).all(), k


# ---- the same files through the C ABI (atn_sbvh_read / atn_sbvh_write: what a C++ host calls; no GPU involved) ----------
def test_c_abi_reader_and_writer_agree_with_the_python_ones(tmp_path):
    from aten_b200 import api
    nodes = _nodes(33)
    nodes["meshid"][4] = 7
    p = str(tmp_path / "c.sbvh"); q = str(tmp_path / "py.sbvh")
    api.sbvh_write(p, nodes, 9, [0, 0, 0], [1, 2, 3], {7: "floor", 2: "wall_x"})
    sbvh_io.write_sbvh(q, nodes, 9, [0, 0, 0], [1, 2, 3], {7: "floor", 2: "wall_x"})
    assert open(p, "rb").read() == open(q, "rb").read()
    got, info = api.sbvh_read(p)
    assert got.tobytes() == nodes.tobytes() and info["max_depth"] == 9 and info["voxel_materials"] == 2
    assert info["boxmax"].tolist() == [1, 2, 3] and info["version"] == (0, 0, 0, 1)
    # importTree semantics: triangle offset on leaves only, material ids through the name table -> index into the names
    got, _ = api.sbvh_read(p, offset_tri_idx=100, material_names=["a", "b", "c", "floor", "d", "wall_x"])
    want = sbvh_io.read_sbvh(q, offset_tri_idx=100, material_index_by_name={"floor": 3, "wall_x": 5}.get)
    assert got.tobytes() == want.nodes.tobytes()
    with pytest.raises(api.AtnError, match="not found in the scene"):
        api.sbvh_read(p, material_names=["only_this"])
    (tmp_path / "bad.sbvh").write_bytes(b"NOPE" + b"\0" * 60)
    with pytest.raises(api.AtnError, match="bad magic"):
        api.sbvh_read(str(tmp_path / "bad.sbvh"))
    (tmp_path / "short.sbvh").write_bytes(open(p, "rb").read()[:-10])
    with pytest.raises(api.AtnError, match="truncated"):
        api.sbvh_read(str(tmp_path / "short.sbvh"))
    with pytest.raises(api.AtnError, match="cannot open"):
        api.sbvh_read(str(tmp_path / "missing.sbvh"))


@pytest.mark.skipif(not os.path.exists(ASSET), reason="reference asset sponza_lod.sbvh not present")
def test_c_abi_reads_the_shipped_sponza_tree(golden):
    from aten_b200 import api
    got, info = api.sbvh_read(ASSET)
    want = sbvh_io.read_sbvh(ASSET)
    assert got.tobytes() == want.nodes.tobytes() and info["max_depth"] == 23 and len(got) == 37999


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(ASSET), reason="reference asset sponza_lod.sbvh not present")
def test_sbvh_file_to_upload_to_trace_and_encoded_cache(golden, tmp_path):
    """File -> reader -> upload -> trace: the shipped sponza_lod.sbvh, read through the C ABI and put in place of the
    scene's own Sponza list, reproduces the reference's golden hits bit for bit; and an upload from the encoded-tree
    cache gives byte-identical results to the fresh upload that wrote it (hits, and a rendered frame)."""
    from aten_b200 import PathTracing, api
    from tests.test_gpu_parity import assert_hits_equal
    from tests.conftest import bits
    name = "sponza_atrium"
    scene = golden.scene(name)
    nodes, info = api.sbvh_read(ASSET)
    k = [i for i, l in enumerate(scene.node_lists) if len(l) == len(nodes)][0]
    scene.node_lists[k] = nodes                       # the tree as read from the file
    rays, tmax = golden.rays(name)
    cache = str(tmp_path / "sponza.atnenc")
    W, H = scene.width, scene.height
    out = []
    for expect_hit in (False, True):
        with PathTracing(0) as r:
            r.UpdateSceneData(scene, cache_path=cache)
            assert r.cache_hit is expect_hit
            hits = r.trace_closest(rays, tmax)
            assert_hits_equal(hits, golden.hits(name))
            r.set_gather_mode(api.GATHER_SUM)
            r.render(W, H, 2, 5, 3)
            out.append((hits.tobytes(), r.readback().copy()))
    assert out[0][0] == out[1][0] and (bits(out[0][1]) == bits(out[1][1])).all()
    # a cache file of another scene is not used: it is replaced
    with PathTracing(0) as r:
        r.UpdateSceneData(golden.scene("cornell"), cache_path=cache)
        assert r.cache_hit is False
    with PathTracing(0) as r:
        r.UpdateSceneData(scene, cache_path=cache)
        assert r.cache_hit is False
        assert r.trace_closest(rays, tmax).tobytes() == out[0][0]
