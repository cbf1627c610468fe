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
