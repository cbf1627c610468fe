"""CPU checks of the product's QuadNode encoder (aten_b200/csrc/atn_bvh.h, encode_wide): the traversal
kernel's per-ray logic, emulated sequentially over that encoding (tests/emu_traversal.cpp), must report
the reference's own hits.  The wide nodes are an exact collapse of the reference's binary trees -- no
approximate box test decides a visit inside a tree.  The front-to-back walk alone may differ from the reference on near
ties (assert_hits_equal(..., allow_ties=True)); with the kernel's near-tie rule -- tied rays are flagged and decided by the
reference-order walk -- every hit is the reference's, bit for bit."""
import numpy as np
import pytest

from aten_b200 import scene_io as sio
from tests import emu_lib
from tests import oracle_lib as ol
from tests.test_gpu_parity import assert_hits_equal

SCENES = ["cornell", "cornell_mtrl", "instanced", "sponza_atrium", "cornell_sphere", "cornell_alpha"]


@pytest.mark.parametrize("name", SCENES)
def test_wide_encoding_golden_rays(golden, name):
    """4096 incoherent rays (half with a finite t_max) against the hits the reference itself reported."""
    scene = golden.scene(name)
    rays, tmax = golden.rays(name)
    got, st = emu_lib.trace_closest(scene, rays, tmax)
    assert_hits_equal(got, golden.hits(name), allow_ties=True)
    assert st["max_sp"] <= st["stack_bound"]
    got, st = emu_lib.trace_closest_resolved(scene, rays, tmax)
    assert_hits_equal(got, golden.hits(name))
    assert st["flagged"] <= 0.02 * len(rays)


@pytest.mark.parametrize("name", SCENES)
def test_wide_encoding_reference_path_rays(golden, name):
    """The rays the reference traced at every bounce of its own paths."""
    ref = golden.trace(name)
    valid = (ref["flags"] & 1) != 0
    rays = np.ascontiguousarray(ref["ray"][valid])
    got, st = emu_lib.trace_closest_resolved(golden.scene(name), rays)
    want = np.ascontiguousarray(ref["hit"][valid]).copy()
    miss = want["objid"] < 0
    want["primid"][miss] = -1; want["mtrlid"][miss] = -1; want["meshid"][miss] = -1
    assert_hits_equal(got, want)
    assert st["max_sp"] <= st["stack_bound"] and st["flagged"] <= 0.02 * len(rays)


def test_wide_encoding_random_rays_vs_oracle_and_visit_order(golden):
    """100k random rays through Sponza: same hits as the oracle's reference-order walk; the visit order
    inside a node (entry distance vs leaves first) and the instance pre-filter change the work, never the result."""
    scene = golden.scene("sponza_atrium")
    cfg = scene.config[0]
    lo, hi = np.array(cfg["scene_bbox_min"], np.float64), np.array(cfg["scene_bbox_max"], np.float64)
    rng = np.random.default_rng(5)
    n = 100000
    rays = np.zeros(n, sio.RAY_DTYPE)
    rays["org"] = (lo + (hi - lo) * rng.random((n, 3))).astype(np.float32)
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    rays["dir"] = d.astype(np.float32)
    want = ol.trace_closest(scene, rays)
    miss = want["objid"] < 0
    want = want.copy(); want["primid"][miss] = -1; want["mtrlid"][miss] = -1; want["meshid"][miss] = -1
    a, sa = emu_lib.trace_closest(scene, rays)
    b, sb = emu_lib.trace_closest(scene, rays, leaves_first=True)
    c, sc = emu_lib.trace_closest(scene, rays, entry_filter=False)
    assert_hits_equal(a, want, allow_ties=True)
    for other in (b, c):
        assert (other["primid"] == a["primid"]).all() and (other["objid"] == a["objid"]).all()
        assert (other["t"].view(np.uint32) == a["t"].view(np.uint32)).all()
    # the near-tie rule: no ray that differs from the reference goes unflagged, and few are flagged
    d, flags, sd = emu_lib.trace_closest_flagged(scene, rays, emu_lib.KERNEL_BAND)
    differ = (d["objid"] != want["objid"]) | (d["primid"] != want["primid"]) | (d["t"].view(np.uint32) != want["t"].view(np.uint32))
    assert not (differ & (flags == 0)).any()
    assert flags.mean() < 0.01
    assert sc["entries"] >= sa["entries"] and sa["entries_filtered"] > 0
    # the collapse roughly halves the node steps of the binary tree and uses most slots
    assert sa["slots_used"] >= 2.5 * sa["node_count"]
    assert sa["not_nested"] <= 0.01 * sa["node_count"]


def test_deep_chain_tree_stack_bound(golden):
    """A maximally unbalanced tree (tests/synth_scenes.py): the traversal really holds ~depth entries on its stack, the
    encoder's worst-case bound covers it, and the hits are the reference-order walk's.  A list deeper than the encoder
    follows (96 levels) is refused, not truncated."""
    from tests import synth_scenes as ss
    base = golden.scene("cornell")
    scene = ss.chain_scene(base, 88)
    rays = ss.chain_rays(scene, 20000)
    want = ol.trace_closest(scene, rays)
    miss = want["objid"] < 0
    want = want.copy(); want["primid"][miss] = -1; want["mtrlid"][miss] = -1; want["meshid"][miss] = -1
    got, st = emu_lib.trace_closest(scene, rays)
    assert got.tobytes() == want.tobytes()
    assert 64 < st["max_sp"] <= st["stack_bound"] <= 16 + 240
    with pytest.raises(RuntimeError):
        emu_lib.trace_closest(ss.chain_scene(base, 120), rays[:8])


@pytest.mark.parametrize("name", ["sponza_atrium", "instanced"])
def test_encoded_tree_cache_round_trip(golden, tmp_path, name):
    """The encoded-tree file (aten_b200/csrc/atn_cache.h, behind atn_upload_scene_cached): what comes back from it, with the
    top layer encoded again, is byte-identical to a fresh encoding; a file of another scene or hash is refused."""
    rc = emu_lib.encoded_cache_round_trip(golden.scene(name), str(tmp_path / "t.atnenc"), golden.scene("cornell"))
    assert rc == 0, rc
