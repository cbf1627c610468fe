"""N1 (ReSTIR): the CPU restatement (oracle/pt_oracle.cpp + aten_b200/csrc/atn_restir.cuh) against the reference's own
per-pixel steps -- ReSTIRRenderer::Shade / EvaluateVisibility / ApplyTemporalReuse / ApplySpatialReuse /
ComputePixelColor, run by oracle/ref_harness.cpp in the order of idaten::ReSTIRPathTracing::OnRender because the stock
host driver is in a debugging state (DESIGN.md "ReSTIR").  Golden: contribution images of four frames with a constant
motion vector (frames 2 and 3 use temporal reuse) and the final reservoirs, bit for bit."""
import numpy as np
import pytest

from tests import oracle_lib as ol

SCENES = ["cornell_mtrl", "instanced", "sponza_atrium"]


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("name", SCENES)
def test_restir_restatement_matches_reference_steps(golden, name):
    scene = golden.scene(name)
    imgs, res, motion = golden.restir(name)
    n = scene.width * scene.height
    md = np.zeros((n, 4), np.float32); md[:, 0] = motion[0]; md[:, 1] = motion[1]
    for f in range(len(imgs)):
        img, r = ol.render_restir(scene, frames=f + 1, motion=md)
        assert (bits(img) == bits(imgs[f])).all(), f"frame {f}"
    assert (bits(r) == bits(res)).all(), "reservoirs"
    valid = r[:, 2] >= 0
    assert valid.any() and (r[valid, 3] > 0).all()          # a kept sample has a positive weight W


def test_restir_reuse_changes_the_estimate_not_its_mean(golden):
    """Spatial reuse: M grows to the neighbourhood's sample count; the frame mean stays near the path tracer's."""
    scene = golden.scene("cornell_mtrl")
    img, r = ol.render_restir(scene, frames=1)
    assert r[:, 1].max() <= 9 and r[:, 1].max() >= 4
    pt, _, _ = ol.render(scene, frames=1)
    a, b = img[..., :3].mean(), pt[..., :3].mean()
    assert abs(a - b) < 0.35 * b


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (no /root/reference on this machine)")
def test_restir_live_reference_other_size():
    W, H = 72, 40
    scene = ol.ref_scene("cornell", W, H)
    imgs, res = ol.ref_render_restir("cornell", W, H, frames=3)
    img, r = ol.render_restir(scene, frames=3)
    assert (bits(img) == bits(imgs[2])).all() and (bits(r) == bits(res)).all()
