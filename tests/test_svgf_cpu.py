"""N1 (SVGF): the CPU restatement (oracle/pt_oracle.cpp + aten_b200/csrc/atn_svgf.cuh) against the stock
aten::SVGFRenderer -- golden film images of three frames with a constant motion vector and the AOV set
of the last frame (tests/golden/*.svgf.f32.gz, made by tests/golden/make_golden.py svgf), bit for bit."""
import numpy as np
import pytest

from tests import oracle_lib as ol

SCENES = ["cornell_mtrl", "instanced", "sponza_atrium"]


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("name", SCENES)
def test_svgf_restatement_matches_stock_renderer(golden, name):
    scene = golden.scene(name)
    imgs, aov, motion = golden.svgf(name)
    n = scene.width * scene.height
    md = np.zeros((n, 4), np.float32); md[:, 0] = motion[0]; md[:, 1] = motion[1]
    state = None
    for f in range(len(imgs)):
        img, state = ol.render_svgf(scene, frames=1, first_frame=f, state=state, motion=md, host_order=True)
        assert (bits(img) == bits(imgs[f])).all(), f"film of frame {f} differs"
    mine = ol.svgf_last_aov(state, len(imgs))
    for k, what in enumerate(("normal_depth", "albedo_meshid", "color_variance", "moment_temporalweight")):
        assert (bits(mine[k]) == bits(aov[k])).all(), what


def test_svgf_one_call_equals_frame_by_frame(golden):
    scene = golden.scene("instanced")
    a, sa = ol.render_svgf(scene, frames=3)
    st = None
    for f in range(3):
        b, st = ol.render_svgf(scene, frames=1, first_frame=f, state=st)
    assert (bits(a) == bits(b)).all() and (bits(sa) == bits(st)).all()


def test_svgf_order_independent_form_is_close_to_host_order(golden):
    """The GPU path estimates the variance from the unfiltered colours of all pixels (the reference filters in
    place, so its result depends on the pixel order: pixels processed later see already filtered neighbours).
    On a noisy 1-sample frame the two stay within a few per cent RMSE of each other."""
    scene = golden.scene("cornell_mtrl")
    a, _ = ol.render_svgf(scene, frames=2, host_order=True)
    b, _ = ol.render_svgf(scene, frames=2, host_order=False)
    rel = np.sqrt(((a - b)[..., :3] ** 2).mean()) / a[..., :3].mean()
    assert 0 < rel < 0.05


@pytest.mark.skipif(not ol.have_ref(), reason="oracle/_ref not built (no /root/reference on this machine)")
def test_svgf_live_reference_other_size_and_iterations():
    W, H = 80, 45
    scene = ol.ref_scene("cornell", W, H)
    for iters in (1, 3):
        imgs, aov = ol.ref_render_svgf("cornell", W, H, frames=2, iters=iters)
        img, st = ol.render_svgf(scene, frames=2, host_order=True, iters=iters)
        assert (bits(img) == bits(imgs[1])).all()
