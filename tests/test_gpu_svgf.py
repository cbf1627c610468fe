"""N1 (SVGF) on the GPU: atn_render_svgf against the CPU restatement (oracle_render_svgf, itself bit-exact
against the stock aten::SVGFRenderer -- tests/test_svgf_cpu.py) in the order-independent variance form, frame by
frame with a motion vector; against the stock renderer's golden AOVs where the two forms coincide (the path-traced
AOVs); and the properties that hold at full size."""
import numpy as np
import pytest

from aten_b200.api import PathTracing, AtnError
from tests import oracle_lib as ol

pytestmark = pytest.mark.gpu
SCENES = ["cornell_mtrl", "instanced", "sponza_atrium"]

# expf / powf / tanf differ between libm and CUDA by an ulp or two, and one frame feeds the next through the
# colour history: relative tolerance on the film and on the float AOV planes
TOL = 1e-3


def rel_err(a, b, floor=1e-3):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)


def plane_err(a, b):
    """Error of a float4 plane relative to max(|value|, 1 % of the plane's mean level per channel): a path whose
    radiance is a thousandth of its neighbours' may differ in the 4th digit (sin / cos / exp of libm vs CUDA)
    without that meaning anything for the filter."""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    floor = np.maximum(0.01 * np.abs(b).mean(axis=0, keepdims=True), 1e-6)
    return np.abs(a - b) / np.maximum(np.maximum(np.abs(a), np.abs(b)), floor)


def near_tie_pixels(nd_gpu, nd_ref):
    """Pixels whose first hit differs between the GPU and the oracle (documented near ties)."""
    return np.abs(nd_gpu[:, 3] - nd_ref[:, 3]) > 1e-4 * np.maximum(np.abs(nd_ref[:, 3]), 1.0)


@pytest.mark.parametrize("name", SCENES)
def test_svgf_frames_match_restatement(golden, name):
    scene = golden.scene(name)
    W, H = scene.width, scene.height
    n = W * H
    md = np.zeros((n, 4), np.float32); md[:, 0] = 0.02; md[:, 1] = -0.03
    r = PathTracing(0)
    try:
        r.UpdateSceneData(scene)
        r.set_motion_depth(md)
        state = None
        for f in range(4):
            want, state = ol.render_svgf(scene, frames=1, first_frame=f, state=state, motion=md, host_order=False)
            r.render_svgf(W, H, 5, 3)
            got = r.readback()
            aov = r.readback_svgf()
            ref_aov = ol.svgf_last_aov(state, f + 1)
            # path-traced AOVs: normal, depth, albedo, material id -- identical except on near-tie pixels
            tie = near_tie_pixels(aov[0], ref_aov[0])
            assert tie.mean() <= 2e-3
            ok = ~tie
            assert (aov[1][ok, 3] == ref_aov[1][ok, 3]).all(), "material ids"
            assert rel_err(aov[0][ok], ref_aov[0][ok]).max() <= 1e-5
            assert rel_err(aov[1][ok], ref_aov[1][ok]).max() <= 1e-5
            # film and filter state: compare away from pixels whose 5-iteration footprint (radius 2 * 31) saw a tie
            if tie.any():
                continue
            assert plane_err(got.reshape(n, 4)[:, :3], want.reshape(n, 4)[:, :3]).max() <= TOL, f"film, frame {f}"
            assert plane_err(aov[2][:, :3], ref_aov[2][:, :3]).max() <= TOL, f"colour history, frame {f}"
            # the moments hold the RAW luminance of single paths: one in a thousand may differ visibly when its
            # radiance is tiny (see plane_err); all others agree
            me = plane_err(aov[3], ref_aov[3]).max(axis=1)
            assert (me > TOL).mean() <= 1e-3 and np.median(me) <= 1e-5, f"moments, frame {f}"
            # variance = E[l^2] - E[l]^2 cancels: its error scales with E[l^2], not with the variance itself
            second_moment = ref_aov[3][:, 0] / ref_aov[3][:, 2]
            dv = np.abs(aov[2][:, 3].astype(np.float64) - ref_aov[2][:, 3])
            assert (dv <= TOL * np.maximum(ref_aov[2][:, 3], second_moment) + 1e-7).all(), f"variance, frame {f}"
        assert r.GetFrameCount() == 4
    finally:
        r.close()


def test_svgf_path_traced_aovs_match_stock_renderer(golden):
    """The AOV planes written by the path-tracing half do not depend on the variance form: compare with the stock
    renderer's own dump (golden, frame 2)."""
    name = "cornell_mtrl"
    scene = golden.scene(name)
    imgs, aov, motion = golden.svgf(name)
    n = scene.width * scene.height
    md = np.zeros((n, 4), np.float32); md[:, 0] = motion[0]; md[:, 1] = motion[1]
    r = PathTracing(0)
    try:
        r.UpdateSceneData(scene)
        r.set_motion_depth(md)
        for f in range(3):
            r.render_svgf(scene.width, scene.height, 5, 3)
        got = r.readback_svgf()
    finally:
        r.close()
    tie = near_tie_pixels(got[0], aov[0])
    assert tie.mean() <= 2e-3
    ok = ~tie
    assert (got[1][ok, 3] == aov[1][ok, 3]).all()
    assert rel_err(got[0][ok], aov[0][ok]).max() <= 1e-5 and rel_err(got[1][ok], aov[1][ok]).max() <= 1e-5


def test_svgf_static_sequence_properties_full_size(golden):
    """1920x1080, 6 frames, static camera: deterministic; background pixels pass through; the frame counter of the
    moments reaches the number of frames where the history is accepted; variance shrinks as history accumulates."""
    scene = golden.scene("sponza_atrium").with_size(1920, 1080)
    W, H = 1920, 1080
    r = PathTracing(0)
    try:
        r.UpdateSceneData(scene)
        films, var = [], []
        for f in range(6):
            r.render_svgf(W, H, 5, 3)
            films.append(r.readback().copy())
            aov = r.readback_svgf()
            var.append(float(aov[2][:, 3].mean()))
        frames_acc = aov[3][:, 2]
        assert np.isfinite(films[-1]).all()
        assert (frames_acc >= 1).all() and frames_acc.max() == 6
        assert (frames_acc == 6).mean() > 0.9            # static camera: almost every pixel keeps its history
        assert var[-1] < 0.5 * var[0]
        # same sequence again: bit-identical
        r.set_frame(0)
        for f in range(6):
            r.render_svgf(W, H, 5, 3)
        again = r.readback()
        assert (again.view(np.uint32) == films[-1].view(np.uint32)).all()
        st = r.stats()
        assert st["closest_rays"] > W * H and st["kernel_launches"] >= 1 + 5 * 5 + 8
    finally:
        r.close()


def test_svgf_api_errors(golden):
    scene = golden.scene("instanced")
    r = PathTracing(0)
    try:
        r.UpdateSceneData(scene)
        with pytest.raises(AtnError):
            r.readback_svgf()                      # nothing rendered yet
        r.set_frame(3)
        with pytest.raises(AtnError):
            r.render_svgf(scene.width, scene.height, 5, 3)   # no history for frame 3
        r.set_frame(0)
        with pytest.raises(AtnError):
            r.render_svgf(scene.width + 1, scene.height, 5, 3)
        with pytest.raises(AtnError):
            r.set_motion_depth(np.zeros((5, 4), np.float32))
        r.render_svgf(scene.width, scene.height, 5, 3, atrous_iters=0)     # 0 -> the reference's 5
        assert r.GetFrameCount() == 1
    finally:
        r.close()
