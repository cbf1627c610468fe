"""GPU checks (-m gpu) of the boundary's bookkeeping rather than of the arithmetic: the three gather modes, the
asynchronous render call and its counters, and the large-stack instantiation of the traversal kernel."""
import numpy as np
import pytest

from aten_b200 import AtnError, PathTracing
from aten_b200.api import GATHER_OVERWRITE, GATHER_PROGRESSIVE, GATHER_SUM
from tests import oracle_lib as ol
from tests import synth_scenes as ss
from tests.conftest import bits

pytestmark = pytest.mark.gpu


def test_gather_modes(golden):
    """OVERWRITE = the reference GPU renderer with progressive off (pathtracing_impl.cu:313-316): every call leaves the
    mean of ITS samples with w = 1, whatever the image held; PROGRESSIVE = FilmProgressive::put per sample;
    SUM = sum + count.  All three hold the same samples, so they agree up to float rounding."""
    name = "cornell_mtrl"; m = golden.meta[name]
    W, H = m["w"], m["h"]
    with PathTracing(0) as r:
        r.UpdateSceneData(golden.scene(name))
        r.set_gather_mode(GATHER_SUM)
        r.render(W, H, 3, 5, 3)
        total = r.readback().copy()
        assert (total[..., 3] == 3).all()
        r.reset(); r.set_gather_mode(GATHER_PROGRESSIVE)
        r.render(W, H, 3, 5, 3)
        prog = r.readback().copy()
        assert (prog[..., 3] == 3).all()
        assert np.allclose(prog[..., :3], total[..., :3] / 3, rtol=1e-5, atol=1e-7)
        # overwrite: two calls in a row, the second one's image does not depend on the first
        r.reset(); r.SetEnableProgressive(False)
        r.render(W, H, 3, 5, 3)
        first = r.readback().copy()
        assert (first[..., 3] == 1).all()
        assert (bits(first[..., :3]) == bits(total[..., :3] / np.float32(3))).all()
        r.render(W, H, 2, 5, 3)                                    # frames 3, 4
        second = r.readback().copy()
        assert (second[..., 3] == 1).all()
        r.reset(); r.set_frame(3); r.set_gather_mode(GATHER_SUM)
        r.render(W, H, 2, 5, 3)
        alone = r.readback()
        assert (bits(second[..., :3]) == bits(alone[..., :3] / np.float32(2))).all()
        with pytest.raises(AtnError, match="atn_set_gather_mode"):
            r.set_gather_mode(7)


def test_render_is_asynchronous_and_stats_follow(golden):
    """atn_render returns without waiting; atn_get_stats waits for the call it reports on.  Two renders queued back to
    back give the counters of the second and the image of both."""
    name = "sponza_atrium"; m = golden.meta[name]
    W, H = m["w"], m["h"]
    with PathTracing(0) as r:
        r.UpdateSceneData(golden.scene(name))
        r.set_gather_mode(GATHER_SUM)
        r.render(W, H, 1, 5, 3)
        one = r.stats()
        r.reset(); r.set_frame(0)
        r.render(W, H, 1, 5, 3)
        r.render(W, H, 2, 5, 3)
        st = r.stats()
        assert st["camera_samples"] == 2 * W * H and one["camera_samples"] == W * H
        assert st["closest_rays"] > one["closest_rays"] and st["device_ms"] > 0
        img = r.readback()
        assert (img[..., 3] == 3).all()


def test_large_stack_instantiation_on_a_deep_chain(golden):
    """Stack need 93 > 16 shared + 48 local entries: the upload picks the 240-entry instantiation; hits equal the
    reference-order walk's bit for bit.  A node list deeper than the encoder follows is refused with a message."""
    base = golden.scene("cornell")
    scene = ss.chain_scene(base, 88)
    rays = ss.chain_rays(scene, 100000)
    want = ol.trace_closest(scene, rays).copy()
    miss = want["objid"] < 0
    want["primid"][miss] = -1; want["mtrlid"][miss] = -1; want["meshid"][miss] = -1
    with PathTracing(0) as r:
        r.UpdateSceneData(scene)
        got = r.trace_closest(rays)
        assert got.tobytes() == want.tobytes()
        assert (got["objid"] >= 0).mean() > 0.4
    with PathTracing(0) as r:
        with pytest.raises(AtnError, match="deeper than 96 levels"):
            r.UpdateSceneData(ss.chain_scene(base, 120))


def test_static_query_reports_its_own_counters(golden):
    scene = golden.scene("cornell")
    rays, _ = golden.rays("cornell")
    with PathTracing(0) as r:
        r.UpdateSceneData(scene)
        r.render(scene.width, scene.height, 1, 5, 3)
        r.trace_closest_static(rays[:777])
        st = r.stats()
        assert st["closest_rays"] == 777 and st["shadow_rays"] == 0 and st["camera_samples"] == 0 and st["kernel_launches"] == 1


def test_tile_set_render_equals_full_frame(golden):
    """atn_render_tiles: the interleaved tile sets of 3 ranks (the last column and row of tiles are partly outside the
    frame), each rendered as one wavefront into the same image, give the full-frame render bit for bit -- the property the
    multi-GPU tile partition rests on -- and a rank's counters cover its own pixels only."""
    name = "sponza_atrium"; m = golden.meta[name]
    W, H = 200, 120                                   # not a multiple of the 64-pixel tile
    scene = golden.scene(name).with_size(W, H)
    with PathTracing(0) as r:
        r.UpdateSceneData(scene)
        r.set_gather_mode(GATHER_SUM)
        r.render(W, H, 3, 5, 3)
        full = r.readback().copy()
        whole = r.stats()
        r.reset(); r.set_frame(0)
        samples = rays = 0
        for rank in range(3):
            r.render_tiles(W, H, 64, 3, rank, 0, 3, 5, 3)
            st = r.stats()
            samples += st["camera_samples"]; rays += st["closest_rays"] + st["shadow_rays"]
        tiles = r.readback()
        assert (bits(tiles) == bits(full)).all()
        assert samples == whole["camera_samples"] == W * H * 3
        assert rays == whole["closest_rays"] + whole["shadow_rays"]
        with pytest.raises(AtnError, match="atn_render_tiles"):
            r.render_tiles(W, H, 64, 2, 2, 0, 1, 5, 3)


def test_wavefront_size_does_not_change_the_image(golden):
    """atn_set_wavefront_paths bounds the path-state memory (samples per wavefront); the image and the ray counts are those
    of the default wavefront, bit for bit, whether 1, 2 or all 5 samples of the frame share a wavefront -- and with the
    scanline slot order of a frame whose size is not a multiple of the 8 x 4 pixel blocks."""
    name = "sponza_atrium"
    for W, H in ((96, 56), (90, 50)):
        scene = golden.scene(name).with_size(W, H)
        images, rays = [], []
        for cap in (None, W * H, 2 * W * H):
            with PathTracing(0) as r:
                if cap is not None:
                    r.set_wavefront_paths(cap)
                r.UpdateSceneData(scene)
                r.set_gather_mode(GATHER_SUM)
                r.render(W, H, 5, 5, 3)
                images.append(r.readback().copy())
                st = r.stats()
                rays.append((st["closest_rays"], st["shadow_rays"], st["camera_samples"]))
        assert (bits(images[1]) == bits(images[0])).all() and (bits(images[2]) == bits(images[0])).all()
        assert rays[0] == rays[1] == rays[2] and rays[0][2] == W * H * 5
    with PathTracing(0) as r:
        with pytest.raises(AtnError, match="atn_set_wavefront_paths"):
            r.set_wavefront_paths(0)


def test_area_light_early_exit_and_side_stream_do_not_change_the_image(golden, monkeypatch):
    """Two things that are allowed to change the schedule only: shadow rays towards an area light end at the first hit
    clearly in front of the light object's bounds (instead of running the full closest hit), and the exact walk of
    near-tie rays runs on a side stream next to the shade kernels.  With either switched off (ATEN_B200_EXP, read when
    the handle is created) the image is the same bit for bit -- scenes with quad, sphere and instanced area lights."""
    for name in ("cornell", "cornell_mtrl", "cornell_sphere", "sponza_atrium", "instanced"):
        m = golden.meta[name]
        images = []
        for exp in ("0", "8", "1", "9"):
            monkeypatch.setenv("ATEN_B200_EXP", exp)
            with PathTracing(0) as r:
                r.UpdateSceneData(golden.scene(name))
                r.render(m["w"], m["h"], 3, m["depth"], m["rr"])
                images.append(r.readback().copy())
        for img in images[1:]:
            assert (bits(img) == bits(images[0])).all(), name
