"""Texture fetch modes (atn_set_texture_fetch).  The default -- the reference's host rule aten::texture::at --
is what every parity test runs.  The device rule (the GPU renderer's tex2DLod with each texture's filter and address mode)
has no runnable reference on this box (PARITY UNPINNED); it is pinned here to an independent numpy restatement of the
texture-fetch formulas of the CUDA programming guide, and the GPU path to the CPU restatement of the same text."""
import numpy as np
import pytest

from tests import oracle_lib as ol


def fetch_numpy(tex, filter_mode, address_mode, uv):
    """Independent restatement: normalised coordinates; point = T[floor(N u)]; linear = bilinear about floor(N u - 0.5)
    with 8-bit weights; wrap / clamp / mirror / border addressing."""
    h, w, _ = tex.shape
    uv = uv.astype(np.float32)

    def coord(c):
        if address_mode == 0:
            return c - np.floor(c)
        if address_mode == 2:
            f = np.floor(c); r = c - f
            return np.where(f.astype(np.int64) & 1, np.float32(1) - r, r)
        return c

    def texel(i, j):
        i = i.astype(np.int64); j = j.astype(np.int64)
        if address_mode == 0:
            return tex[j % h, i % w]
        if address_mode == 3:
            ok = (i >= 0) & (i < w) & (j >= 0) & (j < h)
            out = tex[np.clip(j, 0, h - 1), np.clip(i, 0, w - 1)].copy()
            out[~ok] = 0
            return out
        return tex[np.clip(j, 0, h - 1), np.clip(i, 0, w - 1)]

    x = (coord(uv[:, 0]) * np.float32(w)).astype(np.float32)
    y = (coord(uv[:, 1]) * np.float32(h)).astype(np.float32)
    if filter_mode != 1:
        return texel(np.floor(x), np.floor(y))
    xb = x - np.float32(0.5); yb = y - np.float32(0.5)
    i = np.floor(xb); j = np.floor(yb)
    a = (np.floor((xb - i) * np.float32(256) + np.float32(0.5)) / np.float32(256)).astype(np.float32)[:, None]
    b = (np.floor((yb - j) * np.float32(256) + np.float32(0.5)) / np.float32(256)).astype(np.float32)[:, None]
    one = np.float32(1)
    return ((one - a) * (one - b) * texel(i, j) + a * (one - b) * texel(i + 1, j)
            + (one - a) * b * texel(i, j + 1) + a * b * texel(i + 1, j + 1)).astype(np.float32)


@pytest.mark.parametrize("filter_mode", [0, 1])
@pytest.mark.parametrize("address_mode", [0, 1, 2, 3])
def test_device_fetch_formula(filter_mode, address_mode):
    rng = np.random.default_rng(7 + 10 * filter_mode + address_mode)
    tex = rng.random((13, 29, 4)).astype(np.float32)
    uv = rng.uniform(-2.2, 3.1, (20000, 2)).astype(np.float32)
    uv[:64] = np.array([[0, 0], [1, 1], [0.5, 0.5], [1, 0], [-1, 2], [0.999999, 0.0], [1e-7, 1 - 1e-7], [2.5, -0.5]] * 8, np.float32)
    got = ol.tex_fetch(tex, filter_mode, address_mode, uv)
    want = fetch_numpy(tex, filter_mode, address_mode, uv)
    assert np.allclose(got, want, rtol=0, atol=2e-7), np.abs(got - want).max()
    if filter_mode == 0:
        assert (got == want).all()
    # a texel centre returns the texel itself in both filters
    cu = (np.arange(29) + 0.5) / 29; cv = (np.arange(13) + 0.5) / 13
    centres = np.stack(np.meshgrid(cu, cv), axis=-1).reshape(-1, 2).astype(np.float32)
    c = ol.tex_fetch(tex, filter_mode, address_mode, centres).reshape(13, 29, 4)
    assert np.allclose(c, tex, atol=1e-6)


def test_restatement_switch_changes_a_textured_render(golden):
    """The switch reaches the render: with the device rule the environment map of sponza_atrium is read bilinearly / at
    other texels, the image changes slightly, and switching back restores the default result bit for bit."""
    scene = golden.scene("sponza_atrium")
    base, _, _ = ol.render(scene, frames=1)
    ol.set_texture_fetch(True)
    try:
        dev, _, _ = ol.render(scene, frames=1)
    finally:
        ol.set_texture_fetch(False)
    again, _, _ = ol.render(scene, frames=1)
    assert (base.view(np.uint32) == again.view(np.uint32)).all()
    assert not (base.view(np.uint32) == dev.view(np.uint32)).all()
    assert abs(dev[..., :3].mean() - base[..., :3].mean()) < 0.1 * base[..., :3].mean()


@pytest.mark.gpu
def test_gpu_device_fetch_vs_restatement(golden):
    """GPU with atn_set_texture_fetch(1) against the restatement with the same rule (same text, nvcc vs g++), on the scene
    with an environment map; the default stays the host rule."""
    from aten_b200 import PathTracing
    name = "sponza_atrium"; m = golden.meta[name]
    scene = golden.scene(name)
    for tex, flt, adr in scene.textures:
        assert tex.shape[0] > 1
    scene.textures = [(t, 1, 0) for (t, _, _) in scene.textures]          # Linear + Wrap, as the Sponza scenes set them (scenedefs.cpp:823-826)
    with PathTracing(0) as r:
        r.UpdateSceneData(scene)
        r.set_texture_fetch(True)
        r.render(m["w"], m["h"], 2, 5, 3)
        got = r.readback().copy()
        r.set_texture_fetch(False)
        r.reset(); r.set_frame(0)
        r.render(m["w"], m["h"], 2, 5, 3)
        host_rule = r.readback().copy()
    ol.set_texture_fetch(True)
    try:
        want, _, _ = ol.render(scene, frames=2)
    finally:
        ol.set_texture_fetch(False)
    want_host, _, _ = ol.render(scene, frames=2)
    def close(a, b):
        return (np.abs(a[..., :3] - b[..., :3]) <= 1e-4 * np.maximum(np.abs(b[..., :3]), 1e-3)).all(axis=2).mean()
    assert close(got, want) > 0.98, close(got, want)
    assert close(host_rule, want_host) > 0.98
    assert close(got, want_host) < 0.9            # the two rules really differ on this scene
