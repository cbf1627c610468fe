"""The C++ host side on the GPU: `oracle/_ref/device_host device ...` is the reference's device_renderer frame loop
(src/device_renderer/main.cpp:133-148, 185-202) with `idaten_b200::PathTracing` (include/idaten_b200_shim.h) in the
place of `idaten::PathTracing` -- scene built by the reference's own loaders and SBVH builder, handed to the C ABI
through the shim class, rendered on the GPU, image read back.  Checked against the ctypes host path on the same
scene (bit-identical: both end in the same library calls) and against the CPU restatement (relative RMSE <= 2e-3,
the image tolerance of tests/test_gpu_parity.py)."""
import json
import os
import subprocess
import tempfile

import numpy as np
import pytest

from aten_b200 import PathTracing
from tests import oracle_lib as ol
from tests.conftest import bits

pytestmark = pytest.mark.gpu

DEVICE_HOST = os.path.join(ol.ORACLE_DIR, "_ref", "device_host")


def run_device_host(name, W, H, frames, spp, depth):
    with tempfile.TemporaryDirectory() as d:
        out = os.path.join(d, "img.f32")
        r = subprocess.run([DEVICE_HOST, ol.REF_ASSET, "device", name, str(W), str(H), str(frames), str(spp), str(depth), out],
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        info = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
        return np.fromfile(out, np.float32).reshape(H, W, 4), info


@pytest.mark.skipif(not os.path.exists(DEVICE_HOST), reason="oracle/_ref/device_host not built (needs /root/reference at build time)")
@pytest.mark.parametrize("name,W,H,frames,spp", [("cornell", 256, 256, 3, 1), ("instanced", 320, 180, 2, 2)])
def test_cpp_shim_renders_like_the_ctypes_host_and_the_restatement(name, W, H, frames, spp):
    depth = 5
    img, info = run_device_host(name, W, H, frames, spp, depth)
    assert info["frame_count"] == frames * spp
    assert np.isfinite(img).all() and (img[..., 3] == frames * spp).all()

    scene = ol.ref_scene(name, W, H)
    scene.camera["znear"] = 0.1; scene.camera["zfar"] = 10000.0        # device_renderer sets these before the upload
    with PathTracing(0) as r:
        r.UpdateSceneData(scene)
        r.SetEnableProgressive(True)
        for _ in range(frames):
            r.render(W, H, spp, depth, 3)
        py = r.readback()
    assert (bits(img) == bits(py)).all()

    want, _, _ = ol.render(scene, frames=frames * spp, max_depth=depth, rr_depth=3)
    rmse = np.sqrt(np.mean((img[..., :3] - want[..., :3]) ** 2)); scale = np.sqrt(np.mean(want[..., :3] ** 2))
    assert rmse <= 2e-3 * scale, (rmse, scale)


@pytest.mark.skipif(not os.path.exists(DEVICE_HOST), reason="oracle/_ref/device_host not built (needs /root/reference at build time)")
def test_cpp_shim_update_bvh_equals_fresh_upload():
    """`idaten_b200::PathTracing::updateBVH` (the app's per-frame call for moving instances, src/device_renderer/main.cpp:
    440-470): after handing over the top layer, objects and matrices of the scene with one instance displaced, the
    renderer draws what a renderer freshly given that scene draws, bit for bit -- and not what it drew before."""
    W = H = 128
    with tempfile.TemporaryDirectory() as d:
        pa, pb = os.path.join(d, "a.f32"), os.path.join(d, "b.f32")
        r = subprocess.run([DEVICE_HOST, ol.REF_ASSET, "device_update", "cornell", str(W), str(H), "2", "5", pa, pb],
                           stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        a = np.fromfile(pa, np.float32).reshape(H, W, 4); b = np.fromfile(pb, np.float32).reshape(H, W, 4)
    assert (bits(a) == bits(b)).all() and np.isfinite(a).all() and a[..., :3].max() > 0
    base, _ = run_device_host("cornell", W, H, 1, 2, 5)
    assert (bits(base) != bits(a)).any()
