import gzip
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # GPU tests are skipped (not failed) where there is no device, e.g. in the CPU-only container.
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def _gz(name):
    with gzip.open(os.path.join(GOLDEN, name), "rb") as fp:
        return fp.read()


class Golden:
    """Golden vectors produced by the reference itself (tests/golden/make_golden.py)."""

    def __init__(self):
        from aten_b200 import scene_io as sio
        self.sio = sio
        self.meta = json.load(open(os.path.join(GOLDEN, "meta.json")))

    @property
    def scenes(self):
        return sorted(self.meta)

    def scene(self, name):
        return self.sio.load_scene(os.path.join(GOLDEN, f"{name}.atnscene.gz"))

    def trace(self, name):
        m = self.meta[name]
        return np.frombuffer(_gz(f"{name}.trace.gz"), self.sio.BOUNCE_DTYPE).reshape(m["w"] * m["h"], m["depth"])

    def image(self, name):
        m = self.meta[name]
        return np.frombuffer(_gz(f"{name}.image.f32.gz"), np.float32).reshape(m["h"], m["w"], 4)

    def ao(self, name):
        m = self.meta[name]
        return np.frombuffer(_gz(f"{name}.ao.f32.gz"), np.float32).reshape(m["h"], m["w"], 4)

    def rays(self, name):
        buf = np.frombuffer(_gz(f"{name}.rays.gz"), np.float32).reshape(-1, 7)
        rays = np.zeros(len(buf), self.sio.RAY_DTYPE)
        rays["org"] = buf[:, :3]
        rays["dir"] = buf[:, 3:6]
        return rays, buf[:, 6].copy()

    def svgf(self, name, frames=3):
        """Stock aten::SVGFRenderer: ([frames, H, W, 4] film images, [4, H*W, 4] AOV set of the last frame, motion)."""
        m = self.meta[name]
        n = m["w"] * m["h"]
        buf = np.frombuffer(_gz(f"{name}.svgf.f32.gz"), np.float32)
        imgs = buf[:frames * n * 4].reshape(frames, m["h"], m["w"], 4)
        aov = buf[frames * n * 4:].reshape(4, n, 4)
        return imgs, aov, (0.02, -0.03)

    def restir(self, name, frames=4):
        """The reference's ReSTIR steps: ([frames, H, W, 4] contribution images, [H*W, 5] final reservoirs, motion)."""
        m = self.meta[name]
        n = m["w"] * m["h"]
        buf = np.frombuffer(_gz(f"{name}.restir.f32.gz"), np.float32)
        return buf[:frames * n * 4].reshape(frames, m["h"], m["w"], 4), buf[frames * n * 4:].reshape(n, 5), (0.02, -0.03)

    def hits(self, name):
        return np.frombuffer(_gz(f"{name}.hits.gz"), self.sio.HIT_DTYPE)

    def cmj(self):
        return np.frombuffer(_gz("cmj.f32.gz"), np.float32)


@pytest.fixture(scope="session")
def golden():
    return Golden()


def bits(a):
    """uint32 view of a float32 array for bit-exact comparison."""
    return np.ascontiguousarray(a, np.float32).view(np.uint32)
