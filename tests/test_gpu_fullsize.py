"""Image parity at BASELINE.json's full frame size (-m gpu): the CUDA path against the CPU restatement (itself bit-exact
against the reference, tests/test_oracle_cpu.py) with the same scene, camera, seeds, samples and bounces.

The metric, declared once (DESIGN.md "Exactness"), and why it is not a plain RMSE: the two sides run the same arithmetic
except for libm vs CUDA libdevice in the shading functions (last-ulp differences of sinf / cosf / atan2f / acosf / expf /
logf).  On almost every pixel that changes the radiance in the 7th digit.  On a few pixels per 10^4 it flips a DISCRETE
decision -- Russian roulette, the light pick, a texel of a point-sampled texture, a primitive at a secondary hit -- and
from that bounce on the two sides follow different random paths: both are valid samples of the same integrand, but the
pixel values differ by a sample's variance, and with an environment light a handful of them are fireflies that dominate
any un-normalised norm of the whole frame.  So per config:
  equal      share of pixels whose three channels agree within 1e-4 relative (floor 1e-3)       >= the config's bound
  rmse_equal relative RMSE over those pixels                                                    <= 1e-5
  relmse     mean of (a - b)^2 / (b^2 + 1e-2) over the whole frame (fireflies are bounded by b) <= the config's bound
  mean       whole-frame mean radiance                                                          within 2e-3 relative
  rays       closest + shadow rays traced                                                       within 1e-5 relative
and non-finite pixels (the reference's Disney material produces a few) are the same pixels on both sides."""
import os

import numpy as np
import pytest

from aten_b200 import PathTracing
from tests import oracle_lib as ol

pytestmark = pytest.mark.gpu

W, H = 1920, 1080
CASES = [
    # name,               spp, equal >=, relmse <=   (bounds ~3x the measured values, profiles/r02_full_size_image_parity.json)
    ("sponza_atrium",      16, 0.999,    1e-5),      # configs[1] stand-in at the bench size
    ("sponza_tex",          4, 0.97,     6e-2),      # configs[2] stand-in: albedo + normal maps (point-sampled texels amplify ulps)
    ("instanced_dragon",    4, 0.999,    3e-4),      # configs[3] stand-in: transformed instances, Disney, refraction
]


def compare(got, want):
    fin_g = np.isfinite(got[..., :3]).all(axis=2); fin_w = np.isfinite(want[..., :3]).all(axis=2)
    finite = fin_g & fin_w
    a = np.where(finite[..., None], got[..., :3], 0).astype(np.float64)
    b = np.where(finite[..., None], want[..., :3], 0).astype(np.float64)
    same = (np.abs(a - b) <= 1e-4 * np.maximum(np.abs(b), 1e-3)).all(axis=2)
    return {
        "same_non_finite": bool((fin_g == fin_w).all()), "non_finite": int((~finite).sum()),
        "equal": float(same.mean()),
        "rmse_equal": float(np.sqrt(np.mean((a[same] - b[same]) ** 2)) / np.sqrt(np.mean(b[same] ** 2))),
        "relmse": float(np.mean((a - b) ** 2 / (b ** 2 + 1e-2))),
        "mean_rel": float(abs(a.mean() - b.mean()) / b.mean()),
    }


@pytest.mark.parametrize("name,spp,equal_min,relmse_max", CASES)
def test_full_size_image_vs_restatement(golden, name, spp, equal_min, relmse_max):
    if name == "sponza_atrium":
        scene = golden.scene(name).with_size(W, H)
    else:
        if not ol.have_ref():
            pytest.skip("scene is built by the reference's loaders (oracle/_ref did not travel)")
        scene = ol.ref_scene(name, W, H)
    with PathTracing(0) as r:
        r.UpdateSceneData(scene)
        r.SetEnableProgressive(True)
        r.render(W, H, spp, 5, 3)
        got = r.readback()
        st = r.stats()
    want, _, stats = ol.render(scene, frames=spp, max_depth=5, rr_depth=3, threads=os.cpu_count() or 1)
    m = compare(got, want)
    rays_g = st["closest_rays"] + st["shadow_rays"]; rays_w = int(stats[0] + stats[1])
    print({"scene": name, "spp": spp, **m, "gpu_rays": rays_g, "restatement_rays": rays_w, "exact_rays": st["exact_rays"]})
    assert m["same_non_finite"]
    assert m["equal"] >= equal_min, m
    assert m["rmse_equal"] <= 1e-5, m
    assert m["relmse"] <= relmse_max, m
    assert m["mean_rel"] <= 2e-3, m
    assert abs(rays_g - rays_w) <= 1e-5 * rays_w, (rays_g, rays_w)
    assert (got[..., 3] == spp).all()
