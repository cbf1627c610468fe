"""What a smaller wavefront (atn_set_wavefront_paths: less path-state memory) costs on the bench frame."""
import os, sys, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
from aten_b200 import PathTracing, scene_io as sio
W, H, SPP = 1920, 1080, 16
scene = sio.load_scene(os.path.join(ROOT, "tests/golden/sponza_atrium.atnscene.gz")).with_size(W, H)
for cap in (32 << 20, 16 << 20, 8 << 20, 4 << 20, 2 << 20):
    with PathTracing(0) as r:
        r.set_wavefront_paths(cap)
        r.UpdateSceneData(scene)
        best = 1e9
        for it in range(4):
            r.reset(); r.synchronize(); t0 = time.perf_counter()
            r.render(W, H, SPP, 5, 3); r.synchronize()
            if it: best = min(best, time.perf_counter() - t0)
        st = r.stats()
        print(json.dumps({"wavefront_paths": cap, "samples_per_wavefront": max(1, min(16, round(cap / (W * H)))), "path_state_gb": round(265 * W * H * max(1, min(16, round(cap / (W * H)))) / 1e9, 2),
                          "mrays_per_s": round((st["closest_rays"] + st["shadow_rays"]) / best / 1e6, 1), "ms": round(best * 1e3, 2)}), flush=True)
