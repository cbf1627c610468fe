"""What the encoded-tree cache (atn_upload_scene_cached) saves on the 10 M-triangle scene (config C5: 100 distinct dragon meshes):
upload without a cache, upload that writes the cache, upload that reads it."""
import os, sys, time, json, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
from aten_b200 import PathTracing
from tests import oracle_lib as ol
import flat_scene as fs
scene = fs.flatten_instances(ol.ref_scene("dragon_grid10", 8, 8))
path = os.path.join(tempfile.mkdtemp(), "c5.atnenc")
out = {"triangles": int(len(scene.triangles))}
def upload(cache):
    with PathTracing(0) as r:
        t0 = time.perf_counter(); r.UpdateSceneData(scene, cache_path=cache); r.synchronize()
        return round(time.perf_counter() - t0, 3), r.cache_hit
upload(None)                                   # warm-up (CUDA context, allocator)
out["no_cache_s"], _ = upload(None)
out["write_cache_s"], hit = upload(path); assert hit is False
out["cache_file_mb"] = round(os.path.getsize(path) / 1e6, 1)
out["read_cache_s"], hit = upload(path); assert hit is True
print(json.dumps(out), flush=True)
