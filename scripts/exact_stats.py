import os, sys, json
sys.path.insert(0, os.getcwd())
import numpy as np
from aten_b200 import PathTracing, scene_io as sio
scene = sio.load_scene("tests/golden/sponza_atrium.atnscene.gz").with_size(1920, 1080)
r = PathTracing(0)
r.UpdateSceneData(scene, seeds=sio.mt19937_seeds(1920*1080))
r.set_gather_mode(2)
for spp in (4, 4):
    r.reset(); r.render_region(1920,1080,0,0,1920,1080,0,spp,5,3)
    st = r.stats()
print(json.dumps(st))
print("exact share", st["exact_rays"]/(st["closest_rays"]+st["shadow_rays"]))
r.close()
