"""Three SVGF frames at 1080p on the golden Sponza scene (the command the ncu captures of the image-space kernels run)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aten_b200 import PathTracing, scene_io as sio  # noqa: E402

sc = sio.load_scene(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "sponza_atrium.atnscene.gz")).with_size(1920, 1080)
r = PathTracing(0)
r.UpdateSceneData(sc)
for f in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    r.render_svgf(1920, 1080, 5, 3)
print("frames", r.GetFrameCount())
