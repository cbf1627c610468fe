import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from tests.conftest import Golden
from tests import oracle_lib as ol
from aten_b200 import PathTracing
g = Golden()
for name in g.scenes:
    r = PathTracing(0); sc = g.scene(name); r.UpdateSceneData(sc)
    ref = g.trace(name); valid = (ref["flags"] & 1) != 0
    rays = np.ascontiguousarray(ref["ray"][valid]); want = np.ascontiguousarray(ref["hit"][valid])
    got = r.trace_closest(rays)
    hit = want["objid"] >= 0
    bad = np.where((got["objid"] != want["objid"]) | (hit & (got["primid"] != want["primid"])) | (got["t"].view(np.uint32) != want["t"].view(np.uint32)))[0]
    print(name, "path rays", len(rays), "mismatches", len(bad))
    for i in bad[:10]:
        print("  ray", rays[i], "\n   got", got[i], "\n  want", want[i])
        # is the other primitive at the same t ?
    r.close()
