import sys, os, concurrent.futures as cf
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from aten_b200 import build
V = dict(a.split("=", 1) for a in sys.argv[1:])   # name=DEF1:DEF2
def one(kv):
    name, defs = kv
    out = os.path.join(build.HERE, "variants", f"lib_{name}.so")
    build.build(force=True, defines=[d for d in defs.split(":") if d], out=out)
    return out
with cf.ThreadPoolExecutor(8) as ex:
    for o in ex.map(one, V.items()): print(o)
