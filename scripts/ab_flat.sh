# A/B of traversal variants on the C5 scene (100 distinct dragon meshes, 30 M surface rays):
# the default build, and the build variants under aten_b200/variants/ (scripts/build_variants.py).
mkdir -p gpurun_out
R=${RAYS:-30000000}
ATEN_B200_STREAM=0 python scripts/bench_configs.py --only flat --rays $R > gpurun_out/abf_control_stream0.json 2> gpurun_out/abf_control.err
python scripts/bench_configs.py --only flat --rays $R > gpurun_out/abf_default.json 2> gpurun_out/abf_default.err
for f in aten_b200/variants/lib_*.so; do
  [ -e "$f" ] || continue
  v=$(basename $f .so); v=${v#lib_}
  ATEN_B200_LIB=$f python scripts/bench_configs.py --only flat --rays $R > gpurun_out/abf_$v.json 2> gpurun_out/abf_$v.err
done
for f in gpurun_out/abf_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    if "closest" in d: print("  closest", d["closest"]["mrays_per_s"], "any", d["any"]["mrays_per_s"], "nodes/ray", d["closest"]["quad_nodes_per_ray"], "parity", d["parity_vs_restatement"]["differing"])
    else: print("  bench", d["value"])
except Exception as e: print("  failed", e)
PY
done
