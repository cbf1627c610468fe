# A/B of build variants (aten_b200/variants/lib_<name>.so, scripts/build_variants.py) on the bench workload.
mkdir -p gpurun_out
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ab_default.json 2> gpurun_out/ab_default.err
for f in aten_b200/variants/lib_*.so; do
  [ -e "$f" ] || continue
  v=$(basename $f .so); v=${v#lib_}
  ATEN_B200_LIB=$f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
done
for f in gpurun_out/ab_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); print(sys.argv[1], d["value"], d["own_layout"])
except Exception as e: print(sys.argv[1], "failed", e)
PY
done
