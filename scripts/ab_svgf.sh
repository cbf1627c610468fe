# A/B of build variants (aten_b200/variants/lib_<name>.so) on the SVGF sequence bench.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_svgf.py -x -q -m gpu 2>&1 | tail -3
echo default; python scripts/bench_configs.py --only svgf 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_frame'], d['filter_ms'])"
for f in aten_b200/variants/lib_*.so; do
  [ -e "$f" ] || continue
  echo $f; ATEN_B200_LIB=$f python scripts/bench_configs.py --only svgf 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_frame'], d['filter_ms'])"
done
