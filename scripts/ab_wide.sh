# A/B of traversal build variants (aten_b200/variants/lib_<name>.so) on the bench workload.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -15 > gpurun_out/ab_tests.log
python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ab_default.json 2> gpurun_out/ab_default.err
for f in aten_b200/variants/lib_*.so; do
  [ -e "$f" ] || continue
  v=$(basename $f .so); v=${v#lib_}
  ATEN_B200_LIB=$f python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err
done
cat gpurun_out/ab_tests.log
