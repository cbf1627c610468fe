run() { ATEN_B200_EXACT_TIES=0 ATEN_B200_LIB=$PWD/$1 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extra 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('exact-off $1', d['value'], d['ms_per_step'])"; }
run aten_b200/libaten_b200.so
for v in "$@"; do run aten_b200/variants/lib_$v.so; done
