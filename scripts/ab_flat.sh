# A/B of kernel variants on config C5 (100 distinct dragon meshes, streaming form of the traversal kernel)
run() { ATEN_B200_LIB=$PWD/$1 python scripts/bench_configs.py --only flat --rays 40000000 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['closest']['mrays_per_s'], d['any']['mrays_per_s'])"; }
run aten_b200/libaten_b200.so
for v in "$@"; do run aten_b200/variants/lib_$v.so; done
