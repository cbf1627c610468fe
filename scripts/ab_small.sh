run() { ATEN_B200_LIB=$PWD/$1 python scripts/bench_configs.py --only small 2>/dev/null | python -c "
import json,sys
out=[]
for l in sys.stdin:
    d=json.loads(l); out.append(str(d.get('ms_per_step', d.get('ms_per_frame'))))
print('$1', ' '.join(out))"; }
run aten_b200/libaten_b200.so
for v in "$@"; do run aten_b200/variants/lib_$v.so; done
