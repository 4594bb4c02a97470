run() { name=$1; shift; env "$@" python bench.py --steps 200 --warmup 10 --no-others --no-cpu --no-parity 2>/dev/null | python -c "
import sys,json
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('$name', round(d['value']/1e6,1), 'M ex/s', round(d['ms_per_step']*1e3,1), 'us/step', {k: round(v*1e3,1) for k,v in d['kernel_ms'].items()})"; }
run base X=1
for v in r64 r64g12 r64g16 r64i8 r56; do run $v WH_LIB=$PWD/wormhole_b200/variants/lib_$v.so; done
run base2 X=1
