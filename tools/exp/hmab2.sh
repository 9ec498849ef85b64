run() { ME_B200_TRACE_LIB=$2 python bench.py --steps ${5:-20} --warmup 5 --no-extra --no-cpu --workload $3 --kind $4 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);r=d['roofline'];print('$1 $3 $4', round(d['value'],1),round(d['e2e']['value'],1),r['kernel'],round(r['frac'],4),{k:round(v['ms_per_step'],3) for k,v in r['kernels'].items()},d['parity']['equal'])"; }
for v in "$@"; do run $v tools/exp/lib_$v.so 1080p-m8 motion; run $v tools/exp/lib_$v.so 1080p-m4 motion; done
