for v in c5 st10000 st25000 st50000; do echo $v; ME_B200_TRACE_LIB=tools/exp/lib_$v.so python tools/profile_step.py 5 motion 3840:2160:4:2:2 2>&1 | grep "^poc" | tail -3 | cut -c1-100; ME_B200_TRACE_LIB=tools/exp/lib_$v.so python tools/profile_step.py 5 motion 1080p-m4 2>&1 | grep "^poc" | tail -3 | cut -c1-100; done
run() { ME_B200_TRACE_LIB=$2 python bench.py --steps 20 --warmup 5 --no-extra --no-cpu --workload $3 --kind $4 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('$1 $3 $4', round(d['value'],1),round(d['e2e']['value'],1),{k:round(v['ms_per_step'],3) for k,v in d['roofline']['kernels'].items()},d['parity']['equal'])"; }
for v in c5 st10000 st25000; do run $v tools/exp/lib_$v.so 1080p-m4 motion; done
