run() { ME_B200_TRACE_LIB=$2 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('$1', round(d['value']),round(d['e2e']['value']),d['roofline']['kernels']['k_hme']['ms_per_step'],d['parity']['equal'])"; }
for v in 128 96 80 64; do run regs$v tools/exp/lib$v.so; done
