run() { python bench.py --steps 10 --warmup 3 --no-extra --no-cpu 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('$1', round(d['value']),round(d['e2e']['value']),d['roofline']['kernels']['k_hme']['ms_per_step'],d['parity']['equal'])"; }
run base
CUDA_DEVICE_MAX_CONNECTIONS=32 run conn32
BENCH_WINDOW=6 run win6
BENCH_WINDOW=15 CUDA_DEVICE_MAX_CONNECTIONS=32 run win15conn32
python tools/host_overhead.py 2>&1 | tail -8
