run() { ME_B200_TRACE_LIB=$2 python bench.py --steps 10 --warmup 3 --no-extra --no-cpu --workload $3 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('$1 $3', round(d['value'],1),round(d['e2e']['value'],1),d['roofline']['kernels']['k_hme']['ms_per_step'],d['parity']['equal'])"; }
for wl in 1080p-m4 1080p-m8 4k-m2 4k-m0; do run v1 tools/exp/libv1.so $wl; run v2 svt-av1-psyex_b200/libme_gpu.so $wl; done
