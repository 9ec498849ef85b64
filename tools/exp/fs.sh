timeout 1400 python -m pytest tests -x -q -m gpu --deselect tests/test_gpu_encoder.py 2>&1 | tail -4
run() { python bench.py --steps 10 --warmup 3 --no-extra --no-cpu --workload $1 --kind $2 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);r=d['roofline'];print('$1 $2', round(d['value'],1),round(d['e2e']['value'],1),r['kernel'],round(r['frac'],3),{k:round(v['ms_per_step'],3) for k,v in r['kernels'].items()},round(r['whole_path_frac'],3),d['parity']['equal'])"; }
run 1080p-m4 iid; run 4k-m0 iid; run 4k-m2 iid; run 1080p-m4 motion; run 4k-m2 motion
