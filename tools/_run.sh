(time timeout 900 python -m pytest tests -m gpu -x -q) > gpurun_out/r9_pytest.log 2>&1; tail -5 gpurun_out/r9_pytest.log
show() { python -c "
import json,sys
d=json.load(open(sys.argv[1]))
print(sys.argv[1], round(d['value'],1), round(d['e2e']['value'],1), d['roofline']['kernel'], round(d['roofline']['frac'],4), {k:round(v['ms_per_step'],3) for k,v in d['roofline']['kernels'].items()})
" $1; }
for a in "1080p-m4 iid" "1080p-m4 motion" "4k-m0 iid" "4k-m2 iid" "1080p-m8 motion"; do
  set -- $a
  timeout 600 python bench.py --workload $1 --kind $2 --no-cpu > gpurun_out/r9_$1_$2.json 2> gpurun_out/r9_$1_$2.err; show gpurun_out/r9_$1_$2.json
done
