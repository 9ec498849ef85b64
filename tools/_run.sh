for c in 2 3 4 6 8 12 15; do
for wl in 1080p-m4; do
BENCH_WINDOW=$c timeout 300 python bench.py --no-cpu --no-extra --steps 10 --workload $wl > gpurun_out/r23_${wl}_$c.json 2>/dev/null
python -c "
import json,sys
d=json.load(open(sys.argv[1])); print(sys.argv[1], round(d['value'],1), round(d['e2e']['value'],1))
" gpurun_out/r23_${wl}_$c.json
done; done
