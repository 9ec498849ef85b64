N=${1:-8}
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29656 bench.py --gpus $N --steps 10 --warmup 3 > $O/r2_${N}gpu_bench_default.json 2> $O/r2_${N}gpu_bench_default.err
$TR --master-port 29657 bench.py --gpus $N --workload 4k-m0 --kind iid --steps 4 --warmup 3 > $O/r2_${N}gpu_bench_4k-m0_iid.json 2> $O/r2_${N}gpu_bench_4k-m0_iid.err
for f in $O/r2_${N}gpu_bench_*.json; do tail -1 $f | python -c "
import json,sys
d=json.loads(sys.stdin.read())
s=d.get('sharded_4k') or {}
print('$f', d['scaling'], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['parity']['equal'], 'nvlink', d.get('nvlink_bytes_per_step'), d.get('peer_copy_verified'), '| sharded_4k', s.get('value'), (s.get('e2e') or {}).get('value'), (s.get('parity') or {}).get('equal'), s.get('nvlink_bytes_per_step'), s.get('peer_copy_verified'))"; done
grep -h "Error" $O/r2_${N}gpu_bench_*.err | sort | uniq -c | head
