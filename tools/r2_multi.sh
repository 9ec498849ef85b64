#!/bin/bash
# Round 2's multi-GPU recipe: tools/r2_multi.sh N   (run under gpurun --gpus N). Outputs in gpurun_out/r2_<N>gpu_*.
N=${1:-2}
O=gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi topo -m > $O/r2_${N}gpu_topo.txt 2>&1
# host<->device copy rate per rank, one rank and N ranks at once
python tools/h2d_scale.py > $O/r2_${N}gpu_h2d_1rank.json 2> $O/r2_${N}gpu_h2d_1rank.err
$TR --master-port 29655 tools/h2d_scale.py > $O/r2_${N}gpu_h2d.json 2> $O/r2_${N}gpu_h2d.err
# the driver's command: 1080p replicas (weak) + one 4K preset-2 stream sharded over the ranks (strong) in the same line
$TR --master-port 29656 bench.py --gpus $N --steps 10 --warmup 3 > $O/r2_${N}gpu_bench_default.json 2> $O/r2_${N}gpu_bench_default.err
# BASELINE configs[4]: 4K preset 0 sharded by picture, worst-case input
$TR --master-port 29657 bench.py --gpus $N --workload 4k-m0 --kind iid --steps 4 --warmup 3 > $O/r2_${N}gpu_bench_4k-m0_iid.json 2> $O/r2_${N}gpu_bench_4k-m0_iid.err
# two-process IPC test and the in-process two-device test (skipped by the 1-GPU suite)
python -m pytest tests/test_gpu_two_processes.py "tests/test_gpu_parity_own_engine.py::test_two_devices_in_one_process" -x -q -m gpu > $O/r2_${N}gpu_tests.log 2>&1
tail -2 $O/r2_${N}gpu_tests.log
cat $O/r2_${N}gpu_h2d_1rank.json $O/r2_${N}gpu_h2d.json
for f in $O/r2_${N}gpu_bench_*.json; do tail -1 $f | python -c "
import json,sys
d=json.loads(sys.stdin.read())
s=d.get('sharded_4k') or {}
print('$f', d['scaling'], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), d['parity']['equal'], 'nvlink', d.get('nvlink_bytes_per_step'), d.get('peer_copy_verified'), '| sharded_4k', s.get('value'), (s.get('e2e') or {}).get('value'), (s.get('parity') or {}).get('equal'), s.get('nvlink_bytes_per_step'), s.get('peer_copy_verified'))"; done
