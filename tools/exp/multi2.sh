python -m pytest tests/test_gpu_two_processes.py "tests/test_gpu_parity_own_engine.py::test_two_devices_in_one_process" -x -q -m gpu 2>&1 | tail -5
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --workload 4k-m2 --steps 10 --warmup 3 2> gpurun_out/multi2.err | tail -1 | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(json.dumps({k:d.get(k) for k in ('value','scaling','n_gpus','ms_per_step','e2e','parity','nvlink_bytes_per_step','peer_copy_verified','pictures_per_rank_per_step')})[:1500])"
python bench.py --workload 4k-m2 --steps 10 --warmup 3 --no-cpu --no-extra 2>/dev/null | tail -1 | python -c "
import json,sys;d=json.loads(sys.stdin.read());print('N=1', json.dumps({k:d.get(k) for k in ('value','ms_per_step','e2e')})[:600])"
