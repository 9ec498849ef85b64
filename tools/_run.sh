env | grep -i nccl; nproc; lscpu | grep -i "numa\|socket\|model name" 
for i in 0 1 2 3 4 5 6 7; do python - <<PY
import torch
p=torch.cuda.get_device_properties($i)
path=f"/sys/bus/pci/devices/{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0/"
import os
print($i, path, open(path+"local_cpulist").read().strip() if os.path.exists(path+"local_cpulist") else "missing", open(path+"numa_node").read().strip() if os.path.exists(path+"numa_node") else "")
PY
done 2>&1 | head -10
run() { n=$1; wl=$2; kind=$3; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --workload $wl --kind $kind --no-cpu > gpurun_out/r14_n${n}_${wl}_${kind}.json 2> gpurun_out/r14_n${n}_${wl}_${kind}.err; python -c "
import json,sys
for line in open(sys.argv[1]):
    if line.startswith('{'):
        d=json.loads(line); print(sys.argv[1], 'value', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'n', d['n_gpus'], d.get('host'))
    else: print('non-json', line[:80])
" gpurun_out/r14_n${n}_${wl}_${kind}.json || tail -5 gpurun_out/r14_n${n}_${wl}_${kind}.err; }
run 8 1080p-m4 motion
