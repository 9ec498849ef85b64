"""Host<->device copy rate per rank with N ranks copying at once (the limiter of the end-to-end number at N=8: VERDICT 4c).
Every rank copies picture-sized chunks (2 MB up, 1 MB down: a 1080p luma plane and its result slab) from / to pinned
memory on two streams at the same time, all ranks between the same two barriers.
    python tools/h2d_scale.py                       (one rank)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29655 tools/h2d_scale.py
Prints one JSON line: per-rank and aggregate GB/s, up, down, and both at once."""
import json
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
UP, DOWN, N = 2 * 1024 * 1024, 1024 * 1024, 400
hup = torch.empty(16 * UP, dtype=torch.uint8).pin_memory()
hdn = torch.empty(16 * DOWN, dtype=torch.uint8).pin_memory()
dup = torch.empty(16 * UP, dtype=torch.uint8, device="cuda")
ddn = torch.empty(16 * DOWN, dtype=torch.uint8, device="cuda")
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()


def run(up, down):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s_up.wait_event(e0)
    s_dn.wait_event(e0)
    for i in range(N):
        k = i % 16
        if up:
            with torch.cuda.stream(s_up):
                dup[k * UP:(k + 1) * UP].copy_(hup[k * UP:(k + 1) * UP], non_blocking=True)
        if down:
            with torch.cuda.stream(s_dn):
                hdn[k * DOWN:(k + 1) * DOWN].copy_(ddn[k * DOWN:(k + 1) * DOWN], non_blocking=True)
    torch.cuda.current_stream().wait_stream(s_up)
    torch.cuda.current_stream().wait_stream(s_dn)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    s = ms.item() * 1e-3
    return {"up_gbs_per_rank": N * UP / s / 1e9 if up else 0.0, "down_gbs_per_rank": N * DOWN / s / 1e9 if down else 0.0, "ms": ms.item()}


for _ in range(2):
    run(True, True)
res = {"ranks": world, "chunk_up_bytes": UP, "chunk_down_bytes": DOWN, "copies": N,
       "up_only": run(True, False), "down_only": run(False, True), "both": run(True, True)}
res["aggregate_both_gbs"] = world * (res["both"]["up_gbs_per_rank"] + res["both"]["down_gbs_per_rank"])
if rank == 0:
    print(json.dumps(res))
if world > 1:
    dist.destroy_process_group()
