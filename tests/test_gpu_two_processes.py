"""SURVEY 8(e) as north_star states it: one process per GPU, pyramids replicated GPU to GPU. Two processes on two GPUs: rank 0
owns (uploads) the reference pictures, rank 1 owns the picture to search; rank 1 pulls the references' pyramids out of rank
0's pool through the CUDA IPC mapping (device-to-device copy over NVLink, ordered by rank 0's interprocess event), reads one
back (identical to the CPU-built pyramid) and runs ME against them: bit-exact with the oracle. Needs two GPUs: skipped on
the one-GPU box; the log of the two-GPU lease is kept under profiles/."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

W, H, N, PRESET, POC = 640, 360, 17, 4, 4


def _worker(rank, world, port, ret):
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, here)
    sys.path.insert(0, os.path.join(here, "..", "svt-av1-psyex_b200", "python"))
    import torch
    import torch.distributed as dist
    from checkers import Oracle
    from helpers import PictureSet, derive, used_refs
    from me_b200 import planes
    from me_b200.engine import Engine
    torch.cuda.set_device(rank)
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    oracle = Oracle()
    ps = PictureSet(W, H, N, "motion", oracle)
    cfg, p, l0, l1 = derive(PRESET, W, H, POC, N)
    refs = sorted(set(l0 + l1))
    eng = Engine(W, H, 4, 3, pictures_in_flight=12, devices=[rank])
    try:
        handles = [None] * world
        dist.all_gather_object(handles, eng.ipc_export_pool())
        for r, (h, sb) in enumerate(handles):
            if r != rank:
                eng.ipc_open_peer(r, h, sb)
        for rep in range(3):  # the same pool slots and interprocess events are used again and again
            base = 1000 * rep
            mine = {}
            if rank == 0:
                for k in refs:
                    eng.register_picture(base + k, ps.get(k).planes[0], W, H)
                    mine[k] = eng.ipc_picture(base + k)
            infos = [None] * world
            dist.all_gather_object(infos, mine)
            if rank == 1:
                for k in refs:
                    slot, evh = infos[0][k]
                    eng.register_picture_from_peer(base + k, 0, slot, evh, W, H)
                eng.register_picture(base + POC, ps.get(POC).planes[0], W, H)
                pyr = eng.read_pyramid(base + refs[0], W, H)
                ok_pyr = all(np.array_equal(pyr[lvl], ps.get(refs[0]).planes[lvl]) for lvl in range(3))
                q = type(p).from_buffer_copy(p)
                q.picture_number = base + POC
                got = eng.me_picture(base + POC, W, H, q, [[base + k for k in l0], [base + k for k in l1]])
                exp = oracle.me_picture(ps.get(POC), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
                bad = planes.compare_outputs(exp, got, used_refs=used_refs(l0, l1))
                ret[f"rep{rep}"] = (ok_pyr, bad, eng.ipc_bytes_pulled())
                for k in refs + [POC]:
                    eng.release_picture(base + k)
            dist.barrier()  # rank 0 keeps its pictures until rank 1 is done with them
            if rank == 0:
                for k in refs:
                    eng.release_picture(base + k)
    finally:
        eng.close()
        dist.destroy_process_group()


def test_pyramids_pulled_from_another_process_over_nvlink():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(2, 29500 + os.getpid() % 2000, ret), nprocs=2, join=True)
        assert sorted(ret.keys()) == ["rep0", "rep1", "rep2"]
        pulled = 0
        for rep in range(3):
            ok_pyr, bad, nbytes = ret[f"rep{rep}"]
            assert ok_pyr and not bad, (rep, ok_pyr, bad)
            assert nbytes > pulled  # bytes really came through the peer mapping
            pulled = nbytes
