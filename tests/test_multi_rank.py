"""N > 1 host logic on the CPU: two gloo ranks shard the pictures of a mini-GOP, each runs ME of its shard (the
control-logic emulation stands in for the GPU), results are gathered and must equal the single-process run."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from helpers import Emulator, PictureSet, derive
from me_b200 import shard, synth
from checkers import Oracle

W, H, N, PRESET = 136, 72, 33, 8
POCS = [32 if g == 0 else 16 + g for g in synth.RA5_DECODE_ORDER]


def _checksum(poc, ps, emul):
    cfg, p, l0, l1 = derive(PRESET, W, H, poc, N)
    out = emul.me_picture(ps.get(poc), [[ps.get(k) for k in l0], [ps.get(k) for k in l1]], p)
    return int(out["me_mv_array"].astype(np.uint64).sum() * 31 + out["me_candidate_array"].astype(np.uint64).sum()
               + out["me_8x8_distortion"].astype(np.uint64).sum())


def _worker(rank, world, port, ret):
    sys.path.insert(0, os.path.dirname(__file__))
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    ps, emul = PictureSet(W, H, N, "motion", Oracle()), Emulator()
    mine = shard.pictures_for_rank(POCS, rank, world)
    local = torch.zeros(len(POCS), dtype=torch.int64)
    for poc in mine:
        local[POCS.index(poc)] = _checksum(poc, ps, emul)
    dist.all_reduce(local, op=dist.ReduceOp.SUM)  # every picture is owned by exactly one rank
    count = torch.tensor([len(mine)])
    dist.all_reduce(count)
    if rank == 0:
        ret["sums"] = local.tolist()
        ret["count"] = int(count.item())
    dist.destroy_process_group()


def test_two_ranks_cover_the_mini_gop_once():
    a, b = shard.pictures_for_rank(POCS, 0, 2), shard.pictures_for_rank(POCS, 1, 2)
    assert sorted(a + b) == sorted(POCS) and not set(a) & set(b)
    need = shard.pyramids_needed(a, lambda poc: derive(PRESET, W, H, poc, N)[2:])
    assert set(a) <= set(need) and all(0 <= k < N for k in need)


def test_cost_aware_dealing_balances_every_mini_gop():
    """shard.owners with a cost: every picture owned exactly once, every rank computes the same map, and within a mini-GOP no
    rank carries more than the lightest one plus one picture's cost (longest-processing-time-first bound)"""
    order = synth.decode_order(N)
    cost = {poc: 3.0 + (poc * 7 % 5) for poc in order}  # 3..7 "references"
    for world in (2, 3, 4, 8):
        own = shard.owners(order[1:], world, cost=lambda poc: cost[poc])
        assert sorted(own) == sorted(order[1:]) and set(own.values()) <= set(range(world))
        assert own == shard.owners(order[1:], world, cost=lambda poc: cost[poc])
        load = [0.0] * world
        for g0 in range(0, len(order) - 1, 16):
            for poc in order[1 + g0:1 + g0 + 16]:
                load[own[poc]] += cost[poc]
            assert max(load) - min(load) <= max(cost.values()), (world, load)
    # without a cost: the decode order dealt round-robin
    assert shard.owners(order, 4) == {poc: i % 4 for i, poc in enumerate(order)}


def test_gloo_world2_matches_single_process():
    ps, emul = PictureSet(W, H, N, "motion", Oracle()), Emulator()
    expect = [_checksum(poc, ps, emul) for poc in POCS]
    with mp.Manager() as m:
        ret = m.dict()
        port = 29500 + (os.getpid() % 2000)
        mp.spawn(_worker, args=(2, port, ret), nprocs=2, join=True)
        assert ret["count"] == len(POCS)
        assert ret["sums"] == expect
