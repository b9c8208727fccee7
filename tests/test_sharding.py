"""CPU, world_size 2 over gloo: container round-robin + all-gather of compressed sizes -> output offsets."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cramtools_b200 import sharding


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, sizes, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    mine = sharding.local_ordinals(len(sizes), rank, world)
    local = torch.tensor([sizes[g] for g in mine], dtype=torch.int64)
    offs = sharding.allgather_offsets(local, rank, world, base=26)
    out[rank] = (mine, offs.tolist())
    dist.destroy_process_group()


def test_round_robin_ownership():
    assert [sharding.container_owner(k, 4) for k in range(6)] == [0, 1, 2, 3, 0, 1]
    assert sharding.local_ordinals(7, 1, 2) == [1, 3, 5]


def test_offsets_single_rank():
    s = torch.tensor([5, 7, 11], dtype=torch.int64)
    assert sharding.allgather_offsets(s, 0, 1, base=100).tolist() == [100, 105, 112]


def test_allgather_offsets_world2_gloo():
    world = 2
    rng = np.random.default_rng(1)
    sizes = [int(x) for x in rng.integers(1000, 900000, size=12)]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), sizes, out), nprocs=world, join=True)
    expect = np.concatenate([[0], np.cumsum(sizes)[:-1]]) + 26
    for rank in range(world):
        mine, offs = out[rank]
        assert offs == [int(expect[g]) for g in mine]
