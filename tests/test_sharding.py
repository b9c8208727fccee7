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


def _worker_abi(rank, world, port, sizes, recs, out):
    os.environ['MASTER_ADDR'] = '127.0.0.1'
    os.environ['MASTER_PORT'] = str(port)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    mine = sharding.local_ordinals(len(sizes), rank, world)          # 7 containers over 2 ranks: the ranks hold 4 and 3
    off, rec, tot = sharding.allgather_sizes([sizes[g] for g in mine], [recs[g] for g in mine], [150 * recs[g] for g in mine],
                                             rank, world, base_offset=26, base_record=1000)
    out[rank] = (mine, off.tolist(), rec.tolist(), tot)
    dist.destroy_process_group()


def test_allgather_sizes_c_abi_world2_gloo():
    """cramgpu_allgather_sizes (the C export a JVM host binds): container.offset and globalRecordCounter of every rank's
    containers, uneven container counts"""
    world = 2
    rng = np.random.default_rng(2)
    sizes = [int(x) for x in rng.integers(1000, 900000, size=7)]
    recs = [int(x) for x in rng.integers(1, 10001, size=7)]
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_abi, args=(world, _free_port(), sizes, recs, out), nprocs=world, join=True)
    eo = np.concatenate([[0], np.cumsum(sizes)[:-1]]) + 26
    er = np.concatenate([[0], np.cumsum(recs)[:-1]]) + 1000
    for rank in range(world):
        mine, off, rec, tot = out[rank]
        assert off == [int(eo[g]) for g in mine] and rec == [int(er[g]) for g in mine]
        assert tot == {'compressed_bytes': sum(sizes), 'n_records': sum(recs), 'n_bases': 150 * sum(recs)}


def test_allgather_sizes_single_rank():
    off, rec, tot = sharding.allgather_sizes([5, 7, 11], [2, 3, 4], [20, 30, 40], 0, 1, base_offset=100, base_record=1)
    assert off.tolist() == [100, 105, 112] and rec.tolist() == [1, 3, 6] and tot['n_bases'] == 90
