"""CPU, world_size 2 over gloo: the replica partition and the cross-rank timing reduction bench.py uses at N > 1."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rebvio_b200 import replicas


def test_partition_and_seeds():
    assert replicas.sequence_seeds(0)[:2] == [1000, 1001] and replicas.sequence_seeds(3)[0] == 1000 + 192
    all_seeds = [s for r in range(8) for s in replicas.sequence_seeds(r)]
    assert len(set(all_seeds)) == 512                       # no sequence is replayed twice
    assert replicas.owner_of(0, 8) == (0, 0) and replicas.owner_of(130, 8) == (2, 2)
    with pytest.raises(ValueError):
        replicas.owner_of(512, 8)
    parts = replicas.partition(10, 4)
    assert [len(p) for p in parts] == [3, 3, 2, 2] and [i for p in parts for i in p] == list(range(10))
    assert replicas.reduce_scalar(3.5, "max") == 3.5        # no process group: identity


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    seeds = replicas.sequence_seeds(rank, lanes=4)
    # ranks "process" different amounts of time; the job rate uses the slowest rank
    rate = replicas.whole_job_rate(units_this_rank=len(seeds) * 10, seconds_this_rank=1.0 + rank)
    gathered = [None] * world
    dist.all_gather_object(gathered, seeds)
    dist.barrier()
    if rank == 0:
        out.put((rate, gathered))
    dist.destroy_process_group()


def test_two_rank_replicas_over_gloo():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs: p.start()
    rate, gathered = q.get(timeout=120)
    for p in procs: p.join(timeout=60)
    assert all(p.exitcode == 0 for p in procs)
    assert rate == pytest.approx(80 / 2.0)                  # 2 ranks x 4 lanes x 10 frames / max(1 s, 2 s)
    assert gathered[0] == [1000, 1001, 1002, 1003] and gathered[1] == [1064, 1065, 1066, 1067]
    assert not set(gathered[0]) & set(gathered[1])
