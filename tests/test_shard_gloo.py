"""N > 1 host logic on CPU: world_size-2 gloo processes shard studies, reduce timings and gather records in study order."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from totalsegmentator2d_b200 import shard


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_items, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert shard.rank_world() == (rank, world, rank)
        mine = shard.shard_indices(n_items, rank, world)
        records = [{"study": i, "rank": rank, "labels": i * 7 % 5} for i in mine]
        everything = shard.gather_in_order(records, n_items)
        t_max = shard.reduce_max(10.0 + rank)            # the step time is the slowest rank's
        n_sum = shard.reduce_sum(len(mine))
        q.put((rank, mine, everything, t_max, n_sum))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_items", [7, 8, 1])
def test_two_ranks_shard_and_gather(n_items):
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_items, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    seen = sorted(i for _, mine, _, _, _ in out for i in mine)
    assert seen == list(range(n_items)), "every study is processed exactly once"
    for rank, mine, everything, t_max, n_sum in out:
        assert [r["study"] for r in everything] == list(range(n_items))
        assert all(r["rank"] == r["study"] % world for r in everything)
        assert t_max == 11.0 and n_sum == n_items
        assert abs(len(mine) - n_items / world) <= 0.5 + 1e-9


def test_single_process_fallbacks():
    assert shard.shard_indices(5, 0, 1) == [0, 1, 2, 3, 4]
    assert shard.shard(list("abcde"), 1, 2) == ["b", "d"]
    assert shard.reduce_max(3.5) == 3.5 and shard.reduce_sum(2) == 2.0
    assert shard.gather_in_order([1, 2], 2) == [1, 2]
    with pytest.raises(ValueError):
        shard.shard_indices(3, 2, 2)
