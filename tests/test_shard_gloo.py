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


def _latency_worker(rank, world, port, home, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        widths = [18, 23, 24, 26, 26]                      # ts2d-v2 heads per group model
        H, W = 20, 12
        img = torch.arange(2 * H * W, dtype=torch.float32).reshape(2, H, W) if rank == home else None
        img, meta = shard.broadcast_image(img, home, meta=(1.5, 0.75))
        mine = shard.owned_models(len(widths), rank, world, home)
        # "model" k marks its window with k + 1 wherever the broadcast image is odd
        odd = (img[0].to(torch.int64) % 2 == 1)
        local = {k: (odd[..., None] * (k + 1)).expand(H, W, widths[k]).to(torch.uint8).contiguous() for k in mine}
        present = {k: torch.full((widths[k],), k + 1, dtype=torch.uint8) for k in mine}
        mask = shard.gather_windows(local, len(widths), widths, (H, W), home, dtype=torch.uint8)
        pres = shard.gather_windows(present, len(widths), widths, (), home, dtype=torch.uint8)
        q.put((rank, mine, meta, None if mask is None else mask.clone(), None if pres is None else pres.clone(), tuple(img.shape)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,home", [(2, 0), (2, 1), (3, 1)])
def test_latency_mode_broadcast_and_gather(world, home):
    """one study over `world` ranks: image broadcast from the home rank, per-model windows gathered back in model order"""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_latency_worker, args=(r, world, port, home, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = {r[0]: r for r in (q.get(timeout=120) for _ in range(world))}
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    widths = [18, 23, 24, 26, 26]
    assert sorted(k for r in out.values() for k in r[1]) == [0, 1, 2, 3, 4], "every group model runs exactly once"
    assert 0 in out[home][1], "the home rank runs the first model itself"
    for rank, (_, mine, meta, mask, pres, shape) in out.items():
        assert shape == (2, 20, 12) and meta[:2] == [1.5, 0.75]
        assert (mask is None) == (rank != home)
    mask, pres = out[home][3], out[home][4]
    odd = (torch.arange(20 * 12).reshape(20, 12) % 2 == 1)
    off = 0
    for k, w in enumerate(widths):
        assert torch.equal(mask[..., off:off + w], (odd[..., None] * (k + 1)).expand(20, 12, w).to(torch.uint8))
        assert torch.equal(pres[off:off + w], torch.full((w,), k + 1, dtype=torch.uint8))
        off += w


def test_model_owner_round_robin():
    assert [shard.model_owner(k, 8, 3) for k in range(5)] == [3, 4, 5, 6, 7]
    assert shard.owned_models(5, 0, 2) == [0, 2, 4] and shard.owned_models(5, 1, 2) == [1, 3]
    assert shard.owned_models(5, 0, 1) == [0, 1, 2, 3, 4]
    with pytest.raises(ValueError):
        shard.model_owner(0, 2, 2)


def test_gpu_numa_binding_helpers(tmp_path, monkeypatch):
    """bind_to_gpu: CPUs of the GPU's NUMA node from (a fake) sysfs, an even split when sysfs is silent, off with TS2D_BIND=0"""
    import os
    from totalsegmentator2d_b200 import shard
    assert shard._parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    dev = tmp_path / "bus/pci/devices/0000:1b:00.0"
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices/system/node/node1"
    node.mkdir(parents=True)
    (node / "cpulist").write_text("4-7\n")
    assert shard.gpu_numa_cpus("0000:1B:00.0", str(tmp_path)) == [4, 5, 6, 7]
    assert shard.gpu_numa_cpus("1b:00.0", str(tmp_path)) == [4, 5, 6, 7]
    assert shard.gpu_numa_cpus("0000:99:00.0", str(tmp_path)) == []
    (dev / "numa_node").write_text("-1\n")
    assert shard.gpu_numa_cpus("0000:1b:00.0", str(tmp_path)) == []
    if not hasattr(os, "sched_setaffinity"):
        return
    before = sorted(os.sched_getaffinity(0))
    try:
        assert shard.bind_to_gpu(0, 1) == []                                   # a single local rank is left alone
        monkeypatch.setenv("TS2D_BIND", "0")
        assert shard.bind_to_gpu(1, 2) == []
        monkeypatch.delenv("TS2D_BIND")
        if len(before) >= 2:
            got = shard.bind_to_gpu(1, 2, "0000:1b:00.0", str(tmp_path))       # sysfs silent -> even split
            per = len(before) // 2
            assert got == before[per:2 * per] and sorted(os.sched_getaffinity(0)) == got
    finally:
        os.sched_setaffinity(0, before)
