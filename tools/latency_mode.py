#!/usr/bin/env python
"""Latency mode on N GPUs of one box: ONE study spread over the ranks (home rank projects, 4 MB image broadcast, group
models round-robin over the ranks, label masks gathered on the home rank over NCCL).  Run under torchrun:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/latency_mode.py

Checks that the gathered result equals the single-GPU result bit for bit, then reports the p50 device-resident latency of
one study (volume already in the home GPU's HBM; wall clock on the home rank between two device synchronisations) next
to the single-GPU latency of the same study.
"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist

import bench


def main():
    from totalsegmentator2d_b200 import shard
    from totalsegmentator2d_b200.tool import TS2D
    rank, world, local = shard.rank_world()
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    small = "--small" in sys.argv
    shape = (96, 64, 160) if small else bench.VOL_SHAPE
    tmp = tempfile.TemporaryDirectory()
    bench.write_zoo(tmp.name, small, True)
    ts = TS2D(key="ts2d", models_root=tmp.name, param={"b200.precision": os.environ.get("TS2D_PRECISION", "fp16")})
    home = 0
    vol = torch.from_numpy(bench.synthetic_volume(20260000, shape)).cuda() if rank == home else None

    def sharded():
        return ts.predict_volume_sharded(vol, bench.SPACING, None, home=home)

    # parity: sharded == single GPU on the home rank
    res = sharded()
    ok = True
    if rank == home:
        ref = ts.predict_volume_device(vol, bench.SPACING, None)
        ok = bool(torch.equal(res["mask"], ref["mask"]) and torch.equal(res["present"], ref["present"]))
    for _ in range(3):
        sharded()
    lat = []
    for _ in range(9):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sharded()
        torch.cuda.synchronize()
        lat.append(1000.0 * (time.perf_counter() - t0))
    single = []
    if rank == home:
        from totalsegmentator2d_b200 import ops
        for _ in range(3):
            ts.predict_volume_device(vol, bench.SPACING, None)
        for _ in range(9):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = ts.predict_volume_device(vol, bench.SPACING, None)
            ops.merge_present_order(r["mask"], r["present"], r["offsets"])
            torch.cuda.synchronize()
            single.append(1000.0 * (time.perf_counter() - t0))
        print(json.dumps({"mode": "latency (one study over the GPUs, NCCL mask gather)", "n_gpus": world, "bit_identical_to_single_gpu": ok,
                          "latency_p50_ms": float(np.median(lat)), "single_gpu_latency_p50_ms": float(np.median(single)),
                          "models_per_rank": [len(shard.owned_models(len(ts.models), r, world, home)) for r in range(world)],
                          "volume": list(shape), "data": "synthetic, resident in the home GPU's HBM"}))
    ts.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not ok:
        sys.exit(1)


if __name__ == "__main__":
    main()
