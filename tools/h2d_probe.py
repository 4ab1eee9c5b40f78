#!/usr/bin/env python
"""Host->device copy ceiling of the box: N ranks (torchrun) each looping pinned cudaMemcpyAsync of one study-sized buffer
(524 MB int16 volume, bench.py's per-study upload), all ranks at once.  Prints per-rank and aggregate GB/s for
  whole   one cudaMemcpyAsync per study on one stream
  chunk4  4 chunks per study alternating over 2 streams
  d2h     the 59.9 MB merged mask going the other way at the same time as `whole` (full-duplex check)
so that bench.py's e2e studies/s x 524 MB can be read as a fraction of what the machine can move.

    python tools/h2d_probe.py                      # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/h2d_probe.py
"""
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    nbytes = 1000 * 512 * 512 * 2
    mbytes = 1000 * 512 * 117
    wc = os.environ.get("H2D_WC", "0") == "1"          # write-combined upload buffers (ops.UploadBuffer) instead of torch's pinned memory
    if wc:
        from totalsegmentator2d_b200 import ops
        import numpy as np
        keep = [ops.UploadBuffer((nbytes,), np.uint8, write_combined=True) for _ in range(2)]
        for k in keep:
            k.array[:] = 3
        host = [torch.from_numpy(k.array) for k in keep]
    else:
        host = [torch.empty(nbytes, dtype=torch.uint8).pin_memory() for _ in range(2)]
        for h in host:
            h.fill_(3)
    dev = [torch.empty(nbytes, dtype=torch.uint8, device="cuda") for _ in range(2)]
    mdev = torch.zeros(mbytes, dtype=torch.uint8, device="cuda")
    mhost = torch.empty(mbytes, dtype=torch.uint8).pin_memory()
    s = [torch.cuda.Stream() for _ in range(3)]
    reps = int(os.environ.get("H2D_REPS", "24"))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize()

    def run(mode):
        barrier()
        t0 = time.perf_counter()
        for r in range(reps):
            if mode in ("whole", "d2h"):
                with torch.cuda.stream(s[0]):
                    dev[r & 1].copy_(host[r & 1], non_blocking=True)
                if mode == "d2h":
                    with torch.cuda.stream(s[2]):
                        mhost.copy_(mdev, non_blocking=True)
            else:
                q = nbytes // 4
                for c in range(4):
                    with torch.cuda.stream(s[c & 1]):
                        dev[r & 1][c * q:(c + 1) * q].copy_(host[r & 1][c * q:(c + 1) * q], non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        return reps * nbytes / dt / 1e9

    out = {}
    for mode in ("whole", "chunk4", "d2h"):
        run(mode)                       # warm-up
        mine = run(mode)
        if world > 1:
            import torch.distributed as dist
            t = torch.tensor([mine], device="cuda", dtype=torch.float64)
            allv = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allv, t)
            per = [float(v.item()) for v in allv]
        else:
            per = [mine]
        # ranks run concurrently for (about) the same wall time: aggregate = sum of the per-rank rates
        out[mode] = {"per_rank_gbs": [round(v, 2) for v in per], "aggregate_gbs": round(sum(per), 2), "min_rank_gbs": round(min(per), 2)}
    if rank == 0:
        print(json.dumps({"probe": "h2d_ceiling", "write_combined": wc, "ranks": world, "bytes_per_copy": nbytes, "reps": reps, **out}))
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
