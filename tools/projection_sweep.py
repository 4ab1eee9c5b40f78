#!/usr/bin/env python
"""projection (coronal mean + max, one pass) over device-resident synthetic volumes, 256^3 .. 512x512x2000, int16 and
fp32: achieved GB/s = algorithmic bytes (X*Y*Z*sizeof(voxel) + 2*X*Z*4) / median kernel time (CUDA events), against the measured HBM peak.

    python tools/projection_sweep.py [--out profiles/x.txt]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    ap.add_argument("--reps", type=int, default=15)
    a = ap.parse_args()
    # under torchrun (SURVEY 8d config 5 at 2 / 4 / 8 GPUs): every rank sweeps its own volumes on its own GPU at the same time (the
    # path has no exchange step); rank 0 prints the per-rank GB/s and their sum
    world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
    from totalsegmentator2d_b200 import ops
    pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
    hbm = float(json.load(open(pk))["hbm_gbs"]) if os.path.exists(pk) else 6650.0
    lines = ["# projection sweep (mean + max in one pass, reduce axis = y), device-resident, measured HBM peak {:.1f} GB/s".format(hbm),
             "%-16s %-6s %10s %10s %9s %7s" % ("volume z,y,x", "dtype", "MB(alg)", "ms(median)", "GB/s", "%hbm")]
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")        # > L2: the 256^3 volume would otherwise be served from L2
    for (z, y, x) in [(256, 256, 256), (500, 512, 512), (1000, 512, 512), (2000, 512, 512)]:
        for dt in (torch.int16, torch.float32):
            g = torch.Generator(device="cuda").manual_seed(z + y)
            vol = (torch.randn((z, y, x), device="cuda", generator=g) * 450 - 300).clamp_(-1024, 3071).to(dt)
            out = torch.empty((2, z, x), dtype=torch.float32, device="cuda")
            ms = []
            if world > 1:
                dist.barrier()
            for _ in range(a.reps):
                flush.fill_(1)
                e0.record()
                ops.project(vol, ["mean", "max"], reduce_dim=1, h_dim=0, out=out)
                e1.record()
                torch.cuda.synchronize()
                ms.append(e0.elapsed_time(e1))
            t = float(np.median(ms[3:]))
            nbytes = z * y * x * vol.element_size() + 2 * z * x * 4
            gbs = nbytes / (t * 1e-3) / 1e9
            # spot check against torch on the device
            ref_max = vol.float().amax(dim=1)
            assert torch.equal(out[1], ref_max)
            if world > 1:
                tt = torch.tensor([gbs], device="cuda", dtype=torch.float64)
                allv = [torch.zeros_like(tt) for _ in range(world)]
                dist.all_gather(allv, tt)
                per = [float(v.item()) for v in allv]
                lines.append("%-16s %-6s %10.1f %10.4f %9.0f %7.1f   %d GPUs: sum %.0f GB/s, min rank %.0f, max rank %.0f" % (
                    f"{z},{y},{x}", str(dt).replace("torch.", ""), nbytes / 1e6, t, gbs, 100 * gbs / hbm, world, sum(per), min(per), max(per)))
            else:
                lines.append("%-16s %-6s %10.1f %10.4f %9.0f %7.1f" % (f"{z},{y},{x}", str(dt).replace("torch.", ""), nbytes / 1e6, t, gbs, 100 * gbs / hbm))
            del vol, out
    text = "\n".join(lines)
    if rank == 0:
        print(text)
        if a.out:
            open(a.out, "w").write(text + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
