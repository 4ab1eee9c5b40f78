#!/usr/bin/env python
"""Host time to issue ONE study's launches (queue empty at the start, no sync inside) against the device time of the same study.
    python tools/host_enqueue.py [--reps 5]"""
import argparse, os, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    from totalsegmentator2d_b200.tool import TS2D
    tmp = tempfile.TemporaryDirectory()
    bench.write_zoo(tmp.name, False, True)
    ts = TS2D(key="ts2d", models_root=tmp.name, param={"b200.precision": os.environ.get("TS2D_PRECISION", "fp16")})
    n_heads = ts.head_offsets()[-1]
    vol = torch.from_numpy(bench.synthetic_volume(20260000)).cuda()
    H, W = bench.VOL_SHAPE[0], bench.VOL_SHAPE[2]
    mask = torch.zeros((H, W, n_heads), dtype=torch.uint8, device="cuda")
    present = torch.zeros(n_heads, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        ts.predict_volume_device(vol, bench.SPACING, None, mask=mask, present=present)
    torch.cuda.synchronize()
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record()
        ts.predict_volume_device(vol, bench.SPACING, None, mask=mask, present=present)
        e1.record()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print("host enqueue {:.2f} ms, device {:.2f} ms".format(1000 * (t1 - t0), e0.elapsed_time(e1)))
    import cProfile, pstats
    pr = cProfile.Profile()
    torch.cuda.synchronize()
    pr.enable()
    ts.predict_volume_device(vol, bench.SPACING, None, mask=mask, present=present)
    pr.disable()
    torch.cuda.synchronize()
    pstats.Stats(pr).sort_stats("tottime").print_stats(18)


if __name__ == "__main__":
    main()
