#!/usr/bin/env python
"""Per-launch timing of one U-Net forward (CUDA events on the launching stream, warm) against both rooflines.

    python tools/layer_profile.py [--batch 12] [--heads 26] [--precision bf16] [--reps 3] [--out profiles/x.txt]

One CS-512 group model (ribs-shaped, 26 heads), `batch` tiles of 512x512x2.
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
    ap.add_argument("--batch", type=int, default=12)
    ap.add_argument("--heads", type=int, default=26)
    ap.add_argument("--precision", default="fp16")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--patch", type=int, default=512)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from totalsegmentator2d_b200 import synth
    from totalsegmentator2d_b200.core.inference.predictor import B200nnUNetPredictor
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0, "bf16_tflops_sustained": 1400.0}
    arch = synth.cs512_arch()
    labels = [f"l{i}" for i in range(a.heads)]
    plans = synth.make_plans(arch, (a.patch, a.patch), (1.5, 1.5), ["ZScoreNormalization"] * 2)
    dj = synth.make_dataset_json(["mean", "max"], labels)
    sd = synth.init_state_dict(arch, 2, a.heads, 1003)
    p = B200nnUNetPredictor(precision=a.precision, max_batch=a.batch)
    p.manual_initialization(plans, "2d", dj, [sd], (0, 1))
    dt = torch.bfloat16 if a.precision == "bf16" else torch.float16
    x = torch.randn(a.batch, a.patch, a.patch, 2, device="cuda").to(dt)
    out = torch.empty((a.batch, a.heads, a.patch, a.patch), dtype=torch.float32, device="cuda")
    for _ in range(2):
        p.predict_tile_logits(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    from totalsegmentator2d_b200 import _lib as L
    st = L.stream_ptr()
    tot = []
    for _ in range(a.reps):
        e0.record()
        L.check(p._lib.ts2d_unet_forward(p._handles[0], x.data_ptr(), a.batch, st))
        e1.record()
        torch.cuda.synchronize()
        tot.append(e0.elapsed_time(e1))
    p.set_profiling(True)
    acc = None
    for _ in range(a.reps):
        L.check(p._lib.ts2d_unet_forward(p._handles[0], x.data_ptr(), a.batch, st))
        torch.cuda.synchronize()
        prof = p.layer_profile()
        if acc is None:
            acc = [[n, k, [ms], fl, by] for n, k, ms, fl, by in prof]
        else:
            for r, (n, k, ms, fl, by) in zip(acc, prof):
                r[2].append(ms)
    p.set_profiling(False)
    lines = []
    hbm, tf = peaks["hbm_gbs"], peaks["bf16_tflops_sustained"]
    lines.append(f"# one CS-{a.patch} forward, batch {a.batch}, {a.precision}; whole forward (no profiling events) {np.median(tot):.3f} ms; "
                 f"peaks: HBM {hbm} GB/s, bf16 sustained {tf} TFLOP/s")
    lines.append(f"{'launch':<16}{'kind':<6}{'ms':>9}{'GFLOP':>10}{'TFLOP/s':>10}{'%tensor':>9}{'MB(alg)':>10}{'GB/s':>9}{'%hbm':>7}{'t_min_ms':>10}{'x t_min':>8}")
    sums = {0: 0.0, 1: 0.0, 2: 0.0}
    tmin_sum = 0.0
    for n, k, msl, fl, by in acc:
        ms = float(np.median(msl))
        sums[k] += ms
        tfl = fl / ms / 1e9 if ms > 0 else 0
        gbs = by / ms / 1e6 if ms > 0 else 0
        tmin = max(fl / (tf * 1e12), by / (hbm * 1e9)) * 1e3
        tmin_sum += tmin
        lines.append(f"{n:<16}{['conv','up','norm'][k]:<6}{ms:9.4f}{fl/1e9:10.2f}{tfl:10.1f}{100*tfl/tf:9.1f}{by/1e6:10.1f}{gbs:9.0f}{100*gbs/hbm:7.1f}{tmin:10.4f}{ms/tmin if tmin>0 else 0:8.1f}")
    lines.append(f"# sum conv {sums[0]:.3f} ms, transposed conv {sums[1]:.3f} ms, norm-apply {sums[2]:.3f} ms; sum of per-launch roofline minima {tmin_sum:.3f} ms")
    txt = "\n".join(lines)
    print(txt)
    if a.out:
        os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
        open(a.out, "w").write(txt + "\n")


if __name__ == "__main__":
    main()
