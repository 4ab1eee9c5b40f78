#!/usr/bin/env python
"""Smallest program that runs the U-Net engine the way bench.py does: one CS-512 group model, `--batch` tiles of
512x512x2, `--forwards` forwards through the C ABI.  Used as the ncu target (B200_PROFILING.md recipe):

    python tools/ncu_forward.py && ncu --set full --clock-control none --import-source on \
        -k regex:conv_ -s 37 -c 37 -o gpurun_out/prof python tools/ncu_forward.py
(37 conv/transposed-conv launches per forward: skip the first forward, capture the second.)
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=12)
    ap.add_argument("--heads", type=int, default=26)
    ap.add_argument("--forwards", type=int, default=2)
    ap.add_argument("--precision", default="fp16")
    a = ap.parse_args()
    from totalsegmentator2d_b200 import _lib as L
    from totalsegmentator2d_b200 import synth
    from totalsegmentator2d_b200.core.inference.predictor import B200nnUNetPredictor
    arch = synth.cs512_arch()
    plans = synth.make_plans(arch, (512, 512), (1.5, 1.5), ["ZScoreNormalization"] * 2)
    dj = synth.make_dataset_json(["mean", "max"], [f"l{i}" for i in range(a.heads)])
    sd = synth.init_state_dict(arch, 2, a.heads, 1003)
    p = B200nnUNetPredictor(precision=a.precision, max_batch=a.batch)
    p.manual_initialization(plans, "2d", dj, [sd], (0, 1))
    dt = torch.bfloat16 if a.precision == "bf16" else torch.float16
    x = torch.randn(a.batch, 512, 512, 2, device="cuda").to(dt)
    st = L.stream_ptr()
    for _ in range(a.forwards):
        L.check(p._lib.ts2d_unet_forward(p._handles[0], x.data_ptr(), a.batch, st))
    torch.cuda.synchronize()
    print("ok: {} forwards at batch {}".format(a.forwards, a.batch))


if __name__ == "__main__":
    main()
