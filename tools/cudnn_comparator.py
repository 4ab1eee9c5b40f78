#!/usr/bin/env python
"""Kernel-quality bar (SURVEY 2.1 / BASELINE.md section 5): the SAME random-init CS-512 PlainConvUNet, batch 12
(3 tiles x 4 mirror passes) of 512 x 512 x 2, run by PyTorch + cuDNN in fp16 channels-last on the same B200, timed with CUDA events
next to this repo's engine.  Tools only: nothing in the package imports torch.nn or cuDNN.

    python tools/cudnn_comparator.py [--batch 12] [--reps 20]
"""
import argparse
import json
import os
import sys

import torch
import torch.nn as nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class Block(nn.Sequential):
    def __init__(self, cin, cout, stride):
        super().__init__(nn.Conv2d(cin, cout, 3, stride, 1, bias=True), nn.InstanceNorm2d(cout, eps=1e-5, affine=True), nn.LeakyReLU(0.01, inplace=True))


class PlainConvUNet(nn.Module):
    """dynamic_network_architectures PlainConvUNet, wired like oracle/nnunet_net.py (upstream state-dict names are not needed here)"""

    def __init__(self, feats, in_ch, heads):
        super().__init__()
        self.enc = nn.ModuleList()
        c = in_ch
        for s, f in enumerate(feats):
            self.enc.append(nn.Sequential(Block(c, f, 1 if s == 0 else 2), Block(f, f, 1)))
            c = f
        self.up, self.dec = nn.ModuleList(), nn.ModuleList()
        for d in range(len(feats) - 1):
            below, skip = feats[-(d + 1)], feats[-(d + 2)]
            self.up.append(nn.ConvTranspose2d(below, skip, 2, 2, bias=True))
            self.dec.append(nn.Sequential(Block(2 * skip, skip, 1), Block(skip, skip, 1)))
        self.head = nn.Conv2d(feats[0], heads, 1)

    def forward(self, x):
        skips = []
        for e in self.enc:
            x = e(x)
            skips.append(x)
        for d, (u, dec) in enumerate(zip(self.up, self.dec)):
            x = dec(torch.cat((u(x), skips[-(d + 2)]), 1))
        return self.head(x)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=12)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--engine", type=int, default=1, help="also time this repo's engine on the same batch")
    a = ap.parse_args()
    torch.backends.cudnn.benchmark = True
    feats = [32, 64, 128, 256, 512, 512, 512, 512]
    net = PlainConvUNet(feats, 2, 26).cuda().half().to(memory_format=torch.channels_last).eval()
    x = torch.randn(a.batch, 2, 512, 512, device="cuda", dtype=torch.float16).to(memory_format=torch.channels_last)
    out = {"batch": a.batch, "torch": torch.__version__, "cudnn": torch.backends.cudnn.version()}

    def timeit(fn, reps):
        for _ in range(5):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    with torch.no_grad():
        out["cudnn_fp16_channels_last_ms"] = timeit(lambda: net(x), a.reps)
        g = torch.cuda.CUDAGraph()
        sx = x.clone()
        with torch.cuda.graph(g):
            sy = net(sx)
        out["cudnn_fp16_channels_last_cudagraph_ms"] = timeit(lambda: g.replay(), a.reps)
        xn = x.contiguous(memory_format=torch.contiguous_format)
        netn = PlainConvUNet(feats, 2, 26).cuda().half().eval()
        out["cudnn_fp16_nchw_ms"] = timeit(lambda: netn(xn), a.reps)
    if a.engine:
        from totalsegmentator2d_b200 import _lib as L
        from totalsegmentator2d_b200 import synth
        from totalsegmentator2d_b200.core.inference.predictor import B200nnUNetPredictor
        arch = synth.cs512_arch()
        p = B200nnUNetPredictor(precision="fp16", max_batch=a.batch)
        p.manual_initialization(synth.make_plans(arch, (512, 512), (1.5, 1.5), ["ZScoreNormalization"] * 2),
                                "2d", synth.make_dataset_json(["mean", "max"], [f"l{i}" for i in range(26)]),
                                [synth.init_state_dict(arch, 2, 26, 1003)], (0, 1))
        xd = x.permute(0, 2, 3, 1).contiguous()
        st = L.stream_ptr()
        h = p._handles[0]
        out["engine_fp16_forward_ms"] = timeit(lambda: L.check(p._lib.ts2d_unet_forward(h, xd.data_ptr(), a.batch, st)), a.reps)
        out["engine_flops"] = p.flops(a.batch)
        p.close()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
