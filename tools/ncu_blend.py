#!/usr/bin/env python
"""ncu target for the blend / head / threshold kernel (K8/K9): one ribs-shaped CS-512 model on a preprocessed 1000 x 512 image
(3 tiles x 4 mirror passes), mask output into a 117-channel buffer exactly as TS2D.predict does it.

    python tools/ncu_blend.py && ncu --set full --clock-control none --import-source on -k regex:blend_head -s 1 -c 1 -o gpurun_out/blend python tools/ncu_blend.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch


def main():
    from totalsegmentator2d_b200 import synth
    from totalsegmentator2d_b200.core.inference.predictor import B200nnUNetPredictor
    heads = 26
    arch = synth.cs512_arch()
    plans = synth.make_plans(arch, (512, 512), (1.5, 1.5), ["ZScoreNormalization"] * 2)
    dj = synth.make_dataset_json(["mean", "max"], [f"l{i}" for i in range(heads)])
    sd = synth.init_state_dict(arch, 2, heads, 1003)
    p = B200nnUNetPredictor(precision="fp16", max_batch=12, use_mirroring=True)
    p.manual_initialization(plans, "2d", dj, [sd], (0, 1))
    g = torch.Generator().manual_seed(7)
    data = torch.randn(2, 1000, 512, generator=g).cuda()
    tiles, layout = p.make_tiles(data)
    mask = torch.empty((1000, 512, 117), dtype=torch.uint8, device="cuda")
    present = torch.zeros(117, dtype=torch.uint8, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for it in range(3):
        p.run_tiles(tiles, layout, (1000, 512), mask=mask, crop_origin=(0, 0), ch_off=65, present=present)
    torch.cuda.synchronize()
    print("ok: mask sum", int(mask[..., 65:91].sum()))


if __name__ == "__main__":
    main()
