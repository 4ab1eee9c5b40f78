#!/usr/bin/env python
"""Per-stage device time of one study at bench shapes (CUDA events on the launching stream, single stream, warm):
projection, preprocess + tiling, and per group model the U-Net forward and the blend/head kernel; then whole-study
time for several TS2D_MODEL_STREAMS settings.  Output goes to stdout (tee it under gpurun_out/).
"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench


def ev():
    e = torch.cuda.Event(enable_timing=True)
    e.record()
    return e


def main():
    from totalsegmentator2d_b200 import _lib as L
    from totalsegmentator2d_b200 import ops
    from totalsegmentator2d_b200.core.inference.preprocess import B200Preprocessor
    from totalsegmentator2d_b200.tool import TS2D
    tmp = tempfile.TemporaryDirectory()
    bench.write_zoo(tmp.name, False, True)
    os.environ["TS2D_MODEL_STREAMS"] = "1"
    ts = TS2D(key="ts2d", models_root=tmp.name, param={"b200.precision": os.environ.get("TS2D_PRECISION", "fp16")})
    vol = torch.from_numpy(bench.synthetic_volume(20260000)).cuda()
    for _ in range(3):
        ts.predict_volume_device(vol, bench.SPACING)
    torch.cuda.synchronize()
    # ---- stage by stage (same calls predict_volume_device makes) ----
    ids = list(ts.models.keys())
    reps = 5
    acc = {}
    for _ in range(reps):
        marks = [("start", ev())]
        out = ops.project(vol, ["mean", "max"], reduce_dim=1, h_dim=0, flip_h=False, flip_w=False)
        marks.append(("project", ev()))
        p0 = ts.models[ids[0]].predictor
        data, props = B200Preprocessor().run_case_device(out, (1.5, 1.5), p0.plans_manager, p0.configuration_manager, None)
        marks.append(("preprocess", ev()))
        x, layout = p0.make_tiles(data)
        marks.append(("extract_tiles", ev()))
        H, W = int(data.shape[1]), int(data.shape[2])
        offs = ts.head_offsets()
        mask = torch.zeros((H, W, offs[-1]), dtype=torch.uint8, device="cuda")
        present = torch.zeros(offs[-1], dtype=torch.uint8, device="cuda")
        marks.append(("mask_zero", ev()))
        st = L.stream_ptr()
        pads, padded, tiles = layout
        flat = L.int_array([v for yx in tiles for v in yx])
        for k, id in enumerate(ids):
            p = ts.models[id].predictor
            T, M = len(tiles), ops.mirror_count(p.mirror_mask)
            L.check(p._lib.ts2d_unet_forward(p._handles[0], x.data_ptr(), T * M, st))
            marks.append(("forward", ev()))
            L.check(p._lib.ts2d_unet_blend_head(p._handles[0], flat, T, M, p.mirror_mask, pads[0], pads[1], H, W, p._gaussian.data_ptr(), None,
                                                mask.data_ptr(), H, W, 0, 0, offs[k], offs[-1], present.data_ptr(), st))
            marks.append(("blend_head", ev()))
        merged, perm = ops.merge_present_order(mask, present, offs)
        marks.append(("merge_present_order", ev()))
        torch.cuda.synchronize()
        for (_, a), (name, b) in zip(marks[:-1], marks[1:]):
            acc.setdefault(name, []).append(a.elapsed_time(b))
    total = 0.0
    print("# per-study device time by stage (ms, mean of %d studies; forward / blend_head are sums over the 5 group models)" % reps)
    for name, v in acc.items():
        per_study = sum(v) / reps
        total += per_study
        print("%-22s %8.3f ms   (%d launches-groups per study, %.3f ms each)" % (name, per_study, len(v) // reps, per_study / (len(v) // reps)))
    print("%-22s %8.3f ms  -> %.1f studies/s single stream" % ("sum", total, 1000.0 / total))
    # ---- whole study, model streams ----
    for ns in (1, 2, 3, 5):
        os.environ["TS2D_MODEL_STREAMS"] = str(ns)
        for _ in range(2):
            ts.predict_volume_device(vol, bench.SPACING)
        torch.cuda.synchronize()
        a = ev()
        n = 6
        for _ in range(n):
            ts.predict_volume_device(vol, bench.SPACING)
        b = ev()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / n
        print("TS2D_MODEL_STREAMS=%d: %.3f ms per study -> %.1f studies/s" % (ns, ms, 1000.0 / ms))
    ts.close()


if __name__ == "__main__":
    main()
