#!/usr/bin/env python
"""BASELINE.md section 6 rows that bench.py's headline line does not cover (VERDICT r01 "measurement rows left blank"), on ONE GPU,
device-resident (volumes / images already in HBM), CUDA events around the whole loop, inputs larger than L2 or rotated over 4 buffers:

  A-noTTA   synthetic 512 x 512 x 1000 int16 study, 1.5 mm, five CS-512 group models, mirror TTA off
  A-TTA     same, mirror TTA on (bench.py's workload, repeated here so that the rows share one process and one clock state)
  B-TTA     variant B: the same voxels declared 0.75 mm -> projected image 1000 x 512 resampled to the 1.5 mm plan (500 x 256, cubic
            B-spline, K3), one tile per model, masks resampled back by the export (order-1 logits)
  model     one ribs-shaped group model (26 heads) on a preprocessed (2, 1, 1000, 512) image: 3 tiles x 4 mirror passes (config 2)
  xray      tsxr-style single-channel model on the 320 x 320 chexpert asset, plan spacing 1.25 (no resampling) and 1.0 (400 x 400)

    python tools/config_table.py [--out profiles/r02_config_table.txt]
"""
import argparse
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def timed(fn, reps, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def latency(fn, reps=7):
    out = []
    for _ in range(reps):
        torch.cuda.synchronize()
        t = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        out.append(1000.0 * (time.perf_counter() - t))
    return float(np.median(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import bench
    from totalsegmentator2d_b200 import ops, synth
    from totalsegmentator2d_b200.core.inference.predictor import B200nnUNetPredictor
    from totalsegmentator2d_b200 import nrrd
    from totalsegmentator2d_b200.tool import TS2D
    lines = ["# BASELINE.md section 6 rows, one B200, device-resident inputs, fp16, CUDA events over the whole loop (ms per study / image), p50 = median",
             "# wall time of a single study with a device sync on both sides; per-study throughput = 1000 / ms",
             "%-10s %-58s %10s %10s %10s" % ("row", "workload", "ms", "per s", "p50 ms")]
    S = 4
    vols = [torch.from_numpy(bench.synthetic_volume(20260000 + i)).cuda() for i in range(S)]
    H, W = bench.VOL_SHAPE[0], bench.VOL_SHAPE[2]
    for row, mirror, spacing in (("A-noTTA", False, (1.5, 1.5, 1.5)), ("A-TTA", True, (1.5, 1.5, 1.5)), ("B-noTTA", False, (0.75, 0.75, 0.75)),
                                 ("B-TTA", True, (0.75, 0.75, 0.75))):
        with tempfile.TemporaryDirectory() as tmp:
            bench.write_zoo(tmp, False, mirror)
            ts = TS2D(key="ts2d", models_root=tmp, param={"b200.precision": "fp16"})
            n_heads = ts.head_offsets()[-1]
            offs = ts.head_offsets()
            masks = [torch.zeros((H, W, n_heads), dtype=torch.uint8, device="cuda") for _ in range(S)]
            merged = [torch.empty_like(m) for m in masks]
            pres = [torch.zeros(n_heads, dtype=torch.uint8, device="cuda") for _ in range(S)]
            k = [0]

            def one():
                i = k[0] % S
                k[0] += 1
                ts.predict_volume_device(vols[i], spacing, None, mask=masks[i], present=pres[i])
                ops.merge_present_order(masks[i], pres[i], offs, out=merged[i])
            ms = timed(one, 12)
            p50 = latency(one)
            what = "512x512x1000 int16 @ %.2f mm, 5 models, mirror %s%s" % (spacing[0], "on" if mirror else "off",
                                                                            "" if spacing[0] == 1.5 else " (resample 1000x512 -> 500x256)")
            lines.append("%-10s %-58s %10.3f %10.1f %10.2f" % (row, what, ms, 1000.0 / ms, p50))
            ts.close()
            del ts, masks, merged
            torch.cuda.empty_cache()
    del vols
    torch.cuda.empty_cache()
    # ---- config 2: one group model on a preprocessed image ----
    arch = synth.cs512_arch()
    for mirror in (False, True):
        plans = synth.make_plans(arch, (512, 512), (1.5, 1.5), ["ZScoreNormalization"] * 2)
        dj = synth.make_dataset_json(["mean", "max"], [f"l{i}" for i in range(26)])
        sd = synth.init_state_dict(arch, 2, 26, 1003)
        p = B200nnUNetPredictor(precision="fp16", max_batch=12, use_mirroring=mirror)
        p.manual_initialization(plans, "2d", dj, [sd], (0, 1))
        g = torch.Generator().manual_seed(7)
        datas = [torch.randn(2, 1, 1000, 512, generator=g).cuda() for _ in range(4)]
        k = [0]

        def one():
            k[0] += 1
            return p.predict_logits_from_preprocessed_data(datas[k[0] % 4])
        ms = timed(one, 12)
        lines.append("%-10s %-58s %10.3f %10.1f %10.2f" % ("model", "ribs-shaped CS-512 model, (2,1,1000,512), 3 tiles, mirror %s, fp32 logits out" % ("on" if mirror else "off"),
                                                            ms, 1000.0 / ms, latency(one)))
        p.close()
    # ---- config 3: X-ray, one channel ----
    img = nrrd.read_nrrd(os.path.join(ROOT, "tests", "golden", "sample_chexpert.nrrd"))
    for plan_sp in (1.25, 1.0):
        with tempfile.TemporaryDirectory() as tmp:
            synth.write_model_dir(tmp, "tsxr-v2-ep1000b2", "ribs", synth.TS2D_V2_GROUPS["ribs"], arch, (512, 512), (plan_sp, plan_sp), seed=1200,
                                  mirror_axes=(0, 1), channel_names=("xray",))
            ts = TS2D(key="tsxr-v2-ep1000b2_ribs", models_root=tmp, param={"b200.precision": "fp16"})
            x = torch.from_numpy(np.ascontiguousarray(np.asarray(img.array, dtype=np.float32).reshape(1, 320, 320))).cuda()

            def one():
                return ts.predict_image_device({m: x for m in ts.models}, img.GetSpacing()[:2])
            ms = timed(one, 12)
            lines.append("%-10s %-58s %10.3f %10.1f %10.2f" % ("xray", "chexpert 320x320 uint8 @1.25 mm, 1 model (26 heads), plan %.2f mm%s" % (
                plan_sp, "" if plan_sp == 1.25 else " (-> 400x400)"), ms, 1000.0 / ms, latency(one)))
            ts.close()
    text = "\n".join(lines)
    print(text)
    if a.out:
        open(a.out, "w").write(text + "\n")


if __name__ == "__main__":
    main()
