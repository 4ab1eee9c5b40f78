#!/usr/bin/env python
"""cProfile of the public API path (TS2D.predict on a pinned host volume + Result.get_segmentation) at bench shapes:
where do the milliseconds between the device-resident number and the e2e number go?"""
import cProfile
import os
import pstats
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import bench


def main():
    from totalsegmentator2d_b200.image import Image
    from totalsegmentator2d_b200.tool import TS2D
    tmp = tempfile.TemporaryDirectory()
    bench.write_zoo(tmp.name, False, True)
    ts = TS2D(key="ts2d", models_root=tmp.name, param={"b200.precision": os.environ.get("TS2D_PRECISION", "fp16")})
    vol = torch.from_numpy(bench.synthetic_volume(20260000)).pin_memory()
    img = Image(vol.numpy(), bench.SPACING, (0.0, 0.0, 0.0), np.eye(3).reshape(-1))
    for _ in range(3):
        ts.predict(img).get_segmentation()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        ts.predict(img).get_segmentation()
    torch.cuda.synchronize()
    print("e2e ms per study:", (time.perf_counter() - t0) / 3 * 1e3)
    pr = cProfile.Profile()
    pr.enable()
    for _ in range(3):
        ts.predict(img).get_segmentation()
    torch.cuda.synchronize()
    pr.disable()
    pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
    ts.close()


if __name__ == "__main__":
    main()
