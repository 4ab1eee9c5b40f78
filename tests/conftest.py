import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100) GPU; run with -m gpu on the GPU box")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def synthetic_ct(shape_zyx, seed, dtype=np.int16):
    """clip(round(N(-300,450)), -1024, 3071) with a centred ellipsoid 'body' of N(300,200)."""
    rng = np.random.default_rng(seed)
    nz, ny, nx = shape_zyx
    vol = np.clip(np.round(rng.normal(-300, 450, size=shape_zyx)), -1024, 3071)
    zz, yy, xx = np.ogrid[:nz, :ny, :nx]
    body = ((zz - nz / 2) / (0.4 * nz)) ** 2 + ((yy - ny / 2) / (0.35 * ny)) ** 2 + ((xx - nx / 2) / (0.4 * nx)) ** 2 < 1
    vol[body] = np.clip(np.round(rng.normal(300, 200, size=int(body.sum()))), -1024, 3071)
    return vol.astype(dtype)


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN
