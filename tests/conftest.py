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


@pytest.fixture(scope="session")
def cs512_zoo(tmp_path_factory):
    """the BASELINE zoo (five CS-512 group models, heads 18/23/24/26/26, seeds 1000+g as bench.py) written once per session"""
    import bench
    root = str(tmp_path_factory.mktemp("cs512_zoo"))
    groups, patch = bench.write_zoo(root, False, True)
    return root, groups


def oracle_models(root, model_key, groups):
    from oracle import pipeline as opipe
    return {f"{model_key}_{g}": opipe.OracleModel(opipe.find_training_dir(os.path.join(root, f"{model_key}_{g}", "r001"))) for g in groups}
