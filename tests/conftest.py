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


@pytest.fixture(scope="session")
def cs512_trained_zoo(tmp_path_factory):
    """the same five CS-512 group models with TRAINED-LIKE weights (tools/make_trained_like_zoo.py: a few hundred Adam steps on
    synthetic projections with organ-like targets, ~10 s per model on the GPU box), written once per session; the north_star gates
    (99.9 % masks, 1e-2 logits) are asserted on this zoo"""
    import bench
    from tools import make_trained_like_zoo as mk
    from totalsegmentator2d_b200 import synth
    root = str(tmp_path_factory.mktemp("cs512_trained_zoo"))
    mk.write_trained_like_zoo(root, bench.MODEL_KEY, synth.TS2D_V2_GROUPS, synth.cs512_arch(), steps=int(os.environ.get("TS2D_TRAIN_STEPS", "400")))
    return root, synth.TS2D_V2_GROUPS
