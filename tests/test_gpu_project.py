"""GPU parity: K1 projection vs the CPU oracle (bit-exact max; mean bit-exact for integer input, 1e-5 rel for float)."""
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import geometry
from tests.conftest import GOLDEN, synthetic_ct

pytestmark = pytest.mark.gpu


def _gpu_project(vol, modes, reduce_dim, h_dim, flip_h=False, flip_w=False):
    from totalsegmentator2d_b200 import ops
    t = torch.from_numpy(vol.view(np.int16) if vol.dtype == np.uint16 else vol).cuda()
    if vol.dtype == np.uint16:
        t = t.view(torch.uint16)
    return ops.project(t, modes, reduce_dim, h_dim, flip_h, flip_w).cpu().numpy()


def _oracle(vol, reduce_dim, h_dim, flip_h, flip_w):
    mx = vol.max(axis=reduce_dim).astype(np.float32)
    mn = vol.min(axis=reduce_dim).astype(np.float32)
    mean = geometry.project_mean(vol, reduce_dim).astype(np.float32)
    kept = [d for d in range(3) if d != reduce_dim]
    out = []
    for a in (mean, mx, mn):
        if h_dim == kept[1]:
            a = a.T
        if flip_h:
            a = a[::-1]
        if flip_w:
            a = a[:, ::-1]
        out.append(np.ascontiguousarray(a))
    return np.stack(out)


@pytest.mark.parametrize("shape", [(37, 64, 128), (16, 33, 512), (5, 7, 24), (40, 300, 96), (3, 9, 53)])
@pytest.mark.parametrize("reduce_dim", [0, 1, 2])
def test_project_int16_all_axes(shape, reduce_dim):
    vol = synthetic_ct(shape, seed=sum(shape) + reduce_dim)
    kept = [d for d in range(3) if d != reduce_dim]
    for h_dim in kept:
        for fh, fw in ((False, False), (True, False), (True, True)):
            got = _gpu_project(vol, ["mean", "max", "min"], reduce_dim, h_dim, fh, fw)
            ref = _oracle(vol, reduce_dim, h_dim, fh, fw)
            assert got.shape == ref.shape
            assert np.array_equal(got[1], ref[1]), "max not bit-exact"
            assert np.array_equal(got[2], ref[2]), "min not bit-exact"
            assert np.array_equal(got[0], ref[0]), "integer mean not bit-exact"


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.int32, np.float32, np.float64])
def test_project_other_dtypes(dtype):
    rng = np.random.default_rng(11)
    if np.issubdtype(dtype, np.integer):
        info = np.iinfo(dtype)
        vol = rng.integers(max(info.min, -100000), min(info.max, 100000), size=(9, 40, 72)).astype(dtype)
    else:
        vol = (rng.normal(size=(9, 40, 72)) * 500).astype(dtype)
    for reduce_dim in (0, 1, 2):
        kept = [d for d in range(3) if d != reduce_dim]
        got = _gpu_project(vol, ["mean", "max", "min"], reduce_dim, kept[0])
        ref = _oracle(vol, reduce_dim, kept[0], False, False)
        assert np.array_equal(got[1], ref[1]) and np.array_equal(got[2], ref[2])
        if np.issubdtype(dtype, np.integer):
            assert np.array_equal(got[0], ref[0])
        else:
            assert np.allclose(got[0], ref[0], rtol=1e-5, atol=1e-5 * np.abs(ref[0]).max())


def test_project_extreme_values_and_single_slice():
    vol = np.full((4, 1, 16), -32768, dtype=np.int16)
    vol[1] = 32767
    got = _gpu_project(vol, ["mean", "max"], 1, 0)
    assert np.array_equal(got[0], vol[:, 0].astype(np.float32)) and np.array_equal(got[1], vol[:, 0].astype(np.float32))
    big = np.full((2, 4096, 64), 32767, dtype=np.int16)            # sum = 134 M per column: exceeds fp32 exactness, not int64
    got = _gpu_project(big, ["mean", "max"], 1, 0)
    assert np.all(got[0] == 32767.0) and np.all(got[1] == 32767.0)


def test_project_golden_asset_s0521():
    from totalsegmentator2d_b200 import nrrd
    from totalsegmentator2d_b200.core.util.image import project_channels
    img = nrrd.read_nrrd(os.path.join(GOLDEN, "sample_s0521.nrrd"))
    res = project_channels(img, ["mean", "max"])
    with open(os.path.join(GOLDEN, "s0521_projection.json")) as f:
        gold = json.load(f)
    mean = np.asarray(res["mean"].array).reshape(133, 53)
    mx = np.asarray(res["max"].array).reshape(133, 53)
    assert hashlib.sha256(np.ascontiguousarray(mean).tobytes()).hexdigest() == gold["mean_sha256"]
    assert hashlib.sha256(np.ascontiguousarray(mx).tobytes()).hexdigest() == gold["max_sha256"]
    assert res["max"].GetSize() == (53, 1, 133) and res["max"].GetSpacing() == (1.5, 180.0, 1.5)
    assert np.allclose(res["max"].direction_matrix(), -np.eye(3))


def test_project_permuted_orientation_matches_reorient_then_project():
    """DICOMOrient folded into the kernel == materialised reorientation followed by a coronal projection"""
    from totalsegmentator2d_b200.core.util.image import project_channels
    from totalsegmentator2d_b200.image import Image
    vol = synthetic_ct((24, 40, 56), seed=9)
    for d in (np.diag([1.0, 1.0, 1.0]), np.array([[0, 0, 1.0], [1.0, 0, 0], [0, -1.0, 0]]), np.array([[0, -1.0, 0], [0, 0, -1.0], [1.0, 0, 0]]),
              np.array([[0.99, 0.1, 0], [-0.1, 0.99, 0.05], [0, -0.05, -1.0]])):
        img = Image(vol, (1.0, 2.0, 3.0), (5, 6, 7), d.reshape(-1))
        perm, flip = geometry.orientation_remap(d, "RAI")
        rai = geometry.reorient_array(vol, perm, flip)
        ref = geometry.project_channels(rai, ["mean", "max"])
        res = project_channels(img, ["mean", "max"])
        H, W = ref.shape[1:]
        assert np.array_equal(np.asarray(res["mean"].array).reshape(H, W), ref[0])
        assert np.array_equal(np.asarray(res["max"].array).reshape(H, W), ref[1])


def test_project_full_size_properties():
    """BASELINE size 512x512x1000 int16: size-independent properties instead of an element-wise oracle run."""
    from totalsegmentator2d_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(1)
    vol = torch.randint(-1024, 3072, (1000, 512, 512), dtype=torch.int16, device="cuda", generator=g)
    out = ops.project(vol, ["mean", "max", "min"], 1, 0)
    assert torch.equal(out[1], vol.amax(dim=1).float()) and torch.equal(out[2], vol.amin(dim=1).float())
    ref_mean = (vol.sum(dim=1, dtype=torch.int64).double() / 512.0).float()
    assert torch.equal(out[0], ref_mean)
    # linearity of the mean: mean(v + c) == mean(v) + c exactly for an integer shift inside the int16 range
    out2 = ops.project(vol + 7, ["mean"], 1, 0)
    assert torch.equal(out2[0], ((vol.sum(dim=1, dtype=torch.int64) + 7 * 512).double() / 512.0).float())
    # idempotence: projecting the (X,1,Z) result again changes nothing
    again = ops.project(out[1].to(torch.int16)[:, None, :].contiguous(), ["max"], 1, 0)
    assert torch.equal(again[0], out[1])
