"""GPU parity: crop / normalise / resample / tiling vs the CPU oracle."""
import numpy as np
import pytest
import torch

from oracle import nnunet_inf, nnunet_pre

pytestmark = pytest.mark.gpu


def _img(seed, shape=(2, 70, 55), zero_border=True):
    rng = np.random.default_rng(seed)
    a = (rng.normal(size=shape) * 300 + 50).astype(np.float32)
    if zero_border:
        a[:, :6] = 0
        a[:, -3:] = 0
        a[:, :, :4] = 0
        a[:, :, -9:] = 0
        a[0, 20, 30] = 0
    return a


def test_bbox_and_stats():
    from totalsegmentator2d_b200 import ops
    a = _img(1)
    d = torch.from_numpy(a).cuda()
    bbox = ops.bbox_nonzero(d).tolist()
    _, ref = nnunet_pre.crop_to_nonzero(a[:, None])
    assert bbox == [ref[1][0], ref[1][1], ref[2][0], ref[2][1]]
    assert ops.bbox_nonzero(torch.zeros((2, 9, 11), device="cuda")).tolist() == [0, 9, 0, 11]     # empty mask keeps the frame
    one = torch.zeros((1, 9, 11), device="cuda")
    one[0, 4, 7] = -1
    assert ops.bbox_nonzero(one).tolist() == [4, 5, 7, 8]
    st = ops.channel_stats(d, bbox).cpu().numpy()
    crop = a[:, bbox[0]:bbox[1], bbox[2]:bbox[3]].astype(np.float64)
    assert np.allclose(st[:, 0], crop.sum(axis=(1, 2)), rtol=1e-12)
    assert np.allclose(st[:, 1], (crop ** 2).sum(axis=(1, 2)), rtol=1e-12)
    assert np.array_equal(st[:, 2], crop.min(axis=(1, 2))) and np.array_equal(st[:, 3], crop.max(axis=(1, 2)))


@pytest.mark.parametrize("schemes", [["ZScoreNormalization", "ZScoreNormalization"], ["CTNormalization", "NoNormalization"],
                                     ["RescaleTo01Normalization", "RGBTo01Normalization"]])
@pytest.mark.parametrize("spacing_xy,plan", [((1.5, 1.5), (1.5, 1.5)), ((0.75, 0.75), (1.5, 1.5)), ((1.25, 1.25), (1.0, 1.0)), ((1.5, 1.0), (1.2, 1.3))])
def test_run_case_matches_oracle(schemes, spacing_xy, plan):
    from totalsegmentator2d_b200.core.inference.plans import PlansManager
    from totalsegmentator2d_b200.core.inference.preprocess import B200Preprocessor
    from totalsegmentator2d_b200 import synth
    a = _img(7)
    props = {"0": {"percentile_00_5": -400.0, "percentile_99_5": 600.0, "mean": 20.0, "std": 250.0, "max": 0, "min": 0, "median": 0},
             "1": {"percentile_00_5": -300.0, "percentile_99_5": 500.0, "mean": 10.0, "std": 200.0, "max": 0, "min": 0, "median": 0}}
    plans = synth.make_plans(synth.make_arch_kwargs([8, 16]), (32, 32), plan, schemes, props)
    pm = PlansManager(plans)
    cm = pm.get_configuration("2d")
    data, p = B200Preprocessor().run_case_device(torch.from_numpy(a).cuda(), spacing_xy, pm, cm)
    ref, rp = nnunet_pre.run_case_npy(a[:, None], nnunet_pre.sitkio_spacing_2d(spacing_xy), list(plan), schemes, props)
    assert p["bbox_used_for_cropping"] == rp["bbox_used_for_cropping"]
    assert tuple(p["shape_after_cropping_and_before_resampling"]) == tuple(rp["shape_after_cropping_and_before_resampling"])
    got = data.cpu().numpy()
    assert got.shape == ref[:, 0].shape
    scale = max(1.0, float(np.abs(ref).max()))
    assert np.abs(got - ref[:, 0]).max() <= 2e-5 * scale, np.abs(got - ref[:, 0]).max()


@pytest.mark.parametrize("order", [1, 3])
@pytest.mark.parametrize("new", [(35, 27), (140, 111), (70, 20), (9, 200)])
def test_resample_matches_skimage_restatement(order, new):
    from totalsegmentator2d_b200 import ops
    a = _img(3, zero_border=False)
    got = ops.resample2d(torch.from_numpy(a).cuda(), new, order).cpu().numpy()
    ref = np.stack([nnunet_pre.skimage_resize_2d(a[c], new, order) for c in range(a.shape[0])]).astype(np.float32)
    assert np.abs(got - ref).max() <= 1e-4 * np.abs(ref).max()


@pytest.mark.parametrize("hw,patch", [((70, 55), (32, 32)), ((20, 90), (32, 48)), ((32, 32), (32, 32))])
@pytest.mark.parametrize("mirror", [0, 1, 2, 3])
def test_extract_tiles_matches_pad_slice_flip(hw, patch, mirror):
    from totalsegmentator2d_b200 import _lib as L, ops
    from totalsegmentator2d_b200.core.inference.predictor import sliding_window_layout
    rng = np.random.default_rng(5)
    a = rng.normal(size=(2, *hw)).astype(np.float32)
    pads, padded, tiles = sliding_window_layout(hw, patch, 0.5)
    x = ops.extract_tiles(torch.from_numpy(a).cuda(), pads, patch, tiles, mirror, L.PREC_F32).cpu()
    pt, _, rpads = nnunet_inf.pad_to_patch(torch.from_numpy(a)[:, None], patch)
    assert tuple(pt.shape[2:]) == tuple(padded) and (rpads[0][0], rpads[1][0]) == tuple(pads)
    steps = nnunet_inf.compute_steps_for_sliding_window(padded, patch, 0.5)
    assert tiles == [(y, xx) for y in steps[0] for xx in steps[1]]
    axes = [m for m in (0, 1) if mirror & (1 << m)]
    combos = [()] + [tuple(c) for c in nnunet_inf.mirror_axes_combinations(axes)]
    M = len(combos)
    for t, (ty, tx) in enumerate(tiles):
        tile = pt[:, 0, ty:ty + patch[0], tx:tx + patch[1]][None]
        for m, ax in enumerate(combos):
            ref = torch.flip(tile, ax) if ax else tile
            assert torch.equal(x[t * M + m].permute(2, 0, 1), ref[0])


def test_visual_kernels_match_create_visual_restatement():
    """csrc/visual.cu (label flatten + nearest-neighbour uniform resampling + LabelToRGB; cubic B-spline uniform resampling +
    minmax window + IntensityWindowing + uint8 truncation) vs oracle/visual.py, the CPU restatement of reference
    ts2d/core/util/image.py:383-481; through the host mirror create_visual on isotropic and anisotropic images."""
    from oracle import visual as ovis
    from totalsegmentator2d_b200 import ops
    from totalsegmentator2d_b200.core.util import visual as vis
    from totalsegmentator2d_b200.core.util.meta import set_annotation_meta
    from totalsegmentator2d_b200.image import Image
    rng = np.random.default_rng(11)
    # ---- labels: 13 overlapping channels, anisotropic spacing (1.5, 0.75) -> uniform 0.75 ----
    m = (rng.random((90, 70, 13)) > 0.8).astype(np.uint8)
    m[:10] = 0
    size_xy, sp_xy = (70, 90), (1.5, 0.75)
    new_size, ratios, changed = vis.uniform_geometry(size_xy, sp_xy)
    assert changed and new_size == [140, 90]
    cmap = vis.to_palette({i + 1: (int(rng.integers(0, 255)), int(rng.integers(0, 255)), int(rng.integers(0, 255))) for i in range(13) if i != 4})
    got = ops.visual_labels(torch.from_numpy(m).cuda(), (new_size[1], new_size[0]), (ratios[1], ratios[0]), np.asarray(cmap, np.uint8)).cpu().numpy()
    want = ovis.visual_labels(m, (new_size[1], new_size[0]), (ratios[1], ratios[0]), cmap)
    assert got.shape == (90, 140, 3) and np.array_equal(got, want)
    seg = Image(m, sp_xy, (0.0, 0.0), None, is_vector=True)
    set_annotation_meta(seg, names={i + 1: f"n{i}" for i in range(13)}, colors={f"n{i}": tuple(cmap[i + 1]) for i in range(13)})
    # the palette comes back from the Segment{N}_Color metadata: "r g b" floats with 3 decimals, x 255 and TRUNCATED by to_color
    # (reference color.py:72-85, meta.py:213-230): 37 -> "0.145" -> 36.  The expectation takes the same round trip on the host.
    rt = {i + 1: tuple(int(float(np.format_float_positional(c / 255.0, precision=3, unique=False)) * 255) for c in cmap[i + 1]) for i in range(13)}
    want_rt = ovis.visual_labels(m, (new_size[1], new_size[0]), (ratios[1], ratios[0]), vis.to_palette(rt))
    assert np.array_equal(vis.create_visual(seg, labels=True, axis="coronal"), want_rt)
    # ---- intensity: anisotropic (B-spline resampling) and isotropic (window only) ----
    a = rng.normal(100, 300, size=(60, 50)).astype(np.float32)
    new_size, ratios, changed = vis.uniform_geometry((50, 60), (1.0, 2.5))
    assert changed and new_size == [50, 150]
    got = ops.visual_intensity(torch.from_numpy(a).cuda(), (150, 50), (ratios[1], ratios[0])).cpu().numpy()
    want = ovis.visual_intensity(a, (150, 50), (ratios[1], ratios[0]))
    d = np.abs(got.astype(int) - want.astype(int))
    assert got.shape == (150, 50) and d.max() <= 1 and (d == 0).mean() >= 0.995, (d.max(), (d == 0).mean())     # truncation of values within 1e-9 of an integer
    got = vis.create_visual(Image(a, (1.5, 1.5), (0.0, 0.0), None))
    want = ovis.visual_intensity(a, a.shape, (1.0, 1.0), resample=False)
    assert np.array_equal(got, want) and got.min() == 0 and got.max() == 255
    got = ops.visual_intensity(torch.from_numpy(a).cuda(), a.shape, (1.0, 1.0), (-200.0, 400.0)).cpu().numpy()
    assert np.array_equal(got, ovis.visual_intensity(a, a.shape, (1.0, 1.0), (-200.0, 400.0), resample=False))
    # ---- a 3-D projected image container (X, 1, Z) as Result.save passes it ----
    proj = Image(a.reshape(60, 1, 50), (1.5, 403.5, 1.5), (0, 0, 0), np.diag([-1.0, -1.0, -1.0]).reshape(-1))
    assert np.array_equal(vis.create_visual(proj), want if False else ovis.visual_intensity(a, a.shape, (1.0, 1.0), resample=False))
