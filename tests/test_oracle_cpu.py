"""CPU suite: pins the oracle against the reference's asset files and against the scipy/torch primitives that
upstream nnU-Net itself calls.  (The reference's own tests hold no numeric vectors: parity is otherwise unpinned.)"""
import ctypes
import hashlib
import json
import os

import numpy as np
import pytest
import torch

from oracle import geometry, nnunet_inf, nnunet_net, nnunet_pre, pipeline
from totalsegmentator2d_b200 import nrrd, synth
from tests.conftest import GOLDEN, ROOT, synthetic_ct


def _read(name):
    return nrrd.read_nrrd(os.path.join(GOLDEN, name))


def test_asset_s0616_is_reference_projection_output():
    """sample_s0616 is a real output of the reference's projection stage: ch1 integral (max), ch0 = k/261 in
    double (mean = double sum / N), max >= mean -> channel order {0: mean, 1: max} and double accumulation."""
    img = _read("sample_s0616.nrrd")
    a = np.asarray(img.array)
    assert a.dtype == np.float64 and a.shape == (644, 337, 2)
    mean, mx = a[..., 0], a[..., 1]
    assert np.all(mx == np.round(mx))
    assert np.all(mx >= mean)
    k = mean * 261.0
    assert np.max(np.abs(k - np.round(k))) < 1e-9          # exact multiples of 1/261 in double
    assert np.abs(mean - np.round(mean)).max() > 0.1        # and genuinely fractional


def test_asset_s0332_geometry_of_project():
    """sample_s0332 carries project()'s output geometry: size (X,1,Z), spacing_y = Ny*sy (403.5 = 269*1.5), RAI."""
    img = _read("sample_s0332.nrrd")
    assert img.GetSize() == (269, 1, 505) and img.GetNumberOfComponentsPerPixel() == 2
    assert img.GetSpacing() == (1.5, 403.5, 1.5)
    assert np.allclose(img.direction_matrix(), -np.eye(3))
    a = np.asarray(img.array)
    assert np.all(a[..., 1] == np.round(a[..., 1])) and np.all(a[..., 1] >= a[..., 0])
    size, spacing, origin, d = geometry.projected_geometry((269, 269, 505), (1.5, 1.5, 1.5), (0, 0, 0), -np.eye(3), 1)
    assert size == [269, 1, 505] and spacing == [1.5, 403.5, 1.5]


def test_orientation_remap_s0521():
    img = _read("sample_s0521.nrrd")
    perm, flip = geometry.orientation_remap(img.direction_matrix(), "RAI")
    assert perm == (0, 1, 2) and flip == (False, False, True)          # RAS-like asset: only z flips
    # permuted / oblique
    d = np.array([[0.0, 0.0, 1.0], [0.98, -0.2, 0.0], [0.2, 0.98, 0.0]])      # axis0->P, axis1->S(ish), axis2->L
    perm, flip = geometry.orientation_remap(d, "RAI")
    assert perm == (2, 0, 1) and flip == (True, True, True)
    vol = np.arange(2 * 3 * 4).reshape(2, 3, 4).astype(np.int16)
    out = geometry.reorient_array(vol, perm, flip)
    assert out.shape == (3, 4, 2)[::-1][::-1] or out.size == vol.size


def test_projection_known_answer_s0521():
    """freeze the oracle projection of the only 3-D asset (known-answer test)."""
    img = _read("sample_s0521.nrrd")
    perm, flip = geometry.orientation_remap(img.direction_matrix(), "RAI")
    rai = geometry.reorient_array(np.asarray(img.array), perm, flip)
    proj = geometry.project_channels(rai, ["mean", "max"])
    assert proj.shape == (2, 133, 53) and proj.dtype == np.float32
    with open(os.path.join(GOLDEN, "s0521_projection.json")) as f:
        gold = json.load(f)
    assert hashlib.sha256(proj[0].tobytes()).hexdigest() == gold["mean_sha256"]
    assert hashlib.sha256(proj[1].tobytes()).hexdigest() == gold["max_sha256"]
    # the C restatement agrees bit for bit
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "liboracle_projection.so"))
    out = np.empty((133, 53), dtype=np.float32)
    for fn, ref in (("oracle_project_mean_i16", proj[0]), ("oracle_project_max_i16", proj[1])):
        getattr(lib, fn)(rai.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(133), ctypes.c_int64(120), ctypes.c_int64(53),
                         out.ctypes.data_as(ctypes.c_void_p))
        assert np.array_equal(out, ref)


def test_crop_normalize_resample_against_primitives():
    rng = np.random.default_rng(3)
    data = rng.normal(size=(2, 1, 40, 30)).astype(np.float32)
    data[:, :, :5] = 0
    data[:, :, :, 27:] = 0
    out, props = nnunet_pre.run_case_npy(data, [1498.5, 1.5, 1.5], [1.5, 1.5], ["ZScoreNormalization", "CTNormalization"],
                                         {"1": {"percentile_00_5": -1.0, "percentile_99_5": 1.2, "mean": 0.1, "std": 0.8}})
    assert props["bbox_used_for_cropping"] == [[0, 1], [5, 40], [0, 27]]
    assert out.shape == (2, 1, 35, 27)
    assert abs(out[0].mean()) < 1e-5 and abs(out[0].std() - 1) < 1e-5
    c = np.clip(data[1, :, 5:, :27], -1.0, 1.2)
    assert np.allclose(out[1], (c - 0.1) / 0.8, atol=1e-6)
    # resampling active: 0.75 mm data on a 1.5 mm plan halves the grid
    out2, _ = nnunet_pre.run_case_npy(data, [749.25, 0.75, 0.75], [1.5, 1.5], ["NoNormalization"] * 2)
    assert out2.shape == (2, 1, 18, 14)


def test_sliding_window_layout_matches_survey():
    assert nnunet_inf.compute_steps_for_sliding_window((1000, 512), (512, 512), 0.5) == [[0, 244, 488], [0]]
    assert len(nnunet_inf.compute_steps_for_sliding_window((2000, 512), (512, 512), 0.5)[0]) == 7
    g = nnunet_inf.compute_gaussian((64, 64))
    assert g.dtype == torch.float16 and float(g.max()) == 10.0 and float(g.min()) > 0


def test_unet_oracle_equals_torch_modules():
    """the functional oracle forward == an nn.Module PlainConvUNet assembled from torch layers with the same weights"""
    arch = synth.make_arch_kwargs([4, 8, 16])
    sd = synth.init_state_dict(arch, 2, 3, seed=1)
    spec = nnunet_net.ArchSpec(arch, 2, 3)
    x = torch.randn(2, 2, 16, 24, generator=torch.Generator().manual_seed(0))
    y = nnunet_net.unet_forward(x, sd, spec)
    assert y.shape == (2, 3, 16, 24)
    import torch.nn as nn

    def block(cin, cout, stride, p):
        m = nn.Sequential(nn.Conv2d(cin, cout, 3, stride, 1), nn.InstanceNorm2d(cout, eps=1e-5, affine=True), nn.LeakyReLU(0.01))
        m[0].weight.data, m[0].bias.data = sd[p + ".conv.weight"], sd[p + ".conv.bias"]
        m[1].weight.data, m[1].bias.data = sd[p + ".norm.weight"], sd[p + ".norm.bias"]
        return m
    skips, cur, cin = [], x, 2
    for s, f in enumerate([4, 8, 16]):
        for i in range(2):
            cur = block(cin, f, (1 if s == 0 else 2) if i == 0 else 1, f"encoder.stages.{s}.0.convs.{i}")(cur)
            cin = f
        skips.append(cur)
    for d in range(2):
        up = nn.ConvTranspose2d(skips[-(d + 1)].shape[1] if d == 0 else cur.shape[1], skips[-(d + 2)].shape[1], 2, 2)
        up.weight.data, up.bias.data = sd[f"decoder.transpconvs.{d}.weight"], sd[f"decoder.transpconvs.{d}.bias"]
        cur = torch.cat((up(cur if d else skips[-1]), skips[-(d + 2)]), 1)
        f = skips[-(d + 2)].shape[1]
        for i in range(2):
            cur = block(cur.shape[1], f, 1, f"decoder.stages.{d}.convs.{i}")(cur)
    head = nn.Conv2d(4, 3, 1)
    head.weight.data, head.bias.data = sd["decoder.seg_layers.1.weight"], sd["decoder.seg_layers.1.bias"]
    with torch.no_grad():
        assert torch.allclose(head(cur), y, atol=1e-5)
    assert nnunet_net.conv_flops(nnunet_net.ArchSpec(synth.cs512_arch(), 2, 26), 512, 512) == pytest.approx(119.7e9, rel=2e-3)


def test_mirror_and_blend_properties():
    """a flip-equivariant 'network' must be unchanged by mirror TTA; blending a constant gives the constant"""
    net = lambda x: x[:, :1] * 2.0 + 1.0
    x = torch.randn(1, 1, 70, 50, generator=torch.Generator().manual_seed(2))
    lg = nnunet_inf.predict_sliding_window(net, x.permute(1, 0, 2, 3), (32, 32), 1, 0.5, (0, 1), True, "ideal")
    assert torch.allclose(lg[0, 0], x[0, 0] * 2 + 1, atol=1e-5)
    lg16 = nnunet_inf.predict_sliding_window(net, x.permute(1, 0, 2, 3), (32, 32), 1, 0.5, (0, 1), True, "faithful")
    # upstream keeps fp16 accumulators and an fp16 Gaussian: near tile corners (tiny weights) that costs ~2e-2
    assert torch.allclose(lg16, lg, atol=6e-2) and float((lg16 - lg).abs().mean()) < 3e-3


def test_combine_order_is_present_first():
    a = np.zeros((4, 4, 3), np.uint8)
    a[0, 0, 2] = 1
    b = np.zeros((4, 4, 2), np.uint8)
    b[1, 1, 0] = 1
    merged, names = pipeline.combine_segmentations([a, b], [["x", "y", "z"], ["p", "q"]])
    assert names == ["z", "x", "y", "p", "q"] and merged.shape == (4, 4, 5) and merged[0, 0, 0] == 1


def test_combine_order_absent_labels_follow_std_map_key_order():
    """reference meta.py:254-286 walks sitk GetMetaDataKeys() (a std::map): absent labels come out in the lexicographic order of
    "Segment{id}_" (0, 10, 11, 1, 2, ...), not ascending.  12 labels, channels 2 and 11 present -> 2, 11, then 0, 10, 1, 3, 4, ...
    Pinned for the oracle and for the host mirror of combine_segmentations (same Image metadata semantics)."""
    from totalsegmentator2d_b200.core.util.image import combine_segmentations
    from totalsegmentator2d_b200.core.util.meta import get_annotation_labels, set_annotation_meta
    from totalsegmentator2d_b200.image import Image
    a = np.zeros((4, 4, 12), np.uint8)
    a[0, 0, 2] = 1
    a[1, 1, 11] = 1
    nm = [f"n{i}" for i in range(12)]
    want = ["n2", "n11", "n0", "n10", "n1", "n3", "n4", "n5", "n6", "n7", "n8", "n9"]
    merged, names = pipeline.combine_segmentations([a], [nm])
    assert names == want and merged[0, 0, 0] == 1 and merged[1, 1, 1] == 1
    seg = Image(a, is_vector=True)
    set_annotation_meta(seg, names={i + 1: n for i, n in enumerate(nm)}, colors={})
    assert seg.GetMetaDataKeys()[0] == "Segment0_ID" and seg.GetMetaDataKeys().index("Segment10_ID") < seg.GetMetaDataKeys().index("Segment1_ID")
    out = combine_segmentations([seg])
    got = [v["name"] for v in sorted(get_annotation_labels(out, fetch=False).values(), key=lambda v: v["value"])]
    assert got == want
    assert np.array_equal(np.asarray(out.array), merged)


def test_orientation_remap_against_hand_derived_vectors():
    """sitk.DICOMOrient(img, 'RAI') (reference ts2d/core/util/image.py:32-43) as literal expectations worked out by hand, so that the
    product's host logic and the oracle are not only checked against each other.  DICOMOrient letters say where an axis points TO in
    LPS (+x = L, +y = P, +z = S); column r of the direction matrix is where image axis r points.
      diag(-1,-1,+1) (the reference's sample_s0521): x -> R, y -> A already; z -> S must become I: z flip only
      identity (LPS): all three axes reversed
      columns (+y, -x, -z): R is image axis 1 as it is, A is axis 0 reversed (it points P), I is axis 2 as it is
      columns (+x, +z, -y): R is axis 0 reversed, A is axis 2 as it is, I is axis 1 reversed
      diag(-1,-1,-1) turned 0.3 rad about z (oblique): the dominant cosines decide: nothing to do"""
    from oracle import geometry
    from totalsegmentator2d_b200.core.util.image import orientation_remap
    th = 0.3
    rot = np.array([[np.cos(th), -np.sin(th), 0], [np.sin(th), np.cos(th), 0], [0, 0, 1.0]])
    cases = [
        (np.diag([-1.0, -1.0, 1.0]), (0, 1, 2), (False, False, True)),
        (np.eye(3), (0, 1, 2), (True, True, True)),
        (np.array([[0.0, -1.0, 0.0], [1.0, 0.0, 0.0], [0.0, 0.0, -1.0]]), (1, 0, 2), (False, True, False)),
        (np.array([[1.0, 0.0, 0.0], [0.0, 0.0, -1.0], [0.0, 1.0, 0.0]]), (0, 2, 1), (True, False, True)),
        (rot @ np.diag([-1.0, -1.0, -1.0]), (0, 1, 2), (False, False, False)),
    ]
    for D, perm, flip in cases:
        for fn in (orientation_remap, geometry.orientation_remap):
            got = fn(D * 1.5 if fn is orientation_remap else D, "RAI")       # a spacing factor on the matrix must not matter to the product's rule
            assert (tuple(got[0]), tuple(bool(f) for f in got[1])) == (perm, flip), (fn.__module__, D.tolist(), got)
    # the same rule read the other way: for target 'LPS' every flip is the opposite one
    for D, perm, flip in cases:
        got = orientation_remap(D, "LPS")
        assert tuple(got[0]) == perm and tuple(bool(f) for f in got[1]) == tuple(not f for f in flip)


def test_default_compute_type_is_the_references_autocast_type():
    """fp16 (what nnU-Net's autocast computes in on CUDA) is the default storage type of the engine, the bench and the tools: it is the
    16-bit type that meets the 99.9 % mask gate against the fp32 oracle (bf16: 0.9939 on random-init weights, tests/test_gpu_pipeline.py)"""
    import inspect
    import importlib.util
    import os
    from totalsegmentator2d_b200.core.inference.predictor import B200nnUNetPredictor
    assert inspect.signature(B200nnUNetPredictor.__init__).parameters["precision"].default == "fp16"
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = open(os.path.join(root, "bench.py")).read()
    assert 'os.environ.get("TS2D_BENCH_PRECISION", "fp16")' in src
    src = open(os.path.join(root, "totalsegmentator2d_b200", "core", "inference", "nnu.py")).read()
    assert '"b200.precision", "fp16"' in src


def test_c_abi_library_exports_every_declared_symbol():
    """the drop-in boundary: libts2d_b200.so loads without a GPU and exports every entry point include/ts2d_b200.h declares
    (no compute call here); the product loader refuses to run without it (no CPU fallback)"""
    import re
    from totalsegmentator2d_b200 import _lib
    hdr = open(os.path.join(ROOT, "include", "ts2d_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    hdr = re.sub(r"//[^\n]*", "", hdr)
    names = sorted(set(re.findall(r"\b(ts2d_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 25 and "ts2d_project" in names and "ts2d_unet_forward" in names and "ts2d_unet_blend_head" in names
    path = os.path.join(ROOT, "totalsegmentator2d_b200", "libts2d_b200.so")
    assert os.path.exists(path), "build the extension first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(path)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    lib.ts2d_abi_version.restype = ctypes.c_int
    assert lib.ts2d_abi_version() == 1
    lib.ts2d_last_error.restype = ctypes.c_char_p
    assert isinstance(lib.ts2d_last_error(), (bytes, type(None)))
    # the product path fails loudly when the library is absent, and without a GPU (this container has none)
    import pytest
    import torch
    saved_lib, saved_path = _lib._lib, _lib.LIB_PATH
    try:
        _lib._lib, _lib.LIB_PATH = None, os.path.join(ROOT, "does_not_exist.so")
        with pytest.raises(_lib.TS2DError, match="no CPU fallback"):
            _lib.load()
    finally:
        _lib._lib, _lib.LIB_PATH = saved_lib, saved_path
    if not torch.cuda.is_available():
        with pytest.raises(_lib.TS2DError, match="no CUDA device"):
            _lib.load(check_device=True)


def test_nifti_and_metaimage_readers_round_trip(tmp_path):
    """reference ts2d/main.py:25 whitelists nii / nii.gz / mha / mhd and reads them with sitk.ReadImage (image.py:225-238): the
    restated readers must return the same [z][y][x] array and LPS geometry for every writer variant, incl. an oblique direction,
    a flipped axis (negative determinant -> qfac = -1), big-endian data, scl_slope / scl_inter and detached .mhd data."""
    import gzip
    import struct
    from totalsegmentator2d_b200 import formats, nrrd
    from totalsegmentator2d_b200.core.util.image import read_image
    from totalsegmentator2d_b200.image import Image
    rng = np.random.default_rng(5)
    arr = rng.integers(-1024, 3071, size=(7, 5, 9)).astype(np.int16)
    th = 0.3
    rot = np.array([[np.cos(th), -np.sin(th), 0], [np.sin(th), np.cos(th), 0], [0, 0, 1.0]])
    for D in (np.eye(3), np.diag([-1.0, -1.0, 1.0]), rot, rot @ np.diag([1.0, -1.0, 1.0]), np.eye(3)[[2, 0, 1]]):
        img = Image(arr, (0.7, 1.5, 2.25), (10.0, -20.0, 30.5), D.reshape(-1))
        for ext in ("nii", "nii.gz", "mha", "mhd", "nrrd"):
            fp = str(tmp_path / f"v.{ext}")
            (nrrd.write_nrrd(img, fp, True) if ext == "nrrd" else formats.write_any(img, fp))
            back = read_image(fp)
            assert np.array_equal(back.array, arr) and back.array.dtype == np.int16, ext
            assert np.allclose(back.GetSpacing(), img.GetSpacing(), atol=1e-5), (ext, back.GetSpacing())
            assert np.allclose(back.GetOrigin(), img.GetOrigin(), atol=1e-4), (ext, back.GetOrigin())
            assert np.allclose(back.direction_matrix(), D, atol=1e-5), (ext, back.direction_matrix())
    # sform-only, big-endian, scaled NIfTI written by hand: x fastest, RAS affine diag(2,3,4) + offset -> LPS flips x and y
    a = np.arange(2 * 3 * 4, dtype=">i2").reshape(4, 3, 2)
    hdr = bytearray(352)
    struct.pack_into(">i", hdr, 0, 348)
    struct.pack_into(">8h", hdr, 40, 3, 2, 3, 4, 1, 1, 1, 1)
    struct.pack_into(">2h", hdr, 70, 4, 16)
    struct.pack_into(">8f", hdr, 76, 1, 2, 3, 4, 0, 0, 0, 0)
    struct.pack_into(">3f", hdr, 108, 352.0, 0.5, -3.0)
    struct.pack_into(">2h", hdr, 252, 0, 1)
    struct.pack_into(">12f", hdr, 280, 2, 0, 0, 5, 0, 3, 0, 6, 0, 0, 4, 7)
    hdr[344:348] = b"n+1\0"
    fp = str(tmp_path / "be.nii.gz")
    with gzip.open(fp, "wb") as f:
        f.write(bytes(hdr) + a.tobytes())
    back = read_image(fp)
    assert back.array.dtype == np.float32 and np.allclose(back.array, a.astype(np.float64) * 0.5 - 3.0)
    assert np.allclose(back.GetSpacing(), (2, 3, 4)) and np.allclose(back.GetOrigin(), (-5, -6, 7))
    assert np.allclose(back.direction_matrix(), np.diag([-1.0, -1.0, 1.0]))
    # vector MetaImage (2-D, 3 channels) and the NRRD metadata escape round trip (backslash followed by 'n')
    v = rng.integers(0, 255, size=(6, 4, 3)).astype(np.uint8)
    vi = Image(v, (1.0, 2.0), (0.0, 1.0), None, is_vector=True)
    formats.write_any(vi, str(tmp_path / "vec.mha"))
    vb = read_image(str(tmp_path / "vec.mha"))
    assert vb.is_vector and np.array_equal(vb.array, v) and vb.GetSize() == (4, 6)
    vi.SetMetaData("k", "a\\nb\nc\\\\")
    nrrd.write_nrrd(vi, str(tmp_path / "m.nrrd"), False)
    assert nrrd.read_nrrd(str(tmp_path / "m.nrrd")).GetMetaData("k") == "a\\nb\nc\\\\"


def test_nifti_vector_masks_and_sform_over_qform(tmp_path):
    """(1) Result.save(ext='nii.gz') writes multi-label masks: vector pixels go out the way ITK writes them (5-D file, dim[5] =
    components, intent vector) and read back; the nnU-Net export seam honours dataset.json's file_ending through the same dispatcher.
    (2) a file whose sform and qform disagree is read with the sform when that is a rotation x scale (ITK 5.3+ = SimpleITK 2.5.x, the
    reference's pin), with the qform when the sform is sheared, and spacing is pixdim."""
    import struct
    from totalsegmentator2d_b200 import formats
    from totalsegmentator2d_b200.core.util.image import read_image
    from totalsegmentator2d_b200.image import Image
    rng = np.random.default_rng(9)
    for shape, sp, org in [((5, 4, 6, 7), (0.7, 1.5, 2.25), (1.0, -2.0, 3.0)), ((4, 6, 3), (1.5, 2.0), (0.5, 1.0))]:
        v = (rng.random(shape) > 0.5).astype(np.uint8)
        vi = Image(v, sp, org, None, is_vector=True)
        for ext in ("nii", "nii.gz"):
            fp = str(tmp_path / f"vec{len(shape)}.{ext}")
            formats.write_any(vi, fp)
            raw = formats.gzip.open(fp, "rb").read() if ext.endswith("gz") else open(fp, "rb").read()
            dim = struct.unpack("<8h", raw[40:56])
            assert dim[0] == 5 and dim[4] == 1 and dim[5] == shape[-1] and struct.unpack("<h", raw[68:70])[0] == 1007
            back = read_image(fp)
            assert back.is_vector and back.array.dtype == np.uint8
            assert np.array_equal(np.asarray(back.array).reshape(shape), v)
            assert np.allclose(back.GetSpacing()[:len(sp)], sp, atol=1e-6) and np.allclose(back.GetOrigin()[:len(org)], org, atol=1e-5)
    # the dispatcher also serves .nrrd, so export.py can route every file_ending through it
    formats.write_any(vi, str(tmp_path / "vec.nrrd"))
    assert np.array_equal(formats.read_any(str(tmp_path / "vec.nrrd")).array, v)

    def nifti(qform_code, sform_code, srow, quat=(0.0, 0.0, 0.0), qoff=(1.0, 2.0, 3.0), pixdim=(1.0, 2.0, 3.0, 4.0)):
        a = np.arange(2 * 3 * 4, dtype="<i2").reshape(4, 3, 2)
        hdr = bytearray(352)
        struct.pack_into("<i", hdr, 0, 348)
        struct.pack_into("<8h", hdr, 40, 3, 2, 3, 4, 1, 1, 1, 1)
        struct.pack_into("<2h", hdr, 70, 4, 16)
        struct.pack_into("<8f", hdr, 76, *pixdim, 0, 0, 0, 0)
        struct.pack_into("<3f", hdr, 108, 352.0, 1.0, 0.0)
        struct.pack_into("<2h", hdr, 252, qform_code, sform_code)
        struct.pack_into("<6f", hdr, 256, *quat, *qoff)
        struct.pack_into("<12f", hdr, 280, *srow)
        hdr[344:348] = b"n+1\0"
        fp = str(tmp_path / f"q{qform_code}s{sform_code}_{abs(hash(tuple(srow))) % 9999}.nii")
        with open(fp, "wb") as f:
            f.write(bytes(hdr) + a.tobytes())
        return fp

    # qform: identity rotation (RAS), offset (1,2,3); sform: 90 degree turn about z with offset (5,6,7) -> the sform must win
    turned = (0, -3, 0, 5, 2, 0, 0, 6, 0, 0, 4, 7)           # columns: x -> +y (2 mm), y -> -x (3 mm), z -> +z (4 mm)
    back = read_image(nifti(1, 1, turned))
    assert np.allclose(back.GetSpacing(), (2, 3, 4))
    assert np.allclose(back.GetOrigin(), (-5, -6, 7))
    assert np.allclose(back.direction_matrix(), np.diag([-1.0, -1.0, 1.0]) @ np.array([[0, -1, 0], [1, 0, 0], [0, 0, 1.0]]))
    # the same file with a sheared sform: not a rotation x scale, so the qform (identity, offset (1,2,3)) is used
    sheared = (2, 1.5, 0, 5, 0, 3, 0, 6, 0, 0, 4, 7)
    back = read_image(nifti(1, 1, sheared))
    assert np.allclose(back.GetOrigin(), (-1, -2, 3)) and np.allclose(back.direction_matrix(), np.diag([-1.0, -1.0, 1.0]))
    assert np.allclose(back.GetSpacing(), (2, 3, 4))
    # qform only / sform only keep working
    assert np.allclose(read_image(nifti(1, 0, turned)).GetOrigin(), (-1, -2, 3))
    assert np.allclose(read_image(nifti(0, 2, turned)).GetOrigin(), (-5, -6, 7))


def test_visual_restatement_properties():
    """oracle/visual.py (create_visual restatement): identity geometry is the identity, label = highest non-zero component,
    background black, 2x finer uniform grid replicates pixels around the kept centre index, window endpoints map to 0 / 255."""
    from oracle import visual as ovis
    m = np.zeros((4, 6, 3), np.uint8)
    m[1, 2, 0] = 1
    m[1, 2, 2] = 1
    m[3, 5, 1] = 7
    cmap = [[255, 255, 255], [10, 20, 30], [40, 50, 60], [70, 80, 90]]
    out = ovis.visual_labels(m, (4, 6), (1.0, 1.0), cmap)
    assert out[1, 2].tolist() == [70, 80, 90] and out[3, 5].tolist() == [40, 50, 60] and out[0, 0].tolist() == [0, 0, 0]
    fine = ovis.visual_labels(m, (8, 6), (0.5, 1.0), cmap)               # rows twice as fine: input index 2 stays at output index 4,
    # output rows 1, 2 sample input indices 0.5 (rounds half up to 1) and 1.0; row 3 samples 1.5 -> 2
    assert fine.shape == (8, 6, 3) and fine[1, 2].tolist() == [70, 80, 90] and fine[2, 2].tolist() == [70, 80, 90] and fine[3, 2].tolist() == [0, 0, 0]
    a = np.array([[0.0, 50.0], [100.0, 25.0]], np.float32)
    assert ovis.visual_intensity(a, a.shape, (1, 1), resample=False).tolist() == [[0, 127], [255, 63]]
    assert np.array_equal(ovis.visual_intensity(a, a.shape, (1.0, 1.0)), ovis.visual_intensity(a, a.shape, (1, 1), resample=False))


def test_unpack_mask_bits_host_is_numpy_unpackbits():
    """host half of the bit-packed mask transport (no GPU call): ts2d_unpack_mask_bits_host == np.unpackbits(bitorder='little'), for
    lengths that are not multiples of 8 and for a study-sized mask with several threads"""
    from totalsegmentator2d_b200 import ops
    rng = np.random.default_rng(12)
    for n, threads in ((0, 1), (5, 1), (8, 2), (117 * 13, 4), (1000 * 512 * 3, 4), (1 << 21, 8)):
        m = (rng.random(n) > 0.7).astype(np.uint8)
        packed = np.packbits(m, bitorder="little")
        assert np.array_equal(ops.unpack_mask_bits_host(packed, (n,), threads=threads), m), n
    with pytest.raises(Exception):
        ops.unpack_mask_bits_host(np.zeros(3, np.uint8), (100,))
