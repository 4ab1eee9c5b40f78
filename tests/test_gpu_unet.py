"""GPU parity: the U-Net engine vs the oracle's torch-CPU PlainConvUNet (fp32 CUDA-core path <= 1e-4; bf16/fp16
tcgen05 path <= the stated mixed-precision tolerance), layer by layer and end to end through blend/export."""
import numpy as np
import pytest
import torch

from oracle import nnunet_inf, nnunet_net

pytestmark = pytest.mark.gpu


def _make(features, patch, heads=5, in_ch=2, precision="fp32", tc=True, seed=3, strides=None, mirror=True, n_conv=2):
    from totalsegmentator2d_b200 import synth
    from totalsegmentator2d_b200.core.inference.predictor import B200nnUNetPredictor
    arch = synth.make_arch_kwargs(features, strides, n_conv, n_conv)
    labels = [f"l{i}" for i in range(heads)]
    plans = synth.make_plans(arch, patch, (1.5, 1.5), ["ZScoreNormalization"] * in_ch)
    dj = synth.make_dataset_json([f"c{i}" for i in range(in_ch)], labels)
    sd = synth.init_state_dict(arch, in_ch, heads, seed, with_aliases=True)
    p = B200nnUNetPredictor(use_mirroring=mirror, precision=precision, use_tensor_cores=tc, max_batch=4)
    p.manual_initialization(plans, "2d", dj, [sd], (0, 1))
    spec = nnunet_net.ArchSpec(arch, in_ch, heads)
    return p, sd, spec


def _tiles(n, c, patch, seed, dtype):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(n, c, *patch, generator=g)
    return x, x.permute(0, 2, 3, 1).contiguous().to(dtype).cuda()


@pytest.mark.parametrize("features,patch,strides", [([8, 16, 32], (32, 48), None), ([4, 8], (16, 16), None),
                                                     ([8, 16, 16], (24, 40), [[1, 1], [2, 2], [2, 1]])])
def test_fp32_direct_path_logits(features, patch, strides):
    p, sd, spec = _make(features, patch, precision="fp32", strides=strides)
    x, xd = _tiles(3, 2, patch, 1, torch.float32)
    got = p.predict_tile_logits(xd).cpu()
    with torch.no_grad():
        ref = nnunet_net.unet_forward(x, sd, spec)
    assert float((got - ref).abs().max()) <= 1e-4, float((got - ref).abs().max())


def _layer_errors(p, sd, spec, x, xd):
    coll = {}
    with torch.no_grad():
        nnunet_net.unet_forward(x, sd, spec, collect=coll)
    errs = {}
    for name, ref in coll.items():
        got = p.forward_debug(xd, name, tuple(ref.shape[1:])).cpu()
        errs[name] = float((got - ref).abs().max())
    return errs


@pytest.mark.parametrize("precision,tol", [("bf16", 0.2), ("fp16", 0.025)])
def test_tensor_core_layers(precision, tol):
    """every conv layer of a TC-eligible net (KC=32 and 64, stride-2 parity maps, several BN) against the fp32 oracle.
    Activations are O(1) after InstanceNorm; bf16 has 8 mantissa bits (2^-9 = 2e-3 per rounding), fp16 11 bits."""
    dt = torch.bfloat16 if precision == "bf16" else torch.float16
    p, sd, spec = _make([32, 64, 128, 256], (64, 64), heads=7, precision=precision, seed=5)
    x, xd = _tiles(4, 2, (64, 64), 2, dt)
    x = xd.float().cpu().permute(0, 3, 1, 2).contiguous()      # the oracle sees the rounded input
    errs = _layer_errors(p, sd, spec, x, xd)
    print(precision, {k: round(v, 4) for k, v in errs.items()})
    assert max(errs.values()) <= tol, errs


@pytest.mark.parametrize("precision,tol", [("fp16", 0.08), ("bf16", 0.6)])      # 7 stages deep: ~1.25x per layer (fp16 0.062, bf16 0.45 measured)
def test_tensor_core_anisotropic_strides(precision, tol, monkeypatch):
    """SURVEY 8d's second plan, CS-640x320 (7 stages, non-square patch 640 x 320, one [2,1] stride; here also a [1,2] one) on the
    TENSOR-CORE path: every axis of the 5-D parity view strides independently, the transposed convs have 2 taps.  No layer may
    fall to conv_direct_kernel (direct_layer_count == 0), every layer is checked against the fp32 oracle, and the tensor-core
    path agrees with the CUDA-core path at the same storage precision.  conv_tc2_kernel is forced on (TS2D_TC2_MIN_UNITS=1) so
    its producer takes the anisotropic view too."""
    dt = torch.bfloat16 if precision == "bf16" else torch.float16
    strides = [[1, 1], [2, 2], [2, 1], [2, 2], [1, 2], [2, 2], [2, 1]]
    feats = [32, 64, 128, 128, 256, 256, 256]
    monkeypatch.setenv("TS2D_TC2_MIN_UNITS", "1")
    p, sd, spec = _make(feats, (640, 320), heads=6, precision=precision, seed=31, strides=strides)
    assert p.direct_layer_count() == 0
    x, xd = _tiles(2, 2, (640, 320), 12, dt)
    x = xd.float().cpu().permute(0, 3, 1, 2).contiguous()
    errs = _layer_errors(p, sd, spec, x, xd)
    print(precision, {k: round(v, 4) for k, v in errs.items()})
    assert max(errs.values()) <= tol, errs
    a = p.predict_tile_logits(xd).cpu()
    p.close()
    pd, _, _ = _make(feats, (640, 320), heads=6, precision=precision, seed=31, strides=strides, tc=False)
    b = pd.predict_tile_logits(xd).cpu()
    pd.close()
    assert float((a - b).abs().max()) <= tol, float((a - b).abs().max())


def test_tensor_core_equals_direct_same_precision():
    """tcgen05 path vs CUDA-core path at the same storage precision: only fp32 accumulation order differs"""
    pt, sd, spec = _make([32, 64, 128], (64, 96), heads=6, precision="fp16", tc=True, seed=8)
    pd, _, _ = _make([32, 64, 128], (64, 96), heads=6, precision="fp16", tc=False, seed=8)
    x, xd = _tiles(3, 2, (64, 96), 4, torch.float16)
    a = pt.predict_tile_logits(xd).cpu()
    b = pd.predict_tile_logits(xd).cpu()
    assert float((a - b).abs().max()) <= 2e-2, float((a - b).abs().max())


@pytest.mark.parametrize("precision,tol", [("bf16", 0.03), ("fp16", 0.004)])
@pytest.mark.parametrize("patch", [(40, 72), (16, 8), (48, 36)])
def test_first_conv_tensor_core(precision, tol, patch, monkeypatch):
    """conv_first_tc_kernel (im2col rows built by the workers, bias in two extra K columns): first layer against the fp32
    oracle on patches whose 8 x 16 pixel tiles overhang the image, whose border tiles are the majority, and with 64 output
    channels (two channel blocks); and against the CUDA-core first-layer kernel (TS2D_FIRST_TC=0)."""
    dt = torch.bfloat16 if precision == "bf16" else torch.float16
    outs = []
    for first_tc in ("1", "0"):
        monkeypatch.setenv("TS2D_FIRST_TC", first_tc)
        p, sd, spec = _make([64, 64], patch, heads=3, precision=precision, seed=21, strides=[[1, 1], [2, 2]])
        x, xd = _tiles(3, 2, patch, 5, dt)
        x = xd.float().cpu().permute(0, 3, 1, 2).contiguous()
        coll = {}
        with torch.no_grad():
            nnunet_net.unet_forward(x, sd, spec, collect=coll)
        ref = coll["enc0.0"]
        got = p.forward_debug(xd, "enc0.0", tuple(ref.shape[1:])).cpu()
        outs.append(got)
        assert float((got - ref).abs().max()) <= tol, (first_tc, float((got - ref).abs().max()))
        p.close()
    # the two kernels round the same fp32 sums independently: at most one storage ulp apart (2^-5 for |x| in [4, 8) in bf16)
    assert float((outs[0] - outs[1]).abs().max()) <= 2 * tol


def test_tensor_core_ragged_tiles_and_small_levels():
    """W=160/80/40/20 (tiles overhang the image), deepest level 5x3.. covers masked rows and multi-image tiles"""
    p, sd, spec = _make([32, 32, 64, 64, 64], (48, 160), heads=4, precision="fp16", seed=9)
    x, xd = _tiles(3, 2, (48, 160), 6, torch.float16)
    x = xd.float().cpu().permute(0, 3, 1, 2).contiguous()
    errs = _layer_errors(p, sd, spec, x, xd)
    assert max(errs.values()) <= 0.05, errs          # 5 stages deep, fp16 storage: ~1.3x per level (see test_tensor_core_layers)


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-4), ("fp16", 2e-2), ("bf16", 1.5e-1)])
@pytest.mark.parametrize("mirror", [False, True])
def test_sliding_window_logits_and_mask(precision, tol, mirror):
    """predict_logits_from_preprocessed_data (tiling + mirror TTA + Gaussian blend + head) vs the oracle, and the
    thresholded masks: >= 99.9% agreement for fp32/fp16, reported for bf16."""
    feats = [32, 64, 128] if precision != "fp32" else [8, 16, 32]
    p, sd, spec = _make(feats, (64, 64), heads=6, precision=precision, seed=11, mirror=mirror)
    g = torch.Generator().manual_seed(13)
    data = torch.randn(2, 1, 150, 100, generator=g)
    got = p.predict_logits_from_preprocessed_data(data).cpu()
    net = lambda t: nnunet_net.unet_forward(t, sd, spec)
    with torch.no_grad():
        ref = nnunet_inf.predict_sliding_window(net, data, (64, 64), 6, 0.5, (0, 1) if mirror else None, True, "ideal")
    err = float((got - ref).abs().max())
    agree = float(((got > 0) == (ref > 0)).float().mean())
    print(f"{precision} mirror={mirror}: max |dlogit| {err:.4g}, mask agreement {agree:.5f}")
    assert got.shape == ref.shape and err <= tol
    if precision != "bf16":
        assert agree >= 0.999


@pytest.mark.parametrize("precision", ["fp16", "bf16", "fp32"])
def test_no_kernel_stores_outside_its_buffer(precision, monkeypatch):
    """Stand-in for compute-sanitizer memcheck (closed on the GPU pool): with TS2D_ARENA_GUARD every workspace buffer of the engine
    is followed by a 64 KiB canary band; after forwards that exercise every conv kernel (first-layer tcgen05, halo + fused transform,
    conv_tc with ragged / multi-image tiles and split-K, conv_tc2, anisotropic strides and 2-tap transposed convs, the CUDA-core
    kernels in fp32) on shapes whose tiles overhang the image, no band may be damaged.  Outputs handed in by the caller (logits, mask)
    get their own canaries."""
    monkeypatch.setenv("TS2D_ARENA_GUARD", "65536")
    monkeypatch.setenv("TS2D_TC2_MIN_UNITS", "1")
    dt = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[precision]
    cases = [([32, 32, 64, 64, 64], (48, 160), None), ([32, 64, 128, 256, 256], (80, 96), None),
             ([32, 64, 128, 128], (80, 48), [[1, 1], [2, 2], [2, 1], [1, 2]])]
    if precision == "fp32":
        cases = [([8, 16, 32], (40, 72), None), ([8, 16, 16], (24, 40), [[1, 1], [2, 2], [2, 1]])]
    for feats, patch, strides in cases:
        p, sd, spec = _make(feats, patch, heads=5, precision=precision, seed=3, strides=strides)
        x, xd = _tiles(3, 2, patch, 7, dt)
        logits = p.predict_tile_logits(xd)
        n, bad = p.check_guards()
        assert n > 0 and bad == 0, (feats, patch, n, bad)
        # blend + head + mask export into canary-framed outputs
        g = torch.Generator().manual_seed(5)
        data = torch.randn(2, patch[0] + 21, patch[1] + 13, generator=g).cuda()
        tiles, layout = p.make_tiles(data)
        h, w = data.shape[1:]
        buf = torch.full((h + 4, w, 9), 0x5A, dtype=torch.uint8, device="cuda")
        p.run_tiles(tiles, layout, (h, w), mask=buf[2:-2], crop_origin=(0, 0), ch_off=2)
        assert int((buf[:2] != 0x5A).sum()) == 0 and int((buf[-2:] != 0x5A).sum()) == 0
        assert int((buf[2:-2, :, :2] != 0x5A).sum()) == 0 and int((buf[2:-2, :, 7:] != 0x5A).sum()) == 0
        n, bad = p.check_guards()
        assert bad == 0, (feats, patch, n, bad)
        assert torch.isfinite(logits).all()
        p.close()


@pytest.mark.parametrize("precision", ["fp16", "bf16"])
@pytest.mark.parametrize("feats,patch", [([32, 64, 128], (96, 80)), ([64, 64, 64], (72, 104)), ([32, 32, 64], (64, 48))])
def test_stride2_halo_kernel(precision, feats, patch, monkeypatch):
    """conv_halo_s2_kernel (four parity-plane halos per tile, taps as shifted descriptors, resident weights) against conv_tc_kernel on the
    same strided layer (TS2D_CONV_HALO_S2=0): with one K-chunk per tap both accumulate the same products in the same order, so the conv
    outputs agree bit for bit wherever the InstanceNorm statistics (grouped differently) round alike; and against the fp32 oracle.
    Patches whose 8 x 16 output tiles overhang the image, 32 / 64 input channels (KC = 32 / 64), 32 / 64 output channels."""
    dt = torch.bfloat16 if precision == "bf16" else torch.float16
    outs = {}
    for s2 in ("1", "0"):
        monkeypatch.setenv("TS2D_CONV_HALO_S2", s2)
        p, sd, spec = _make(feats, patch, heads=4, precision=precision, seed=29)
        x, xd = _tiles(3, 2, patch, 8, dt)
        x = xd.float().cpu().permute(0, 3, 1, 2).contiguous()
        coll = {}
        with torch.no_grad():
            nnunet_net.unet_forward(x, sd, spec, collect=coll)
        outs[s2] = {n: p.forward_debug(xd, n, tuple(coll[n].shape[1:])).cpu() for n in ("enc1.0", "enc1.1", "enc2.0")}
        if s2 == "1":
            errs = {n: float((outs[s2][n] - coll[n]).abs().max()) for n in outs[s2]}
            assert max(errs.values()) <= (0.15 if precision == "bf16" else 0.02), errs
        p.close()
    ulp = 2.0 ** -5 if precision == "bf16" else 2.0 ** -8
    d = (outs["1"]["enc1.0"] - outs["0"]["enc1.0"]).abs()
    assert float(d.max()) <= ulp and float((d == 0).float().mean()) >= 0.98, (float(d.max()), float((d == 0).float().mean()))
    for n in ("enc1.1", "enc2.0"):
        assert float((outs["1"][n] - outs["0"][n]).abs().max()) <= 2 * ulp


def test_errors_are_loud():
    from totalsegmentator2d_b200 import _lib as L
    p, sd, spec = _make([8, 16], (16, 16))
    with pytest.raises(L.TS2DError):
        p.predict_logits_from_preprocessed_data(torch.zeros(2, 16, 16))
    with pytest.raises(L.TS2DError):
        _make([8, 16], (17, 17))                      # patch not divisible by the strides


@pytest.mark.parametrize("precision", ["bf16", "fp16"])
@pytest.mark.parametrize("fuse", ["2", "3"])
def test_deferred_norm_matches_materialised(precision, fuse, monkeypatch):
    """conv_halo_kernel<FUSE> (the producer's InstanceNorm + LeakyReLU applied to the consumer's halo in shared memory by dedicated
    transform warps, same fma / leaky arithmetic as norm_act_kernel) against the materialised norm_act_kernel path.  The fused kernel
    has ONE epilogue warpgroup (the plain kernel two), so the fp32 partial sums of ITS InstanceNorm statistics are grouped differently:
    the first fused layer's input transform is exact (enc0.1's output is bit-equal up to those statistics, i.e. rare 1-ulp flips after
    its own normalisation), later layers inherit the flips.  TS2D_FUSE_NORM=3 fuses only the 32-channel consumers."""
    dt = torch.bfloat16 if precision == "bf16" else torch.float16
    outs = []
    for f in ("0", fuse):
        monkeypatch.setenv("TS2D_FUSE_NORM", f)
        p, sd, spec = _make([32, 64, 128, 128], (64, 96), heads=6, precision=precision, seed=11)
        x, xd = _tiles(4, 2, (64, 96), 9, dt)
        tile_logits = p.predict_tile_logits(xd).cpu()
        first = p.forward_debug(xd, "enc0.1", (32, 64, 96)).cpu()
        mid = p.forward_debug(xd, "enc1.1", (64, 32, 48)).cpu()
        g = torch.Generator().manual_seed(4)
        data = torch.randn(2, 1, 100, 130, generator=g)
        blended = p.predict_logits_from_preprocessed_data(data).cpu()
        outs.append((tile_logits, first, mid, blended))
        p.close()
    (tl_a, f_a, mid_a, bl_a), (tl_b, f_b, mid_b, bl_b) = outs
    ulp = 2.0 ** -5 if precision == "bf16" else 2.0 ** -8            # one storage ulp for |x| in [4, 8)
    d = (f_a - f_b).abs()
    assert float(d.max()) <= ulp and float((d == 0).float().mean()) >= 0.99, (float(d.max()), float((d == 0).float().mean()))
    d = (mid_a - mid_b).abs()
    assert float(d.max()) <= 2 * ulp and float((d == 0).float().mean()) >= 0.90, (float(d.max()), float((d == 0).float().mean()))
    tol = 5e-2 if precision == "bf16" else 1e-2
    assert float((tl_a - tl_b).abs().max()) <= tol
    # the blend kernel normalises the last decoder conv in fp32 and feeds the fp32 values to the blend (the materialised
    # path rounds them to the storage type first): one storage rounding of the 32 features apart
    assert float((bl_a - bl_b).abs().max()) <= tol


@pytest.mark.parametrize("precision,tol", [("bf16", 0.25), ("fp16", 0.04)])      # five levels deep: a little above the 4-level test
def test_partial_mode_tiny_levels(precision, tol, monkeypatch):
    """4x4 / 8x8 levels (several images per 128-pixel tile): conv_tc_kernel stores fp32 accumulators ("partial" mode, optionally with K
    split over 3 or 9 tap groups on separate CTAs) and splitk_finish_kernel = slice sum + bias + InstanceNorm statistics + normalise +
    LeakyReLU, against the epilogue-atomics + norm_act_kernel path (TS2D_KSPLIT=0) and the fp32 oracle; stride-2 and stride-1 convs."""
    dt = torch.bfloat16 if precision == "bf16" else torch.float16
    outs = {}
    for mode in ("0", "1", "3", "9"):
        monkeypatch.setenv("TS2D_KSPLIT", mode)
        p, sd, spec = _make([32, 64, 128, 256, 256], (64, 64), heads=7, precision=precision, seed=5)
        x, xd = _tiles(3, 2, (64, 64), 2, dt)
        x = xd.float().cpu().permute(0, 3, 1, 2).contiguous()
        errs = _layer_errors(p, sd, spec, x, xd)
        assert max(errs.values()) <= tol, (mode, errs)
        outs[mode] = (p.forward_debug(xd, "enc4.1", (256, 4, 4)).cpu(), p.predict_tile_logits(xd).cpu())
        p.close()
    for mode in ("1", "3", "9"):
        # same products, different fp32 summation order (and statistics from fp32 sums instead of per-value double atomics)
        assert float((outs[mode][0] - outs["0"][0]).abs().max()) <= (6e-2 if precision == "bf16" else 8e-3)
        assert float((outs[mode][1] - outs["0"][1]).abs().max()) <= tol
    # deterministic: slices are added in a fixed order
    monkeypatch.setenv("TS2D_KSPLIT", "9")
    p, sd, spec = _make([32, 64, 128, 256, 256], (64, 64), heads=7, precision=precision, seed=5)
    again = p.predict_tile_logits(xd).cpu()
    assert torch.equal(again, outs["9"][1])


@pytest.mark.parametrize("precision", ["bf16", "fp16"])
def test_multicast_pairs_bit_identical(precision, monkeypatch):
    """conv_tc_kernel as clusters of two CTAs (TS2D_MC=1: each CTA loads half of the weight tile, TMA multicast into both shared
    memories, tcgen05.commit multicast frees the slot for both producers) computes exactly what the single-CTA kernel computes."""
    dt = torch.bfloat16 if precision == "bf16" else torch.float16
    outs = []
    for mc in ("0", "1"):
        monkeypatch.setenv("TS2D_MC", mc)
        p, sd, spec = _make([32, 64, 128, 256], (64, 96), heads=6, precision=precision, seed=13)
        x, xd = _tiles(4, 2, (64, 96), 6, dt)
        outs.append((p.predict_tile_logits(xd).cpu(), p.forward_debug(xd, "dec0.0", (128, 16, 24)).cpu()))
        p.close()
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])


@pytest.mark.parametrize("precision", ["fp16", "bf16"])
@pytest.mark.parametrize("feats,mirror", [([32, 64, 128], True), ([16, 32, 64], False), ([64, 64], True)])
def test_blend_v2_equals_v1(precision, feats, mirror, monkeypatch):
    """blend_head_v2_kernel (F/8 lanes per pixel, head through shared memory, whole-image grid that also writes the zeros outside the
    crop box) against blend_head_kernel: logits equal up to the re-association of the head's fp32 dot product, mask bytes equal
    wherever |logit| is not tiny, and the v2 mask needs no memset (the buffer starts as 0xFF)."""
    from totalsegmentator2d_b200 import ops
    p, sd, spec = _make(feats, (64, 64), heads=7, precision=precision, seed=17, mirror=mirror)
    g = torch.Generator().manual_seed(2)
    data = torch.randn(2, 100, 150, generator=g).cuda()
    x, layout = p.make_tiles(data)
    outs = {}
    for v2 in ("0", "1"):
        monkeypatch.setenv("TS2D_BLEND_V2", v2)
        logits = torch.empty((7, 100, 150), dtype=torch.float32, device="cuda")
        p.run_tiles(x, layout, (100, 150), logits=logits)
        mask = torch.full((120, 170, 20), 255 if v2 == "1" else 0, dtype=torch.uint8, device="cuda")
        present = torch.zeros(20, dtype=torch.uint8, device="cuda")
        p.run_tiles(x, layout, (100, 150), mask=mask, crop_origin=(11, 13), ch_off=5, present=present)
        outs[v2] = (logits.cpu(), mask.cpu(), present.cpu())
    assert p.blend_covers_image
    la, lb = outs["0"][0], outs["1"][0]
    assert float((la - lb).abs().max()) <= 2e-5, float((la - lb).abs().max())
    ma, mb = outs["0"][1], outs["1"][1]
    assert torch.equal(mb[..., :5], torch.full_like(mb[..., :5], 255)) and torch.equal(mb[..., 12:], torch.full_like(mb[..., 12:], 255))     # other models' windows untouched
    win = mb[..., 5:12]
    assert int(win[:11].sum()) == 0 and int(win[111:].sum()) == 0 and int(win[:, :13].sum()) == 0 and int(win[:, 163:].sum()) == 0      # zeros outside the crop box
    sure = (la.abs() > 1e-4).permute(1, 2, 0)
    assert torch.equal(win[11:111, 13:163][sure], ma[11:111, 13:163, 5:12][sure])
    assert torch.equal(win[11:111, 13:163], (lb > 0).permute(1, 2, 0).to(torch.uint8))
    assert torch.equal(outs["0"][2], outs["1"][2])


@pytest.mark.parametrize("precision", ["fp16", "bf16"])
def test_conv_tc2_two_tiles_per_weight_block(precision, monkeypatch):
    """conv_tc2_kernel (two 128-pixel tiles share every weight block; BN = 128 double-buffered and BN = 256 single-buffered TMEM,
    stride 1 and stride 2) against conv_tc_kernel on the same layers: the MMAs accumulate the same k-chunks in the same order, only
    the fp32 partial sums of the InstanceNorm statistics are grouped differently, so outputs agree to one storage ulp; and against
    the fp32 oracle layer by layer."""
    dt = torch.bfloat16 if precision == "bf16" else torch.float16
    outs = {}
    layers = ["enc2.0", "enc2.1", "enc3.0", "enc3.1", "dec0.0", "dec0.1", "dec1.0"]
    for tc2 in ("0", "1"):
        monkeypatch.setenv("TS2D_TC2", tc2)
        monkeypatch.setenv("TS2D_TC2_MIN_UNITS", "1")
        monkeypatch.setenv("TS2D_CONV_HALO", "0")          # every 3x3 layer goes through conv_tc / conv_tc2
        p, sd, spec = _make([32, 64, 128, 256], (128, 160), heads=5, precision=precision, seed=23)
        p._ensure_batch(6)
        x, xd = _tiles(6, 2, (128, 160), 7, dt)
        x = xd.float().cpu().permute(0, 3, 1, 2).contiguous()
        coll = {}
        with torch.no_grad():
            nnunet_net.unet_forward(x, sd, spec, collect=coll)
        outs[tc2] = {n: p.forward_debug(xd, n, tuple(coll[n].shape[1:])).cpu() for n in layers}
        outs[tc2]["logits"] = p.predict_tile_logits(xd).cpu()
        if tc2 == "1":
            errs = {n: float((outs[tc2][n] - coll[n]).abs().max()) for n in layers}
            print(precision, "conv_tc2 vs fp32 oracle:", {k: round(v, 4) for k, v in errs.items()})
            assert max(errs.values()) <= (0.2 if precision == "bf16" else 0.03), errs
        p.close()
    ulp = 2.0 ** -5 if precision == "bf16" else 2.0 ** -8            # one storage ulp for |x| in [4, 8)
    for n in layers + ["logits"]:
        d = (outs["0"][n] - outs["1"][n]).abs()
        assert float(d.max()) <= 2 * ulp, (n, float(d.max()))
        # statistics grouped differently: rare 1-ulp flips in the first conv_tc2 layer (same input bits); every flipped input value
        # then moves all the fp32 sums it feeds, so the share of bit-equal outputs decays layer by layer (measured: 0.99 -> 0.87 -> 0.65; fp32 logits 0.37): only the 2-ulp bound is asserted there
        if n == layers[0]:
            assert float((d == 0).float().mean()) >= 0.95, (n, float((d == 0).float().mean()))
