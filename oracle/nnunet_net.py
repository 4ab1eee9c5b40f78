"""Oracle (test infrastructure): ``PlainConvUNet`` forward on torch-CPU.

Restates dynamic_network_architectures (third party, pulled in by nnunetv2ml;
not vendored under /root/reference):
  architectures/unet.py                PlainConvUNet
  building_blocks/plain_conv_encoder.py PlainConvEncoder (stages of StackedConvBlocks)
  building_blocks/unet_decoder.py       UNetDecoder (transpconv -> cat(up, skip) -> convs; seg head 1x1)
  building_blocks/simple_conv_blocks.py ConvDropoutNormReLU = Conv2d -> InstanceNorm2d -> LeakyReLU
The network is built at reference ts2d/core/inference/nnu.py:151-166 through
``nnUNetPredictor.initialize_from_trained_model_folder``.

State-dict names are the upstream ones, so a real ``checkpoint_final.pth`` loads.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import torch
import torch.nn.functional as F


def _pair(v):
    return (int(v), int(v)) if not isinstance(v, (list, tuple)) else (int(v[0]), int(v[1]))


class ArchSpec:
    """The ``arch_kwargs`` of a 2-D PlainConvUNet plan."""

    def __init__(self, arch_kwargs: dict, in_channels: int, num_heads: int):
        k = arch_kwargs
        self.n_stages = int(k["n_stages"])
        self.features = [int(f) for f in k["features_per_stage"]]
        self.kernel_sizes = [_pair(s) for s in k["kernel_sizes"]]
        self.strides = [_pair(s) for s in k["strides"]]
        self.n_conv_enc = [int(n) for n in k["n_conv_per_stage"]]
        self.n_conv_dec = [int(n) for n in k["n_conv_per_stage_decoder"]]
        self.conv_bias = bool(k.get("conv_bias", True))
        nk = k.get("norm_op_kwargs") or {}
        self.eps = float(nk.get("eps", 1e-5))
        self.affine = bool(nk.get("affine", True))
        nl = k.get("nonlin_kwargs") or {}
        self.negative_slope = float(nl.get("negative_slope", 0.01))
        self.in_channels = int(in_channels)
        self.num_heads = int(num_heads)
        assert len(self.features) == self.n_stages == len(self.strides) == len(self.kernel_sizes)


def _block(x, sd, prefix, stride, pad, spec: ArchSpec):
    x = F.conv2d(x, sd[prefix + ".conv.weight"], sd.get(prefix + ".conv.bias"), stride=stride, padding=pad)
    x = F.instance_norm(x, weight=sd.get(prefix + ".norm.weight"), bias=sd.get(prefix + ".norm.bias"), eps=spec.eps)
    return F.leaky_relu(x, spec.negative_slope)


def unet_forward(x: torch.Tensor, sd: Dict[str, torch.Tensor], spec: ArchSpec, collect=None) -> torch.Tensor:
    """x: (N, C, H, W) float32 -> logits (N, heads, H, W). ``collect`` (a dict) receives
    intermediate activations by layer name for layer-wise parity tests."""
    skips: List[torch.Tensor] = []
    for s in range(spec.n_stages):
        kh, kw = spec.kernel_sizes[s]
        pad = ((kh - 1) // 2, (kw - 1) // 2)
        for i in range(spec.n_conv_enc[s]):
            stride = spec.strides[s] if i == 0 else (1, 1)
            x = _block(x, sd, f"encoder.stages.{s}.0.convs.{i}", stride, pad, spec)
            if collect is not None:
                collect[f"enc{s}.{i}"] = x
        skips.append(x)
    lres = skips[-1]
    n_dec = spec.n_stages - 1
    for d in range(n_dec):
        stride = spec.strides[-(d + 1)]
        up = F.conv_transpose2d(lres, sd[f"decoder.transpconvs.{d}.weight"],
                                sd.get(f"decoder.transpconvs.{d}.bias"), stride=stride)
        x = torch.cat((up, skips[-(d + 2)]), dim=1)
        kh, kw = spec.kernel_sizes[-(d + 2)]
        pad = ((kh - 1) // 2, (kw - 1) // 2)
        for i in range(spec.n_conv_dec[d]):
            x = _block(x, sd, f"decoder.stages.{d}.convs.{i}", (1, 1), pad, spec)
            if collect is not None:
                collect[f"dec{d}.{i}"] = x
        lres = x
    head = n_dec - 1
    return F.conv2d(lres, sd[f"decoder.seg_layers.{head}.weight"], sd.get(f"decoder.seg_layers.{head}.bias"))


def conv_flops(spec: ArchSpec, h: int, w: int) -> float:
    """multiply-add = 2 flops; conv 2*H*W*k*k*Cin*Cout, transposed conv 2*Hout*Wout*Cin*Cout."""
    fl = 0.0
    cin = spec.in_channels
    hh, ww = h, w
    sizes = []
    for s in range(spec.n_stages):
        kh, kw = spec.kernel_sizes[s]
        for i in range(spec.n_conv_enc[s]):
            if i == 0:
                hh, ww = -(-hh // spec.strides[s][0]), -(-ww // spec.strides[s][1])
            fl += 2.0 * hh * ww * kh * kw * cin * spec.features[s]
            cin = spec.features[s]
        sizes.append((hh, ww))
    for d in range(spec.n_stages - 1):
        cskip = spec.features[-(d + 2)]
        hh, ww = sizes[-(d + 2)]
        fl += 2.0 * hh * ww * cin * cskip
        kh, kw = spec.kernel_sizes[-(d + 2)]
        c_in = 2 * cskip
        for i in range(spec.n_conv_dec[d]):
            fl += 2.0 * hh * ww * kh * kw * c_in * cskip
            c_in = cskip
        cin = cskip
    fl += 2.0 * h * w * cin * spec.num_heads
    return fl
