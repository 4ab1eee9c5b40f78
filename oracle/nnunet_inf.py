"""Oracle (test infrastructure): nnU-Net sliding-window inference and export.

Reference call sites: ts2d/core/inference/prediction_worker.py:206-209
(``predict_logits_from_preprocessed_data``) and :215-221
(``export_prediction_from_logits``); predictor kwargs at
ts2d/core/inference/nnu.py:151-166 (``tile_step_size``, ``use_mirroring``).
Restated from the published nnU-Net v2.6 sources (nnunetv2ml 2.6.2 is not vendored):
  inference/predict_from_raw_data.py    predict_sliding_window_return_logits,
                                        _internal_maybe_mirror_and_predict
  inference/sliding_window_prediction.py compute_gaussian, compute_steps_for_sliding_window
  acvl_utils pad_nd_image
  inference/export_prediction.py        convert_predicted_logits_to_segmentation_with_correct_shape
Two modes: ``faithful`` keeps upstream's fp16 accumulators / fp16 Gaussian;
``ideal`` accumulates in fp32 (what the CUDA path does).
"""
from __future__ import annotations

import itertools
from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch
from scipy.ndimage import gaussian_filter

from .nnunet_pre import skimage_resize_2d


def compute_steps_for_sliding_window(image_size: Sequence[int], tile_size: Sequence[int], tile_step_size: float) -> List[List[int]]:
    assert all(i >= j for i, j in zip(image_size, tile_size)), "image size must be as large or larger than patch_size"
    assert 0 < tile_step_size <= 1
    target = [i * tile_step_size for i in tile_size]
    num_steps = [int(np.ceil((i - k) / j)) + 1 for i, j, k in zip(image_size, target, tile_size)]
    steps = []
    for dim in range(len(tile_size)):
        max_step = image_size[dim] - tile_size[dim]
        actual = max_step / (num_steps[dim] - 1) if num_steps[dim] > 1 else 99999999999
        steps.append([int(np.round(actual * i)) for i in range(num_steps[dim])])
    return steps


def compute_gaussian(tile_size: Sequence[int], sigma_scale: float = 1.0 / 8, value_scaling_factor: float = 10.0,
                     dtype=torch.float16) -> torch.Tensor:
    tmp = np.zeros(tile_size)
    center = [i // 2 for i in tile_size]
    sigmas = [i * sigma_scale for i in tile_size]
    tmp[tuple(center)] = 1
    g = gaussian_filter(tmp, sigmas, 0, mode="constant", cval=0)
    g = torch.from_numpy(g)
    g /= (torch.max(g) / value_scaling_factor)
    g = g.to(dtype=dtype)
    mask = g == 0
    g[mask] = torch.min(g[~mask])
    return g


def pad_to_patch(data: torch.Tensor, patch: Sequence[int]):
    """pad_nd_image(data, patch, 'constant', {'value': 0}, True): centred zero padding of
    the trailing len(patch) axes up to at least the patch size. returns (padded, slicer)."""
    shape = list(data.shape)
    nd = len(patch)
    pads, slicer = [], [slice(None)] * (len(shape) - nd)
    for a in range(nd):
        old = shape[len(shape) - nd + a]
        new = max(old, int(patch[a]))
        diff = new - old
        below, above = diff // 2, diff // 2 + diff % 2
        pads.append((below, above))
        slicer.append(slice(below, below + old))
    tpad = []
    for below, above in reversed(pads):
        tpad += [below, above]
    padded = torch.nn.functional.pad(data, tpad, mode="constant", value=0.0) if any(p != (0, 0) for p in pads) else data
    return padded, tuple(slicer), pads


def mirror_axes_combinations(mirror_axes: Optional[Sequence[int]]) -> List[Tuple[int, ...]]:
    if mirror_axes is None or len(mirror_axes) == 0:
        return []
    axes = [m + 2 for m in mirror_axes]          # on (N, C, H, W)
    return [c for i in range(len(axes)) for c in itertools.combinations(axes, i + 1)]


def maybe_mirror_and_predict(net: Callable[[torch.Tensor], torch.Tensor], x: torch.Tensor,
                             mirror_axes: Optional[Sequence[int]]) -> torch.Tensor:
    pred = net(x)
    combos = mirror_axes_combinations(mirror_axes)
    for axes in combos:
        pred = pred + torch.flip(net(torch.flip(x, axes)), axes)
    if combos:
        pred = pred / (len(combos) + 1)
    return pred


def predict_sliding_window(net: Callable[[torch.Tensor], torch.Tensor], data: torch.Tensor, patch: Sequence[int],
                           num_heads: int, tile_step_size: float = 0.5, mirror_axes: Optional[Sequence[int]] = (0, 1),
                           use_gaussian: bool = True, mode: str = "ideal") -> torch.Tensor:
    """data: (C, 1, H, W) float32 (2-D configuration: the first spatial axis is the
    slice axis and is iterated).  returns logits (heads, 1, H, W) float32 (``ideal``)
    or float16 values widened to float32 (``faithful``)."""
    assert data.ndim == 4 and mode in ("ideal", "faithful")
    acc_dtype = torch.float32 if mode == "ideal" else torch.float16
    padded, slicer, _ = pad_to_patch(data, patch)
    image_size = padded.shape[1:]
    steps = compute_steps_for_sliding_window(image_size[1:], patch, tile_step_size)
    acc = torch.zeros((num_heads, *image_size), dtype=acc_dtype)
    npred = torch.zeros(tuple(image_size), dtype=acc_dtype)
    if use_gaussian:
        g16 = compute_gaussian(tuple(patch), 1.0 / 8, 10.0, dtype=torch.float16)
        g = g16 if mode == "faithful" else g16.to(torch.float32)
    for d in range(image_size[0]):
        for sy in steps[0]:
            for sx in steps[1]:
                tile = padded[:, d, sy:sy + patch[0], sx:sx + patch[1]][None]
                pred = maybe_mirror_and_predict(net, tile, mirror_axes)[0]
                if use_gaussian:
                    pred = pred * g          # fp32 * fp16 -> fp32, like upstream on CPU
                acc[:, d, sy:sy + patch[0], sx:sx + patch[1]] += pred.to(acc_dtype)
                npred[d, sy:sy + patch[0], sx:sx + patch[1]] += (g if use_gaussian else 1)
    acc = acc / npred
    if torch.any(torch.isinf(acc)):
        raise RuntimeError("Encountered inf in predicted array. Aborting...")
    return acc[(slice(None), *slicer[1:])].to(torch.float32)


def resample_logits(logits: np.ndarray, new_shape: Sequence[int]) -> np.ndarray:
    """resampling_fn_probabilities: order-1 skimage resize per (head, slice); no-op on equal shapes."""
    if tuple(logits.shape[1:]) == tuple(int(v) for v in new_shape):
        return logits
    out = np.zeros((logits.shape[0], *[int(v) for v in new_shape]), dtype=logits.dtype)
    work = logits.astype(float, copy=False)
    for c in range(logits.shape[0]):
        for s in range(logits.shape[1]):
            out[c, s] = skimage_resize_2d(work[c, s], new_shape[1:], order=1).astype(logits.dtype)
    return out


def export_multilabel(logits: np.ndarray, props: dict, mode: str = "ideal") -> np.ndarray:
    """logits (heads, 1, H, W) -> uint8 (1, H0, W0, heads): resample to the cropped shape,
    sigmoid > 0.5, one binary channel per foreground label (multilabel fork), paste
    into zeros(shape_before_cropping) at the crop bbox."""
    lg = resample_logits(np.asarray(logits, dtype=np.float32), props["shape_after_cropping_and_before_resampling"])
    if mode == "faithful":
        seg = (torch.sigmoid(torch.from_numpy(lg).float()) > 0.5).numpy()
    else:
        seg = lg > 0
    full = np.zeros((lg.shape[0], *props["shape_before_cropping"]), dtype=np.uint8)
    bbox = props["bbox_used_for_cropping"]
    sl = (slice(None),) + tuple(slice(lo, hi) for lo, hi in bbox)
    full[sl] = seg.astype(np.uint8)
    return np.ascontiguousarray(np.moveaxis(full, 0, -1))
