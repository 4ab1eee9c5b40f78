"""Oracle (test infrastructure): nnU-Net ``DefaultPreprocessor.run_case_npy``.

The reference calls it at ts2d/core/inference/prediction_worker.py:194-199.  The
implementation lives in the third-party package nnunetv2ml==2.6.2 (fork of
nnU-Net v2.6; reference pyproject.toml:25), which is not vendored; this file
restates the published upstream algorithm:
  preprocessing/preprocessors/default_preprocessor.py   run_case_npy
  preprocessing/cropping/cropping.py                     crop_to_nonzero
  preprocessing/normalization/default_normalization_schemes.py
  preprocessing/resampling/default_resampling.py         resample_data_or_seg_to_shape
  (skimage.transform.resize == scipy.ndimage.zoom(grid_mode=True) + clip)
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
from scipy import ndimage as ndi

ANISO_THRESHOLD = 3.0


# --- cropping ---------------------------------------------------------------------------
def create_nonzero_mask(data: np.ndarray) -> np.ndarray:
    """OR over channels of (data != 0), then binary_fill_holes. data: (C, D, H, W)."""
    m = data[0] != 0
    for c in range(1, data.shape[0]):
        m |= data[c] != 0
    return ndi.binary_fill_holes(m)


def bbox_from_mask(mask: np.ndarray) -> List[List[int]]:
    """acvl_utils get_bbox_from_mask: per axis first/last index with any True, half-open;
    an empty mask keeps the full extent."""
    bbox = []
    for ax in range(mask.ndim):
        other = tuple(a for a in range(mask.ndim) if a != ax)
        nz = np.nonzero(mask.any(axis=other))[0]
        if nz.size == 0:
            bbox.append([0, int(mask.shape[ax])])
        else:
            bbox.append([int(nz[0]), int(nz[-1]) + 1])
    return bbox


def crop_to_nonzero(data: np.ndarray) -> Tuple[np.ndarray, List[List[int]]]:
    mask = create_nonzero_mask(data)
    bbox = bbox_from_mask(mask)
    sl = (slice(None),) + tuple(slice(lo, hi) for lo, hi in bbox)
    return data[sl], bbox


# --- normalisation ------------------------------------------------------------------------
def normalize_channel(img: np.ndarray, scheme: str, props: Optional[Dict[str, float]] = None) -> np.ndarray:
    """one channel, float32 in / float32 out; ``scheme`` is the class name stored in
    plans.json ``normalization_schemes``."""
    img = img.astype(np.float32, copy=True)
    if scheme == "ZScoreNormalization":
        mean = img.mean()
        std = img.std()
        img -= mean
        img /= max(std, 1e-8)
    elif scheme == "CTNormalization":
        assert props is not None, "CTNormalization needs foreground intensity properties"
        np.clip(img, props["percentile_00_5"], props["percentile_99_5"], out=img)
        img -= np.float32(props["mean"])
        img /= np.float32(max(props["std"], 1e-8))
    elif scheme == "NoNormalization":
        pass
    elif scheme == "RescaleTo01Normalization":
        img -= img.min()
        img /= np.clip(img.max(), a_min=1e-8, a_max=None)
    elif scheme == "RGBTo01Normalization":
        img /= np.float32(255.0)
    else:
        raise RuntimeError("unknown normalization scheme: {}".format(scheme))
    return img


# --- resampling -----------------------------------------------------------------------------
def compute_new_shape(old_shape: Sequence[int], old_spacing: Sequence[float], new_spacing: Sequence[float]) -> List[int]:
    assert len(old_spacing) == len(old_shape) == len(new_spacing)
    return [int(round(i / j * k)) for i, j, k in zip(old_spacing, new_spacing, old_shape)]


def skimage_resize_2d(img: np.ndarray, new_hw: Sequence[int], order: int) -> np.ndarray:
    """skimage.transform.resize(img, new_hw, order, mode='edge', anti_aliasing=False)
    for a float64 2-D image: ndi.zoom(grid_mode=True, mode='nearest') then clip to the
    input range (skimage ``_clip_warp_output`` with clip=True)."""
    img = np.asarray(img, dtype=np.float64)
    zoom = [n / o for n, o in zip(new_hw, img.shape)]
    out = ndi.zoom(img, zoom, order=order, mode="nearest", cval=0.0, grid_mode=True)
    assert tuple(out.shape) == tuple(int(v) for v in new_hw), (out.shape, new_hw)
    np.clip(out, img.min(), img.max(), out=out)
    return out


def _do_separate_z(spacing: Sequence[float]) -> bool:
    return (max(spacing) / min(spacing)) > ANISO_THRESHOLD


def resample_data_to_shape(data: np.ndarray, new_shape: Sequence[int], current_spacing: Sequence[float],
                           new_spacing: Sequence[float], order: int = 3, order_z: int = 0) -> np.ndarray:
    """resample_data_or_seg_to_shape(is_seg=False, force_separate_z=None) restricted to
    what a 2-D configuration produces: data (C, 1, H, W) whose dummy axis has spacing
    999*max (SimpleITKIO), hence the separate-z branch with axis 0 and an unchanged
    number of slices -> per (channel, slice) 2-D resize."""
    assert data.ndim == 4
    shape = np.array(data.shape[1:])
    new_shape = np.array([int(v) for v in new_shape])
    if not np.any(shape != new_shape):
        return data
    dtype_data = data.dtype
    sep = _do_separate_z(current_spacing) or _do_separate_z(new_spacing)
    if not (sep and int(np.argmax(current_spacing if _do_separate_z(current_spacing) else new_spacing)) == 0
            and shape[0] == new_shape[0]):
        raise NotImplementedError("only the 2-D configuration path (separate z, axis 0) is restated")
    work = data.astype(float, copy=False)
    out = np.zeros((data.shape[0], *new_shape), dtype=dtype_data)
    for c in range(data.shape[0]):
        for s in range(shape[0]):
            out[c, s] = skimage_resize_2d(work[c, s], new_shape[1:], order).astype(dtype_data)
    return out


# --- run_case ---------------------------------------------------------------------------------
def run_case_npy(data: np.ndarray, spacing: Sequence[float], plan_spacing: Sequence[float],
                 schemes: Sequence[str], intensity_props: Optional[Dict[str, Dict[str, float]]] = None,
                 use_mask_for_norm: Optional[Sequence[bool]] = None):
    """data: (C, 1, H, W) any dtype; spacing: (999*max, sy, sx) as SimpleITKIO reports
    for 2-D images.  Order of operations is upstream's: crop -> normalise -> resample."""
    if use_mask_for_norm is not None and any(use_mask_for_norm):
        raise NotImplementedError("use_mask_for_norm=True is not restated")
    data = data.astype(np.float32)
    props = {"spacing": [float(s) for s in spacing], "shape_before_cropping": tuple(data.shape[1:])}
    data, bbox = crop_to_nonzero(data)
    props["bbox_used_for_cropping"] = bbox
    props["shape_after_cropping_and_before_resampling"] = tuple(data.shape[1:])
    target = list(plan_spacing)
    if len(target) < data.ndim - 1:
        target = [spacing[0]] + target
    new_shape = compute_new_shape(data.shape[1:], spacing, target)
    chans = []
    for c in range(data.shape[0]):
        p = None if intensity_props is None else intensity_props.get(str(c))
        chans.append(normalize_channel(data[c], schemes[c], p))
    data = np.stack(chans, axis=0)
    data = resample_data_to_shape(data, new_shape, spacing, target, order=3, order_z=0)
    return np.ascontiguousarray(data, dtype=np.float32), props


def sitkio_spacing_2d(spacing_xy: Sequence[float]) -> List[float]:
    """SimpleITKIO.read_images for a 2-D image: (999*max, sy, sx)."""
    sx, sy = float(spacing_xy[0]), float(spacing_xy[1])
    return [max(sx, sy) * 999.0, sy, sx]
