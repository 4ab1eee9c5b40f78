"""Image helpers on the hot path, backed by the CUDA kernels.

Host-side mirror of reference ts2d/core/util/image.py: ``axis_name_to_index`` (:16-29),
``reorient_image`` (:32-43), ``project`` (:46-101), ``read_image`` (:225-238), ``reduce_dimensions``
(:241-258), ``get_actual_dimension`` (:483-487), ``combine_segmentations`` (:490-510) and
``split_channels`` (:512-520).  Same names, argument meaning and errors; the arithmetic runs in
libts2d_b200.so.  ``project_channels`` is the fused form ``TS2D.predict`` uses: the volume crosses PCIe
once, DICOMOrient is folded into the kernel's addressing and every channel comes out of one pass.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ... import _lib as L
from ... import ops
from ...image import Image, from_sitk
from ...nrrd import read_nrrd
from .meta import copy_image_meta, get_annotation_labels, get_label_mask, set_annotation_meta


def axis_name_to_index(name: str) -> int:
    """Maps an axis name to its axis index in RAI coordinates"""
    return {"a": 2, "ax": 2, "axial": 2, "s": 0, "sag": 0, "sagittal": 0, "c": 1, "cor": 1, "coronal": 1}[name.lower()]


_LETTERS = {"L": (0, 1), "R": (0, -1), "P": (1, 1), "A": (1, -1), "S": (2, 1), "I": (2, -1)}


def orientation_remap(direction, target: str = "RAI") -> Tuple[Tuple[int, ...], Tuple[bool, ...]]:
    """What ``sitk.DICOMOrient(img, target)`` does to the axes: returns (perm, flip) with output image axis k
    = input image axis perm[k], reversed when flip[k].  Follows ITK's DICOMOrientation::DirectionCosinesToOrientation:
    the largest |direction cosine| claims its (anatomical axis, image axis) pair, greedily; DICOMOrient letters
    name where an axis points TO in LPS ('L','P','S' = +x,+y,+z)."""
    d = np.asarray(direction, dtype=np.float64).reshape(3, 3)
    order = sorted(((abs(d[c, r]), c * 3 + r, c, r) for c in range(3) for r in range(3)), reverse=True)
    axis_phys: Dict[int, Tuple[int, int]] = {}
    used_c, used_r = set(), set()
    for _, _, c, r in order:
        if c in used_c or r in used_r:
            continue
        used_c.add(c)
        used_r.add(r)
        axis_phys[r] = (c, 1 if d[c, r] > 0 else -1)
    perm, flip = [], []
    for ch in target.upper():
        pax, sign = _LETTERS[ch]
        r = [a for a, (c, _) in axis_phys.items() if c == pax][0]
        perm.append(r)
        flip.append(axis_phys[r][1] != sign)
    return tuple(perm), tuple(flip)


def _reoriented_geometry(img: Image, perm, flip):
    d = img.direction_matrix()
    sp = np.asarray(img.GetSpacing(), dtype=np.float64)
    size = img.GetSize()
    org = np.asarray(img.GetOrigin(), dtype=np.float64)
    nd = np.zeros((3, 3))
    for k in range(3):
        col = d[:, perm[k]]
        if flip[k]:
            org = org + col * sp[perm[k]] * (size[perm[k]] - 1)
            col = -col
        nd[:, k] = col
    return [size[perm[k]] for k in range(3)], [float(sp[perm[k]]) for k in range(3)], [float(v) for v in org], nd


def _to_device(arr: np.ndarray) -> torch.Tensor:
    L.load(check_device=True)
    if isinstance(arr, torch.Tensor):
        return arr.cuda(non_blocking=True).contiguous()
    a = np.ascontiguousarray(arr)
    view = a.view(np.int16) if a.dtype == np.uint16 else a
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore", UserWarning)          # read-only arrays (np.frombuffer of a decoded file): the upload never writes
        src = torch.from_numpy(view)
    if view.nbytes >= _STAGED_UPLOAD_MIN_BYTES and not src.is_pinned():
        # pageable memory (what read_image returns): staged through the library's page-locked ring by several CPU threads
        lib = L.load()
        dst = torch.empty(src.shape, dtype=src.dtype, device="cuda")
        L.check(lib.ts2d_upload_pageable(dst.data_ptr(), view.ctypes.data, view.nbytes, _upload_threads(), L.stream_ptr()), "ts2d_upload_pageable")
    else:
        dst = src.cuda(non_blocking=True)                     # asynchronous when the array lives in pinned memory
    return dst.view(torch.uint16) if a.dtype == np.uint16 else dst


_STAGED_UPLOAD_MIN_BYTES = 16 << 20


def _upload_threads() -> int:
    """CPU threads of the staged upload: $TS2D_UPLOAD_THREADS, else 8 shared between the ranks on this host"""
    import os
    env = os.environ.get("TS2D_UPLOAD_THREADS")
    if env:
        return max(1, int(env))
    world = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1"))))
    return max(1, min(8, (os.cpu_count() or 8) // world))


def reorient_image(img: Image, orient: str = "RAI") -> Image:
    """DICOMOrient: pure index remap (no interpolation); 2-D images are returned unchanged."""
    if img.GetDimension() > 2:
        perm, flip = orientation_remap(img.direction_matrix(), orient)
        np_axes = [2 - perm[2 - a] for a in range(3)]
        arr = np.transpose(img.array, np_axes + ([3] if img.is_vector else []))
        for k in range(3):
            if flip[k]:
                arr = np.flip(arr, axis=2 - k)
        _, sp, org, nd = _reoriented_geometry(img, perm, flip)
        out = Image(np.ascontiguousarray(arr), sp, org, nd.reshape(-1), is_vector=img.is_vector)
        return copy_image_meta(out, img)
    return img


def _projected_image(arr2d: np.ndarray, size, spacing, origin, direction, axis: int) -> Image:
    """geometry of ITK's ProjectionImageFilter output + the reference's origin reset (image.py:97-100)."""
    spacing = list(spacing)
    spacing[axis] = spacing[axis] * size[axis]
    shape = [size[2], size[1], size[0]]
    shape[2 - axis] = 1
    return Image(arr2d.reshape(shape), spacing, origin, np.asarray(direction).reshape(-1))


def project_channels(img: Image, modes: Sequence[str], axis="coronal", orient: str = "RAI", keep_device: bool = False,
                     volume_dev: Optional[torch.Tensor] = None):
    """reorient_image(img, orient) + project(mode, axis) + Cast(float32) for every mode in ONE pass.
    returns ({mode: Image (X,1,Z) float32}, device tensor float32 [len(modes)][H][W] if keep_device)."""
    ax = axis_name_to_index(axis) if isinstance(axis, str) else list(range(3))[axis]
    if img.GetDimension() != 3 or img.is_vector:
        raise RuntimeError("project_channels expects a scalar 3-D image")
    perm, flip = orientation_remap(img.direction_matrix(), orient)
    size, sp, org, nd = _reoriented_geometry(img, perm, flip)
    # output image axes kept: the two that are not `ax`; rows = the higher image axis (z for coronal), cols = the lower
    kept = [k for k in range(3) if k != ax]
    w_ax, h_ax = kept[0], kept[1]
    mem = lambda image_axis: 2 - image_axis            # memory dim of an input image axis
    vol = volume_dev if volume_dev is not None else _to_device(img.array)
    out = ops.project(vol, modes, reduce_dim=mem(perm[ax]), h_dim=mem(perm[h_ax]), flip_h=flip[h_ax], flip_w=flip[w_ax])
    res = {}
    host = out.cpu().numpy()
    for i, m in enumerate(modes):
        res[m] = _projected_image(host[i], size, sp, org, nd, ax)
    return (res, out) if keep_device else res


def project(img: Image, mode="max", axis=-1) -> Image:
    """Projects an image along ``axis`` ('max'/'mip', 'mean'/'avg', 'min'); other reference modes
    (first/median/std/multiclass/slice) are not on the TS2D.predict path and are not implemented."""
    ax = axis_name_to_index(axis) if isinstance(axis, str) else list(range(img.GetDimension()))[axis]
    mode = str(mode).lower().strip()
    if mode.split(":")[0] not in ("max", "mip", "mean", "avg", "min"):
        if mode.split(":")[0] in ("first", "median", "std", "depth", "multiclass", "slice", "xr"):
            raise NotImplementedError("projection mode '{}' is not part of the accelerated hot path".format(mode))
        raise RuntimeError("Unsupported filter mode: {}".format(mode))
    if img.GetDimension() != 3 or img.is_vector:
        raise RuntimeError("project expects a scalar 3-D image")
    vol = _to_device(img.array)
    kept = [k for k in range(3) if k != ax]
    out = ops.project(vol, [mode], reduce_dim=2 - ax, h_dim=2 - kept[1], flip_h=False, flip_w=False)
    arr = out[0].cpu().numpy()
    if mode in ("max", "mip", "min"):
        arr = arr.astype(img.array.dtype)          # ITK keeps the pixel type for max/min; mean is real-valued
    return _projected_image(arr, img.GetSize(), img.GetSpacing(), img.GetOrigin(), img.GetDirection(), ax)


def read_image(fp: str) -> Image:
    low = fp.lower()
    if low.endswith(".nrrd"):
        return read_nrrd(fp)
    if low.endswith((".nii", ".nii.gz", ".mha", ".mhd")):
        from ...formats import read_any
        return read_any(fp)
    raise RuntimeError("unsupported image format for {} (supported: nrrd, nii, nii.gz, mha, mhd)".format(fp))


def reduce_dimensions(img: Image, strategy=None, min=None) -> Image:
    """Collapses dimensions of size 1 (sitk.Extract with DIRECTIONCOLLAPSETOGUESS)."""
    size = img.GetSize()
    keep = [i for i, s in enumerate(size) if s > 1]
    if min:
        refill = min - len(keep)
        for idx in range(len(size) - 1, -1, -1):
            if refill <= 0:
                break
            if idx not in keep:
                keep.append(idx)
                refill -= 1
        keep = sorted(keep)
    if len(keep) == len(size):
        return img
    shape = [size[i] for i in reversed(keep)] + ([img.GetNumberOfComponentsPerPixel()] if img.is_vector else [])
    d = img.direction_matrix()
    sub = d[np.ix_(keep, keep)]
    if abs(abs(np.linalg.det(sub)) - 1.0) > 1e-6:      # ITK: fall back to identity if the submatrix is not a rotation
        sub = np.eye(len(keep))
    out = Image(img.array.reshape(shape), [img.GetSpacing()[i] for i in keep], [img.GetOrigin()[i] for i in keep], sub.reshape(-1),
                is_vector=img.is_vector)
    return copy_image_meta(out, img)


def get_actual_dimension(img: Image) -> int:
    return sum(s > 1 for s in img.GetSize())


def split_channels(img: Image) -> List[Image]:
    n = img.GetNumberOfComponentsPerPixel()
    if n == 1 and not img.is_vector:
        return [img]
    return [Image(np.ascontiguousarray(img.array[..., i]), img.GetSpacing(), img.GetOrigin(), img.GetDirection()) for i in range(n)]


def compose(chs: Sequence[Image]) -> Image:
    """sitk.Compose: per-pixel vector of the given scalar images."""
    arr = np.stack([c.array for c in chs], axis=-1)
    return Image(arr, chs[0].GetSpacing(), chs[0].GetOrigin(), chs[0].GetDirection(), is_vector=True)


def merge_order(segs: Sequence[Image], presents: Optional[Sequence[Optional[np.ndarray]]] = None):
    """channel order of ``combine_segmentations``: per model, labels present in the image first (ascending),
    then the absent ones in metadata order (reference meta.py:280-286). returns ([(seg index, channel)], names, colors)."""
    order, names, colors = [], dict(), dict()
    for si, seg in enumerate(segs):
        pres = presents[si] if presents is not None else None
        for name, info in get_annotation_labels(seg, present=pres).items():
            idx = len(order)
            names[idx + 1] = name
            if info.get("color") is not None:
                colors[name] = info["color"]
            order.append((si, info["value"] - 1))
    return order, names, colors


def combine_segmentations(segs: List[Image]) -> Image:
    """combines multiple (multilabel) segmentations into a single one (host path, any Images)."""
    order, names, colors = merge_order(segs)
    chans = [get_label_mask(segs[si], ch + 1).array for si, ch in order]
    res = Image(np.stack(chans, axis=-1), segs[0].GetSpacing(), segs[0].GetOrigin(), segs[0].GetDirection(), is_vector=True)
    set_annotation_meta(res, names=names, colors=colors)
    return res
