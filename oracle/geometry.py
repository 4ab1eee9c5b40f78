"""Oracle (test infrastructure): reorientation and projection, restated on numpy.

Follows reference ts2d/core/util/image.py:32-43 (``reorient_image`` ->
``sitk.DICOMOrient(img, 'RAI')``), :46-101 (``project`` -> ITK
Maximum/MeanProjectionImageFilter + origin reset), ts2d/tool.py:152-160,181-185
(per-channel projection, ``sitk.Cast`` to float32, ``sitk.Compose`` in ascending
channel index) and image.py:241-258 (``reduce_dimensions``).

ITK behaviour restated from its published sources (not under /root/reference):
* ``itk::DICOMOrientation::DirectionCosinesToOrientation`` - greedy assignment of
  the largest |direction cosine| to an (anatomical axis, image axis) pair.
* ``itk::ProjectionImageFilter`` with ``MaximumAccumulator`` (running max in the
  pixel type) and ``MeanAccumulator`` (sum in ``NumericTraits::RealType`` =
  double, divided by the line length, written as double).
"""
from __future__ import annotations

from typing import Sequence, Tuple

import numpy as np


def orientation_remap(direction: np.ndarray, target: str = "RAI") -> Tuple[Tuple[int, int, int], Tuple[bool, bool, bool]]:
    """returns (perm, flip): output image axis k is input image axis perm[k],
    traversed backwards when flip[k].

    ``direction`` is the 3x3 matrix whose *columns* are the physical (LPS)
    direction of each image axis.  DICOMOrient letters name the direction an axis
    points *to*: 'L','P','S' are +x,+y,+z in LPS, 'R','A','I' are -x,-y,-z.
    """
    d = np.asarray(direction, dtype=np.float64).reshape(3, 3)
    cand = sorted(((abs(d[c, r]), c * 3 + r, c, r) for c in range(3) for r in range(3)))
    phys_of_axis = [None, None, None]      # image axis -> (physical axis, sign)
    used_c, used_r = set(), set()
    while cand:
        v, _, c, r = cand.pop()            # largest |value|; ties: last inserted (c-major order)
        if c in used_c or r in used_r:
            continue
        used_c.add(c)
        used_r.add(r)
        phys_of_axis[r] = (c, 1 if d[c, r] > 0 else (-1 if d[c, r] < 0 else 0))
    letters = {"L": (0, 1), "R": (0, -1), "P": (1, 1), "A": (1, -1), "S": (2, 1), "I": (2, -1)}
    perm, flip = [], []
    for k, ch in enumerate(target.upper()):
        pax, psign = letters[ch]
        src = [r for r in range(3) if phys_of_axis[r][0] == pax]
        assert len(src) == 1
        r = src[0]
        perm.append(r)
        flip.append(phys_of_axis[r][1] != psign)
    return tuple(perm), tuple(flip)


def reorient_array(vol: np.ndarray, perm: Sequence[int], flip: Sequence[bool]) -> np.ndarray:
    """vol is indexed [z][y][x] (numpy axis = 2 - image axis)."""
    np_axes = [2 - perm[2 - a] for a in range(3)]     # output numpy axis a <- input numpy axis
    out = np.transpose(vol, np_axes)
    for k in range(3):
        if flip[k]:
            out = np.flip(out, axis=2 - k)
    return np.ascontiguousarray(out)


def reorient_geometry(size, spacing, origin, direction, perm, flip):
    """geometry of the image produced by ``reorient_array`` (ITK PermuteAxes + Flip semantics)."""
    d = np.asarray(direction, dtype=np.float64).reshape(3, 3)
    sp = np.asarray(spacing, dtype=np.float64)
    org = np.asarray(origin, dtype=np.float64)
    new_size = [int(size[perm[k]]) for k in range(3)]
    new_sp = [float(sp[perm[k]]) for k in range(3)]
    new_d = np.zeros((3, 3))
    new_org = org.copy()
    for k in range(3):
        col = d[:, perm[k]]
        if flip[k]:
            new_org = new_org + col * sp[perm[k]] * (size[perm[k]] - 1)
            col = -col
        new_d[:, k] = col
    return new_size, new_sp, [float(v) for v in new_org], new_d


def project_max(vol: np.ndarray, np_axis: int = 1) -> np.ndarray:
    """MaximumProjectionImageFilter: same pixel type as the input (image.py:75,90-93)."""
    return vol.max(axis=np_axis)


def project_mean(vol: np.ndarray, np_axis: int = 1) -> np.ndarray:
    """MeanProjectionImageFilter: double sum / N, written as double (image.py:79,90-93).

    For integer input the double sum is exact (|v| < 2^15, N < 2^37), so this equals
    the exact rational rounded once to double.  For float input ITK adds in line
    order; numpy's pairwise float64 sum differs from that by < 1e-13 relative,
    far below the float32 cast that follows (tool.py:184).
    """
    n = vol.shape[np_axis]
    if np.issubdtype(vol.dtype, np.integer):
        s = vol.sum(axis=np_axis, dtype=np.int64).astype(np.float64)
    else:
        s = vol.sum(axis=np_axis, dtype=np.float64)
    return s / float(n)


def project_channels(vol_rai: np.ndarray, channel_names: Sequence[str]) -> np.ndarray:
    """TS2D._predict_model for a 3-D input (tool.py:152-160): one projection per
    channel *name* in ascending channel index, each cast to float32, stacked planar
    -> (C, H=Z, W=X) float32."""
    out = []
    for name in channel_names:
        mode = str(name).lower().strip()
        if mode in ("max", "mip"):
            p = project_max(vol_rai, 1)
        elif mode in ("mean", "avg"):
            p = project_mean(vol_rai, 1)
        elif mode == "min":
            p = vol_rai.min(axis=1)
        else:
            raise RuntimeError("Unsupported filter mode: {}".format(mode))
        out.append(p.astype(np.float32))            # sitk.Cast(res, sitkFloat32), tool.py:184
    return np.stack(out, axis=0)


def projected_geometry(size, spacing, origin, direction, axis: int = 1):
    """Output geometry of ``project`` (image.py:90-100): size 1 and spacing N*s on the
    collapsed axis; ITK centres the origin on that axis, the reference resets it."""
    size = list(size)
    spacing = list(spacing)
    spacing[axis] = spacing[axis] * size[axis]
    size[axis] = 1
    return size, spacing, list(origin), np.asarray(direction, dtype=np.float64).reshape(3, 3)
