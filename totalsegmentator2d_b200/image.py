"""Minimal image value object used instead of ``sitk.Image``.

SimpleITK is not part of this image, and the hot path does not need it: every
ITK call the reference makes on the path (``DICOMOrient``, the projection
filters, ``Cast``, ``Compose``, ``Extract``, ``VectorIndexSelectionCast``;
reference ts2d/core/util/image.py:42,75-93,258 and ts2d/tool.py:160,184) is
either an index remap or a reduction that the CUDA kernels do themselves.

``Image`` mirrors the subset of the ``sitk.Image`` interface that the reference
host code touches (``GetSize``/``GetSpacing``/``GetOrigin``/``GetDirection``/
``GetDimension``/``GetNumberOfComponentsPerPixel`` and the metadata dictionary),
so the host-side mirror of the reference reads like the reference.

Memory convention (same as ``sitk.GetArrayFromImage``): ``array`` is indexed
``[z][y][x]`` (x fastest) with an optional trailing component axis, while
``GetSize()`` reports ``(X, Y, Z)``.
"""
from __future__ import annotations

from typing import Dict, Iterable, Optional, Sequence, Tuple

import numpy as np


class Image:
    __slots__ = ("array", "_ncomp", "_vec", "_spacing", "_origin", "_direction", "_meta")

    def __init__(self, array: np.ndarray, spacing: Optional[Sequence[float]] = None,
                 origin: Optional[Sequence[float]] = None,
                 direction: Optional[Sequence[float]] = None,
                 is_vector: bool = False, meta: Optional[Dict[str, str]] = None):
        array = np.asarray(array)
        self._ncomp = int(array.shape[-1]) if is_vector else 1
        self._vec = bool(is_vector)
        dim = array.ndim - (1 if is_vector else 0)
        if dim < 1:
            raise ValueError("an image needs at least one spatial dimension")
        self.array = array
        self._spacing = tuple(float(s) for s in (spacing if spacing is not None else (1.0,) * dim))
        self._origin = tuple(float(o) for o in (origin if origin is not None else (0.0,) * dim))
        if direction is None:
            direction = np.eye(dim).reshape(-1)
        self._direction = tuple(float(d) for d in np.asarray(direction, dtype=np.float64).reshape(-1))
        if len(self._spacing) != dim or len(self._origin) != dim or len(self._direction) != dim * dim:
            raise ValueError("geometry does not match image dimension {}".format(dim))
        self._meta: Dict[str, str] = dict(meta) if meta else {}

    # --- sitk.Image-like geometry accessors ---------------------------------
    def GetDimension(self) -> int:
        return self.array.ndim - (1 if self.is_vector else 0)

    def GetSize(self) -> Tuple[int, ...]:
        shape = self.array.shape[:-1] if self.is_vector else self.array.shape
        return tuple(int(s) for s in reversed(shape))

    def GetSpacing(self) -> Tuple[float, ...]:
        return self._spacing

    def GetOrigin(self) -> Tuple[float, ...]:
        return self._origin

    def GetDirection(self) -> Tuple[float, ...]:
        return self._direction

    def SetSpacing(self, v: Iterable[float]):
        v = tuple(float(x) for x in v)
        assert len(v) == self.GetDimension()
        self._spacing = v

    def SetOrigin(self, v: Iterable[float]):
        v = tuple(float(x) for x in v)
        assert len(v) == self.GetDimension()
        self._origin = v

    def SetDirection(self, v: Iterable[float]):
        v = tuple(float(x) for x in v)
        assert len(v) == self.GetDimension() ** 2
        self._direction = v

    def GetNumberOfComponentsPerPixel(self) -> int:
        return self._ncomp

    @property
    def is_vector(self) -> bool:
        return self._vec

    def GetPixelIDTypeAsString(self) -> str:
        return ("vector of " if self._ncomp > 1 else "") + str(self.array.dtype)

    @property
    def dtype(self):
        return self.array.dtype

    # --- metadata dictionary -------------------------------------------------
    def GetMetaDataKeys(self):
        # itk::MetaDataDictionary is a std::map<std::string, ...>: keys come out in byte-wise lexicographic order
        # ("Segment0_", "Segment10_", ..., "Segment19_", "Segment1_", "Segment20_", ...).  combine_segmentations relies on
        # that order for the labels that are absent from the image (reference ts2d/core/util/meta.py:254-286).
        return tuple(sorted(self._meta.keys(), key=lambda k: k.encode("utf-8")))

    def GetMetaData(self, key: str) -> str:
        return self._meta[key]

    def HasMetaDataKey(self, key: str) -> bool:
        return key in self._meta

    def SetMetaData(self, key: str, value: str):
        self._meta[str(key)] = str(value)

    def EraseMetaData(self, key: str):
        self._meta.pop(key, None)

    # --- helpers ---------------------------------------------------------------
    def direction_matrix(self) -> np.ndarray:
        d = self.GetDimension()
        return np.asarray(self._direction, dtype=np.float64).reshape(d, d)

    def copy_geometry_from(self, other: "Image") -> "Image":
        self.SetOrigin(other.GetOrigin())
        self.SetDirection(other.GetDirection())
        self.SetSpacing(other.GetSpacing())
        return self

    def __bool__(self):
        return True

    def __repr__(self):
        return "Image(size={}, ncomp={}, dtype={}, spacing={})".format(
            self.GetSize(), self._ncomp, self.array.dtype, self._spacing)


def make_vector_image(array: np.ndarray, like: Optional[Image] = None, **kw) -> Image:
    img = Image(array, is_vector=True, **kw)
    if like is not None:
        img.copy_geometry_from(like)
    return img


def is_image(obj) -> bool:
    return isinstance(obj, Image)


def from_sitk(obj) -> Image:
    """Accept a real ``sitk.Image`` opportunistically when SimpleITK is importable."""
    if isinstance(obj, Image):
        return obj
    try:  # pragma: no cover - SimpleITK is absent in this image
        import SimpleITK as sitk
        if isinstance(obj, sitk.Image):
            arr = sitk.GetArrayFromImage(obj)
            img = Image(arr, obj.GetSpacing(), obj.GetOrigin(), obj.GetDirection(),
                        is_vector=obj.GetNumberOfComponentsPerPixel() > 1)
            for k in obj.GetMetaDataKeys():
                img.SetMetaData(k, obj.GetMetaData(k))
            return img
    except ImportError:
        pass
    raise RuntimeError(f"input must be a string path or an image, found: {type(obj).__name__}")
