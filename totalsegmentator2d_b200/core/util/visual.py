"""PNG visuals for ``Result.save(content='visual')``: reference ts2d/core/util/image.py:383-451 ``create_visual`` (with
``get_auto_window`` :454-481, ``image_vector_flatten_max`` :266-290, ``resample_uniform`` :374-379, ``to_palette``
color.py:87-103) as called from ts2d/tool.py:272-284.  The raster arithmetic (label flattening, uniform resampling, LabelToRGB,
windowing, cast) runs in csrc/visual.cu; this module is the host mirror of the function (same arguments, same decisions) plus a
small PNG encoder.  3-D inputs are reoriented to RAI and projected with the projection kernel (``mode``, ``axis``) like the
reference does."""
from __future__ import annotations

import struct
import zlib
from typing import Optional

import numpy as np
import torch

from ... import ops
from ...image import Image
from .color import default_color, to_color
from .meta import get_annotation_labels


def write_png(path: str, rgb: np.ndarray):
    rgb = np.ascontiguousarray(rgb.astype(np.uint8))
    h, w = rgb.shape[:2]
    ctype = 2 if rgb.ndim == 3 else 0
    raw = b"".join(b"\x00" + rgb[y].tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, ctype, 0, 0, 0)) +
                chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def to_palette(v) -> list:
    """reference color.py:87-103: index = label value; entry 0 is a white placeholder for the background"""
    if isinstance(v, dict):
        if any(not isinstance(k, (int, np.integer)) or k < 0 for k in v.keys()):
            raise RuntimeError("Dictionary palette must consist of non-negative integer keys!")
        lim = max(v.keys()) if v.keys() else 0
        res = [[255, 255, 255]]
        for idx in range(1, lim + 1):
            c = v.get(idx)
            res.append(list(to_color(c)) if c is not None else list(default_color(idx)))
        return res
    return [list(to_color(c)) for c in v]


def uniform_geometry(size_xy, spacing_xy):
    """resample_uniform -> resample_filter (image.py:319-379): target spacing = min spacing; returns (new size (x, y), ratios
    new_sp / old_sp per axis (x, y), changed)"""
    sp = float(min(spacing_xy))
    new_size = [int(0.5 + size_xy[i] * spacing_xy[i] / sp) for i in range(2)]
    ratios = [sp / float(spacing_xy[i]) for i in range(2)]
    changed = (not np.allclose([sp, sp], spacing_xy)) or list(new_size) != list(size_xy)
    return new_size, ratios, changed


def _to_2d(img: Image, mode: str, axis) -> Image:
    """reorient_image + the reduce_dimensions / project loop of create_visual (image.py:415-425)"""
    from .image import axis_name_to_index, project, reduce_dimensions, reorient_image
    if img.GetDimension() >= 3:
        img = reorient_image(img)
    _axis = axis_name_to_index(axis) if isinstance(axis, str) else (-1 if axis is None else axis)
    while True:
        img = reduce_dimensions(img, min=2)
        dim = img.GetDimension()
        if dim <= 2:
            return img
        _axis = -1 if abs(_axis) > dim else _axis
        img = project(img, mode=mode, axis=_axis)


def create_visual(img: Image, mode: str = "max", axis=-1, window=None, labels: Optional[bool] = None, palette=None) -> np.ndarray:
    """returns the uint8 raster ([H][W][3] for label images, [H][W] for intensity images)"""
    if labels is None:
        # reference: default(labels, palette or is_label_image(img)); sitk label pixel types + plain (scalar) UInt8
        labels = bool(palette) or (img.array.dtype == np.uint8 and not img.is_vector)
    if labels and not palette:
        palette = dict()
        try:
            for k, v in get_annotation_labels(img).items():
                value, color = v.get("value"), v.get("color")
                if value is not None and color is not None:
                    palette[int(value)] = color
        except Exception:
            palette = dict()
    img = _to_2d(img, mode, axis)
    size, spacing = img.GetSize(), img.GetSpacing()
    new_size, ratios, changed = uniform_geometry(size, spacing)
    arr = np.asarray(img.array)
    if labels:
        cmap = to_palette(palette) if palette is not None else []
        if not img.is_vector:
            # a single-channel label map: LabelToRGB on the values themselves (one pseudo component per value is not needed)
            vals = arr.astype(np.int64)
            mask = np.zeros(arr.shape + (int(vals.max()) if vals.size else 0,), dtype=np.uint8)
            for v in np.unique(vals):
                if v > 0:
                    mask[..., int(v) - 1] = vals == v
            arr = mask if mask.shape[-1] else np.zeros(arr.shape + (1,), np.uint8)
        dev = torch.from_numpy(np.ascontiguousarray(arr.astype(np.uint8))).cuda()
        out = ops.visual_labels(dev, (new_size[1], new_size[0]), (ratios[1], ratios[0]), np.asarray(cmap, dtype=np.uint8).reshape(-1, 3))
        return out.cpu().numpy()
    if img.is_vector:
        arr = np.sqrt((arr.astype(np.float64) ** 2).sum(-1))       # sitk.VectorMagnitude (never reached from Result.save: channels are split first)
    if arr.dtype == np.uint8 and changed:
        raise RuntimeError("SimpleITK does not permit interpolation of UInt8 images (the reference falls back to nearest neighbour): not implemented")
    if isinstance(window, str):
        if window.lower() != "minmax":
            raise RuntimeError("only the 'minmax' auto window is implemented on the device (got '{}')".format(window))
        window = None
    lo, hi = (None, None) if window is None else window
    dev = torch.from_numpy(np.ascontiguousarray(arr.astype(np.float32))).cuda()
    hw = (new_size[1], new_size[0]) if changed else arr.shape
    out = ops.visual_intensity(dev, hw, (ratios[1], ratios[0]) if changed else (1.0, 1.0), (lo, hi))
    return out.cpu().numpy()


def write_visual(img: Image, base_path: str, suffix: str, labels: bool):
    """reference tool.py:272-284 (_export_image, 'visual' content)"""
    if labels:
        write_png(f"{base_path}{suffix}.png", create_visual(img, labels=True, axis="coronal"))
        return
    nch = img.GetNumberOfComponentsPerPixel()
    if nch == 1:
        write_png(f"{base_path}{suffix}.png", create_visual(img, labels=False, axis="coronal"))
    else:
        for c in range(nch):
            ch = Image(np.ascontiguousarray(img.array[..., c]), img.GetSpacing(), img.GetOrigin(), img.GetDirection())
            write_png(f"{base_path}-ch{c}{suffix}.png", create_visual(ch, labels=False, axis="coronal"))
