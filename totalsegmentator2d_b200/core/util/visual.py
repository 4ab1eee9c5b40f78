"""PNG visuals for ``Result.save(content='visual')`` (reference ts2d/core/util/image.py:383-481 create_visual /
get_auto_window and ts2d/tool.py:272-284).  Not on the hot path: a small numpy + zlib PNG writer with the same file
naming; intensity images get a percentile window, label images a colour per segment (max over overlapping labels)."""
from __future__ import annotations

import struct
import zlib

import numpy as np

from ...image import Image
from .color import default_color, tuple_to_color
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


def _plane(arr: np.ndarray) -> np.ndarray:
    a = np.asarray(arr)
    while a.ndim > 2:
        ax = [i for i, s in enumerate(a.shape) if s == 1]
        a = np.squeeze(a, axis=ax[0]) if ax else a.max(axis=1)      # coronal MIP for true 3-D content
    return a


def create_visual(img: Image, labels: bool = False) -> np.ndarray:
    if labels:
        arr = np.asarray(img.array)
        nch = img.GetNumberOfComponentsPerPixel()
        arr = arr.reshape([s for s in arr.shape[:-1] if s > 1] + [nch]) if img.is_vector else _plane(arr)[..., None]
        out = np.zeros(arr.shape[:2] + (3,), dtype=np.uint8)
        info = get_annotation_labels(img, fetch=False) if img.is_vector else {}
        by_value = {v["value"]: v for v in info.values()}
        for c in range(arr.shape[-1]):
            m = arr[..., c] > 0
            if not m.any():
                continue
            col = (by_value.get(c + 1) or {}).get("color") or default_color(c + 1)
            out[m] = np.maximum(out[m], np.asarray(tuple_to_color(col), dtype=np.uint8))
        return out
    a = _plane(img.array).astype(np.float64)
    lo, hi = np.percentile(a, [1, 99]) if a.size else (0, 1)
    if hi <= lo:
        hi = lo + 1
    return (np.clip((a - lo) / (hi - lo), 0, 1) * 255).astype(np.uint8)


def write_visual(img: Image, base_path: str, suffix: str, labels: bool):
    if labels or img.GetNumberOfComponentsPerPixel() == 1:
        write_png(f"{base_path}{suffix}.png", create_visual(img, labels=labels))
    else:
        for c in range(img.GetNumberOfComponentsPerPixel()):
            ch = Image(np.ascontiguousarray(img.array[..., c]), img.GetSpacing(), img.GetOrigin(), img.GetDirection())
            write_png(f"{base_path}-ch{c}{suffix}.png", create_visual(ch))
