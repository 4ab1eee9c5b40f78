"""Thin torch-tensor wrappers over the C ABI (one function per entry point).

PyTorch is plumbing here: it allocates device memory and provides the stream; the
arithmetic is in libts2d_b200.so.  Each wrapper cites the reference call it replaces.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib as L

_NP_TO_TS = {np.dtype("uint8"): L.U8, np.dtype("int16"): L.I16, np.dtype("uint16"): L.U16, np.dtype("int32"): L.I32,
             np.dtype("float32"): L.F32, np.dtype("float64"): L.F64}
_TORCH_TO_TS = {torch.uint8: L.U8, torch.int16: L.I16, torch.uint16: L.U16, torch.int32: L.I32, torch.float32: L.F32,
                torch.float64: L.F64}
_PREC_TO_TORCH = {L.PREC_F32: torch.float32, L.PREC_BF16: torch.bfloat16, L.PREC_F16: torch.float16}
_MODE_IDS = {"mean": L.PROJ_MEAN, "avg": L.PROJ_MEAN, "max": L.PROJ_MAX, "mip": L.PROJ_MAX, "min": L.PROJ_MIN}


def _dev(t: torch.Tensor, what: str):
    if not t.is_cuda or not t.is_contiguous():
        raise L.TS2DError(f"{what} must be a contiguous CUDA tensor")
    return t.data_ptr()


def project(vol: torch.Tensor, modes: Sequence[str], reduce_dim: int = 1, h_dim: int = 0, flip_h: bool = False,
            flip_w: bool = False, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K1. vol: CUDA tensor [n0][n1][n2]; returns float32 [len(modes)][H][W].
    Replaces reorient_image + project + Cast + Compose (reference ts2d/core/util/image.py:32-101, ts2d/tool.py:154-160,181-185)."""
    lib = L.load()
    if vol.dim() != 3:
        raise L.TS2DError("project expects a 3-D volume")
    if vol.dtype not in _TORCH_TO_TS:
        raise L.TS2DError(f"unsupported voxel type {vol.dtype}")
    ids = []
    for m in modes:
        key = str(m).lower().strip()
        if key not in _MODE_IDS:
            raise RuntimeError("Unsupported filter mode: {}".format(key))
        ids.append(_MODE_IDS[key])
    kept = [d for d in range(3) if d != reduce_dim]
    w_dim = kept[0] if kept[1] == h_dim else kept[1]
    H, W = int(vol.shape[h_dim]), int(vol.shape[w_dim])
    if out is None:
        out = torch.empty((len(ids), H, W), dtype=torch.float32, device=vol.device)
    L.check(lib.ts2d_project(_dev(vol, "vol"), _TORCH_TO_TS[vol.dtype], vol.shape[0], vol.shape[1], vol.shape[2], reduce_dim, h_dim,
                             int(flip_h), int(flip_w), L.int_array(ids), len(ids), _dev(out, "out"), L.stream_ptr()), "ts2d_project")
    return out


def bbox_nonzero(img: torch.Tensor) -> torch.Tensor:
    """crop_to_nonzero bbox (nnunetv2 cropping.py) -> int32[4] device tensor {h0,h1,w0,w1}."""
    lib = L.load()
    C, H, W = img.shape
    out = torch.empty(4, dtype=torch.int32, device=img.device)
    L.check(lib.ts2d_bbox_nonzero(_dev(img, "img"), C, H, W, out.data_ptr(), L.stream_ptr()), "ts2d_bbox_nonzero")
    return out


def channel_stats(img: torch.Tensor, bbox: Tuple[int, int, int, int]) -> torch.Tensor:
    lib = L.load()
    C, H, W = img.shape
    out = torch.empty((C, 4), dtype=torch.float64, device=img.device)
    L.check(lib.ts2d_channel_stats(_dev(img, "img"), C, H, W, *[int(b) for b in bbox], out.data_ptr(), L.stream_ptr()), "ts2d_channel_stats")
    return out


def normalize(img: torch.Tensor, bbox, schemes: Sequence[int], params: Optional[Sequence[Sequence[float]]], stats: Optional[torch.Tensor]) -> torch.Tensor:
    """DefaultPreprocessor._normalize on the cropped image (nnunetv2 default_normalization_schemes.py)."""
    lib = L.load()
    C, H, W = img.shape
    h0, h1, w0, w1 = [int(b) for b in bbox]
    out = torch.empty((C, h1 - h0, w1 - w0), dtype=torch.float32, device=img.device)
    flat = None
    if params is not None:
        flat = L.float_array([v for row in params for v in row])
    L.check(lib.ts2d_normalize(_dev(img, "img"), C, H, W, h0, h1, w0, w1, L.int_array(list(schemes)), flat,
                               stats.data_ptr() if stats is not None else None, out.data_ptr(), L.stream_ptr()), "ts2d_normalize")
    return out


def resample2d(img: torch.Tensor, new_hw: Sequence[int], order: int) -> torch.Tensor:
    """skimage.transform.resize(order, mode='edge', anti_aliasing=False) per channel (nnunetv2 default_resampling.py)."""
    lib = L.load()
    C, h, w = img.shape
    h2, w2 = int(new_hw[0]), int(new_hw[1])
    out = torch.empty((C, h2, w2), dtype=torch.float32, device=img.device)
    scratch = None
    if order == 3:
        scratch = torch.empty(C * (h + 24) * (w + 24) + 2 * C, dtype=torch.float64, device=img.device)
    L.check(lib.ts2d_resample2d(_dev(img, "img"), C, h, w, out.data_ptr(), h2, w2, order,
                                scratch.data_ptr() if scratch is not None else None, L.stream_ptr()), "ts2d_resample2d")
    return out


def mirror_count(mirror_mask: int) -> int:
    return 4 if mirror_mask == 3 else (2 if mirror_mask else 1)


def extract_tiles(img: torch.Tensor, pads: Tuple[int, int], patch: Sequence[int], tile_yx: Sequence[Tuple[int, int]], mirror_mask: int,
                  precision: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """pad_nd_image + slicing + torch.flip of predict_sliding_window_return_logits -> NHWC tiles [T*M][Ph][Pw][C]."""
    lib = L.load()
    C, h, w = img.shape
    T, M = len(tile_yx), mirror_count(mirror_mask)
    if out is None:
        out = torch.empty((T * M, int(patch[0]), int(patch[1]), C), dtype=_PREC_TO_TORCH[precision], device=img.device)
    flat = [v for yx in tile_yx for v in yx]
    L.check(lib.ts2d_extract_tiles(_dev(img, "img"), C, h, w, int(pads[0]), int(pads[1]), int(patch[0]), int(patch[1]), L.int_array(flat), T,
                                   mirror_mask, precision, out.data_ptr(), L.stream_ptr()), "ts2d_extract_tiles")
    return out


def threshold_logits(logits: torch.Tensor, new_hw, mask: torch.Tensor, h0: int, w0: int, ch_off: int):
    """export_prediction_from_logits for a resampling plan: order-1 resize -> sigmoid>0.5 -> paste."""
    lib = L.load()
    heads, h, w = logits.shape
    H0, W0, ch_total = mask.shape
    L.check(lib.ts2d_threshold_logits(_dev(logits, "logits"), heads, h, w, int(new_hw[0]), int(new_hw[1]), _dev(mask, "mask"), H0, W0, h0, w0,
                                      ch_off, ch_total, L.stream_ptr()), "ts2d_threshold_logits")


def permute_channels(src: torch.Tensor, perm: Sequence[int], out: Optional[torch.Tensor] = None) -> torch.Tensor:
    lib = L.load()
    Cc = src.shape[-1]
    if out is None:
        out = torch.empty_like(src)
    L.check(lib.ts2d_permute_channels(_dev(src, "src"), _dev(out, "out"), src.numel() // Cc, Cc, L.int_array(list(perm)), L.stream_ptr()),
            "ts2d_permute_channels")
    return out


def merge_present_order(src: torch.Tensor, present: torch.Tensor, offsets: Sequence[int], out: Optional[torch.Tensor] = None):
    """combine_segmentations channel order (labels present in the image first, per model) computed and applied on the device.
    Returns (merged mask, perm int32[C] on the device)."""
    lib = L.load()
    Cc = src.shape[-1]
    if out is None:
        out = torch.empty_like(src)
    perm = torch.empty(Cc, dtype=torch.int32, device=src.device)
    L.check(lib.ts2d_merge_present_order(_dev(src, "src"), _dev(out, "out"), src.numel() // Cc, Cc, _dev(present, "present"),
                                         L.int_array(list(offsets)), len(offsets) - 1, perm.data_ptr(), L.stream_ptr()), "ts2d_merge_present_order")
    return out, perm


def _release_upload_memory(lib, ptr):
    try:
        lib.ts2d_host_free(ptr)
    except Exception:          # interpreter shutdown: the driver reclaims page-locked memory with the process
        pass


class UploadBuffer:
    """Page-locked (optionally write-combined) host array for a volume that is decoded once and uploaded once:
    ``buf = UploadBuffer((1000, 512, 512), np.int16); buf.array[...] = volume; Image(buf.array, ...)``.  torch recognises the memory as
    pinned, so ``TS2D.predict`` uploads it with an asynchronous copy.

    The memory belongs to ``array`` (and to every view, ``Image`` or ``Result['input']`` made from it), not to this wrapper: it is
    released when the last of them is garbage-collected, so ``buf`` itself may be dropped or rebound at any time - e.g. in a generator that
    feeds ``TS2D.predict_batch`` and allocates a fresh buffer per study while earlier results are still being consumed.
    ``ts2d_host_free`` waits for the device before it unpins, so an upload still in flight is not cut short either.
    ``close()`` only drops this wrapper's own reference."""

    def __init__(self, shape, dtype=np.int16, write_combined: bool = True):
        import ctypes
        import weakref
        lib = L.load(check_device=True)
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        ptr = lib.ts2d_host_alloc(self.nbytes, int(bool(write_combined)))
        if not ptr:
            raise L.TS2DError("ts2d_host_alloc: " + L.last_error())
        owner = (ctypes.c_uint8 * self.nbytes).from_address(ptr)      # the ndarray's base: lives as long as any view of the memory
        self._finalizer = weakref.finalize(owner, _release_upload_memory, lib, ptr)
        self.array = np.frombuffer(owner, dtype=dtype).reshape(shape)

    @property
    def released(self) -> bool:
        """True once the page-locked memory has gone back to the driver (no array or view of it is left)"""
        return not self._finalizer.alive

    def close(self):
        self.array = None


def pack_mask_bits(mask: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """0/1 uint8 mask (any shape, contiguous) -> ceil(numel / 8) bytes on the device, bit k of byte i = mask byte 8i + k"""
    lib = L.load()
    n = mask.numel()
    if out is None:
        out = torch.empty((n + 7) // 8, dtype=torch.uint8, device=mask.device)
    L.check(lib.ts2d_pack_mask_bits(_dev(mask, "mask"), _dev(out, "out"), n, L.stream_ptr()), "ts2d_pack_mask_bits")
    return out


def unpack_mask_bits_host(packed: np.ndarray, shape, threads: int = 4) -> np.ndarray:
    """host side of pack_mask_bits: packed uint8 array -> uint8 array of ``shape`` holding 0 / 1 (no GPU involved)"""
    lib = L.load(check_device=False)
    n = int(np.prod(shape))
    packed = np.ascontiguousarray(packed, dtype=np.uint8)
    if packed.size < (n + 7) // 8:
        raise L.TS2DError("unpack_mask_bits_host: {} packed bytes cannot hold {} mask bytes".format(packed.size, n))
    out = np.empty(n, dtype=np.uint8)
    L.check(lib.ts2d_unpack_mask_bits_host(packed.ctypes.data, out.ctypes.data, n, int(threads)), "ts2d_unpack_mask_bits_host")
    return out.reshape(shape)


def visual_labels(mask: torch.Tensor, new_hw, ratio_yx, palette: np.ndarray) -> torch.Tensor:
    """create_visual for a multi-component mask (reference ts2d/core/util/image.py:431-440): [H][W][C] uint8 -> RGB uint8 [H2][W2][3]."""
    lib = L.load()
    H, W, Cc = mask.shape
    H2, W2 = int(new_hw[0]), int(new_hw[1])
    pal = np.ascontiguousarray(np.asarray(palette, dtype=np.uint8).reshape(-1, 3))
    out = torch.empty((H2, W2, 3), dtype=torch.uint8, device=mask.device)
    L.check(lib.ts2d_visual_labels(_dev(mask, "mask"), Cc, H, W, H2, W2, float(ratio_yx[0]), float(ratio_yx[1]), pal.tobytes(), int(pal.shape[0]),
                                   out.data_ptr(), L.stream_ptr()), "ts2d_visual_labels")
    return out


def visual_intensity(img: torch.Tensor, new_hw, ratio_yx, window=(None, None)) -> torch.Tensor:
    """create_visual for an intensity image (reference ts2d/core/util/image.py:441-451): float32 [H][W] -> uint8 [H2][W2]."""
    lib = L.load()
    H, W = img.shape
    H2, W2 = int(new_hw[0]), int(new_hw[1])
    resample = (H2, W2) != (H, W) or not np.allclose(ratio_yx, (1.0, 1.0))
    scratch = torch.empty(H * W + 2, dtype=torch.float64, device=img.device)
    res = torch.empty((H2, W2), dtype=torch.float32, device=img.device)
    out = torch.empty((H2, W2), dtype=torch.uint8, device=img.device)
    lo, hi = window
    L.check(lib.ts2d_visual_intensity(_dev(img, "img"), H, W, H2, W2, float(ratio_yx[0]), float(ratio_yx[1]), int(resample),
                                      float(lo if lo is not None else 0.0), float(hi if hi is not None else 0.0), int(lo is None), int(hi is None),
                                      scratch.data_ptr(), res.data_ptr(), out.data_ptr(), L.stream_ptr()), "ts2d_visual_intensity")
    return out


def torch_dtype_of(precision: int):
    return _PREC_TO_TORCH[precision]
