"""ctypes binding of ``libts2d_b200.so`` (the C ABI declared in include/ts2d_b200.h).

There is deliberately no fallback: if the library is missing or the device is not
an sm_100 GPU the product path raises.  PyTorch only owns device memory and streams;
every pointer handed to the library is ``tensor.data_ptr()``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libts2d_b200.so")

TS2D_MAX_STAGES = 12

# enums (include/ts2d_b200.h)
U8, I16, U16, I32, F32, F64 = range(6)
PROJ_MEAN, PROJ_MAX, PROJ_MIN = range(3)
NORM_NONE, NORM_ZSCORE, NORM_CT, NORM_RESCALE01, NORM_RGB01 = range(5)
PREC_F32, PREC_BF16, PREC_F16 = range(3)

SCHEME_IDS = {
    "NoNormalization": NORM_NONE,
    "ZScoreNormalization": NORM_ZSCORE,
    "CTNormalization": NORM_CT,
    "RescaleTo01Normalization": NORM_RESCALE01,
    "RGBTo01Normalization": NORM_RGB01,
}
PRECISION_IDS = {"fp32": PREC_F32, "float32": PREC_F32, "bf16": PREC_BF16, "bfloat16": PREC_BF16,
                 "fp16": PREC_F16, "float16": PREC_F16, "half": PREC_F16}


class UNetPlan(C.Structure):
    _fields_ = [
        ("n_stages", C.c_int32), ("in_channels", C.c_int32), ("num_heads", C.c_int32),
        ("features", C.c_int32 * TS2D_MAX_STAGES),
        ("strides", (C.c_int32 * 2) * TS2D_MAX_STAGES),
        ("kernel_sizes", (C.c_int32 * 2) * TS2D_MAX_STAGES),
        ("n_conv_enc", C.c_int32 * TS2D_MAX_STAGES),
        ("n_conv_dec", C.c_int32 * TS2D_MAX_STAGES),
        ("eps", C.c_float), ("negative_slope", C.c_float),
        ("patch_h", C.c_int32), ("patch_w", C.c_int32),
        ("precision", C.c_int32), ("max_batch", C.c_int32), ("use_tensor_cores", C.c_int32),
    ]


class TS2DError(RuntimeError):
    pass


_lib: Optional[C.CDLL] = None
_checked_devices = set()

_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float
_ip = C.POINTER(C.c_int)
_PROTOS = {
    "ts2d_last_error": (C.c_char_p, []),
    "ts2d_abi_version": (_i, []),
    "ts2d_device_check": (_i, [_i]),
    "ts2d_launch_count": (_i64, []),
    "ts2d_project": (_i, [_vp, _i, _i64, _i64, _i64, _i, _i, _i, _i, _ip, _i, _vp, _vp]),
    "ts2d_bbox_nonzero": (_i, [_vp, _i, _i, _i, _vp, _vp]),
    "ts2d_channel_stats": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "ts2d_normalize": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _ip, C.POINTER(C.c_float), _vp, _vp, _vp]),
    "ts2d_resample2d": (_i, [_vp, _i, _i, _i, _vp, _i, _i, _i, _vp, _vp]),
    "ts2d_resample2d_static": (_i, [_vp, _i, _i, _i, _vp, _i, _i, _i, _vp]),
    "ts2d_resample2d_static_capacity": (_i64, []),
    "ts2d_extract_tiles": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _ip, _i, _i, _i, _vp, _vp]),
    "ts2d_unet_create": (_i, [C.POINTER(UNetPlan), C.POINTER(_vp)]),
    "ts2d_unet_set_weight": (_i, [_vp, C.c_char_p, _vp, _i64]),
    "ts2d_unet_finalize": (_i, [_vp]),
    "ts2d_unet_resize": (_i, [_vp, _i]),
    "ts2d_unet_workspace_bytes": (C.c_size_t, [_vp]),
    "ts2d_unet_forward": (_i, [_vp, _vp, _i, _vp]),
    "ts2d_unet_tile_logits": (_i, [_vp, _i, _vp, _vp]),
    "ts2d_unet_forward_debug": (_i, [_vp, _vp, _i, C.c_char_p, _vp, _vp]),
    "ts2d_unet_blend_head": (_i, [_vp, _ip, _i, _i, _i, _i, _i, _i, _i, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "ts2d_threshold_logits": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "ts2d_permute_channels": (_i, [_vp, _vp, _i64, _i, _ip, _vp]),
    "ts2d_visual_labels": (_i, [_vp, _i, _i, _i, _i, _i, C.c_double, C.c_double, C.c_char_p, _i, _vp, _vp]),
    "ts2d_visual_intensity": (_i, [_vp, _i, _i, _i, _i, C.c_double, C.c_double, _i, C.c_double, C.c_double, _i, _i, _vp, _vp, _vp, _vp]),
    "ts2d_merge_present_order": (_i, [_vp, _vp, _i64, _i, _vp, _ip, _i, _vp, _vp]),
    "ts2d_unet_launch_count": (_i64, [_vp]),
    "ts2d_pack_mask_bits": (_i, [_vp, _vp, _i64, _vp]),
    "ts2d_host_alloc": (_vp, [C.c_size_t, _i]),
    "ts2d_host_free": (_i, [_vp]),
    "ts2d_upload_pageable": (_i, [_vp, _vp, C.c_size_t, _i, _vp]),
    "ts2d_unpack_mask_bits_host": (_i, [_vp, _vp, _i64, _i]),
    "ts2d_unet_direct_layer_count": (_i, [_vp]),
    "ts2d_unet_check_guards": (_i, [_vp]),
    "ts2d_unet_guard_count": (_i, [_vp]),
    "ts2d_unet_blend_covers_image": (_i, [_vp]),
    "ts2d_unet_flops": (C.c_double, [_vp, _i]),
    "ts2d_unet_set_profiling": (_i, [_vp, _i]),
    "ts2d_unet_last_conv_ms": (C.c_double, [_vp]),
    "ts2d_unet_last_forward_ms": (C.c_double, [_vp]),
    "ts2d_unet_profile_count": (_i, [_vp]),
    "ts2d_unet_profile_entry": (_i, [_vp, _i, C.c_char_p, _i, _ip, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "ts2d_unet_destroy": (_i, [_vp]),
}
EXPORTED_SYMBOLS = tuple(_PROTOS.keys())


def load(check_device: bool = False) -> C.CDLL:
    """Loads the library (once).  Raises when it has not been built: there is no CPU path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise TS2DError(
                "libts2d_b200.so is missing at {} - build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C totalsegmentator2d_b200/csrc`; there is no CPU fallback for the TS2D hot path.".format(LIB_PATH))
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    if check_device:
        import torch
        if not torch.cuda.is_available():
            raise TS2DError("no CUDA device: the TS2D hot path runs on an sm_100 (B200) GPU only")
        dev = torch.cuda.current_device()
        if dev not in _checked_devices:
            check(_lib.ts2d_device_check(dev))
            _checked_devices.add(dev)
    return _lib


def last_error() -> str:
    return (_lib.ts2d_last_error() or b"").decode("utf-8", "replace") if _lib is not None else ""


def check(rc: int, what: str = ""):
    if rc < 0:
        raise TS2DError("{}{} (status {})".format(what + ": " if what else "", last_error(), rc))
    return rc


def int_array(values):
    arr = (C.c_int * len(values))(*[int(v) for v in values])
    return arr


def float_array(values):
    return (C.c_float * len(values))(*[float(v) for v in values])


def stream_ptr(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)
