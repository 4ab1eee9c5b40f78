"""``B200nnUNetPredictor``: the operator the reference worker drives, on hand-written sm_100a kernels.

The reference builds ``nnunetv2.inference.predict_from_raw_data.nnUNetPredictor`` in
``NNUProcessModel._lazy_load_predictor`` (reference ts2d/core/inference/nnu.py:149-166) and its worker
checks ``'nnUNetPredictor' in type(predictor).__name__`` (prediction_worker.py:103,111) before calling
``configuration_manager.preprocessor_class(...).run_case``, ``predict_logits_from_preprocessed_data`` and
``export_prediction_from_logits`` (:194-221).  This class has the same constructor keywords, the same
attributes and the same methods, so it can be handed to that worker unchanged; ``TS2D.predict`` calls it
in-process instead (no subprocess, no temp files).

Sliding window, mirror TTA and Gaussian blending follow upstream nnU-Net v2.6
(predict_sliding_window_return_logits, compute_steps_for_sliding_window, compute_gaussian).
"""
from __future__ import annotations

import ctypes as C
import itertools
import os
import re
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ... import _lib as L
from ... import ops
from .plans import ConfigurationManager, PlansManager, channel_names, foreground_labels, read_json


def compute_steps_for_sliding_window(image_size: Sequence[int], tile_size: Sequence[int], tile_step_size: float) -> List[List[int]]:
    assert all(i >= j for i, j in zip(image_size, tile_size)), "image size must be as large or larger than patch_size"
    assert 0 < tile_step_size <= 1, "step_size must be larger than 0 and smaller or equal to 1"
    target = [i * tile_step_size for i in tile_size]
    num_steps = [int(np.ceil((i - k) / j)) + 1 for i, j, k in zip(image_size, target, tile_size)]
    steps = []
    for dim in range(len(tile_size)):
        max_step = image_size[dim] - tile_size[dim]
        actual = max_step / (num_steps[dim] - 1) if num_steps[dim] > 1 else 99999999999
        steps.append([int(np.round(actual * i)) for i in range(num_steps[dim])])
    return steps


def compute_gaussian(tile_size: Sequence[int], sigma_scale: float = 1.0 / 8, value_scaling_factor: float = 10.0) -> np.ndarray:
    """upstream compute_gaussian: scipy gaussian_filter of a centred delta (truncate 4 sigma, mode 'constant'),
    scaled to max 10, rounded to fp16, zeros replaced by the smallest positive value. Returned as float32 of
    the fp16 values (the kernels blend in fp32)."""
    from scipy.ndimage import gaussian_filter
    tmp = np.zeros(tile_size)
    tmp[tuple(i // 2 for i in tile_size)] = 1
    g = gaussian_filter(tmp, [i * sigma_scale for i in tile_size], 0, mode="constant", cval=0)
    g = g / (g.max() / value_scaling_factor)
    g = g.astype(np.float16)
    g[g == 0] = g[g != 0].min()
    return g.astype(np.float32)


def sliding_window_layout(hw: Sequence[int], patch: Sequence[int], tile_step_size: float):
    """returns (pads (top,left), padded size, tile origins [(y,x)...]) for an image of size hw."""
    pads, padded = [], []
    for a in range(2):
        new = max(int(hw[a]), int(patch[a]))
        diff = new - int(hw[a])
        pads.append(diff // 2)
        padded.append(new)
    steps = compute_steps_for_sliding_window(padded, patch, tile_step_size)
    tiles = [(sy, sx) for sy in steps[0] for sx in steps[1]]
    return tuple(pads), tuple(padded), tiles


class _LabelManager:
    def __init__(self, dataset_json: dict):
        self.foreground_labels = foreground_labels(dataset_json)
        self.num_segmentation_heads = len(self.foreground_labels)


class B200nnUNetPredictor:
    def __init__(self, tile_step_size: float = 0.5, use_gaussian: bool = True, use_mirroring: bool = True,
                 perform_everything_on_device: bool = True, device=None, verbose: bool = False,
                 verbose_preprocessing: bool = False, allow_tqdm: bool = False, precision: str = "fp16",
                 use_tensor_cores: bool = True, max_batch: int = 12):
        self.tile_step_size = tile_step_size
        self.use_gaussian = use_gaussian
        self.use_mirroring = use_mirroring
        self.verbose = verbose
        self.allow_tqdm = allow_tqdm
        if not torch.cuda.is_available():
            raise L.TS2DError("B200nnUNetPredictor needs an sm_100 GPU: there is no CPU path (the reference's CPU "
                              "implementation is upstream nnU-Net, not this package)")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if self.device.type != "cuda":
            raise L.TS2DError("device must be a CUDA device")
        if not use_gaussian:
            raise L.TS2DError("use_gaussian=False is not implemented")
        self.precision = L.PRECISION_IDS[str(precision).lower()]
        self.use_tensor_cores = bool(use_tensor_cores)
        self.max_batch = int(max_batch)
        self.plans_manager: Optional[PlansManager] = None
        self.configuration_manager: Optional[ConfigurationManager] = None
        self.dataset_json: Optional[dict] = None
        self.label_manager: Optional[_LabelManager] = None
        self.allowed_mirroring_axes = None
        self._handles: List[int] = []
        self._gaussian: Optional[torch.Tensor] = None
        self._lib = L.load(check_device=True)

    # ---- loading ---------------------------------------------------------------------------------------
    def initialize_from_trained_model_folder(self, model_training_output_dir: str, use_folds=None,
                                             checkpoint_name: str = "checkpoint_final.pth"):
        dataset_json = read_json(os.path.join(model_training_output_dir, "dataset.json"))
        plans = read_json(os.path.join(model_training_output_dir, "plans.json"))
        if use_folds is None:
            use_folds = sorted(int(p.split("_")[1]) for p in os.listdir(model_training_output_dir) if re.match(r"fold_[0-9]+$", p))
        if isinstance(use_folds, (int, str)):
            use_folds = [use_folds]
        m = re.search(r"__([^_]+(?:_[^_]+)*)$", os.path.basename(os.path.normpath(model_training_output_dir)))
        configuration = os.path.basename(os.path.normpath(model_training_output_dir)).split("__")[-1]
        state_dicts, mirror = [], None
        for f in use_folds:
            ck = torch.load(os.path.join(model_training_output_dir, f"fold_{f}", checkpoint_name), map_location="cpu", weights_only=False)
            state_dicts.append(ck["network_weights"])
            mirror = ck.get("inference_allowed_mirroring_axes", None)
        self.manual_initialization(plans, configuration, dataset_json, state_dicts, mirror)

    def manual_initialization(self, plans: dict, configuration: str, dataset_json: dict, state_dicts: Sequence[dict], inference_allowed_mirroring_axes):
        self.close()
        for key in ("transpose_forward", "transpose_backward"):
            t = [int(v) for v in plans.get(key, [0, 1, 2])]
            if t != sorted(t):
                raise L.TS2DError(f"plans.json {key} = {t}: only identity transposes are implemented (a 2-D configuration of a "
                                  "SimpleITKIO dataset never has another one)")
        self.plans_manager = PlansManager(plans)
        self.configuration_manager = self.plans_manager.get_configuration(configuration)
        self.dataset_json = dataset_json
        self.label_manager = _LabelManager(dataset_json)
        self.allowed_mirroring_axes = tuple(inference_allowed_mirroring_axes) if inference_allowed_mirroring_axes is not None else None
        cm = self.configuration_manager
        if len(cm.patch_size) != 2:
            raise L.TS2DError("only 2-D configurations are supported (patch_size {})".format(cm.patch_size))
        k = cm.arch_kwargs
        n = int(k["n_stages"])
        if n > L.TS2D_MAX_STAGES:
            raise L.TS2DError("too many stages")
        if str(k.get("norm_op", "")).split(".")[-1] != "InstanceNorm2d" or str(k.get("nonlin", "")).split(".")[-1] != "LeakyReLU":
            raise L.TS2DError("only InstanceNorm2d + LeakyReLU PlainConvUNets are implemented")
        if k.get("dropout_op") not in (None, "None"):
            pass   # dropout is the identity at inference
        plan = L.UNetPlan()
        plan.n_stages = n
        plan.in_channels = len(channel_names(dataset_json))
        plan.num_heads = self.label_manager.num_segmentation_heads
        pair = lambda v: (int(v), int(v)) if not isinstance(v, (list, tuple)) else (int(v[0]), int(v[1]))
        for s in range(n):
            plan.features[s] = int(k["features_per_stage"][s])
            plan.strides[s][0], plan.strides[s][1] = pair(k["strides"][s])
            plan.kernel_sizes[s][0], plan.kernel_sizes[s][1] = pair(k["kernel_sizes"][s])
            nce = k["n_conv_per_stage"]
            plan.n_conv_enc[s] = int(nce[s] if isinstance(nce, (list, tuple)) else nce)
        for d in range(n - 1):
            ncd = k["n_conv_per_stage_decoder"]
            plan.n_conv_dec[d] = int(ncd[d] if isinstance(ncd, (list, tuple)) else ncd)
        plan.eps = float((k.get("norm_op_kwargs") or {}).get("eps", 1e-5))
        plan.negative_slope = float((k.get("nonlin_kwargs") or {}).get("negative_slope", 0.01))
        plan.patch_h, plan.patch_w = cm.patch_size
        plan.precision = self.precision
        plan.max_batch = self.max_batch
        plan.use_tensor_cores = int(self.use_tensor_cores)
        self._plan = plan
        with torch.cuda.device(self.device):
            for sd in state_dicts:
                h = C.c_void_p()
                L.check(self._lib.ts2d_unet_create(C.byref(plan), C.byref(h)), "ts2d_unet_create")
                self._handles.append(h)
                for name, t in sd.items():
                    if not isinstance(t, torch.Tensor):
                        continue
                    a = t.detach().to(torch.float32).contiguous().cpu()
                    L.check(self._lib.ts2d_unet_set_weight(h, name.encode(), a.data_ptr(), a.numel()), "ts2d_unet_set_weight")
                L.check(self._lib.ts2d_unet_finalize(h), "ts2d_unet_finalize")
            self._gaussian = torch.from_numpy(compute_gaussian(tuple(cm.patch_size))).to(self.device).contiguous()

    # ---- properties the host code needs ------------------------------------------------------------------
    @property
    def num_heads(self) -> int:
        return self.label_manager.num_segmentation_heads

    @property
    def mirror_mask(self) -> int:
        if not self.use_mirroring or not self.allowed_mirroring_axes:
            return 0
        mask = 0
        for a in self.allowed_mirroring_axes:
            if a not in (0, 1):
                raise L.TS2DError("mirror axis {} is not valid for a 2-D network".format(a))
            mask |= 1 << int(a)
        return mask

    @property
    def blend_covers_image(self) -> bool:
        """run_tiles(mask=...) writes this model's channel window of every mask pixel (zeros outside the crop box)"""
        return len(self._handles) == 1 and bool(self._lib.ts2d_unet_blend_covers_image(self._handles[0]))

    def workspace_bytes(self) -> int:
        return sum(int(self._lib.ts2d_unet_workspace_bytes(h)) for h in self._handles)

    def launch_count(self) -> int:
        return sum(int(self._lib.ts2d_unet_launch_count(h)) for h in self._handles)

    def direct_layer_count(self) -> int:
        """layers of a reduced-precision tensor-core plan that run on the CUDA-core parity kernel (0 = every conv is on tcgen05)"""
        return sum(int(self._lib.ts2d_unet_direct_layer_count(h)) for h in self._handles)

    def check_guards(self):
        """(canary bands, damaged bands) of the engines' workspaces; bands exist when TS2D_ARENA_GUARD was set at initialisation"""
        n = sum(int(self._lib.ts2d_unet_guard_count(h)) for h in self._handles)
        bad = [int(self._lib.ts2d_unet_check_guards(h)) for h in self._handles]
        if any(b < 0 for b in bad):
            raise L.TS2DError("ts2d_unet_check_guards failed (CUDA error)")
        return n, sum(bad)

    def flops(self, batch: int) -> float:
        return float(self._lib.ts2d_unet_flops(self._handles[0], batch)) * len(self._handles)

    def set_profiling(self, enable: bool):
        for h in self._handles:
            L.check(self._lib.ts2d_unet_set_profiling(h, int(enable)))

    def last_conv_ms(self) -> float:
        return sum(float(self._lib.ts2d_unet_last_conv_ms(h)) for h in self._handles)

    def last_forward_ms(self) -> float:
        """all launches of the last profiled forward (convs + norm-apply + split-K finish)"""
        return sum(float(self._lib.ts2d_unet_last_forward_ms(h)) for h in self._handles)

    def layer_profile(self, fold: int = 0):
        """[(name, kind, ms, algorithmic flops, algorithmic bytes)] of the last profiled forward"""
        h = self._handles[fold]
        out = []
        for i in range(int(self._lib.ts2d_unet_profile_count(h))):
            buf = C.create_string_buffer(64)
            kind, ms, fl, by = C.c_int(), C.c_double(), C.c_double(), C.c_double()
            L.check(self._lib.ts2d_unet_profile_entry(h, i, buf, 64, C.byref(kind), C.byref(ms), C.byref(fl), C.byref(by)))
            out.append((buf.value.decode(), int(kind.value), float(ms.value), float(fl.value), float(by.value)))
        return out

    def _ensure_batch(self, n: int):
        if n > self.max_batch:
            for h in self._handles:
                L.check(self._lib.ts2d_unet_resize(h, n), "ts2d_unet_resize")
            self.max_batch = n

    # ---- inference ---------------------------------------------------------------------------------------
    def make_tiles(self, data: torch.Tensor):
        """data (C, h, w) float32 CUDA -> (tiles NHWC, layout) ; shared by all models with the same preprocessing."""
        cm = self.configuration_manager
        pads, padded, tiles = sliding_window_layout(data.shape[1:], cm.patch_size, self.tile_step_size)
        x = ops.extract_tiles(data.contiguous(), pads, cm.patch_size, tiles, self.mirror_mask, self.precision)
        return x, (pads, padded, tiles)

    def run_tiles(self, x: torch.Tensor, layout, hw: Tuple[int, int], logits: Optional[torch.Tensor] = None,
                  mask: Optional[torch.Tensor] = None, crop_origin=(0, 0), ch_off: int = 0, present: Optional[torch.Tensor] = None):
        """forward + fused blend/head for one image. Writes float32 logits (heads, h, w) and/or the uint8 mask
        window [H0][W0][ch_total] (channels ch_off.., pixels at crop_origin)."""
        pads, padded, tiles = layout
        mm = self.mirror_mask
        M = ops.mirror_count(mm)
        T = len(tiles)
        self._ensure_batch(T * M)
        nfold = len(self._handles)
        if nfold > 1 and mask is not None and logits is None:
            logits = torch.empty((self.num_heads, hw[0], hw[1]), dtype=torch.float32, device=x.device)
        st = L.stream_ptr()
        flat = L.int_array([v for yx in tiles for v in yx])
        acc = None
        for h in self._handles:
            L.check(self._lib.ts2d_unet_forward(h, x.data_ptr(), T * M, st), "ts2d_unet_forward")
            direct_mask = mask is not None and nfold == 1
            lg = logits if nfold == 1 else torch.empty((self.num_heads, hw[0], hw[1]), dtype=torch.float32, device=x.device)
            H0, W0, ctot = (mask.shape if mask is not None else (0, 0, 0))
            L.check(self._lib.ts2d_unet_blend_head(
                h, flat, T, M, mm, pads[0], pads[1], int(hw[0]), int(hw[1]), self._gaussian.data_ptr(),
                lg.data_ptr() if lg is not None else None, mask.data_ptr() if direct_mask else None, int(H0), int(W0),
                int(crop_origin[0]), int(crop_origin[1]), int(ch_off), int(ctot),
                present.data_ptr() if (present is not None and direct_mask) else None, st), "ts2d_unet_blend_head")
            if nfold > 1:
                acc = lg if acc is None else acc + lg
        if nfold > 1:
            logits.copy_(acc / nfold)       # upstream: prediction /= len(list_of_parameters)
            if mask is not None:
                ops.threshold_logits(logits, hw, mask, int(crop_origin[0]), int(crop_origin[1]), ch_off)
                if present is not None:
                    present[ch_off:ch_off + self.num_heads] |= (logits > 0).flatten(1).any(1).to(torch.uint8)
        return logits

    @torch.no_grad()
    def predict_logits_from_preprocessed_data(self, data: torch.Tensor) -> torch.Tensor:
        """data: float32 (C, 1, h, w) (host or device) -> logits float32 (heads, 1, h, w) on ``self.device``
        (reference prediction_worker.py:209 calls ``.cpu()`` on the result)."""
        if data.dim() != 4 or data.shape[1] != 1:
            raise L.TS2DError("expected preprocessed data of shape (C, 1, H, W) for a 2-D configuration")
        with torch.cuda.device(self.device):
            d = data.to(device=self.device, dtype=torch.float32)[:, 0].contiguous()
            x, layout = self.make_tiles(d)
            logits = torch.empty((self.num_heads, d.shape[1], d.shape[2]), dtype=torch.float32, device=self.device)
            self.run_tiles(x, layout, (d.shape[1], d.shape[2]), logits=logits)
        return logits[:, None]

    def predict_tile_logits(self, x: torch.Tensor, fold: int = 0) -> torch.Tensor:
        """NHWC tiles -> per-tile logits (N, heads, Ph, Pw): the plain network forward (parity tests)."""
        n = x.shape[0]
        self._ensure_batch(n)
        out = torch.empty((n, self.num_heads, x.shape[1], x.shape[2]), dtype=torch.float32, device=x.device)
        st = L.stream_ptr()
        L.check(self._lib.ts2d_unet_forward(self._handles[fold], x.data_ptr(), n, st), "ts2d_unet_forward")
        L.check(self._lib.ts2d_unet_tile_logits(self._handles[fold], n, out.data_ptr(), st), "ts2d_unet_tile_logits")
        return out

    def forward_debug(self, x: torch.Tensor, layer: str, shape_chw: Tuple[int, int, int], fold: int = 0) -> torch.Tensor:
        n = x.shape[0]
        self._ensure_batch(n)
        out = torch.empty((n, *shape_chw), dtype=torch.float32, device=x.device)
        L.check(self._lib.ts2d_unet_forward_debug(self._handles[fold], x.data_ptr(), n, layer.encode(), out.data_ptr(), L.stream_ptr()),
                "ts2d_unet_forward_debug")
        return out

    def close(self):
        for h in self._handles:
            self._lib.ts2d_unet_destroy(h)
        self._handles = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
