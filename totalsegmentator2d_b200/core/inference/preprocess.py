"""``DefaultPreprocessor`` on the device.

Duck-types ``configuration_manager.preprocessor_class(verbose=).run_case(files, None, plans_manager,
configuration_manager, dataset_json) -> (data, seg, properties)`` as the reference worker consumes it
(reference ts2d/core/inference/prediction_worker.py:194-199), and adds ``run_case_device`` - the same
computation on an image that already lives in HBM (what ``TS2D.predict`` uses, so that the projected
image never travels through a temp file, cf. nnu.py:196-198).

Order of operations is upstream's (nnunetv2 default_preprocessor.py run_case_npy):
crop_to_nonzero -> normalise -> resample (order 3, per 2-D slice).
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from ... import _lib as L
from ... import ops
from ...image import Image
from ...nrrd import read_nrrd


def compute_new_shape(old_shape: Sequence[int], old_spacing: Sequence[float], new_spacing: Sequence[float]) -> List[int]:
    assert len(old_spacing) == len(old_shape) == len(new_spacing)
    return [int(round(i / j * k)) for i, j, k in zip(old_spacing, new_spacing, old_shape)]


def sitkio_spacing_2d(spacing_xy: Sequence[float]) -> List[float]:
    """SimpleITKIO.read_images for a 2-D image: (999 * max, sy, sx)."""
    sx, sy = float(spacing_xy[0]), float(spacing_xy[1])
    return [max(sx, sy) * 999.0, sy, sx]


class B200Preprocessor:
    def __init__(self, verbose: bool = False):
        self.verbose = verbose

    # ---- device path ---------------------------------------------------------------------------
    def run_case_device(self, img: torch.Tensor, spacing_xy: Sequence[float], plans_manager, configuration_manager,
                        dataset_json=None, bbox_future=None) -> Tuple[torch.Tensor, dict]:
        """img: float32 CUDA tensor (C, H, W). returns (data (C, h, w) float32 CUDA, properties).
        ``bbox_future``: (pinned int32[4], event, ...) of a crop box that was already computed for ``img`` on another stream."""
        L.load(check_device=False)
        if img.dtype != torch.float32:
            img = img.to(torch.float32)
        img = img.contiguous()
        C, H, W = img.shape
        schemes = configuration_manager.normalization_schemes
        if len(schemes) != C:
            raise RuntimeError(f"the plans define {len(schemes)} normalisation schemes but the image has {C} channels")
        if any(configuration_manager.use_mask_for_norm):
            raise RuntimeError("use_mask_for_norm=True is not supported by the B200 preprocessor")
        spacing = sitkio_spacing_2d(spacing_xy)
        props = {"spacing": spacing, "shape_before_cropping": (1, H, W)}
        if bbox_future is not None:
            bbox_future[1].synchronize()
            bbox = [int(v) for v in bbox_future[0].tolist()]
        else:
            bbox = [int(v) for v in ops.bbox_nonzero(img).tolist()]      # one 16-byte D2H per image (host sync)
        h0, h1, w0, w1 = bbox
        props["bbox_used_for_cropping"] = [[0, 1], [h0, h1], [w0, w1]]
        props["shape_after_cropping_and_before_resampling"] = (1, h1 - h0, w1 - w0)
        ids, params = [], []
        fip = plans_manager.foreground_intensity_properties_per_channel or {}
        for c, name in enumerate(schemes):
            if name not in L.SCHEME_IDS:
                raise RuntimeError("unknown normalization scheme: {}".format(name))
            ids.append(L.SCHEME_IDS[name])
            p = fip.get(str(c), fip.get(c)) if isinstance(fip, dict) else None
            if name == "CTNormalization":
                if p is None:
                    raise RuntimeError("CTNormalization needs foreground_intensity_properties_per_channel in plans.json")
                params.append([p["percentile_00_5"], p["percentile_99_5"], p["mean"], p["std"]])
            else:
                params.append([0.0, 0.0, 0.0, 1.0])
        need_stats = any(i in (L.NORM_ZSCORE, L.NORM_RESCALE01) for i in ids)
        stats = ops.channel_stats(img, bbox) if need_stats else None
        data = ops.normalize(img, bbox, ids, params, stats)
        target = list(configuration_manager.spacing)
        if len(target) < 3:
            target = [spacing[0]] + target
        new_shape = compute_new_shape((1, h1 - h0, w1 - w0), spacing, target)
        if tuple(new_shape) != (1, h1 - h0, w1 - w0):
            if new_shape[0] != 1:
                raise RuntimeError("only 2-D configurations are supported")
            data = ops.resample2d(data, new_shape[1:], configuration_manager.resampling_order_data)
        return data, props

    # ---- reference-shaped host path ----------------------------------------------------------------
    def run_case(self, image_files: List[str], seg_file, plans_manager, configuration_manager, dataset_json):
        """files -> (data float32 (C,1,h,w) numpy, None, properties), as nnU-Net returns it."""
        if isinstance(image_files, str):
            image_files = [image_files]
        imgs = [read_nrrd(f) for f in image_files]
        arrs, spacing_xy, geo = [], None, None
        for im in imgs:
            a = np.asarray(im.array)
            if im.is_vector:
                a = np.moveaxis(a, -1, 0)
            else:
                a = a[None]
            while a.ndim > 3:                      # (C, 1, H, W) containers
                sq = [i for i in range(1, a.ndim) if a.shape[i] == 1]
                if not sq:
                    raise RuntimeError("only 2-D images are supported by the 2d configuration")
                a = np.squeeze(a, axis=sq[0])
            arrs.append(a)
            sp = [s for s, n in zip(im.GetSpacing(), im.GetSize()) if n > 1] if im.GetDimension() > 2 else list(im.GetSpacing())
            spacing_xy = sp[:2]
            geo = {"spacing": im.GetSpacing(), "origin": im.GetOrigin(), "direction": im.GetDirection(), "size": im.GetSize()}
        chw = np.concatenate(arrs, axis=0).astype(np.float32)
        dev = torch.from_numpy(np.ascontiguousarray(chw)).cuda()
        data, props = self.run_case_device(dev, spacing_xy, plans_manager, configuration_manager, dataset_json)
        props["sitk_stuff"] = geo
        return data[:, None].cpu().numpy(), None, props
