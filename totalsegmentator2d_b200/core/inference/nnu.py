"""Model-level boundary: ``NNUModel.apply`` on the device.

Mirror of reference ts2d/core/inference/nnu.py:16-241 and wrapper.py:53-162,249-280: same properties
(``channels``, ``multilabel``, ``folds``, ``revision``, ``name``), same ``apply`` contract (single in ->
single out; list -> dict ``image{i+1}``; paths in -> paths out; mixed -> RuntimeError) and the same
``nnu.*`` parameters.  What is gone is the worker process, the Manager IPC and the gzip-NRRD temp-file
round trip (nnu.py:188-221, predictor.py:42-188): the predictor lives in this process and the image stays in HBM.
"""
from __future__ import annotations

import os
import re
from typing import Dict, List, Optional, Union

import numpy as np
import torch

from ...image import Image, from_sitk
from ...nrrd import read_nrrd, write_nrrd
from ..util.log import warn
from ..util.meta import copy_image_meta, set_annotation_meta
from .database import _revision_str
from .plans import channel_names, foreground_labels, read_json
from .predictor import B200nnUNetPredictor
from .preprocess import B200Preprocessor


def _get(param: dict, key: str, default=None, dtype=None):
    v = param.get(key, default)
    if v is None or dtype is None:
        return v
    if dtype is bool and isinstance(v, str):
        return v.strip().lower() in ("1", "true", "yes", "on")
    return dtype(v)


def nnu_find_datasets(root: str) -> Dict[int, str]:
    tasks = dict()
    for dn in sorted(os.listdir(root)):
        m = re.match(r"Dataset(\d+)", dn)
        if m:
            tasks[int(m.group(1))] = dn
    return tasks


class NNUModel:
    """Interface providing access to a pretrained model for inference"""

    def __init__(self, config: dict):
        self._config = config
        self._root = str(config["root"])
        if not self.revision.isidentifier():
            raise RuntimeError('Invalid model revision "{}": no valid python identifier'.format(self.revision))
        param = dict(config.get("param") or {})
        if config.get("folds"):
            param["nnu.folds"] = config["folds"]
        self._param = param
        self.task_id = _get(param, "nnu.task", None, int)
        self.plans = _get(param, "nnu.plans", "nnUNetPlans", str)
        self.configuration = _get(param, "nnu.configuration", "3d_fullres", str)
        self.trainer = _get(param, "nnu.trainer", "nnUNetTrainer", str)
        self.checkpoint_name = _get(param, "nnu.predict.checkpoint", "final", str)
        self.augment = _get(param, "nnu.predict.augment", True, bool)
        self.stepsize = _get(param, "nnu.predict.stepsize", None, float)
        self.verbose = _get(param, "nnu.verbose", False, bool)
        self.precision = _get(param, "b200.precision", "fp16", str)
        self.use_tensor_cores = _get(param, "b200.tensor_cores", True, bool)
        self._colors = param.get("nnu.result.colors", "ts2d")
        self._configure()

    def _configure(self):
        tasks = nnu_find_datasets(self._root)
        if not tasks:
            raise RuntimeError("The specified task was not found - no existing tasks identified!")
        if self.task_id is None:
            if len(tasks) > 1:
                raise RuntimeError("The task id is ambiguous as multiple tasks are available ({}): specify 'nnu.task'".format(sorted(tasks)))
            self.task_id = list(tasks.keys())[0]
        if self.task_id not in tasks:
            raise RuntimeError("The selected task is not available: {}".format(self.task_id))
        if "2d" not in self.configuration.lower():
            raise RuntimeError("only 2-D nnU-Net configurations are supported by the B200 path (model.json says '{}')".format(self.configuration))
        self.training_dir = os.path.join(self._root, tasks[self.task_id], "__".join([self.trainer, self.plans, self.configuration]))
        self.dataset_json = read_json(os.path.join(self.training_dir, "dataset.json"))
        self.multilabel = bool(self.dataset_json.get("multilabel", self.dataset_json.get("multiclass", False)))
        fold_ids = sorted(int(p.split("_")[1]) for p in os.listdir(self.training_dir) if re.match(r"fold_[0-9]+$", p))
        folds = self._param.get("nnu.folds")
        self._folds = [int(f) for f in folds] if folds else fold_ids

    # ---- reference properties -----------------------------------------------------------------------
    @property
    def name(self):
        return self._config["model"]

    @property
    def revision(self):
        r = self._config["revision"]
        return _revision_str(r) if isinstance(r, int) else r

    @property
    def folds(self):
        return tuple(self._folds)

    @property
    def channels(self) -> Dict[int, str]:
        return channel_names(self.dataset_json)

    def get_labels(self) -> Dict[int, str]:
        """{0: 'background', 1: first label, ...} (reference wrapper.py:270-271)"""
        return dict(enumerate(self.dataset_json["labels"].keys()))

    def get_colors(self):
        colors = self._colors
        if isinstance(colors, str):
            from ..util.color import named_palette
            labels = [ln for lv, ln in sorted(self.get_labels().items())]
            colors = dict(zip(labels, named_palette(colors, len(labels))))
        return colors

    def apply(self, inputs, result_dir: Optional[str] = None, override: bool = True):
        raise NotImplementedError()


class NNUDeviceModel(NNUModel):
    """In-process replacement of the reference's NNUProcessModel."""

    def __init__(self, config: dict):
        super().__init__(config)
        self._predictor: Optional[B200nnUNetPredictor] = None

    # the reference starts a worker pool here; loading the engine is the equivalent
    def start(self, wait: bool = True):
        if self._predictor is None:
            kwargs = dict()
            if self.stepsize is not None:
                kwargs["tile_step_size"] = self.stepsize
            if self.augment is not None:
                kwargs["use_mirroring"] = self.augment
            p = B200nnUNetPredictor(verbose=self.verbose, allow_tqdm=self.verbose, precision=self.precision,
                                    use_tensor_cores=self.use_tensor_cores, **kwargs)
            p.initialize_from_trained_model_folder(self.training_dir, self.folds, f"checkpoint_{self.checkpoint_name}.pth")
            self._predictor = p

    def await_startup(self):
        self.start()

    def stop(self):
        if self._predictor is not None:
            self._predictor.close()
            self._predictor = None

    @property
    def predictor(self) -> B200nnUNetPredictor:
        self.start()
        return self._predictor

    @property
    def label_names(self) -> List[str]:
        return foreground_labels(self.dataset_json)

    def preprocess_key(self):
        """models that agree on this key share the preprocessed image and the tiles"""
        p = self.predictor
        cm = p.configuration_manager
        return (tuple(sorted(self.channels.items())), tuple(cm.normalization_schemes), tuple(cm.spacing), tuple(cm.patch_size),
                repr(p.plans_manager.foreground_intensity_properties_per_channel) if "CTNormalization" in cm.normalization_schemes else "",
                p.tile_step_size, p.mirror_mask, p.precision, cm.resampling_order_data)

    def annotate(self, seg: Image) -> Image:
        labels = self.get_labels()
        set_annotation_meta(seg, {int(k): v for k, v in labels.items()}, self.get_colors())
        return seg

    def predict_device(self, img_dev: torch.Tensor, spacing_xy, mask: Optional[torch.Tensor] = None, ch_off: int = 0,
                       present: Optional[torch.Tensor] = None, cache: Optional[dict] = None, cache_key=(),
                       prepare_only: bool = False, bbox_future=None) -> torch.Tensor:
        """(C, H, W) float32 CUDA image -> uint8 mask [H][W][ch_total] on the device (this model's channel window).
        Stage order = reference prediction_worker.py:194-221: preprocess -> predict -> export.
        ``prepare_only`` stops after preprocessing + tiling (fills ``cache``): the caller then runs the models on
        separate streams."""
        p = self.predictor
        C, H, W = img_dev.shape
        own_mask = mask is None
        if own_mask:
            mask = torch.empty((H, W, p.num_heads), dtype=torch.uint8, device=img_dev.device)
        key = self.preprocess_key() + tuple(cache_key)
        entry = cache.get(key) if cache is not None else None
        if entry is None:
            data, props = B200Preprocessor().run_case_device(img_dev, spacing_xy, p.plans_manager, p.configuration_manager, self.dataset_json,
                                                             bbox_future=bbox_future)
            x, layout = p.make_tiles(data)
            entry = (data.shape, props, x, layout)
            if cache is not None:
                cache[key] = entry
        if prepare_only:
            return mask
        shape, props, x, layout = entry
        (_, (h0, h1), (w0, w1)) = props["bbox_used_for_cropping"]
        hw = (int(shape[1]), int(shape[2]))
        direct = hw == (h1 - h0, w1 - w0)
        if not (direct and p.blend_covers_image):
            # this path only writes the crop box: clear the model's channel window first
            (mask if own_mask else mask[..., ch_off:ch_off + p.num_heads]).zero_()
        if direct:
            p.run_tiles(x, layout, hw, mask=mask, crop_origin=(h0, w0), ch_off=ch_off, present=present)
        else:
            from ... import ops
            logits = torch.empty((p.num_heads, hw[0], hw[1]), dtype=torch.float32, device=img_dev.device)
            p.run_tiles(x, layout, hw, logits=logits)
            ops.threshold_logits(logits, (h1 - h0, w1 - w0), mask, h0, w0, ch_off)
            if present is not None:
                present[ch_off:ch_off + p.num_heads] |= mask[h0:h1, w0:w1, ch_off:ch_off + p.num_heads].flatten(0, 1).any(0).to(torch.uint8)
        return mask

    def _apply_image(self, img: Image) -> Image:
        if img.GetDimension() != 2:
            raise RuntimeError("the model expects a 2-D (optionally multi-channel) image, found dimension {}".format(img.GetDimension()))
        arr = np.asarray(img.array)
        chw = np.moveaxis(arr, -1, 0) if img.is_vector else arr[None]
        dev = torch.from_numpy(np.ascontiguousarray(chw.astype(np.float32))).cuda()
        mask = self.predict_device(dev, img.GetSpacing())
        seg = Image(mask.cpu().numpy(), img.GetSpacing(), img.GetOrigin(), img.GetDirection(), is_vector=True)
        return self.annotate(seg)

    def apply(self, inputs: Union[str, Image, List, Dict], result_dir: Optional[str] = None, override: bool = True):
        single = False
        if isinstance(inputs, (str, Image)) or not isinstance(inputs, (list, tuple, dict)):
            inputs, single = [inputs], True
        if isinstance(inputs, (list, tuple)):
            inputs = dict((f"image{i + 1}", img) for i, img in enumerate(inputs))
        results, res_paths = dict(), None
        try:
            for name, img in inputs.items():
                if isinstance(img, str):
                    path_input = True
                    image = read_nrrd(img)
                    if image.GetDimension() > 2:
                        from ..util.image import reduce_dimensions
                        image = reduce_dimensions(image)
                else:
                    try:
                        image = from_sitk(img)
                    except RuntimeError:
                        raise RuntimeError("Invalid input type: {}".format(type(img)))
                    path_input = False
                if res_paths is not None and res_paths != path_input:
                    raise RuntimeError("Mixed input types are not supported, choose either images or paths!")
                res_paths = path_input
                seg = self._apply_image(image)
                if path_input or result_dir is not None:
                    odir = result_dir if result_dir is not None else os.path.dirname(os.path.abspath(img))
                    os.makedirs(odir, exist_ok=True)
                    ofile = os.path.join(odir, f"{name}.nrrd")
                    if override or not os.path.exists(ofile):
                        write_nrrd(seg, ofile, True)
                    results[name] = ofile if path_input else seg
                else:
                    results[name] = seg
        except RuntimeError as ex:
            warn(f"Prediction failed: {ex}")
            raise RuntimeError(f"Prediction failed for: {name}") from ex
        return next(iter(results.values())) if single else results
