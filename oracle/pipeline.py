"""Oracle (test infrastructure): the whole ``TS2D.predict`` hot path on the CPU.

Stage order follows reference ts2d/tool.py:85-179 (per-model loop, projection
cache, compose, reduce, apply, merge) and
ts2d/core/inference/prediction_worker.py:177-242 (preprocess -> predict ->
export), with the stage names the reference timestamps (:200,:210,:222).
"""
from __future__ import annotations

import json
import os
import re
import time
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import geometry, nnunet_inf, nnunet_pre
from .nnunet_net import ArchSpec, conv_flops, unet_forward


class OracleModel:
    """What nnUNetPredictor.initialize_from_trained_model_folder reads from
    ``<root>/Dataset*/<trainer>__<plans>__<cfg>/`` (reference wrapper.py:113-162)."""

    def __init__(self, training_dir: str, folds: Sequence[int] = (0,), checkpoint: str = "checkpoint_final.pth",
                 configuration: str = "2d"):
        with open(os.path.join(training_dir, "dataset.json")) as f:
            self.dataset_json = json.load(f)
        with open(os.path.join(training_dir, "plans.json")) as f:
            self.plans = json.load(f)
        cfg = self.plans["configurations"][configuration]
        self.patch_size = [int(p) for p in cfg["patch_size"]]
        self.spacing = [float(s) for s in cfg["spacing"]]
        self.schemes = list(cfg["normalization_schemes"])
        self.use_mask_for_norm = list(cfg.get("use_mask_for_norm", [False] * len(self.schemes)))
        self.intensity_props = self.plans.get("foreground_intensity_properties_per_channel")
        self.channel_names = [n for _, n in sorted(((int(k), v) for k, v in self.dataset_json["channel_names"].items()))]
        labels = self.dataset_json["labels"]
        self.label_names = [k for k, v in sorted(labels.items(), key=lambda kv: kv[1] if isinstance(kv[1], int) else kv[1][0])
                            if k != "background"]
        self.num_heads = len(self.label_names)
        self.spec = ArchSpec(cfg["architecture"]["arch_kwargs"], len(self.channel_names), self.num_heads)
        self.state_dicts = []
        self.mirror_axes = None
        for f in folds:
            ck = torch.load(os.path.join(training_dir, f"fold_{f}", checkpoint), map_location="cpu", weights_only=False)
            self.state_dicts.append({k: v.float() for k, v in ck["network_weights"].items()})
            self.mirror_axes = ck.get("inference_allowed_mirroring_axes", None)

    def net(self, sd):
        return lambda x: unet_forward(x, sd, self.spec)

    def flops_per_tile(self) -> float:
        return conv_flops(self.spec, self.patch_size[0], self.patch_size[1])


def find_training_dir(model_root: str, trainer="nnUNetTrainer", plans="nnUNetPlans", configuration="2d") -> str:
    for dn in sorted(os.listdir(model_root)):
        if re.match(r"Dataset\d+", dn):
            return os.path.join(model_root, dn, "__".join([trainer, plans, configuration]))
    raise RuntimeError("no Dataset* directory under {}".format(model_root))


def predict_model(image_chw: np.ndarray, spacing_xy: Sequence[float], model: OracleModel, tile_step_size: float = 0.5,
                  use_mirroring: bool = True, mode: str = "ideal", timestamps: Optional[dict] = None,
                  return_logits: bool = False):
    """image_chw: (C, H, W) (the vector image nnU-Net reads as (C,1,H,W)).
    returns uint8 (H, W, heads) multilabel mask [and final logits (heads,1,H',W')]."""
    data = np.asarray(image_chw)[:, None]
    sp = nnunet_pre.sitkio_spacing_2d(spacing_xy)
    with torch.no_grad():
        pre, props = nnunet_pre.run_case_npy(data, sp, model.spacing, model.schemes, model.intensity_props,
                                             model.use_mask_for_norm)
        if timestamps is not None:
            timestamps["preprocessed"] = time.time()
        x = torch.from_numpy(pre)
        mirror = model.mirror_axes if use_mirroring else None
        logits = None
        for sd in model.state_dicts:
            lg = nnunet_inf.predict_sliding_window(model.net(sd), x, model.patch_size, model.num_heads,
                                                   tile_step_size, mirror, True, mode)
            logits = lg if logits is None else logits + lg
        if len(model.state_dicts) > 1:
            logits = logits / len(model.state_dicts)
        if timestamps is not None:
            timestamps["predicted"] = time.time()
        seg = nnunet_inf.export_multilabel(logits.numpy(), props, mode)
        if timestamps is not None:
            timestamps["exported"] = time.time()
    seg = seg[0]                                   # (H, W, heads)
    return (seg, logits.numpy(), props) if return_logits else seg


def project_study(vol_zyx: np.ndarray, direction, channel_names: Sequence[str]) -> np.ndarray:
    perm, flip = geometry.orientation_remap(direction, "RAI")
    rai = geometry.reorient_array(vol_zyx, perm, flip)
    return geometry.project_channels(rai, channel_names)


def combine_segmentations(segs: List[np.ndarray], names: List[List[str]]):
    """reference image.py:490-510: per model, labels present in the mask first
    (ascending, meta.py:15-49), then the absent ones in metadata order (meta.py:254-286).
    Metadata order = sitk ``GetMetaDataKeys()`` = keys of a std::map, i.e. the segment ids
    sorted by the STRING "Segment{id}_" (0, 10, 11, ..., 19, 1, 20, ...; '_' sorts after digits)."""
    chans, out_names = [], []
    for seg, nm in zip(segs, names):
        present = [i for i in range(seg.shape[-1]) if seg[..., i].any()]
        lex = sorted(range(seg.shape[-1]), key=lambda i: f"Segment{i}_".encode())
        absent = [i for i in lex if i not in present]
        for i in present + absent:
            chans.append((seg[..., i] > 0).astype(np.uint8))
            out_names.append(nm[i])
    return np.stack(chans, axis=-1), out_names


def predict_study(vol_zyx: np.ndarray, spacing_xyz, direction, models: Dict[str, OracleModel], tile_step_size=0.5,
                  use_mirroring=True, mode="ideal", stage_times: Optional[dict] = None, project_fn=None):
    """3-D study -> {'projections': (C,H,W), 'models': {id: mask}, 'segmentation': merged, 'names': [...]}
    ``project_fn(vol_rai, name) -> float32 (H, W)`` replaces the numpy projection (bench.py passes the C + OpenMP
    restatement, which stands in for ITK's multi-threaded filters when the CPU path is TIMED; results are identical)."""
    t0 = time.time()
    ids = sorted(models.keys())
    first = models[ids[0]]
    perm, flip = geometry.orientation_remap(direction, "RAI")
    rai = geometry.reorient_array(vol_zyx, perm, flip)
    cache: Dict[str, np.ndarray] = {}
    spacing_rai = [spacing_xyz[perm[k]] for k in range(3)]
    res = {"models": {}}
    t_proj = 0.0
    t_stage = {"preprocessed": 0.0, "predicted": 0.0, "exported": 0.0}
    for mid in ids:
        m = models[mid]
        ta = time.time()
        for name in m.channel_names:
            if name not in cache:
                cache[name] = project_fn(rai, name) if project_fn is not None else geometry.project_channels(rai, [name])[0]
        img = np.stack([cache[n] for n in m.channel_names], axis=0)
        tb = time.time()
        t_proj += tb - ta
        ts = {}
        seg = predict_model(img, (spacing_rai[0], spacing_rai[2]), m, tile_step_size, use_mirroring, mode, ts)
        t_stage["preprocessed"] += ts["preprocessed"] - tb
        t_stage["predicted"] += ts["predicted"] - ts["preprocessed"]
        t_stage["exported"] += ts["exported"] - ts["predicted"]
        res["models"][mid] = seg
    tc = time.time()
    merged, names = combine_segmentations([res["models"][i] for i in ids], [models[i].label_names for i in ids])
    res["segmentation"], res["names"] = merged, names
    res["projections"] = cache
    if stage_times is not None:
        stage_times.update({"projected": t_proj + 0.0, **t_stage, "merged": time.time() - tc, "total": time.time() - t0})
    return res
