"""Readers for nnU-Net's ``plans.json`` / ``dataset.json`` (the plan that drives every kernel launch).

Duck-types the small part of nnunetv2's ``PlansManager`` / ``ConfigurationManager``
that the reference touches through its predictor (reference
ts2d/core/inference/prediction_worker.py:76-78,194-199,215-221; wrapper.py:123-134):
``patch_size``, ``spacing``, ``normalization_schemes``, ``preprocessor_class``,
``foreground_intensity_properties_per_channel``, ``transpose_forward/backward``.
"""
from __future__ import annotations

import json
import os
from typing import Dict, List, Optional


def _pair(v):
    return [int(v), int(v)] if not isinstance(v, (list, tuple)) else [int(v[0]), int(v[1])]


class ConfigurationManager:
    def __init__(self, configuration: dict, name: str):
        self.configuration = configuration
        self.name = name

    @property
    def patch_size(self) -> List[int]:
        return [int(p) for p in self.configuration["patch_size"]]

    @property
    def spacing(self) -> List[float]:
        return [float(s) for s in self.configuration["spacing"]]

    @property
    def normalization_schemes(self) -> List[str]:
        return list(self.configuration["normalization_schemes"])

    @property
    def use_mask_for_norm(self) -> List[bool]:
        return list(self.configuration.get("use_mask_for_norm", [False] * len(self.normalization_schemes)))

    @property
    def resampling_order_data(self) -> int:
        return int((self.configuration.get("resampling_fn_data_kwargs") or {}).get("order", 3))

    @property
    def resampling_order_probabilities(self) -> int:
        return int((self.configuration.get("resampling_fn_probabilities_kwargs") or {}).get("order", 1))

    @property
    def arch_kwargs(self) -> dict:
        arch = self.configuration.get("architecture")
        if arch is None:
            raise RuntimeError("plans.json has no 'architecture' entry (nnU-Net >= 2.3 plans are required)")
        cls = str(arch.get("network_class_name", ""))
        if not cls.endswith("PlainConvUNet"):
            raise RuntimeError("unsupported network class '{}': only PlainConvUNet is implemented".format(cls))
        return arch["arch_kwargs"]

    @property
    def preprocessor_class(self):
        from .preprocess import B200Preprocessor
        return B200Preprocessor


class PlansManager:
    def __init__(self, plans: dict):
        self.plans = plans

    @property
    def transpose_forward(self) -> List[int]:
        return [int(v) for v in self.plans.get("transpose_forward", [0, 1, 2])]

    @property
    def transpose_backward(self) -> List[int]:
        return [int(v) for v in self.plans.get("transpose_backward", [0, 1, 2])]

    @property
    def foreground_intensity_properties_per_channel(self) -> Optional[Dict[str, Dict[str, float]]]:
        return self.plans.get("foreground_intensity_properties_per_channel")

    def get_configuration(self, name: str) -> ConfigurationManager:
        cfgs = self.plans["configurations"]
        if name not in cfgs:
            raise RuntimeError("configuration '{}' is not part of the plans (available: {})".format(name, ", ".join(cfgs)))
        cfg = dict(cfgs[name])
        seen = {name}
        while "inherits_from" in cfg:          # nnU-Net lets a configuration inherit from another
            parent = cfg.pop("inherits_from")
            if parent in seen:
                raise RuntimeError("circular inherits_from in plans.json")
            seen.add(parent)
            base = dict(cfgs[parent])
            base.update(cfg)
            cfg = base
        return ConfigurationManager(cfg, name)


def foreground_labels(dataset_json: dict) -> List[str]:
    """label names in ascending label value without background: channel c of the multilabel
    output <-> dataset label c+1 (reference wrapper.py:270-271, meta.py:213)."""
    labels = dataset_json["labels"]

    def key(kv):
        v = kv[1]
        return int(v) if not isinstance(v, (list, tuple)) else int(v[0])
    vals = [key(kv) for kv in labels.items()]
    # the reference names channel c by the (c+1)-th KEY of dataset.json (wrapper.py:270-271: dict(enumerate(labels.keys()))),
    # nnU-Net orders the heads by label VALUE: both agree only when the file lists background = 0 first and the values ascend
    if vals != list(range(len(vals))) or next(iter(labels)) != "background":
        raise RuntimeError("dataset.json labels must be listed as background = 0 followed by 1, 2, ... in ascending order "
                           "(found values {})".format(vals))
    return [k for k, v in labels.items() if k != "background"]


def channel_names(dataset_json: dict) -> Dict[int, str]:
    return dict((int(str(c).strip()), n) for c, n in dataset_json["channel_names"].items())


def read_json(path: str) -> dict:
    with open(path, "r") as f:
        return json.load(f)
