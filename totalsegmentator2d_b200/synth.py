"""Synthetic nnU-Net model directories: seeded random-init weights on a canonical plan.

The published TS2D checkpoints are Zenodo downloads (reference ts2d/data/shared.json:1-34)
and there is no network here, so tests, ``smoke()`` and ``bench.py`` run on models generated
by this module.  The on-disk layout is the reference's
(``<root>/<model>_<group>/rNNN/{model.json, Dataset*/<trainer>__<plans>__<cfg>/{dataset.json,
plans.json, fold_N/checkpoint_final.pth}}``; reference database.py:127-142, wrapper.py:113-162)
and the state-dict names are upstream's, so a real checkpoint drops in unchanged.

Canonical plan CS-512: patch 512x512, spacing 1.5 mm, PlainConvUNet with 8
stages, features 32..512, two convs per stage on both sides, InstanceNorm + LeakyReLU.
"""
from __future__ import annotations

import json
import os
from collections import OrderedDict
from typing import Dict, List, Optional, Sequence

import torch

# ts2d-v2 group -> foreground label names (117 in total: 18 / 23 / 24 / 26 / 26 heads in sorted-id order).
# Names follow reference ts2d/data/label-colors.csv; the split is synthetic (the real dataset.json files are not available offline).
_VESSELS = ["aorta", "brachiocephalic-trunk", "brachiocephalic-vein-left", "brachiocephalic-vein-right",
            "common-carotid-artery-left", "common-carotid-artery-right", "iliac-artery-left", "iliac-artery-right",
            "iliac-vena-left", "iliac-vena-right", "inferior-vena-cava", "portal-vein-and-splenic-vein", "pulmonary-vein",
            "subclavian-artery-left", "subclavian-artery-right", "superior-vena-cava", "pulmonary-artery", "heart"]
_MUSCLES = ["autochthon-left", "gluteus-maximus-left", "gluteus-medius-left", "gluteus-minimus-left", "iliopsoas-left",
            "quadriceps-femoris-left", "autochthon-right", "gluteus-maximus-right", "gluteus-medius-right",
            "gluteus-minimus-right", "iliopsoas-right", "quadriceps-femoris-right",
            "skull", "humerus-left", "humerus-right", "scapula-left", "scapula-right", "clavicula-left", "clavicula-right",
            "femur-left", "femur-right", "hip-left", "hip-right"]
_ORGANS = ["lung-upper-lobe-left", "lung-lower-lobe-left", "lung-upper-lobe-right", "lung-middle-lobe-right",
           "lung-lower-lobe-right", "esophagus", "trachea", "small-bowel", "duodenum", "colon", "urinary-bladder",
           "brain", "spinal-cord", "spleen", "kidney-left", "kidney-right", "gallbladder", "liver", "stomach", "pancreas",
           "adrenal-gland-left", "adrenal-gland-right", "thyroid-gland", "prostate"]
_RIBS = [f"rib-left-{i}" for i in range(1, 13)] + [f"rib-right-{i}" for i in range(1, 13)] + ["sternum", "costal-cartilages"]
_VERTEBRAE = ([f"vertebrae-C{i}" for i in range(1, 8)] + [f"vertebrae-T{i}" for i in range(1, 13)] +
              [f"vertebrae-L{i}" for i in range(1, 6)] + ["vertebrae-S1", "sacrum"])
TS2D_V2_GROUPS: Dict[str, List[str]] = OrderedDict(
    [("cardiac", _VESSELS), ("muscles", _MUSCLES), ("organs", _ORGANS), ("ribs", _RIBS), ("vertebrae", _VERTEBRAE)])


def make_arch_kwargs(features: Sequence[int], strides: Optional[Sequence[Sequence[int]]] = None, n_conv_enc=2, n_conv_dec=2,
                     kernel=3) -> dict:
    n = len(features)
    if strides is None:
        strides = [[1, 1]] + [[2, 2]] * (n - 1)
    enc = [n_conv_enc] * n if isinstance(n_conv_enc, int) else list(n_conv_enc)
    dec = [n_conv_dec] * (n - 1) if isinstance(n_conv_dec, int) else list(n_conv_dec)
    return {
        "n_stages": n, "features_per_stage": [int(f) for f in features], "conv_op": "torch.nn.modules.conv.Conv2d",
        "kernel_sizes": [[kernel, kernel]] * n, "strides": [list(s) for s in strides],
        "n_conv_per_stage": enc, "n_conv_per_stage_decoder": dec, "conv_bias": True,
        "norm_op": "torch.nn.modules.instancenorm.InstanceNorm2d", "norm_op_kwargs": {"eps": 1e-05, "affine": True},
        "dropout_op": None, "dropout_op_kwargs": None, "nonlin": "torch.nn.LeakyReLU", "nonlin_kwargs": {"inplace": True},
    }


def cs512_arch() -> dict:
    return make_arch_kwargs([32, 64, 128, 256, 512, 512, 512, 512])


def make_plans(arch_kwargs: dict, patch_size: Sequence[int], spacing: Sequence[float], schemes: Sequence[str],
               intensity_props: Optional[dict] = None, dataset_name: str = "Dataset001_TS2D") -> dict:
    n_ch = len(schemes)
    if intensity_props is None:
        intensity_props = {str(c): {"max": 3071.0, "mean": 100.0 + 50.0 * c, "median": 80.0, "min": -1024.0,
                                    "percentile_00_5": -900.0, "percentile_99_5": 1200.0 + 300.0 * c, "std": 350.0 + 25.0 * c}
                           for c in range(n_ch)}
    return {
        "dataset_name": dataset_name, "plans_name": "nnUNetPlans",
        "original_median_spacing_after_transp": [999.0, float(spacing[0]), float(spacing[1])],
        "image_reader_writer": "SimpleITKIO", "transpose_forward": [0, 1, 2], "transpose_backward": [0, 1, 2],
        "configurations": {"2d": {
            "data_identifier": "nnUNetPlans_2d", "preprocessor_name": "DefaultPreprocessor", "batch_size": 2,
            "patch_size": [int(p) for p in patch_size], "spacing": [float(s) for s in spacing],
            "normalization_schemes": list(schemes), "use_mask_for_norm": [False] * n_ch,
            "resampling_fn_data": "resample_data_or_seg_to_shape", "resampling_fn_seg": "resample_data_or_seg_to_shape",
            "resampling_fn_data_kwargs": {"is_seg": False, "order": 3, "order_z": 0, "force_separate_z": None},
            "resampling_fn_seg_kwargs": {"is_seg": True, "order": 1, "order_z": 0, "force_separate_z": None},
            "resampling_fn_probabilities": "resample_data_or_seg_to_shape",
            "resampling_fn_probabilities_kwargs": {"is_seg": False, "order": 1, "order_z": 0, "force_separate_z": None},
            "architecture": {"network_class_name": "dynamic_network_architectures.architectures.unet.PlainConvUNet",
                             "arch_kwargs": arch_kwargs, "_kw_requires_import": ["conv_op", "norm_op", "dropout_op", "nonlin"]},
            "batch_dice": True}},
        "experiment_planner_used": "ExperimentPlanner", "label_manager": "LabelManager",
        "foreground_intensity_properties_per_channel": intensity_props,
    }


def make_dataset_json(channel_names: Sequence[str], labels: Sequence[str], file_ending: str = ".nrrd") -> dict:
    lab = OrderedDict([("background", 0)])
    for i, n in enumerate(labels):
        lab[n] = i + 1
    return {"channel_names": {str(i): n for i, n in enumerate(channel_names)}, "labels": lab, "numTraining": 0,
            "file_ending": file_ending, "multilabel": True}


def init_state_dict(arch_kwargs: dict, in_channels: int, num_heads: int, seed: int, with_aliases: bool = False) -> "OrderedDict[str, torch.Tensor]":
    """Seeded random init under the upstream parameter names.  Conv weights are Kaiming-normal
    (a=1e-2, like nnU-Net's InitWeights_He); biases and the InstanceNorm affine are drawn non-trivially
    so that every parameter is exercised by the parity tests."""
    g = torch.Generator(device="cpu")
    g.manual_seed(int(seed))
    feats = arch_kwargs["features_per_stage"]
    ks = arch_kwargs["kernel_sizes"]
    n = arch_kwargs["n_stages"]
    sd: "OrderedDict[str, torch.Tensor]" = OrderedDict()

    def conv(prefix, cout, cin, kh, kw):
        fan_in = cin * kh * kw
        std = (2.0 / ((1 + 0.01 ** 2) * fan_in)) ** 0.5
        sd[prefix + ".conv.weight"] = torch.randn((cout, cin, kh, kw), generator=g) * std
        sd[prefix + ".conv.bias"] = torch.randn((cout,), generator=g) * 0.1
        sd[prefix + ".norm.weight"] = 0.75 + 0.5 * torch.rand((cout,), generator=g)
        sd[prefix + ".norm.bias"] = torch.randn((cout,), generator=g) * 0.1
        if with_aliases:
            sd[prefix + ".all_modules.0.weight"] = sd[prefix + ".conv.weight"]
            sd[prefix + ".all_modules.0.bias"] = sd[prefix + ".conv.bias"]
            sd[prefix + ".all_modules.1.weight"] = sd[prefix + ".norm.weight"]
            sd[prefix + ".all_modules.1.bias"] = sd[prefix + ".norm.bias"]

    cin = in_channels
    for s in range(n):
        for i in range(arch_kwargs["n_conv_per_stage"][s]):
            conv(f"encoder.stages.{s}.0.convs.{i}", feats[s], cin, ks[s][0], ks[s][1])
            cin = feats[s]
    if with_aliases:
        for k in [k for k in sd.keys() if k.startswith("encoder.")]:
            sd["decoder." + k] = sd[k]
    for d in range(n - 1):
        below, skip = feats[-(d + 1)], feats[-(d + 2)]
        st = arch_kwargs["strides"][-(d + 1)]
        std = (1.0 / (below)) ** 0.5
        sd[f"decoder.transpconvs.{d}.weight"] = torch.randn((below, skip, st[0], st[1]), generator=g) * std
        sd[f"decoder.transpconvs.{d}.bias"] = torch.randn((skip,), generator=g) * 0.1
        c = 2 * skip
        k = ks[-(d + 2)]
        for i in range(arch_kwargs["n_conv_per_stage_decoder"][d]):
            conv(f"decoder.stages.{d}.convs.{i}", skip, c, k[0], k[1])
            c = skip
        sd[f"decoder.seg_layers.{d}.weight"] = torch.randn((num_heads, skip, 1, 1), generator=g) * (1.0 / skip) ** 0.5
        sd[f"decoder.seg_layers.{d}.bias"] = torch.randn((num_heads,), generator=g) * 0.1
    return sd


def write_model_dir(models_root: str, model: str, group: str, labels: Sequence[str], arch_kwargs: dict, patch_size, spacing,
                    channel_names: Sequence[str] = ("mean", "max"), schemes: Optional[Sequence[str]] = None, seed: int = 0,
                    revision: int = 1, mirror_axes=(0, 1), param: Optional[dict] = None, with_aliases: bool = False) -> str:
    """Creates ``<models_root>/<model>_<group>/r001/...`` and returns the model id."""
    schemes = list(schemes) if schemes is not None else ["ZScoreNormalization"] * len(channel_names)
    mid = f"{model}_{group}"
    root = os.path.join(models_root, mid, "r{:03d}".format(revision))
    tdir = os.path.join(root, "Dataset001_TS2D", "nnUNetTrainer__nnUNetPlans__2d")
    os.makedirs(os.path.join(tdir, "fold_0"), exist_ok=True)
    with open(os.path.join(root, "model.json"), "w") as f:
        p = {"nnu.configuration": "2d"}
        p.update(param or {})
        json.dump({"model": mid, "folds": [0], "param": p}, f, indent=1)
    with open(os.path.join(tdir, "dataset.json"), "w") as f:
        json.dump(make_dataset_json(channel_names, labels), f, indent=1)
    with open(os.path.join(tdir, "plans.json"), "w") as f:
        json.dump(make_plans(arch_kwargs, patch_size, spacing, schemes), f, indent=1)
    sd = init_state_dict(arch_kwargs, len(channel_names), len(labels), seed, with_aliases)
    torch.save({"network_weights": sd, "inference_allowed_mirroring_axes": tuple(mirror_axes) if mirror_axes is not None else None,
                "trainer_name": "nnUNetTrainer", "init_args": {"configuration": "2d", "fold": 0}},
               os.path.join(tdir, "fold_0", "checkpoint_final.pth"))
    return mid


def write_model_zoo(models_root: str, model: str = "ts2d-v2-ep4000b2", arch_kwargs: Optional[dict] = None, patch_size=(512, 512),
                    spacing=(1.5, 1.5), groups: Optional[Dict[str, List[str]]] = None, base_seed: int = 1000, **kw) -> List[str]:
    """Five group models (seeds base_seed+g), heads 18/23/24/26/26 like ts2d-v2."""
    arch_kwargs = arch_kwargs or cs512_arch()
    groups = groups or TS2D_V2_GROUPS
    ids = []
    for gi, (group, labels) in enumerate(groups.items()):
        ids.append(write_model_dir(models_root, model, group, labels, arch_kwargs, patch_size, spacing, seed=base_seed + gi, **kw))
    return ids
