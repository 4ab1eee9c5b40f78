"""``export_prediction_from_logits``: the free function of the operator seam.

The reference worker imports it from ``nnunetv2.inference.export_prediction`` and calls it right after
``predict_logits_from_preprocessed_data`` (reference ts2d/core/inference/prediction_worker.py:214-221):

    export_prediction_from_logits(prediction, data_properties, predictor.configuration_manager,
                                  predictor.plans_manager, predictor.dataset_json, ofile_truncated,
                                  task.save_probabilities)

Same positional signature here.  Upstream (nnunetv2 inference/export_prediction.py,
``convert_predicted_logits_to_segmentation_with_correct_shape``; multilabel fork): resample the logits to
``shape_after_cropping_and_before_resampling`` with ``resampling_fn_probabilities`` (order 1; no-op on equal
shapes) -> sigmoid -> > 0.5 -> one uint8 channel per foreground label -> paste into
zeros(``shape_before_cropping``) at ``bbox_used_for_cropping`` -> ``transpose_backward`` -> ``rw.write_seg``
(a vector-uint8 image with the input's geometry).  The arithmetic runs in ``ts2d_threshold_logits``
(csrc/preprocess.cu) on the device; this function only moves the tensor and writes the file.
"""
from __future__ import annotations

import pickle
from typing import Union

import numpy as np
import torch

from ... import _lib as L
from ... import ops
from ...image import Image
from .plans import foreground_labels, read_json


def convert_predicted_logits_to_segmentation_with_correct_shape(predicted_logits, plans_manager, configuration_manager, label_manager,
                                                                properties_dict: dict, return_probabilities: bool = False):
    """(heads, 1, h, w) logits -> uint8 (1, H0, W0, heads) multilabel mask [and fp32 probabilities (heads, 1, H0, W0)]."""
    L.load(check_device=True)
    if list(plans_manager.transpose_backward) != sorted(plans_manager.transpose_backward):
        raise RuntimeError("plans with a non-identity transpose_backward are not supported by the B200 path")
    lg = predicted_logits if isinstance(predicted_logits, torch.Tensor) else torch.from_numpy(np.asarray(predicted_logits))
    if lg.dim() != 4 or lg.shape[1] != 1:
        raise RuntimeError("expected logits of shape (heads, 1, h, w) for a 2-D configuration, found {}".format(tuple(lg.shape)))
    dev = lg.to(device="cuda", dtype=torch.float32)[:, 0].contiguous()
    heads = int(dev.shape[0])
    _, H0, W0 = [int(v) for v in properties_dict["shape_before_cropping"]]
    (_, (h0, h1), (w0, w1)) = properties_dict["bbox_used_for_cropping"]
    target = [int(v) for v in properties_dict["shape_after_cropping_and_before_resampling"]][1:]
    if target != [h1 - h0, w1 - w0]:
        raise RuntimeError("properties are inconsistent: cropped shape {} vs bbox {}".format(target, ((h0, h1), (w0, w1))))
    if configuration_manager.resampling_order_probabilities != 1 and tuple(dev.shape[1:]) != tuple(target):
        raise RuntimeError("only order-1 resampling of the logits is implemented (plans say order {})".format(
            configuration_manager.resampling_order_probabilities))
    mask = torch.zeros((H0, W0, heads), dtype=torch.uint8, device=dev.device)
    ops.threshold_logits(dev, target, mask, int(h0), int(w0), 0)
    seg = mask.cpu().numpy()[None]
    if not return_probabilities:
        return seg
    probs = torch.zeros((heads, 1, H0, W0), dtype=torch.float32)
    rs = dev if tuple(dev.shape[1:]) == tuple(target) else ops.resample2d(dev, target, 1)
    probs[:, 0, h0:h1, w0:w1] = torch.sigmoid(rs).cpu()
    return seg, probs.numpy()


def export_prediction_from_logits(predicted_array_or_file: Union[np.ndarray, torch.Tensor, str], properties_dict: dict, configuration_manager,
                                  plans_manager, dataset_json_dict_or_file: Union[dict, str], output_file_truncated: str,
                                  save_probabilities: bool = False):
    if isinstance(dataset_json_dict_or_file, str):
        dataset_json_dict_or_file = read_json(dataset_json_dict_or_file)
    if isinstance(predicted_array_or_file, str):
        predicted_array_or_file = np.load(predicted_array_or_file)
    dj = dataset_json_dict_or_file
    heads = len(foreground_labels(dj))
    n = int(predicted_array_or_file.shape[0])
    if n != heads:
        raise RuntimeError(f"the logits have {n} channels but dataset.json defines {heads} foreground labels")
    ret = convert_predicted_logits_to_segmentation_with_correct_shape(predicted_array_or_file, plans_manager, configuration_manager, None,
                                                                      properties_dict, return_probabilities=save_probabilities)
    if save_probabilities:
        seg, probs = ret
        np.savez_compressed(output_file_truncated + ".npz", probabilities=probs)
        with open(output_file_truncated + ".pkl", "wb") as f:
            pickle.dump(properties_dict, f)
    else:
        seg = ret
    geo = properties_dict.get("sitk_stuff") or {}
    sp, org, dr = geo.get("spacing"), geo.get("origin"), geo.get("direction")
    arr = seg[0]                                        # SimpleITKIO.write_seg: 2-D images drop the dummy axis
    if sp is not None and len(sp) == 3:                 # a (X, 1, Z)-style 3-D container was read: give the axis back
        size = [int(v) for v in geo.get("size", ())] or None
        shape = list(reversed(size)) if size else None
        if shape is None or int(np.prod(shape)) != arr.shape[0] * arr.shape[1]:
            raise RuntimeError("cannot restore the 3-D container shape of the input image")
        arr = arr.reshape(shape + [arr.shape[-1]])
    img = Image(arr, sp, org, dr, is_vector=True)
    # dataset.json's file_ending names the container (upstream: the reader/writer class is chosen by it): .nrrd, .nii(.gz), .mha/.mhd
    from ...formats import write_any
    write_any(img, output_file_truncated + dj["file_ending"])
