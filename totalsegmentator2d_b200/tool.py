"""Public API: ``TS2D(key).predict(input) -> TS2D.Result`` / ``Result.save(...)``.

Drop-in for reference ts2d/tool.py:19-311 (same constructor, context manager, ``predict`` signature, result
dictionary keys, ``Result`` accessors, ``save`` arguments, file naming and error messages).  What changed is
underneath: the volume is uploaded once, reoriented + projected by one kernel (instead of DICOMOrient per model
and one ITK filter per channel, tool.py:154-159), preprocessed once for all models that share a plan, and the
five group models write straight into one merged mask in HBM (instead of five subprocess round trips through
gzip-NRRD temp files and 117 SimpleITK calls in combine_segmentations, image.py:490-510).
"""
from __future__ import annotations

import os
import traceback
from typing import Dict, List, Optional

import numpy as np
import torch

from . import _lib as L
from . import ops
from .core.inference.database import decompose_model_key
from .core.inference.nnu import NNUDeviceModel, NNUModel
from .core.inference.zoo import NNUZoo
from .core.util.config import get_label_colors
from .core.util.image import (_projected_image, _reoriented_geometry, _to_device, axis_name_to_index, combine_segmentations, compose,
                              get_actual_dimension, merge_order, orientation_remap, read_image, reduce_dimensions, split_channels)
from .core.util.log import log, warn
from .core.util.meta import copy_image_geo, copy_image_meta, set_annotation_meta
from .image import Image, from_sitk
from .nrrd import write_nrrd


def _as_list(v):
    if v is None:
        return []
    return list(v) if isinstance(v, (list, tuple, set)) else [v]


class _LazySeg:
    """per-model segmentation that is still in HBM; ``Result.get_segmentation(model)`` brings it to the host"""

    def __init__(self, mask_dev: torch.Tensor, lo: int, hi: int, finish):
        self.mask_dev, self.lo, self.hi, self.finish = mask_dev, lo, hi, finish
        self._img: Optional[Image] = None

    def materialize(self) -> Image:
        if self._img is None:
            arr = self.mask_dev[..., self.lo:self.hi].contiguous().cpu().numpy()
            self._img = self.finish(arr)
            self.mask_dev = None
        return self._img


class TS2D:
    def __init__(self, key: str = "ts2d", use_remote: bool = True, fetch_remote: bool = True, models_root: Optional[str] = None,
                 param: Optional[dict] = None):
        """
        :param key: the model key to use, defaults to "ts2d"
        :param use_remote / fetch_remote: accepted for compatibility; models are resolved from the local database only
               (``~/.ts2d/models`` or ``$TS2D_MODELS``) because downloading is outside the accelerated path
        :param models_root: overrides the local database root
        :param param: extra ``nnu.*`` / ``b200.*`` parameters merged over each model.json
        """
        colors = get_label_colors()
        p = {"server.workers": 1, "nnu.result.colors": colors}
        p.update(param or {})
        self._param = dict(p)
        L.load(check_device=True)       # fail loudly: no CPU fallback
        self.zoo = NNUZoo(remote=False, root=models_root)
        self.models: Dict[str, NNUDeviceModel] = dict()
        ids = self.zoo.resolve(key, unique_model=True)
        if not ids:
            raise RuntimeError(f"No models were resolved for key: {key}")
        if len(ids) > 1:
            log(f"The model key '{key}' was resolved to {len(ids)} models: {', '.join(ids)}.")
        for id in ids:
            try:
                model = self.zoo.load(id, interface="process", param=p)
                model.start(wait=False)
                if not model.multilabel:
                    warn(f"The loaded model {id} is not configured for multilabel inference - this should not be the case in TS2D and may lead to unexpected results.")
                self.models[id] = model
            except Exception:
                traceback.print_exc()
                raise RuntimeError(f"Failed to load model {id}" + (f" (resolved from {key})" if key != id else ""))
        for model in self.models.values():
            model.await_startup()

    def __enter__(self):
        return self

    def __exit__(self, exc_type, exc_val, exc_tb):
        self.close()

    def close(self):
        for model in self.models.values():
            model.stop()
        self.models = dict()

    def __del__(self):
        try:
            if getattr(self, "models", None):
                self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------------------------------------
    # device-level core (what ``predict`` runs between the H2D copy of the input and the D2H copy of the mask)
    def _channel_defs(self):
        defs = dict()
        for id, model in self.models.items():
            ch = model.channels
            if ch is None:
                raise RuntimeError(f"Model {id} does not have a channel definition, cannot project the input image.")
            defs[id] = sorted(ch.items(), key=lambda x: x[0])          # reference tool.py:145-148
        return defs

    def head_offsets(self):
        heads = [self.models[id].predictor.num_heads for id in self.models]
        return [int(v) for v in np.concatenate([[0], np.cumsum(heads)])]

    def predict_image_device(self, imgs_dev: Dict[str, torch.Tensor], spacing_xy, mask: Optional[torch.Tensor] = None,
                             present: Optional[torch.Tensor] = None, bbox_futures: Optional[dict] = None) -> dict:
        """imgs_dev: {model id: float32 CUDA (C, H, W)} -> {'mask': uint8 [H][W][sum heads] (model-id order, label order),
        'present': uint8 [sum heads], 'offsets': channel window of each model}.  Everything stays in HBM.
        Models whose preprocessing agrees share the preprocessed image and its tiles (the reference preprocesses
        per model inside each worker, prediction_worker.py:194-199)."""
        ids = list(self.models.keys())
        offs = self.head_offsets()
        first = imgs_dev[ids[0]]
        H, W = int(first.shape[1]), int(first.shape[2])
        # no memset: every model writes its whole channel window, zeros outside its crop box included (ts2d_unet_blend_covers_image;
        # the models that cannot - fp32 engines, folds, resampled export - clear their own window in predict_device)
        if mask is None:
            mask = torch.empty((H, W, offs[-1]), dtype=torch.uint8, device=first.device)
        if present is None:
            present = torch.zeros(offs[-1], dtype=torch.uint8, device=first.device)
        else:
            present.zero_()
        cache = dict()
        side = self._model_streams(len(ids))
        try:
            if not side:
                for k, id in enumerate(ids):
                    key = (imgs_dev[id].data_ptr(),)
                    self.models[id].predict_device(imgs_dev[id], spacing_xy, mask=mask, ch_off=offs[k], present=present, cache=cache, cache_key=key,
                                                   bbox_future=(bbox_futures or {}).get(imgs_dev[id].data_ptr()))
            else:
                # the group models of one study are independent (reference tool.py:110-112 loops over them): preprocess and
                # tile once on the caller's stream, then fan the U-Nets out over side streams so that the latency-bound
                # low-resolution layers of one model fill the SMs the other models leave idle; join before returning
                # (every tensor the side streams touch is still referenced here, so the caching allocator cannot recycle it)
                cur = torch.cuda.current_stream()
                for k, id in enumerate(ids):
                    key = (imgs_dev[id].data_ptr(),)
                    self.models[id].predict_device(imgs_dev[id], spacing_xy, mask=mask, ch_off=offs[k], present=present, cache=cache,
                                                   cache_key=key, prepare_only=True, bbox_future=(bbox_futures or {}).get(imgs_dev[id].data_ptr()))
                fork = torch.cuda.Event()
                fork.record(cur)
                joins = []
                for k, id in enumerate(ids):
                    s = side[k % len(side)]
                    if k < len(side):
                        s.wait_event(fork)
                    with torch.cuda.stream(s):
                        key = (imgs_dev[id].data_ptr(),)
                        self.models[id].predict_device(imgs_dev[id], spacing_xy, mask=mask, ch_off=offs[k], present=present, cache=cache,
                                                       cache_key=key)
                for s in side[:len(ids)]:
                    ev = torch.cuda.Event()
                    ev.record(s)
                    cur.wait_event(ev)
        except L.TS2DError as ex:
            raise RuntimeError(f"Prediction failed for: {id} ({ex})") from ex
        return {"mask": mask, "present": present, "offsets": offs}

    def _model_streams(self, n_models: int):
        """side streams the group models of one study are spread over (``b200.model_streams`` / $TS2D_MODEL_STREAMS;
        0 or 1 = everything on the caller's stream)."""
        want = int(os.environ.get("TS2D_MODEL_STREAMS", self._param.get("b200.model_streams", 3)))
        want = min(want, n_models)
        if want <= 1:
            return []
        pool = getattr(self, "_side_streams", None)
        if pool is None or len(pool) != want:
            pool = self._side_streams = [torch.cuda.Stream() for _ in range(want)]
        return pool

    def project_volume_device(self, vol_dev: torch.Tensor, spacing_xyz, direction=None, bbox_async: bool = False) -> dict:
        """vol_dev: CUDA tensor [z][y][x] (sitk.GetArrayFromImage order) -> every projection channel any model asks for, from
        ONE pass over the volume (reference: one ITK filter per channel and model, tool.py:156-159); reorient_image('RAI')
        (tool.py:154) is folded into the kernel's addressing.  With ``bbox_async`` the crop_to_nonzero bounding box of each
        model input is computed right away and copied to pinned host memory without waiting (the caller synchronises on the
        event when it needs the numbers), so that no host sync sits between the upload and the U-Nets."""
        names = self._projection_names()
        d = np.eye(3) if direction is None else np.asarray(direction, dtype=np.float64).reshape(3, 3)
        perm, flip = orientation_remap(d, "RAI")
        ax = axis_name_to_index("coronal")
        out = ops.project(vol_dev, names, reduce_dim=2 - perm[ax], h_dim=2 - perm[2], flip_h=flip[2], flip_w=flip[0])
        imgs = self._model_inputs(out, names)
        sp = [float(spacing_xyz[perm[k]]) for k in range(3)]
        st = {"projections": out, "channel_names": names, "imgs": imgs, "spacing_xy": (sp[0], sp[2]), "bbox_futures": None}
        if bbox_async:
            st["bbox_futures"] = self._bbox_futures({t.data_ptr(): t for t in imgs.values()}.values())
        return st

    def _projection_names(self) -> List[str]:
        """every projection channel any model asks for, in first-use order (reference tool.py:145-159)"""
        chan_defs = self._channel_defs()
        names: List[str] = []
        for id in self.models:
            for _, n in chan_defs[id]:
                if n not in names:
                    names.append(n)
        return names

    def _model_inputs(self, out: torch.Tensor, names: List[str]) -> Dict[str, torch.Tensor]:
        """per model the channels it asks for, ascending channel index (reference tool.py:148,160); models that agree share the tensor"""
        chan_defs = self._channel_defs()
        imgs, stacked = dict(), dict()
        for id in self.models:
            idx = tuple(names.index(n) for _, n in chan_defs[id])
            if idx not in stacked:
                stacked[idx] = out if idx == tuple(range(len(names))) else torch.stack([out[i] for i in idx], dim=0)
            imgs[id] = stacked[idx]
        return imgs

    @staticmethod
    def _bbox_futures(images) -> dict:
        futures = dict()
        for img in images:
            bb = ops.bbox_nonzero(img.contiguous())
            host = torch.empty(4, dtype=torch.int32, pin_memory=True)
            host.copy_(bb, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record()
            futures[img.data_ptr()] = (host, ev, bb)
        return futures

    def predict_volume_device(self, vol_dev: torch.Tensor, spacing_xyz, direction=None, mask: Optional[torch.Tensor] = None,
                              present: Optional[torch.Tensor] = None, projected: Optional[dict] = None) -> dict:
        """vol_dev: CUDA tensor [z][y][x] -> projections + merged mask, all in HBM (``projected``: the result of an earlier
        ``project_volume_device`` call for this volume, e.g. issued on the upload stream)."""
        st = projected if projected is not None else self.project_volume_device(vol_dev, spacing_xyz, direction)
        res = self.predict_image_device(st["imgs"], st["spacing_xy"], mask=mask, present=present, bbox_futures=st["bbox_futures"])
        res["projections"], res["channel_names"] = st["projections"], st["channel_names"]
        return res

    def predict_volume_sharded(self, vol_dev: Optional[torch.Tensor], spacing_xyz=None, direction=None, home: int = 0) -> Optional[dict]:
        """Latency mode (collective: every rank of the process group calls it): ONE study spread over the GPUs.
        The home rank holds the volume and projects it; the 4 MB projected image is broadcast; rank r runs the group models
        ``shard.owned_models`` assigns to it (they are independent, reference tool.py:110-112); the per-model label masks and
        present flags are gathered on the home rank (the path's only exchange step), which merges them in the reference's
        channel order.  Returns the same dict as ``predict_volume_device`` plus 'merged' / 'perm' on the home rank, None elsewhere.
        With one process it degenerates to ``predict_volume_device``."""
        from . import shard
        rank, world, _ = shard.rank_world()
        import torch.distributed as dist
        if world == 1 or not dist.is_initialized():
            res = self.predict_volume_device(vol_dev, spacing_xyz, direction)
            res["merged"], res["perm"] = ops.merge_present_order(res["mask"], res["present"], res["offsets"])
            return res
        ids = list(self.models.keys())
        offs = self.head_offsets()
        widths = [offs[k + 1] - offs[k] for k in range(len(ids))]
        st = None
        if rank == home:
            st = self.project_volume_device(vol_dev, spacing_xyz, direction)
            out, meta = shard.broadcast_image(st["projections"], home, meta=st["spacing_xy"])
        else:
            out, meta = shard.broadcast_image(None, home)
        names = st["channel_names"] if st is not None else self._projection_names()
        imgs = self._model_inputs(out, names)
        H, W = int(out.shape[1]), int(out.shape[2])
        local_masks, local_present = dict(), dict()
        cache = dict()
        for k in shard.owned_models(len(ids), rank, world, home):
            id = ids[k]
            m = torch.zeros((H, W, widths[k]), dtype=torch.uint8, device=out.device)
            pr = torch.zeros(widths[k], dtype=torch.uint8, device=out.device)
            try:
                self.models[id].predict_device(imgs[id], (meta[0], meta[1]), mask=m, ch_off=0, present=pr, cache=cache,
                                               cache_key=(imgs[id].data_ptr(),))
            except L.TS2DError as ex:
                raise RuntimeError(f"Prediction failed for: {id} ({ex})") from ex
            local_masks[k], local_present[k] = m, pr
        mask = shard.gather_windows(local_masks, len(ids), widths, (H, W), home, dtype=torch.uint8)
        present = shard.gather_windows(local_present, len(ids), widths, (), home, dtype=torch.uint8)
        if rank != home:
            return None
        merged, perm = ops.merge_present_order(mask, present, offs)
        return {"mask": mask, "present": present, "offsets": offs, "projections": out, "channel_names": names, "merged": merged, "perm": perm}

    # ------------------------------------------------------------------------------------------------
    def predict(self, input, collapse: bool = False, merge: bool = True):
        """
        Predict the segmentation for the given input image.
        :param input: the input image (Image / sitk.Image) or a path to an image file
        :param collapse: collapse the output image and segmentation to 2D (discards the 3D orientation)
        :param merge: merge the individual model segmentations into a single segmentation image
        """
        return self._finish(self._submit(input, collapse, merge))

    def predict_batch(self, inputs, collapse: bool = False, merge: bool = True, depth: int = 2, lookahead: int = 1):
        """Generator over ``predict`` results for a sequence of inputs (the reference CLI loops over its inputs one by one,
        ts2d/main.py:80-83).  Software pipeline over CUDA streams: input i+1.. is uploaded and projected (and its crop box is
        on its way to the host) while study i computes and the mask of study i-1 is downloaded and wrapped on the host; the
        host never waits for the GPU between the upload and the U-Nets.  Results come out in input order."""
        staged, pending = [], []
        failure = None
        for inp in inputs:
            try:
                staged.append(self._stage_upload(inp, collapse, merge))
            except Exception as ex:       # an unreadable input: the studies already in flight are finished and yielded first
                failure = ex
                break
            if len(staged) > max(0, lookahead):
                pending.append(self._stage_compute(staged.pop(0)))
            if len(pending) >= max(1, depth):
                yield self._finish(pending.pop(0))
        while staged:
            pending.append(self._stage_compute(staged.pop(0)))
            if len(pending) >= max(1, depth):
                yield self._finish(pending.pop(0))
        while pending:
            yield self._finish(pending.pop(0))
        if failure is not None:
            raise failure

    def _streams(self):
        if getattr(self, "_h2d_stream", None) is None:
            self._h2d_stream = torch.cuda.Stream()
            # ``b200.copy_streams`` / $TS2D_COPY_STREAMS: 2 = uploads and downloads on separate streams (full duplex), 1 = one copy
            # stream (a mask download never overlaps a volume upload), auto (default) = 1 when >= 4 ranks share the host.  Measured on
            # the 8-GPU box (profiles/r02_h2d_ceiling_n8.json): with all ranks uploading, the slow ranks move 23.3 GB/s host -> device
            # alone but 18.6 GB/s while 60 MB masks go the other way at the same time - 584 MB per study in 25.1 ms serialised against
            # 28.2 ms overlapped; alone on the host a GPU has PCIe bandwidth to spare and full duplex costs nothing.
            want = str(os.environ.get("TS2D_COPY_STREAMS", self._param.get("b200.copy_streams", "auto"))).lower()
            if want == "auto":
                want = "1" if int(os.environ.get("WORLD_SIZE", "1")) >= 4 else "2"
            self._d2h_stream = self._h2d_stream if want == "1" else torch.cuda.Stream()
        return self._h2d_stream, self._d2h_stream

    def _pack_mask(self) -> bool:
        """``b200.pack_mask`` / $TS2D_PACK_MASK (default 0): download the merged mask bit-packed (7.5 MB instead of 59.9 MB) and unpack it
        on the host (~13 ms of one core per study): pays on hosts whose PCIe path is shared by >= 4 uploading GPUs, costs latency alone."""
        return str(os.environ.get("TS2D_PACK_MASK", self._param.get("b200.pack_mask", "0"))).lower() in ("1", "true", "on")

    @staticmethod
    def _pinned_like(t: torch.Tensor) -> torch.Tensor:
        return torch.empty(t.shape, dtype=t.dtype, pin_memory=True)

    def _submit(self, input, collapse: bool, merge: bool) -> dict:
        """enqueues upload, device pipeline and download of one input"""
        return self._stage_compute(self._stage_upload(input, collapse, merge))

    def _stage_upload(self, input, collapse: bool, merge: bool) -> dict:
        """upload stream: H2D copy of the input, projection (3-D inputs) and the crop box of every model input; nothing waits"""
        if isinstance(input, str):
            input = read_image(input)
        try:
            input = from_sitk(input)
        except RuntimeError:
            raise RuntimeError(f"input must be a string path or a SimpleITK image, found: {type(input).__name__}")
        ids = list(self.models.keys())
        is3d = get_actual_dimension(input) > 2
        chan_defs = self._channel_defs()
        h2d, d2h = self._streams()
        job = {"input": input, "collapse": collapse, "merge": merge, "is3d": is3d, "chan_defs": chan_defs, "keep": []}
        if is3d:
            with torch.cuda.stream(h2d):
                vol = _to_device(input.array)
                job["projected"] = self.project_volume_device(vol, input.GetSpacing(), input.direction_matrix(), bbox_async=True)
                ev = torch.cuda.Event()
                ev.record(h2d)
            job["vol"] = vol
        else:
            base2d = input if input.GetDimension() < 3 else reduce_dimensions(input)
            input_nch = input.GetNumberOfComponentsPerPixel()
            for id in ids:
                if len(chan_defs[id]) != input_nch:
                    raise RuntimeError(f"The number of channels in the input image does not match the models channel definition ({len(chan_defs[id])} vs {input_nch}).")
            arr = np.asarray(base2d.array, dtype=np.float32)
            chw = np.moveaxis(arr, -1, 0) if base2d.is_vector else arr[None]
            with torch.cuda.stream(h2d):
                img_dev_2d = torch.from_numpy(np.ascontiguousarray(chw)).cuda(non_blocking=True)
                job["bbox_futures"] = self._bbox_futures([img_dev_2d])
                ev = torch.cuda.Event()
                ev.record(h2d)
            job["img2d"] = img_dev_2d
            job["base2d"] = base2d
        job["uploaded"] = ev
        return job

    def _stage_compute(self, job: dict) -> dict:
        """compute stream: the five U-Nets, blend / threshold and the merged channel order; download stream: results to pinned memory"""
        input, merge, is3d = job["input"], job["merge"], job["is3d"]
        ids = list(self.models.keys())
        h2d, d2h = self._streams()
        cur = torch.cuda.current_stream()
        cur.wait_event(job.pop("uploaded"))
        if is3d:
            vol, st = job.pop("vol"), job.pop("projected")
            vol.record_stream(cur)
            for t in [st["projections"]] + list(st["imgs"].values()) + [f[2] for f in (st["bbox_futures"] or {}).values()]:
                t.record_stream(cur)
            dev = self.predict_volume_device(vol, input.GetSpacing(), input.direction_matrix(), projected=st)
            del vol
        else:
            img_dev_2d = job.pop("img2d")
            img_dev_2d.record_stream(cur)
            futs = job.pop("bbox_futures")
            for f in futs.values():
                f[2].record_stream(cur)
            dev = self.predict_image_device({id: img_dev_2d for id in ids}, job["base2d"].GetSpacing(), bbox_futures=futs)
        mask, present, offs = dev["mask"], dev["present"], dev["offsets"]
        job.update(mask=mask, offs=offs, channel_names=dev.get("channel_names"))
        merged_dev = perm_dev = None
        if merge and len(ids) > 1:
            # channel order of combine_segmentations (present labels first) computed and applied on the device
            merged_dev, perm_dev = ops.merge_present_order(mask, present, offs)
        ev_c = torch.cuda.Event()
        ev_c.record(cur)
        with torch.cuda.stream(d2h):
            d2h.wait_event(ev_c)
            if is3d:
                job["proj_host"] = self._pinned_like(dev["projections"])
                job["proj_host"].copy_(dev["projections"], non_blocking=True)
                dev["projections"].record_stream(d2h)
            if merged_dev is not None:
                if self._pack_mask():
                    # 8 label bytes -> 1 byte on the device, 7.5 MB instead of 59.9 MB over PCIe; _finish unpacks on the host
                    packed = ops.pack_mask_bits(merged_dev)
                    job["merged_packed_host"] = self._pinned_like(packed)
                    job["merged_packed_host"].copy_(packed, non_blocking=True)
                    job["merged_shape"] = tuple(merged_dev.shape)
                    packed.record_stream(d2h)
                else:
                    job["merged_host"] = self._pinned_like(merged_dev)
                    job["merged_host"].copy_(merged_dev, non_blocking=True)
                job["perm_host"] = self._pinned_like(perm_dev)
                job["perm_host"].copy_(perm_dev, non_blocking=True)
                merged_dev.record_stream(d2h)
                perm_dev.record_stream(d2h)
            elif merge:
                lo, hi = int(offs[0]), int(offs[1])
                single = mask[..., lo:hi].contiguous() if (hi - lo) != mask.shape[-1] else mask
                job["merged_host"] = self._pinned_like(single)
                job["merged_host"].copy_(single, non_blocking=True)
                single.record_stream(d2h)
            ev_d = torch.cuda.Event()
            ev_d.record(d2h)
        job["done"] = ev_d
        return job

    def _finish(self, job: dict):
        """waits for the download of one submitted input and wraps it like reference tool.py:108-129"""
        input, collapse, merge, is3d, chan_defs = job["input"], job["collapse"], job["merge"], job["is3d"], job["chan_defs"]
        ids = list(self.models.keys())
        mask, offs = job["mask"], job["offs"]
        job["done"].synchronize()
        if "merged_packed_host" in job:
            # into a pinned buffer of torch's caching host allocator: no page faults on 60 MB of fresh memory per study
            shape = job.pop("merged_shape")
            host = torch.empty(shape, dtype=torch.uint8, pin_memory=True)
            pk = job.pop("merged_packed_host")
            L.check(L.load().ts2d_unpack_mask_bits_host(pk.data_ptr(), host.data_ptr(), host.numel(), 4), "ts2d_unpack_mask_bits_host")
            job["merged_host"] = host
            job["d2h_bytes"] = int(pk.numel())
        result = dict()
        projections: Dict[str, Image] = dict()
        if is3d:
            perm, flip = orientation_remap(input.direction_matrix(), "RAI")
            size, sp, org, nd = _reoriented_geometry(input, perm, flip)
            ax = axis_name_to_index("coronal")
            host = job["proj_host"].numpy()
            for i, n in enumerate(job["channel_names"]):
                projections[n] = _projected_image(host[i], size, sp, org, nd, ax)
        else:
            chs = split_channels(input)
            projections.update((f"ch{i}", ch) for i, ch in enumerate(chs))

        model_inputs = dict()
        composed = dict()
        for k, id in enumerate(ids):
            model = self.models[id]
            res = {"id": id, "revision": model.revision}
            res["model"], res["group"] = decompose_model_key(id)
            if is3d:
                names = tuple(n for _, n in chan_defs[id])
                if names not in composed:      # models with the same channel definition share one composed input image
                    ch_list = [projections[n] for n in names]
                    composed[names] = compose(ch_list) if len(ch_list) > 1 else ch_list[0]
                minput = composed[names]
            else:
                minput = input
            native_2d = minput.GetDimension() < 3
            input2d = minput if native_2d else reduce_dimensions(minput)
            model_inputs[id] = (minput, input2d, native_2d)
            res["input"] = input2d if collapse else minput
            result.setdefault("models", dict())[id] = res

        def finisher(id):
            minput, input2d, native_2d = model_inputs[id]

            def fin(arr):
                seg = Image(arr, input2d.GetSpacing(), input2d.GetOrigin(), input2d.GetDirection(), is_vector=True)
                self.models[id].annotate(seg)
                return seg if collapse or native_2d else self._restore_dimension(seg, minput)
            return fin

        for k, id in enumerate(ids):
            result["models"][id]["segmentation"] = _LazySeg(mask, int(offs[k]), int(offs[k + 1]), finisher(id))

        if merge:
            if len(ids) == 1:
                seg = result["models"][ids[0]]["segmentation"]
                seg._img = seg.finish(job["merged_host"].numpy())
                seg.mask_dev = None
                result["segmentation"] = seg._img
            else:
                perm = [int(v) for v in job["perm_host"].numpy()]
                label_of, color_of = dict(), dict()
                for k, id in enumerate(ids):
                    m = self.models[id]
                    lab = m.label_names
                    col = m.get_colors() or {}
                    for c in range(len(lab)):
                        label_of[int(offs[k]) + c] = lab[c]
                        if col.get(lab[c]) is not None:
                            color_of[lab[c]] = col[lab[c]]
                names = dict((j + 1, label_of[c]) for j, c in enumerate(perm))
                colors = dict((n, color_of[n]) for n in names.values() if n in color_of)
                minput, input2d, native_2d = model_inputs[ids[0]]
                seg = Image(job["merged_host"].numpy(), input2d.GetSpacing(), input2d.GetOrigin(), input2d.GetDirection(), is_vector=True)
                set_annotation_meta(seg, names=names, colors=colors)
                result["segmentation"] = seg if collapse or native_2d else self._restore_dimension(seg, minput)

        result["input"] = input
        if projections:
            result["projections"] = projections
        if "d2h_bytes" in job:
            result["d2h_bytes"] = job["d2h_bytes"]
        return TS2D.Result(result)

    @staticmethod
    def _project(img: Image, mode: str) -> Image:
        from .core.util.image import project
        res = project(img, mode=mode, axis="coronal")
        return Image(res.array.astype(np.float32), res.GetSpacing(), res.GetOrigin(), res.GetDirection())

    @staticmethod
    def _restore_dimension(img: Image, ref: Image) -> Image:
        shape = list(reversed(ref.GetSize())) + [-1]
        out = Image(np.reshape(img.array, shape), ref.GetSpacing(), ref.GetOrigin(), ref.GetDirection(), is_vector=True)
        copy_image_meta(out, img)
        return out

    # ------------------------------------------------------------------------------------------------
    class Result:
        def __init__(self, data: dict):
            self.data = data

        @property
        def models(self) -> List[str]:
            return sorted(self.data.get("models", dict()).keys())

        def get_input(self, model: Optional[str] = None):
            if model is not None:
                return self.data.get("models", dict()).get(model, dict()).get("input")
            return self.data.get("input")

        def get_segmentation(self, model: Optional[str] = None):
            if model is not None:
                entry = self.data.get("models", dict()).get(model, dict())
                seg = entry.get("segmentation")
                if isinstance(seg, _LazySeg):
                    seg = seg.materialize()
                    entry["segmentation"] = seg
                return seg
            return self.data.get("segmentation")

        def get_projection(self, channel: Optional[str] = None):
            projections = self.data.get("projections", dict())
            if channel is not None:
                return projections.get(channel)
            return projections

        def save(self, dest: str, name: str = "result", ext: str = "nrrd", models="final", targets="all", content: str = "all",
                 naming: str = "group"):
            assert ext.lower() != "png", "PNG is not a valid export format for the 'file' content type."
            assert naming in {"group", "model"}, f"Invalid naming scheme '{naming}', must be one of 'group' or 'model'."
            assert content in {"file", "visual", "all"}, f"Invalid export type '{content}', must be one of 'file', 'visual' or 'all'."
            content = {"visual", "file"} if content == "all" else {content}
            models = set(t.strip().lower() if isinstance(t, str) else t for t in _as_list(models))
            if "all" in models:
                models |= set(self.models) | {None}
            if "final" in models:
                models |= {None}
            models -= {"all", "final"}
            targets = set(t.strip().lower() for t in _as_list(targets))

            def _make_filename(base, key):
                if key is not None and naming == "group":
                    _, group = decompose_model_key(key)
                    return f"{base}-{group}"
                return base if key is None else f"{base}-{key}"

            def _write(img, filepath):
                if ext.lower() == "nrrd":
                    write_nrrd(img, filepath, True)
                elif ext.lower() in ("nii", "nii.gz", "mha", "mhd"):
                    from .formats import write_any
                    write_any(img, filepath)
                else:
                    raise RuntimeError(f"unsupported export format '{ext}' (supported: nrrd, nii, nii.gz, mha, mhd)")

            def _export_image(img, base_name, suffix="", labels=False):
                if "file" in content:
                    _write(img, os.path.join(dest, f"{base_name}{suffix}.{ext}"))
                if "visual" in content:
                    from .core.util.visual import write_visual
                    write_visual(img, os.path.join(dest, base_name), suffix, labels)

            os.makedirs(dest, exist_ok=True)
            ordered = sorted(models, key=lambda k: (k is not None, str(k)))
            if {"all", "input"} & targets:
                for key in ordered:
                    img = self.get_input(key)
                    if img is not None:
                        _export_image(img, _make_filename(name, key))
            if {"all", "segmentation"} & targets:
                for key in ordered:
                    img = self.get_segmentation(key)
                    if img is not None:
                        _export_image(img, _make_filename(name, key), suffix=".seg", labels=True)
            if {"all", "projection"} & targets:
                for channel, img in self.get_projection().items():
                    _export_image(img, f"{name}_{channel}")
