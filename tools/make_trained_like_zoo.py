"""Test infrastructure: a model zoo whose weights LOOK TRAINED, for asserting the north_star accuracy gates at the benchmarked plan.

Why.  The reference ships trained checkpoints (Zenodo downloads, ts2d/data/shared.json:1-34; Dice 0.72-0.96,
doc/available_models.md:7-11) that cannot be fetched offline.  Seeded random weights put most logits of a 30-layer
InstanceNorm U-Net within a few hundredths of the decision threshold, so ANY 16-bit-operand arithmetic (the reference's own
torch.autocast(float16) included) flips ~0.3 % of the pixels against an fp32 evaluation, and the "99.9 % mask agreement"
gate cannot be asserted.  A trained network is confident almost everywhere - that is what training does - so the gate is
meaningful on it.  This script produces such weights in about ten seconds per model on the GPU box:

  * start from the same seeded state dicts as the random zoo (``synth.init_state_dict``, upstream parameter names),
  * run a few hundred Adam steps of the oracle's own functional network (``oracle.nnunet_net.unet_forward`` under autograd,
    bf16 autocast, cuDNN) on synthetic coronal projections with organ-like multi-label targets,
  * write upstream-named ``checkpoint_final.pth`` files into the usual zoo layout.

Inputs imitate what the pipeline feeds the network for ``bench.synthetic_volume``: channel 0 is the mean projection of an
ellipsoid body in noisy air (smooth chord-length image), channel 1 the max projection (noise-dominated), both z-scored over
the 1000 x 512 projection, then cropped to a 512 x 512 tile at a random row; targets are ellipses at fixed positions in body
coordinates, one per label of a synthetic anatomy of 117 ellipses that all five group models share; each ellipse is denser (or
lighter) than its surroundings along the whole y column, so both projections show it.  Tests paint the same anatomy into their
synthetic studies (``footprint``), so the models find most of their labels there (multi-label, thresholded at logit 0 like ts2d's
region heads).

Nothing here is product code: only ``tests/`` (the ``cs512_trained_zoo`` fixture) and this script's CLI use it.  PyTorch/cuDNN do
the training; the product's kernels are not involved.  Seeds are fixed; cuDNN's backward is not bit-reproducible, which does not
matter because every test compares the engine and the oracle on the SAME written checkpoint.
"""
from __future__ import annotations

import os
import sys
import time
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PROJ_HW = (1000, 512)      # coronal projection of the BASELINE study (z, x)


ANATOMY_SEED = 777
N_LABELS = 117


def anatomy_layout(total: int = N_LABELS, seed: int = ANATOMY_SEED) -> np.ndarray:
    """One synthetic anatomy shared by all group models: per label (uc, vc, a, b, delta) - an ellipse in body coordinates u, v in
    [-1, 1] (u along x, v along z) that is ``delta`` HU denser than its surroundings along the whole y column, so that both projections
    carry it.  Group model g owns labels [off_g, off_g + heads_g) of it; the others are distractors it must leave alone."""
    rng = np.random.default_rng(seed)
    sign = np.where(rng.random(total) < 0.25, -1.0, 1.0)
    return np.stack([rng.uniform(-0.6, 0.6, total), rng.uniform(-0.75, 0.75, total), rng.uniform(0.06, 0.22, total),
                     rng.uniform(0.04, 0.16, total), sign * rng.uniform(80.0, 260.0, total)], 1)


def footprint(hw=PROJ_HW, layout: Optional[np.ndarray] = None, x_radius: float = 0.4) -> np.ndarray:
    """float32 (H, W): sum of the deltas of every label whose ellipse covers the pixel, for the centred body of
    ``bench.synthetic_volume`` (radii 0.4 H, ``x_radius`` W); tests add it to every y column of a synthetic study"""
    layout = anatomy_layout() if layout is None else layout
    H, W = hw
    v = ((np.arange(H, dtype=np.float32) - H / 2) / (0.4 * H))[:, None]
    u = ((np.arange(W, dtype=np.float32) - W / 2) / (x_radius * W))[None, :]
    inside = (u * u + v * v) < 1.0
    fp = np.zeros(hw, np.float32)
    for uc, vc, a, b, delta in layout:
        fp += np.float32(delta) * ((((u - uc) / a) ** 2 + ((v - vc) / b) ** 2 < 1.0) & inside)
    return fp


def synth_projection(n: int, gen: torch.Generator, device, hw=PROJ_HW, layout: Optional[np.ndarray] = None):
    """n z-scored 2-channel projections (n, 2, H, W) and the body coordinates (u, v, inside) they were drawn from"""
    layout = anatomy_layout() if layout is None else layout
    H, W = hw
    z = torch.arange(H, device=device, dtype=torch.float32)[None, :, None]
    x = torch.arange(W, device=device, dtype=torch.float32)[None, None, :]
    r = lambda lo, hi: (lo + (hi - lo) * torch.rand(n, 1, 1, generator=gen, device=device))
    cz, cx = H / 2 + r(-30, 30), W / 2 + r(-15, 15)
    rz, rx = 0.4 * H * r(0.9, 1.1), 0.4 * W * r(0.85, 1.1)
    v, u = ((z - cz) / rz).expand(n, H, W), ((x - cx) / rx).expand(n, H, W)
    rr = u * u + v * v
    inside = rr < 1.0
    chord = 0.7 * torch.sqrt(torch.clamp(1.0 - rr, min=0.0))                     # fraction of the y column inside the body
    bump = torch.zeros(n, H, W, device=device)
    for uc, vc, a, b, delta in layout.tolist():
        bump += delta * ((((u - uc) / a) ** 2 + ((v - vc) / b) ** 2 < 1.0) & inside)
    mean = -300.0 + 600.0 * chord + bump + 20.0 * torch.randn(n, H, W, generator=gen, device=device)
    mx = 1050.0 + 150.0 * torch.randn(n, H, W, generator=gen, device=device) - 150.0 * chord + bump
    img = torch.stack([mean, mx], 1)
    img = (img - img.mean((2, 3), keepdim=True)) / img.std((2, 3), keepdim=True).clamp_min(1e-8)     # ZScoreNormalization
    return img, u, v, inside


def targets_for(u, v, inside, layout: np.ndarray) -> torch.Tensor:
    lay = torch.as_tensor(layout[:, :4], dtype=torch.float32, device=u.device)
    uc, vc, a, b = [lay[:, k][None, :, None, None] for k in range(4)]
    hit = ((u[:, None] - uc) / a) ** 2 + ((v[:, None] - vc) / b) ** 2 < 1.0
    return (hit & inside[:, None]).float()


def train_state_dict(sd: Dict[str, torch.Tensor], spec, heads: int, seed: int, steps: int = 600, batch: int = 2, patch=(512, 512),
                     lr: float = 1e-3, device="cuda", log=None, label_offset: int = 0) -> Dict[str, torch.Tensor]:
    """returns a float32 CPU state dict under the same names; parameters the inference graph never reads (the deep-supervision
    heads of the lower decoder levels) keep their initial values.  The model learns labels [label_offset, label_offset + heads) of
    the shared anatomy with nnU-Net's loss for region heads (soft Dice + BCE)."""
    from oracle import nnunet_net
    gen = torch.Generator(device=device)
    gen.manual_seed(int(seed))
    anatomy = anatomy_layout()
    own = anatomy[label_offset:label_offset + heads]
    assert own.shape[0] == heads, "the shared anatomy holds fewer labels than label_offset + heads"
    params = {k: v.detach().to(device=device, dtype=torch.float32).clone().requires_grad_(True) for k, v in sd.items()}
    # Adam moves a parameter by at most ~lr per step: the 1x1 head gets 10 x the rate so that its bias reaches the strongly negative
    # background level of a converged model within the few hundred steps spent here
    head = [v for k, v in params.items() if ".seg_layers." in k]
    body = [v for k, v in params.items() if ".seg_layers." not in k]
    opt = torch.optim.Adam([{"params": body, "lr": lr, "scale": 1.0}, {"params": head, "lr": 10 * lr, "scale": 10.0}])
    Ph, Pw = patch
    t0 = time.perf_counter()
    last = (None, None)
    for it in range(steps):
        img, u, v, inside = synth_projection(batch, gen, device, layout=anatomy)
        rows = torch.randint(0, img.shape[2] - Ph + 1, (batch,), generator=gen, device=device).tolist()
        cols = torch.randint(0, img.shape[3] - Pw + 1, (batch,), generator=gen, device=device).tolist()
        crop = lambda t, i, r0, c0: t[i][..., r0:r0 + Ph, c0:c0 + Pw]
        xb = torch.stack([crop(img, i, r0, c0) for i, (r0, c0) in enumerate(zip(rows, cols))])
        ub, vb, ib = [torch.stack([crop(t, i, r0, c0) for i, (r0, c0) in enumerate(zip(rows, cols))]) for t in (u, v, inside)]
        yb = targets_for(ub, vb, ib, own)
        flips = torch.randint(0, 4, (1,), generator=gen, device=device).item()        # the mirroring augmentation the TTA relies on
        dims = [d for d, bit in ((2, 1), (3, 2)) if flips & bit]
        if dims:
            xb, yb = torch.flip(xb, dims), torch.flip(yb, dims)
        for g in opt.param_groups:
            g["lr"] = g["scale"] * lr * (1.0 - it / steps) ** 0.9                      # nnU-Net's poly schedule
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=str(device).startswith("cuda")):
            logits = nnunet_net.unet_forward(xb, params, spec)
        logits = logits.float()
        bce = F.binary_cross_entropy_with_logits(logits, yb)
        pr = torch.sigmoid(logits)
        inter, denom = (pr * yb).sum((0, 2, 3)), pr.sum((0, 2, 3)) + yb.sum((0, 2, 3))
        dice = ((2 * inter + 1.0) / (denom + 1.0)).mean()
        loss = bce + (1.0 - dice)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(list(params.values()), 12.0)
        opt.step()
        last = (float(bce.detach()), float(dice.detach()))
    if log is not None:
        log(f"trained-like: {steps} steps, batch {batch}, final BCE {last[0]:.4f} soft Dice {last[1]:.3f}, {time.perf_counter() - t0:.1f} s")
    return {k: v.detach().float().cpu() for k, v in params.items()}


def write_trained_like_zoo(root: str, model_key: str, groups: Dict[str, List[str]], arch: dict, patch=(512, 512), spacing=(1.5, 1.5),
                           base_seed: int = 1000, steps: int = 600, mirror_axes=(0, 1), only: Optional[Sequence[str]] = None, log=print):
    """the zoo of ``bench.write_zoo`` / ``synth.write_model_zoo`` with every checkpoint replaced by its trained-like version"""
    from oracle import nnunet_net
    from totalsegmentator2d_b200 import synth
    ids = []
    offsets, off = {}, 0
    for group, labels in groups.items():
        offsets[group], off = off, off + len(labels)
    for gi, (group, labels) in enumerate(groups.items()):
        if only is not None and group not in only:
            continue
        mid = synth.write_model_dir(root, model_key, group, labels, arch, patch, spacing, seed=base_seed + gi, mirror_axes=mirror_axes)
        ckpt = os.path.join(root, mid, "r001", "Dataset001_TS2D", "nnUNetTrainer__nnUNetPlans__2d", "fold_0", "checkpoint_final.pth")
        blob = torch.load(ckpt, map_location="cpu", weights_only=False)
        spec = nnunet_net.ArchSpec(arch, 2, len(labels))
        blob["network_weights"] = train_state_dict(blob["network_weights"], spec, len(labels), base_seed + gi, steps=steps, patch=patch, label_offset=offsets[group],
                                                   log=(lambda m, g=group: log(f"[{g}] {m}")) if log else None)
        torch.save(blob, ckpt)
        ids.append(mid)
    return ids


if __name__ == "__main__":
    import argparse
    import bench
    from totalsegmentator2d_b200 import synth
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("out")
    ap.add_argument("--steps", type=int, default=600)
    a = ap.parse_args()
    write_trained_like_zoo(a.out, bench.MODEL_KEY, synth.TS2D_V2_GROUPS, synth.cs512_arch(), steps=a.steps)
