"""Oracle (test infrastructure): the raster arithmetic of ``create_visual`` on the CPU.

Follows reference ts2d/core/util/image.py:383-451 (create_visual), :266-290 (image_vector_flatten_max(index=True)),
:319-379 (resample_filter / resample_uniform: output index j samples input index c_in + (j - c_out) * new_sp / old_sp with
c = int(size / 2); default value 0 outside [-0.5, size - 0.5)), :454-481 (get_auto_window 'minmax'), and the ITK filters they call:
NearestNeighborInterpolate (round half up), BSplineInterpolate order 3 (mirror boundaries = scipy ``mode='mirror'``),
LabelToRGB(background 0 -> black, label L -> colormap[L % n]), IntensityWindowing(lower, upper, 0, 255), Cast(uint8) = truncation.
SimpleITK is not installed, so this restatement is unpinned like the rest of the oracle."""
import numpy as np
from scipy import ndimage


def _src_index(n_in, n_out, ratio):
    return (n_in // 2) + (np.arange(n_out, dtype=np.float64) - (n_out // 2)) * float(ratio)


def visual_labels(mask_hwc: np.ndarray, new_hw, ratio_yx, colormap) -> np.ndarray:
    H, W, C = mask_hwc.shape
    nz = mask_hwc != 0
    label = np.where(nz.any(-1), C - np.argmax(nz[..., ::-1], axis=-1), 0)                  # 1 + highest non-zero component
    sy, sx = _src_index(H, new_hw[0], ratio_yx[0]), _src_index(W, new_hw[1], ratio_yx[1])
    iy, ix = np.floor(sy + 0.5).astype(int), np.floor(sx + 0.5).astype(int)
    oky, okx = (sy >= -0.5) & (sy < H - 0.5), (sx >= -0.5) & (sx < W - 0.5)
    res = label[np.clip(iy, 0, H - 1)][:, np.clip(ix, 0, W - 1)]
    res = np.where(oky[:, None] & okx[None, :], res, 0)
    cmap = np.asarray(colormap, dtype=np.uint8).reshape(-1, 3)
    out = np.zeros(res.shape + (3,), np.uint8)
    if len(cmap):
        fg = res != 0
        out[fg] = cmap[res[fg] % len(cmap)]
    return out


def visual_intensity(img_hw: np.ndarray, new_hw, ratio_yx, window=(None, None), resample=True) -> np.ndarray:
    a = np.asarray(img_hw, dtype=np.float32)
    H, W = a.shape
    if resample:
        sy, sx = _src_index(H, new_hw[0], ratio_yx[0]), _src_index(W, new_hw[1], ratio_yx[1])
        yy, xx = np.meshgrid(sy, sx, indexing="ij")
        r = ndimage.map_coordinates(a.astype(np.float64), [yy, xx], order=3, mode="mirror")
        inside = ((sy >= -0.5) & (sy < H - 0.5))[:, None] & ((sx >= -0.5) & (sx < W - 0.5))[None, :]
        a = np.where(inside, r, 0.0).astype(np.float32)
    lo = float(a.min()) if window[0] is None else float(window[0])
    hi = float(a.max()) if window[1] is None else float(window[1])
    factor = 255.0 / (hi - lo)
    offset = -lo * factor
    y = np.where(a < lo, 0.0, np.where(a > hi, 255.0, (a.astype(np.float64) * factor + offset).astype(np.float32)))
    return y.astype(np.uint8)
