"""Label colours (reference ts2d/core/util/color.py:11-103), without the matplotlib/seaborn dependency:
hex strings and the six default colour names are resolved here; other palette names are not available."""
from __future__ import annotations

import random
from typing import Optional, Union

import numpy as np

_DEFAULT_NAMES = ["red", "green", "blue", "yellow", "cyan", "magenta"]
_NAMED = {"red": (1.0, 0.0, 0.0), "green": (0.0, 0.5, 0.0), "blue": (0.0, 0.0, 1.0), "yellow": (1.0, 1.0, 0.0),
          "cyan": (0.0, 1.0, 1.0), "magenta": (1.0, 0.0, 1.0), "white": (1.0, 1.0, 1.0), "black": (0.0, 0.0, 0.0)}


def format_array(a, p=3, sep=", "):
    a = tuple(a) if isinstance(a, (list, tuple, np.ndarray)) else (a,)
    return sep.join(np.format_float_positional(v, precision=p, unique=False) for v in a)


def _name_to_color(name: str):
    name = name.strip()
    if name.startswith("#") and len(name) in (7, 9):
        return tuple(int(name[i:i + 2], 16) / 255.0 for i in (1, 3, 5))
    if name.lower() in _NAMED:
        return _NAMED[name.lower()]
    raise RuntimeError("unknown colour name: {}".format(name))


def tuple_to_color(v):
    v = tuple(v)
    if len(v) != 3:
        raise RuntimeError("Color tuples need to be of length 3 (found: {})".format(len(v)))
    if any(not isinstance(c, (int, np.integer)) for c in v):
        v = np.multiply(np.clip(v, a_min=0, a_max=1.0), 255).astype(int)
    else:
        v = np.clip(v, a_min=0, a_max=255).astype(int)
    return tuple(int(c) for c in v)


def default_color(index: int) -> tuple:
    assert index >= 0
    if index < len(_DEFAULT_NAMES):
        return to_color(_DEFAULT_NAMES[index])
    rnd = random.Random(index)
    return tuple(rnd.randint(32, 200) for _ in range(3))


def to_color(v: Union[int, tuple, list, str]):
    if isinstance(v, str):
        v = _name_to_color(v)
    elif np.isscalar(v):
        if isinstance(v, (int, np.integer)):
            return default_color(int(v))
        v = (v,) * 3
    return tuple_to_color(v)


def to_color_str_rgb_floats(v, sep=", "):
    v = to_color(v)
    v = np.clip(np.asarray(v, dtype=float) / 255, a_min=0, a_max=1).tolist()
    return format_array(v, p=3, sep=sep)


def named_palette(name: Optional[str] = None, size: Optional[int] = None):
    if name is None or name in {"ts2d", "default"}:
        size = len(_DEFAULT_NAMES) if size is None else size
        return [default_color(i) for i in range(size)]
    raise RuntimeError("named palette '{}' needs seaborn, which is not available".format(name))
