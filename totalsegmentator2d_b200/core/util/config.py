"""Packaged configuration (reference ts2d/core/util/config.py:13-68, ts2d/data/config.json)."""
import csv
import json
import os

_DATA = os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "data")
_CONFIG = None
_LABEL_COLORS = None


def get_label_colors():
    global _LABEL_COLORS
    if _LABEL_COLORS is None:
        colors = dict()
        with open(os.path.join(_DATA, "label-colors.csv"), newline="") as f:
            for row in csv.DictReader(f):
                if row.get("Label"):
                    colors[row["Label"].lower()] = row["Color"]
        _LABEL_COLORS = colors
    return _LABEL_COLORS


def get_config():
    global _CONFIG
    if _CONFIG is None:
        with open(os.path.join(_DATA, "config.json")) as f:
            _CONFIG = json.load(f)
    return _CONFIG


def get_default_model():
    return get_config()["default-model"]


def get_test_model_single():
    return get_config()["default-test-model-single"]


def get_test_model_single_xr():
    return get_config()["default-test-model-single-xr"]


def get_test_model_multi():
    return get_config()["default-test-model-multi"]


def get_model_resolve_map():
    return get_config()["default-resolve"]


def get_local_models_root():
    """``~/.ts2d/models`` (reference path.py:12-16); ``TS2D_MODELS`` overrides it (offline model zoos, tests)."""
    return os.environ.get("TS2D_MODELS") or os.path.join(os.path.expanduser("~"), ".ts2d", "models")
