"""Key -> model ids -> loaded models (reference ts2d/core/inference/zoo.py:54-166), local database only."""
from __future__ import annotations

import json
import os
from typing import List, Optional

from ..util.config import get_local_models_root, get_model_resolve_map
from .database import FileDataBase, _revision_str


class NNUZoo:
    def __init__(self, remote=None, root: Optional[str] = None):
        self._local = FileDataBase(root or get_local_models_root())
        self._remote = None          # remote download is out of scope (no network); see database.py

    @property
    def local(self):
        return self._local

    def resolve(self, key: str, unique_model: bool = False) -> List[str]:
        mapping = get_model_resolve_map()
        while key in mapping:
            key = mapping[key]
        ids = self._local.ids(key=key)
        if len(ids) > 1 and unique_model:
            models = self._local.models(key=key)
            if not models:
                raise RuntimeError(f"No models resolved for key '{key}'")
            return self._local.ids(model=models[0])
        return ids

    def access(self, id: str, revision=None) -> dict:
        ids = self.resolve(id)
        if len(ids) > 1:
            raise RuntimeError(f"The model id '{id}' is ambiguous (matches {', '.join(ids)})")
        if not ids:
            raise RuntimeError("No pretrained model '{}' in the local database!".format(id))
        if revision is None:
            revision = self._local.latest(key=ids[0])
        return self._local.get(key=ids[0], revision=revision)

    def load(self, id: str, interface: str = "process", param: Optional[dict] = None, **kwargs):
        from .nnu import NNUDeviceModel
        try:
            config = self.access(id=id, **kwargs)
            with open(os.path.join(config["root"], "model.json")) as f:
                config.update(json.load(f))
            merged = dict(config.get("param") or {})
            merged.update(param or {})
            config["param"] = merged
            config["model"] = config.get("model", id)
            return NNUDeviceModel(config)
        except Exception as ex:
            raise RuntimeError("Failed to load a pretrained model: {}".format(ex)) from ex
