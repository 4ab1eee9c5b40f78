"""Local model database: ``<root>/<model>_<group>/rNNN`` (reference ts2d/core/inference/database.py:17-22,
87-142).  The remote (Zenodo) database of the reference is out of scope - there is no network - so
``use_remote`` is accepted and ignored; the on-disk contract is kept so real checkpoints drop in."""
from __future__ import annotations

import os
from glob import glob
from typing import Optional


def decompose_model_key(key: str):
    model, group = key.rsplit("_", maxsplit=1) if "_" in key else (key, None)
    return model, group


def _revision_str(revision):
    return "r{:03d}".format(revision) if isinstance(revision, int) else revision


def _parse_revision(rn):
    try:
        return int(rn) if isinstance(rn, int) else int(str(rn)[1:] if str(rn).startswith("r") else str(rn))
    except ValueError:
        return None


class FileDataBase:
    def __init__(self, root: str):
        self._root = root

    @property
    def root(self):
        return self._root

    def _enumerate(self):
        for dn in glob(os.path.join(self._root, "*", "r*")):
            rdn = os.path.relpath(dn, self._root)
            model, rn = os.path.split(rdn)
            rn = _parse_revision(rn)
            if rn is None or "_" not in model:
                continue
            model, group = model.rsplit("_", maxsplit=1)
            yield model, group, rn, dn

    @staticmethod
    def _match_model_str(match: Optional[str], model: str) -> bool:
        if match is None:
            return True
        if "-" in model:
            mparts, parts = match.split("-"), model.split("-")
            for i in range(len(parts)):
                if i < len(mparts) and mparts[i] and mparts[i] != parts[i]:
                    return False
            return True
        return model == match

    def list(self, model=None, group=None, key=None, revision=None):
        if key is not None:
            model, group = decompose_model_key(key)
        revision = _parse_revision(revision) if isinstance(revision, str) else revision
        res = dict()
        for m, g, r, p in self._enumerate():
            if self._match_model_str(model, m) and (revision is None or revision == r) and (group is None or group == g):
                res[(m, g, r)] = p
        return res

    def ids(self, **kw):
        return sorted(set(f"{m}_{g}" for (m, g, r) in self.list(**kw)))

    def models(self, **kw):
        return sorted(set(m for (m, g, r) in self.list(**kw)))

    def latest(self, **kw) -> Optional[int]:
        revs = sorted(set(r for (m, g, r) in self.list(**kw)))
        return revs[-1] if revs else None

    def get(self, key: str, revision=None) -> dict:
        items = sorted(self.list(key=key, revision=revision).items(), key=lambda t: f"{t[0][0]}_{t[0][1]}")
        (m, g, r), p = items[0]
        return {"id": f"{m}_{g}", "model": m, "group": g, "revision": r, "path": p, "root": p}
