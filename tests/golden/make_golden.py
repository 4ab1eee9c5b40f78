"""Generates tests/golden/s0521_projection.json: SHA-256 of the oracle's mean/max projection (float32) of the
reference's only 3-D asset (assets/sample_s0521.nrrd, copied next to this script).  Run from the repo root:
    python tests/golden/make_golden.py
"""
import hashlib, json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import geometry
from totalsegmentator2d_b200 import nrrd
here = os.path.dirname(os.path.abspath(__file__))
img = nrrd.read_nrrd(os.path.join(here, "sample_s0521.nrrd"))
perm, flip = geometry.orientation_remap(img.direction_matrix(), "RAI")
rai = geometry.reorient_array(np.asarray(img.array), perm, flip)
proj = geometry.project_channels(rai, ["mean", "max"])
out = {"shape": list(proj.shape), "mean_sha256": hashlib.sha256(proj[0].tobytes()).hexdigest(),
       "max_sha256": hashlib.sha256(proj[1].tobytes()).hexdigest(),
       "mean_sum": float(proj[0].astype(np.float64).sum()), "max_sum": float(proj[1].astype(np.float64).sum())}
json.dump(out, open(os.path.join(here, "s0521_projection.json"), "w"), indent=1)
print(out)
