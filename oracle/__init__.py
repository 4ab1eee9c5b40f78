"""CPU oracle for the TS2D inference hot path.  TEST INFRASTRUCTURE ONLY.

This package is a CPU restatement (numpy + scipy.ndimage + torch-CPU) of the
algorithm the reference executes on its hot path.  It is *not* part of the
product: only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import it, and only as the checker
or as the timed CPU baseline.  Nothing under ``totalsegmentator2d_b200/`` imports
it, and the product path raises when the CUDA library is missing.

PARITY STATUS: **parity unpinned**.  The arithmetic of the reference lives in
third-party packages that are neither vendored under /root/reference nor
installable offline (SimpleITK ~2.5.2, nnunetv2ml 2.6.2, scikit-image,
dynamic_network_architectures; reference pyproject.toml:17-26), and the
reference's own tests hold no numeric golden vectors (reference
test/test_020_predict_default.py:37-38 asserts types only).  The oracle
therefore restates the *published* algorithms of those packages and is pinned
only by (a) the structural signature of the reference's asset files
(``sample_s0332``/``sample_s0616`` are real outputs of the projection stage:
channel 0 = double-precision mean, channel 1 = integral max), and (b) scipy /
torch-CPU primitives that the upstream code itself calls
(``scipy.ndimage.zoom``, ``gaussian_filter``, ``binary_fill_holes``,
``torch.nn.functional.conv2d`` ...).  See tests/test_oracle_*.py.

Module map (each function cites the reference file:line it follows):
  geometry.py   DICOMOrient('RAI'), max/mean projection, compose/reduce
  nnunet_pre.py crop_to_nonzero, normalisation schemes, spline resampling
  nnunet_net.py PlainConvUNet forward from an nnU-Net state dict
  nnunet_inf.py sliding window, mirror TTA, Gaussian blending, export
  pipeline.py   TS2D.predict for one study / five group models + merge
"""
