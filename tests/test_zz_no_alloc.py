"""``ts2d_resample2d_static``: the cubic resampler for callers that cannot allocate device memory.

``ts2d_resample2d`` wants a double-precision work area from its caller.  Not every caller can provide one: a plugin invoked
from inside a host application's callback must not call ``cudaMalloc`` there (it locks the whole context and can stall the
application's streams), and a binding that is given only an input and an output pointer owns no pool to carve one from.
``ts2d_resample2d_static`` keeps the work area in an array compiled into the library instead (16 MiB + 512 B per GPU) and
allocates nothing.  This file checks, through the C ABI only, that

* it returns exactly what ``ts2d_resample2d`` returns (same kernels, same order of operations, so the comparison is
  bit for bit, tighter than the 1e-4 the oracle comparison of ``ts2d_resample2d`` itself allows),
* a second call right after the first is still right although the built-in area is left dirty: the second image is smaller
  than the first, so stale padded pixels and a stale clip range of the first call lie exactly where a bug would read them,
* sizes beyond the built-in area are refused with a message that names the limit, and the refusal does not disturb later calls.

The file name sorts after every other test file on purpose: ``pytest -x`` stops at the first failure, and a failure here
should not hide the parity tests.  It has to pass alone in a fresh process (first call does the per-GPU address look-up) and
after the rest of the suite (library already warm).
"""
import ctypes

import numpy as np
import pytest

torch = pytest.importorskip("torch")

pytestmark = pytest.mark.gpu


def _image(shape, seed):
    rng = np.random.default_rng(seed)
    c, h, w = shape
    yy, xx = np.mgrid[:h, :w]
    base = np.stack([np.sin(yy / (3.0 + k)) * np.cos(xx / (5.0 + k)) * (100.0 + 50 * k) for k in range(c)])
    return (base + rng.normal(scale=10.0, size=shape)).astype(np.float32)


def _caller_scratch(lib, L, x, new_hw, order):
    """the existing entry point; the work area comes from torch here, i.e. from a caller that is able to allocate"""
    c, h, w = x.shape
    out = torch.full((c, *new_hw), float("nan"), dtype=torch.float32, device=x.device)
    scratch = torch.empty(c * (h + 24) * (w + 24) + 2 * c, dtype=torch.float64, device=x.device) if order == 3 else None
    L.check(lib.ts2d_resample2d(x.data_ptr(), c, h, w, out.data_ptr(), new_hw[0], new_hw[1], order,
                                scratch.data_ptr() if scratch is not None else None, L.stream_ptr()), "ts2d_resample2d")
    torch.cuda.synchronize()
    return out.cpu().numpy()


def _built_in_scratch(lib, L, x, out, new_hw, order):
    """the entry point under test: inputs and outputs exist before the call, nothing is allocated around or inside it"""
    c, h, w = x.shape
    rc = lib.ts2d_resample2d_static(x.data_ptr(), c, h, w, out.data_ptr(), new_hw[0], new_hw[1], order, L.stream_ptr())
    return rc


def test_static_resampler_equals_caller_scratch_resampler_twice_in_a_row():
    from totalsegmentator2d_b200 import _lib as L
    lib = L.load(check_device=True)
    assert lib.ts2d_resample2d_static_capacity() == (1 << 21) + 64

    # call 1 is large (fills most of the built-in area), call 2 small with a different value range and channel count
    cases = [((2, 500, 256), (1000, 512), 11), ((3, 41, 67), (90, 33), 12)]
    xs = [torch.from_numpy(_image(shape, seed) * (1.0 if i == 0 else 1e-3)).cuda() for i, (shape, _, seed) in enumerate(cases)]
    outs = [torch.full((shape[0], *new), float("nan"), dtype=torch.float32, device="cuda") for shape, new, _ in cases]
    want = [_caller_scratch(lib, L, x, new, 3) for x, (_, new, _) in zip(xs, cases)]

    # everything the two calls touch exists now; from here to the synchronize the caching allocator must stay untouched
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    for x, out, (_, new, _) in zip(xs, outs, cases):
        L.check(_built_in_scratch(lib, L, x, out, new, 3), "ts2d_resample2d_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before

    for out, ref in zip(outs, want):
        got = out.cpu().numpy()
        assert np.isfinite(got).all()
        assert np.array_equal(got, ref), float(np.abs(got - ref).max())

    # the same pair again, in the opposite order, on the now dirty area
    for i in (1, 0):
        outs[i].fill_(float("nan"))
    for i in (1, 0):
        L.check(_built_in_scratch(lib, L, xs[i], outs[i], cases[i][1], 3), "ts2d_resample2d_static")
    torch.cuda.synchronize()
    for out, ref in zip(outs, want):
        assert np.array_equal(out.cpu().numpy(), ref)

    # order 1 takes no work area; it must agree as well
    x, new = xs[1], cases[1][1]
    lin = torch.empty((x.shape[0], *new), dtype=torch.float32, device="cuda")
    L.check(_built_in_scratch(lib, L, x, lin, new, 1), "ts2d_resample2d_static")
    torch.cuda.synchronize()
    assert np.array_equal(lin.cpu().numpy(), _caller_scratch(lib, L, x, new, 1))


def test_static_resampler_refuses_what_does_not_fit():
    from totalsegmentator2d_b200 import _lib as L
    lib = L.load(check_device=True)
    cap = lib.ts2d_resample2d_static_capacity()
    x = torch.zeros(16, dtype=torch.float32, device="cuda")          # never read: the size check comes first
    out = torch.zeros(16, dtype=torch.float32, device="cuda")
    # 2 x 1000 x 1000 is the documented largest case; one more row does not fit
    assert 2 * 1024 * 1024 + 4 <= cap < 2 * 1025 * 1024 + 4
    rc = lib.ts2d_resample2d_static(x.data_ptr(), 2, 1001, 1000, out.data_ptr(), 4, 4, 3, L.stream_ptr())
    assert rc == -1, rc                                                # TS2D_ERR_ARG
    msg = L.last_error()
    assert str(cap) in msg and "ts2d_resample2d" in msg, msg
    rc = lib.ts2d_resample2d_static(x.data_ptr(), 1, 100000, 100000, out.data_ptr(), 4, 4, 3, L.stream_ptr())
    assert rc == -1, rc                                                # no 32-bit wrap-around in the size computation
    with pytest.raises(L.TS2DError, match="order must be 1 or 3"):
        L.check(lib.ts2d_resample2d_static(x.data_ptr(), 1, 4, 4, out.data_ptr(), 4, 4, 2, L.stream_ptr()))
    torch.cuda.synchronize()

    # a refused call leaves the library usable: the exact-fit case right at the limit still matches
    big = torch.from_numpy(_image((2, 1000, 1000), 13)).cuda()
    res = torch.empty((2, 300, 700), dtype=torch.float32, device="cuda")
    L.check(_built_in_scratch(lib, L, big, res, (300, 700), 3), "ts2d_resample2d_static")
    torch.cuda.synchronize()
    assert np.array_equal(res.cpu().numpy(), _caller_scratch(lib, L, big, (300, 700), 3))
