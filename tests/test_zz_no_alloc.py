"""`css_peak_normalize_noalloc`: the peak normalisation (css/executor/beamformer.py:122-123) with its scratch in a
file-scope device array of the library instead of a caller buffer -- an entry point that allocates nothing (it has no
counterpart in the reference).  Collected last on purpose (file name).

Checked through the C ABI against the existing `css_peak_normalize` (same kernels -> bit-equal; the DSP tolerance of
tests/test_gpu_kernels.py, 1e-5 relative, is asserted as well against the reference formula), on two calls in a row with
different data so that a stale maximum left in the shared array by the first call would show in the second.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _nv():
    from css_b200 import _native as nv

    return nv


def _both(nv, x):
    """-> (existing entry point's result, no-alloc entry point's result) on copies of x."""
    a, b = x.clone(), x.clone()
    rows, n = x.shape
    scratch = torch.empty(rows, dtype=torch.int32, device=x.device)
    nv.call("css_peak_normalize", nv.ptr(a), nv.c_int(rows), nv.c_long(n), nv.c_float(0.9), nv.ptr(scratch))
    before = torch.cuda.memory_allocated(x.device)
    nv.call("css_peak_normalize_noalloc", nv.ptr(b), nv.c_int(rows), nv.c_long(n), nv.c_float(0.9))
    assert torch.cuda.memory_allocated(x.device) == before
    torch.cuda.synchronize(x.device)
    return a, b


def test_peak_normalize_noalloc_matches_existing_entry_point_twice(cuda):
    nv = _nv()
    g = torch.Generator().manual_seed(11)
    # first call: large peaks; second call: 1000x smaller peaks and another shape -- stale bits from call 1 would win the max
    first = (torch.randn(2, 480_000, generator=g) * 3.0).to(cuda)
    second = (torch.randn(3, 77_777, generator=g) * 3.0e-3).to(cuda)
    for x in (first, second):
        a, b = _both(nv, x)
        assert torch.equal(a, b)
        ref = x * 0.9 / x.abs().amax(dim=1, keepdim=True)
        assert float((b - ref).abs().max()) <= 1e-5 * 0.9
        assert float((b.abs().amax(dim=1) - 0.9).abs().max()) < 1e-6


def test_peak_normalize_noalloc_rejects_more_rows_than_the_builtin_scratch(cuda):
    nv = _nv()
    lib = nv.lib()
    lib.css_peak_normalize_noalloc_max_rows.restype = ctypes.c_int
    limit = lib.css_peak_normalize_noalloc_max_rows()
    assert limit == 64
    x = torch.randn(limit + 1, 128, device=cuda)
    y = x.clone()
    with pytest.raises(nv.NativeError, match="exceed the built-in scratch"):
        nv.call("css_peak_normalize_noalloc", nv.ptr(y), nv.c_int(limit + 1), nv.c_long(128), nv.c_float(0.9))
    torch.cuda.synchronize(cuda)
    assert torch.equal(x, y)                      # rejected before any launch
    z = x[:limit].clone()
    a, b = _both(nv, z)                           # the limit itself is accepted
    assert torch.equal(a, b)
