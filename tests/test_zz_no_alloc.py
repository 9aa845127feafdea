"""pyn_vec_dot_noalloc (include/pynama_b200.h): the allocation-free variant of pyn_vec_dot.

No counterpart in the reference; added for callers that cannot call cudaMalloc.  Checked through the C-ABI against the
existing entry point (same kernels, same grid, same order of additions: bit-identical) on two calls in a row, on sizes
either side of the grid cap, plus the size limit.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def be():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from pynama_b200 import backend
    return backend


def _dot_existing(be, lib, x, y):
    out = C.c_double(0.0)
    lib.check(lib.load().pyn_vec_dot(x.numel(), be.ptr(x), be.ptr(y), C.byref(out), be.stream()))
    return out.value


def _dot_noalloc(be, lib, x, y, d_out):
    lib.check(lib.load().pyn_vec_dot_noalloc(x.numel(), be.ptr(x), be.ptr(y), be.ptr(d_out), be.stream()))
    return float(d_out.cpu().numpy()[0])


@pytest.mark.parametrize("n", [1, 1000, 100003, 4_000_037])
def test_noalloc_dot_equals_existing_entry_point_twice(be, n):
    import torch
    from pynama_b200 import _lib as lib
    rng = np.random.default_rng(n)
    x, y = rng.standard_normal(n), rng.standard_normal(n)
    xd, yd = be.to_device(x), be.to_device(y)
    x2 = be.to_device(rng.standard_normal(n))
    d_out = be.zeros(1)                      # the caller's output buffer, allocated before the calls under test
    ref1, ref2 = _dot_existing(be, lib, xd, yd), _dot_existing(be, lib, x2, yd)
    assert abs(ref1 - x @ y) <= 1e-10 * max(1.0, np.abs(x * y).sum())
    got1 = _dot_noalloc(be, lib, xd, yd, d_out)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    got2 = _dot_noalloc(be, lib, x2, yd, d_out)          # second call: the shared scratch holds the first call's sums
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before   # nothing was taken from (or returned to) the device
    got3 = _dot_noalloc(be, lib, xd, yd, d_out)
    assert got1 == ref1 and got2 == ref2 and got3 == ref1   # bit-identical: same grid, same order of additions


def test_noalloc_dot_rejects_sizes_beyond_the_maximum(be):
    from pynama_b200 import _lib as lib
    x = be.zeros(8)
    d_out = be.zeros(1)
    rc = lib.load().pyn_vec_dot_noalloc((1 << 28) + 1, be.ptr(x), be.ptr(x), be.ptr(d_out), be.stream())
    assert rc != 0
    assert b"exceeds the maximum" in lib.load().pyn_last_error()
    with pytest.raises(lib.PynamaError):
        lib.check(lib.load().pyn_vec_dot_noalloc(0, be.ptr(x), be.ptr(x), be.ptr(d_out), be.stream()))
