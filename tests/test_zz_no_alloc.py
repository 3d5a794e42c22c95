"""nb_reduce_static — the reduction entry point for callers that cannot allocate device memory.

Who needs it: hosts that drive libnumsb200.so with buffers they merely borrow (a cgo / JNI / ctypes
wrapper around another runtime's memory pool) and plug-ins inside services whose HBM budget is fixed at
start-up.  They cannot supply the split-reduction scratch nb_reduce asks for, so nb_reduce_static keeps a
1 MiB scratch array inside the library.

The test calls it through the C ABI and compares with nb_reduce (caller-provided scratch) on the same
input, twice in a row: the second call runs on an array that still holds the first call's partials,
which must not matter.  Same kernels in the same order, so the results are expected to be identical;
the asserted tolerance is the suite's float64 bar (rel-err 1e-10).  The file sorts last on purpose, so a
run that stops at the first failure has already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
FTOL = 1e-10


def _stream(dev):
    import torch
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def _run_pair(a, outer, length, inner, red, acc, out_dtype):
    """-> (nb_reduce result, nb_reduce_static result) as host arrays."""
    from nums_b200 import _lib, device as dv, hostlogic as hl
    lib = _lib.load()
    dev = 0
    src = dv.from_numpy(a, dev)
    want = dv.empty((outer, inner), out_dtype, dev)
    got = dv.empty((outer, inner), out_dtype, dev)
    need = lib.nb_reduce_scratch_bytes(outer, length, inner)
    assert 0 < need <= (1 << 20)
    sc = dv.scratch(need, dev)
    _lib.check(lib.nb_reduce(dev, hl.RED_CODE[red], outer, length, inner, src.ptr, src.code, want.ptr, want.code, acc,
                             sc.data_ptr(), sc.numel(), _stream(dev)))
    _lib.check(lib.nb_reduce_static(dev, hl.RED_CODE[red], outer, length, inner, src.ptr, src.code, got.ptr, got.code,
                                    acc, _stream(dev)))
    return dv.to_numpy(want), dv.to_numpy(got)


def test_reduce_static_matches_reduce_twice_in_a_row():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from nums_b200 import hostlogic as hl
    rng = np.random.default_rng(2024)
    # call 1: full float64 sum of 3e6 elements (split over 592 CTAs)
    a = rng.standard_normal(3_000_000)
    want, got = _run_pair(a, 1, a.size, 1, "sum", hl.NB_F64, np.float64)
    assert abs(got[0, 0] - want[0, 0]) <= FTOL * abs(want[0, 0])
    assert abs(got[0, 0] - a.sum()) <= FTOL * np.abs(a).sum()
    # call 2, immediately after, different shape and op: column max of a tall-skinny matrix; the built-in
    # scratch still contains call 1's partial sums
    b = rng.standard_normal((200_000, 16))
    want, got = _run_pair(b, 1, b.shape[0], b.shape[1], "max", hl.NB_F64, np.float64)
    np.testing.assert_allclose(got, want, rtol=FTOL, atol=0)
    np.testing.assert_array_equal(got[0], b.max(axis=0))
    # and once more with integers (bit-exact)
    c = rng.integers(-1000, 1000, 1_000_000)
    want, got = _run_pair(c, 1, c.size, 1, "sum", hl.NB_I64, np.int64)
    assert int(got[0, 0]) == int(want[0, 0]) == int(c.sum())


def test_reduce_static_rejects_oversize():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from nums_b200 import _lib, device as dv, hostlogic as hl
    lib = _lib.load()
    outer, length, inner = 1, 60_000, 4096          # needs about 1.2 MB of partials > 1 MiB
    assert lib.nb_reduce_scratch_bytes(outer, length, inner) > (1 << 20)
    src = dv.empty((length, inner), np.float64, 0)
    out = dv.empty((inner,), np.float64, 0)
    rc = lib.nb_reduce_static(0, hl.RED_CODE["sum"], outer, length, inner, src.ptr, src.code, out.ptr, out.code,
                              hl.NB_F64, _stream(0))
    assert rc != 0
    assert b"built-in buffer" in lib.nb_last_error()
