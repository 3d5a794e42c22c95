"""Fused GLM operators, registered with the kernel manager by name (`km.register`, the door the
reference's own fusion uses: nums/experimental/optimizer/graph.py:801-810, invoked like :957-959).

Each takes/returns DeviceBlocks and calls one C-ABI entry point of libnumsb200.so.
"""
import math
import os

import numpy as np
import torch

from . import _lib
from . import device as dv
from . import hostlogic as hl
from .device import DeviceBlock


NB_GLM_INT8_HESSIAN = 4
NB_GLM_INT8_NO_LEVEL5 = 32
NB_GLM_INT8_FOUR_PLANES = 64
_HESSIAN_MODES = ("fp64", "int8", "int8!")


def default_hessian():
    """Which kernel forms X^T diag(w) X for 128 < d <= 256: "int8" (tcgen05 integer tensor path, error-free digit
    split, csrc/glm_ozaki.cu) or "fp64" (DMMA, csrc/glm_fused.cu).  NUMS_B200_HESSIAN overrides the default."""
    mode = os.environ.get("NUMS_B200_HESSIAN", "int8")
    if mode not in _HESSIAN_MODES:
        raise ValueError("NUMS_B200_HESSIAN must be one of %s" % (_HESSIAN_MODES,))
    return mode


def int8_hessian_applies(x, family):
    n, d = x.shape
    return (family in ("logistic", "logistic_irls", "linear") and n > 0 and 128 < d <= 256 and d % 4 == 0
            and x.strides[1] == 1 and x.strides[0] % 4 == 0 and x.ptr % 32 == 0)


INT8_MAX_RANGE_LOG2 = 5     # largest entry of a column at most 32x its r.m.s.: fixed-point error of H below 1e-10


def int8_flags(found, rows):
    """Mode bits of the int8 Hessian for a prepared block of `rows` rows: the level-5 digit products are left out when their
    contribution (zero-mean, ~ 2^(2r - 38.6) / sqrt(n) relative to H, with a factor 5 for the largest of 65536 entries) is
    below 1e-11 -- a tenth of the parity bar."""
    flags = NB_GLM_INT8_HESSIAN
    if rows > 0 and 5.0 * 2.0 ** (2 * found.range_log2 - 38.6) / math.sqrt(rows) <= 1e-11:
        flags |= NB_GLM_INT8_NO_LEVEL5
    # four planes instead of five: x sqrt(w) keeps 31 bits, with a dithered rounding whose errors are zero-mean and
    # independent over the rows whatever the data.  Largest error of the 65536 entries of H, relative to sqrt(H_ii H_jj):
    # ~ 6 * 2^(r - 31) / sqrt(n) measured (tests/test_gpu_glm.py::test_hessian_int8_four_planes); taken as 8 * ..., it
    # must stay below 2e-11, a fifth of the parity bar -- blocks of >= 2.2e6 rows for Gaussian columns (r = 3)
    planes = os.environ.get("NUMS_B200_INT8_PLANES", "auto")          # "4" / "5" pin the choice (tests, measurements)
    if planes == "4" or (planes != "5" and rows > 0 and
                         8.0 * 2.0 ** (found.range_log2 - 31) / math.sqrt(rows) <= 2e-11):
        flags |= NB_GLM_INT8_FOUR_PLANES
    return flags


class BlockPrep:
    """What nb_glm_prepare found in an X block: the device buffer the kernels read, and the verdict read back once."""

    __slots__ = ("buf", "bad", "range_log2")

    def __init__(self, buf, bad, range_log2):
        self.buf, self.bad, self.range_log2 = buf, bad, range_log2

    @property
    def usable(self):
        return not self.bad and self.range_log2 <= INT8_MAX_RANGE_LOG2


def glm_prepare(x):
    """Per-column power-of-two scales of an X block for the int8 Hessian (nb_glm_prepare): one pass over the block
    and one 8-byte read-back, kept with the block (blocks are immutable by convention, numpy_kernel.py:131-135)."""
    if x.prep is None:
        lib = _lib.load()
        dev = x.device_index
        n, d = x.shape
        buf = dv.scratch(lib.nb_glm_prep_bytes(d), dev)
        _lib.check(lib.nb_glm_prepare(dev, x.ptr, n, d, x.strides[0], buf.data_ptr(), buf.numel(), dv.stream_ptr(dev)))
        verdict = buf[1024:1032].view(torch.int32).cpu()
        x.prep = BlockPrep(buf, bool(int(verdict[0])), int(verdict[1]))
    return x.prep


def glm_stats(x, y, beta, family, want_hess, need_obj=True, hessian=None):
    """Packed [H (p*p) | g (p) | objective (1)] of one row block (see nb_glm_stats).  need_obj=False (with
    want_hess) leaves the objective out: the Newton loop never reads it.  hessian: "int8" | "fp64" | None (default);
    "int8" falls back to FP64 for a block whose columns have a large dynamic range (BlockPrep.usable)."""
    lib = _lib.load()
    dev = x.device_index
    if x.ndim != 2 or x.dtype != np.dtype(np.float64):
        raise TypeError("glm_stats needs a 2-d float64 block, got %s %s" % (x.shape, x.dtype))
    if x.strides[1] != 1:
        x = dv.contiguous(x)
    y = dv.contiguous(y)
    if y.dtype == np.dtype(np.bool_):
        raise TypeError("glm_stats: y must be float64 or int64")
    beta = dv.contiguous(dv.to_device(beta, dev))
    n, d = x.shape
    if beta.shape != (d + 1,) or y.shape != (n,):
        raise ValueError("glm_stats: shapes X%s y%s beta%s do not agree" % (x.shape, y.shape, beta.shape))
    out = dv.empty((lib.nb_glm_stats_len(d),), np.float64, dev)
    if not want_hess:   # the H part is not written by the kernel: keep the buffer defined for the all-reduce
        _lib.check(lib.nb_fill(dev, out.size, out.ptr, out.code, 0.0, 0, dv.stream_ptr(dev)))
    mode = 3 if (want_hess and not need_obj) else int(bool(want_hess))      # 3 = NB_GLM_HESS_NO_OBJECTIVE
    prep = None
    choice = hessian or default_hessian()
    if want_hess and choice in ("int8", "int8!") and int8_hessian_applies(x, family):
        found = glm_prepare(x)
        if found.usable or choice == "int8!":     # "int8!": whatever the dynamic range of the columns (tests)
            prep = found.buf.data_ptr()
            mode |= int8_flags(found, n)
    sc = dv.scratch(lib.nb_glm_scratch_bytes(dev, n, d, mode), dev)
    _lib.check(lib.nb_glm_stats_prepared(dev, hl.FAMILY_CODE[family], x.ptr, n, d, x.strides[0], y.ptr, y.code,
                                         beta.ptr, mode, prep, out.ptr, sc.data_ptr(), sc.numel(),
                                         dv.stream_ptr(dev)))
    return out


def glm_newton_update(stats, l2vec, beta):
    """(beta - inv(H + diag(l2)) @ (g + l2*beta), max|g + l2*beta|) from packed, already reduced stats.
    Consumes `stats` (H and g are updated in place with the penalty)."""
    lib = _lib.load()
    dev = stats.device_index
    beta = dv.contiguous(dv.to_device(beta, dev))
    p = beta.shape[0]
    if stats.shape != (p * p + p + 1,):
        raise ValueError("stats length %s does not match p=%d" % (stats.shape, p))
    l2 = dv.contiguous(dv.to_device(l2vec, dev)) if l2vec is not None else None
    out = dv.empty((p,), np.float64, dev)
    gmax = dv.empty((), np.float64, dev)
    info = dv.empty((1,), np.int32, dev)
    sc = dv.scratch(lib.nb_glm_newton_scratch_bytes(p), dev)
    _lib.check(lib.nb_glm_newton_update(dev, p, stats.ptr, l2.ptr if l2 is not None else None, beta.ptr, out.ptr,
                                        gmax.ptr, info.ptr, sc.data_ptr(), sc.numel(), dv.stream_ptr(dev)))
    return out, gmax, info


def glm_newton_step(stats, l2vec, beta):
    """glm_newton_update with its two scalars in ONE 16-byte block (max|g| as float64, then info as int32), so that
    the convergence test of an iteration is a single read-back: -> (beta_new, verdict)."""
    lib = _lib.load()
    dev = stats.device_index
    beta = dv.contiguous(dv.to_device(beta, dev))
    p = beta.shape[0]
    if stats.shape != (p * p + p + 1,):
        raise ValueError("stats length %s does not match p=%d" % (stats.shape, p))
    l2 = dv.contiguous(dv.to_device(l2vec, dev)) if l2vec is not None else None
    out = dv.empty((p,), np.float64, dev)
    verdict = dv.empty((2,), np.float64, dev)
    sc = dv.scratch(lib.nb_glm_newton_scratch_bytes(p), dev)
    _lib.check(lib.nb_glm_newton_update(dev, p, stats.ptr, l2.ptr if l2 is not None else None, beta.ptr, out.ptr,
                                        verdict.ptr, verdict.ptr + 8, sc.data_ptr(), sc.numel(), dv.stream_ptr(dev)))
    return out, verdict


def glm_newton_fit(family, beta, l2vec, tol, max_iter, hessian, *blocks):
    """The whole Newton loop in one library call (nb_glm_newton_fit).  blocks = X row blocks followed by the matching y
    blocks, all on one device.  -> (beta_new, [iterations, max|g|, info])."""
    lib = _lib.load()
    nb = len(blocks) // 2
    xs, ys = list(blocks[:nb]), [dv.contiguous(b) for b in blocks[nb:]]
    dev = xs[0].device_index
    d = xs[0].shape[1]
    p = d + 1
    for i, x in enumerate(xs):
        if x.ndim != 2 or x.dtype != np.dtype(np.float64) or x.shape[1] != d or x.device_index != dev:
            raise TypeError("glm_newton_fit: row blocks must be 2-d float64 of one width on one device")
        if x.strides[1] != 1:
            xs[i] = dv.contiguous(x)
        if ys[i].shape != (xs[i].shape[0],) or ys[i].dtype != ys[0].dtype or ys[i].dtype == np.dtype(np.bool_):
            raise TypeError("glm_newton_fit: y blocks must match the row blocks (float64 or int64)")
    out = dv.empty((p,), np.float64, dev)
    dv.copy_into(out.t, dv.contiguous(dv.to_device(beta, dev)).t)      # blocks are immutable: iterate on a copy
    l2 = dv.contiguous(dv.to_device(l2vec, dev)) if l2vec is not None else None
    mode = 3                                                            # NB_GLM_HESS_NO_OBJECTIVE
    preps = [None] * nb
    flags = []
    choice = hessian or default_hessian()
    if choice in ("int8", "int8!"):
        for i, x in enumerate(xs):
            if int8_hessian_applies(x, family):
                found = glm_prepare(x)
                if found.usable or choice == "int8!":
                    preps[i] = found.buf.data_ptr()
                    flags.append(int8_flags(found, x.shape[0]))
        if flags:
            mode |= NB_GLM_INT8_HESSIAN
            for bit in (NB_GLM_INT8_NO_LEVEL5, NB_GLM_INT8_FOUR_PLANES):
                if all(f & bit for f in flags):
                    mode |= bit
    import ctypes as C
    VP, I64 = C.c_void_p * nb, C.c_int64 * nb
    rows = [x.shape[0] for x in xs]
    sc = dv.scratch(lib.nb_glm_newton_fit_scratch_bytes(dev, nb, max(rows), d, mode), dev)
    iters, gmax, info = C.c_int32(0), C.c_double(0.0), C.c_int32(0)
    _lib.check(lib.nb_glm_newton_fit(
        dev, hl.FAMILY_CODE[family], nb, VP(*[x.ptr for x in xs]), I64(*rows), d, I64(*[x.strides[0] for x in xs]),
        VP(*[y.ptr for y in ys]), ys[0].code, VP(*preps), mode, l2.ptr if l2 is not None else None, out.ptr, float(tol),
        int(max_iter), C.byref(iters), C.byref(gmax), C.byref(info), sc.data_ptr(), sc.numel(), dv.stream_ptr(dev)))
    return out, [int(iters.value), float(gmax.value), int(info.value)]


def read_verdict(values):
    """(max|g|, info) from the host copy of a glm_newton_step verdict."""
    values = np.ascontiguousarray(values, dtype=np.float64)
    return float(values[0]), int(values[1:2].view(np.int32)[0])


def glm_predict(x, beta, family):
    lib = _lib.load()
    dev = x.device_index
    if x.strides[1] != 1:
        x = dv.contiguous(x)
    beta = dv.contiguous(dv.to_device(beta, dev))
    n, d = x.shape
    out = dv.empty((n,), np.float64, dev)
    _lib.check(lib.nb_glm_predict(dev, hl.FAMILY_CODE[family], x.ptr, n, d, x.strides[0], beta.ptr, out.ptr,
                                  dv.stream_ptr(dev)))
    return out


def stats_slice(stats, p, what):
    """Views into the packed stats block: 'H' (p,p), 'g' (p,), 'obj' ()."""
    t = stats.t
    if what == "H":
        return DeviceBlock(t[: p * p].view(p, p))
    if what == "g":
        return DeviceBlock(t[p * p: p * p + p])
    return DeviceBlock(t[p * p + p])


def stats_add_g(stats, gvec):
    """A copy of the packed stats whose gradient part has `gvec` added (blocks are immutable by convention)."""
    from .cuda_kernel import KernelCls
    dev = stats.device_index
    gvec = dv.contiguous(dv.to_device(gvec, dev))
    p = gvec.shape[0]
    out = dv.empty(stats.shape, np.float64, dev)
    dv.copy_into(out.t, stats.t)
    summed = KernelCls().bop("add", DeviceBlock(stats.t[p * p: p * p + p]), gvec, False, False, {})
    dv.copy_into(out.t[p * p: p * p + p], summed.t)
    return out


def register(km):
    km.register("stats_add_g", stats_add_g, {})
    km.register("glm_stats", glm_stats, {})
    km.register("glm_newton_update", glm_newton_update, {})
    km.register("glm_newton_step", glm_newton_step, {})
    km.register("glm_newton_fit", glm_newton_fit, {})
    km.register("glm_predict", glm_predict, {})
    km.register("stats_slice", stats_slice, {})
