"""The B200 kernel module: `KernelCls` + `RNG`, the pair NumS's KernelManager asks a kernel module
for (nums/core/kernel/kernel_manager.py:60-90).  Method names and parameter names are those of the
`Kernel` interface (nums/core/kernel/kernel_interface.py:6-163, minus `syskwargs`), so
`check_implementation` (nums/core/backends/utils.py:74-87) accepts this class unchanged.

Every array argument/result is a `DeviceBlock` living in HBM; every computation is a call into
libnumsb200.so.  Methods outside the hot path exist but raise
NotImplementedError when called.  There is no host fallback for any arithmetic.
"""
import ctypes as C
import random

import numpy as np
from numpy.random import Generator, PCG64

from . import _lib
from . import device as dv
from . import hostlogic as hl
from .device import DeviceBlock
from .grid import ArrayGrid


def block_rng(seed, jump_index):
    """Generator of block `jump_index` (numpy_kernel.py:31-32): PCG64 jumped once per block."""
    return Generator(PCG64(seed).jumped(jump_index))


class RNG:
    """Per-block parallel RNG parameters (numpy_kernel.py:35-83): every new block takes the next jump."""

    def __init__(self, seed=None, jump_index=0):
        self.seed = random.getrandbits(128) if seed is None else seed
        self.rng = PCG64(self.seed)
        self.jump_index = jump_index

    def new_block_rng_params(self):
        params = self.seed, self.jump_index
        self.jump_index += 1
        return params


def _dev(*blocks):
    for b in blocks:
        if isinstance(b, DeviceBlock):
            return b.device_index
    raise TypeError("no device block among the operands")


def _as_block(value, device_index):
    """Operands may arrive as DeviceBlocks (the normal case) or as host scalars/ndarrays that NumS
    passed by value; the latter are uploaded."""
    if isinstance(value, DeviceBlock):
        return dv.to_device(value, device_index)
    return dv.from_numpy(np.asarray(value), device_index)


def _cur(device_index):
    """Creation kernels have no operand to take the device from: the backend selects the CUDA
    device before calling, or passes device_index explicitly."""
    if device_index is not None:
        return int(device_index)
    import torch
    dv.require_cuda()
    return torch.cuda.current_device()


def _unsupported(name):
    raise NotImplementedError(
        "KernelCls.%s is outside the B200 hot path (SURVEY.md section 8) and has no device implementation" % name)

# Helpers live at module level: NumS registers EVERY bound method of KernelCls() as a remote function and looks
# each one up in the Kernel interface (kernel_manager.py:76-84), so the class may hold interface methods only.
def _stream_state(rng_params):
    """NumPy seeds and jumps the bit generator (SeedSequence + PCG64.jumped stay on the host); the device gets
    the block's 128-bit LCG state and increment as {hi, lo} pairs."""
    seed, jump = rng_params
    st = PCG64(seed).jumped(jump).state["state"]
    mask = (1 << 64) - 1
    return ((C.c_uint64 * 2)(st["state"] >> 64, st["state"] & mask),
            (C.c_uint64 * 2)(st["inc"] >> 64, st["inc"] & mask))


def _random_on_device(K, rng_params, size, shape, dev):
    """Uniform doubles generated in HBM, bit-identical to the host stream (nb_pcg64_random)."""
    state, inc = _stream_state(rng_params)
    out = dv.empty((size,), np.float64, dev)
    _lib.check(_lib.load().nb_pcg64_random(dev, size, state, inc, out.ptr, dv.stream_ptr(dev)))
    return DeviceBlock(out.t.view(tuple(shape)))


def _normal_on_device(K, rng_params, size, shape, dev, loc, scale, standard):
    """Generator.normal / standard_normal in HBM, in the host generator's order (nb_pcg64_normal)."""
    lib = _lib.load()
    state, inc = _stream_state(rng_params)
    out = dv.empty((size,), np.float64, dev)
    sc = dv.scratch(lib.nb_pcg64_normal_scratch_bytes(size), dev)
    _lib.check(lib.nb_pcg64_normal(dev, size, state, inc, loc, scale, 1 if standard else 0, out.ptr,
                                   sc.data_ptr(), sc.numel(), dv.stream_ptr(dev)))
    return DeviceBlock(out.t.view(tuple(shape)))


def _integers_on_device(K, rng_params, low, high, size, endpoint, shape, dev):
    """Generator.integers(low, high, size, int64, endpoint) in HBM, bit-identical (nb_pcg64_integers)."""
    if high is None:
        low, high = 0, low
    low, high = int(low), int(high)
    if not endpoint:
        high -= 1
    lo64, hi64 = -(1 << 63), (1 << 63) - 1
    if low < lo64:
        raise ValueError("low is out of bounds for int64")
    if high > hi64:
        raise ValueError("high is out of bounds for int64")
    if low > high:
        raise ValueError("low > high" if endpoint else "low >= high")
    lib = _lib.load()
    state, inc = _stream_state(rng_params)
    out = dv.empty((size,), np.int64, dev)
    rng = high - low
    sc = dv.scratch(lib.nb_pcg64_integers_scratch_bytes(size, rng), dev)
    _lib.check(lib.nb_pcg64_integers(dev, size, state, inc, low, rng, out.ptr, sc.data_ptr(), sc.numel(),
                                     dv.stream_ptr(dev)))
    return DeviceBlock(out.t.view(tuple(shape)))


def _check_int_power(K, exponent):
    # NumPy raises for integer ** negative integer; the device kernel cannot raise, so test first.
    if exponent.size == 0:
        return
    lo = dv.to_numpy(K.reduce_axis("min", exponent, None, False, False))
    if int(lo) < 0:
        raise ValueError("Integers to negative integer powers are not allowed.")


def _matmul(K, a1, a2):
    # np.matmul semantics for <= 2-d operands (numpy_kernel.py:420-424)
    dev = a1.device_index
    lib = _lib.load()
    if a1.ndim == 0 or a2.ndim == 0:
        raise ValueError("matmul: Input operand does not have enough dimensions")
    res_dtype = np.result_type(a1.dtype, a2.dtype)
    if res_dtype == np.dtype(np.bool_):
        raise NotImplementedError("matmul of boolean blocks")
    want = np.dtype(np.float64) if res_dtype.kind == "f" else np.dtype(np.int64)
    if a1.dtype != want:
        a1 = K.astype(a1, want.name)
    if a2.dtype != want:
        a2 = K.astype(a2, want.name)
    # shapes: (m,k)@(k,n), (m,k)@(k,), (k,)@(k,n), (k,)@(k,)
    m = a1.shape[0] if a1.ndim == 2 else 1
    k = a1.shape[-1]
    n = a2.shape[1] if a2.ndim == 2 else 1
    if a2.shape[0] != k:
        raise ValueError("matmul: Input operand 1 has a mismatch in its core dimension 0 (size %d is different "
                         "from %d)" % (a2.shape[0], k))
    if a1.ndim == 2:
        a_rs, a_cs = a1.strides
    else:
        a_rs, a_cs = 0, a1.strides[0]
    if a2.ndim == 2:
        b_rs, b_cs = a2.strides
    else:
        b_rs, b_cs = a2.strides[0], 0
    # the kernels need one unit stride per matrix operand; make odd views contiguous
    if a1.ndim == 2 and m > 1 and k > 1 and 1 not in (a_rs, a_cs):
        a1 = dv.contiguous(a1)
        a_rs, a_cs = a1.strides
    if a2.ndim == 2 and n > 1 and k > 1 and 1 not in (b_rs, b_cs):
        a2 = dv.contiguous(a2)
        b_rs, b_cs = a2.strides
    out_shape = ((m,) if a1.ndim == 2 else ()) + ((n,) if a2.ndim == 2 else ())
    out = dv.empty(out_shape, want, dev)
    if out.size == 0:
        return out
    st = dv.stream_ptr(dev)
    if want == np.dtype(np.float64):
        nbytes = lib.nb_matmul_scratch_bytes(dev, m, n, k)
        sc = dv.scratch(nbytes, dev)
        _lib.check(lib.nb_matmul_f64(dev, m, n, k, a1.ptr, a_rs, a_cs, a2.ptr, b_rs, b_cs, None, out.ptr, n,
                                     sc.data_ptr(), sc.numel(), st))
    else:
        _lib.check(lib.nb_matmul_i64(dev, m, n, k, a1.ptr, a_rs, a_cs, a2.ptr, b_rs, b_cs, out.ptr, n, st))
    return out


def _f64_matrix(K, arr, dev):
    src = dv.contiguous(dv.to_device(arr, dev))
    if src.ndim != 2:
        raise np.linalg.LinAlgError("%d-dimensional array given. Array must be two-dimensional" % src.ndim)
    if src.dtype != np.dtype(np.float64):
        src = K.astype(src, "float64")
    return src


def _concat2d(mats, axis, dev):
    """np.concatenate of float64 matrices along `axis`, by strided device copies (nb_copy_strided)."""
    other = mats[0].shape[1 - axis]
    total = sum(b.shape[axis] for b in mats)
    out = dv.empty((total, other) if axis == 0 else (other, total), np.float64, dev)
    lo = 0
    for b in mats:
        if b.shape[1 - axis] != other:
            raise ValueError("all the input array dimensions except for the concatenation axis must match exactly")
        hi = lo + b.shape[axis]
        dv.copy_into(out.t[lo:hi] if axis == 0 else out.t[:, lo:hi], b.t)
        lo = hi
    return out


def _chol_lower(K, a, dev):
    """Lower Cholesky factor or None when a is not positive definite (one 4-byte read-back)."""
    lib = _lib.load()
    p = a.shape[0]
    out = dv.empty((p, p), np.float64, dev)
    info = dv.empty((1,), np.int32, dev)
    _lib.check(lib.nb_cholesky_f64(dev, p, a.ptr, out.ptr, info.ptr, dv.stream_ptr(dev)))
    return out if int(dv.to_numpy(info)[0]) == 0 else None


def _cholqr_pass(K, Y, dev, shift=0.0):
    """One Cholesky-QR pass: Y = Q R with R = chol(Y^T Y + shift I)^T.  Returns (Q, R) or None."""
    n = Y.shape[1]
    G = K.bop("tensordot", Y, Y, True, False, 1)
    if shift:
        eye = dv.empty((n, n), np.float64, dev)
        _lib.check(_lib.load().nb_eye(dev, n, n, eye.ptr, eye.code, dv.stream_ptr(dev)))
        G = K.bop("add", G, K.bop("mul", eye, dv.from_numpy(np.float64(shift), dev), False, False, {}),
                     False, False, {})
    L = _chol_lower(K, G, dev)
    if L is None:
        return None
    R = dv.contiguous(L.T)
    Rinv = K.inv(R)
    return K.bop("tensordot", Y, Rinv, False, False, 1), R, Rinv


class KernelCls:

    # ---------------------------------------------------------------- I/O, creation, plumbing
    def touch(self, arr):
        # numpy_kernel.py:90 — used to wait for a computation without fetching data.
        if isinstance(arr, DeviceBlock):
            import torch
            torch.cuda.current_stream(arr.device_index).synchronize()
            return True
        return False

    def new_block(self, op_name, grid_entry, grid_meta, device_index=None):
        # numpy_kernel.py:93-101: empty / zeros / ones / eye blocks of an ArrayGrid
        device_index = _cur(device_index)
        grid = ArrayGrid.from_meta(grid_meta)
        shape = grid.get_block_shape(grid_entry)
        dtype = grid.np_dtype()
        out = dv.empty(shape, dtype, device_index)
        lib = _lib.load()
        st = dv.stream_ptr(device_index)
        if op_name == "eye":
            if len(set(grid_entry)) > 1:
                raise AssertionError("eye blocks exist on the diagonal of the grid only")
            _lib.check(lib.nb_eye(device_index, shape[0], shape[1], out.ptr, out.code, st))
        elif op_name in ("zeros", "ones", "empty"):
            v = 1 if op_name == "ones" else 0
            _lib.check(lib.nb_fill(device_index, out.size, out.ptr, out.code, float(v), int(v), st))
        else:
            raise NotImplementedError("new_block op '%s'" % op_name)
        return out

    def random_block(self, rng_params, rfunc_name, rfunc_args, shape, dtype, device_index=None):
        # numpy_kernel.py:103-110.  random / normal / standard_normal (float64) and integers (int64) are generated in
        # HBM in the host generator's order (csrc/rng.cu, csrc/rng_scan.cu); the remaining distributions are drawn
        # by the reference's own host RNG and uploaded; see DESIGN.md "inputs".
        dev = _cur(device_index)
        f64 = np.dtype(dtype) == np.dtype(np.float64)
        if rfunc_name == "random" and f64 and (len(rfunc_args) == 1 or (
                len(rfunc_args) == 2 and np.dtype(rfunc_args[1]) == np.dtype(np.float64))):   # random.py:168
            return _random_on_device(self, rng_params, int(rfunc_args[0]), shape, dev)
        if rfunc_name == "normal" and len(rfunc_args) == 3 and f64 and np.isscalar(rfunc_args[0]) \
                and np.isscalar(rfunc_args[1]):
            loc, scale, size = rfunc_args
            if scale < 0:
                raise ValueError("scale < 0")
            return _normal_on_device(self, rng_params, int(size), shape, dev, float(loc), float(scale), False)
        if rfunc_name == "standard_normal" and len(rfunc_args) == 1 and f64:
            return _normal_on_device(self, rng_params, int(rfunc_args[0]), shape, dev, 0.0, 1.0, True)
        if rfunc_name == "integers" and len(rfunc_args) == 5:
            low, high, size, idt, endpoint = rfunc_args
            if np.dtype(idt) == np.dtype(np.int64) and np.isscalar(low) and (high is None or np.isscalar(high)):
                return _integers_on_device(self, rng_params, low, high, int(size), bool(endpoint), shape, dev)
        rng = block_rng(*rng_params)
        result = getattr(rng, rfunc_name)(*rfunc_args).reshape(shape)
        if rfunc_name not in ("random", "integers"):
            result = result.astype(dtype)
        return dv.from_numpy(result, dev)

    def permutation(self, rng_params, size, device_index=None):
        # numpy_kernel.py:112-114, generated in HBM (csrc/perm.cu), bit-identical to the host generator's shuffle
        dev = _cur(device_index)
        size = int(size)
        if size >= 2 ** 31:
            return dv.from_numpy(block_rng(*rng_params).permutation(size), dev)
        lib = _lib.load()
        state, inc = _stream_state(rng_params)
        out = dv.empty((size,), np.int64, dev)
        if size == 0:
            return out
        sc = dv.scratch(lib.nb_pcg64_permutation_scratch_bytes(size), dev)
        _lib.check(lib.nb_pcg64_permutation(dev, size, state, inc, out.ptr, sc.data_ptr(), sc.numel(),
                                            dv.stream_ptr(dev)))
        return out

    def create_block(self, *src_arrs, src_params, dst_params, dst_shape, dst_shape_bc):
        # numpy_kernel.py:116-129: gather sub-rectangles of several blocks into a new block
        dev = _dev(*src_arrs)
        out = dv.empty(dst_shape, src_arrs[0].dtype, dev)
        assert len(src_params) == len(dst_params)
        for i, src in enumerate(src_arrs):
            src_sel, src_t = src_params[i]
            dst_sel, _ = dst_params[i]
            s = dv.to_device(src, dev).t
            if src_t:
                s = s.permute(*reversed(range(s.dim())))
            d = out.t.view(dst_shape_bc) if dst_shape_bc is not None else out.t
            dv.copy_into(d[dst_sel], s[src_sel])
        return out

    def update_block(self, dst_arr, *src_arrs, src_params, dst_params):
        # numpy_kernel.py:131-151: copy-on-write update of sub-rectangles
        dev = dst_arr.device_index
        base = dv.contiguous(dst_arr)
        out = dv.empty(base.shape, base.dtype, dev)
        dv.copy_into(out.t, base.t)
        assert len(src_params) == len(dst_params)
        _, dst_t = dst_params[0]
        d = out.t.permute(*reversed(range(out.t.dim()))) if dst_t else out.t
        for i, src in enumerate(src_arrs):
            src_sel, src_shape_bc, src_t = src_params[i]
            dst_sel, _ = dst_params[i]
            s = _as_block(src, dev).t
            if src_t:
                s = s.permute(*reversed(range(s.dim())))
            if src_shape_bc is not None:
                s = dv.contiguous(DeviceBlock(s)).t.view(src_shape_bc)
            if s.dtype != d.dtype:
                s = self.astype(DeviceBlock(s), dv._TORCH2NP[d.dtype].name).t
            dv.copy_into(d[dst_sel], s[src_sel])
        return DeviceBlock(d) if dst_t else out

    def update_block_by_index(self, dst_arr, src_arr, index_pairs):
        _unsupported("update_block_by_index")

    def advanced_assign_block_along_axis(self, dst_arr, src_arr, ss, axis, dst_coord, src_coord):
        _unsupported("advanced_assign_block_along_axis")

    def advanced_select_block_along_axis(self, dst_arr, src_arr, ss, dst_axis, src_axis, dst_coord, src_coord):
        _unsupported("advanced_select_block_along_axis")

    def diag(self, arr, offset=0):
        # numpy_kernel.py:285-286 (np.diag, k=offset); offset 0 is what GLM.fit uses (glms.py:163)
        if offset != 0:
            raise NotImplementedError("diag with a non-zero offset")
        dev = arr.device_index
        lib = _lib.load()
        src = dv.contiguous(arr)
        if arr.ndim == 1:
            n = arr.shape[0]
            out = dv.empty((n, n), arr.dtype, dev)
            _lib.check(lib.nb_fill(dev, out.size, out.ptr, out.code, 0.0, 0, dv.stream_ptr(dev)))
            dv.copy_into(out.t.view(-1)[:: n + 1] if n > 0 else out.t.view(-1), src.t)
            return out
        if arr.ndim == 2:
            n = min(arr.shape)
            out = dv.empty((n,), arr.dtype, dev)
            dv.copy_into(out.t, src.t.as_strided((n,), (arr.shape[1] + 1,)))
            return out
        raise ValueError("Input must be 1- or 2-d.")

    def arange(self, start, stop, step, dtype, device_index=None):
        # numpy_kernel.py:288-289
        return dv.from_numpy(np.arange(start, stop, step, dtype), _cur(device_index))

    def transpose(self, arr):
        # numpy_kernel.py:343-344: a view; strides carry the transpose into the consuming kernel
        return arr.T

    def swapaxes(self, arr, axis1, axis2):
        return DeviceBlock(arr.t.transpose(axis1, axis2))

    def reshape(self, arr, shape):
        # numpy_kernel.py:510-511
        return DeviceBlock(dv.contiguous(arr).t.view(tuple(shape)))

    def identity(self, value):
        return value

    def shape_dtype(self, arr):
        return arr.shape, arr.dtype

    def size(self, arr):
        return arr.size

    def split(self, arr, indices_or_sections, axis, transposed):
        _unsupported("split")

    def astype(self, arr, dtype_str):
        # numpy_kernel.py:339-341
        if isinstance(dtype_str, str):
            # the reference does getattr(np, dtype_str) with names such as "int" (glms.py:508, NumPy < 1.24 aliases)
            dtype_str = {"int": "int64", "float": "float64", "bool": "bool_"}.get(dtype_str, dtype_str)
            dtype = np.dtype(getattr(np, dtype_str))
        else:
            dtype = np.dtype(dtype_str)
        dev = arr.device_index
        src = dv.contiguous(arr)
        out = dv.empty(src.shape, dtype, dev)
        _lib.check(_lib.load().nb_astype(dev, src.size, src.ptr, src.code, out.ptr, out.code, dv.stream_ptr(dev)))
        return out

    # ---------------------------------------------------------------- elementwise
    def bop(self, op, a1, a2, a1_T, a2_T, axes):
        # numpy_kernel.py:414-431
        dev = _dev(a1, a2)
        a1 = _as_block(a1, dev)
        a2 = _as_block(a2, dev)
        if a1_T:
            a1 = a1.T
        if a2_T:
            a2 = a2.T
        if op == "tensordot":
            if axes == 1 and max(a1.ndim, a2.ndim) <= 2:
                return _matmul(self, a1, a2)
            raise NotImplementedError("tensordot with axes=%s on %d-d/%d-d blocks" % (axes, a1.ndim, a2.ndim))
        code, cdt, odt = hl.resolve_bop(op, a1.dtype, a2.dtype)
        if hl.canonical_bop(op) == "power" and cdt == hl.NB_I64:
            _check_int_power(self, a2)
        shape, sa, sb = hl.broadcast_geometry(a1.shape, a1.strides, a2.shape, a2.strides)
        cshape, csa, csb = hl.collapse_dims(shape, sa, sb)
        if len(cshape) > 4:
            a1, a2 = dv.contiguous(a1), dv.contiguous(a2)
            shape, sa, sb = hl.broadcast_geometry(a1.shape, a1.strides, a2.shape, a2.strides)
            cshape, csa, csb = hl.collapse_dims(shape, sa, sb)
            if len(cshape) > 4:
                raise NotImplementedError("broadcast of more than 4 non-mergeable axes")
        out = dv.empty(shape, odt, dev)
        if out.size:
            _lib.check(_lib.load().nb_bop(
                dev, code, len(cshape), _lib.i64_array(cshape), a1.ptr, a1.code, _lib.i64_array(csa),
                a2.ptr, a2.code, _lib.i64_array(csb), out.ptr, out.code, cdt, dv.stream_ptr(dev)))
        return out

    def bop_reduce(self, op, a1, a2, a1_T, a2_T):
        # numpy_kernel.py:433-453: pairwise combine of two equal-shape partial results
        dev = _dev(a1, a2)
        a1 = _as_block(a1, dev)
        a2 = _as_block(a2, dev)
        if a1_T:
            a1 = a1.T
        if a2_T:
            a2 = a2.T
        if a1.shape != a2.shape:
            raise AssertionError("bop_reduce operands differ in shape: %s vs %s" % (a1.shape, a2.shape))
        name = hl.canonical_red(op)
        if a1.dtype != a2.dtype:
            common = np.result_type(a1.dtype, a2.dtype)
            a1, a2 = self.astype(a1, common.name), self.astype(a2, common.name)
        a1, a2 = dv.contiguous(a1), dv.contiguous(a2)
        if name in ("all", "any") and a1.dtype != np.dtype(np.bool_):
            a1, a2 = self.astype(a1, "bool_"), self.astype(a2, "bool_")
        out = dv.empty(a1.shape, a1.dtype, dev)
        if out.size:
            _lib.check(_lib.load().nb_bop_reduce(dev, hl.RED_CODE[name], a1.size, a1.ptr, a2.ptr, out.ptr, a1.code,
                                                 dv.stream_ptr(dev)))
        return out

    def sum_reduce(self, *arrs):
        # numpy_kernel.py:291-292
        dev = _dev(*arrs)
        blocks = [dv.contiguous(_as_block(a, dev)) for a in arrs]
        dtype = np.result_type(*[b.dtype for b in blocks])
        if dtype == np.dtype(np.bool_):
            dtype = np.dtype(np.int64)
        blocks = [b if b.dtype == dtype else self.astype(b, dtype.name) for b in blocks]
        shape = blocks[0].shape
        for b in blocks:
            if b.shape != shape:
                raise ValueError("sum_reduce needs equal shapes")
        out = dv.empty(shape, dtype, dev)
        ptrs = (C.c_void_p * len(blocks))(*[b.ptr for b in blocks])
        _lib.check(_lib.load().nb_sum_reduce(dev, len(blocks), out.size, ptrs, out.ptr, out.code, dv.stream_ptr(dev)))
        return out

    def map_uop(self, op_name, arr, args, kwargs):
        # numpy_kernel.py:316-318
        if args or kwargs:
            raise NotImplementedError("map_uop with extra ufunc arguments (%s, %s)" % (args, kwargs))
        dev = arr.device_index
        code, cdt, odt = hl.resolve_uop(op_name, arr.dtype)
        src = dv.contiguous(arr)
        out = dv.empty(src.shape, odt, dev)
        if out.size:
            _lib.check(_lib.load().nb_uop(dev, code, src.size, src.ptr, src.code, out.ptr, out.code, cdt,
                                          dv.stream_ptr(dev)))
        return out

    def xlogy(self, arr_x, arr_y):
        # numpy_kernel.py:336-337
        return self.bop("xlogy", arr_x, arr_y, False, False, {})

    # ---------------------------------------------------------------- reductions
    def reduce_axis(self, op_name, arr, axis, keepdims, transposed):
        # numpy_kernel.py:294-313, including quirk Q1: sum over an int axis of length > 1 without
        # keepdims is computed as matmul(ones_f64, arr) => float64 data even for integer input.
        dev = arr.device_index
        if transposed:
            arr = arr.T
        name = hl.canonical_red(op_name)
        src = dv.contiguous(arr)
        if axis is not None and axis < 0:
            axis += src.ndim
        out_dtype = hl.reduce_out_dtype(name, src.dtype)
        if axis is not None and src.ndim >= 1 and src.shape[axis] > 1 and name == "sum" and not keepdims:
            out_dtype = np.dtype(np.float64)
        if name in ("all", "any"):
            acc = hl.NB_BOOL
        elif name in ("min", "max") and src.dtype == np.dtype(np.bool_):
            acc = hl.NB_BOOL
        else:
            acc = hl.NB_F64 if out_dtype.kind == "f" else hl.NB_I64
        if src.size == 0 or (axis is not None and src.shape[axis] == 0):
            if name in ("min", "max", "nanmin", "nanmax"):
                raise ValueError("zero-size array to reduction operation %s which has no identity" % name)
        outer, length, inner = hl.reduce_geometry(src.shape, axis)
        if axis is None:
            out_shape = (1,) * src.ndim if keepdims else ()
        else:
            out_shape = tuple(1 if i == axis else s for i, s in enumerate(src.shape)) if keepdims else \
                tuple(s for i, s in enumerate(src.shape) if i != axis)
        out = dv.empty(out_shape, out_dtype, dev)
        lib = _lib.load()
        st = dv.stream_ptr(dev)
        if length == 0:
            ident = {"sum": 0, "nansum": 0, "prod": 1, "all": 1, "any": 0}[name]
            _lib.check(lib.nb_fill(dev, out.size, out.ptr, out.code, float(ident), int(ident), st))
            return out
        if out.size == 0:
            return out
        nbytes = lib.nb_reduce_scratch_bytes(outer, length, inner)
        sc = dv.scratch(nbytes, dev)
        _lib.check(lib.nb_reduce(dev, hl.RED_CODE[name], outer, length, inner, src.ptr, src.code, out.ptr, out.code,
                                 acc, sc.data_ptr(), sc.numel(), st))
        return out

    def arg_op(self, op_name, arr, block_slice, other_argoptima=None, other_optima=None):
        # numpy_kernel.py:492-508, including quirk Q2 (a later block wins ties)
        if op_name not in ("argmin", "argmax"):
            raise Exception("Unsupported arg op.")
        dev = arr.device_index
        src = dv.contiguous(arr)
        if src.dtype == np.dtype(np.bool_):
            raise NotImplementedError("arg_op on boolean blocks")
        lib = _lib.load()
        out_idx = dv.empty((), np.int64, dev)
        out_val = dv.empty((), src.dtype, dev)
        oi = ov = None
        if other_optima is not None:
            oi = _as_block(other_argoptima, dev)
            ov = _as_block(other_optima, dev)
            if ov.dtype != src.dtype:
                ov = self.astype(ov, src.dtype.name)
            if oi.dtype != np.dtype(np.int64):
                oi = self.astype(oi, "int64")
        nbytes = lib.nb_argop_scratch_bytes(src.size)
        sc = dv.scratch(nbytes, dev)
        start = int(block_slice.start or 0)
        _lib.check(lib.nb_argop(dev, 1 if op_name == "argmax" else 0, src.size, src.ptr, src.code, start,
                                oi.ptr if oi is not None else None, ov.ptr if ov is not None else None,
                                out_idx.ptr, out_val.ptr, sc.data_ptr(), sc.numel(), dv.stream_ptr(dev)))
        return out_idx, out_val

    def where(self, arr, x, y, block_slice_tuples):
        # numpy_kernel.py:320-334
        dev = arr.device_index
        lib = _lib.load()
        st = dv.stream_ptr(dev)
        cond = dv.contiguous(arr)
        if x is None:
            assert y is None
            if cond.ndim == 0:
                raise ValueError("Calling nonzero on 0d arrays is not allowed.")
            if cond.ndim > 4:
                raise NotImplementedError("where on blocks with more than 4 axes")
            n = cond.size
            nchunks = lib.nb_where_chunks(n)
            counts = dv.empty((nchunks,), np.int64, dev)
            total = dv.empty((1,), np.int64, dev)
            _lib.check(lib.nb_where_count(dev, n, cond.ptr, cond.code, counts.ptr, total.ptr, st))
            nnz = int(dv.to_numpy(total)[0])     # the one host sync (reference: application.py:634)
            outs = [dv.empty((nnz,), np.int64, dev) for _ in range(cond.ndim)]
            if nnz:
                ptrs = (C.c_void_p * cond.ndim)(*[o.ptr for o in outs])
                offs = _lib.i64_array([start for start, _ in block_slice_tuples])
                _lib.check(lib.nb_where_fill(dev, n, cond.ptr, cond.code, cond.ndim, _lib.i64_array(cond.shape), offs,
                                             counts.ptr, ptrs, st))
            return tuple(outs) + ((nnz,),)
        xb, yb = _as_block(x, dev), _as_block(y, dev)
        if not (cond.shape == xb.shape == yb.shape):
            raise AssertionError("where needs equal shapes")
        if xb.dtype != yb.dtype:
            common = np.result_type(xb.dtype, yb.dtype)
            xb, yb = self.astype(xb, common.name), self.astype(yb, common.name)
        xb, yb = dv.contiguous(xb), dv.contiguous(yb)
        out = dv.empty(xb.shape, xb.dtype, dev)
        if out.size:
            _lib.check(lib.nb_where3(dev, out.size, cond.ptr, cond.code, xb.ptr, yb.ptr, out.ptr, xb.code, st))
        return out

    # ---------------------------------------------------------------- linear algebra
    def inv(self, arr):
        # numpy_kernel.py:471-472
        dev = arr.device_index
        if arr.ndim != 2 or arr.shape[0] != arr.shape[1]:
            raise np.linalg.LinAlgError("Last 2 dimensions of the array must be square")
        src = dv.contiguous(arr)
        if src.dtype != np.dtype(np.float64):
            src = self.astype(src, "float64")
        p = src.shape[0]
        lib = _lib.load()
        out = dv.empty((p, p), np.float64, dev)
        info = dv.empty((1,), np.int32, dev)
        sc = dv.scratch(lib.nb_inv_scratch_bytes(p), dev)
        _lib.check(lib.nb_inv_f64(dev, p, src.ptr, out.ptr, info.ptr, sc.data_ptr(), sc.numel(), dv.stream_ptr(dev)))
        if int(dv.to_numpy(info)[0]) != 0:
            raise np.linalg.LinAlgError("Singular matrix")
        return out

    # ---------------------------------------------------------------- factorisations
    def qr(self, *arrays, mode="reduced", axis=None):
        """numpy_kernel.py:455-461: np.linalg.qr of the concatenated operands, modes "r" and "reduced".

        Tall-skinny (m >= n) on the device as CholeskyQR2: R1 = chol(X^T X)^T, Q1 = X R1^-1, R2 = chol(Q1^T Q1)^T,
        R = R2 R1, Q = Q1 R2^-1 -- two passes over X on the FP64 tensor-core Gram/GEMM kernels, orthogonality at
        rounding level for cond(X) < ~1e8; when the first Cholesky breaks down the shifted three-pass variant
        (Fukaya et al. 2020) is used.  LAPACK's sign convention for diag(R) is then restored from the top n x n
        block of Q (nb_qr_signs_f64), so the factors agree with the reference's to rounding, not just up to signs."""
        dev = _dev(*arrays)
        if mode not in ("r", "reduced"):
            raise NotImplementedError("qr mode '%s'" % mode)
        in_dtype = arrays[0].dtype
        if len(arrays) > 1:
            assert axis is not None
            X = _concat2d([_f64_matrix(self, a, dev) for a in arrays], axis, dev)
        else:
            X = _f64_matrix(self, arrays[0], dev)
        m, n = X.shape
        back = (lambda b: b) if in_dtype == np.dtype(np.float64) else (lambda b: self.astype(b, in_dtype.name))
        if m < n:
            # wide block (a column-blocked covariance in linalg.pca): LAPACK reflects the leading m x m part and
            # carries the other columns along, R = [R1, Q^T A2]
            Q, R1 = self.qr(dv.contiguous(DeviceBlock(X.t[:, :m])), mode="reduced")
            R2 = self.bop("tensordot", Q, dv.contiguous(DeviceBlock(X.t[:, m:])), True, False, 1)
            R = _concat2d([R1, R2], 1, dev)
            return back(R) if mode == "r" else (back(Q), back(R))
        lib = _lib.load()
        st = dv.stream_ptr(dev)
        first = _cholqr_pass(self, X, dev)
        passes = []
        if first is None:
            # shifted CholeskyQR3: the shift makes the first Gram matrix safely positive definite
            fro2 = float(dv.to_numpy(self.bop("tensordot", DeviceBlock(X.t.reshape(-1)),
                                             DeviceBlock(X.t.reshape(-1)), False, False, 1)))
            shift = 11.0 * (m * n + n * (n + 1)) * np.finfo(np.float64).eps * fro2
            first = _cholqr_pass(self, X, dev, shift)
            if first is None:
                raise np.linalg.LinAlgError("qr: block is numerically rank deficient beyond the shifted Cholesky QR")
            second = _cholqr_pass(self, first[0], dev)
            if second is None:
                raise np.linalg.LinAlgError("qr: block is numerically rank deficient beyond the shifted Cholesky QR")
            passes = [first, second]
            Y = second[0]
        else:
            passes = [first]
            Y = first[0]
        # last pass: keep R^-1 apart so that mode "r" never forms the final Q
        G = self.bop("tensordot", Y, Y, True, False, 1)
        L = _chol_lower(self, G, dev)
        if L is None:
            raise np.linalg.LinAlgError("qr: Gram matrix of the orthogonalised block is not positive definite")
        R = dv.contiguous(L.T)
        Rinv = self.inv(R)
        for _, Rp, _ in reversed(passes):
            R = self.bop("tensordot", R, Rp, False, False, 1)
        # Householder signs from the top n x n block of Q+
        Ytop = dv.contiguous(DeviceBlock(Y.t[:n]))
        Qtop = self.bop("tensordot", Ytop, Rinv, False, False, 1)
        signs = dv.empty((n,), np.float64, dev)
        sc = dv.scratch(lib.nb_qr_signs_scratch_bytes(n), dev)
        _lib.check(lib.nb_qr_signs_f64(dev, n, Qtop.ptr, n, 1 if m == n else 0, signs.ptr, sc.data_ptr(), sc.numel(), st))
        R = self.bop("mul", DeviceBlock(signs.t.view(n, 1)), R, False, False, {})
        if mode == "r":
            return back(R)
        Rinv_s = self.bop("mul", Rinv, DeviceBlock(signs.t.view(1, n)), False, False, {})
        Q = self.bop("tensordot", Y, Rinv_s, False, False, 1)
        return back(Q), back(R)

    def cholesky(self, arr):
        # numpy_kernel.py:463-464
        dev = arr.device_index
        a = _f64_matrix(self, arr, dev)
        if a.shape[0] != a.shape[1]:
            raise np.linalg.LinAlgError("Last 2 dimensions of the array must be square")
        L = _chol_lower(self, a, dev)
        if L is None:
            raise np.linalg.LinAlgError("Matrix is not positive definite")
        return L if arr.dtype == np.dtype(np.float64) else self.astype(L, arr.dtype.name)

    def svd(self, arr):
        # numpy_kernel.py:466-469 for the square R of a tall-skinny factorisation (linalg.py:221-223)
        dev = arr.device_index
        a = _f64_matrix(self, arr, dev)
        p = a.shape[0]
        if a.shape[1] != p:
            raise NotImplementedError("svd of a non-square block (%d x %d)" % a.shape)
        lib = _lib.load()
        U = dv.empty((p, p), np.float64, dev)
        S = dv.empty((p,), np.float64, dev)
        VT = dv.empty((p, p), np.float64, dev)
        info = dv.empty((1,), np.int32, dev)
        sc = dv.scratch(lib.nb_svd_scratch_bytes(p), dev)
        _lib.check(lib.nb_svd_f64(dev, p, a.ptr, U.ptr, S.ptr, VT.ptr, info.ptr, sc.data_ptr(), sc.numel(),
                                  dv.stream_ptr(dev)))
        if int(dv.to_numpy(info)[0]) < 0:
            raise np.linalg.LinAlgError("SVD did not converge")
        if arr.dtype != np.dtype(np.float64):
            U, S, VT = (self.astype(b, arr.dtype.name) for b in (U, S, VT))
        return U, S, VT

    def einsum(self, subscript, *operands):
        _unsupported("einsum")

    # ---------------------------------------------------------------- boolean helpers
    def array_compare(self, func_name, a, b, args):
        # numpy_kernel.py:476-485 (allclose / array_equal ...), composed from device kernels
        dev = _dev(a, b)
        a, b = _as_block(a, dev), _as_block(b, dev)
        if func_name == "allclose":
            rtol, atol = args
            diff = self.map_uop("abs", self.bop("sub", a, b, False, False, {}), (), {})
            tol = self.bop("add", dv.from_numpy(np.float64(atol), dev),
                           self.bop("mul", dv.from_numpy(np.float64(rtol), dev),
                                    self.map_uop("abs", self.astype(b, "float64"), (), {}), False, False, {}),
                           False, False, {})
            ok = self.bop("le", diff, tol, False, False, {})
            return self.reduce_axis("all", ok, None, False, False)
        if func_name == "array_equal":
            if a.shape != b.shape:
                return dv.from_numpy(np.bool_(False), dev)
            return self.reduce_axis("all", self.bop("eq", a, b, False, False, {}), None, False, False)
        raise NotImplementedError("array_compare '%s'" % func_name)

    def logical_and(self, *bool_list):
        # numpy_kernel.py:489-490: np.all over a list of scalar booleans
        dev = _dev(*bool_list)
        acc = None
        for b in bool_list:
            cur = self.reduce_axis("all", _as_block(b, dev), None, False, False)
            acc = cur if acc is None else self.bop("logical_and", acc, cur, False, False, {})
        return acc

    # ---------------------------------------------------------------- statistics (out of scope)
    def tdigest_chunk(self, arr):
        _unsupported("tdigest_chunk")

    def percentiles_from_tdigest(self, q, *digests):
        _unsupported("percentiles_from_tdigest")

    def select_median(self, arr):
        _unsupported("select_median")

    def weighted_median(self, *arr_and_weights):
        _unsupported("weighted_median")

    def pivot_partition(self, arr, pivot, op):
        _unsupported("pivot_partition")
