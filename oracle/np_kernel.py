"""ORACLE — test infrastructure only.  Never imported by the nums_b200 package.

A NumPy restatement of what the reference's block kernels compute
(nums/core/kernel/numpy_kernel.py, v0.2.8).  The reference delegates all arithmetic to NumPy /
SciPy (un-vendored: setup.cfg pins numpy<=1.20; this image has NumPy 2.3), so the restatement is
"which NumPy call, with which arguments, in which order" — including the reference's quirks:

  Q1  reduce_axis: `sum` over an int axis of length > 1 without keepdims is matmul(ones, arr), i.e.
      float64 data even for integer input                                   (numpy_kernel.py:294-313)
  Q2  arg_op chain: the earlier optimum survives only if STRICTLY better    (numpy_kernel.py:492-508)

Pinned by tests/test_oracle_golden.py against tests/golden/*.npz, which were produced by running the
reference itself (tests/golden/make_golden.py).
"""
import numpy as np
import scipy.special

# nums/core/settings.py:50-63
UFUNC_ALIAS = {
    "truediv": "true_divide", "sub": "subtract", "pow": "power", "mult": "multiply", "mul": "multiply",
    "tensordot": "multiply", "lt": "less", "le": "less_equal", "gt": "greater", "ge": "greater_equal",
    "eq": "equal", "ne": "not_equal",
}


def bop(op, a1, a2, a1_T=False, a2_T=False, axes=None):
    """numpy_kernel.py:414-431"""
    lhs = a1.T if a1_T else a1
    rhs = a2.T if a2_T else a2
    if op == "tensordot":
        if axes == 1 and max(np.ndim(lhs), np.ndim(rhs)) <= 2:
            return np.matmul(lhs, rhs)
        return np.tensordot(lhs, rhs, axes=axes)
    name = UFUNC_ALIAS.get(op, op)
    fn = getattr(np, name, None)
    if fn is None:
        fn = getattr(scipy.special, name)
    return fn(lhs, rhs)


def bop_reduce(op, a1, a2, a1_T=False, a2_T=False):
    """numpy_kernel.py:433-453"""
    lhs = a1.T if a1_T else a1
    rhs = a2.T if a2_T else a2
    if op == "sum":
        return lhs + rhs
    if op == "prod":
        return lhs * rhs
    return getattr(np, op)(np.stack([lhs, rhs], axis=0), axis=0, keepdims=False)


def map_uop(op_name, arr, args=(), kwargs=None):
    """numpy_kernel.py:316-318"""
    return getattr(np, op_name)(arr, *args, **(kwargs or {}))


def sum_reduce(*arrs):
    """numpy_kernel.py:291-292"""
    return np.add.reduce(arrs)


def reduce_axis(op_name, arr, axis, keepdims, transposed):
    """numpy_kernel.py:294-313 (quirk Q1)"""
    data = arr.T if transposed else arr
    if axis is not None and data.shape[axis] > 1 and op_name == "sum" and not keepdims:
        if data.ndim == 2:
            ones = np.ones(data.shape[axis])
            return np.matmul(ones, data) if axis == 0 else np.matmul(data, ones)
        return np.tensordot(np.ones(data.shape[axis]), data, axes=(0, axis))
    return getattr(np, op_name)(data, axis=axis, keepdims=keepdims)


def where(arr, x, y, block_slice_tuples):
    """numpy_kernel.py:320-334"""
    if x is None:
        idx = [ix + start for ix, (start, _stop) in zip(np.where(arr), block_slice_tuples)]
        return tuple(idx) + (idx[0].shape,)
    assert arr.shape == x.shape == y.shape
    return np.where(arr, x, y)


def arg_op(op_name, arr, block_slice, other_argoptima=None, other_optima=None):
    """numpy_kernel.py:492-508 (quirk Q2)"""
    if op_name == "argmax":
        pos = np.argmax(arr)
        keep_other = other_optima is not None and other_optima > arr[pos]
    elif op_name == "argmin":
        pos = np.argmin(arr)
        keep_other = other_optima is not None and other_optima < arr[pos]
    else:
        raise Exception("Unsupported arg op.")
    if keep_other:
        return other_argoptima, other_optima
    return block_slice.start + pos, arr[pos]


def inv(arr):
    """numpy_kernel.py:471-472"""
    return np.linalg.inv(arr)


def astype(arr, dtype_str):
    """numpy_kernel.py:339-341"""
    return arr.astype(getattr(np, dtype_str))


def block_rng(seed, jump_index):
    """numpy_kernel.py:31-32"""
    return np.random.Generator(np.random.PCG64(seed).jumped(jump_index))


def random_block(rng_params, rfunc_name, rfunc_args, shape, dtype):
    """numpy_kernel.py:103-110"""
    out = getattr(block_rng(*rng_params), rfunc_name)(*rfunc_args).reshape(shape)
    return out if rfunc_name in ("random", "integers") else out.astype(dtype)
