"""Pure-host logic of the kernel module: op-name tables, NumPy dtype resolution, broadcasting and
reduction geometry.  No device access here, so all of it is unit-tested on CPU.

dtype rules follow what the reference does at nums/core/array/utils.py:34-51,100-103: it *executes*
the NumPy ufunc on 0-d samples to learn the result dtype.  We ask the ufunc's type resolver
instead (same answer, no arithmetic) and additionally learn the loop's compute dtype.
"""
import numpy as np

# ---- codes (must match include/nums_b200.h) ------------------------------------------------------
NB_F64, NB_I64, NB_BOOL, NB_F32, NB_I32 = 0, 1, 2, 3, 4

_DT_CODE = {
    np.dtype(np.float64): NB_F64,
    np.dtype(np.int64): NB_I64,
    np.dtype(np.bool_): NB_BOOL,
    np.dtype(np.float32): NB_F32,
    np.dtype(np.int32): NB_I32,
}

_BOP_NAMES = [
    "add", "subtract", "multiply", "true_divide", "floor_divide", "remainder", "fmod", "power", "float_power",
    "maximum", "minimum", "fmax", "fmin",
    "less", "less_equal", "greater", "greater_equal", "equal", "not_equal",
    "logical_and", "logical_or", "logical_xor",
    "bitwise_and", "bitwise_or", "bitwise_xor", "left_shift", "right_shift",
    "arctan2", "hypot", "copysign", "nextafter", "ldexp", "heaviside",
    "logaddexp", "logaddexp2", "gcd", "lcm", "xlogy",
]
BOP_CODE = {name: i for i, name in enumerate(_BOP_NAMES)}

# reference aliases: nums/core/settings.py:50-63 (np_ufunc_map) plus NumPy's own synonyms
BOP_ALIAS = {
    "truediv": "true_divide", "divide": "true_divide", "sub": "subtract", "pow": "power",
    "mult": "multiply", "mul": "multiply", "mod": "remainder",
    "lt": "less", "le": "less_equal", "gt": "greater", "ge": "greater_equal", "eq": "equal", "ne": "not_equal",
}

_UOP_NAMES = [
    "abs", "negative", "positive", "sign", "square", "reciprocal", "sqrt", "cbrt",
    "exp", "exp2", "expm1", "log", "log2", "log10", "log1p",
    "sin", "cos", "tan", "arcsin", "arccos", "arctan",
    "sinh", "cosh", "tanh", "arcsinh", "arccosh", "arctanh",
    "ceil", "floor", "trunc", "rint", "deg2rad", "rad2deg",
    "isnan", "isinf", "isfinite", "signbit", "logical_not", "invert", "spacing",
]
UOP_CODE = {name: i for i, name in enumerate(_UOP_NAMES)}
UOP_ALIAS = {
    "absolute": "abs", "fabs": "abs", "conj": "positive", "conjugate": "positive",
    "bitwise_not": "invert", "degrees": "rad2deg", "radians": "deg2rad",
}

_RED_NAMES = ["sum", "prod", "min", "max", "all", "any", "nansum", "nanmin", "nanmax"]
RED_CODE = {name: i for i, name in enumerate(_RED_NAMES)}
RED_ALIAS = {"amin": "min", "amax": "max", "alltrue": "all", "sometrue": "any", "product": "prod"}

FAMILY_CODE = {"linear": 0, "logistic": 1, "poisson": 2, "logistic_irls": 3}


def dtype_code(dtype):
    dtype = np.dtype(dtype)
    if dtype not in _DT_CODE:
        raise TypeError("dtype %s is not supported by the B200 kernel module" % dtype)
    return _DT_CODE[dtype]


def _compute_code(dt):
    dt = np.dtype(dt)
    if dt.kind == "f":
        return NB_F64
    if dt.kind in "iu":
        return NB_I64
    if dt.kind == "b":
        return NB_BOOL
    raise TypeError("no device loop for dtype %s" % dt)


def canonical_bop(op_name):
    name = BOP_ALIAS.get(op_name, op_name)
    if name not in BOP_CODE:
        raise NotImplementedError("binary op '%s' is not implemented on the device" % op_name)
    return name


def resolve_bop(op_name, dtype_a, dtype_b):
    """-> (op_code, compute_code, out_dtype) for a binary ufunc, following NumPy's type resolution."""
    name = canonical_bop(op_name)
    da, db = np.dtype(dtype_a), np.dtype(dtype_b)
    if name == "xlogy":   # scipy.special ufunc (numpy_kernel.py:429-430): floating loops only
        return BOP_CODE[name], NB_F64, np.dtype(np.float64)
    ufunc = getattr(np, name)
    ra, rb, ro = ufunc.resolve_dtypes((da, db, None))   # raises TypeError exactly when NumPy would
    if name == "ldexp":
        cdt = NB_F64
    else:
        ca, cb = _compute_code(ra), _compute_code(rb)
        if ca != cb:
            raise TypeError("mixed loop dtypes %s,%s for %s" % (ra, rb, name))
        cdt = ca
    out = np.dtype(ro)
    if out not in _DT_CODE or out.itemsize not in (1, 8):
        raise NotImplementedError("%s(%s, %s) -> %s: result dtype not supported on the device" % (name, da, db, out))
    return BOP_CODE[name], cdt, out


def canonical_uop(op_name):
    name = UOP_ALIAS.get(op_name, op_name)
    if name not in UOP_CODE:
        raise NotImplementedError("unary op '%s' is not implemented on the device" % op_name)
    return name


def resolve_uop(op_name, dtype):
    """-> (op_code, compute_code, out_dtype) for a unary ufunc."""
    name = canonical_uop(op_name)
    src = {"abs": "absolute"}.get(name, name)
    # the resolver must see the op the user named (fabs/conj resolve differently from abs/positive)
    ufunc = getattr(np, op_name if hasattr(np, op_name) and isinstance(getattr(np, op_name), np.ufunc) else src)
    ri, ro = ufunc.resolve_dtypes((np.dtype(dtype), None))
    out = np.dtype(ro)
    if out not in _DT_CODE or out.itemsize not in (1, 8):
        raise NotImplementedError("%s(%s) -> %s: result dtype not supported on the device" % (op_name, dtype, out))
    return UOP_CODE[name], _compute_code(ri), out


def canonical_red(op_name):
    name = RED_ALIAS.get(op_name, op_name)
    if name not in RED_CODE:
        raise NotImplementedError("reduction '%s' is not implemented on the device" % op_name)
    return name


def reduce_out_dtype(op_name, dtype):
    """Result dtype of np.<op>(array of dtype) — nums/core/array/utils.py:100-103."""
    name = canonical_red(op_name)
    dtype = np.dtype(dtype)
    if name in ("all", "any"):
        return np.dtype(np.bool_)
    if name in ("sum", "prod", "nansum") and dtype.kind in "bi":
        return np.dtype(np.int64)
    return dtype


def contiguous_strides(shape):
    strides, acc = [], 1
    for dim in reversed(shape):
        strides.append(acc)
        acc *= dim
    return tuple(reversed(strides))


def broadcast_geometry(shape_a, strides_a, shape_b, strides_b):
    """NumPy broadcasting of two strided operands -> (out_shape, sa, sb) with 0 strides on
    broadcast axes.  Raises ValueError like NumPy on mismatch."""
    nd = max(len(shape_a), len(shape_b))
    sha = (1,) * (nd - len(shape_a)) + tuple(shape_a)
    shb = (1,) * (nd - len(shape_b)) + tuple(shape_b)
    sta = (0,) * (nd - len(strides_a)) + tuple(strides_a)
    stb = (0,) * (nd - len(strides_b)) + tuple(strides_b)
    out, sa, sb = [], [], []
    for da, db, xa, xb in zip(sha, shb, sta, stb):
        if da == db:
            dim = da
        elif da == 1:
            dim = db
        elif db == 1:
            dim = da
        else:
            raise ValueError("operands could not be broadcast together with shapes %s %s" % (shape_a, shape_b))
        out.append(dim)
        sa.append(xa if da == dim and dim != 1 else 0)
        sb.append(xb if db == dim and dim != 1 else 0)
    return tuple(out), tuple(sa), tuple(sb)


def collapse_dims(shape, sa, sb):
    """Merges adjacent axes that are jointly contiguous for both operands so that <= 4 remain."""
    dims = [(n, a, b) for n, a, b in zip(shape, sa, sb) if n != 1]
    if not dims:
        return (1,), (0,), (0,)
    merged = [dims[0]]
    for n, a, b in dims[1:]:
        pn, pa, pb = merged[-1]
        if pa == a * n and pb == b * n:
            merged[-1] = (pn * n, a, b)
        else:
            merged.append((n, a, b))
    shape2, sa2, sb2 = zip(*merged)
    return tuple(shape2), tuple(sa2), tuple(sb2)


def reduce_geometry(shape, axis):
    """Views a C-contiguous array as [outer, len, inner] for a reduction over `axis` (None = all)."""
    if axis is None:
        total = int(np.prod(shape, dtype=np.int64)) if len(shape) else 1
        return 1, total, 1
    nd = len(shape)
    if axis < 0:
        axis += nd
    if not 0 <= axis < nd:
        raise np.exceptions.AxisError(axis, nd) if hasattr(np, "exceptions") else ValueError("bad axis")
    outer = int(np.prod(shape[:axis], dtype=np.int64)) if axis > 0 else 1
    inner = int(np.prod(shape[axis + 1:], dtype=np.int64)) if axis + 1 < nd else 1
    return outer, int(shape[axis]), inner


def row_shard(n, world_size, rank):
    """Rows [start, stop) of rank `rank` when n rows are dealt in ceil(n/world) blocks —
    exactly ArrayGrid's slices for grid (world, 1) (nums/core/grid/grid.py:27-106)."""
    per = -(-n // world_size)
    start = min(n, rank * per)
    return start, min(n, start + per)
