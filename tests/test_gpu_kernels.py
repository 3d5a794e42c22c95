"""GPU parity tests of the block kernels (KernelCls -> ctypes -> libnumsb200.so) against the golden
vectors produced by the reference and against the oracle on seeded inputs.
Bar: bit-exact for integer / bool / index results; rel-err <= 1e-10 for float64 (tolerance from
BASELINE.json's north_star)."""
import numpy as np
import pytest

from conftest import load_golden, rel_err
from oracle import np_kernel as ok

pytestmark = pytest.mark.gpu
FTOL = 1e-10


@pytest.fixture(scope="module")
def K(app):
    from nums_b200 import cuda_kernel
    return cuda_kernel.KernelCls()


def up(a, dev=0):
    from nums_b200 import device as dv
    return dv.from_numpy(np.asarray(a), dev)


def down(b):
    from nums_b200 import device as dv
    if isinstance(b, tuple):
        return tuple(down(x) for x in b)
    return dv.to_numpy(b)


def same_float(a, b, tol=FTOL):
    """equal NaN/inf pattern, finite values within tol (relative to magnitude, elementwise)."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    assert np.array_equal(np.isnan(a), np.isnan(b)), (a, b)
    fin = np.isfinite(b)
    assert np.array_equal(a[~fin & ~np.isnan(b)], b[~fin & ~np.isnan(b)]), (a, b)
    if fin.any():
        err = np.abs(a[fin] - b[fin]) / np.maximum(1.0, np.abs(b[fin]))
        assert err.max() <= tol, (err.max(), a, b)


INT_OPS = ["add", "sub", "mul", "truediv", "floor_divide", "mod", "fmod", "maximum", "minimum", "lt", "le", "gt",
           "ge", "eq", "ne", "bitwise_and", "bitwise_or", "bitwise_xor", "left_shift", "right_shift", "gcd", "lcm",
           "logical_and", "logical_or", "logical_xor"]


def test_int64_binary_ops_bit_exact(K):
    g = load_golden("kernels")
    a, b = up(g["ia"]), up(g["ib"])
    for op in INT_OPS:
        r = down(K.bop(op, a, b, False, False, {}))
        want = g["i_" + op]
        assert r.dtype == want.dtype, op
        if want.dtype.kind == "f":
            same_float(r, want)
        else:
            np.testing.assert_array_equal(r, want, err_msg=op)
    r = down(K.bop("pow", up(np.array([3, -2, 2, 7, 0], dtype=np.int64)),
                   up(np.array([41, 3, 62, 0, 0], dtype=np.int64)), False, False, {}))
    np.testing.assert_array_equal(r, g["i_pow"])
    with pytest.raises(ValueError):
        K.bop("pow", a, up(np.full(len(g["ia"]), -1, dtype=np.int64)), False, False, {})


def test_float64_binary_ops(K):
    g = load_golden("kernels")
    a, b = up(g["fa"]), up(g["fb"])
    for key in [k for k in g if k.startswith("f_")]:
        r = down(K.bop(key[2:], a, b, False, False, {}))
        want = g[key]
        assert r.dtype == want.dtype, key
        if want.dtype == np.bool_:
            np.testing.assert_array_equal(r, want, err_msg=key)
        else:
            same_float(r, want)
    same_float(down(K.bop("add", up(g["fa"][:5]), up(g["ia"][:5]), False, False, {})), g["mixed_add"])
    np.testing.assert_array_equal(down(K.bop("lt", up(g["ia"][:5]), up(g["fa"][:5]), False, False, {})), g["mixed_lt"])
    same_float(down(K.bop("ldexp", a, up(np.arange(-6, 6, dtype=np.int64)), False, False, {})), g["ldexp"])


def test_unary_ops(K):
    g = load_golden("kernels")
    a = up(g["ua"])
    for key in [k for k in g if k.startswith("u_")]:
        r = down(K.map_uop(key[2:], a, (), {}))
        want = g[key]
        assert r.dtype == want.dtype, key
        if want.dtype == np.bool_:
            np.testing.assert_array_equal(r, want, err_msg=key)
        else:
            same_float(r, want, 1e-13)
    ai = up(g["ui"])
    for key in [k for k in g if k.startswith("ui_")]:
        r = down(K.map_uop(key[3:], ai, (), {}))
        want = g[key]
        assert r.dtype == want.dtype, key
        if want.dtype.kind == "f":
            same_float(r, want, 1e-13)
        else:
            np.testing.assert_array_equal(r, want, err_msg=key)


def test_broadcasting_shapes_of_the_newton_loop(K):
    g = load_golden("kernels")
    col, mat = up(g["col"]), up(g["mat"])
    np.testing.assert_array_equal(down(K.bop("mul", col, mat, False, False, {})), g["bc_col_mat"])
    r = K.bop("truediv", up(np.float64(1.0)), up(g["mat"][:, 0] + 8), False, False, {})
    same_float(down(r), g["bc_scalar"])
    np.testing.assert_array_equal(down(K.bop("add", mat, up(g["mat"].T.copy()), False, True, {})), g["bc_T"])


def test_reductions_and_quirk_q1(K):
    g = load_golden("kernels")
    imat, mat, nanv, t3 = up(g["imat"]), up(g["mat"]), up(g["nanv"]), up(g["t3"])
    cases = [("red_sum_ax0_int", "sum", imat, 0, False, False), ("red_sum_ax1_int", "sum", imat, 1, False, False),
             ("red_sum_all_int", "sum", imat, None, False, False), ("red_sum_ax0_keep", "sum", imat, 0, True, False),
             ("red_sum_ax0_T", "sum", imat, 0, False, True), ("red_max_ax1", "max", mat, 1, False, False),
             ("red_min_all", "min", mat, None, False, False), ("red_max_nan", "max", nanv, None, False, False),
             ("red_nanmax", "nanmax", nanv, None, False, False), ("red_nansum", "nansum", nanv, None, False, False),
             ("red_t3_ax1", "sum", t3, 1, False, False)]
    for key, op, arr, axis, keep, tr in cases:
        r = down(K.reduce_axis(op, arr, axis, keep, tr))
        want = g[key]
        assert r.dtype == want.dtype, key        # Q1: float64 data for the int axis-sum
        assert r.shape == want.shape, key
        if want.dtype.kind == "f":
            same_float(r, want)
        else:
            np.testing.assert_array_equal(r, want, err_msg=key)
    np.testing.assert_array_equal(down(K.reduce_axis("any", up(g["mat"] > 10), 0, False, False)), g["red_any"])
    np.testing.assert_array_equal(down(K.reduce_axis("all", up(g["mat"] > -8), None, False, False)), g["red_all"])
    np.testing.assert_array_equal(down(K.bop_reduce("sum", mat, up(g["mat"] * 2), False, False)), g["bopred_sum"])
    same_float(down(K.bop_reduce("max", nanv, up(g["nanv"][::-1].copy()), False, False)), g["bopred_max"])
    np.testing.assert_array_equal(down(K.sum_reduce(mat, up(g["mat"] * 2), up(g["mat"] * 3))), g["sumred"])


def test_arg_op_chain_quirk_q2_and_nan(K):
    g = load_golden("kernels")
    r0 = K.arg_op("argmax", up(np.array([5.0, 1.0])), slice(0, 2))
    r1 = K.arg_op("argmax", up(np.array([5.0, 0.0])), slice(2, 4), *r0)
    assert [int(down(r0[0])), int(down(r1[0]))] == list(g["arg_chain"])
    np.testing.assert_array_equal([float(down(r0[1])), float(down(r1[1]))], g["arg_chain_val"])
    nanb = up(np.array([1.0, np.nan, 7.0, np.nan]))
    got = [int(down(K.arg_op("argmax", nanb, slice(0, 4))[0])), int(down(K.arg_op("argmin", nanb, slice(0, 4))[0]))]
    assert got == list(g["arg_nan"])
    r0 = K.arg_op("argmin", up(np.array([4, 2, 2], dtype=np.int64)), slice(10, 13))
    r1 = K.arg_op("argmin", up(np.array([9, 2, 3], dtype=np.int64)), slice(13, 16), *r0)
    assert [int(down(r0[0])), int(down(r1[0]))] == list(g["arg_chain_min_int"])


def test_where(K):
    g = load_golden("kernels")
    w = K.where(up(g["where_cond"]), None, None, [(10, 15), (100, 106)])
    np.testing.assert_array_equal(down(w[0]), g["where_rows"])
    np.testing.assert_array_equal(down(w[1]), g["where_cols"])
    assert tuple(w[2]) == tuple(g["where_count"])
    xs, ys = np.arange(30, dtype=float).reshape(5, 6), -np.arange(30, dtype=float).reshape(5, 6)
    np.testing.assert_array_equal(down(K.where(up(g["where_cond"]), up(xs), up(ys), None)), g["where3"])
    # empty and all-true
    w = K.where(up(np.zeros(5000, dtype=bool)), None, None, [(0, 5000)])
    assert down(w[0]).shape == (0,) and w[1] == (0,)
    w = K.where(up(np.ones(5000, dtype=bool)), None, None, [(7, 5007)])
    np.testing.assert_array_equal(down(w[0]), np.arange(7, 5007))


def test_tensordot_and_inv_golden(K):
    g = load_golden("kernels")
    A, Bm = up(g["A"]), up(g["Bm"])
    same_float(down(K.bop("tensordot", A, Bm, False, False, 1)), g["mm"])
    same_float(down(K.bop("tensordot", A, A, True, False, 1)), g["mm_T"])
    same_float(down(K.bop("tensordot", A, up(g["Bm"][:, 0].copy()), False, False, 1)), g["mv"])
    same_float(down(K.bop("tensordot", A, up(np.arange(7, dtype=float)), True, False, 1)), g["vm_T"])
    same_float(down(K.bop("tensordot", up(g["A"][:, 0].copy()), up(g["A"][:, 1].copy()), False, False, 1)), g["dot"])
    np.testing.assert_array_equal(down(K.bop("tensordot", up(g["imat"].T.copy()), up(g["imat"]), False, False, 1)),
                                  g["imm"])
    assert rel_err(down(K.inv(up(g["S"]))), g["S_inv"]) < FTOL


# ------------------------------------------------------------------------------------------------
# seeded comparisons with the oracle at sizes that cross the kernels' internal thresholds
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n", [1, 31, 1000, 4097, 300001, 2 ** 22 + 3])
def test_elementwise_sizes(K, n):
    rng = np.random.default_rng(n)
    a, b = rng.standard_normal(n), rng.standard_normal(n)
    same_float(down(K.bop("mul", up(a), up(b), False, False, {})), a * b, 0)
    same_float(down(K.map_uop("exp", up(a), (), {})), np.exp(a), 1e-14)
    ia = rng.integers(-10 ** 12, 10 ** 12, n)
    ib = rng.integers(-50, 50, n)
    np.testing.assert_array_equal(down(K.bop("floor_divide", up(ia), up(ib), False, False, {})),
                                  ok.bop("floor_divide", ia, ib))
    np.testing.assert_array_equal(down(K.bop("eq", up(ib), up(np.int64(7)), False, False, {})), ib == 7)
    # comparisons of 8-byte operands pack eight results per store above 4096 elements; NaN compares false
    a[::17] = np.nan
    for op, fn in (("lt", np.less), ("ge", np.greater_equal), ("ne", np.not_equal)):
        np.testing.assert_array_equal(down(K.bop(op, up(a), up(b), False, False, {})), fn(a, b))
        np.testing.assert_array_equal(down(K.bop(op, up(np.float64(0.25)), up(b), False, False, {})), fn(0.25, b))
    np.testing.assert_array_equal(down(K.bop("le", up(ia), up(ib), False, False, {})), ia <= ib)


@pytest.mark.parametrize("shape,axis", [((1,), None), ((5000,), None), ((3_000_017,), None), ((2000, 3), 0),
                                        ((2000, 3), 1), ((200_003, 16), 0), ((7, 100_000), 1), ((300, 400), 0),
                                        ((300, 400), 1), ((4, 5, 6), 1), ((50_000, 300), 0), ((70, 3000), 1)])
def test_reduce_shapes(K, shape, axis):
    rng = np.random.default_rng(sum(shape))
    a = rng.standard_normal(shape)
    for op in ("sum", "max", "min"):
        got = down(K.reduce_axis(op, up(a), axis, False, False))
        want = ok.reduce_axis(op, a, axis, False, False)
        assert got.shape == np.shape(want)
        assert rel_err(got, want) < FTOL, (op, shape, axis)
    ia = rng.integers(-1000, 1000, shape)
    got = down(K.reduce_axis("sum", up(ia), axis, True, False))
    np.testing.assert_array_equal(got, np.sum(ia, axis=axis, keepdims=True))
    assert got.dtype == np.int64


def test_argop_large_with_ties(K):
    rng = np.random.default_rng(5)
    a = rng.integers(0, 1000, 3_000_000).astype(np.float64)
    for op, ref in (("argmax", np.argmax), ("argmin", np.argmin)):
        idx, val = K.arg_op(op, up(a), slice(100, 100 + a.size))
        assert int(down(idx)) == 100 + int(ref(a))          # first occurrence within a block, like NumPy
        assert float(down(val)) == a[ref(a)]
    a[1_234_567] = np.nan
    a[2_000_000] = np.nan
    assert int(down(K.arg_op("argmax", up(a), slice(0, a.size))[0])) == 1_234_567
    p = rng.permutation(5_000_000)
    assert int(down(K.arg_op("argmax", up(p), slice(0, p.size))[0])) == int(np.argmax(p))


def test_where_large(K):
    rng = np.random.default_rng(11)
    a = rng.integers(0, 10, (1500, 700))
    w = K.where(up(a == 3), None, None, [(5, 1505), (0, 700)])
    r, c = np.where(a == 3)
    np.testing.assert_array_equal(down(w[0]), r + 5)
    np.testing.assert_array_equal(down(w[1]), c)
    f = rng.standard_normal(2_000_003)
    w = K.where(up(f), None, None, [(0, f.size)])     # float condition: non-zero test
    assert w[1] == (f.size,)


MM_SHAPES = [(5, 3, 7), (32, 32, 64), (33, 17, 1000), (64, 64, 4096), (100, 90, 513), (128, 128, 256),
             (257, 129, 300), (300, 300, 77)]


@pytest.mark.parametrize("m,n,k", MM_SHAPES)
@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_matmul_layouts(K, m, n, k, ta, tb):
    rng = np.random.default_rng(m * 1000 + n + k)
    A = rng.standard_normal((k, m) if ta else (m, k))
    B = rng.standard_normal((n, k) if tb else (k, n))
    got = down(K.bop("tensordot", up(A), up(B), ta, tb, 1))
    want = ok.bop("tensordot", A, B, ta, tb, 1)
    assert rel_err(got, want) < FTOL


@pytest.mark.parametrize("n,d", [(200_000, 2), (100_003, 10), (50_000, 64), (20_000, 100), (40_000, 256),
                                 (10_000, 300), (3_000, 1025)])
def test_tall_skinny_gram_and_vectors(K, n, d):
    """x^T x, X^T X (symmetric split-k path), X @ b, X^T r at tall-skinny shapes."""
    rng = np.random.default_rng(d)
    X = rng.standard_normal((n, d))
    b, r = rng.standard_normal(d), rng.standard_normal(n)
    Xd = up(X)
    G = down(K.bop("tensordot", Xd, Xd, True, False, 1))
    assert rel_err(G, X.T @ X) < FTOL
    assert np.array_equal(G, G.T)
    assert rel_err(down(K.bop("tensordot", Xd, up(b), False, False, 1)), X @ b) < FTOL
    assert rel_err(down(K.bop("tensordot", Xd, up(r), True, False, 1)), X.T @ r) < FTOL
    assert rel_err(down(K.bop("tensordot", up(r), up(r), False, False, 1)), r @ r) < FTOL


@pytest.mark.parametrize("n,d1,d2", [(60_000, 384, 384), (40_002, 512, 300), (30_000, 258, 1024), (9_000, 640, 640),
                                     (2_000, 1024, 1024)])
def test_tall_skinny_products_split_k_tma_path(K, n, d1, d2):
    """A^T B of two tall blocks (few 128 x 128 output tiles, long contraction): the warp-specialised TMA kernel cut
    along k, partial tiles summed in chunk order; the symmetric case computes the upper-triangular tiles only; the
    weighted case X^T diag(w) X is what the Newton statistics use for d > 256."""
    import ctypes as C
    from nums_b200 import _lib, device as dv
    rng = np.random.default_rng(n + d1)
    A, B = rng.standard_normal((n, d1)), rng.standard_normal((n, d2))
    Ad, Bd = up(A), up(B)
    assert rel_err(down(K.bop("tensordot", Ad, Bd, True, False, 1)), A.T @ B) < FTOL
    if d1 == d2:
        G = down(K.bop("tensordot", Ad, Ad, True, False, 1))
        assert rel_err(G, A.T @ A) < FTOL and np.array_equal(G, G.T)
        w = rng.random(n) + 0.1
        wd, out = up(w), dv.empty((d1, d1), np.float64, 0)
        lib = _lib.load()
        sc = dv.scratch(lib.nb_matmul_scratch_bytes(0, d1, d1, n), 0)
        _lib.check(lib.nb_matmul_f64(0, d1, d1, n, Ad.ptr, 1, d1, Ad.ptr, d1, 1, wd.ptr, out.ptr, d1,
                                     sc.data_ptr(), sc.numel(), dv.stream_ptr(0)))
        H = down(out)
        assert rel_err(H, (A * w[:, None]).T @ A) < FTOL and np.array_equal(H, H.T)
    # B^T-layout operand (k contiguous on both sides): (d1 x n) @ (n x d2) with A^T materialised
    At = up(np.ascontiguousarray(A.T))
    assert rel_err(down(K.bop("tensordot", At, Bd, False, False, 1)), A.T @ B) < FTOL


@pytest.mark.parametrize("p", [1, 2, 3, 17, 32, 33, 100, 257, 600])
def test_inv_sizes(K, p):
    rng = np.random.default_rng(p)
    A = rng.standard_normal((p, p)) + np.eye(p) * 0.5
    got = down(K.inv(up(A)))
    want = np.linalg.inv(A)
    assert rel_err(got, want) < 1e-9 * max(1.0, np.linalg.cond(A) / 1e3)
    assert rel_err(got @ A, np.eye(p)) < 1e-9 * max(1.0, np.linalg.cond(A) / 1e3)
    S = A.T @ A + np.eye(p)        # the SPD Hessian case
    assert rel_err(down(K.inv(up(S))), np.linalg.inv(S)) < FTOL * max(1.0, np.linalg.cond(S))


def test_inv_singular_raises(K):
    with pytest.raises(np.linalg.LinAlgError):
        K.inv(up(np.zeros((4, 4))))


def test_plumbing_blocks(K):
    rng = np.random.default_rng(0)
    A = rng.standard_normal((37, 53))
    Ad = up(A)
    np.testing.assert_array_equal(down(K.transpose(Ad)), A.T)
    np.testing.assert_array_equal(down(K.astype(Ad, "int64")), A.astype(np.int64))
    np.testing.assert_array_equal(down(K.reshape(Ad, (53, 37))), A.reshape(53, 37))
    np.testing.assert_array_equal(down(K.diag(up(A[0]), 0)), np.diag(A[0]))
    out = K.create_block(Ad, Ad, src_params=[((slice(0, 10), slice(3, 9)), False), ((slice(0, 10), slice(5, 11)), True)],
                         dst_params=[((slice(0, 10), slice(0, 6)), False), ((slice(0, 10), slice(6, 12)), False)],
                         dst_shape=(10, 12), dst_shape_bc=None)
    np.testing.assert_array_equal(down(out), np.hstack([A[0:10, 3:9], A.T[0:10, 5:11]]))
    upd = K.update_block(Ad, up(np.ones((4, 1))), src_params=[((slice(0, 4), slice(0, 1)), None, False)],
                         dst_params=[((slice(2, 6), slice(7, 8)), False)])
    want = A.copy()
    want[2:6, 7:8] = 1.0
    np.testing.assert_array_equal(down(upd), want)
    np.testing.assert_array_equal(down(Ad), A)      # inputs are never mutated (numpy_kernel.py:131-135)
    ident = K.new_block("eye", (0, 0), {"shape": (6, 6), "block_shape": (6, 6), "dtype": "float64"}, device_index=0)
    np.testing.assert_array_equal(down(ident), np.eye(6))


@pytest.mark.parametrize("n,jump", [(1, 0), (1000, 3), (1_000_003, 5), (10_000_000, 17)])
def test_device_block_rng_uniform_is_bit_identical(K, n, jump):
    """random_block("random") is generated in HBM and must equal the reference's host stream
    Generator(PCG64(seed).jumped(jump)).random(n) bit for bit (numpy_kernel.py:31-32,103-110)."""
    got = down(K.random_block((1337, jump), "random", (n,), (n,), np.float64, device_index=0))
    want = ok.random_block((1337, jump), "random", (n,), (n,), np.float64)
    np.testing.assert_array_equal(got, want)
    mat = down(K.random_block((42, jump), "random", (6 * 7,), (6, 7), np.float64, device_index=0))
    np.testing.assert_array_equal(mat, ok.random_block((42, jump), "random", (6 * 7,), (6, 7), np.float64))
    # other distributions still come from the host stream, unchanged
    # a distribution that stays on the host generator (uploaded): same stream by construction
    z = down(K.random_block((1337, jump), "exponential", (2.0, 257), (257,), np.float64, device_index=0))
    np.testing.assert_array_equal(z, ok.random_block((1337, jump), "exponential", (2.0, 257), (257,), np.float64))


@pytest.mark.parametrize("n,jump", [(1, 0), (63, 1), (64, 2), (65, 3), (1000, 4), (100_003, 5), (5_000_000, 17)])
def test_device_block_rng_normal_follows_the_host_ziggurat(K, n, jump):
    """random_block("normal") is generated in HBM (csrc/rng_scan.cu) in the order of the reference's
    Generator(PCG64(seed).jumped(jump)).normal(loc, scale, n) (numpy_kernel.py:31-32,103-110; random.py:158-195).
    Bit-identical, tail draws included: the 0.03 % drawn from the tail go through the C library's log1p on the host,
    which the device reproduces rounding by rounding (csrc/libm_exact.cuh).  One misplaced rejection would shift every
    later element, so element-wise equality over millions of draws also pins the word-consumption logic."""
    for args, name in (((0.0, 1.0, n), "normal"), ((-3.5, 2.25, n), "normal"), ((n,), "standard_normal")):
        got = down(K.random_block((1337, jump), name, args, (n,), np.float64, device_index=0))
        want = ok.random_block((1337, jump), name, args, (n,), np.float64)
        assert got.shape == want.shape and got.dtype == want.dtype
        np.testing.assert_array_equal(got, want)
    mat = down(K.random_block((7, jump), "normal", (1.0, 0.5, 12 * 5), (12, 5), np.float64, device_index=0))
    np.testing.assert_array_equal(mat, ok.random_block((7, jump), "normal", (1.0, 0.5, 12 * 5), (12, 5), np.float64))


INT_RANGES = [
    (0, 10, False), (-5, 5, True), (3, 3, True), (0, 2 ** 31 + 12345, False), (0, 2 ** 32, False),
    (-7, 2 ** 32 + 5, False), (-2 ** 62, 2 ** 62, False), (0, 3 * 2 ** 61, False), (-2 ** 63, 2 ** 63 - 1, True),
    (10 ** 8, None, False), (0, 2 ** 20, False),
]


@pytest.mark.parametrize("low,high,endpoint", INT_RANGES)
def test_device_block_rng_integers_is_bit_identical(K, low, high, endpoint):
    """random_block("integers", dtype=int64): Lemire rejection on 32-bit half-words / 64-bit words, constant, raw 32
    and raw 64-bit ranges, all in the host generator's order (random.py:173-178 argument order)."""
    for n, jump in ((1, 0), (129, 1), (1_000_003, 2)):
        args = (low, high, n, np.int64, endpoint)
        got = down(K.random_block((99, jump), "integers", args, (n,), np.int64, device_index=0))
        want = ok.random_block((99, jump), "integers", args, (n,), np.int64)
        assert got.dtype == want.dtype == np.int64
        np.testing.assert_array_equal(got, want)


@pytest.mark.parametrize("n", [1, 2, 3, 5, 17, 1000, 4097, 100_003, 2 ** 20 - 1, 2 ** 20, 2 ** 20 + 1, 3_000_000])
def test_device_block_rng_permutation_is_bit_identical(K, n):
    """KernelCls.permutation: NumPy's Fisher-Yates (masked-rejection draws on 32-bit half-words, swaps from the top)
    restated as prefix sums and pointer jumping (csrc/perm.cu); one wrong acceptance would change every later draw."""
    for jump in (0, 7):
        got = down(K.permutation((1337, jump), n, device_index=0))
        want = ok.block_rng(1337, jump).permutation(n)
        assert got.dtype == np.int64
        np.testing.assert_array_equal(got, want)


def test_device_block_rng_integers_rejects_bad_bounds(K):
    with pytest.raises(ValueError):
        K.random_block((1, 0), "integers", (5, 5, 10, np.int64, False), (10,), np.int64, device_index=0)
    with pytest.raises(ValueError):
        K.random_block((1, 0), "integers", (0, 2 ** 63 + 1, 10, np.int64, False), (10,), np.int64, device_index=0)


@pytest.mark.parametrize("m,n,k", [(1664, 1664, 512), (1666, 1730, 1002), (2048, 1540, 70)])
@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_matmul_large_tma_path(K, m, n, k, ta, tb):
    """Sizes that fill the machine with 128x128 tiles take the warp-specialised TMA-fed kernel (matmul_ws.cu),
    including ragged m / n / k tails."""
    rng = np.random.default_rng(m + n + k)
    A = rng.standard_normal((k, m) if ta else (m, k))
    B = rng.standard_normal((n, k) if tb else (k, n))
    got = down(K.bop("tensordot", up(A), up(B), ta, tb, 1))
    want = ok.bop("tensordot", A, B, ta, tb, 1)
    assert rel_err(got, want) < FTOL
