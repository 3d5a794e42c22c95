"""GPU parity tests of the factorisation kernels (KernelCls.qr / cholesky / svd -> libnumsb200.so) against NumPy's
LAPACK, which is what the reference's kernels call (nums/core/kernel/numpy_kernel.py:455-469), and of the
tall-skinny routines of nums_b200/linalg.py against the oracle restatement of nums/core/linalg.py and the golden
vectors produced by the reference.  Tolerance: rel-err <= 1e-10 (BASELINE.json north_star) on well-conditioned
inputs; factorisation invariants (Q R = X, Q^T Q = I, U S V^T = A) everywhere."""
import numpy as np
import pytest

from conftest import rel_err

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


@pytest.mark.parametrize("m,n", [(1, 1), (9, 9), (50, 4), (2345, 9), (300, 300), (100_000, 64), (40_000, 256),
                                 (30_001, 33)])
def test_qr_matches_lapack_including_signs(K, m, n):
    rng = np.random.default_rng(m * 7 + n)
    X = rng.standard_normal((m, n)) + (0.5 if m > 1 else 0.0)
    q0, r0 = np.linalg.qr(X)
    Q, R = down(K.qr(up(X), mode="reduced"))
    assert Q.shape == q0.shape and R.shape == r0.shape
    assert np.array_equal(np.tril(R, -1), np.zeros_like(R))
    assert rel_err(R, r0) <= FTOL, rel_err(R, r0)
    assert rel_err(Q, q0) <= FTOL, rel_err(Q, q0)
    assert np.abs(Q.T @ Q - np.eye(n)).max() <= 1e-13
    assert rel_err(Q @ R, X) <= 1e-13
    Ronly = down(K.qr(up(X), mode="r"))
    np.testing.assert_array_equal(Ronly, R)


def test_qr_concatenates_its_operands(K):
    rng = np.random.default_rng(3)
    a, b, c = rng.standard_normal((400, 7)), rng.standard_normal((300, 7)), rng.standard_normal((7, 7))
    R = down(K.qr(up(a), up(b), up(c), mode="r", axis=0))
    assert rel_err(R, np.linalg.qr(np.concatenate([a, b, c], axis=0), mode="r")) <= FTOL
    l, r = rng.standard_normal((500, 4)), rng.standard_normal((500, 5))
    Q, R = down(K.qr(up(l), up(r), mode="reduced", axis=1))
    q0, r0 = np.linalg.qr(np.concatenate([l, r], axis=1))
    assert rel_err(Q, q0) <= FTOL and rel_err(R, r0) <= FTOL
    # stacked triangular factors, the second level of linalg.py:161-171
    r1, r2 = np.linalg.qr(a, mode="r"), np.linalg.qr(b, mode="r")
    Q2, R2 = down(K.qr(up(r1), up(r2), mode="reduced", axis=0))
    q0, r0 = np.linalg.qr(np.concatenate([r1, r2], axis=0))
    assert rel_err(Q2, q0) <= FTOL and rel_err(R2, r0) <= FTOL


def test_qr_ill_conditioned_block_takes_the_shifted_path(K):
    rng = np.random.default_rng(11)
    m, n = 20_000, 24
    U, _ = np.linalg.qr(rng.standard_normal((m, n)))
    V, _ = np.linalg.qr(rng.standard_normal((n, n)))
    X = (U * np.logspace(0, -11, n)) @ V.T            # cond = 1e11: chol(X^T X) breaks down
    Q, R = down(K.qr(up(X), mode="reduced"))
    assert np.abs(Q.T @ Q - np.eye(n)).max() <= 1e-10
    assert rel_err(Q @ R, X) <= 1e-12
    assert np.array_equal(np.tril(R, -1), np.zeros_like(R))


def test_qr_float32_blocks_come_back_as_float32(K):
    rng = np.random.default_rng(5)
    X = rng.standard_normal((1000, 6)).astype(np.float32)
    Q, R = down(K.qr(up(X), mode="reduced"))
    assert Q.dtype == np.float32 and R.dtype == np.float32
    q0, r0 = np.linalg.qr(X)
    assert rel_err(R, r0) <= 1e-5 and rel_err(Q, q0) <= 1e-5


@pytest.mark.parametrize("p", [1, 2, 5, 64, 257, 600])
def test_cholesky(K, p):
    rng = np.random.default_rng(p)
    A = rng.standard_normal((p + 5, p))
    A = A.T @ A + np.eye(p)
    L = down(K.cholesky(up(A)))
    assert np.array_equal(np.triu(L, 1), np.zeros_like(L))
    assert rel_err(L, np.linalg.cholesky(A)) <= FTOL


def test_cholesky_not_positive_definite_raises(K):
    A = np.array([[1.0, 2.0], [2.0, 1.0]])
    with pytest.raises(np.linalg.LinAlgError):
        K.cholesky(up(A))
    with pytest.raises(np.linalg.LinAlgError):
        K.cholesky(up(np.ones((3, 4))))


@pytest.mark.parametrize("p", [1, 2, 9, 64, 257])
def test_svd_of_square_blocks(K, p):
    rng = np.random.default_rng(p + 100)
    A = rng.standard_normal((p, p))
    U, S, VT = down(K.svd(up(A)))
    u0, s0, vt0 = np.linalg.svd(A)
    assert np.all(np.diff(S) <= 0)
    assert rel_err(S, s0) <= 1e-12
    assert rel_err((U * S) @ VT, A) <= 1e-12
    assert np.abs(U.T @ U - np.eye(p)).max() <= 1e-12 and np.abs(VT @ VT.T - np.eye(p)).max() <= 1e-12
    # the factors agree with LAPACK's up to the sign of each singular pair
    sg = np.sign(np.sum(U * u0, axis=0))
    assert rel_err(U * sg, u0) <= 1e-8 and rel_err(VT * sg[:, None], vt0) <= 1e-8


def test_svd_of_a_rank_deficient_block_reconstructs(K):
    rng = np.random.default_rng(9)
    B = rng.standard_normal((12, 3))
    A = B @ B.T
    U, S, VT = down(K.svd(up(A)))
    assert rel_err((U * S) @ VT, A) <= 1e-12
    assert rel_err(S, np.linalg.svd(A, compute_uv=False)) <= 1e-12


# ---------------------------------------------------------------- nums_b200.linalg vs the reference's outputs
def _sign_fix(got_cols, want_cols):
    """Singular pairs are defined up to a common sign; align each column of `got` with `want`."""
    sg = np.sign(np.sum(got_cols * want_cols, axis=0))
    sg[sg == 0] = 1.0
    return sg


@pytest.fixture(scope="module")
def G():
    from conftest import load_golden
    return load_golden("linalg")


def test_wide_qr_block(K):
    rng = np.random.default_rng(21)
    X = rng.standard_normal((4, 9))
    Q, R = down(K.qr(up(X), mode="reduced"))
    q0, r0 = np.linalg.qr(X)
    assert rel_err(Q, q0) <= FTOL and rel_err(R, r0) <= FTOL
    a, b = rng.standard_normal((1, 4)), rng.standard_normal((1, 5))
    Q, R = down(K.qr(up(a), up(b), mode="reduced", axis=1))
    q0, r0 = np.linalg.qr(np.concatenate([a, b], axis=1))
    assert rel_err(Q, q0) <= FTOL and rel_err(R, r0) <= FTOL


def test_tsqr_family_matches_reference_golden(app, G):
    from nums_b200 import linalg
    X, rb, cb = G["X"], int(G["row_block"]), int(G["col_block"])
    XB = app.array(X, block_shape=(rb, cb))
    R = linalg.indirect_tsr(app, XB)
    assert R.block_shape == (cb, cb)
    assert rel_err(R.get(), G["tsr_R"]) <= FTOL
    for fn, tag in ((linalg.indirect_tsqr, "itsqr"), (linalg.direct_tsqr, "dtsqr")):
        Q, R = fn(app, XB)
        assert Q.block_shape == (rb, cb) and R.block_shape == (cb, cb)
        assert rel_err(Q.get(), G[tag + "_Q"]) <= FTOL, tag
        assert rel_err(R.get(), G[tag + "_R"]) <= FTOL, tag
        assert rel_err(Q.get() @ R.get(), X) <= 1e-13            # tests/core/test_linalg.py:155-161
    assert linalg.qr(app, XB)[1].shape == (9, 9)
    XT = app.array(G["X_tall"], block_shape=(512, 32))
    Q, R = linalg.direct_tsqr(app, XT)
    assert rel_err(Q.get(), G["tall_dtsqr_Q"]) <= FTOL and rel_err(R.get(), G["tall_dtsqr_R"]) <= FTOL
    assert rel_err(linalg.indirect_tsr(app, XT).get(), G["tall_tsr_R"]) <= FTOL


def test_svd_and_pca_match_reference_golden_up_to_pair_signs(app, G):
    from nums_b200 import linalg
    X, rb, cb = G["X"], int(G["row_block"]), int(G["col_block"])
    XB = app.array(X, block_shape=(rb, cb))
    U, S, VT = linalg.svd(app, XB)
    u, s, vt = U.get(), S.get(), VT.get()
    assert rel_err(s, G["svd_S"]) <= FTOL
    assert rel_err((u * s) @ vt, X) <= 1e-13                      # tests/core/test_linalg.py:164-168
    sg = _sign_fix(u, G["svd_U"])
    assert rel_err(u * sg, G["svd_U"]) <= FTOL and rel_err(vt * sg[:, None], G["svd_VT"]) <= FTOL
    C = app.cov(XB, rowvar=False)
    assert rel_err(C.get(), G["cov"]) <= FTOL
    V, _, VT2 = linalg.svd(app, C)
    assert np.allclose(V.get(), VT2.get().T)                     # tests/core/test_linalg.py:171-181
    pc = linalg.pca(app, XB).get()
    assert rel_err(pc, (XB @ V).get()) <= 1e-13
    sg = _sign_fix(pc, G["pca"])
    assert rel_err(pc * sg, G["pca"]) <= 1e-9


def test_regressions_inverses_and_cholesky_match_reference_golden(app, G):
    from nums_b200 import linalg
    X, y, rb, cb = G["X"], G["y"], int(G["row_block"]), int(G["col_block"])
    XB, yB = app.array(X, block_shape=(rb, cb)), app.array(y, block_shape=(rb,))
    assert rel_err(linalg.linear_regression(app, XB, yB).get(), G["lr_theta"]) <= FTOL
    assert rel_err(linalg.fast_linear_regression(app, XB, yB).get(), G["fast_lr_theta"]) <= FTOL
    assert rel_err(linalg.ridge_regression(app, XB, yB, lamb=0.0).get(), G["rr_theta_0"]) <= FTOL
    assert rel_err(linalg.ridge_regression(app, XB, yB, lamb=10.0).get(), G["rr_theta_10"]) <= FTOL
    A = app.array(G["spd"], block_shape=(cb, cb))
    L = linalg.cholesky(app, A)
    assert L.block_shape == (cb, cb) and rel_err(L.get(), G["chol_L"]) <= FTOL
    assert rel_err(linalg.inv(app, A).get(), G["inv_spd"]) <= FTOL
    Rinv = linalg.inv_uppertri(app, app.array(G["uppertri"], block_shape=(3, 3)))
    assert Rinv.block_shape == (3, 3)
    assert rel_err(Rinv.get(), G["inv_uppertri"]) <= 1e-9        # cond ~ 1e4: rounding of a different elimination order


def test_tsqr_on_a_large_tall_skinny_array_against_the_oracle(app):
    """Bigger than the golden: 200k x 64 in 4 row blocks, checked against the oracle restatement (LAPACK)."""
    from nums_b200 import linalg
    from oracle import linalg as ol
    n, d, rb = 200_000, 64, 50_000
    X = np.random.default_rng(8).standard_normal((n, d))
    XB = app.array(X, block_shape=(rb, d))
    Q, R = linalg.direct_tsqr(app, XB)
    q0, r0 = ol.direct_tsqr(X, rb, d)
    assert rel_err(R.get(), r0) <= FTOL and rel_err(Q.get(), q0) <= FTOL
    assert rel_err(linalg.indirect_tsr(app, XB).get(), ol.indirect_tsr(X, rb, d)) <= FTOL
    U, S, VT = linalg.svd(app, XB)
    assert rel_err(S.get(), np.linalg.svd(X, compute_uv=False)) <= FTOL


def test_reblocking_round_trip(app):
    x = np.arange(23 * 11, dtype=np.float64).reshape(23, 11)
    B = app.array(x, block_shape=(5, 4))
    C = B.reshape(block_shape=(7, 11))
    assert C.block_shape == (7, 11) and np.array_equal(C.get(), x)
    D = C.reshape(block_shape=(23, 3)).transpose(defer=True).reshape(block_shape=(2, 9))
    assert np.array_equal(D.get(), x.T)
    assert np.array_equal(B.copy().get(), x)
