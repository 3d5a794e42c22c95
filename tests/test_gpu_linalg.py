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
