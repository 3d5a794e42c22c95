"""Tall-skinny linear algebra on device blocks: the host-side mirror of nums/core/linalg.py.

Same functions, arguments and block structure as the reference -- one `km.qr` per row of blocks, the FIFO pairwise
reduction of the R factors (linalg.py:27-52), the second-level QR of the stacked R's (linalg.py:160-171), `km.svd`
of the d x d factor (linalg.py:221-223) -- because the block structure fixes the factors (including the signs of
diag(R), which the device QR takes from LAPACK's convention) that the oracle is compared with.  Every `km.*` call
lands in KernelCls (cuda_kernel.py) and from there in libnumsb200.so; nothing is computed on the host.
"""
import numpy as np

from .blockarray import BlockArray
from .grid import ArrayGrid


def qr(app, X):
    # linalg.py:23-24
    return indirect_tsqr(app, X)


def _sys(entry, grid_shape, num_returns=1):
    return {"grid_entry": entry, "grid_shape": grid_shape, "options": {"num_returns": num_returns}}


def _row_oids(X, i):
    return [X.blocks[i, j].oid for j in range(X.grid.grid_shape[1])]


def _single_block(app, oid, shape, dtype):
    out = BlockArray(ArrayGrid(shape=shape, block_shape=shape, dtype=np.dtype(dtype).name), app.km)
    out.blocks[(0,) * len(shape)].oid = oid
    return out


def _qr_tree_reduce(app, items, result_entry, result_grid_shape):
    """linalg.py:27-52: R factors are merged two at a time, oldest first; the last merge lands on the result's
    grid entry."""
    if len(items) == 1:
        return items[0][0]
    queue = list(items)
    while len(queue) > 1:
        a_oid, a_entry, a_gs = queue.pop(0)
        b_oid, _, _ = queue.pop(0)
        entry, gs = (result_entry, result_grid_shape) if not queue else (a_entry, a_gs)
        merged = app.km.qr(a_oid, b_oid, mode="r", axis=0, syskwargs=_sys(entry, gs))
        queue.append((merged, entry, gs))
    oid, entry, gs = queue.pop(0)
    assert entry == result_entry and gs == result_grid_shape
    return oid


def _check_untransposed(X):
    for entry in X.grid.get_entry_iterator():
        if X.blocks[entry].transposed:
            raise NotImplementedError("tall-skinny factorisations take an array without deferred transposes")


def indirect_tsr(app, X, reshape_output=True):
    """linalg.py:55-103: R of X = QR from per-row-block R factors and their pairwise reduction."""
    assert X.ndim == 2
    assert X.block_shape[0] >= X.shape[1]
    _check_untransposed(X)
    rows = X.grid.grid_shape[0]
    items = []
    for i in range(rows):
        entry, gs = (i, 0), (rows, 1)
        items.append((app.km.qr(*_row_oids(X, i), mode="r", axis=1, syskwargs=_sys(entry, gs)), entry, gs))
    d = X.shape[1]
    R = _single_block(app, _qr_tree_reduce(app, items, (0, 0), (1, 1)), (d, d), X.dtype)
    R_block_shape = (X.block_shape[1], X.block_shape[1])
    if (d, d) != R_block_shape and reshape_output:
        return R.reshape((d, d), block_shape=R_block_shape)
    return R


def indirect_tsqr(app, X, reshape_output=True):
    """linalg.py:106-127: Q = X R^-1 with R from indirect_tsr."""
    d = X.shape[1]
    R_shape, R_block_shape = (d, d), (X.block_shape[1], X.block_shape[1])
    tsR = indirect_tsr(app, X, reshape_output=False)
    tsR_inv = inv(app, tsR)
    if R_shape != R_block_shape:
        R_inv = tsR_inv.reshape(R_shape, block_shape=R_block_shape)
        R = tsR.reshape(R_shape, block_shape=R_block_shape) if reshape_output else tsR
    else:
        R_inv, R = tsR_inv, tsR
    return X @ R_inv, R


def direct_tsqr(app, X, reshape_output=True):
    """linalg.py:130-208: reduced QR per row of blocks, one QR of the stacked R factors, Q_i @ Q2_i."""
    assert X.ndim == 2
    _check_untransposed(X)
    shape, block_shape = X.shape, X.block_shape
    rows = X.grid.grid_shape[0]
    d = shape[1]
    Q_oids, R_oids, ks = [], [], []
    for i in range(rows):
        m = X.grid.get_block_shape((i, 0))[0]
        ks.append(min(m, d))
        q_oid, r_oid = app.km.qr(*_row_oids(X, i), mode="reduced", axis=1, syskwargs=_sys((i, 0), (rows, 1), 2))
        Q_oids.append(q_oid)
        R_oids.append(r_oid)
    Q2_oid, R2_oid = app.km.qr(*R_oids, mode="reduced", axis=0, syskwargs=_sys((0, 0), (1, 1), 2))
    Q2 = app.vec_from_oids([Q2_oid], shape=(int(sum(ks)), d), block_shape=(ks[0], d), dtype=X.dtype)
    Q2_oids = [b.oid for b in Q2.blocks.flatten()]
    Q = app.zeros(shape=shape, block_shape=(block_shape[0], d), dtype=X.dtype)
    for i, entry in enumerate(Q.grid.get_entry_iterator()):
        Q.blocks[entry].oid = app.km.bop("tensordot", Q_oids[i], Q2_oids[i], a1_T=False, a2_T=False, axes=1,
                                         syskwargs={"grid_entry": entry, "grid_shape": Q.grid.grid_shape})
    R_shape, R_block_shape = (d, d), (block_shape[1], block_shape[1])
    R = app.vec_from_oids([R2_oid], shape=R_shape, block_shape=R_shape, dtype=X.dtype)
    if R_shape != R_block_shape and reshape_output:
        R = R.reshape(R_shape, block_shape=R_block_shape)
    if Q.shape != block_shape or not reshape_output:
        Q = Q.reshape(shape, block_shape=block_shape)
    return Q, R


def svd(app, X):
    """linalg.py:211-229: U = Q @ U_R with Q, R from direct_tsqr and R = U_R S V^T on one block."""
    assert X.ndim == 2
    d = X.shape[1]
    R_shape, R_block_shape = (d, d), (X.block_shape[1], X.block_shape[1])
    Q, R = direct_tsqr(app, X, reshape_output=False)
    assert R.shape == R.block_shape
    u_oid, s_oid, vt_oid = app.km.svd(R.blocks[0, 0].oid, syskwargs=_sys((0, 0), (1, 1), 3))
    R_U = app.vec_from_oids([u_oid], R_shape, R_block_shape, X.dtype)
    S = app.vec_from_oids([s_oid], R_shape[:1], R_block_shape[:1], X.dtype)
    VT = app.vec_from_oids([vt_oid], R_shape, R_block_shape, X.dtype)
    return Q @ R_U, S, VT


def pca(app, X):
    # linalg.py:232-235
    C = app.cov(X, rowvar=False)
    U, _, _ = svd(app, C)
    return X @ U


def _on_single_block(app, X, kernel_name):
    assert X.ndim == 2 and X.shape[0] == X.shape[1]
    single = X.shape == X.block_shape
    out = X.copy() if single else X.reshape(block_shape=X.shape)
    blk = out.blocks[0, 0]
    oid = blk.oid
    if blk.transposed:
        oid = app.km.transpose(oid, syskwargs=blk.syskwargs())
        blk.transposed = False
    blk.oid = getattr(app.km, kernel_name)(oid, syskwargs=_sys((0, 0), (1, 1)))
    return out if single else out.reshape(block_shape=X.block_shape)


def inv(app, X):
    # linalg.py:238-261: gathered into one block, inverted there, re-blocked
    return _on_single_block(app, X, "inv")


def cholesky(app, X):
    # linalg.py:444-465
    return _on_single_block(app, X, "cholesky")


def inv_uppertri(app, X):
    """linalg.py:264-441: blocked inverse of an upper-triangular matrix.  The leading block R_TL already holds its
    inverse; each step inverts the next diagonal block R_11 and replaces the column above it by
    -inv(R_TL) @ R_01 @ inv(R_11)."""
    assert X.ndim == 2 and X.shape[0] == X.shape[1], "This function only accepts square matrices"
    if X.shape == X.block_shape or X.block_shape[0] != X.block_shape[1]:
        return inv(app, X)
    gs = X.grid.grid_shape
    R = X.copy()
    zeros = app.zeros(X.shape, X.block_shape, X.dtype)
    R.blocks[0, 0].oid = app.km.inv(R.blocks[0, 0].oid, syskwargs={"grid_entry": (0, 0), "grid_shape": gs})
    for k in range(1, gs[0]):
        r11_inv = app.km.inv(R.blocks[k, k].oid, syskwargs={"grid_entry": (k, k), "grid_shape": gs})
        R.blocks[k, k].oid = r11_inv
        above = [R.blocks[i, k].oid for i in range(k)]
        first = []
        for i in range(k):
            terms = []
            for j in range(k):
                sk = {"grid_entry": (i, j), "grid_shape": gs}
                neg = app.km.bop("subtract", zeros.blocks[i, j].oid, R.blocks[i, j].oid, False, False, axes=1,
                                 syskwargs=sk)
                terms.append(app.km.bop("tensordot", neg, above[j], False, False, axes=1,
                                        syskwargs={"grid_entry": (j, k), "grid_shape": gs}))
            first.append(app.km.sum_reduce(*terms, syskwargs={"grid_entry": (i, k), "grid_shape": gs}))
        for i in range(k):
            R.blocks[i, k].oid = app.km.bop("tensordot", first[i], r11_inv, False, False, axes=1,
                                            syskwargs={"grid_entry": (i, k), "grid_shape": gs})
    return R.reshape(block_shape=X.block_shape)


def _theta_from_qr(app, X, y, Q, R):
    d = X.shape[1]
    R_shape, R_block_shape = (d, d), (X.block_shape[1], X.block_shape[1])
    R_inv = inv(app, R)
    if R_shape != R_block_shape:
        R_inv = R_inv.reshape(R_shape, block_shape=R_block_shape)
    return R_inv @ (Q.transpose(defer=True) @ y)


def fast_linear_regression(app, X, y):
    # linalg.py:468-480
    assert X.ndim == 2 and y.ndim == 1
    Q, R = indirect_tsqr(app, X, reshape_output=False)
    return _theta_from_qr(app, X, y, Q, R)


def linear_regression(app, X, y):
    # linalg.py:483-496
    assert X.ndim == 2 and y.ndim == 1
    Q, R = direct_tsqr(app, X, reshape_output=False)
    return _theta_from_qr(app, X, y, Q, R)


def ridge_regression(app, X, y, lamb):
    # linalg.py:499-515: theta = (lamb I + R^T R)^-1 X^T y with R from indirect_tsr
    assert X.ndim == 2 and y.ndim == 1
    assert lamb >= 0
    d = X.shape[1]
    R_block_shape = (X.block_shape[1], X.block_shape[1])
    R = indirect_tsr(app, X)
    lamb_eye = app.array(lamb * np.eye(d), block_shape=R_block_shape)
    return inv(app, lamb_eye + R.transpose(defer=True) @ R) @ (X.transpose(defer=True) @ y)
