"""ORACLE — test infrastructure only.  Never imported by the nums_b200 package.

A NumPy restatement of the tall-skinny linear algebra of the reference (nums/core/linalg.py, v0.2.8) on explicit
lists of row blocks: which LAPACK-backed NumPy call (np.linalg.qr / svd / inv / cholesky, as in
nums/core/kernel/numpy_kernel.py:455-472) runs on which concatenation of blocks, in which order.  Matrix products
between blocked operands follow BlockArray.tensordot's FIFO pairwise reduction (blockarray.py:643-689), restated
in oracle/glm.py:tree_sum.

Pinned by tests/test_oracle_golden.py against tests/golden/linalg.npz, produced by running the reference itself
(tests/golden/make_golden.py).
"""
import numpy as np

from . import np_kernel
from .glm import split_rows, tree_sum


def col_blocks(arr, col_block):
    return [arr[:, j:j + col_block] for j in range(0, arr.shape[1], col_block)]


def kernel_qr(*arrays, mode="reduced", axis=None):
    """numpy_kernel.py:455-461"""
    arr = np.concatenate(arrays, axis=axis) if len(arrays) > 1 else arrays[0]
    return np.linalg.qr(arr, mode=mode)


def qr_tree_reduce(rs):
    """linalg.py:27-52"""
    queue = list(rs)
    while len(queue) > 1:
        a = queue.pop(0)
        b = queue.pop(0)
        queue.append(kernel_qr(a, b, mode="r", axis=0))
    return queue[0]


def blocked_matmul_rows(row_blocks, B, inner_block):
    """(row-blocked A) @ B with A's columns cut at inner_block: per row block the partial products over the
    column blocks are summed pairwise, oldest first (blockarray.py:643-689)."""
    out = []
    for A in row_blocks:
        parts = [A[:, k:k + inner_block] @ B[k:k + inner_block] for k in range(0, A.shape[1], inner_block)]
        out.append(tree_sum(parts))
    return out


def blocked_tn(row_blocks_a, row_blocks_b):
    """A^T @ B for two arrays blocked along the rows only."""
    return tree_sum([a.T @ b for a, b in zip(row_blocks_a, row_blocks_b)])


def indirect_tsr(X, row_block, col_block):
    """linalg.py:55-103"""
    rs = [kernel_qr(*col_blocks(blk, col_block), mode="r", axis=1) for blk in split_rows(X, row_block)]
    return qr_tree_reduce(rs)


def indirect_tsqr(X, row_block, col_block):
    """linalg.py:106-127"""
    R = indirect_tsr(X, row_block, col_block)
    R_inv = np.linalg.inv(R)
    Q = np.concatenate(blocked_matmul_rows(split_rows(X, row_block), R_inv, col_block), axis=0)
    return Q, R


def direct_tsqr(X, row_block, col_block):
    """linalg.py:130-208"""
    qs, rs = [], []
    for blk in split_rows(X, row_block):
        q, r = kernel_qr(*col_blocks(blk, col_block), mode="reduced", axis=1)
        qs.append(q)
        rs.append(r)
    q2, r2 = kernel_qr(*rs, mode="reduced", axis=0)
    k = rs[0].shape[0]
    q2_blocks = [q2[i:i + k] for i in range(0, q2.shape[0], k)]
    Q = np.concatenate([q @ q2b for q, q2b in zip(qs, q2_blocks)], axis=0)
    return Q, r2


def svd(X, row_block, col_block):
    """linalg.py:211-229"""
    Q, R = direct_tsqr(X, row_block, col_block)
    u, s, vt = np.linalg.svd(R)
    # Q keeps the row blocks of X and a single column block; R_U is cut at col_block
    U = np.concatenate(blocked_matmul_rows(split_rows(Q, row_block), u, col_block), axis=0)
    return U, s, vt


def mean_axis0(X, row_block):
    """application.py:499-529 with axis=0, keepdims=True: per-block sums, pairwise reduced, divided by the count."""
    return tree_sum([np_kernel.reduce_axis("sum", blk, 0, True, False) for blk in split_rows(X, row_block)]) / X.shape[0]


def cov(X, row_block, col_block):
    """application.py:531-540 with rowvar=False"""
    Xc = X - mean_axis0(X, row_block)
    blocks = split_rows(Xc, row_block)
    return blocked_tn(blocks, blocks) / (X.shape[0] - 1)


def pca(X, row_block, col_block):
    """linalg.py:232-235: the covariance keeps X's column blocking on both axes"""
    C = cov(X, row_block, col_block)
    U, _, _ = svd(C, col_block, col_block)
    return np.concatenate(blocked_matmul_rows(split_rows(X, row_block), U, col_block), axis=0)


def _theta(X, y, Q, R, row_block, col_block):
    R_inv = np.linalg.inv(R)
    qty = blocked_tn(split_rows(Q, row_block), split_rows(y, row_block))
    parts = [R_inv[:, k:k + col_block] @ qty[k:k + col_block] for k in range(0, R_inv.shape[1], col_block)]
    return tree_sum(parts)


def fast_linear_regression(X, y, row_block, col_block):
    """linalg.py:468-480"""
    Q, R = indirect_tsqr(X, row_block, col_block)
    return _theta(X, y, Q, R, row_block, col_block)


def linear_regression(X, y, row_block, col_block):
    """linalg.py:483-496"""
    Q, R = direct_tsqr(X, row_block, col_block)
    return _theta(X, y, Q, R, row_block, col_block)


def ridge_regression(X, y, lamb, row_block, col_block):
    """linalg.py:499-515"""
    R = indirect_tsr(X, row_block, col_block)
    d = X.shape[1]
    rtr = tree_sum([R[k:k + col_block].T @ R[k:k + col_block] for k in range(0, d, col_block)])
    M_inv = np.linalg.inv(lamb * np.eye(d) + rtr)
    xty = blocked_tn(split_rows(X, row_block), split_rows(y, row_block))
    return tree_sum([M_inv[:, k:k + col_block] @ xty[k:k + col_block] for k in range(0, d, col_block)])


def inv_uppertri(R, block):
    """linalg.py:264-441: the result equals inv(R); the blocked recurrence is restated for the summation order."""
    n = R.shape[0]
    if block >= n:
        return np.linalg.inv(R)
    g = (n + block - 1) // block
    sl = [slice(i * block, min((i + 1) * block, n)) for i in range(g)]
    out = R.copy()
    out[sl[0], sl[0]] = np.linalg.inv(R[sl[0], sl[0]])
    for k in range(1, g):
        r11_inv = np.linalg.inv(out[sl[k], sl[k]])
        out[sl[k], sl[k]] = r11_inv
        above = [out[sl[i], sl[k]].copy() for i in range(k)]
        for i in range(k):
            acc = None
            for j in range(k):                       # sum_reduce is a plain left-to-right sum (numpy_kernel.py:291)
                term = (0.0 - out[sl[i], sl[j]]) @ above[j]
                acc = term if acc is None else acc + term
            out[sl[i], sl[k]] = acc @ r11_inv
    return out
