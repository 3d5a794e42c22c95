"""ORACLE — test infrastructure only.  Never imported by the nums_b200 package.

Exact integer restatement of the int8 Hessian (nums_b200/csrc/glm_ozaki.cu) for the LINEAR family, where the weights
are 1 and every step of the device path is integer or exactly representable arithmetic: the power-of-two column
scales, the rounding of x 2^q to a 39-bit integer (ties to even, with the dither byte of the four-plane variant), the
balanced base-256 digits, the plane products, their levels, and the conversion back to float64.  What the reference
computes there is X^T X (nums/models/glms.py:318-327, the Gram matrix of LinearRegression.hessian); this module says
which float64 the device must return for it BIT FOR BIT, so the parity bar "bit-exact for integer work" applies to the
whole digit pipeline, not only a tolerance against the FP64 result.
"""
import numpy as np

M32 = 0xFFFFFFFF

# products (p, q), p <= q, kept by each variant: level p + q <= 5 (eleven), <= 4 (nine), four planes (eight)
PRODUCTS = {
    (5, False): [(p, q) for p in range(5) for q in range(p, 5) if p + q <= 5],
    (5, True): [(p, q) for p in range(5) for q in range(p, 5) if p + q <= 4],
    (4, True): [(p, q) for p in range(4) for q in range(p, 4) if p + q <= 4],
}


def column_exponents(X):
    """nb_glm_prepare: max|x[:, i]| < 2^e_i (frexp of the column maximum); None for an all-zero column."""
    m = np.max(np.abs(X), axis=0)
    e = np.frexp(m)[1].astype(np.int64)
    return [int(v) if mv != 0.0 else None for v, mv in zip(e, m)]


def dither_bytes(n, d, seed=0):
    """Four planes: byte hash(row) ^ key(lane) added to the lowest byte before it is cut off; lane = the thread that owns
    the column (columns 4 l .. 4 l + 3 and 128 + 4 l .. 128 + 4 l + 3)."""
    rows = (np.arange(n, dtype=np.uint64) + np.uint64(seed)) & np.uint64(M32)
    h = (rows * np.uint64(0x9E3779B1)) & np.uint64(M32)
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x85EBCA77)) & np.uint64(M32)
    rowkey = (h >> np.uint64(24)).astype(np.int64)
    lanes = (np.arange(d, dtype=np.uint64) % np.uint64(128)) // np.uint64(4)
    lanekey = (((lanes * np.uint64(0x9E3779B1)) & np.uint64(M32)) >> np.uint64(24)).astype(np.int64)
    return rowkey[:, None] ^ lanekey[None, :]


def digit_planes(X, nplanes, ce=None, seed=0):
    """-> planes [nplanes][n][d] (float64 holding integers in [-128, 127], plane 0 most significant) of the rows X under
    the column exponents `ce` of the whole block (default: X's own)."""
    n, d = X.shape
    ce = column_exponents(X) if ce is None else ce
    q = np.array([38 - c if c is not None else 0 for c in ce], dtype=np.int64)      # ew = 0: sqrt(w) = 1
    live = np.array([c is not None for c in ce])
    y = np.where(live[None, :], np.ldexp(X, q[None, :]), 0.0)     # exact: |y| < 2^38
    if nplanes == 5:
        w = np.rint(y).astype(np.int64) + 0x8080808080            # fma against an even magic constant: ties to even
        digs = [((w >> (8 * s)) & 255) - 128 for s in range(5)]   # digit s, least significant first
    else:
        u = dither_bytes(n, d, seed)
        # y + u is exact in float64; the fma rounds it to an integer, ties to even; the bias sits on bytes 1..4 only
        w = np.rint(y + u.astype(np.float64)).astype(np.int64) + 0x8080808000
        digs = [((w >> (8 * s)) & 255) - 128 for s in range(1, 5)]
    return np.stack(digs[::-1]).astype(np.float64)


def exact_gram_linear(X, nplanes=5, no_level5=True, chunk_rows=None):
    """The d x d corner of H that nb_glm_stats(... | NB_GLM_INT8_HESSIAN) returns for the linear family.
    chunk_rows: rows per chunk of the device pass (the dither seed of a chunk is its first row / 64)."""
    n, d = X.shape
    products = PRODUCTS[(nplanes, True if nplanes == 4 else bool(no_level5))]
    top = 2 * (nplanes - 1)
    vbits = 38 if nplanes == 5 else 30
    ce = column_exponents(X)
    lev = [np.zeros((d, d), dtype=np.int64) for _ in range(top + 1)]
    step = chunk_rows or n
    for c0 in range(0, n, step):
        planes = digit_planes(X[c0:c0 + step], nplanes, ce, seed=c0 // 64)
        for (p, r) in products:
            P = np.rint(planes[p].T @ planes[r]).astype(np.int64)         # float64 BLAS on small integers: exact below 2^53
            lev[p + r] += P if p == r else P + P.T
    acc = np.zeros((d, d))
    for L in range(top, -1, -1):                                  # smallest terms first, each exact in float64
        acc = acc + np.ldexp(lev[L].astype(np.float64), 8 * (top - L))
    live = np.array([c is not None for c in ce])
    back = np.array([-(vbits - c) if c is not None else 0 for c in ce], dtype=np.int64)
    H = np.ldexp(np.ldexp(acc, back[:, None]), back[None, :])
    return np.where(live[:, None] & live[None, :], H, 0.0)
