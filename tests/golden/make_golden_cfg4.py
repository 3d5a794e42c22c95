"""Golden vectors on the BENCHMARKED kernel path (last row): a 2*10^5 x 256 slice of BASELINE
config 4 and the iteration counts of the README fit, produced by the REFERENCE ITSELF (oracle/_ref, the unmodified
nums package on its SerialBackend).  Run in the build container:  python tests/golden/make_golden_cfg4.py

Inputs are not stored (410 MB): they are the reference's own block-RNG stream — nps.random.seed(1337), block j drawn
from Generator(PCG64(1337).jumped(j)) — which the tests regenerate bit for bit.  Stored: beta, beta0, the number of
Newton iterations, and g / H / objective at beta = 0 through the reference's own model methods.
"""
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import refharness  # noqa: E402


class CountedInv:
    def __init__(self, app):
        self.km, self.n = app.km, 0

    def __enter__(self):
        inner = self.km.methods["inv"]

        def counted(*a, **k):
            self.n += 1
            return inner(*a, **k)
        self._inner = inner
        self.km.methods["inv"] = counted
        return self

    def __exit__(self, *exc):
        self.km.methods["inv"] = self._inner


def readme_data(nps, size, d, shift):
    X1 = nps.random.randn(size // 2, d)
    y1 = nps.zeros(shape=(size // 2,), dtype=int)
    X2 = nps.random.randn(size // 2, d) + shift
    y2 = nps.ones(shape=(size // 2,), dtype=int)
    return nps.concatenate([X1, X2], axis=0), nps.concatenate([y1, y2], axis=0)


def main():
    warnings.simplefilter("ignore")
    app = refharness.use(refharness.serial_app(1))
    import nums.numpy as nps
    from nums.models.glms import LogisticRegression
    out = {}
    # ---- config 4 slice: README generalised to d = 256, delta = 2 sqrt(2/256)
    n, d = 200_000, 256
    nps.random.seed(1337)
    X, y = readme_data(nps, n, d, 2.0 * np.sqrt(2.0 / d))
    m = LogisticRegression(solver="newton-cg", penalty="l2", C=10)
    with CountedInv(app) as c:
        m.fit(X, y)
    out["cfg4_n"], out["cfg4_d"] = np.int64(n), np.int64(d)
    out["cfg4_row_blocks"] = np.array([b.shape[0] for b in X.blocks[:, 0]], dtype=np.int64)
    out["cfg4_beta"], out["cfg4_beta0"], out["cfg4_iters"] = m._beta.get(), m._beta0.get(), np.int64(c.n)
    out["cfg4_x_checksum"] = np.float64(nps.sum(X).get())
    free = LogisticRegression(solver="newton", penalty="none")
    free.fit(X[:1000], y[:1000])      # sets the penalty members; the statistics below are of the full slice
    Xa = app.concatenate([X, app.ones(shape=(n, 1), block_shape=(X.block_shape[0], 1), dtype=X.dtype)], axis=1,
                         axis_block_size=X.block_shape[1])
    beta0 = app.zeros((d + 1,), (d,), dtype=X.dtype)
    mu = free.forward(Xa, beta0)
    out["cfg4_g0"] = free.gradient(Xa, y, mu, beta=beta0).get()
    out["cfg4_H0"] = free.hessian(Xa, y, mu).get()
    out["cfg4_obj0"] = free.objective(Xa, y, beta0, mu).get()
    # ---- README fit: iteration counts at the golden's size and at the README's own size
    for size in (20_000, 1_000_000):
        nps.random.seed(1337)
        Xr, yr = readme_data(nps, size, 2, 2.0)
        mr = LogisticRegression(solver="newton-cg", penalty="l2", C=10)
        with CountedInv(app) as c:
            mr.fit(Xr, yr)
        out["readme_%d_iters" % size] = np.int64(c.n)
        out["readme_%d_beta" % size] = np.concatenate([mr._beta.get(), [mr._beta0.get()]])
    path = os.path.join(HERE, "glm_cfg4_d256.npz")
    np.savez_compressed(path, **out)
    print("wrote %s (%.1f KB): cfg4 iterations %d, README iterations %s" % (
        path, os.path.getsize(path) / 1024, out["cfg4_iters"], [int(out["readme_%d_iters" % s]) for s in (20_000, 1_000_000)]))


if __name__ == "__main__":
    main()
