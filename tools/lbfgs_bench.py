"""BASELINE configs[4]: LinearRegression / PoissonRegression with solver lbfgs on a tall 1024-column shard.
Times a fixed number of L-BFGS iterations and counts the statistics passes they issue."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import device as dv, fused
from nums_b200.app import make_app, set_instance
from nums_b200.blockarray import BlockArray
from nums_b200 import glms

app = set_instance(make_app(num_gpus=1))
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 3_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 10
rng = np.random.default_rng(0)
chunk = 1 << 16
Xc = rng.standard_normal((chunk, d)) / np.sqrt(d)
w = np.linspace(-1, 1, d)
src = dv.from_numpy(Xc, 0)
X = dv.empty((n, d), np.float64, 0)
for lo in range(0, n, chunk):
    hi = min(n, lo + chunk)
    dv.copy_into(X.t[lo:hi], src.t[: hi - lo])
passes = [0]
orig = fused.glm_stats
def counted(*a, **k):
    passes[0] += 1
    return orig(*a, **k)
fused.glm_stats = counted
app.km.backend._remote_functions["glm_stats"] = counted
XB = BlockArray.from_oid(X, (n, d), np.float64, app.km)
for name, cls, yc in (("linear", glms.LinearRegression, Xc @ w + 0.1 * rng.standard_normal(chunk)),
                      ("poisson", glms.PoissonRegression, rng.poisson(np.exp(Xc @ w)).astype(np.float64))):
    ysrc = dv.from_numpy(yc, 0)
    y = dv.empty((n,), np.float64, 0)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        dv.copy_into(y.t[lo:hi], ysrc.t[: hi - lo])
    yB = BlockArray.from_oid(y, (n,), np.float64, app.km)
    for rep in range(2):
        passes[0] = 0
        m = cls(app=app, solver="lbfgs", tol=0.0, max_iter=iters)
        torch.cuda.synchronize(); t0 = time.perf_counter()
        m.fit(XB, yB)
        b = m._beta.get()
        torch.cuda.synchronize(); dt = time.perf_counter() - t0
    nbytes = n * (8 * d + 8)
    print("%-8s n=%d d=%d: %d iterations, %d statistics passes, %.1f ms total, %.2f ms per pass incl. host, "
          "%.0f GB/s per pass" % (name, n, d, m.n_iter_, passes[0], dt * 1e3, dt * 1e3 / passes[0],
                                  passes[0] * nbytes / dt / 1e9), flush=True)
