"""Where the wall time of the README fit (10^6 x 2, 8 Newton iterations) goes: cProfile of one warm fit."""
import cProfile, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200.app import make_app, set_instance
from nums_b200.glms import LogisticRegression

app = set_instance(make_app(num_gpus=1))
Xr = np.random.default_rng(3).standard_normal((10 ** 6, 2)); Xr[500_000:] += 2.0
yr = np.concatenate([np.zeros(500_000, dtype=np.int64), np.ones(500_000, dtype=np.int64)])
Xb, yb = app.array(Xr, block_shape=(500_000, 2)), app.array(yr, block_shape=(500_000,))


def fit():
    m = LogisticRegression(app=app, solver="newton", penalty="l2", C=10)
    m.fit(Xb, yb)
    return m


for _ in range(3):
    fit()
torch.cuda.synchronize()
ts = []
for _ in range(10):
    t0 = time.perf_counter(); m = fit(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
print("README fit: min %.3f ms, median %.3f ms, %d iterations" % (min(ts), sorted(ts)[5], m.n_iter_))
pr = cProfile.Profile(); pr.enable(); fit(); torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
