import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from nums_b200 import device as dv, fused
from nums_b200.app import make_app, set_instance
set_instance(make_app(num_gpus=1))
n, d = 12_500_000, 256
g = torch.Generator(device="cuda:0"); g.manual_seed(1)
X = dv.DeviceBlock(torch.randn((n, d), dtype=torch.float64, device="cuda:0", generator=g))
y = dv.DeviceBlock((torch.rand((n,), dtype=torch.float64, device="cuda:0", generator=g) < 0.5).to(torch.int64))
beta = dv.from_numpy(np.random.default_rng(0).standard_normal(d + 1) * 0.01, 0)
def timed(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
for fam in ("logistic", "poisson", "linear"):
    ms = timed(lambda: fused.glm_stats(X, y, beta, fam, False))
    s = dv.to_numpy(fused.glm_stats(X, y, beta, fam, False))
    print("%s gradient pass (%s): %.3f ms  %.0f GB/s  checksum %.12e" % (fam, "ring", ms, n * (8 * d + 8) / ms / 1e6, float(np.sum(s))), flush=True)
