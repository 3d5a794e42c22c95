"""Time of the Newton update (penalty, max|g|, blocked Cholesky solve, beta - delta) at p = d + 1, CUDA events.
Usage: python tools/newton_update_time.py [d]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import device as dv, fused
from nums_b200.app import make_app, set_instance

set_instance(make_app(num_gpus=1))
d = int(sys.argv[1]) if len(sys.argv) > 1 else 256
p = d + 1
rng = np.random.default_rng(0)
A = rng.standard_normal((4 * p, p))
H = A.T @ A
g = rng.standard_normal(p)
stats_h = np.concatenate([H.reshape(-1), g, [0.0]])
beta = dv.from_numpy(rng.standard_normal(p) * 0.01, 0)
l2 = dv.from_numpy(np.full(p, 0.1), 0)
ref = dv.to_numpy(beta) - np.linalg.solve(H + np.diag(np.full(p, 0.1)), g + 0.1 * dv.to_numpy(beta))


def run():
    stats = dv.from_numpy(stats_h, 0)          # the update consumes its statistics block
    return fused.glm_newton_step(stats, l2, beta)


out, verdict = run()
print("max |beta_new - numpy| = %.3e" % np.max(np.abs(dv.to_numpy(out) - ref)))
stats_list = [dv.from_numpy(stats_h, 0) for _ in range(24)]
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for s in stats_list[:4]:
    fused.glm_newton_step(s, l2, beta)
torch.cuda.synchronize()
e0.record()
for s in stats_list[4:]:
    fused.glm_newton_step(s, l2, beta)
e1.record()
torch.cuda.synchronize()
print("Newton update at p = %d: %.3f ms per call (20 calls back to back, host overhead included)" % (p, e0.elapsed_time(e1) / 20))
