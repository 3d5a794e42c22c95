"""Times nb_glm_stats (gradient pass and Newton pass) over a range of feature counts; one line per (d, pass)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import device as dv, fused
from nums_b200.app import make_app, set_instance

set_instance(make_app(num_gpus=1))
ds = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [8, 16, 28, 32, 48, 64, 128, 256]
total = int(float(sys.argv[2])) if len(sys.argv) > 2 else 4 * 10 ** 8      # doubles of X
hbm = json.load(open("MEASURED_PEAKS.json")).get("hbm_gbs", 6555.5) if os.path.exists("MEASURED_PEAKS.json") else 6555.5
rng = np.random.default_rng(0)
for d in ds:
    n = total // d
    base = dv.from_numpy(rng.standard_normal(1 << 22), 0)
    X = dv.empty((n * d,), np.float64, 0)
    for lo in range(0, n * d, base.size):
        hi = min(n * d, lo + base.size)
        dv.copy_into(X.t[lo:hi], base.t[: hi - lo])
    X = dv.DeviceBlock(X.t.view(n, d))
    y = dv.from_numpy((rng.random(n) < 0.5).astype(np.int64), 0)
    beta = dv.from_numpy(rng.standard_normal(d + 1) * 0.01, 0)
    for hess in (False, True):
        for _ in range(2):
            fused.glm_stats(X, y, beta, "logistic", hess)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            fused.glm_stats(X, y, beta, "logistic", hess)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        gbs = n * (8 * d + 8) / ms / 1e6
        tf = n * d * (d + 1) / ms / 1e9 if hess else 0.0
        print("d=%4d n=%9d %-8s %8.3f ms  %7.0f GB/s (%4.1f%% of HBM)  %5.2f TFLOP/s" % (
            d, n, "newton" if hess else "gradient", ms, gbs, 100 * gbs / hbm, tf), flush=True)
    del X, y
