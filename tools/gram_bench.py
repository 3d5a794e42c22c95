"""Times X^T diag(w) X through nb_matmul_f64 for tall-skinny X (with / without weights)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import _lib, device as dv
from nums_b200.app import make_app, set_instance
set_instance(make_app(num_gpus=1))
lib = _lib.load()
rng = np.random.default_rng(0)
ds = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [256, 384, 512, 1024]
total = int(float(sys.argv[2])) if len(sys.argv) > 2 else 4 * 10 ** 8
for d in ds:
    n = total // d
    base = dv.from_numpy(rng.standard_normal(1 << 22), 0)
    X = dv.empty((n * d,), np.float64, 0)
    for lo in range(0, n * d, base.size):
        hi = min(n * d, lo + base.size)
        dv.copy_into(X.t[lo:hi], base.t[: hi - lo])
    w = dv.from_numpy(rng.random(n) + 0.5, 0)
    out = dv.empty((d, d), np.float64, 0)
    sc = dv.scratch(lib.nb_matmul_scratch_bytes(0, d, d, n), 0)
    for wp, tag in ((None, "plain"), (w.ptr, "weighted")):
        def run():
            _lib.check(lib.nb_matmul_f64(0, d, d, n, X.ptr, 1, d, X.ptr, d, 1, wp, out.ptr, d, sc.data_ptr(), sc.numel(),
                                         dv.stream_ptr(0)))
        for _ in range(2): run()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): run()
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        print("d=%4d n=%8d %-8s %8.3f ms  %6.2f TFLOP/s (n d (d+1))" % (d, n, tag, ms, n * d * (d + 1) / ms / 1e9), flush=True)
    del X
