"""Times the square FP64 matmul (BASELINE config 3) in the four operand layouts."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import cuda_kernel, device as dv
from nums_b200.app import make_app, set_instance
set_instance(make_app(num_gpus=1))
K = cuda_kernel.KernelCls()
m = int(sys.argv[1]) if len(sys.argv) > 1 else 10 ** 4
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
rng = np.random.default_rng(0)
base = dv.from_numpy(rng.standard_normal(1 << 22), 0)
def big(count):
    out = dv.empty((count,), np.float64, 0)
    for lo in range(0, count, base.size):
        hi = min(count, lo + base.size)
        dv.copy_into(out.t[lo:hi], base.t[: hi - lo])
    return out
A = dv.DeviceBlock(big(m * m).t.view(m, m)); B = dv.DeviceBlock(big(m * m).t.view(m, m))
for ta, tb in ((False, False), (True, False), (False, True), (True, True)):
    f = lambda: K.bop("tensordot", A, B, ta, tb, 1)
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): f()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print("A%s @ B%s  %d^3  %.2f ms  %.2f TFLOP/s" % ("^T" if ta else "", "^T" if tb else "", m, ms, 2.0 * m ** 3 / ms / 1e9), flush=True)
