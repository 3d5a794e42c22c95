"""Times the HBM-bound vector kernels of BASELINE config 2 (dot, X@b, X^T r, reductions)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import cuda_kernel, device as dv
from nums_b200.app import make_app, set_instance
set_instance(make_app(num_gpus=1))
K = cuda_kernel.KernelCls()
n = 10 ** 8
rng = np.random.default_rng(0)
base = dv.from_numpy(rng.standard_normal(1 << 22), 0)
x = dv.empty((n,), np.float64, 0)
for lo in range(0, n, base.size):
    hi = min(n, lo + base.size)
    dv.copy_into(x.t[lo:hi], base.t[: hi - lo])
def timeit(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
def show(name, ms, nbytes): print("%-44s %.3f ms %6.0f GB/s" % (name, ms, nbytes / ms / 1e6), flush=True)
show("dot x.x (1e8)", timeit(lambda: K.bop("tensordot", x, x, False, False, 1)), 8 * n)
for d in (64, 256, 1024):
    rows = n // d
    X = dv.DeviceBlock(x.t[: rows * d].view(rows, d))
    b = dv.from_numpy(rng.standard_normal(d), 0)
    r = dv.DeviceBlock(x.t[:rows])
    show("X@b   (%d x %d)" % (rows, d), timeit(lambda: K.bop("tensordot", X, b, False, False, 1)), 8 * rows * d)
    show("X^T@r (%d x %d)" % (rows, d), timeit(lambda: K.bop("tensordot", X, r, True, False, 1)), 8 * rows * d)
    show("sum axis=0 (%d x %d)" % (rows, d), timeit(lambda: K.reduce_axis("sum", X, 0, False, False)), 8 * rows * d)
for d in (64, 256):
    rows = n // d
    X = dv.DeviceBlock(x.t[: rows * d].view(rows, d))
    col = dv.DeviceBlock(x.t[:rows].view(rows, 1))
    rowv = dv.DeviceBlock(x.t[:d].view(1, d))
    show("(n,1)*(n,%d) broadcast mul" % d, timeit(lambda: K.bop("mul", col, X, False, False, {})), 16 * rows * d + 8 * rows)
    show("(n,%d)-(1,%d) broadcast sub" % (d, d), timeit(lambda: K.bop("sub", X, rowv, False, False, {})), 16 * rows * d)
xi = K.astype(dv.DeviceBlock(x.t[:n]), "int64")
ci = dv.from_numpy(np.int64(3), 0)
show("x == c  (int64 -> bool, 1e8)", timeit(lambda: K.bop("eq", xi, ci, False, False, {})), 9 * n)
show("x < y   (f64, f64 -> bool, 1e8)", timeit(lambda: K.bop("lt", x, x, False, False, {})), 17 * n)
show("argmax int64 1e8", timeit(lambda: K.arg_op("argmax", xi, slice(0, n))), 8 * n)
from nums_b200 import fused
for d in (8, 64, 256):
    rows = n // d
    X = dv.DeviceBlock(x.t[: rows * d].view(rows, d))
    bfull = dv.from_numpy(rng.standard_normal(d + 1) * 0.1, 0)
    show("glm_predict logistic (%d x %d)" % (rows, d), timeit(lambda: fused.glm_predict(X, bfull, "logistic")), 8 * rows * d + 8 * rows)
