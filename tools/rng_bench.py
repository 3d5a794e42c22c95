"""Times the on-device block RNG (uniform / normal / integers) with CUDA events; prints one line per case."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import cuda_kernel
from nums_b200.app import make_app, set_instance

app = set_instance(make_app(num_gpus=1))
K = cuda_kernel.KernelCls()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 10 ** 8
for name, args, dt in (("random", (n,), np.float64), ("normal", (0.0, 1.0, n), np.float64),
                       ("integers", (0, 10 ** 6, n, np.int64, False), np.int64),
                       ("integers", (0, 3 * 2 ** 61, n, np.int64, False), np.int64)):
    for _ in range(2):
        K.random_block((1337, 3), name, args, (n,), dt, device_index=0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        K.random_block((1337, 3), name, args, (n,), dt, device_index=0)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("%-10s %-28s n=%d  %.3f ms  %.1f GB/s written" % (name, str(args[:2]), n, ms, 8 * n / ms / 1e6))

import time
from numpy.random import PCG64 as _P, Generator as _G
for pn in (10 ** 6, 10 ** 7, n):
    for _ in range(2):
        K.permutation((1337, 3), pn, device_index=0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        out = K.permutation((1337, 3), pn, device_index=0)
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) / 3 * 1e3
    t0 = time.perf_counter()
    ref = _G(_P(1337).jumped(3)).permutation(pn)
    host_ms = (time.perf_counter() - t0) * 1e3
    from nums_b200 import device as _dv
    same = np.array_equal(_dv.to_numpy(out), ref)
    print("permutation n=%d  device %.2f ms (wall, incl. convergence syncs)  host NumPy %.0f ms  identical=%s" % (pn, ms, host_ms, same))
