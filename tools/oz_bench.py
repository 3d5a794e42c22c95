"""Int8-tensor-path Hessian (csrc/glm_ozaki.cu) against the FP64 DMMA path: accuracy on an oracle-sized block and the
time of the statistics pass at the benchmark shard.  Usage: python tools/oz_bench.py [rows] [d]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import device as dv, fused
from nums_b200.app import make_app, set_instance

set_instance(make_app(num_gpus=1))
n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 12_500_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 256
rng = np.random.default_rng(0)

# accuracy: fp64 reference in extended precision on the host is slow; compare both paths with a float128 Gram of a sample
m = 20000
Xs = rng.standard_normal((m, d)) + 0.2
ys = (rng.random(m) < 0.5).astype(np.int64)
bs = rng.standard_normal(d + 1) * 0.05
eta = Xs @ bs[:d] + bs[d]
mu = 1 / (1 + np.exp(-eta)); w = mu * (1 - mu)
Xl = Xs.astype(np.longdouble)
H_ref = np.asarray((Xl * w[:, None].astype(np.longdouble)).T @ Xl, dtype=np.float64)
for mode in ("int8/5", "int8/4", "fp64"):
    os.environ["NUMS_B200_INT8_PLANES"] = mode[5:] or "auto"
    mode = mode[:4]
    s = dv.to_numpy(fused.glm_stats(dv.from_numpy(Xs, 0), dv.from_numpy(ys, 0), dv.from_numpy(bs, 0), "logistic", True,
                                    hessian=mode))
    H = s[: (d + 1) ** 2].reshape(d + 1, d + 1)[:d, :d]
    print("accuracy %s (planes %s): max|dH|/max|H| = %.3e" % (mode, os.environ["NUMS_B200_INT8_PLANES"], np.max(np.abs(H - H_ref)) / np.max(np.abs(H_ref))), flush=True)

base = dv.from_numpy(rng.standard_normal(1 << 22), 0)
X = dv.empty((n * d,), np.float64, 0)
for lo in range(0, n * d, base.size):
    hi = min(n * d, lo + base.size)
    dv.copy_into(X.t[lo:hi], base.t[: hi - lo])
X = dv.DeviceBlock(X.t.view(n, d))
y = dv.from_numpy((rng.random(n) < 0.5).astype(np.int64), 0)
beta = dv.from_numpy(rng.standard_normal(d + 1) * 0.01, 0)


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def prep():
    X.prep = None
    fused.glm_prepare(X)


print("prepare (column maxima, once per block): %.3f ms" % timed(prep), flush=True)
for mode in ("int8/5", "int8/4", "int8/auto", "fp64"):
    os.environ["NUMS_B200_INT8_PLANES"] = mode[5:] or "auto"
    planes, mode = mode[5:], mode[:4]
    ms = timed(lambda: fused.glm_stats(X, y, beta, "logistic", True, need_obj=False, hessian=mode))
    print("stats pass %s %s: %.3f ms  %.0f GB/s algorithmic" % (mode, planes, ms, n * (8 * d + 8) / ms / 1e6), flush=True)
from nums_b200 import _lib
lib = _lib.load()
stats_d = dv.empty((lib.nb_glm_stats_len(d),), np.float64, 0)
prep_found = fused.glm_prepare(X)
for planes in ("5", "4"):
    os.environ["NUMS_B200_INT8_PLANES"] = planes
    flags = 3 | fused.int8_flags(prep_found, n)
    sc = dv.scratch(lib.nb_glm_scratch_bytes(0, n, d, flags), 0)
    for name, extra in (("rows", 8), ("products", 16)):
        ms = timed(lambda: _lib.check(lib.nb_glm_stats_prepared(0, 1, X.ptr, n, d, d, y.ptr, y.code, beta.ptr, flags | extra,
                                                                prep_found.buf.data_ptr(), stats_d.ptr, sc.data_ptr(), sc.numel(),
                                                                dv.stream_ptr(0))))
        print("  %s planes, %s only: %.3f ms" % (planes, name, ms), flush=True)
    del sc
# per-kernel times inside a back-to-back loop of whole passes (CUPTI activity records: the kernels at the clocks of the loop)
from torch.profiler import profile, ProfilerActivity
for planes in ("5", "4"):
    os.environ["NUMS_B200_INT8_PLANES"] = planes
    run = lambda: fused.glm_stats(X, y, beta, "logistic", True, need_obj=False, hessian="int8")
    for _ in range(3):
        run()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(10):
            run()
        torch.cuda.synchronize()
    for ev in sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:4]:
        print("  in-loop, %s planes: %-50s %8.3f ms x %d" % (planes, ev.key[:50], ev.device_time_total / ev.count / 1e3, ev.count), flush=True)
b = dv.to_numpy(fused.glm_stats(X, y, beta, "logistic", True, hessian="fp64"))
for planes in ("5", "4"):
    os.environ["NUMS_B200_INT8_PLANES"] = planes
    a = dv.to_numpy(fused.glm_stats(X, y, beta, "logistic", True, hessian="int8"))
    print("full size, %s planes: max|int8 - fp64| / max|fp64| = %.3e" % (planes, np.max(np.abs(a - b)) / np.max(np.abs(b))))
