#!/usr/bin/env python
"""Achieved bandwidth / throughput of every hot-path kernel family at BASELINE sizes,
through the kernel module (KernelCls -> C ABI), timed with CUDA events.  Prints a markdown table and a
JSON blob; `profiles/rNN_kernels.md` is this script's output.

    python tools/kernel_bench.py [--n 100000000] [--out profiles/r01_kernels.md]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=10 ** 8)
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    import torch
    from nums_b200 import cuda_kernel, device as dv
    from nums_b200.app import make_app, set_instance
    app = set_instance(make_app(num_gpus=1))
    K = cuda_kernel.KernelCls()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    fp64 = None
    try:
        fp64 = json.load(open(os.path.join(ROOT, "profiles", "fp64_peak.json")))["dmma_tflops"]
    except Exception:
        pass
    n = args.n
    rng = np.random.default_rng(0)
    base = rng.random(1 << 22)

    def big(count, dtype=np.float64):
        """count elements on the device, tiled from a 4M-element host draw (upload stays small)."""
        out = dv.empty((count,), dtype, 0)
        src = dv.from_numpy(base.astype(dtype) if dtype == np.float64 else (base * 1e6).astype(dtype), 0)
        step = src.size
        for lo in range(0, count, step):
            hi = min(count, lo + step)
            dv.copy_into(out.t[lo:hi], src.t[: hi - lo])
        return out

    def timeit(fn, reps=5, warm=2):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    rows = []

    def add(name, ref, ms, nbytes=None, flops=None):
        rec = {"kernel": name, "reference": ref, "ms": ms}
        if nbytes is not None:
            rec["GB/s"] = nbytes / (ms * 1e-3) / 1e9
            rec["frac_hbm"] = rec["GB/s"] / hbm
        if flops is not None:
            rec["TFLOP/s"] = flops / (ms * 1e-3) / 1e12
            if fp64:
                rec["frac_fp64"] = rec["TFLOP/s"] / fp64
        rows.append(rec)

    x = big(n)
    y = big(n)
    xi = big(n, np.int64)
    one = dv.from_numpy(np.float64(1.0), 0)
    c_i = dv.from_numpy(np.int64(n // 2), 0)
    add("bop add f64+f64 (n=%d)" % n, "numpy_kernel.py:426-431", timeit(lambda: K.bop("add", x, y, False, False, {})), 24 * n)
    add("bop truediv scalar/f64", "glms.py:457-458 (1/(1+e))", timeit(lambda: K.bop("truediv", one, x, False, False, {})), 16 * n)
    add("bop eq i64==scalar -> bool", "blockarray.py:997 (x == c)", timeit(lambda: K.bop("eq", xi, c_i, False, False, {})), 9 * n)
    add("bop floor_divide i64", "Appendix A", timeit(lambda: K.bop("floor_divide", xi, c_i, False, False, {})), 24 * n)
    add("map_uop exp f64", "numpy_kernel.py:316-318", timeit(lambda: K.map_uop("exp", x, (), {})), 16 * n)
    add("map_uop negative f64", "numpy_kernel.py:316-318", timeit(lambda: K.map_uop("negative", x, (), {})), 16 * n)
    rows_m, d = n // 64, 64
    mat = dv.DeviceBlock(x.t.view(rows_m, d))
    col = dv.DeviceBlock(y.t[:rows_m].view(rows_m, 1))
    add("bop mul (n,1)*(n,%d) broadcast (the s*X of glms.py:496)" % d, "glms.py:495-496",
        timeit(lambda: K.bop("mul", col, mat, False, False, {})), 16 * n + 8 * rows_m)
    add("reduce_axis sum f64 (all)", "numpy_kernel.py:294-313", timeit(lambda: K.reduce_axis("sum", x, None, False, False)), 8 * n)
    add("reduce_axis max f64 (all)", "numpy_kernel.py:294-313", timeit(lambda: K.reduce_axis("max", x, None, False, False)), 8 * n)
    add("reduce_axis sum i64 (all)", "numpy_kernel.py:294-313", timeit(lambda: K.reduce_axis("sum", xi, None, False, False)), 8 * n)
    add("reduce_axis sum axis=0 of (n/64, 64)", "numpy_kernel.py:298-312", timeit(lambda: K.reduce_axis("sum", mat, 0, False, False)), 8 * n)
    add("reduce_axis sum axis=1 of (n/64, 64)", "numpy_kernel.py:298-312", timeit(lambda: K.reduce_axis("sum", mat, 1, False, False)), 8 * n)
    add("arg_op argmax i64", "numpy_kernel.py:492-508", timeit(lambda: K.arg_op("argmax", xi, slice(0, n))), 8 * n)
    cond = K.bop("eq", xi, c_i, False, False, {})
    add("where (1-arg) on bool, ~0 hits", "numpy_kernel.py:320-330", timeit(lambda: K.where(cond, None, None, [(0, n)])), 2 * n)
    add("where (3-arg) f64 (cond mostly false: one operand is read)", "numpy_kernel.py:331-334", timeit(lambda: K.where(cond, x, y, None)), 17 * n)
    add("tensordot x.x (dot, n=%d)" % n, "numpy_kernel.py:420-424", timeit(lambda: K.bop("tensordot", x, x, False, False, 1)), 8 * n)
    del y, cond
    rows_g, dg = n // 256, 256
    Xg = dv.DeviceBlock(x.t[: rows_g * dg].view(rows_g, dg))
    bvec = dv.from_numpy(rng.standard_normal(dg), 0)
    rvec = dv.DeviceBlock(xi.t[:rows_g].view(-1))
    rvec = K.astype(rvec, "float64")
    add("tensordot X@b (%d x 256)" % rows_g, "blockarray.py:837-848", timeit(lambda: K.bop("tensordot", Xg, bvec, False, False, 1)), 8 * rows_g * dg)
    add("tensordot X^T@r (%d x 256)" % rows_g, "blockarray.py:837-848", timeit(lambda: K.bop("tensordot", Xg, rvec, True, False, 1)), 8 * rows_g * dg)
    add("tensordot X^T@X (%d x 256, symmetric split-k GEMM)" % rows_g, "glms.py:318", timeit(lambda: K.bop("tensordot", Xg, Xg, True, False, 1), reps=3),
        8 * rows_g * dg, flops=rows_g * dg * (dg + 1))
    del xi, rvec
    m = 10 ** 4
    A = dv.DeviceBlock(x.t[: m * m].view(m, m))
    Bm = big(m * m)
    Bm = dv.DeviceBlock(Bm.t.view(m, m))
    add("tensordot A@B 10^4 x 10^4 (FP64 DMMA)", "BASELINE config 3", timeit(lambda: K.bop("tensordot", A, Bm, False, False, 1), reps=2, warm=1),
        flops=2.0 * m ** 3)
    add("tensordot A^T@B 10^4 x 10^4", "BASELINE config 3", timeit(lambda: K.bop("tensordot", A, Bm, True, False, 1), reps=2, warm=1),
        flops=2.0 * m ** 3)
    # GLM passes (the L-BFGS / gradient evaluation is pass 1 alone; Newton adds pass 2)
    from nums_b200 import fused
    del A, Bm, Xg
    for dd in (64, 256, 1024):
        nr = (n * 4) // dd if dd >= 256 else n // dd
        Xd = big(nr * dd)
        Xd = dv.DeviceBlock(Xd.t.view(nr, dd))
        yd = dv.from_numpy((rng.random(1 << 20) < 0.5).astype(np.int64), 0)
        yfull = dv.empty((nr,), np.int64, 0)
        for lo in range(0, nr, yd.size):
            hi = min(nr, lo + yd.size)
            dv.copy_into(yfull.t[lo:hi], yd.t[: hi - lo])
        bd = dv.from_numpy(rng.standard_normal(dd + 1) * 0.01, 0)
        nbytes = nr * (8 * dd + 8)
        for fam in ("logistic", "linear", "poisson") if dd == 1024 else ("logistic",):
            add("glm_stats gradient+objective pass, %s, %d x %d (L-BFGS evaluation)" % (fam, nr, dd), "glms.py:457-489 / lbfgs.py:32-35",
                timeit(lambda: fused.glm_stats(Xd, yfull, bd, fam, False), reps=3), nbytes)
        add("glm_stats Newton pass (g + H), logistic, %d x %d" % (nr, dd), "glms.py:660-676",
            timeit(lambda: fused.glm_stats(Xd, yfull, bd, "logistic", True), reps=3), nbytes, flops=nr * dd * (dd + 1))
        del Xd, yfull
    S = rng.standard_normal((257, 257))
    S = dv.from_numpy(S.T @ S + np.eye(257), 0)
    add("inv 257 x 257", "numpy_kernel.py:471", timeit(lambda: K.inv(S), reps=5))

    # factorisations: one block of a tall-skinny QR, and the p x p ends
    nq = (n * 4) // 256
    Xq = big(nq * 256)
    Xq = dv.DeviceBlock(Xq.t.view(nq, 256))
    pert = dv.from_numpy(rng.standard_normal((256,)), 0)
    Xq = K.bop("add", Xq, dv.DeviceBlock(pert.t.view(1, 256)), False, False, {})     # tiled draws alone are rank deficient
    qr_flops = 2.0 * nq * 256 * 257 + 2.0 * nq * 256 * 256        # two Gram passes (symmetric) + one n x d x d product
    add("qr mode='r' (CholeskyQR2 + Householder signs), %d x 256" % nq, "numpy_kernel.py:455-461 / linalg.py:55-103",
        timeit(lambda: K.qr(Xq, mode="r"), reps=2, warm=1), 8 * nq * 256, flops=qr_flops)
    add("qr mode='reduced', %d x 256" % nq, "numpy_kernel.py:455-461 / linalg.py:130-208",
        timeit(lambda: K.qr(Xq, mode="reduced"), reps=2, warm=1), 8 * nq * 256, flops=qr_flops + 2.0 * nq * 256 * 256)
    del Xq
    Ssp = dv.from_numpy(np.asarray(dv.to_numpy(S))[:256, :256].copy(), 0)
    add("cholesky 256 x 256", "numpy_kernel.py:463-464", timeit(lambda: K.cholesky(Ssp), reps=5))
    add("svd 256 x 256 (one-sided Jacobi)", "numpy_kernel.py:466-469", timeit(lambda: K.svd(Ssp), reps=3))

    # block RNG in HBM: same values as the host generator, which is timed on 10^7 draws beside it
    import time as _t
    from numpy.random import PCG64, Generator
    for name, rargs, dt_ in (("random", (n,), np.float64), ("normal", (0.0, 1.0, n), np.float64),
                             ("integers", (0, 10 ** 6, n, np.int64, False), np.int64),
                             ("integers", (0, 3 * 2 ** 61, n, np.int64, False), np.int64)):
        hn = 10 ** 7
        hargs = tuple(hn if a == n else a for a in rargs)
        g = Generator(PCG64(1337).jumped(3))
        t0 = _t.perf_counter()
        getattr(g, name)(*hargs)
        host_ms = (_t.perf_counter() - t0) * 1e3 * (n / hn)
        label = "random_block %s%s n=%d (host NumPy: %.0f ms, extrapolated from 10^7)" % (
            name, "" if name != "integers" else " [0,%.3g)" % rargs[1], n, host_ms)
        add(label, "numpy_kernel.py:103-110", timeit(lambda: K.random_block((1337, 3), name, rargs, (n,), dt_, device_index=0), reps=3), 8 * n)

    # BASELINE config 1: the README fit (10^6 x 2, newton-cg, l2, C=10), wall clock incl. host control flow
    import time
    import warnings
    from nums_b200.glms import LogisticRegression
    half = 500_000
    r2 = np.random.default_rng(1337)
    X1 = np.concatenate([r2.standard_normal((half, 2)), r2.standard_normal((half, 2)) + 2.0])
    y1 = np.concatenate([np.zeros(half, dtype=np.int64), np.ones(half, dtype=np.int64)])
    Xb1, yb1 = app.array(X1, block_shape=(half, 2)), app.array(y1, block_shape=(half,))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for _ in range(2):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            m1 = LogisticRegression(app=app, solver="newton-cg", penalty="l2", C=10)
            m1.fit(Xb1, yb1)
            b1 = m1._beta.get()
            dt1 = time.perf_counter() - t0
    rows.append({"kernel": "README fit: LogisticRegression newton-cg 10^6 x 2, %d iterations, beta=%s (wall clock)"
                           % (m1.n_iter_, np.round(b1, 4).tolist()), "reference": "README.md:115-130", "ms": dt1 * 1e3})

    lines = ["| kernel | reference | ms | GB/s | of HBM peak (%.0f GB/s measured) | TFLOP/s | of FP64 DMMA peak |" % hbm,
             "|---|---|---:|---:|---:|---:|---:|"]
    for r in rows:
        lines.append("| %s | `%s` | %.3f | %s | %s | %s | %s |" % (
            r["kernel"], r["reference"], r["ms"],
            "%.0f" % r["GB/s"] if "GB/s" in r else "", "%.0f%%" % (100 * r["frac_hbm"]) if "frac_hbm" in r else "",
            "%.2f" % r["TFLOP/s"] if "TFLOP/s" in r else "", "%.0f%%" % (100 * r["frac_fp64"]) if "frac_fp64" in r else ""))
    text = "\n".join(lines)
    print(text)
    print(json.dumps(rows))
    if args.out:
        with open(args.out, "w") as fh:
            fh.write("# Per-kernel achieved bandwidth / throughput on B200 (tools/kernel_bench.py, CUDA events, n=%d)\n\n" % n)
            fh.write("Algorithmic bytes = inputs read once + outputs written once.\n\n")
            fh.write(text + "\n")


if __name__ == "__main__":
    main()
