#!/usr/bin/env python
"""bench.py — LogisticRegression Newton on tall-skinny float64 X, row-sharded one block per GPU.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[3], weak scaling): every GPU holds one row block of 1.25e7 x 256
float64 (25.6 GB) — at N=8 that is exactly the 10^8 x 256 configuration; 10^8 x 256 (204.8 GB) does
not fit one 180 GB GPU, so N=1 runs the per-GPU shard.  A "step" is one Newton iteration
(nums/models/glms.py:660-676): fused statistics pass over the shard, one all-reduce of the packed
p*p+p+1 doubles, the replicated Newton update, and the 8-byte convergence read-back.
metric = algorithmic HBM bytes of the pass, n*(8d+8) summed over ranks, per second.

Besides the headline the line carries, measured in the same run: the kernels of the step timed apart
(`roofline`, `tensor`, `fp64`), the other BASELINE configurations (`extras`: configs 2, 3, 5 at their own
sizes), at N > 1 a parity check of the sharded fit against the same rows on one rank (`parity_multi`) and
the real 10^8 x 256 array split N ways (`strong`), the end-to-end fit from host buffers (`e2e`) and the
reference's own CPU implementation on a bounded sample (`cpu_baseline`).
"""
import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "tall-skinny X^T.D.X HBM GB/s (LogReg Newton iteration, algorithmic bytes n*(8d+8))"
UNIT = "GB/s"
D_DEFAULT = 256
ROWS_DEFAULT = 12_500_000
CHUNK_ROWS = 1 << 19          # rows drawn from the reference's block RNG, then tiled over the shard
L2_NOTE = "inputs (25.6 GB/GPU) far exceed the 126 MB L2; no flush needed"
INT8_TOPS_NOMINAL = 4500.0    # dense int8 tensor peak of a B200 (no measured int8 figure in MEASURED_PEAKS.json)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=ROWS_DEFAULT, help="rows per GPU (weak scaling)")
    ap.add_argument("--d", type=int, default=D_DEFAULT)
    ap.add_argument("--cpu-rows", type=int, default=400_000, help="rows of the bounded CPU-baseline sample")
    ap.add_argument("--cpu-steps", type=int, default=4)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    ap.add_argument("--no-strong", action="store_true")
    ap.add_argument("--config", default="newton256", choices=["newton256", "lbfgs1024"],
                    help="newton256: the headline (BASELINE configs[3]); lbfgs1024: configs[4], Linear / Poisson lbfgs on a "
                         "1.25e7 x 1024 shard per GPU (102 GB resident), 10 iterations")
    return ap.parse_args()


def host_chunk(rank, rows, d):
    """Synthetic rows in the shape of BASELINE config 4: reference block RNG stream (block `rank`),
    second half shifted by delta = 2*sqrt(2/d)."""
    rng = np.random.Generator(np.random.PCG64(1337).jumped(rank))   # numpy_kernel.py:31-32
    X = rng.normal(0.0, 1.0, rows * d).reshape(rows, d)
    X[rows // 2:] += 2.0 * np.sqrt(2.0 / d)
    y = np.concatenate([np.zeros(rows // 2, dtype=np.int64), np.ones(rows - rows // 2, dtype=np.int64)])
    return X, y


def algorithmic_bytes(n, d):
    return n * (8 * d + 8)


def host_threads():
    import psutil
    cores = psutil.cpu_count(logical=False) or os.cpu_count() or 1
    return max(1, min(cores, len(os.sched_getaffinity(0))))


# ------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the reference's own code (oracle/_ref) on the host cores, else the oracle port
# ------------------------------------------------------------------------------------------------
def cpu_newton_steps(rows, d, steps, warmup=1):
    """Times `steps` Newton iterations of the reference on a bounded sample: the unmodified nums package
    (oracle/_ref: nums.models.glms.LogisticRegression on its SerialBackend, NumPy/OpenBLAS kernels) when it travelled
    with the repository, else the oracle port of the same algorithm.  Returns (GB/s, s/step, threads, kind)."""
    try:
        from threadpoolctl import threadpool_limits
    except Exception:  # pragma: no cover
        threadpool_limits = None
    cores = host_threads()
    X, y = host_chunk(0, rows, d)
    from oracle import refharness
    if refharness.present():
        import io
        import warnings
        warnings.simplefilter("ignore")
        app = refharness.use(refharness.serial_app(1))
        from nums.models.glms import LogisticRegression
        rb = max(1, rows // 4)
        Xb, yb = app.array(X, block_shape=(rb, d)), app.array(y, block_shape=(rb,))

        def run(k):
            m = LogisticRegression(solver="newton", penalty="l2", C=10, tol=-1.0, max_iter=k)
            with contextlib.redirect_stdout(io.StringIO()):
                m.fit(Xb, yb)
            return m._beta.get()
        kind = "reference"
    else:
        from oracle import glm as oglm
        model = oglm.BlockedGLM("logistic", oglm.split_rows(X, max(1, rows // 4)), oglm.split_rows(y, max(1, rows // 4)),
                                "l2", 0.1)

        def run(k):
            beta = np.zeros(d + 1)
            for _ in range(k):
                fwd = model.forward(beta)
                g = model.gradient(beta, fwd)
                H = model.hessian(beta, fwd)
                beta = beta - np.linalg.inv(H) @ g
                float(np.max(np.abs(g)))
            return beta
        kind = "port"
    with (threadpool_limits(limits=cores) if threadpool_limits else contextlib.nullcontext()):
        run(warmup)
        t0 = time.perf_counter()
        run(steps)
        dt = (time.perf_counter() - t0) / steps
    return algorithmic_bytes(rows, d) / dt / 1e9, dt, cores, kind


def cpu_sample_text(rows, d, steps, cores, kind, dt):
    what = ("the unmodified reference (oracle/_ref: nums.models.glms.LogisticRegression.fit on SerialBackend, NumPy/OpenBLAS)"
            if kind == "reference" else "the oracle port of the reference algorithm (oracle/glm.py, NumPy/OpenBLAS)")
    return "%d x %d float64 rows (4 row blocks), %d Newton iterations of %s on %d threads, %.2f s/iteration" % (
        rows, d, steps, what, cores, dt)


def run_reference(args, result):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    gbs, dt, cores, kind = cpu_newton_steps(args.cpu_rows, args.d, args.steps, max(1, min(args.warmup, 2)))
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "LogisticRegression newton, tall-skinny %dx%d float64 per GPU (BASELINE configs[3])"
                               % (args.rows, args.d),
                   "rows_per_gpu": args.rows, "d": args.d, "penalty": "l2", "C": 10},
        "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": cpu_sample_text(args.cpu_rows, args.d, args.steps, cores, kind, dt)},
        "e2e": {"value": gbs, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    result["line"] = line
    return 0


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self._stop = threading.Event()
        self._th = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.FIELDS,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                parts = [p.strip() for p in out.strip().split(",")]
                if len(parts) >= 6:
                    self.rows.append(parts)
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._th = threading.Thread(target=self._loop, daemon=True)
        self._th.start()

    def stop(self):
        self._stop.set()
        if self._th:
            self._th.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = []
        for name, col in (("hw_slowdown", 2), ("hw_thermal_slowdown", 3), ("sw_thermal_slowdown", 4),
                          ("sw_power_cap", 5)):
            if any(r[col].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.rows)}


def event_ms(torch, fn, reps, warm=2):
    """Mean device time of fn() over `reps` calls, CUDA events on the current stream (the one the library launches on)."""
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


def event_ms_alternating(torch, fns, reps, warm=2):
    """Mean device time of each fn when they run alternately (f0 f1 f0 f1 ...), every call between its own pair of CUDA
    events: the kernels of one step, at the clocks and power state of the step loop.  (A loop over ONE of them runs in
    another power regime: the board is at its power cap in steady state.)"""
    for _ in range(warm):
        for fn in fns:
            fn()
    torch.cuda.synchronize()
    marks = []
    for _ in range(reps):
        for i, fn in enumerate(fns):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            marks.append((i, e0, e1))
    torch.cuda.synchronize()
    out = [0.0] * len(fns)
    for i, e0, e1 in marks:
        out[i] += e0.elapsed_time(e1) / reps
    return out


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def tiled_shard(dv, K_, rank, n, d, local):
    """Resident synthetic shard: a chunk of the reference's block-RNG stream tiled over n rows (scaled per tile)."""
    chunk_rows = min(CHUNK_ROWS, n)
    Xc, yc = host_chunk(rank, chunk_rows, d)
    X = dv.empty((n, d), np.float64, local)
    y = dv.empty((n,), np.int64, local)
    Xc_d, yc_d = dv.from_numpy(Xc, local), dv.from_numpy(yc, local)
    for t, lo in enumerate(range(0, n, chunk_rows)):
        hi = min(n, lo + chunk_rows)
        scaled = K_.bop("mul", Xc_d, dv.from_numpy(np.float64(1.0 + 1e-3 * (t % 7)), local), False, False, {})
        dv.copy_into(X.t[lo:hi], scaled.t[: hi - lo])
        dv.copy_into(y.t[lo:hi], yc_d.t[: hi - lo])
    return X, y, Xc, yc


def parity_multi(app, dist, world, rank, local, d):
    """Sharded fit against the same rows on one rank: 100000*world x d rows drawn per rank from the block RNG, fitted
    (a) row-sharded with the NCCL all-reduce of the packed statistics, (b) gathered on every rank and fitted without
    the collective.  Returns max |beta_a - beta_b| and whether the iteration counts agree."""
    import torch
    from nums_b200 import device as dv
    from nums_b200.glms import LogisticRegression
    rows = 100_000
    Xs, ys = host_chunk(100 + rank, rows, d)
    Xd, yd = dv.from_numpy(Xs, local), dv.from_numpy(ys, local)
    from nums_b200.blockarray import BlockArray
    from nums_b200.grid import ArrayGrid
    km = app.km
    m = LogisticRegression(app=app, solver="newton", penalty="l2", C=10)
    m.fit(BlockArray.from_oid(Xd, (rows, d), np.float64, km), BlockArray.from_oid(yd, (rows,), np.int64, km))
    sharded = m._beta_full.get()
    it_sharded = m.n_iter_
    allX = [torch.empty_like(Xd.t) for _ in range(world)]
    ally = [torch.empty_like(yd.t) for _ in range(world)]
    dist.all_gather(allX, Xd.t)
    dist.all_gather(ally, yd.t)
    m2 = LogisticRegression(app=app, solver="newton", penalty="l2", C=10, sharded=False)
    Xall = BlockArray(ArrayGrid((rows * world, d), (rows, d), "float64"), km)
    yall = BlockArray(ArrayGrid((rows * world,), (rows,), "int64"), km)
    for r in range(world):
        Xall.blocks[(r, 0)].oid = dv.DeviceBlock(allX[r])
        yall.blocks[(r,)].oid = dv.DeviceBlock(ally[r])
    m2.fit(Xall, yall)
    single = m2._beta_full.get()
    err = float(np.max(np.abs(sharded - single)))
    t = torch.tensor([err], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return {"rows": rows * world, "d": d, "max_abs_dbeta": float(t.item()), "iters_sharded": it_sharded,
            "iters_single": m2.n_iter_, "iters_equal": it_sharded == m2.n_iter_,
            "what": "LogisticRegression newton (l2, C=10) to convergence: row-sharded over %d ranks with the NCCL "
                    "all-reduce vs the same %d rows as %d blocks on every rank without it" % (world, rows * world, world)}


def extras_single_gpu(torch, app, dv, fused, lib, _lib, local, hbm_peak, fp64_peak):
    """BASELINE configs 2, 3 and 5 at their own sizes on one GPU (device-timed, CUDA events)."""
    K_ = __import__("nums_b200.cuda_kernel", fromlist=["KernelCls"]).KernelCls()
    out = {}
    n = 10 ** 8
    # ---- config 2: x.T @ x, where(p == c), argmax(p), sum(p) on 10^8 elements
    x = K_.random_block((1337, 0), "random", (n,), (n,), np.float64, device_index=local)
    p = K_.permutation((1337, 1), n, device_index=local)
    ms = event_ms(torch, lambda: K_.bop("tensordot", x, x, True, False, 1), 10)
    out["dot_1e8"] = {"ms": ms, "gbs": 8 * n / ms / 1e6, "frac_hbm": 8 * n / ms / 1e6 / hbm_peak}
    c = dv.from_numpy(np.int64(n // 2), local)

    def where_eq():
        m = K_.bop("eq", p, c, False, False, {})
        return K_.where(m, None, None, [(0, n)])
    ms = event_ms(torch, where_eq, 10)
    out["where_eq_1e8"] = {"ms": ms, "gbs": (8 * n + 2 * n) / ms / 1e6, "frac_hbm": (8 * n + 2 * n) / ms / 1e6 / hbm_peak,
                           "what": "p == c (8n read, n written) then where (n read twice): 10n bytes"}
    ms = event_ms(torch, lambda: K_.arg_op("argmax", p, slice(0, n)), 10)
    out["argmax_1e8"] = {"ms": ms, "gbs": 8 * n / ms / 1e6, "frac_hbm": 8 * n / ms / 1e6 / hbm_peak}
    ms = event_ms(torch, lambda: K_.reduce_axis("sum", p, None, False, False), 10)
    out["sum_1e8"] = {"ms": ms, "gbs": 8 * n / ms / 1e6, "frac_hbm": 8 * n / ms / 1e6 / hbm_peak}
    del x, p
    torch.cuda.empty_cache()
    # ---- config 3: 10^4 x 10^4 @ 10^4 x 10^4 float64
    m = 10 ** 4
    A = K_.random_block((1337, 2), "random", (m * m,), (m, m), np.float64, device_index=local)
    B = K_.random_block((1337, 3), "random", (m * m,), (m, m), np.float64, device_index=local)
    for tag, tb in (("nn", False), ("nt", True)):
        ms = event_ms(torch, lambda: K_.bop("tensordot", A, B, False, tb, 1), 3, warm=1)
        tf = 2.0 * m ** 3 / ms / 1e9
        out["gemm_1e4_" + tag] = {"ms": ms, "tflops": tf, "frac_dmma": (tf / fp64_peak) if fp64_peak else None}
    del A, B
    torch.cuda.empty_cache()
    # ---- config 5: the L-BFGS evaluation (gradient + objective pass) on a 1.25e7 x 1024 shard (102.4 GB)
    n5, d5 = 12_500_000, 1024
    free, _ = torch.cuda.mem_get_info()
    if free > n5 * d5 * 8 + (8 << 30):
        base = dv.from_numpy(np.random.default_rng(0).standard_normal(1 << 22), local)
        X5 = dv.empty((n5 * d5,), np.float64, local)
        for lo in range(0, n5 * d5, base.size):
            hi = min(n5 * d5, lo + base.size)
            dv.copy_into(X5.t[lo:hi], base.t[: hi - lo])
        X5 = dv.DeviceBlock(X5.t.view(n5, d5))
        y5 = dv.from_numpy(np.random.default_rng(1).poisson(1.0, n5).astype(np.float64), local)
        b5 = dv.from_numpy(np.random.default_rng(2).standard_normal(d5 + 1) * 1e-3, local)
        for fam in ("linear", "poisson"):
            ms = event_ms(torch, lambda: fused.glm_stats(X5, y5, b5, fam, False), 3, warm=1)
            out["lbfgs_pass_1024_" + fam] = {"rows": n5, "d": d5, "ms": ms, "gbs": n5 * (8 * d5 + 8) / ms / 1e6,
                                            "frac_hbm": n5 * (8 * d5 + 8) / ms / 1e6 / hbm_peak}
        del X5, y5, b5
        torch.cuda.empty_cache()
    else:
        out["lbfgs_pass_1024"] = {"skipped": "not enough free HBM (%d GB)" % (free >> 30)}
    # ---- config 1: README fit (10^6 x 2) wall time
    from nums_b200.glms import LogisticRegression
    Xr = np.random.default_rng(3).standard_normal((10 ** 6, 2))
    Xr[500_000:] += 2.0
    yr = np.concatenate([np.zeros(500_000, dtype=np.int64), np.ones(500_000, dtype=np.int64)])
    Xb, yb = app.array(Xr, block_shape=(500_000, 2)), app.array(yr, block_shape=(500_000,))

    def readme():
        mm = LogisticRegression(app=app, solver="newton", penalty="l2", C=10)
        mm.fit(Xb, yb)
        return mm
    readme()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    mm = readme()
    torch.cuda.synchronize()
    out["readme_fit_ms"] = {"ms": (time.perf_counter() - t0) * 1e3, "iterations": mm.n_iter_}
    return out


def run_lbfgs1024(args, result):
    """BASELINE configs[4]: LinearRegression / PoissonRegression, solver lbfgs, 10^8 x 1024 float64 at 8 GPUs = a
    1.25e7 x 1024 shard per rank (weak scaling for fewer ranks).  Every evaluation is one fused gradient + objective pass
    and one all-reduce of p + 1 doubles (packed with the unused Hessian slot); the optimiser state lives on the device."""
    import torch
    import torch.distributed as dist
    from nums_b200 import device as dv, fused
    from nums_b200.app import make_app, set_instance
    from nums_b200.blockarray import BlockArray
    from nums_b200 import glms
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        # one process per GPU: keep this rank's host threads and the pages it touches first next to its GPU
        dv.bind_host_to_gpu(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    app = set_instance(make_app(device_ids=[local]))
    km = app.km
    n, d, iters = args.rows, 1024, 10
    rng = np.random.Generator(np.random.PCG64(1337).jumped(rank))
    chunk = 1 << 16
    Xc = rng.normal(0.0, 1.0, chunk * d).reshape(chunk, d) / np.sqrt(d)
    w_true = np.random.Generator(np.random.PCG64(7)).normal(0.0, 1.0, d)
    src = dv.from_numpy(Xc, local)
    X = dv.empty((n, d), np.float64, local)
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        dv.copy_into(X.t[lo:hi], src.t[: hi - lo])
    XB = BlockArray.from_oid(X, (n, d), np.float64, km)
    passes = [0]
    inner = km.backend._remote_functions["glm_stats"] if "glm_stats" in km.backend._remote_functions else None
    fused.register(km)
    inner = km.backend._remote_functions["glm_stats"]

    def counted(*a, **k):
        passes[0] += 1
        return inner(*a, **k)
    km.backend._remote_functions["glm_stats"] = counted
    out = {}
    for name, cls, yc in (("linear", glms.LinearRegression, Xc @ w_true + 0.1 * rng.normal(0.0, 1.0, chunk)),
                          ("poisson", glms.PoissonRegression, np.exp(Xc @ w_true))):      # tests/models/test_glms.py:245
        ysrc = dv.from_numpy(yc, local)
        y = dv.empty((n,), np.float64, local)
        for lo in range(0, n, chunk):
            hi = min(n, lo + chunk)
            dv.copy_into(y.t[lo:hi], ysrc.t[: hi - lo])
        yB = BlockArray.from_oid(y, (n,), np.float64, km)
        for rep in range(2):            # the first fit warms allocator and caches
            passes[0] = 0
            m = cls(app=app, solver="lbfgs", tol=0.0, max_iter=iters)
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            m.fit(XB, yB)
            coef = m._beta.get()
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        out[name] = {"iterations": m.n_iter_, "passes": passes[0], "seconds": dt, "ms_per_pass_incl_host": dt * 1e3 / passes[0],
                     "gbs": world * passes[0] * n * (8 * d + 8) / dt / 1e9, "coef_checksum": float(np.sum(np.abs(coef)))}
    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm = float(peaks.get("hbm_gbs", 6650.0))
        lin = out["linear"]
        result["line"] = {
            "metric": "L-BFGS evaluation HBM GB/s (LinearRegression lbfgs, algorithmic bytes n*(8d+8) per pass, host overhead included)",
            "value": lin["gbs"], "unit": "GB/s", "n_gpus": world, "steps": iters, "warmup": 1,
            "ms_per_step": lin["seconds"] * 1e3 / iters, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic (reference block-RNG chunk of %d rows tiled over each shard)" % chunk,
            "config": {"workload": "LinearRegression / PoissonRegression lbfgs, %dx1024 float64 per GPU (BASELINE configs[4]: "
                                   "10^8x1024 = 8 such shards)" % n, "rows_per_gpu": n, "d": d, "max_iter": iters},
            "frac_of_hbm_per_gpu": lin["gbs"] / world / hbm, "linear": lin, "poisson": out["poisson"], "gpu_launches": 2 * lin["passes"]}
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_ours(args, result):
    if args.config == "lbfgs1024":
        return run_lbfgs1024(args, result)
    import torch
    import torch.distributed as dist
    from nums_b200 import _lib, device as dv, fused
    from nums_b200.app import make_app, set_instance
    from nums_b200.blockarray import BlockArray
    from nums_b200.glms import LogisticRegression, newton

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun (one process per GPU)" % args.gpus)
    torch.cuda.set_device(local)
    bound_cpus = None
    if world > 1:
        # one process per GPU: keep this rank's host threads and the pages it touches first next to its GPU
        bound_cpus = dv.bind_host_to_gpu(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    app = set_instance(make_app(device_ids=[local]))
    km = app.km
    lib = _lib.load()
    n, d, p = args.rows, args.d, args.d + 1
    K, W = args.steps, max(args.warmup, 3)
    K_ = __import__("nums_b200.cuda_kernel", fromlist=["KernelCls"]).KernelCls()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- N > 1: the sharded fit must give the single-rank coefficients before anything is timed
    pm = None
    if world > 1:
        pm = parity_multi(app, dist, world, rank, local, d)
        if pm["max_abs_dbeta"] > 1e-8 or not pm["iters_equal"]:
            if rank == 0:
                sys.stderr.write("parity_multi FAILED: %s\n" % json.dumps(pm))
            dist.destroy_process_group()
            return 4

    # ---- synthetic shard, resident in HBM: reference-RNG chunk tiled over the shard
    X, y, Xc, yc = tiled_shard(dv, K_, rank, n, d, local)
    chunk_rows = Xc.shape[0]
    Xb = BlockArray.from_oid(X, (n, d), np.float64, km)
    yb = BlockArray.from_oid(y, (n,), np.int64, km)
    hessian_mode = fused.default_hessian()
    prep = fused.glm_prepare(X) if fused.int8_hessian_applies(X, "logistic") else None
    int8_used = bool(prep is not None and prep.usable and hessian_mode == "int8")
    model = LogisticRegression(app=app, solver="newton", penalty="l2", C=10, tol=-1.0, max_iter=K)
    model._setup_penalties(p, p)

    def newton_steps(k):
        beta = app.zeros((p,), (p,), np.float64)
        return newton(model, beta, Xb, yb, -1.0, k)

    newton_steps(W)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record()
    beta = newton_steps(K)
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms_total], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / K
    value = world * algorithmic_bytes(n, d) / (ms_step * 1e-3) / 1e9

    # ---- a real fit to convergence on the resident shard(s): the "LogReg Newton fit sec" half of the metric
    fit_model = LogisticRegression(app=app, solver="newton", penalty="l2", C=10)      # default tol 1e-4, max_iter 100
    fit_model._setup_penalties(p, p)
    barrier()
    t_fit = time.perf_counter()
    newton(fit_model, app.zeros((p,), (p,), np.float64), Xb, yb, fit_model._tol, fit_model._max_iter)
    torch.cuda.synchronize()
    fit_sec = time.perf_counter() - t_fit
    fit_iters = fit_model.n_iter_

    # ---- the kernels of one step, timed apart through the C ABI with CUDA events on the launching stream
    st = dv.stream_ptr(local)
    beta_d = dv.from_numpy(np.full(p, 0.01), local)
    stats_d = dv.empty((lib.nb_glm_stats_len(d),), np.float64, local)
    reps = max(3, min(K, 10))
    parts = {}

    def stats_call(mode, prep_ptr, scratch):
        _lib.check(lib.nb_glm_stats_prepared(local, 1, X.ptr, n, d, d, y.ptr, y.code, beta_d.ptr, mode, prep_ptr,
                                             stats_d.ptr, scratch.data_ptr(), scratch.numel(), st))

    if int8_used:
        INT8, ROWS, PROD = fused.int8_flags(prep, n), 8, 16          # what the step runs (level 5 left out at this n)
        sc8 = dv.scratch(lib.nb_glm_scratch_bytes(local, n, d, 3 | INT8), local)
        pp = prep.buf.data_ptr()
        parts["stats_int8_ms"] = event_ms(torch, lambda: stats_call(3 | INT8, pp, sc8), reps)
        parts["rows_ms"], parts["products_ms"] = event_ms_alternating(
            torch, [lambda: stats_call(3 | INT8 | ROWS, pp, sc8), lambda: stats_call(3 | INT8 | PROD, pp, sc8)], reps)
        del sc8
        torch.cuda.empty_cache()
    sc64 = dv.scratch(lib.nb_glm_scratch_bytes(local, n, d, 3), local)
    parts["stats_fp64_ms"] = event_ms(torch, lambda: stats_call(3, None, sc64), 3, warm=1)
    del sc64
    # the FP64 Hessian kernel alone (k_tri_gram + its finish) through nb_matmul_f64, which routes the symmetric
    # X^T diag(w) X to the kernel nb_glm_stats uses
    wbuf = dv.empty((n,), np.float64, local)
    _lib.check(lib.nb_fill(local, n, wbuf.ptr, wbuf.code, 0.25, 0, st))
    Gm = dv.empty((d, d), np.float64, local)
    sc2 = dv.scratch(lib.nb_matmul_scratch_bytes(local, d, d, n), local)
    tri_ms = event_ms(torch, lambda: _lib.check(lib.nb_matmul_f64(local, d, d, n, X.ptr, 1, d, X.ptr, d, 1, wbuf.ptr, Gm.ptr,
                                                                   d, sc2.data_ptr(), sc2.numel(), st)), 3, warm=1)
    del wbuf, Gm, sc2
    # the gradient / objective pass (what an L-BFGS evaluation runs)
    grad_ms = event_ms(torch, lambda: fused.glm_stats(X, y, beta_d, "logistic", False), reps)
    prep_ms = None
    if prep is not None:
        def redo_prep():
            X.prep = None
            fused.glm_prepare(X)
        prep_ms = event_ms(torch, redo_prep, 3, warm=1)
    clocks = sampler.stop() if rank == 0 else None
    del stats_d

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    fp64_peak = None
    try:
        fp64_peak = json.load(open(os.path.join(ROOT, "profiles", "fp64_peak.json"))).get("dmma_tflops")
    except Exception:
        pass

    # ---- N >= 2: the real 10^8 x 256 array split N ways (strong scaling of BASELINE configs[3])
    strong = None
    if world > 1 and not args.no_strong:
        total_rows = 10 ** 8
        rows_s = total_rows // world
        if rows_s == n:
            strong = {"rows_total": total_rows, "rows_per_gpu": rows_s, "ms_per_step": ms_step, "fit_sec": fit_sec,
                      "fit_iterations": fit_iters, "note": "at this rank count the weak-scaling shard IS 10^8 x 256 / N"}
        else:
            del Xb, yb
            X = y = None
            torch.cuda.empty_cache()
            free, _ = torch.cuda.mem_get_info()
            # X + one bounded buffer of digit planes (<= 12.52e6 rows) + 2 MiB of int32 partials per 32768 rows + slack
            need = rows_s * d * 8 + min(rows_s, 12_517_376) * 256 * 5 + (rows_s // 32768 + 1) * (2 << 20) + (8 << 30)
            mode_s = "int8" if free > need else "fp64"
            os.environ["NUMS_B200_HESSIAN"] = mode_s
            X, y, _, _ = tiled_shard(dv, K_, rank, rows_s, d, local)
            Xb = BlockArray.from_oid(X, (rows_s, d), np.float64, km)
            yb = BlockArray.from_oid(y, (rows_s,), np.int64, km)
            ms_model = LogisticRegression(app=app, solver="newton", penalty="l2", C=10, tol=-1.0, max_iter=5)
            ms_model._setup_penalties(p, p)
            newton(ms_model, app.zeros((p,), (p,), np.float64), Xb, yb, -1.0, 2)
            barrier()
            s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s0.record()
            newton(ms_model, app.zeros((p,), (p,), np.float64), Xb, yb, -1.0, 5)
            s1.record()
            barrier()
            tt = torch.tensor([s0.elapsed_time(s1) / 5], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            fm = LogisticRegression(app=app, solver="newton", penalty="l2", C=10)
            fm._setup_penalties(p, p)
            barrier()
            t0 = time.perf_counter()
            newton(fm, app.zeros((p,), (p,), np.float64), Xb, yb, fm._tol, fm._max_iter)
            torch.cuda.synchronize()
            strong = {"rows_total": total_rows, "rows_per_gpu": rows_s, "hessian": mode_s, "ms_per_step": float(tt.item()),
                      "gbs": total_rows * (8 * d + 8) / float(tt.item()) / 1e6, "fit_sec": time.perf_counter() - t0,
                      "fit_iterations": fm.n_iter_}
            os.environ["NUMS_B200_HESSIAN"] = hessian_mode

    # ---- end to end through the public API with HOST buffers (upload + K iterations + read-back)
    e2e = None
    if not args.no_e2e:
        import psutil
        avail = psutil.virtual_memory().available
        per_rank_budget = avail // (3 * max(1, world))
        e_rows = int(min(n, max(chunk_rows, per_rank_budget // (8 * d + 8))))
        Xb = yb = None
        X = y = None
        torch.cuda.empty_cache()
        try:
            Xh = torch.empty((e_rows, d), dtype=torch.float64, pin_memory=True).numpy()
            yh = torch.empty((e_rows,), dtype=torch.int64, pin_memory=True).numpy()
        except RuntimeError:
            e_rows = chunk_rows
            Xh = torch.empty((e_rows, d), dtype=torch.float64, pin_memory=True).numpy()
            yh = torch.empty((e_rows,), dtype=torch.int64, pin_memory=True).numpy()
        for lo in range(0, e_rows, chunk_rows):
            hi = min(e_rows, lo + chunk_rows)
            Xh[lo:hi] = Xc[: hi - lo]
            yh[lo:hi] = yc[: hi - lo]

        def fit_from_host():
            m = LogisticRegression(app=app, solver="newton", penalty="l2", C=10, tol=-1.0, max_iter=K)
            m.fit(app.array(Xh, block_shape=(e_rows, d)), app.array(yh, block_shape=(e_rows,)))
            return m._beta_full.get()

        fit_from_host()                      # warm-up (allocator, pinned path)
        # the upload alone, to attribute the end-to-end time (PCIe Gen5 x16: ~55 GB/s per GPU when the host can feed it)
        barrier()
        t0 = time.perf_counter()
        up = app.array(Xh, block_shape=(e_rows, d))
        torch.cuda.synchronize()
        upload_s = time.perf_counter() - t0
        del up
        if world > 1:
            tu = torch.tensor([upload_s], device="cuda", dtype=torch.float64)
            dist.all_reduce(tu, op=dist.ReduceOp.MAX)
            upload_s = float(tu.item())
        barrier()
        t0 = time.perf_counter()
        fit_from_host()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], device="cuda", dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": world * K * algorithmic_bytes(e_rows, d) / dt / 1e9, "unit": UNIT,
               "h2d_bytes_per_step": (e_rows * (8 * d + 8)) // K, "d2h_bytes_per_step": 8 + 4 + (8 * p) // K,
               "rows_per_gpu": e_rows, "seconds": dt, "upload_seconds": upload_s,
               "h2d_gbs_per_gpu": e_rows * d * 8 / upload_s / 1e9, "h2d_gbs_aggregate": world * e_rows * d * 8 / upload_s / 1e9,
               "compute_seconds_expected": K * ms_step * 1e-3,
               "host_cpus_bound_rank0": len(bound_cpus) if bound_cpus else None,
               "what": "LogisticRegression.fit(host X, host y): upload from pinned host memory + %d Newton "
                       "iterations + coefficient read-back, per rank" % K}
        del Xh, yh

    extras = None
    if world == 1 and not args.no_extras:
        Xb = yb = None
        X = y = None
        torch.cuda.empty_cache()
        extras = extras_single_gpu(torch, app, dv, fused, lib, _lib, local, hbm_peak, fp64_peak)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    gram_bytes = algorithmic_bytes(n, d)
    traffic = {}
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if int(tr.get("rows", -1)) == n and int(tr.get("d", -1)) == d:
            traffic = tr
    except Exception:
        pass
    nblk = (d + 31) // 32
    # tile products k_tri_gram executes per 4-row step: 16 per off-diagonal 32x32 block, 10 per diagonal block
    # (its lower 8x8 tiles are skipped), 2 * 8 * 8 * 4 flops each
    executed_flops = n * 128.0 * (16 * (nblk * (nblk - 1) // 2) + 10 * nblk)
    useful_flops = float(n) * d * (d + 1)
    if int8_used:
        dominant = "rows" if parts["rows_ms"] >= parts["products_ms"] else "products"
        roof_ms = parts["rows_ms"]
        i8flags = fused.int8_flags(prep, n)
        nplanes = 4 if (i8flags & fused.NB_GLM_INT8_FOUR_PLANES) else 5
        moved = n * d * (8 + nplanes) + 8 * n
        roofline = {
            "bound": "hbm",
            "kernel": "k_oz_rows (pass A of the int8 Hessian: link, gradient, u and the %d int8 digit planes; reads X "
                      "once, writes %d bytes per element) + k_glm_rows_finish, timed alone (NB_GLM_INT8_ROWS_ONLY) with CUDA "
                      "events on its stream, alternating with pass B as in a step" % (nplanes, nplanes),
            "achieved": gram_bytes / (roof_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
            "frac": gram_bytes / (roof_ms * 1e-3) / 1e9 / hbm_peak, "traffic": traffic.get("k_oz_rows_dram_bytes"),
            "peak_source": peak_src, "ms_per_launch": roof_ms, "algorithmic_bytes_per_launch": gram_bytes,
            "moved_bytes_per_launch": moved,
            "frac_of_moved_bytes": moved / (roof_ms * 1e-3) / 1e9 / hbm_peak,
            "dominant_of_step": dominant,
            "note": "algorithmic bytes = n*(8d+8) (X and y once); the kernel also writes the digit planes (%d B per element), "
                    "so it moves %d/8 of that: frac_of_moved_bytes is its HBM efficiency" % (nplanes, 8 + nplanes)}
        nprod = 8 if nplanes == 4 else (9 if (i8flags & fused.NB_GLM_INT8_NO_LEVEL5) else 11)
        int8_ops = nprod * 2.0 * 256 * 256 * n
        tensor = {
            "bound": "tensor", "kernel": "k_oz_gemm (pass B: %d exact int8 plane products," % nprod + " 72 CTA pairs, tcgen05.mma "
                                         "kind::i8 cta_group::2, TMA tensor copies, TMEM accumulators) + k_oz_reduce + k_oz_finish, "
                                         "timed alone (NB_GLM_INT8_PRODUCTS_ONLY), alternating with pass A as in a step",
            "note": "ncu: the tensor pipe is busy 94-96 % of the kernel's active cycles (profiles/r02_ncu_oz_full.csv); the board "
                    "runs this kernel at ~1.5 GHz of its 1.965 GHz, which is the distance to the nominal rate",
            "ms_per_launch": parts["products_ms"], "achieved": int8_ops / (parts["products_ms"] * 1e-3) / 1e12,
            "peak": INT8_TOPS_NOMINAL, "unit": "TOP/s", "frac": int8_ops / (parts["products_ms"] * 1e-3) / 1e12 / INT8_TOPS_NOMINAL,
            "peak_source": "nominal dense int8 (4.5 POP/s); MEASURED_PEAKS.json has no int8 figure",
            "traffic": traffic.get("k_oz_gemm_dram_bytes"),
            "equivalent_fp64_tflops": useful_flops / (parts["products_ms"] * 1e-3) / 1e12}
    else:
        roofline = {
            "bound": "fp64", "kernel": "k_tri_gram (warp-specialised upper-triangular X^T diag(w) X, FP64 DMMA fed by TMA bulk "
                                       "copies) + k_fused_finish, timed alone with CUDA events on its stream",
            "achieved": useful_flops / (tri_ms * 1e-3) / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
            "frac": (useful_flops / (tri_ms * 1e-3) / 1e12 / fp64_peak) if fp64_peak else None,
            "peak_source": "builder-measured DMMA peak (profiles/fp64_peak.json, tools/fp64_peak.cu)",
            "traffic": traffic.get("k_tri_gram_dram_bytes"), "ms_per_launch": tri_ms,
            "hbm_frac": gram_bytes / (tri_ms * 1e-3) / 1e9 / hbm_peak}
        tensor = None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic (reference block-RNG chunk of %d rows tiled over each shard)" % chunk_rows,
        "config": {"workload": "LogisticRegression newton, tall-skinny %dx%d float64 per GPU (BASELINE configs[3]; "
                               "10^8x256 = 8 such shards)" % (n, d),
                   "rows_per_gpu": n, "d": d, "penalty": "l2", "C": 10, "l2_flush": L2_NOTE,
                   "hessian": "int8 tensor path (exact digit split of x*sqrt(w), csrc/glm_ozaki.cu)" if int8_used
                              else "fp64 DMMA (csrc/glm_fused.cu)",
                   "step": "one Newton iteration: fused stats pass + all-reduce(p*p+p+1 f64) + Newton update + 8-byte read-back"},
        "fit_sec_per_iteration": ms_step * 1e-3,
        "fit_sec": fit_sec, "fit_iterations": fit_iters,
        "step_frac_of_hbm": value / world / hbm_peak,
        "roofline": roofline,
        "fp64": {"kernel": "k_tri_gram (the FP64 DMMA Hessian, used for heavy-tailed blocks and d outside 129..256)",
                 "ms_per_launch": tri_ms, "achieved_tflops_useful": useful_flops / (tri_ms * 1e-3) / 1e12,
                 "achieved_tflops_executed": executed_flops / (tri_ms * 1e-3) / 1e12,
                 "peak_tflops": fp64_peak, "peak_source": "builder-measured (profiles/fp64_peak.json)",
                 "frac_useful": (useful_flops / (tri_ms * 1e-3) / 1e12 / fp64_peak) if fp64_peak else None,
                 "frac_executed": (executed_flops / (tri_ms * 1e-3) / 1e12 / fp64_peak) if fp64_peak else None,
                 "stats_pass_ms": parts["stats_fp64_ms"]},
        "stats_pass": {"what": "nb_glm_stats_prepared(NB_GLM_HESS_NO_OBJECTIVE%s): everything a Newton step runs before the update; "
                               "timed in its own loop after the step loop -- the board is at its power cap and the clock it grants "
                               "differs from loop to loop by several per cent, so this can come out above ms_per_step"
                               % (" | NB_GLM_INT8_HESSIAN" if int8_used else ""),
                       "ms_per_launch": parts.get("stats_int8_ms", parts["stats_fp64_ms"]),
                       "achieved": gram_bytes / (parts.get("stats_int8_ms", parts["stats_fp64_ms"]) * 1e-3) / 1e9, "unit": "GB/s"},
        "gradient_pass": {"kernel": "k_oz_rows<0> (eta/link/objective/gradient through the bulk-copy ring: the L-BFGS evaluation at 192 < d <= 256, HBM-bound)",
                          "ms_per_launch": grad_ms, "achieved": gram_bytes / (grad_ms * 1e-3) / 1e9, "peak": hbm_peak,
                          "unit": "GB/s", "frac": gram_bytes / (grad_ms * 1e-3) / 1e9 / hbm_peak},
        "prepare_ms": prep_ms,
        "clocks": clocks,
        # per Newton step (int8 Hessian): memset, k_oz_rows, TMA-fed k_oz_gemm, k_oz_reduce, k_oz_finish, k_glm_rows_finish,
        # k_newton_prep, k_chol_solve and the gated fallback pair k_inv / k_newton_apply (exit at once after a good solve);
        # fp64 Hessian: k_glm_rows, k_glm_rows_finish, k_tri_gram, k_fused_finish + the same four
        "gpu_launches": (9 if int8_used else 8) * K,
        "coef_checksum": float(np.sum(np.abs(km.get(beta.blocks[(0,)].oid)))),
    }
    if tensor is not None:
        line["tensor"] = tensor
    if pm is not None:
        line["parity_multi"] = pm
    if strong is not None:
        line["strong"] = strong
    if e2e is not None:
        line["e2e"] = e2e
    if extras is not None:
        line["extras"] = extras
    if world == 1 and not args.no_cpu_baseline:
        gbs, dt, cores, kind = cpu_newton_steps(args.cpu_rows, d, args.cpu_steps)
        line["cpu_baseline"] = {"value": gbs, "unit": UNIT, "cores": cores, "kind": kind,
                                "sample": cpu_sample_text(args.cpu_rows, d, args.cpu_steps, cores, kind, dt)}
    result["line"] = line
    if world > 1:
        dist.destroy_process_group()
    return 0


class _StdoutToStderr:
    """Libraries (NCCL's version banner, warnings) may write to fd 1; the contract is ONE JSON line on stdout.
    Everything is routed to stderr while the bench runs and the real stdout is restored for the final print."""

    def __enter__(self):
        sys.stdout.flush()
        self._saved = os.dup(1)
        os.dup2(2, 1)
        return self

    def __exit__(self, *exc):
        sys.stdout.flush()
        os.dup2(self._saved, 1)
        os.close(self._saved)
        return False


def main():
    args = parse()
    result = {}
    with _StdoutToStderr():
        rc = run_reference(args, result) if args.impl == "reference" else run_ours(args, result)
    if "line" in result:
        print(json.dumps(result["line"]))
        sys.stdout.flush()
    return rc


if __name__ == "__main__":
    sys.exit(main())
