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
"""
import argparse
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


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=ROWS_DEFAULT, help="rows per GPU (weak scaling)")
    ap.add_argument("--d", type=int, default=D_DEFAULT)
    ap.add_argument("--cpu-rows", type=int, default=400_000, help="rows of the bounded CPU-baseline sample")
    ap.add_argument("--cpu-steps", type=int, default=5)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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


# ------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the oracle port of the reference algorithm on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_newton_steps(rows, d, steps, warmup=1):
    """Times `steps` Newton iterations of the reference algorithm (oracle/glm.py: per-block NumPy calls
    incl. the materialised s*X) on a bounded sample.  Returns (GB/s, seconds per step, threads)."""
    from oracle import glm as oglm
    try:
        from threadpoolctl import threadpool_limits
    except Exception:  # pragma: no cover
        threadpool_limits = None
    import psutil
    cores = psutil.cpu_count(logical=False) or os.cpu_count() or 1
    cores = max(1, min(cores, len(os.sched_getaffinity(0))))
    X, y = host_chunk(0, rows, d)
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

    import contextlib
    with (threadpool_limits(limits=cores) if threadpool_limits else contextlib.nullcontext()):
        run(warmup)
        t0 = time.perf_counter()
        run(steps)
        dt = (time.perf_counter() - t0) / steps
    return algorithmic_bytes(rows, d) / dt / 1e9, dt, cores


def run_reference(args, result):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    gbs, dt, cores = cpu_newton_steps(args.cpu_rows, args.d, args.steps, max(1, min(args.warmup, 2)))
    sample = "%d x %d float64 rows (4 row blocks), %d Newton iterations, NumPy/OpenBLAS on %d threads" % (
        args.cpu_rows, args.d, args.steps, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "LogisticRegression newton, tall-skinny %dx%d float64 per GPU (BASELINE configs[3])"
                               % (args.rows, args.d),
                   "rows_per_gpu": args.rows, "d": args.d, "penalty": "l2", "C": 10},
        "cpu_baseline": {"value": gbs, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
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


# ------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------
def run_ours(args, result):
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
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    app = set_instance(make_app(device_ids=[local]))
    km = app.km
    lib = _lib.load()
    n, d, p = args.rows, args.d, args.d + 1
    K, W = args.steps, max(args.warmup, 3)

    # ---- synthetic shard, resident in HBM: reference-RNG chunk tiled over the shard
    chunk_rows = min(CHUNK_ROWS, n)
    Xc, yc = host_chunk(rank, chunk_rows, d)
    X = dv.empty((n, d), np.float64, local)
    y = dv.empty((n,), np.int64, local)
    Xc_d, yc_d = dv.from_numpy(Xc, local), dv.from_numpy(yc, local)
    K_ = __import__("nums_b200.cuda_kernel", fromlist=["KernelCls"]).KernelCls()
    for t, lo in enumerate(range(0, n, chunk_rows)):
        hi = min(n, lo + chunk_rows)
        scaled = K_.bop("mul", Xc_d, dv.from_numpy(np.float64(1.0 + 1e-3 * (t % 7)), local), False, False, {})
        dv.copy_into(X.t[lo:hi], scaled.t[: hi - lo])
        dv.copy_into(y.t[lo:hi], yc_d.t[: hi - lo])
    del Xc_d, yc_d
    Xb = BlockArray.from_oid(X, (n, d), np.float64, km)
    yb = BlockArray.from_oid(y, (n,), np.int64, km)
    model = LogisticRegression(app=app, solver="newton", penalty="l2", C=10, tol=-1.0, max_iter=K)
    model._setup_penalties(p, p)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

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

    # ---- the statistics pass alone (pass 1 + pass 2 + their finish kernels) through the C ABI, timed with
    # CUDA events on the stream it is launched on; pass 1 alone is timed below, pass 2 is the difference
    st = dv.stream_ptr(local)
    beta_d = dv.from_numpy(np.full(p, 0.01), local)
    stats_d = dv.empty((lib.nb_glm_stats_len(d),), np.float64, local)
    sc = dv.scratch(lib.nb_glm_scratch_bytes(local, n, d, 1), local)

    def gram():
        _lib.check(lib.nb_glm_stats(local, 1, X.ptr, n, d, d, y.ptr, y.code, beta_d.ptr, 3, stats_d.ptr,
                                    sc.data_ptr(), sc.numel(), st))

    for _ in range(3):
        gram()
    torch.cuda.synchronize()
    reps = max(3, min(K, 10))
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record()
    for _ in range(reps):
        gram()
    g1.record()
    torch.cuda.synchronize()
    gram_ms = g0.elapsed_time(g1) / reps
    # ---- the dominant kernel alone: pass 2 (k_tri_gram + its finish) through nb_matmul_f64, which routes the
    # symmetric X^T diag(w) X to the same kernel nb_glm_stats uses
    wbuf = dv.empty((n,), np.float64, local)
    _lib.check(lib.nb_fill(local, n, wbuf.ptr, wbuf.code, 0.25, 0, st))
    Gm = dv.empty((d, d), np.float64, local)
    sc2 = dv.scratch(lib.nb_matmul_scratch_bytes(local, d, d, n), local)

    def tri():
        _lib.check(lib.nb_matmul_f64(local, d, d, n, X.ptr, 1, d, X.ptr, d, 1, wbuf.ptr, Gm.ptr, d,
                                     sc2.data_ptr(), sc2.numel(), st))

    for _ in range(3):
        tri()
    torch.cuda.synchronize()
    t0e, t1e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0e.record()
    for _ in range(reps):
        tri()
    t1e.record()
    torch.cuda.synchronize()
    tri_ms = t0e.elapsed_time(t1e) / reps
    del wbuf, Gm, sc2
    # pass 1 alone (HBM-bound link / gradient pass = the L-BFGS evaluation)
    for _ in range(2):
        fused.glm_stats(X, y, beta_d, "logistic", False)
    torch.cuda.synchronize()
    r0, r1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    r0.record()
    for _ in range(reps):
        fused.glm_stats(X, y, beta_d, "logistic", False)
    r1.record()
    torch.cuda.synchronize()
    rows_ms = r0.elapsed_time(r1) / reps
    clocks = sampler.stop() if rank == 0 else None
    del stats_d, sc

    # ---- end to end through the public API with HOST buffers (upload + K iterations + read-back)
    e2e = None
    if not args.no_e2e:
        import psutil
        avail = psutil.virtual_memory().available
        per_rank_budget = avail // (3 * max(1, world))
        e_rows = int(min(n, max(chunk_rows, per_rank_budget // (8 * d + 8))))
        del Xb, yb
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
               "rows_per_gpu": e_rows, "seconds": dt,
               "what": "LogisticRegression.fit(host X, host y): upload from pinned host memory + %d Newton "
                       "iterations + coefficient read-back, per rank" % K}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    gram_bytes = n * (8 * d + 8)
    gram_flops_sym = n * d * (d + 1)
    fp64_peak = None
    try:
        fp64_peak = json.load(open(os.path.join(ROOT, "profiles", "fp64_peak.json"))).get("dmma_tflops")
    except Exception:
        pass
    traffic = None       # dram__bytes_read.sum + dram__bytes_write.sum of one k_tri_gram launch (ncu --set full)
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if int(tr.get("rows", -1)) == n and int(tr.get("d", -1)) == d:
            traffic = tr.get("k_tri_gram_dram_bytes")
    except Exception:
        pass
    executed_flops = 2.0 * n * 1024 * (((d + 31) // 32) * ((d + 31) // 32 + 1) // 2)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic (reference block-RNG chunk of %d rows tiled over each shard)" % chunk_rows,
        "config": {"workload": "LogisticRegression newton, tall-skinny %dx%d float64 per GPU (BASELINE configs[3]; "
                               "10^8x256 = 8 such shards)" % (n, d),
                   "rows_per_gpu": n, "d": d, "penalty": "l2", "C": 10, "l2_flush": L2_NOTE,
                   "step": "one Newton iteration: fused stats pass + all-reduce(p*p+p+1 f64) + Newton update + 8-byte read-back"},
        "fit_sec_per_iteration": ms_step * 1e-3,
        "fit_sec": fit_sec, "fit_iterations": fit_iters,
        "roofline": {"bound": "hbm",
                     "kernel": "k_tri_gram (pass 2: warp-specialised upper-triangular X^T diag(w) X, FP64 DMMA fed by TMA "
                               "bulk copies) + k_fused_finish, timed alone with CUDA events on its stream",
                     "achieved": gram_bytes / (tri_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                     "frac": gram_bytes / (tri_ms * 1e-3) / 1e9 / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                     "ms_per_launch": tri_ms, "algorithmic_bytes_per_launch": gram_bytes,
                     "note": "FP64-tensor-bound at d=256 (32 flop/B vs a balance of ~5.7): the DMMA ceiling is "
                             "n*d*(d+1)/37.2e12 s = 18% of the HBM peak; see fp64 (tensor roofline) and pass1 (HBM-bound pass)"},
        "fp64": {"kernel": "k_tri_gram", "ms_per_launch": tri_ms,
                 "achieved_tflops_sym": gram_flops_sym / (tri_ms * 1e-3) / 1e12,
                 "achieved_tflops_executed": executed_flops / (tri_ms * 1e-3) / 1e12,
                 "peak_tflops": fp64_peak,
                 "frac_executed": (executed_flops / (tri_ms * 1e-3) / 1e12 / fp64_peak) if fp64_peak else None},
        "stats_pass": {"what": "nb_glm_stats(NB_GLM_HESS_NO_OBJECTIVE), what a Newton step runs = pass 1 + pass 2 + finish kernels", "ms_per_launch": gram_ms,
                       "achieved": gram_bytes / (gram_ms * 1e-3) / 1e9, "unit": "GB/s"},
        "pass1": {"kernel": "k_glm_rows (eta/link/objective/gradient, HBM-bound)", "ms_per_launch": rows_ms,
                  "achieved": gram_bytes / (rows_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                  "frac": gram_bytes / (rows_ms * 1e-3) / 1e9 / hbm_peak},
        "clocks": clocks,
        # per Newton step: k_glm_rows, k_glm_rows_finish, k_tri_gram, k_fused_finish, k_newton_prep, k_chol_solve, and the
        # gated fallback pair k_inv / k_newton_apply (launched, exit at once when the Cholesky step succeeded)
        "gpu_launches": 8 * K,
        "coef_checksum": float(np.sum(np.abs(km.get(beta.blocks[(0,)].oid)))),
    }
    if e2e is not None:
        line["e2e"] = e2e
    if world == 1 and not args.no_cpu_baseline:
        gbs, dt, cores = cpu_newton_steps(args.cpu_rows, d, args.cpu_steps)
        line["cpu_baseline"] = {
            "value": gbs, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "%d x %d rows (4 row blocks), %d Newton iterations of the reference algorithm "
                      "(oracle/glm.py), NumPy/OpenBLAS on %d threads, %.2f s/iteration"
                      % (args.cpu_rows, d, args.cpu_steps, cores, dt)}
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
