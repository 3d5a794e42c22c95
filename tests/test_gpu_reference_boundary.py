"""The reference ITSELF on this repository's plug-in.

oracle/_ref (made by oracle/make_ref.py, shipped to the GPU box with the snapshot) is the unmodified
`nums` package.  Each test runs an unchanged nums.numpy / nums.models.glms program twice in this
process: on NumS's ArrayApplication wired to CudaBackend + cuda_kernel + CyclicDeviceGrid((G, 1), "gpu")
exactly as the reference's tests/conftest.py:145-158 wires its own backends, and on the reference's
SerialBackend + numpy_kernel presenting the same number of devices (same block shapes, same summation
order).  Bars: indices / integer results bit-exact, float64 contractions rel-err <= 1e-10,
coefficients <= 1e-8 with equal iteration counts.

Skipped (not passed) when oracle/_ref is absent.
"""
import warnings

import numpy as np
import pytest

from conftest import rel_err
from oracle import refharness as rh

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not rh.present(), reason="oracle/_ref absent (python oracle/make_ref.py)")]


@pytest.fixture(scope="module")
def apps():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    rh.bootstrap()
    warnings.simplefilter("ignore")
    cuda = rh.cuda_app()
    G = cuda.km.backend.num_gpus
    serial = rh.serial_app(G)
    return {"cuda": cuda, "serial": serial, "G": G}


def both(apps, program):
    """Runs `program(nps, app)` under each application; returns {"cuda": result, "serial": result}."""
    import nums.numpy as nps
    out = {}
    for side in ("serial", "cuda"):
        app = rh.use(apps[side])
        nps.random.seed(1337)
        out[side] = program(nps, app)
    return out


class CountedNewton:
    """Counts the Newton iterations of a fit by counting `inv` calls (one per iteration, glms.py:673)."""

    def __init__(self, app):
        self.km, self.n = app.km, 0

    def __enter__(self):
        inner = self.km.methods["inv"]

        def counted(*a, **k):
            self.n += 1
            return inner(*a, **k)
        self._inner = inner
        self.km.methods["inv"] = counted
        return self

    def __exit__(self, *exc):
        self.km.methods["inv"] = self._inner


def readme_data(nps, size, d=2, shift=2.0):
    X1 = nps.random.randn(size // 2, d)
    y1 = nps.zeros(shape=(size // 2,), dtype=int)
    X2 = nps.random.randn(size // 2, d) + shift
    y2 = nps.ones(shape=(size // 2,), dtype=int)
    return nps.concatenate([X1, X2], axis=0), nps.concatenate([y1, y2], axis=0)


def test_cuda_kernel_registers_in_reference_kernel_manager(apps):
    km = apps["cuda"].km
    from nums.core.kernel.kernel_interface import Kernel
    import inspect
    names = [n for n, _ in inspect.getmembers(Kernel(), predicate=inspect.ismethod)]
    assert sorted(km.methods) == sorted(names)


def test_readme_logistic_regression_unchanged_program(apps):
    """README.md:115-130 at its own size (10^6 x 2), through the generic block kernels."""
    from nums.models.glms import LogisticRegression
    from nums_b200 import nums_plugin

    def program(nps, app):
        nums_plugin.disable_fused_glm()
        X, y = readme_data(nps, 10 ** 6)
        model = LogisticRegression(solver="newton-cg", penalty="l2", C=10)
        with CountedNewton(app) as c:
            model.fit(X, y)
        y_pred = model.predict(X)
        acc = (nps.sum(y == y_pred) / X.shape[0]).get()
        return dict(beta=model._beta.get(), beta0=model._beta0.get(), iters=c.n, acc=acc, pred=y_pred.get(),
                    X=X.get(), y=y.get())

    r = both(apps, program)
    assert np.array_equal(r["cuda"]["X"], r["serial"]["X"])          # block RNG stream is bit-identical
    assert np.array_equal(r["cuda"]["y"], r["serial"]["y"])
    assert r["cuda"]["iters"] == r["serial"]["iters"] == 8
    assert np.max(np.abs(r["cuda"]["beta"] - r["serial"]["beta"])) <= 1e-8
    assert abs(r["cuda"]["beta0"] - r["serial"]["beta0"]) <= 1e-8
    assert np.array_equal(r["cuda"]["pred"], r["serial"]["pred"])
    assert r["cuda"]["acc"] == r["serial"]["acc"]


def test_config2_dot_where_argmax_sum(apps):
    n = 10 ** 7

    def program(nps, app):
        x = nps.random.rand(n)
        dot = (x.T @ x).get()
        p = nps.random.permutation(n)
        idx = nps.where(p == n // 2)
        return dict(dot=dot, where=idx[0].get(), argmax=nps.argmax(p).get(), argmin=nps.argmin(x).get(),
                    sum=nps.sum(p).get(), max=nps.max(x).get(), head=p.get()[:1000], mean=nps.mean(x).get())

    r = both(apps, program)
    c, s = r["cuda"], r["serial"]
    assert rel_err(c["dot"], s["dot"]) <= 1e-10
    assert np.array_equal(c["head"], s["head"])
    assert np.array_equal(c["where"], s["where"]) and c["where"].dtype == s["where"].dtype
    assert c["argmax"] == s["argmax"] and c["argmin"] == s["argmin"]
    assert c["sum"] == s["sum"] == n * (n - 1) // 2
    assert c["max"] == s["max"]
    assert rel_err(c["mean"], s["mean"]) <= 1e-12


def test_block_matmul_and_axis_reductions(apps):
    def program(nps, app):
        A = nps.random.rand(700, 300)
        B = nps.random.rand(300, 200)
        C = A @ B
        G = A.T @ A
        return dict(C=C.get(), G=G.get(), s0=nps.sum(A, axis=0).get(), m1=nps.mean(A, axis=1).get(),
                    e=nps.exp(-A).get(), cmp=(A > 0.5).get(), mx=nps.max(A, axis=0).get())

    r = both(apps, program)
    c, s = r["cuda"], r["serial"]
    for k in ("C", "G", "s0", "m1"):
        assert rel_err(c[k], s[k]) <= 1e-10, k
    assert rel_err(c["e"], s["e"]) <= 1e-13
    assert np.array_equal(c["cmp"], s["cmp"]) and np.array_equal(c["mx"], s["mx"])


@pytest.mark.parametrize("solver,kw", [("lbfgs", dict(penalty="l2", C=2.0, max_iter=8)),
                                       ("newton", dict(penalty="l2", C=2.0)),
                                       ("gd", dict(penalty="l2", C=2.0, max_iter=5, lr=1e-4)),
                                       ("irls", dict(max_iter=5))])
def test_logistic_solvers_of_the_reference(apps, solver, kw):
    from nums.models.glms import LogisticRegression
    from nums_b200 import nums_plugin

    def program(nps, app):
        nums_plugin.disable_fused_glm()
        X, y = readme_data(nps, 20000, d=16, shift=0.5)
        m = LogisticRegression(solver=solver, **kw)
        m.fit(X, y)
        return dict(beta=m._beta.get(), beta0=m._beta0.get(), proba=m.predict_proba(X).get())

    r = both(apps, program)
    assert np.max(np.abs(r["cuda"]["beta"] - r["serial"]["beta"])) <= 1e-8
    assert abs(r["cuda"]["beta0"] - r["serial"]["beta0"]) <= 1e-8
    assert rel_err(r["cuda"]["proba"], r["serial"]["proba"]) <= 1e-8


def test_linear_and_poisson_of_the_reference(apps):
    from nums.models.glms import LinearRegression, PoissonRegression, Ridge
    from nums_b200 import nums_plugin

    def program(nps, app):
        nums_plugin.disable_fused_glm()
        X = nps.random.randn(8000, 8)
        w = nps.array(np.linspace(-0.3, 0.3, 8))
        y = X @ w + 0.1 * nps.random.randn(8000)
        out = {}
        for tag, m in (("lin_newton", LinearRegression(solver="newton")),
                       ("lin_lbfgs", LinearRegression(solver="lbfgs", max_iter=6)),
                       ("ridge", Ridge(alpha=0.5, solver="newton"))):
            m.fit(X, y)
            out[tag] = np.concatenate([m._beta.get(), [m._beta0.get()]])
        yp = nps.exp(X @ w)
        m = PoissonRegression(solver="newton", max_iter=6)
        m.fit(X, yp)
        out["poi_newton"] = np.concatenate([m._beta.get(), [m._beta0.get()]])
        m = PoissonRegression(solver="lbfgs", max_iter=6)
        m.fit(X, yp)
        out["poi_lbfgs"] = np.concatenate([m._beta.get(), [m._beta0.get()]])
        return out

    r = both(apps, program)
    for k in r["serial"]:
        assert np.max(np.abs(r["cuda"][k] - r["serial"][k])) <= 1e-8, k


@pytest.mark.parametrize("d,size,family", [(2, 10 ** 6, "logistic"), (16, 40000, "logistic"), (256, 40000, "logistic"),
                                          (8, 8000, "poisson"), (8, 8000, "ridge")])
def test_reference_glm_fit_with_the_fused_newton_routed_in(apps, d, size, family):
    """nums.models.glms with `newton` routed to the fused pass (nums_plugin.enable_fused_glm): the reference's own
    fit on the serial side, the reference's fit + fused inner loop on the device side."""
    from nums.models.glms import LogisticRegression, PoissonRegression, Ridge
    from nums_b200 import nums_plugin

    def program(nps, app):
        cuda = app is apps["cuda"]
        (nums_plugin.enable_fused_glm if cuda else nums_plugin.disable_fused_glm)()
        try:
            if family == "logistic":
                X, y = readme_data(nps, size, d=d, shift=2.0 * np.sqrt(2.0 / d))
                m = LogisticRegression(solver="newton-cg", penalty="l2", C=10)
            else:
                X = nps.random.randn(size, d)
                w = nps.array(np.linspace(-0.3, 0.3, d))
                y = nps.exp(X @ w) if family == "poisson" else X @ w + 0.1 * nps.random.randn(size)
                m = PoissonRegression(solver="newton", max_iter=8) if family == "poisson" else \
                    Ridge(alpha=0.5, solver="newton")
            with CountedNewton(app) as c:
                m.fit(X, y)
            if cuda:
                assert c.n == 0, "the generic inverse ran: the fused path was not taken"
                assert "glm_stats" in app.km.backend._remote_functions
            return np.concatenate([m._beta.get(), [m._beta0.get()]])
        finally:
            nums_plugin.disable_fused_glm()

    r = both(apps, program)
    assert np.max(np.abs(r["cuda"] - r["serial"])) <= 1e-8


def test_elementwise_and_integer_semantics_through_nums_numpy(apps):
    def program(nps, app):
        a = nps.array(np.array([7, -7, 7, -7, 0, 5, -(2 ** 62)], dtype=np.int64))
        b = nps.array(np.array([2, 2, -2, -2, 3, 0, 3], dtype=np.int64))
        x = nps.random.rand(1000) - 0.5
        return dict(fd=(a // b).get(), md=(a % b).get(), td=(a / (b + 10)).get(), mul=(a * b).get(),
                    lt=(a < b).get(), sq=nps.sqrt(nps.abs(x)).get(), mx=nps.maximum(x, -x).get(),
                    wh=nps.where(x > 0, x, -x).get(), amin=nps.min(x).get(), sm=nps.sum(a).get())

    r = both(apps, program)
    c, s = r["cuda"], r["serial"]
    for k in ("fd", "md", "mul", "lt", "mx", "wh", "amin", "sm"):
        assert np.array_equal(c[k], s[k]) and np.asarray(c[k]).dtype == np.asarray(s[k]).dtype, k
    assert rel_err(c["td"], s["td"]) <= 1e-15 and rel_err(c["sq"], s["sq"]) <= 1e-15
