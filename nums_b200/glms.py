"""Generalised linear models on device-resident block arrays: the API of nums/models/glms.py
(LogisticRegression / LinearRegression / Ridge / PoissonRegression; solvers newton, newton-cg, lbfgs,
gd) with the inner loop replaced by the fused kernels of libnumsb200.so.

Two execution modes, same results:
  fused=True  (default)  per row block ONE `glm_stats` pass (objective, gradient, Hessian; the ones
              column of glms.py:124-135 is implicit), partials summed with the reference's FIFO tree
              (and, when torch.distributed is initialised with world_size > 1, one NCCL all-reduce of
              the packed p*p+p+1 doubles), then `glm_newton_update` on every rank identically — no
              broadcast of beta is needed.
  fused=False the reference algorithm op by op through the generic block kernels (X@beta, exp, X^T r,
              s*X, X^T(sX), inv ...), exactly as glms.py:457-499/660-676 schedules it.  Used to
              cross-check the generic kernels end to end.
Quirks kept on purpose: a non-finite gradient or Hessian ends newton() with beta = NaN, as the reference's loop does
(np.linalg.inv of a NaN matrix returns NaNs, and NaN <= tol never stops it), not with LinAlgError; "newton-cg" is plain
Newton with an explicit inverse (glms.py:179-182,673);
the convergence test uses the gradient from before the update (:674); the objective penalty is the
un-squared norm over all of beta (:242-252) while the gradient penalty skips the intercept (:254-264).
"""
import warnings

import numpy as np

from . import fused as _fused
from .blockarray import BlockArray
from .grid import ArrayGrid


def _dist():
    try:
        import torch.distributed as dist
    except Exception:  # pragma: no cover
        return None
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


class GLM:
    family = None

    def __init__(self, app=None, penalty="none", alpha=1.0, l1_ratio=0.5, tol=0.0001, max_iter=100,
                 solver="newton", lr=0.01, random_state=None, fit_intercept=True, normalize=False,
                 fused=True, sharded=None):
        if fit_intercept is False:
            raise NotImplementedError("fit_incercept=False currently not supported.")
        if normalize is True:
            raise NotImplementedError("normalize=True currently not supported.")
        if app is None:
            from .app import instance
            app = instance()
        self._app = app
        if random_state is None:
            self.rs = app.random
        elif isinstance(random_state, (int, np.integer)):
            from .app import NumsRandomState
            self.rs = NumsRandomState(app.km, int(random_state))
        else:
            self.rs = random_state
        self._penalty = None if penalty == "none" else penalty
        if self._penalty not in (None, "l1", "l2", "elasticnet"):
            raise NotImplementedError("%s penalty not supported" % self._penalty)
        self._lambda = alpha
        self.alpha = l1_ratio
        self._l1penalty = self._l1penalty_vec = None
        self._l2penalty = self._l2penalty_vec = self._l2penalty_diag = None
        self._tol = tol
        self._max_iter = max_iter
        self._opt = solver
        self._lr = lr
        self._beta = None
        self._beta0 = None
        self._fused = fused
        self._sharded = sharded       # None: all-reduce iff torch.distributed is initialised
        self.n_iter_ = None
        _fused.register(app.km)

    # ------------------------------------------------------------------ fit
    def fit(self, X, y):
        app = self._app
        if X.dtype.kind in "iu":
            X = X.astype(np.float64)
        assert X.ndim == 2 and y.ndim == 1
        d = X.shape[1]
        p = d + 1
        self._setup_penalties(p, X.block_shape[1] if not self._fused else p)
        if self._fused:
            if X.grid.grid_shape[1] != 1:
                raise NotImplementedError("the fused path needs X blocked along rows only")
            beta = app.zeros((p,), (p,), np.float64)
            beta = self._solve(beta, X, y)
            self._split_beta(beta, d)
            return self
        Xa = app.concatenate([X, app.ones((X.shape[0], 1), (X.block_shape[0], 1), X.dtype)], axis=1,
                             axis_block_size=X.block_shape[1])
        beta = app.zeros((p,), (X.block_shape[1],), X.dtype)
        beta = self._solve(beta, Xa, y)
        self._split_beta(beta, d)
        return self

    def _solve(self, beta, X, y):
        tol, max_iter = self._tol, self._max_iter
        if self._opt in ("newton", "newton-cg"):
            if self._opt == "newton-cg":
                warnings.warn("Specified newton-cg solver, using newton instead.")
            return newton(self, beta, X, y, tol, max_iter)
        if self._opt == "lbfgs":
            from .lbfgs import LBFGS
            return LBFGS(self, m=10, max_iter=max_iter, thresh=tol).execute(X, y, beta)
        if self._opt == "gd":
            return gd(self, beta, X, y, tol, max_iter, self._lr)
        if self._opt == "sgd":
            return sgd(self, beta, X, y, tol, max_iter, self._lr)
        if self._opt == "block_sgd":
            return block_sgd(self, beta, X, y, tol, max_iter, self._lr)
        if self._opt == "irls":
            assert isinstance(self, LogisticRegression)
            return irls(self, beta, X, y, tol, max_iter)
        raise Exception("Unsupported optimizer specified %s." % self._opt)

    def _setup_penalties(self, p, beta_block):
        app = self._app

        def vec(value):
            v = np.full(p, value, dtype=np.float64)
            v[-1] = 0.0                     # the intercept is not penalised (glms.py:150,155,162,169)
            return v

        if self._penalty == "elasticnet":
            self._l1penalty = self.alpha * self._lambda
            self._l2penalty = 0.5 * (1.0 - self.alpha) * self._lambda
        elif self._penalty == "l2":
            self._l2penalty = 0.5 * self._lambda
        elif self._penalty == "l1":
            self._l1penalty = self._lambda
        if self._l1penalty is not None:
            self._l1penalty_vec = app.array(vec(self._l1penalty), (beta_block,))
        if self._l2penalty is not None:
            l2 = vec(self._l2penalty)
            self._l2penalty_vec = app.array(l2, (beta_block,))
            if not self._fused:
                self._l2penalty_diag = app.array(np.diag(l2), (beta_block, beta_block))

    def _split_beta(self, beta, d):
        km = self._app.km
        one = _gather_single_block(beta)
        sk = {"grid_entry": (0,), "grid_shape": (1,)}
        b = km.create_block(one, src_params=[((slice(0, d),), False)], dst_params=[((slice(0, d),), False)],
                            dst_shape=(d,), dst_shape_bc=None, syskwargs=sk)
        b0 = km.create_block(one, src_params=[((d,), False)], dst_params=[((), False)],
                             dst_shape=(), dst_shape_bc=None, syskwargs={"grid_entry": (), "grid_shape": ()})
        self._beta = BlockArray.from_oid(b, (d,), np.float64, km)
        self._beta0 = BlockArray.from_oid(b0, (), np.float64, km)
        self._beta_full = BlockArray.from_oid(one, (d + 1,), np.float64, km)

    # ------------------------------------------------------------------ fused building blocks
    def _allreduce(self, stats_oid):
        use = self._sharded
        dist = _dist()
        if use is None:
            use = dist is not None
        if not use:
            return stats_oid
        if dist is None:
            raise RuntimeError("sharded=True needs an initialised torch.distributed process group")
        # ONE collective per pass: the packed [H | g | obj] buffer (p*p+p+1 doubles) over NCCL
        dist.all_reduce(stats_oid.t, op=dist.ReduceOp.SUM)
        return stats_oid

    def fused_stats(self, X, y, beta, want_hess, family=None, need_obj=True):
        """Sum over all row blocks (and ranks) of the packed statistics at `beta` (single block)."""
        km = self._app.km
        family = self.family if family is None else family
        b_oid = beta.blocks[(0,)].oid
        items = []
        for i in range(X.grid.grid_shape[0]):
            xb, yb = X.blocks[(i, 0)], y.blocks[(i,)]
            if xb.transposed:
                raise NotImplementedError("fused path on a transposed X")
            oid = km.call("glm_stats", xb.oid, yb.oid, b_oid, family, want_hess, need_obj, syskwargs=xb.syskwargs())
            items.append((oid, xb.grid_entry, xb.grid_shape, False))
        total = X.tree_reduce("sum", items, (0,), (1,))
        return self._allreduce(total)

    # ------------------------------------------------------------------ reference-shaped model API
    def forward(self, X, beta=None):
        if self._fused:
            b = beta if beta is not None else self._beta_full
            km = self._app.km
            grid = ArrayGrid((X.shape[0],), (X.block_shape[0],), "float64")
            out = BlockArray(grid, km)
            for i in range(X.grid.grid_shape[0]):
                xb = X.blocks[(i, 0)]
                out.blocks[(i,)].oid = km.call("glm_predict", xb.oid, b.blocks[(0,)].oid, self.family,
                                               syskwargs=xb.syskwargs())
            return out
        if beta is not None:
            return self.link_inv(X @ beta)
        return self.link_inv(self._beta0 + X @ self._beta)

    def predict(self, X):
        return self.forward(X)

    def obj_penalty(self, beta):
        app = self._app
        if self._penalty == "l1":
            return self._l1penalty * app.norm(beta, order=1)
        if self._penalty == "l2":
            return self._l2penalty * app.norm(beta, order=2)
        if self._penalty == "elasticnet":
            return self._l2penalty * app.norm(beta, order=2) + self._l1penalty * app.norm(beta, order=1)
        raise ValueError("Unexpected call to objective term, penalty=None.")

    def grad_penalty(self, beta):
        app = self._app
        if self._penalty == "l1":
            return self._l1penalty_vec * app.map_uop("sign", beta)
        if self._penalty == "l2":
            return self._l2penalty_vec * beta
        if self._penalty == "elasticnet":
            return self._l2penalty_vec * beta + self._l1penalty_vec * app.map_uop("sign", beta)
        raise ValueError("Unexpected call to objective term, penalty=None.")

    def hessian_penalty(self):
        if self._penalty == "l1":
            return 0.0
        if self._penalty in ("l2", "elasticnet"):
            return self._l2penalty_diag
        raise ValueError("Unexpected call to objective term, penalty=None.")

    def objective_and_gradient(self, X, y, beta):
        """(f(beta), grad f(beta)) from one fused pass (+ penalties), as single-block arrays."""
        km = self._app.km
        p = beta.shape[0]
        stats = self.fused_stats(X, y, beta, False)
        sk = {"grid_entry": (0,), "grid_shape": (1,)}
        f = BlockArray.from_oid(km.call("stats_slice", stats, p, "obj", syskwargs=sk), (), np.float64, km)
        g = BlockArray.from_oid(km.call("stats_slice", stats, p, "g", syskwargs=sk), (p,), np.float64, km)
        if self._penalty is not None:
            # the family decides whether the objective carries the penalty (Poisson's does not, :519-531)
            if self.objective_has_penalty:
                f = f + self.obj_penalty(beta)
            if self.gradient_has_penalty:
                g = g + self.grad_penalty(beta)
        return f, g

    objective_has_penalty = True
    gradient_has_penalty = True

    def deviance_sqr(self, X, y):
        y_pred = self.predict(X)
        dev = self.deviance(y, y_pred)
        dev_null = self.deviance(y, self._app.mean(y))
        return 1 - dev / dev_null


def _gather_single_block(vec):
    """oid of a 1-d block array as ONE block (copies only when it has several)."""
    if vec.grid.grid_shape == (1,):
        return vec.blocks[(0,)].oid
    km = vec.km
    oids, sp, dp = [], [], []
    for entry in vec.grid.get_entry_iterator():
        lo, hi = vec.grid.get_slice_tuples(entry)[0]
        oids.append(vec.blocks[entry].oid)
        sp.append(((slice(0, hi - lo),), False))
        dp.append(((slice(lo, hi),), False))
    return km.create_block(*oids, src_params=sp, dst_params=dp, dst_shape=vec.shape, dst_shape_bc=None,
                           syskwargs={"grid_entry": (0,), "grid_shape": (1,)})


def _to_single_block_matrix(M):
    """(oid, was_single) of a 2-d block array gathered into one block (linalg.py:246-261 reshapes H so)."""
    if M.grid.grid_shape == (1, 1):
        return M.blocks[(0, 0)].oid
    km = M.km
    oids, sp, dp = [], [], []
    for entry in M.grid.get_entry_iterator():
        (r0, r1), (c0, c1) = M.grid.get_slice_tuples(entry)
        blk = M.blocks[entry]
        oids.append(blk.oid)
        sp.append(((slice(0, r1 - r0), slice(0, c1 - c0)), blk.transposed))
        dp.append(((slice(r0, r1), slice(c0, c1)), False))
    return km.create_block(*oids, src_params=sp, dst_params=dp, dst_shape=M.shape, dst_shape_bc=None,
                           syskwargs={"grid_entry": (0, 0), "grid_shape": (1, 1)})


def _split_matrix_like(oid, like):
    """Cuts a single (p,p) block back into the block partition of `like`."""
    km = like.km
    out = BlockArray(ArrayGrid(like.shape, like.block_shape, "float64"), km)
    for entry in out.grid.get_entry_iterator():
        (r0, r1), (c0, c1) = out.grid.get_slice_tuples(entry)
        out.blocks[entry].oid = km.create_block(
            oid, src_params=[((slice(r0, r1), slice(c0, c1)), False)],
            dst_params=[((slice(0, r1 - r0), slice(0, c1 - c0)), False)], dst_shape=(r1 - r0, c1 - c0),
            dst_shape_bc=None, syskwargs={"grid_entry": entry, "grid_shape": out.grid.grid_shape})
    return out


def inv(app, M):
    """linalg.inv (nums/core/linalg.py:246-261): gather to one block, LAPACK-style inverse, re-block."""
    oid = _to_single_block_matrix(M)
    r = app.km.inv(oid, syskwargs={"grid_entry": (0, 0), "grid_shape": (1, 1)})
    return _split_matrix_like(r, M)


# ---------------------------------------------------------------------- families
class LinearRegressionBase(GLM):
    family = "linear"

    def link_inv(self, eta):
        return eta

    def objective(self, X, y, beta=None, mu=None):
        mu = self.forward(X, beta) if mu is None else mu
        r = self._app.sum((y - mu) ** self._app.two)
        if self._penalty is not None:
            r = r + self.obj_penalty(beta)
        return r

    def gradient(self, X, y, mu=None, beta=None):
        if mu is None:
            mu = self.forward(X)
        r = X.transpose(defer=True) @ (mu - y)
        if self._penalty is not None:
            r = r + self.grad_penalty(beta)
        return r

    def hessian(self, X, y, mu=None):
        r = X.transpose(defer=True) @ X
        if self._penalty is not None:
            r = r + self.hessian_penalty()
        return r

    def deviance(self, y, y_pred):
        return self._app.sum((y - y_pred) ** self._app.two)


class LinearRegression(LinearRegressionBase):
    def __init__(self, tol=0.0001, max_iter=100, solver="newton", lr=0.01, **kw):
        super().__init__(tol=tol, max_iter=max_iter, solver=solver, lr=lr, **kw)


class Ridge(LinearRegressionBase):
    def __init__(self, alpha=1.0, tol=0.0001, max_iter=100, solver="newton", lr=0.01, **kw):
        super().__init__(penalty="l2", alpha=alpha, tol=tol, max_iter=max_iter, solver=solver, lr=lr, **kw)


class ElasticNet(LinearRegressionBase):
    def __init__(self, alpha=1.0, l1_ratio=0.5, tol=0.0001, max_iter=100, solver="newton", lr=0.01, **kw):
        super().__init__(penalty="elasticnet", alpha=alpha, l1_ratio=l1_ratio, tol=tol, max_iter=max_iter,
                         solver=solver, lr=lr, **kw)


class Lasso(LinearRegressionBase):
    def __init__(self, alpha=1.0, tol=0.0001, max_iter=100, solver="newton", lr=0.01, **kw):
        super().__init__(penalty="l1", alpha=alpha, tol=tol, max_iter=max_iter, solver=solver, lr=lr, **kw)


class LogisticRegression(GLM):
    family = "logistic"

    def __init__(self, penalty="none", C=1.0, tol=0.0001, max_iter=100, solver="newton", lr=0.01, **kw):
        super().__init__(penalty=penalty, alpha=1.0 / C, tol=tol, max_iter=max_iter, solver=solver, lr=lr, **kw)

    def link_inv(self, eta):
        one = self._app.one
        return one / (one + self._app.exp(-eta))

    def objective(self, X, y, beta=None, mu=None):
        log, one = self._app.log, self._app.one
        mu = self.forward(X, beta) if mu is None else mu
        r = -self._app.sum(y * log(mu) + (one - y) * log(one - mu))
        if self._penalty is not None:
            r = r + self.obj_penalty(beta)
        return r

    def gradient(self, X, y, mu=None, beta=None):
        if mu is None:
            mu = self.forward(X)
        r = X.transpose(defer=True) @ (mu - y)
        if self._penalty is not None:
            r = r + self.grad_penalty(beta)
        return r

    def hessian(self, X, y, mu=None):
        if mu is None:
            mu = self.forward(X)
        s = (mu * (self._app.one - mu)).reshape((mu.shape[0], 1), (mu.block_shape[0], 1))
        r = X.transpose(defer=True) @ (s * X)
        if self._penalty is not None:
            r = r + self.hessian_penalty()
        return r

    def predict(self, X):
        return (self.forward(X) > 0.5).astype(np.int64)

    def predict_proba(self, X):
        # glms.py:507-512: columns (P(y=1), P(y=0))
        pos = self.forward(X).reshape((X.shape[0], 1), (X.block_shape[0], 1))
        neg = 1 - pos
        out = BlockArray(ArrayGrid((X.shape[0], 2), (X.block_shape[0], 2), "float64"), self._app.km)
        for i in range(pos.grid.grid_shape[0]):
            pb, nb = pos.blocks[(i, 0)], neg.blocks[(i, 0)]
            out.blocks[(i, 0)].oid = self._app.km.create_block(
                pb.oid, nb.oid, src_params=[((slice(None), slice(0, 1)), False), ((slice(None), slice(0, 1)), False)],
                dst_params=[((slice(None), slice(0, 1)), False), ((slice(None), slice(1, 2)), False)],
                dst_shape=(pb.shape[0], 2), dst_shape_bc=None, syskwargs=pb.syskwargs())
        return out


class PoissonRegression(GLM):
    family = "poisson"
    objective_has_penalty = False     # glms.py:519-531 and :533-544 never add the penalty terms
    gradient_has_penalty = False

    def link_inv(self, eta):
        return self._app.exp(eta)

    def objective(self, X, y, beta=None, mu=None):
        eta = X @ beta if beta is not None else X @ self._beta + self._beta0
        mu = self._app.exp(eta) if mu is None else mu
        return self._app.sum(mu - y * eta)

    def gradient(self, X, y, mu=None, beta=None):
        if mu is None:
            mu = self.forward(X)
        return X.transpose(defer=True) @ (mu - y)

    def hessian(self, X, y, mu=None):
        if mu is None:
            mu = self.forward(X)
        return (X.transpose(defer=True) * mu) @ X

    def deviance(self, y, y_pred):
        app = self._app
        yf = y.astype(np.float64) if y.dtype.kind != "f" else y
        return app.sum(app.two * yf.__elementwise__("xlogy", yf / y_pred) - y + y_pred)


PoissonRegressor = PoissonRegression


# ---------------------------------------------------------------------- solvers
def newton(model, beta, X, y, tol, max_iter):
    """glms.py:660-676."""
    app = model._app
    km = app.km
    model.n_iter_ = 0
    if model._fused:
        poisson_skip = isinstance(model, PoissonRegression)   # its gradient/hessian carry no penalty
        l2 = None
        if model._l2penalty_vec is not None and not poisson_skip:
            l2 = model._l2penalty_vec.blocks[(0,)].oid
        sk = {"grid_entry": (0,), "grid_shape": (1,)}
        p = beta.shape[0]
        R = X.grid.grid_shape[0]
        sharded = model._sharded if model._sharded is not None else (_dist() is not None)
        one_device = len({X.blocks[(i, 0)].oid.device_index for i in range(R)}) == 1
        if (not sharded and one_device and (model._l1penalty_vec is None or poisson_skip)
                and not any(X.blocks[(i, 0)].transposed for i in range(R)) and max_iter > 0):
            # every row block on one GPU and no collective in the loop: the loop itself runs in the library
            # (nb_glm_newton_fit); the host sees one call per fit instead of ~5 per iteration
            oids = [X.blocks[(i, 0)].oid for i in range(R)] + [y.blocks[(i,)].oid for i in range(R)]
            b_new, (iters, gm, inf) = km.call("glm_newton_fit", model.family, beta.blocks[(0,)].oid, l2, float(tol),
                                               int(max_iter), None, *oids, syskwargs=X.blocks[(0, 0)].syskwargs())
            model.n_iter_ = iters
            if inf != 0:
                if gm != gm:          # non-finite g / H: the reference's loop ends with beta = NaN (see below)
                    return app.array(np.full(p, np.nan), (p,))
                raise np.linalg.LinAlgError("Singular matrix")
            return BlockArray.from_oid(b_new, (p,), np.float64, km)
        for _ in range(max_iter):
            stats = model.fused_stats(X, y, beta, True, need_obj=False)     # newton() never evaluates the objective
            if model._l1penalty_vec is not None and not poisson_skip:
                # the L1 term of grad_penalty (glms.py:254-264) is added to g before the update; its Hessian
                # contribution is zero (:266-272)
                extra = model._l1penalty_vec * app.map_uop("sign", beta)
                stats = km.call("stats_add_g", stats, extra.blocks[(0,)].oid, syskwargs=sk)
            b_new, verdict = km.call("glm_newton_step", stats, l2, beta.blocks[(0,)].oid, syskwargs=sk)
            beta = BlockArray.from_oid(b_new, (p,), np.float64, km)
            model.n_iter_ += 1
            gm, inf = _fused.read_verdict(km.get(verdict))      # the one host sync per iteration (glms.py:674)
            if inf != 0:
                if gm != gm:
                    # a NaN/Inf reached g or H (diverged eta, non-finite X): np.linalg.inv returns NaNs without
                    # raising and the reference's loop runs on with beta = NaN until max_iter (glms.py:673-675);
                    # the result is the same NaN vector, returned at once
                    beta = app.array(np.full(p, np.nan), (p,))
                    break
                raise np.linalg.LinAlgError("Singular matrix")
            if gm <= tol:
                break
        return beta
    tol_ba = app.scalar(tol)
    for _ in range(max_iter):
        mu = model.forward(X, beta)
        g = model.gradient(X, y, mu, beta=beta)
        beta = beta + (-inv(app, model.hessian(X, y, mu))) @ g
        model.n_iter_ += 1
        if app.max(app.abs(g)) <= tol_ba:
            break
    return beta


def gd(model, beta, X, y, tol, max_iter, lr):
    """glms.py:642-657."""
    app = model._app
    model.n_iter_ = 0
    tol_ba = app.scalar(tol)
    lr_ba = app.scalar(lr)
    for _ in range(max_iter):
        if model._fused:
            _, g = model.objective_and_gradient(X, y, beta)
        else:
            mu = model.forward(X, beta)
            g = model.gradient(X, y, mu, beta=beta)
        beta = beta + (-lr_ba * g)
        model.n_iter_ += 1
        if app.max(app.abs(g)) <= tol_ba:
            break
    return beta


def _row_range(arr, start, stop):
    """Rows [start, stop) of a row-blocked array as an array whose blocks are windows on the owning blocks'
    memory (the reference-copy case of view.py:104-128); the range must lie inside one row block."""
    from .device import DeviceBlock
    rb = arr.block_shape[0]
    i = start // rb
    if (stop - 1) // rb != i:
        raise NotImplementedError("row ranges that straddle blocks")
    lo, cnt = start - i * rb, stop - start
    shape = (cnt,) + tuple(arr.shape[1:])
    block_shape = (cnt,) + tuple(arr.block_shape[1:])
    out = BlockArray(ArrayGrid(shape, block_shape, arr.grid.dtype), arr.km)
    for entry in out.grid.get_entry_iterator():
        src = arr.blocks[(i,) + tuple(entry[1:])]
        if src.transposed:
            raise NotImplementedError("row range of a transposed block")
        out.blocks[entry].oid = DeviceBlock(src.oid.t[lo:lo + cnt])
    return out


def _gradient_any(model, X, y, beta):
    if model._fused:
        if _dist() is not None and model._sharded is not False:
            raise NotImplementedError("sgd / block_sgd sample from the global array: not available row-sharded")
        return model.objective_and_gradient(X, y, beta)[1]
    return model.gradient(X, y, model.forward(X, beta), beta=beta)


def sgd(model, beta, X, y, tol, max_iter, lr):
    """glms.py:595-616: one uniformly drawn row per step (host RNG: model.rs.numpy().integers)."""
    app = model._app
    tol_ba, lr_ba = app.scalar(tol), app.scalar(lr)
    model.n_iter_ = 0
    for _ in range(max_iter):
        idx = int(model.rs.numpy().integers(X.shape[0]))
        g = _gradient_any(model, _row_range(X, idx, idx + 1), _row_range(y, idx, idx + 1), beta)
        beta = beta + (-lr_ba * g)
        model.n_iter_ += 1
        if app.max(app.abs(g)) <= tol_ba:
            break
    return beta


def block_sgd(model, beta, X, y, tol, max_iter, lr):
    """glms.py:619-639: one step per row block; the convergence test only leaves the inner loop (as in the
    reference)."""
    app = model._app
    tol_ba, lr_ba = app.scalar(tol), app.scalar(lr)
    model.n_iter_ = 0
    rb = X.block_shape[0]
    for _ in range(max_iter):
        for start in range(0, X.shape[0], rb):
            stop = min(X.shape[0], start + rb)
            g = _gradient_any(model, _row_range(X, start, stop), _row_range(y, start, stop), beta)
            beta = beta + (-lr_ba * g)
            if app.max(app.abs(g)) <= tol_ba:
                break
        model.n_iter_ += 1
    return beta


def irls(model, beta, X, y, tol, max_iter):
    """glms.py:679-700.  Fused: H' = X^T diag(mu(1-mu)+1e-16) X and g0 = X^T(mu - y) from one statistics
    pass; X^T S z = H' beta - g0 because z = eta + (y - mu)/s and eta = X beta."""
    app = model._app
    km = app.km
    model.n_iter_ = 0
    tol_ba = app.scalar(tol)
    if not model._fused:
        for _ in range(max_iter):
            eta = X @ beta
            mu = model.link_inv(eta)
            s = mu * (1 - mu) + 1e-16
            XT_s = X.transpose(defer=True) * s
            XTsX_inv = inv(app, XT_s @ X)
            z = eta + (y - mu) / s
            beta = XTsX_inv @ XT_s @ z
            g = model.gradient(X, y, mu, beta)
            model.n_iter_ += 1
            if app.max(app.abs(g)) <= tol_ba:
                break
        return beta
    p = beta.shape[0]
    sk = {"grid_entry": (0,), "grid_shape": (1,)}
    for _ in range(max_iter):
        stats = model.fused_stats(X, y, beta, True, family="logistic_irls")
        H = BlockArray.from_oid(km.call("stats_slice", stats, p, "H", syskwargs={"grid_entry": (0, 0), "grid_shape": (1, 1)}),
                                (p, p), np.float64, km)
        g0 = BlockArray.from_oid(km.call("stats_slice", stats, p, "g", syskwargs=sk), (p,), np.float64, km)
        new_beta = inv(app, H) @ (H @ beta - g0)
        g = g0 if model._penalty is None else g0 + model.grad_penalty(new_beta)    # :698 uses the NEW beta
        beta = new_beta
        model.n_iter_ += 1
        if app.max(app.abs(g)) <= tol_ba:
            break
    return beta
