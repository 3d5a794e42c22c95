"""L-BFGS with backtracking line search: the algorithm of nums/models/lbfgs.py (two-loop recursion
:86-100, memory :55-62, line search :37-52, driver :102-155; constants m=10, rho=1e-2, c=1e-4,
alpha0=1, alpha_min=1e-30, eps=1e-30).

The p-sized vector algebra runs through the block kernels exactly as the reference schedules it
(including the dense gamma*I "H @ q" product, :78-84,96).  What changes is the cost of an evaluation:
with a fused model every f(theta) is ONE pass over X that returns the objective AND the gradient, so
the accepted trial's gradient is reused instead of the reference's two extra passes (:136-138).
"""
import numpy as np

from .blockarray import BlockArray


class LBFGSMemory:
    def __init__(self, k, s, y):
        self.k = k
        self.s = s
        self.y = y
        ys_inner = s.transpose(defer=True) @ y
        self.rho = 1.0 / (ys_inner + 1e-30)
        self.gamma = ys_inner / (y.transpose(defer=True) @ y + 1e-30)


class BackTrackingLineSearch:
    def __init__(self, model):
        self.app = model._app
        self.model = model

    def f(self, X, y, theta):
        """-> (objective, gradient or None)"""
        m = self.model
        if m._fused:
            return m.objective_and_gradient(X, y, theta)
        return m.objective(X, y, theta, m.forward(X, theta)), None

    def execute(self, X, y, theta, grad, p, f_val=None, rho=1.0e-1, init_alpha=1.0, c=1e-4, min_alpha=1e-10):
        """Returns (alpha, f(theta + alpha p), gradient there or None)."""
        alpha = init_alpha
        if f_val is None:
            f_val, _ = self.f(X, y, theta)
        f_next, g_next = self.f(X, y, theta + alpha * p)
        slope = grad.transpose(defer=True) @ p
        while self.app.isnan(f_next) or f_next > f_val + c * alpha * slope:
            alpha *= rho
            if alpha < min_alpha:
                return min_alpha, None, None
            f_next, g_next = self.f(X, y, theta + alpha * p)
        return max(alpha, min_alpha), f_next, g_next


class LBFGS:
    def __init__(self, model, m=10, max_iter=100, thresh=1e-4, dtype=np.float64):
        self.app = model._app
        self.model = model
        self.m = m
        self.max_iter = max_iter
        self.thresh = thresh
        self.dtype = dtype
        self.k = 0
        self.identity = None
        self.memory = [None] * m
        self.ls = BackTrackingLineSearch(model)

    def get_H(self):
        if self.k == 0:
            return self.identity
        mem = self.memory[-1]
        assert mem.k == self.k - 1
        return mem.gamma * self.identity

    def get_p(self, H, g):
        q = g
        forward_vars = []
        for i in range(-1, -self.m - 1, -1):
            mem_i = self.memory[i]
            if mem_i is None:
                break
            alpha = mem_i.rho * mem_i.s.transpose(defer=True) @ q
            q = q - alpha * mem_i.y
            forward_vars.insert(0, (alpha, mem_i))
        r = H @ q
        for alpha, mem_i in forward_vars:
            beta = mem_i.rho * mem_i.y.transpose(defer=True) @ r
            r = r + mem_i.s * (alpha - beta)
        return r

    def _gradient(self, X, y, theta):
        m = self.model
        if m._fused:
            return m.objective_and_gradient(X, y, theta)
        return None, m.gradient(X, y, m.forward(X, theta), theta)

    def execute(self, X, y, theta):
        if self.k != 0:
            raise Exception("Unexpected state.")
        model = self.model
        bs = theta.block_shape[0]
        self.identity = self.app.eye((theta.shape[0], theta.shape[0]), (bs, bs), dtype=self.dtype)
        f_val, g = self._gradient(X, y, theta)
        thresh = self.app.scalar(self.thresh)
        model.n_iter_ = 0
        while self.k < self.max_iter:
            H = self.get_H()
            p = -self.get_p(H, g)
            alpha, f_next, g_next = self.ls.execute(X, y, theta, g, p, f_val=f_val, rho=1e-2, init_alpha=1.0,
                                                     c=1e-4, min_alpha=1e-30)
            next_theta = theta + alpha * p
            model.n_iter_ += 1
            if self.k + 1 >= self.max_iter:
                theta = next_theta
                break
            if g_next is None:
                f_next, g_next = self._gradient(X, y, next_theta)
            mem = LBFGSMemory(k=self.k, s=next_theta - theta, y=g_next - g)
            self.memory.append(mem)
            self.memory.pop(0)
            self.k += 1
            theta, g, f_val = next_theta, g_next, f_next
            if self.app.max(self.app.abs(g)) <= thresh:
                break
        self.k = 0
        self.identity = None
        self.memory = [None] * self.m
        return theta
