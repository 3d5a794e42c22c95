"""ORACLE — test infrastructure only.  Never imported by the nums_b200 package.

NumPy restatement of the reference's GLM fit (nums/models/glms.py + nums/models/lbfgs.py, v0.2.8)
in the order the reference schedules it on row blocks: every BlockArray operator becomes one NumPy
call per block, contractions are per-block matmuls summed by the FIFO pairwise tree of
blockarray.py:643-689, and GLM.fit's appended ones column (glms.py:124-135) is a real second column
block, so each X-sized temporary the reference materialises (s*X, glms.py:495-496) is materialised
here too.  This is what `bench.py` times as the CPU baseline ("port").
"""
import numpy as np


def split_rows(arr, row_block):
    return [arr[i:i + row_block] for i in range(0, arr.shape[0], row_block)]


def tree_sum(parts):
    """blockarray.py:643-689 with op "sum": pop two from the front, append their sum."""
    q = list(parts)
    while len(q) > 1:
        a = q.pop(0)
        b = q.pop(0)
        q.append(a + b)
    return q[0]


class Family:
    """link_inv / objective / hessian weights of glms.py:275-327 (linear), :430-512 (logistic),
    :515-556 (poisson)."""

    def __init__(self, name):
        if name not in ("linear", "logistic", "poisson"):
            raise ValueError(name)
        self.name = name

    def link_inv(self, eta):
        if self.name == "logistic":
            return 1.0 / (1.0 + np.exp(-eta))          # glms.py:457-458
        if self.name == "poisson":
            return np.exp(eta)                          # glms.py:516-517
        return eta                                      # glms.py:283-284

    def objective_terms(self, y, mu, eta):
        if self.name == "logistic":
            with np.errstate(divide="ignore", invalid="ignore"):
                return -(y * np.log(mu) + (1.0 - y) * np.log(1.0 - mu))    # glms.py:470 (summed by caller)
        if self.name == "poisson":
            return mu - y * eta                         # glms.py:531
        return (y - mu) ** 2.0                          # glms.py:296

    def weights(self, mu):
        if self.name == "logistic":
            return mu * (1.0 - mu)                      # glms.py:495
        if self.name == "poisson":
            return mu                                   # glms.py:548
        return None                                     # glms.py:318: plain Gram


class BlockedGLM:
    """State of one fit: row blocks of X (float64) and y, the implicit ones blocks, penalties."""

    def __init__(self, family, X_blocks, y_blocks, penalty=None, lam=1.0, l1_ratio=0.5):
        self.fam = Family(family)
        self.Xb = X_blocks
        self.yb = y_blocks
        self.ones = [np.ones((xb.shape[0], 1)) for xb in X_blocks]
        self.d = X_blocks[0].shape[1]
        self.p = self.d + 1
        self.penalty = penalty
        self.l1 = self.l2 = None
        # glms.py:141-173
        if penalty == "elasticnet":
            self.l1 = l1_ratio * lam
            self.l2 = 0.5 * (1.0 - l1_ratio) * lam
        elif penalty == "l2":
            self.l2 = 0.5 * lam
        elif penalty == "l1":
            self.l1 = lam
        self.l1vec = self.l2vec = None
        if self.l1 is not None:
            self.l1vec = self.l1 * np.ones(self.p)
            self.l1vec[-1] = 0.0
        if self.l2 is not None:
            self.l2vec = self.l2 * np.ones(self.p)
            self.l2vec[-1] = 0.0
        self.skip_penalty = family == "poisson"   # glms.py:519-548 never add the penalty terms

    # ---- pieces, each as the reference schedules them
    def eta_blocks(self, beta):
        bx, b1 = beta[: self.d], beta[self.d:]
        return [np.matmul(xb, bx) + np.matmul(ob, b1) for xb, ob in zip(self.Xb, self.ones)]

    def forward(self, beta):
        etas = self.eta_blocks(beta)
        return etas, [self.fam.link_inv(e) for e in etas]

    def obj_penalty(self, beta):
        # glms.py:242-252: un-squared norms over ALL of beta (quirk Q3)
        out = 0.0
        if self.l2 is not None:
            out = out + self.l2 * np.sqrt(np.matmul(beta, beta))
        if self.l1 is not None:
            out = out + self.l1 * np.sum(np.abs(beta))
        return out

    def grad_penalty(self, beta):
        # glms.py:254-264
        out = 0.0
        if self.l2vec is not None:
            out = out + self.l2vec * beta
        if self.l1vec is not None:
            out = out + self.l1vec * np.sign(beta)
        return out

    def objective(self, beta, fwd=None):
        etas, mus = fwd if fwd is not None else self.forward(beta)
        parts = [np.sum(self.fam.objective_terms(y, m, e)) for y, m, e in zip(self.yb, mus, etas)]
        r = tree_sum(parts)
        if self.penalty is not None and not self.skip_penalty:
            r = r + self.obj_penalty(beta)
        return r

    def gradient(self, beta, fwd=None):
        _, mus = fwd if fwd is not None else self.forward(beta)
        res = [m - y for m, y in zip(mus, self.yb)]
        gx = tree_sum([np.matmul(xb.T, r) for xb, r in zip(self.Xb, res)])
        g1 = tree_sum([np.matmul(ob.T, r) for ob, r in zip(self.ones, res)])
        g = np.concatenate([gx, g1])
        if self.penalty is not None and not self.skip_penalty:
            g = g + self.grad_penalty(beta)
        return g

    def hessian(self, beta, fwd=None):
        _, mus = fwd if fwd is not None else self.forward(beta)
        h00, h01, h10, h11 = [], [], [], []
        for xb, ob, m in zip(self.Xb, self.ones, mus):
            w = self.fam.weights(m)
            if w is None:
                sx, s1 = xb, ob
            else:
                s = w.reshape((-1, 1))
                sx, s1 = s * xb, s * ob        # the X-sized temporary of glms.py:495-496
            h00.append(np.matmul(xb.T, sx))
            h01.append(np.matmul(xb.T, s1))
            h10.append(np.matmul(ob.T, sx))
            h11.append(np.matmul(ob.T, s1))
        H = np.block([[tree_sum(h00), tree_sum(h01)], [tree_sum(h10), tree_sum(h11)]])
        if self.l2vec is not None and not self.skip_penalty:
            H = H + np.diag(self.l2vec)
        return H


def newton(model, tol=1e-4, max_iter=100, trace=None):
    """glms.py:660-676.  Returns (beta, iterations)."""
    beta = np.zeros(model.p)
    d = model.d
    it = 0
    for _ in range(max_iter):
        fwd = model.forward(beta)
        g = model.gradient(beta, fwd)
        H = model.hessian(beta, fwd)
        if trace is not None:
            trace.append({"beta": beta.copy(), "g": g.copy(), "H": H.copy(), "obj": model.objective(beta, fwd)})
        ninv = -np.linalg.inv(H)                       # linalg.py:246-261 + blockarray __neg__
        # blocked (-inv(H)) @ g with column blocks (d, 1): two partial products per row block, summed
        top = np.matmul(ninv[:d, :d], g[:d]) + np.matmul(ninv[:d, d:], g[d:])
        bot = np.matmul(ninv[d:, :d], g[:d]) + np.matmul(ninv[d:, d:], g[d:])
        beta = beta + np.concatenate([top, bot])
        it += 1
        if np.max(np.abs(g)) <= tol:                   # glms.py:674 (g from before the update)
            break
    return beta, it


def gd(model, tol=1e-4, max_iter=100, lr=0.01):
    """glms.py:642-657"""
    beta = np.zeros(model.p)
    it = 0
    for _ in range(max_iter):
        g = model.gradient(beta)
        beta = beta + (-lr * g)
        it += 1
        if np.max(np.abs(g)) <= tol:
            break
    return beta, it


def sgd(model, X, y, tol=1e-4, max_iter=100, lr=0.01, seed=None):
    """glms.py:595-616; the row index comes from model.rs.numpy().integers(n): a fresh block generator per
    step (random.py:32-36), i.e. PCG64(seed).jumped(step)."""
    from .np_kernel import block_rng
    beta = np.zeros(model.p)
    n = X.shape[0]
    it = 0
    for step in range(max_iter):
        idx = int(block_rng(seed, step).integers(n))
        one = BlockedGLM(model.fam.name, [X[idx:idx + 1].astype(float)], [y[idx:idx + 1]], model.penalty, 1.0)
        one.l1vec, one.l2vec, one.l1, one.l2 = model.l1vec, model.l2vec, model.l1, model.l2
        g = one.gradient(beta)
        beta = beta + (-lr * g)
        it += 1
        if np.max(np.abs(g)) <= tol:
            break
    return beta, it


def block_sgd(model, tol=1e-4, max_iter=100, lr=0.01):
    """glms.py:619-639 (the convergence test leaves only the inner loop)."""
    beta = np.zeros(model.p)
    it = 0
    for _ in range(max_iter):
        for xb, yb in zip(model.Xb, model.yb):
            one = BlockedGLM(model.fam.name, [xb], [yb], model.penalty, 1.0)
            one.l1vec, one.l2vec, one.l1, one.l2 = model.l1vec, model.l2vec, model.l1, model.l2
            g = one.gradient(beta)
            beta = beta + (-lr * g)
            if np.max(np.abs(g)) <= tol:
                break
        it += 1
    return beta, it


def irls(model, tol=1e-4, max_iter=100):
    """glms.py:679-700 (logistic only), on the blocked arrays incl. the p x n temporaries it builds."""
    beta = np.zeros(model.p)
    d = model.d
    it = 0
    for _ in range(max_iter):
        etas, mus = model.forward(beta)
        h00, h01, h10, h11, t0, t1 = [], [], [], [], [], []
        for xb, ob, eta, mu, yb in zip(model.Xb, model.ones, etas, mus, model.yb):
            s = mu * (1 - mu) + 1e-16
            xs, os_ = xb.T * s, ob.T * s                  # X.T(defer) * s, per column block
            z = eta + (yb - mu) / s
            h00.append(np.matmul(xs, xb)); h01.append(np.matmul(xs, ob))
            h10.append(np.matmul(os_, xb)); h11.append(np.matmul(os_, ob))
            t0.append((xs, z)); t1.append((os_, z))
        H = np.block([[tree_sum(h00), tree_sum(h01)], [tree_sum(h10), tree_sum(h11)]])
        Hinv = np.linalg.inv(H)
        # (XTsX_inv @ XT_s) @ z with XT_s blocked (2 x R): form the (p x rb) products per row block first
        acc = []
        for (xs, z), (os_, _) in zip(t0, t1):
            m_top = np.matmul(Hinv[:, :d], xs) + np.matmul(Hinv[:, d:], os_)      # (p, rb)
            acc.append(np.matmul(m_top, z))
        beta = tree_sum(acc)
        g = model.gradient(beta, (etas, mus))
        it += 1
        if np.max(np.abs(g)) <= tol:
            break
    return beta, it


def lbfgs(model, tol=1e-4, max_iter=100, m=10):
    """lbfgs.py:102-159 with its line search :37-52 and memory :55-62."""
    theta = np.zeros(model.p)
    memory = [None] * m
    k = 0
    ident = np.eye(model.p)

    def f(th):
        return model.objective(th, model.forward(th))

    def get_p(H, g):
        q = g
        fw = []
        for i in range(-1, -m - 1, -1):
            mem = memory[i]
            if mem is None:
                break
            alpha = np.matmul(mem["rho"] * mem["s"], q)
            q = q - alpha * mem["y"]
            fw.insert(0, (alpha, mem))
        r = np.matmul(H, q)
        for alpha, mem in fw:
            b = np.matmul(mem["rho"] * mem["y"], r)
            r = r + mem["s"] * (alpha - b)
        return r

    g = model.gradient(theta)
    it = 0
    while k < max_iter:
        H = ident if k == 0 else memory[-1]["gamma"] * ident
        p = -get_p(H, g)
        # BackTrackingLineSearch.execute(rho=1e-2, init_alpha=1, c=1e-4, min_alpha=1e-30)
        alpha = 1.0
        f_val = f(theta)
        f_next = f(theta + alpha * p)
        slope = np.matmul(g, p)
        while np.isnan(f_next) or f_next > f_val + 1e-4 * alpha * slope:
            alpha *= 1e-2
            if alpha < 1e-30:
                alpha = 1e-30
                break
            f_next = f(theta + alpha * p)
        alpha = max(alpha, 1e-30)
        next_theta = theta + alpha * p
        it += 1
        if k + 1 >= max_iter:
            theta = next_theta
            break
        next_g = model.gradient(next_theta)
        s, yv = next_theta - theta, next_g - g
        ys = np.matmul(s, yv)
        memory.append({"k": k, "s": s, "y": yv, "rho": 1.0 / (ys + 1e-30),
                       "gamma": ys / (np.matmul(yv, yv) + 1e-30)})
        memory.pop(0)
        k += 1
        theta, g = next_theta, next_g
        if np.max(np.abs(g)) <= tol:
            break
    return theta, it


def fit(family, X, y, row_block, solver="newton", penalty=None, lam=1.0, tol=1e-4, max_iter=100, lr=0.01,
        trace=None, l1_ratio=0.5, seed=None):
    """GLM.fit (glms.py:115-192) on host arrays cut into row blocks of `row_block` rows."""
    if np.issubdtype(X.dtype, np.integer):
        X = X.astype(float)
    model = BlockedGLM(family, split_rows(X, row_block), split_rows(y, row_block), penalty, lam, l1_ratio)
    if solver == "sgd":
        return sgd(model, X, y, tol, max_iter, lr, seed)
    if solver in ("newton", "newton-cg"):
        return newton(model, tol, max_iter, trace)
    if solver == "lbfgs":
        return lbfgs(model, tol, max_iter)
    if solver == "gd":
        return gd(model, tol, max_iter, lr)
    if solver == "block_sgd":
        return block_sgd(model, tol, max_iter, lr)
    if solver == "irls":
        return irls(model, tol, max_iter)
    raise ValueError(solver)
