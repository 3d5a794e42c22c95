"""L-BFGS with a backtracking line search for the GLM models: what nums/models/lbfgs.py computes (memory of m = 10
pairs, rho = 1e-2, c = 1e-4, alpha0 = 1, alpha_min = 1e-30, the 1e-30 guards), organised for the device.

The optimiser state never leaves HBM.  `DeviceLBFGS` owns a workspace (ring of (s, y) pairs, rho, gamma) and drives
three entry points of libnumsb200 (csrc/lbfgs.cu): the two-loop recursion is ONE launch instead of ~12 block-array
calls per memory slot, a trial point is one launch, and the Armijo test, the memory update and max|g| are one launch
whose 16-byte verdict is the only thing the host reads per trial.  Every evaluation of the objective is one fused pass
over X that also returns the gradient, so the accepted trial's gradient is reused (the reference pays two more passes
per iteration, lbfgs.py:136-138, and one per line search for f(theta), :44).
"""
import numpy as np

from . import _lib
from . import device as dv
from .device import DeviceBlock

MEMORY = 10
SHRINK = 1e-2
ARMIJO_C = 1e-4
ALPHA_MIN = 1e-30


class DeviceLBFGS:
    """Workspace + the three kernels, for contiguous float64 vectors of length p on one device."""

    def __init__(self, p, device_index, m=MEMORY):
        self.p, self.m, self.dev = int(p), int(m), device_index
        self.lib = _lib.load()
        self.ws = dv.scratch(self.lib.nb_lbfgs_ws_bytes(self.p, self.m), device_index)
        self.slope = dv.empty((1,), np.float64, device_index)
        self.verdict = dv.empty((2,), np.float64, device_index)
        _lib.check(self.lib.nb_lbfgs_reset(device_index, self.p, self.m, self.ws.data_ptr(), self.ws.numel(), self._st()))

    def _st(self):
        return dv.stream_ptr(self.dev)

    def direction(self, g):
        d = dv.empty((self.p,), np.float64, self.dev)
        _lib.check(self.lib.nb_lbfgs_direction(self.dev, self.p, self.m, self.ws.data_ptr(), g.ptr, d.ptr, self.slope.ptr,
                                               self._st()))
        return d

    def trial(self, theta, direction, alpha):
        out = dv.empty((self.p,), np.float64, self.dev)
        _lib.check(self.lib.nb_lbfgs_trial(self.dev, self.p, theta.ptr, direction.ptr, float(alpha), out.ptr, self._st()))
        return out

    def judge(self, theta, g, f, theta_t, g_t, f_t, alpha, force, remember):
        """-> (accepted, max|g_t|): the one host read of a trial."""
        _lib.check(self.lib.nb_lbfgs_judge(self.dev, self.p, self.m, self.ws.data_ptr(), theta.ptr, g.ptr, f.ptr, theta_t.ptr,
                                           g_t.ptr, f_t.ptr, self.slope.ptr, float(alpha), ARMIJO_C, int(force), int(remember),
                                           self.verdict.ptr, self._st()))
        v = dv.to_numpy(self.verdict)
        return bool(v[0] != 0.0), float(v[1])


def _packed(vec):
    """A 1-d BlockArray as one contiguous float64 block."""
    from .glms import _gather_single_block
    return dv.contiguous(_gather_single_block(vec))


def _scalar_block(ba):
    """A 0-d BlockArray as a (1,) device block."""
    return DeviceBlock(dv.contiguous(ba.blocks[()].oid).t.reshape(1))


class LBFGS:
    def __init__(self, model, m=MEMORY, max_iter=100, thresh=1e-4):
        self.model, self.m, self.max_iter, self.thresh = model, m, max_iter, thresh

    def _evaluate(self, X, y, theta_block, like):
        """(f, g) at the packed point, as device blocks: one fused pass, or the generic block kernels."""
        model = self.model
        theta = _as_blockarray(theta_block, like)
        if model._fused:
            f, g = model.objective_and_gradient(X, y, theta)
        else:
            mu = model.forward(X, theta)
            f, g = model.objective(X, y, theta, mu), model.gradient(X, y, mu, theta)
        return _scalar_block(f), _packed(g)

    def execute(self, X, y, theta0):
        model = self.model
        theta = _packed(theta0)
        dev = theta.device_index
        opt = DeviceLBFGS(theta.shape[0], dev, self.m)
        f, g = self._evaluate(X, y, theta, theta0)
        model.n_iter_ = 0
        for k in range(self.max_iter):
            last = k + 1 >= self.max_iter
            direction = opt.direction(g)
            alpha = 1.0
            while True:
                trial = opt.trial(theta, direction, alpha)
                f_t, g_t = self._evaluate(X, y, trial, theta0)
                accepted, gmax = opt.judge(theta, g, f, trial, g_t, f_t, alpha, force=False, remember=not last)
                if accepted:
                    break
                alpha *= SHRINK
                if alpha < ALPHA_MIN:
                    # lbfgs.py:49-50: the search gives up and the step alpha_min is taken as it is
                    trial = opt.trial(theta, direction, ALPHA_MIN)
                    f_t, g_t = self._evaluate(X, y, trial, theta0)
                    _, gmax = opt.judge(theta, g, f, trial, g_t, f_t, ALPHA_MIN, force=True, remember=not last)
                    break
            theta, g, f = trial, g_t, f_t
            model.n_iter_ += 1
            if last or gmax <= self.thresh:
                break
        return _as_blockarray(theta, theta0)


def _as_blockarray(block, like):
    """The packed vector as a BlockArray with the grid of `like` (views, no copies)."""
    from .blockarray import BlockArray
    out = BlockArray(like.grid.copy(), like.km)
    for entry in like.grid.get_entry_iterator():
        lo, hi = like.grid.get_slice_tuples(entry)[0]
        out.blocks[entry].oid = DeviceBlock(block.t[lo:hi])
    return out
