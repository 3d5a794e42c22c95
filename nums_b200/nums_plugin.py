"""Plugging the B200 backend into NumS itself (the unmodified `nums` package).

`install()` is the hand wiring of the reference's tests/conftest.py:145-158 with this repository's
Backend + kernel module: after it, unchanged `nums.numpy` / `nums.models.glms` programs run with
every block resident in HBM.  `enable_fused_glm()` additionally routes `nums.models.glms.newton`
(glms.py:660-676) to the fused statistics pass: NumS keeps GLM.fit, the penalty set-up and the
beta split; the loop body becomes one `glm_stats` per row block, the reference's own FIFO
`tree_reduce`, and one `glm_newton_update`.

Nothing here runs on the host: a model or an X layout the fused operators do not cover is handed
back to NumS's own `newton`, i.e. to the generic block kernels on the device.
"""
import numpy as np

from . import fused as _fused
from . import device as dv
from .device import DeviceBlock

_FAMILY = {"LogisticRegression": "logistic", "LinearRegression": "linear", "LinearRegressionBase": "linear",
           "Ridge": "linear", "PoissonRegression": "poisson"}

_state = {"newton": None}


def install(num_gpus=None, device_ids=None, fused_glm=True):
    """Builds NumS's ArrayApplication on CudaBackend + cuda_kernel + CyclicDeviceGrid((G, 1), "gpu", ...)
    and makes it the process-wide instance (application_manager.py:39).  Call before any nps.* /
    nums.models use; returns the application."""
    from nums.core import application_manager
    from nums.core.array.application import ArrayApplication
    from nums.core.filesystem import FileSystem
    from nums.core.grid.grid import CyclicDeviceGrid
    from nums.core.kernel.kernel_manager import KernelManager
    from . import cuda_kernel
    from .cuda_backend import CudaBackend

    backend = CudaBackend(num_gpus=num_gpus, device_ids=device_ids)
    backend.init()
    grid = CyclicDeviceGrid((backend.num_gpus, 1), "gpu", backend.devices())
    km = KernelManager(backend, cuda_kernel, grid)
    app = ArrayApplication(km, FileSystem(km))
    application_manager._instance = app
    if fused_glm:
        enable_fused_glm()
    return app


def pack_vector(*parts):
    """The blocks of a 1-d BlockArray gathered into one contiguous float64 block."""
    dev = parts[0].device_index
    total = sum(p.size for p in parts)
    out = dv.empty((total,), np.float64, dev)
    lo = 0
    for p in parts:
        p = dv.to_device(p, dev)
        dv.copy_into(out.t[lo:lo + p.size], p.t.reshape(-1))
        lo += p.size
    return out


def vector_part(vec, lo, hi):
    return DeviceBlock(vec.t[lo:hi])


def _register(km):
    _fused.register(km)
    km.register("pack_vector", pack_vector, {})
    km.register("vector_part", vector_part, {})


def _family_of(model):
    for cls in type(model).__mro__:
        if cls.__name__ in _FAMILY:
            return _FAMILY[cls.__name__]
    return None


def _fusable(model, X, y, beta):
    """X must be [X_i | ones] as GLM.fit builds it (glms.py:124-135): grid (R, 2), second column block one wide."""
    if _family_of(model) is None or X.ndim != 2 or y.ndim != 1:
        return False
    if X.grid.grid_shape[1] != 2 or X.shape[1] - X.block_shape[1] != 1:
        return False
    if X.dtype != np.float64 or y.dtype not in (np.float64, np.int64, float, int):
        return False
    if X.grid.grid_shape[0] != y.grid.grid_shape[0] or X.block_shape[0] != y.block_shape[0]:
        return False
    if any(X.blocks[i, 0].transposed for i in range(X.grid.grid_shape[0])):
        return False
    return beta.grid.grid_shape == (2,)


def fused_newton(app, model, beta, X, y, tol, max_iter):
    """Replacement for nums.models.glms.newton with its signature and semantics: plain Newton,
    convergence tested on the gradient from before the update (glms.py:674)."""
    if not _fusable(model, X, y, beta):
        return _state["newton"](app, model, beta, X, y, tol, max_iter)
    from nums.core.array.blockarray import BlockArray
    km = app.km
    _register(km)
    family = _family_of(model)
    p = X.shape[1]
    d = p - 1
    R = X.grid.grid_shape[0]
    home = {"grid_entry": (0,), "grid_shape": (1,)}
    poisson = family == "poisson"        # its gradient/hessian carry no penalty terms (glms.py:539-548)

    def packed(vec):
        return km.call("pack_vector", *[vec.blocks[i].oid for i in range(vec.grid.grid_shape[0])], syskwargs=home)

    l2 = packed(model._l2penalty_vec) if (model._l2penalty_vec is not None and not poisson) else None
    l1 = model._l1penalty_vec if (model._l1penalty_vec is not None and not poisson) else None
    b = packed(beta)
    tol_value = float(tol.get()) if hasattr(tol, "get") else float(tol)
    for _ in range(max_iter):
        parts = []
        for i in range(R):
            xb, yb = X.blocks[i, 0], y.blocks[i]
            oid = km.call("glm_stats", xb.oid, yb.oid, b, family, True, False,
                          syskwargs={"grid_entry": xb.grid_entry, "grid_shape": xb.grid_shape})
            parts.append((oid, xb.grid_entry, xb.grid_shape, False))
        stats = X.tree_reduce("sum", parts, (0,), (1,))
        if l1 is not None:
            extra = l1 * app.map_uop("sign", beta)          # L1 term of grad_penalty (glms.py:254-264)
            stats = km.call("stats_add_g", stats, packed(extra), syskwargs=home)
        b, verdict = km.call("glm_newton_step", stats, l2, b, syskwargs={**home, "options": {"num_returns": 2}})
        if l1 is not None:
            beta = _unpack(BlockArray, km, beta, b, d)
        gm, inf = _fused.read_verdict(km.get(verdict))      # the one host sync per iteration
        if inf != 0:
            if gm == gm:
                raise np.linalg.LinAlgError("Singular matrix")
            # non-finite g / H: the reference's inverse returns NaNs and its loop runs on to max_iter with beta = NaN
            b = km.put(np.full(p, np.nan), syskwargs=home)
            break
        if gm <= tol_value:
            break
    return _unpack(BlockArray, km, beta, b, d)


def _unpack(BlockArray, km, like, vec, d):
    """The packed (p,) vector as a BlockArray with the grid of `like` (blocks (d,), (1,))."""
    out = BlockArray(like.grid.copy(), km)
    home = {"grid_entry": (0,), "grid_shape": (1,)}
    lo = 0
    for i in range(like.grid.grid_shape[0]):
        hi = lo + like.grid.get_block_shape((i,))[0]
        out.blocks[i].oid = km.call("vector_part", vec, lo, hi, syskwargs=home)
        lo = hi
    return out


def enable_fused_glm():
    """Routes nums.models.glms.newton to `fused_newton` (the module global GLM.fit looks up, glms.py:181)."""
    import nums.models.glms as ref_glms
    if _state["newton"] is None:
        _state["newton"] = ref_glms.newton
    ref_glms.newton = fused_newton


def disable_fused_glm():
    import nums.models.glms as ref_glms
    if _state["newton"] is not None:
        ref_glms.newton = _state["newton"]
