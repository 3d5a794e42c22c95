"""Imports the reference itself (oracle/_ref, made by oracle/make_ref.py) in this process and wires
applications on it.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): used by tests/, by
tests/golden/make_golden.py and by bench.py's CPU reference arm; never by nums_b200.

What has to be patched is the environment, not the reference: NumPy 2 removed aliases the reference
(pinned to NumPy <= 1.20, setup.cfg) still uses.  They are restored on the `numpy` module before
`import nums`.
"""
import contextlib
import io
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

_state = {"done": False}


def present():
    return os.path.isdir(os.path.join(REF_DIR, "nums"))


def bootstrap():
    """Make `import nums` resolve to oracle/_ref/nums.  Idempotent.  Raises if oracle/_ref is absent."""
    if _state["done"]:
        return REF_DIR
    if not present():
        raise RuntimeError("oracle/_ref is absent: run `python oracle/make_ref.py` where /root/reference exists")
    import numpy as np
    import numpy.ma  # noqa: F401  (import before the aliases are patched in)
    import scipy.linalg  # noqa: F401
    import scipy.special  # noqa: F401
    sys.path.insert(0, os.path.join(REF_DIR, "_stubs"))
    sys.path.insert(0, REF_DIR)
    os.environ.setdefault("NUMS_BACKEND", "serial")
    for name, value in {"int": int, "object": object, "float": float, "bool": bool, "product": np.prod,
                        "NINF": -np.inf, "PZERO": 0.0, "NZERO": -0.0}.items():
        if not hasattr(np, name):
            setattr(np, name, value)
    if "numpy.compat" not in sys.modules or not hasattr(sys.modules["numpy.compat"], "asbytes"):
        compat = types.ModuleType("numpy.compat")
        compat.asbytes = lambda s: s if isinstance(s, bytes) else str(s).encode("latin1")
        compat.asstr = lambda s: s.decode("latin1") if isinstance(s, bytes) else str(s)
        compat.asunicode = compat.asstr
        sys.modules["numpy.compat"] = compat
        np.compat = compat
    np.lib.stride_tricks.broadcast_to = np.broadcast_to
    _state["done"] = True
    return REF_DIR


@contextlib.contextmanager
def _quiet():
    """DeviceGrid.__init__ prints every placement (grid.py:130)."""
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def serial_app(num_devices=1):
    """The reference's own application on its SerialBackend + numpy_kernel, wired like
    tests/conftest.py:145-158 of the reference (KernelManager built directly so that several
    applications can coexist in one process).  num_devices=G presents G cpu devices on a (G, 1)
    cluster, so that the block-shape rule (kernel_manager.py:176-233) cuts arrays exactly as it
    does for G GPUs and both sides sum partial results in the same order."""
    bootstrap()
    from nums.core.array.application import ArrayApplication
    from nums.core.backends import SerialBackend
    from nums.core.filesystem import FileSystem
    from nums.core.grid.grid import CyclicDeviceGrid
    from nums.core.kernel import numpy_kernel
    from nums.core.kernel.kernel_manager import KernelManager
    from nums.core.grid.grid import Device

    class SerialDevices(SerialBackend):
        def devices(self):
            return [Device(0, "localhost", "cpu", g) for g in range(num_devices)]

        def num_cores_total(self):
            return num_devices

    backend = SerialDevices()
    backend.init()
    with _quiet():
        grid = CyclicDeviceGrid((num_devices, 1), "cpu", backend.devices())
    km = KernelManager(backend, numpy_kernel, grid)
    return ArrayApplication(km, FileSystem(km))


def cuda_app(num_gpus=None, device_ids=None, cores_per_gpu=1):
    """The reference's ArrayApplication / KernelManager / FileSystem on this repository's plug-in:
    CudaBackend + cuda_kernel + CyclicDeviceGrid((G, 1), "gpu", ...) — INTEGRATION.md section 2."""
    bootstrap()
    from nums.core.array.application import ArrayApplication
    from nums.core.filesystem import FileSystem
    from nums.core.grid.grid import CyclicDeviceGrid
    from nums.core.kernel.kernel_manager import KernelManager
    from nums_b200 import cuda_kernel
    from nums_b200.cuda_backend import CudaBackend
    backend = CudaBackend(num_gpus=num_gpus, device_ids=device_ids)
    backend.init()
    with _quiet():
        grid = CyclicDeviceGrid((backend.num_gpus, 1), "gpu", backend.devices())
    km = KernelManager(backend, cuda_kernel, grid)
    return ArrayApplication(km, FileSystem(km))


def use(app):
    """Make `app` the process-wide instance that nums.numpy / nums.models look up
    (application_manager.py:39)."""
    from nums.core import application_manager
    application_manager._instance = app
    return app
