"""`CudaBackend`: the NumS `Backend` (nums/core/backends/base.py:20-85) for the GPUs of one box.

Object ids are `DeviceBlock`s resident in HBM (or plain host objects for the non-array values NumS
also `put`s, e.g. subscript tuples, blockarray.py:368).  `call` runs the registered kernel-module
function on the GPU named by `device.device`, first moving any operand that lives on another GPU
with a peer copy over NVLink — NVSwitch makes every pair equally near, which is why the static
placement of `CyclicDeviceGrid` over a (G, 1) cluster replaces the Ray scheduler
(nums/core/backends/ray.py:168-175).  Execution is stream-ordered and asynchronous; `get` is the
synchronisation point, like `ray.get`.

When the real `nums` package is importable this class derives from its `Backend`; otherwise from a
local stand-in with the same 12 methods, so the plug-in also runs where NumS itself is absent.
"""
import os
import sys

import numpy as np

from . import device as dv
from .device import DeviceBlock
from .grid import Device, gpu_devices

try:  # pragma: no cover - depends on the environment
    from nums.core.backends.base import Backend as _BackendBase
    from nums.core.grid.grid import Device as _NumsDevice
except Exception:  # NumS not installed (e.g. on the GPU box)
    _NumsDevice = None

    class _BackendBase:
        """Stand-in with the interface of nums.core.backends.base.Backend."""


_NVTX = bool(os.environ.get("NUMS_B200_NVTX"))


class HostObject:
    """A non-array value stored by `put` (kept on the host, returned as is by `get`)."""

    __slots__ = ("value",)

    def __init__(self, value):
        self.value = value


class CudaBackend(_BackendBase):
    def __init__(self, num_gpus=None, device_ids=None):
        """num_gpus: use CUDA devices 0..num_gpus-1 (default: all visible).
        device_ids: explicit physical CUDA indices; logical device g of the grid is device_ids[g].
        One-process-per-GPU (torchrun) programs pass device_ids=[LOCAL_RANK]."""
        if device_ids is not None:
            num_gpus = len(device_ids)
        self._requested = num_gpus
        self._device_ids = list(device_ids) if device_ids is not None else None
        self.num_gpus = None
        self._remote_functions = {}
        self._actors = {}
        self._devices = None

    # ------------------------------------------------------------------ lifecycle
    def init(self):
        import torch
        dv.require_cuda()
        from . import _lib
        _lib.load()   # fail now, loudly, if the CUDA library is missing
        visible = torch.cuda.device_count()
        self.num_gpus = visible if self._requested is None else int(self._requested)
        if self.num_gpus < 1 or self.num_gpus > visible:
            raise RuntimeError("requested %s GPUs, %d visible" % (self._requested, visible))
        if self._device_ids is None:
            self._device_ids = list(range(self.num_gpus))
        if any(i < 0 or i >= visible for i in self._device_ids):
            raise RuntimeError("device_ids %s out of range (%d visible)" % (self._device_ids, visible))
        dev_cls = _NumsDevice if _NumsDevice is not None else Device
        if dev_cls is Device and "nums.core.grid.grid" in sys.modules:   # NumS imported after this module
            dev_cls = sys.modules["nums.core.grid.grid"].Device
        self._devices = [dev_cls(0, "localhost", "gpu", g) for g in range(self.num_gpus)]

    def shutdown(self):
        self._remote_functions.clear()
        self._actors.clear()

    def devices(self):
        return list(self._devices) if self._devices is not None else gpu_devices(self._requested or 1)

    def num_cores_total(self):
        # The block-shape heuristic (kernel_manager.py:176-233) makes ~num_cores blocks: one per GPU.
        return self.num_gpus if self.num_gpus is not None else (self._requested or 1)

    # ------------------------------------------------------------------ object store
    def physical(self, device):
        """CUDA index of a grid `Device`."""
        return self._device_ids[device.device]

    def put(self, value, device):
        target = self.physical(device)
        if isinstance(value, DeviceBlock):
            return dv.to_device(value, target)
        if isinstance(value, (np.ndarray, np.generic)) and np.asarray(value).dtype.kind in "fiub":
            return dv.from_numpy(value, target)
        if isinstance(value, (bool, int, float)):
            return dv.from_numpy(np.asarray(value), target)
        return HostObject(value)

    def get(self, object_ids):
        if isinstance(object_ids, list):
            return [self.get(o) for o in object_ids]
        if isinstance(object_ids, tuple):
            # a tuple of oids (multi-return) is fetched element-wise; a plain tuple (e.g. the shape `where` returns,
            # application.py:634) is a value and comes back as it is
            if any(isinstance(o, (DeviceBlock, HostObject)) for o in object_ids):
                return tuple(self.get(o) for o in object_ids)
            return object_ids
        if isinstance(object_ids, DeviceBlock):
            arr = dv.to_numpy(object_ids)
            return arr[()] if arr.ndim == 0 else arr
        if isinstance(object_ids, HostObject):
            return object_ids.value
        return object_ids

    # ------------------------------------------------------------------ remote functions
    def remote(self, function, remote_params):
        return function

    def register(self, name, func, remote_params=None):
        if name in self._remote_functions:
            return
        self._remote_functions[name] = self.remote(func, remote_params or {})

    def call(self, name, args, kwargs, device, options):
        import torch
        target = self.physical(device)
        args = [self._localise(a, target) for a in args]
        kwargs = {k: self._localise(v, target) for k, v in kwargs.items()}
        with torch.cuda.device(target):
            if not _NVTX:
                return self._remote_functions[name](*args, **kwargs)
            # NUMS_B200_NVTX=1: one NVTX range per block operation (the tracing hook of SURVEY.md section 5: nsys / ncu
            # --nvtx timelines then show the reference's operation names above the kernels they launch)
            torch.cuda.nvtx.range_push("nums:%s@gpu%d" % (name, target))
            try:
                return self._remote_functions[name](*args, **kwargs)
            finally:
                torch.cuda.nvtx.range_pop()

    @staticmethod
    def _localise(value, target):
        if isinstance(value, DeviceBlock):
            return dv.to_device(value, target)
        if isinstance(value, HostObject):
            return value.value
        return value

    # ------------------------------------------------------------------ actors (host objects)
    def register_actor(self, name, cls):
        if name in self._actors:
            raise AssertionError("actor %s already registered" % name)
        self._actors[name] = cls

    def make_actor(self, name, *args, device=None, **kwargs):
        return self._actors[name](*args, **kwargs)

    def call_actor_method(self, actor, method, *args, **kwargs):
        return getattr(actor, method)(*args, **kwargs)
