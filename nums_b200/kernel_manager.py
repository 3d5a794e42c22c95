"""Host-side dispatch between block arrays and the backend, for environments where the real NumS
`KernelManager` (nums/core/kernel/kernel_manager.py:30-273) is not importable.  Same contract:

  * every public method of the kernel module's `KernelCls` is registered with the backend by name
    (:72-90) and exposed as `km.<name>(*args, syskwargs={...})`;
  * `syskwargs` carries either ("grid_entry", "grid_shape") — resolved to a device by the device
    grid (:133-149) — or an explicit "device", plus optional "options";
  * `get_block_shape` applies the reference's block-shape rule (:176-273) so arrays are cut into the
    same blocks as under the NumPy backend.
"""
import inspect

import numpy as np


class KernelManager:
    def __init__(self, backend, kernel_module, device_grid):
        self.backend = backend
        self.device_grid = device_grid
        self.rng_cls = kernel_module.RNG
        self._block_shape_map = {}
        self._names = set()
        imp = kernel_module.KernelCls()
        for name, fn in inspect.getmembers(imp, predicate=inspect.ismethod):
            if name.startswith("_"):
                continue
            self.backend.register(name, fn, {})
            self._names.add(name)

    def __getattr__(self, name):
        # only reached for names that are not real attributes: the registered kernels
        names = self.__dict__.get("_names", ())
        if name in names:
            return lambda *args, **kwargs: self.call(name, *args, **kwargs)
        raise AttributeError(name)

    def get_rng(self, seed):
        return self.rng_cls(seed)

    # ---- backend interface
    def _resolve(self, syskwargs):
        if "grid_entry" in syskwargs:
            device = self.device_grid.get_device(syskwargs["grid_entry"], syskwargs["grid_shape"])
        elif "device" in syskwargs:
            device = syskwargs["device"]
        else:
            raise Exception("All calls require device or grid_entry and grid_shape.")
        return device, syskwargs.get("options", {})

    def put(self, value, syskwargs):
        device, _ = self._resolve(syskwargs)
        return self.backend.put(value, device)

    def get(self, object_ids):
        return self.backend.get(object_ids)

    def register(self, name, func, remote_params=None):
        self.backend.register(name, func, remote_params)
        self._names.add(name)

    def call(self, name, *args, **kwargs):
        kwargs = dict(kwargs)
        device, options = self._resolve(kwargs.pop("syskwargs"))
        return self.backend.call(name, args, kwargs, device, options)

    def devices(self):
        return self.backend.devices()

    def num_cores_total(self):
        return self.backend.num_cores_total()

    # ---- block shapes
    @staticmethod
    def compute_block_shape_static(shape, dtype, cluster_shape, num_cores):
        """The reference's rule (kernel_manager.py:176-233): arrays under 1e8 bytes stay one block;
        otherwise ~num_cores blocks are laid out with a softmax(cluster_shape) preference per axis,
        the remainder going to the largest axis."""
        shape = tuple(int(s) for s in shape)
        itemsize = np.dtype(dtype).alignment
        if int(np.prod(shape, dtype=np.int64)) * itemsize < 10 ** 8:
            return shape
        cluster = list(cluster_shape[: len(shape)]) + [1] * max(0, len(shape) - len(cluster_shape))
        w = np.exp(np.array(cluster, dtype=float))
        w = w / w.sum()
        frac = np.array(shape, dtype=float) / float(np.sum(shape))
        wf = w * frac
        wf = wf / wf.sum()
        grid_frac = float(num_cores) ** wf
        grid = np.floor(grid_frac)
        grid[int(np.argmax(shape))] += float(np.sum(grid_frac - grid))
        grid = np.ceil(grid).astype(int)
        return tuple(int((s + g - 1) // g) for s, g in zip(shape, grid))

    def compute_block_shape(self, shape, dtype, cluster_shape=None, num_cores=None):
        return self.compute_block_shape_static(
            shape, dtype,
            self.device_grid.grid_shape if cluster_shape is None else cluster_shape,
            self.num_cores_total() if num_cores is None else num_cores)

    def get_block_shape(self, shape, dtype):
        # the first block size seen for a dimension length is reused (kernel_manager.py:253-273)
        block_shape = self.compute_block_shape(shape, dtype)
        out = []
        for dim, bdim in zip(shape, block_shape):
            out.append(self._block_shape_map.setdefault(dim, bdim))
        return tuple(out)
