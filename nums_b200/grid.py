"""Block geometry and static device placement.

Mirrors the behaviour (not the code) of nums/core/grid/grid.py: `ArrayGrid` :27-106 cuts an array
into ceil(shape/block_shape) blocks; `Device` :109-114 names one executor; `CyclicDeviceGrid`
:160-188 places block entry e on cluster cell e % cluster_shape.  With cluster shape (G, 1) and G
row blocks this *is* static row-shard placement: block i lives on GPU i.
"""
import itertools
from dataclasses import dataclass

import numpy as np


@dataclass(frozen=True)
class Device:
    node_id: int
    node_addr: str
    device_type: str
    device: int


def gpu_devices(num_gpus):
    return [Device(0, "localhost", "gpu", g) for g in range(num_gpus)]


class ArrayGrid:
    def __init__(self, shape, block_shape, dtype):
        self.shape = tuple(int(s) for s in shape)
        self.block_shape = tuple(int(b) for b in block_shape)
        self.dtype = dtype if isinstance(dtype, str) else np.dtype(dtype).name
        if len(self.shape) != len(self.block_shape):
            raise ValueError("shape %s and block_shape %s differ in rank" % (self.shape, self.block_shape))
        gs = []
        for s, b in zip(self.shape, self.block_shape):
            if s == 0:
                gs.append(1)
            else:
                if b <= 0:
                    raise ValueError("block_shape entries must be positive")
                gs.append(-(-s // b))
        self.grid_shape = tuple(gs)

    @classmethod
    def from_meta(cls, meta):
        return cls(meta["shape"], meta["block_shape"], meta["dtype"])

    def to_meta(self):
        return {"shape": self.shape, "block_shape": self.block_shape, "dtype": self.dtype}

    def copy(self):
        return ArrayGrid(self.shape, self.block_shape, self.dtype)

    def np_dtype(self):
        name = {"bool": "bool_", "float": "float64", "int": "int64"}.get(self.dtype, self.dtype)
        return np.dtype(getattr(np, name))

    def get_entry_iterator(self):
        return itertools.product(*(range(g) for g in self.grid_shape))

    def get_slice_tuples(self, entry):
        out = []
        for e, b, s in zip(entry, self.block_shape, self.shape):
            lo = e * b
            out.append((lo, min(s, lo + b)))
        return out

    def get_slice(self, entry):
        return tuple(slice(lo, hi) for lo, hi in self.get_slice_tuples(entry))

    def get_block_shape(self, entry):
        return tuple(hi - lo for lo, hi in self.get_slice_tuples(entry))


class CyclicDeviceGrid:
    """entry -> device by entry[axis] % cluster_shape[axis]; axes beyond the cluster rank map to 0."""

    def __init__(self, grid_shape, device_type, devices):
        self.grid_shape = tuple(grid_shape)
        self.device_type = device_type
        ordered = sorted(devices, key=lambda d: (d.device, d.node_id))
        cells = list(itertools.product(*(range(g) for g in self.grid_shape)))
        if len(ordered) < len(cells):
            raise ValueError("cluster shape %s needs %d devices, got %d" % (self.grid_shape, len(cells), len(ordered)))
        self.device_grid = np.empty(self.grid_shape, dtype=object)
        for cell, dev in zip(cells, ordered):
            self.device_grid[cell] = dev

    def get_cluster_entry(self, agrid_entry, agrid_shape=None):
        cell = []
        for axis, dim in enumerate(self.grid_shape):
            cell.append(agrid_entry[axis] % dim if axis < len(agrid_entry) else 0)
        return tuple(cell)

    def get_device(self, agrid_entry, agrid_shape=None):
        return self.device_grid[self.get_cluster_entry(agrid_entry, agrid_shape)]
