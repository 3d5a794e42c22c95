"""Application layer for the hot path: wiring (`make_app`, after tests/conftest.py:119-158 of the
reference) and the handful of `ArrayApplication` entry points the path uses
(nums/core/array/application.py: array :144, zeros/ones :190-225, concatenate :271-335 (block-aligned
case), reduce :499, argop :542-566, where :588-662, map_uop :872) plus the block RNG front end
(nums/core/array/random.py:158-213).
"""
import numpy as np

from . import cuda_kernel
from .blockarray import BlockArray, Block
from .cuda_backend import CudaBackend
from .grid import ArrayGrid, CyclicDeviceGrid
from .kernel_manager import KernelManager


class NumsRandomState:
    """Block RNG front end (random.py): block j of the process draws from PCG64(seed).jumped(j)."""

    def __init__(self, km, seed=1337):
        self._km = km
        self._rng = km.get_rng(seed)

    def numpy(self):
        """A host Generator on the next block stream (random.py:32-36); used by sgd to pick rows."""
        return cuda_kernel.block_rng(*self._rng.new_block_rng_params())

    def _sample(self, rfunc_name, shape, block_shape, dtype, rfunc_args):
        dtype = np.dtype(np.float64 if dtype is None else dtype)
        grid = ArrayGrid(shape, block_shape, dtype.name)
        ba = BlockArray(grid, self._km)
        for entry in grid.get_entry_iterator():
            params = list(self._rng.new_block_rng_params())
            bshape = grid.get_block_shape(entry)
            size = int(np.prod(bshape))
            if rfunc_name == "random":
                args = (size,) + tuple(rfunc_args)
            elif rfunc_name == "integers":
                args = tuple(rfunc_args[:2]) + (size,) + tuple(rfunc_args[2:])
            else:
                args = tuple(rfunc_args) + (size,)
            ba.blocks[entry].oid = self._km.random_block(
                params, rfunc_name, args, bshape, dtype.type,
                syskwargs={"grid_entry": entry, "grid_shape": grid.grid_shape})
        return ba

    def random(self, shape, block_shape, dtype=None):
        return self._sample("random", shape, block_shape, dtype, ())

    def normal(self, loc=0.0, scale=1.0, shape=None, block_shape=None, dtype=None):
        return self._sample("normal", shape, block_shape, dtype, (loc, scale))

    def integers(self, low, high, shape, block_shape, dtype=np.int64, endpoint=False):
        return self._sample("integers", shape, block_shape, dtype, (low, high, np.dtype(dtype).type, endpoint))

    def poisson(self, lam, shape, block_shape, dtype=None):
        return self._sample("poisson", shape, block_shape, np.int64 if dtype is None else dtype, (lam,))

    def permutation(self, size, block_size):
        # ONE block drawn by rng.permutation(size), then re-blocked (random.py:198-213)
        params = list(self._rng.new_block_rng_params())
        oid = self._km.permutation(params, size, syskwargs={"grid_entry": (0,), "grid_shape": (1,)})
        whole = BlockArray.from_oid(oid, (size,), np.int64, self._km)
        if block_size >= size:
            return whole
        grid = ArrayGrid((size,), (block_size,), "int64")
        out = BlockArray(grid, self._km)
        for entry in grid.get_entry_iterator():
            lo, hi = grid.get_slice_tuples(entry)[0]
            out.blocks[entry].oid = self._km.create_block(
                oid, src_params=[((slice(lo, hi),), False)], dst_params=[((slice(0, hi - lo),), False)],
                dst_shape=(hi - lo,), dst_shape_bc=None,
                syskwargs={"grid_entry": entry, "grid_shape": grid.grid_shape})
        return out


class ArrayApplication:
    def __init__(self, km):
        self.km = km
        self.random = NumsRandomState(km, 1337)
        self.one = self.scalar(1.0)
        self.two = self.scalar(2.0)
        self.zero = self.scalar(0.0)
        self.one_half = self.scalar(0.5)

    # ---- creation
    def scalar(self, value):
        return BlockArray.from_scalar(value, self.km)

    def array(self, arr, block_shape=None):
        arr = np.asarray(arr)
        if block_shape is None:
            block_shape = self.km.get_block_shape(arr.shape, arr.dtype)
        return BlockArray.from_np(arr, block_shape, self.km)

    def zeros(self, shape, block_shape, dtype=np.float64):
        return BlockArray.create("zeros", shape, block_shape, dtype, self.km)

    def ones(self, shape, block_shape, dtype=np.float64):
        return BlockArray.create("ones", shape, block_shape, dtype, self.km)

    def eye(self, shape, block_shape, dtype=np.float64):
        return BlockArray.create("eye", shape, block_shape, dtype, self.km)

    def _concatenate_general(self, arrays, axis, axis_block_size):
        """application.py:271-335 for operands whose extents are not multiples of the block size (read_csv parts):
        every destination block gathers the rectangles of the source blocks it overlaps (create_block)."""
        import itertools
        first = arrays[0]
        nd = first.ndim
        shape = list(first.shape)
        shape[axis] = sum(a.shape[axis] for a in arrays)
        block_shape = list(first.block_shape)
        block_shape[axis] = axis_block_size
        for ax in range(nd):
            if ax != axis:
                block_shape[ax] = first.shape[ax]      # off-axis: one block (re-blocked by the caller if needed)
        grid = ArrayGrid(tuple(shape), tuple(block_shape), first.grid.dtype)
        out = BlockArray(grid, self.km)
        # source pieces along the axis: (array, block index along axis, global start, stop)
        pieces, pos = [], 0
        for a in arrays:
            for g in range(a.grid.grid_shape[axis]):
                lo = pos + g * a.block_shape[axis]
                hi = min(pos + (g + 1) * a.block_shape[axis], pos + a.shape[axis])
                pieces.append((a, g, lo, hi))
            pos += a.shape[axis]
        for entry in grid.get_entry_iterator():
            d_lo = entry[axis] * block_shape[axis]
            d_hi = min(d_lo + block_shape[axis], shape[axis])
            oids, src_params, dst_params = [], [], []
            for a, g, lo, hi in pieces:
                s, e = max(lo, d_lo), min(hi, d_hi)
                if s >= e:
                    continue
                # off-axis: the source may be blocked, the destination is not
                off_ranges = []
                for ax in range(nd):
                    if ax == axis:
                        off_ranges.append([g])
                    else:
                        off_ranges.append(list(range(a.grid.grid_shape[ax])))
                for src_entry in itertools.product(*off_ranges):
                    sb = a.blocks[src_entry]
                    src_sel, dst_sel = [], []
                    for ax in range(nd):
                        if ax == axis:
                            src_sel.append(slice(s - lo, e - lo))
                            dst_sel.append(slice(s - d_lo, e - d_lo))
                        else:
                            o = src_entry[ax] * a.block_shape[ax]
                            w = min(a.block_shape[ax], a.shape[ax] - o)
                            src_sel.append(slice(0, w))
                            dst_sel.append(slice(o, o + w))
                    oids.append(sb.oid)
                    src_params.append((tuple(src_sel), sb.transposed))
                    dst_params.append((tuple(dst_sel), False))
            db = out.blocks[entry]
            db.oid = self.km.create_block(*oids, src_params=src_params, dst_params=dst_params, dst_shape=db.shape,
                                          dst_shape_bc=None, syskwargs=db.syskwargs())
        return out

    def concatenate(self, arrays, axis, axis_block_size=None):
        """application.py:271-335.  Block-aligned operands are referenced, not copied (the zero-copy case, view.py:256);
        anything else is gathered block by block."""
        try:
            return self._concatenate_aligned(arrays, axis, axis_block_size)
        except NotImplementedError:
            first = arrays[0]
            if any(a.ndim != first.ndim or a.dtype != first.dtype for a in arrays):
                raise
            for a in arrays:
                for ax in range(first.ndim):
                    if ax != axis and a.shape[ax] != first.shape[ax]:
                        raise
            bs = first.block_shape[axis] if axis_block_size is None else axis_block_size
            return self._concatenate_general(arrays, axis, bs)

    def _concatenate_aligned(self, arrays, axis, axis_block_size=None):
        first = arrays[0]
        nd = first.ndim
        for a in arrays:
            if a.ndim != nd or a.dtype != first.dtype:
                raise NotImplementedError("concatenate needs equal rank and dtype")
            for ax in range(nd):
                if ax != axis and (a.shape[ax] != first.shape[ax] or a.block_shape[ax] != first.block_shape[ax]):
                    raise NotImplementedError("concatenate needs equal off-axis shapes and block shapes")
        bs = first.block_shape[axis] if axis_block_size is None else axis_block_size
        for a in arrays[:-1]:
            if a.shape[axis] % bs != 0 or (a.block_shape[axis] != bs and a.grid.grid_shape[axis] != 1):
                raise NotImplementedError("only block-aligned concatenation is implemented")
        shape = list(first.shape)
        shape[axis] = sum(a.shape[axis] for a in arrays)
        block_shape = list(first.block_shape)
        block_shape[axis] = bs
        for a in arrays:
            if a.grid.grid_shape[axis] == 1 and a.shape[axis] > bs:
                raise NotImplementedError("only block-aligned concatenation is implemented")
        grid = ArrayGrid(tuple(shape), tuple(block_shape), first.grid.dtype)
        out = BlockArray(grid, self.km)
        offset = 0
        for a in arrays:
            for entry in a.grid.get_entry_iterator():
                dst = list(entry)
                dst[axis] += offset
                src = a.blocks[entry]
                out.blocks[tuple(dst)] = Block(tuple(dst), grid.grid_shape, src.shape, src.dtype, src.transposed,
                                               self.km, src.oid)
            offset += a.grid.grid_shape[axis]
        if offset != grid.grid_shape[axis]:
            raise NotImplementedError("only block-aligned concatenation is implemented")
        return out

    # ---- unary maps / reductions
    def map_uop(self, op_name, arr):
        return arr.ufunc(op_name)

    def exp(self, X): return X.ufunc("exp")
    def log(self, X): return X.ufunc("log")
    def abs(self, X): return X.ufunc("abs")

    def sqrt(self, X):
        if X.dtype.kind != "f":
            X = X.astype(np.float64)
        return X.ufunc("sqrt")

    def reduce(self, op_name, X, axis=None, keepdims=False, dtype=None):
        res = X.reduce_axis(op_name, axis, keepdims=keepdims)
        return res if dtype is None else res.astype(dtype)

    def sum(self, X, axis=None, keepdims=False, dtype=None): return self.reduce("sum", X, axis, keepdims, dtype)
    def max(self, X, axis=None, keepdims=False): return self.reduce("max", X, axis, keepdims)
    def min(self, X, axis=None, keepdims=False): return self.reduce("min", X, axis, keepdims)

    def mean(self, X, axis=None, keepdims=False):
        if X.dtype.kind != "f":
            X = X.astype(np.float64)
        count = int(np.prod(X.shape)) if axis is None else X.shape[axis]
        return self.sum(X, axis=axis, keepdims=keepdims) / count

    # ---- ingest (application.py:72-78, 172-190)
    @property
    def fs(self):
        if getattr(self, "_fs", None) is None:
            from .filesystem import FileSystem
            self._fs = FileSystem(self.km)
        return self._fs

    def read_csv(self, filename, dtype=float, delimiter=",", has_header=False, num_workers=None):
        if num_workers is None:
            num_workers = self.km.num_cores_total()
        arrays = self.fs.read_csv(filename, dtype, delimiter, has_header, num_workers)
        shape = tuple(int(v) for v in np.sum([a.shape for a in arrays], axis=0))
        shape = (shape[0],) + tuple(arrays[0].shape[1:])
        out_dtype = np.int64 if dtype is int else dtype
        block_shape = self.km.get_block_shape(shape, out_dtype)
        result = self.concatenate(arrays, axis=0, axis_block_size=block_shape[0])
        del arrays
        if len(block_shape) > 1 and result.block_shape[1] != block_shape[1]:
            result = result.reshape(block_shape=block_shape)
        return result

    def write_fs(self, ba, filename):
        return self.fs.write_fs(ba, filename)

    def read_fs(self, filename):
        return self.fs.read_fs(filename)

    def delete_fs(self, filename):
        return self.fs.delete_fs(filename)

    def cov(self, X, rowvar=True, bias=False, dtype=None):
        # application.py:531-540
        n = X.shape[1] if rowvar else X.shape[0]
        const = n if bias else n - 1
        Xc = X - self.mean(X, axis=(1 if rowvar else 0), keepdims=True)
        r = (Xc @ Xc.transpose(defer=True) if rowvar else Xc.transpose(defer=True) @ Xc) / const
        return r if dtype is None else r.astype(dtype)

    def vec_from_oids(self, oids, shape, block_shape, dtype):
        # application.py:1168-1180: blocks handed over as a vector of object ids, then re-blocked
        shape, block_shape = tuple(shape), tuple(block_shape)
        arr = BlockArray(ArrayGrid(shape=shape, block_shape=shape, dtype=np.dtype(dtype).name), self.km)
        for i, entry in enumerate(arr.grid.get_entry_iterator()):
            arr.blocks[entry].oid = oids[i]
        return arr if block_shape == shape else arr.reshape(block_shape=block_shape)

    def norm(self, X, order=2):
        assert X.ndim == 1 and order in (1, 2)
        if order == 2:
            return self.sqrt(X.transpose(defer=True) @ X)
        return self.sum(self.abs(X))

    def isnan(self, X): return X.ufunc("isnan")

    def argop(self, op_name, arr, axis=None):
        # application.py:542-566: sequential chain over the blocks of a 1-d array
        if arr.ndim > 1:
            raise NotImplementedError("%s currently supports one-dimensional arrays." % op_name)
        result = BlockArray(ArrayGrid((), (), "int64"), self.km)
        red = (None, None)
        for entry in arr.grid.get_entry_iterator():
            block_slice = arr.grid.get_slice(entry)[0]
            red = self.km.arg_op(op_name, arr.blocks[entry].oid, block_slice, *red,
                                 syskwargs={"grid_entry": entry, "grid_shape": arr.grid.grid_shape,
                                            "options": {"num_returns": 2}})
        result.blocks[()].oid = red[0]
        return result

    def argmax(self, arr): return self.argop("argmax", arr)
    def argmin(self, arr): return self.argop("argmin", arr)

    def where(self, condition, x=None, y=None):
        # application.py:588-662
        if x is not None or y is not None:
            assert x is not None and y is not None
            assert condition.shape == x.shape == y.shape
            assert condition.block_shape == x.block_shape == y.block_shape
            out = BlockArray(x.grid.copy(), self.km)
            for entry in condition.grid.get_entry_iterator():
                out.blocks[entry].oid = self.km.where(
                    condition.blocks[entry].oid, x.blocks[entry].oid, y.blocks[entry].oid, None,
                    syskwargs={"grid_entry": entry, "grid_shape": condition.grid.grid_shape,
                               "options": {"num_returns": 1}})
            return out
        num_axes = max(1, condition.ndim)
        per_block = []
        for entry in condition.grid.get_entry_iterator():
            res = self.km.where(condition.blocks[entry].oid, None, None, condition.grid.get_slice_tuples(entry),
                                syskwargs={"grid_entry": entry, "grid_shape": condition.grid.grid_shape,
                                           "options": {"num_returns": num_axes + 1}})
            per_block.append((res[:-1], res[-1]))
        shapes = self.km.get([s for _, s in per_block])
        total = int(sum(int(s[0]) for s in shapes))
        if total == 0:
            return tuple(self.array(np.array([], dtype=np.int64), block_shape=(0,)) for _ in range(num_axes))
        results = []
        for axis in range(num_axes):
            parts = [BlockArray.from_oid(oids[axis], tuple(int(v) for v in shp), np.int64, self.km)
                     for (oids, _), shp in zip(per_block, shapes) if int(shp[0]) > 0]
            results.append(parts[0] if len(parts) == 1 else self._concat_ragged(parts))
        return tuple(results)

    def _concat_ragged(self, parts):
        """Ragged 1-d concatenation on the device: one create_block gathers every part, in order, into a
        single block (the reference re-blocks through concatenate; values and order are the same)."""
        total = sum(p.shape[0] for p in parts)
        src_params, dst_params, oids, pos = [], [], [], 0
        for p in parts:
            n = p.shape[0]
            oids.append(p.blocks[(0,)].oid)
            src_params.append(((slice(0, n),), False))
            dst_params.append(((slice(pos, pos + n),), False))
            pos += n
        oid = self.km.create_block(*oids, src_params=src_params, dst_params=dst_params, dst_shape=(total,),
                                   dst_shape_bc=None, syskwargs={"grid_entry": (0,), "grid_shape": (1,)})
        return BlockArray.from_oid(oid, (total,), np.int64, self.km)


def make_app(num_gpus=None, cluster_shape=None, device_ids=None):
    """Backend -> device grid -> kernel manager -> application, by hand, like the reference's test
    harness (tests/conftest.py:145-158), with GPUs as the devices of a (G, 1) cyclic grid.
    One-process-per-GPU programs (torchrun) pass device_ids=[LOCAL_RANK]."""
    backend = CudaBackend(num_gpus, device_ids=device_ids)
    backend.init()
    g = backend.num_gpus
    grid = CyclicDeviceGrid((g, 1) if cluster_shape is None else cluster_shape, "gpu", backend.devices())
    km = KernelManager(backend, cuda_kernel, grid)
    return ArrayApplication(km)


_instance = None


def instance():
    """Lazily created process-wide application (cf. nums/core/application_manager.py:56-63)."""
    global _instance
    if _instance is None:
        _instance = make_app()
    return _instance


def set_instance(app):
    global _instance
    _instance = app
    return app
