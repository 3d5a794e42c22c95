"""A compact host-side mirror of NumS's `BlockArray` for the hot path, used where the `nums` package
itself is not importable (the GPU box) and as the driver of the parity tests.

It reproduces the *scheduling structure* of nums/core/array/blockarray.py and base.py — one kernel
call per block, contractions as per-block `bop("tensordot")` followed by the FIFO pairwise
`tree_reduce` (:643-689), reductions as per-block `reduce_axis` + tree (:691-769), deferred
transposes as a per-block flag (base.py:94-110) — because those decide the floating-point
summation order the oracle is compared with.  Only what sections 8(a) a1-a9 need is here: no
slicing engine, no advanced indexing, no einsum.
"""
import itertools

import numpy as np

from . import hostlogic as hl
from .grid import ArrayGrid


class Block:
    __slots__ = ("grid_entry", "grid_shape", "shape", "dtype", "transposed", "oid", "km")

    def __init__(self, grid_entry, grid_shape, shape, dtype, transposed, km, oid=None):
        self.grid_entry = tuple(grid_entry)
        self.grid_shape = tuple(grid_shape)
        self.shape = tuple(shape)          # logical shape (after the deferred transpose)
        self.dtype = np.dtype(dtype)
        self.transposed = transposed
        self.km = km
        self.oid = oid

    def true_grid_entry(self):
        return tuple(reversed(self.grid_entry)) if self.transposed else self.grid_entry

    def true_grid_shape(self):
        return tuple(reversed(self.grid_shape)) if self.transposed else self.grid_shape

    def syskwargs(self):
        return {"grid_entry": self.true_grid_entry(), "grid_shape": self.true_grid_shape()}

    def transpose(self, defer=True):
        out = Block(tuple(reversed(self.grid_entry)), tuple(reversed(self.grid_shape)), tuple(reversed(self.shape)),
                    self.dtype, not self.transposed, self.km, self.oid)
        if not defer and len(self.shape) > 1:
            out.transposed = False
            if self.transposed:
                # stored data already has the transposed layout
                out.oid = self.oid
            else:
                out.oid = self.km.transpose(self.oid, syskwargs=self.syskwargs())
        elif len(self.shape) <= 1:
            out.transposed = False
        return out


class BlockArray:
    def __init__(self, grid, km, blocks=None):
        self.grid = grid
        self.km = km
        self.shape = grid.shape
        self.block_shape = grid.block_shape
        self.dtype = grid.np_dtype()
        self.ndim = len(self.shape)
        if blocks is None:
            blocks = np.empty(grid.grid_shape, dtype=object)
            for entry in grid.get_entry_iterator():
                blocks[entry] = Block(entry, grid.grid_shape, grid.get_block_shape(entry), self.dtype, False, km)
        self.blocks = blocks

    # ------------------------------------------------------------------ construction / fetch
    @classmethod
    def from_np(cls, arr, block_shape, km):
        arr = np.asarray(arr)
        grid = ArrayGrid(arr.shape, block_shape, arr.dtype.name)
        out = cls(grid, km)
        for entry in grid.get_entry_iterator():
            sel = grid.get_slice(entry)
            out.blocks[entry].oid = km.put(np.ascontiguousarray(arr[sel]),
                                           syskwargs={"grid_entry": entry, "grid_shape": grid.grid_shape})
        return out

    @classmethod
    def from_scalar(cls, value, km):
        return cls.from_np(np.asarray(value), (), km)

    @classmethod
    def from_oid(cls, oid, shape, dtype, km):
        grid = ArrayGrid(shape, shape, np.dtype(dtype).name)
        out = cls(grid, km)
        out.blocks[(0,) * len(shape)].oid = oid
        return out

    @classmethod
    def create(cls, op_name, shape, block_shape, dtype, km):
        grid = ArrayGrid(shape, block_shape, np.dtype(dtype).name)
        out = cls(grid, km)
        for entry in grid.get_entry_iterator():
            if op_name == "eye" and len(set(entry)) > 1:
                out.blocks[entry].oid = km.new_block("zeros", entry, grid.to_meta(),
                                                     syskwargs={"grid_entry": entry, "grid_shape": grid.grid_shape})
            else:
                out.blocks[entry].oid = km.new_block(op_name, entry, grid.to_meta(),
                                                     syskwargs={"grid_entry": entry, "grid_shape": grid.grid_shape})
        return out

    def get(self):
        """Assembles the array on the host (base.py:352-371): the synchronisation point."""
        entries = list(self.grid.get_entry_iterator())
        values = self.km.get([self.blocks[e].oid for e in entries])
        if self.shape == ():
            v = np.asarray(values[0])
            return v.astype(self.dtype)[()] if v.dtype != self.dtype else v[()]
        out = np.empty(self.shape, dtype=self.dtype)
        for entry, val in zip(entries, values):
            val = np.asarray(val)
            if self.blocks[entry].transposed:
                val = val.T
            out[self.grid.get_slice(entry)] = val
        return out

    def touch(self):
        for entry in self.grid.get_entry_iterator():
            b = self.blocks[entry]
            self.km.touch(b.oid, syskwargs=b.syskwargs())
        return self

    def __bool__(self):
        # blockarray.py:1067-1072: defined for single-element arrays (host sync), else True
        if all(s == 1 for s in self.shape):
            return bool(self.get())
        return True

    def __len__(self):
        return self.shape[0]

    # ------------------------------------------------------------------ layout
    def transpose(self, defer=False):
        grid = ArrayGrid(tuple(reversed(self.shape)), tuple(reversed(self.block_shape)), self.grid.dtype)
        blocks = np.empty(grid.grid_shape, dtype=object)
        for entry in self.grid.get_entry_iterator():
            blocks[tuple(reversed(entry))] = self.blocks[entry].transpose(defer=defer)
        return BlockArray(grid, self.km, blocks)

    @property
    def T(self):
        return self.transpose(defer=False)

    def copy(self):
        """base.py:337-345: a new BlockArray over the same (immutable) block objects."""
        out = BlockArray(self.grid.copy(), self.km)
        for entry in self.grid.get_entry_iterator():
            sb, ob = self.blocks[entry], out.blocks[entry]
            ob.oid, ob.transposed = sb.oid, sb.transposed
        return out

    def reshape(self, shape=None, block_shape=None):
        """blockarray.py:1135-1290 (Reshape) for the cases the path needs: vector <-> column reshapes that keep the
        block boundaries (glms.py:495) and re-blocking at constant shape (linalg.py:96-103, vec_from_oids)."""
        shape = self.shape if shape is None else tuple(shape)
        block_shape = self.block_shape if block_shape is None else tuple(block_shape)
        if shape == self.shape and block_shape == self.block_shape:
            return self
        if shape == self.shape:
            return self._reblock(block_shape)
        grid = ArrayGrid(shape, block_shape, self.grid.dtype)
        if int(np.prod(grid.grid_shape)) != int(np.prod(self.grid.grid_shape)):
            raise NotImplementedError("reshape that changes both the shape and the block partition")
        out = BlockArray(grid, self.km)
        for src, dst in zip(self.grid.get_entry_iterator(), grid.get_entry_iterator()):
            sb, db = self.blocks[src], out.blocks[dst]
            if int(np.prod(sb.shape)) != int(np.prod(db.shape)):
                raise NotImplementedError("reshape that changes both the shape and the block partition")
            oid = sb.oid
            if sb.transposed:
                oid = self.km.transpose(oid, syskwargs=sb.syskwargs())
            db.oid = self.km.reshape(oid, db.shape, syskwargs=db.syskwargs())
        return out

    def _reblock(self, block_shape):
        """Same array, new block partition: every destination block gathers the rectangles of the source blocks it
        overlaps with one create_block call (blockarray.py:1245-1251 -> numpy_kernel.py:116-129)."""
        grid = ArrayGrid(self.shape, block_shape, self.grid.dtype)
        out = BlockArray(grid, self.km)
        nd = len(self.shape)

        def overlaps(ax, lo, hi):
            bs = self.block_shape[ax]
            first, last = lo // bs, (hi - 1) // bs
            for g in range(first, last + 1):
                s_lo, s_hi = g * bs, min((g + 1) * bs, self.shape[ax])
                a, b = max(lo, s_lo), min(hi, s_hi)
                yield g, slice(a - s_lo, b - s_lo), slice(a - lo, b - lo)

        for entry in grid.get_entry_iterator():
            dst_slices = grid.get_slice(entry)
            per_axis = [list(overlaps(ax, dst_slices[ax].start, dst_slices[ax].stop)) for ax in range(nd)]
            oids, src_params, dst_params = [], [], []
            for combo in itertools.product(*per_axis):
                src_entry = tuple(c[0] for c in combo)
                sb = self.blocks[src_entry]
                oids.append(sb.oid)
                src_params.append((tuple(c[1] for c in combo), sb.transposed))
                dst_params.append((tuple(c[2] for c in combo), False))
            db = out.blocks[entry]
            db.oid = self.km.create_block(*oids, src_params=src_params, dst_params=dst_params, dst_shape=db.shape,
                                          dst_shape_bc=None, syskwargs=db.syskwargs())
        return out

    def astype(self, dtype):
        dtype = np.dtype(dtype)
        grid = ArrayGrid(self.shape, self.block_shape, dtype.name)
        out = BlockArray(grid, self.km)
        for entry in self.grid.get_entry_iterator():
            b = self.blocks[entry]
            ob = out.blocks[entry]
            ob.transposed = b.transposed
            ob.oid = self.km.astype(b.oid, dtype.name, syskwargs=b.syskwargs())
        return out

    # ------------------------------------------------------------------ elementwise
    def _convert(self, other):
        if isinstance(other, BlockArray):
            return other
        return BlockArray.from_scalar(other, self.km)

    def ufunc(self, op_name):
        code, cdt, odt = hl.resolve_uop(op_name, self.dtype)
        grid = ArrayGrid(self.shape, self.block_shape, odt.name)
        out = BlockArray(grid, self.km)
        for entry in self.grid.get_entry_iterator():
            b = self.blocks[entry]
            ob = out.blocks[entry]
            ob.transposed = b.transposed
            ob.oid = self.km.map_uop(op_name, b.oid, (), {}, syskwargs=b.syskwargs())
        return out

    def __elementwise__(self, op_name, other):
        other = self._convert(other)
        _, _, odt = hl.resolve_bop(op_name, self.dtype, other.dtype)
        shape, _, _ = hl.broadcast_geometry(self.shape, (0,) * self.ndim, other.shape, (0,) * other.ndim)
        nd = len(shape)

        def padded(ba):
            pad = nd - ba.ndim
            return (1,) * pad + ba.shape, (1,) * pad + ba.block_shape, (1,) * pad + ba.grid.grid_shape, pad

        sa, ba_, ga, pa = padded(self)
        sb, bb_, gb, pb = padded(other)
        block_shape = []
        for ax in range(nd):
            if sa[ax] == shape[ax] and sb[ax] == shape[ax] and sa[ax] != 1:
                if ba_[ax] != bb_[ax]:
                    raise NotImplementedError("elementwise ops need equal block shapes along shared axes")
                block_shape.append(ba_[ax])
            elif sa[ax] == shape[ax]:
                block_shape.append(ba_[ax])
            else:
                block_shape.append(bb_[ax])
        grid = ArrayGrid(shape, tuple(block_shape), odt.name)
        out = BlockArray(grid, self.km)
        for entry in grid.get_entry_iterator():
            ea = tuple(entry[ax] if ga[ax] > 1 else 0 for ax in range(nd))[pa:]
            eb = tuple(entry[ax] if gb[ax] > 1 else 0 for ax in range(nd))[pb:]
            blk_a, blk_b = self.blocks[ea], other.blocks[eb]
            out.blocks[entry].oid = self.km.bop(
                op_name, blk_a.oid, blk_b.oid, blk_a.transposed, blk_b.transposed, axes={},
                syskwargs={"grid_entry": entry, "grid_shape": grid.grid_shape})
        return out

    def __neg__(self): return self.ufunc("negative")
    def __abs__(self): return self.ufunc("abs")
    def __invert__(self): return self.ufunc("invert")
    def __add__(self, o): return self.__elementwise__("add", o)
    def __radd__(self, o): return self._convert(o).__elementwise__("add", self)
    def __sub__(self, o): return self.__elementwise__("sub", o)
    def __rsub__(self, o): return self._convert(o).__elementwise__("sub", self)
    def __mul__(self, o): return self.__elementwise__("mul", o)
    def __rmul__(self, o): return self._convert(o).__elementwise__("mul", self)
    def __truediv__(self, o): return self.__elementwise__("truediv", o)
    def __rtruediv__(self, o): return self._convert(o).__elementwise__("truediv", self)
    def __floordiv__(self, o): return self.__elementwise__("floor_divide", o)
    def __rfloordiv__(self, o): return self._convert(o).__elementwise__("floor_divide", self)
    def __mod__(self, o): return self.__elementwise__("mod", o)
    def __rmod__(self, o): return self._convert(o).__elementwise__("mod", self)
    def __pow__(self, o): return self.__elementwise__("pow", o)
    def __rpow__(self, o): return self._convert(o).__elementwise__("pow", self)
    def __and__(self, o): return self.__elementwise__("bitwise_and", o)
    def __or__(self, o): return self.__elementwise__("bitwise_or", o)
    def __xor__(self, o): return self.__elementwise__("bitwise_xor", o)
    def __lshift__(self, o): return self.__elementwise__("left_shift", o)
    def __rshift__(self, o): return self.__elementwise__("right_shift", o)
    def __lt__(self, o): return self.__elementwise__("lt", o)
    def __le__(self, o): return self.__elementwise__("le", o)
    def __gt__(self, o): return self.__elementwise__("gt", o)
    def __ge__(self, o): return self.__elementwise__("ge", o)
    def __eq__(self, o): return self.__elementwise__("eq", o)
    def __ne__(self, o): return self.__elementwise__("ne", o)
    __hash__ = None

    # ------------------------------------------------------------------ reductions
    def tree_reduce(self, op_name, items, result_entry, result_grid_shape):
        """FIFO pairwise reduction (blockarray.py:643-689): pop two, combine on the left operand's
        device, append; the last combine is placed at the result entry."""
        q = list(items)
        if len(q) == 1:
            return q[0][0]
        while len(q) > 1:
            a_oid, a_ge, a_gs, a_t = q.pop(0)
            b_oid, _, _, b_t = q.pop(0)
            ge, gs = (result_entry, result_grid_shape) if not q else (a_ge, a_gs)
            c = self.km.bop_reduce(op_name, a_oid, b_oid, a_t, b_t, syskwargs={"grid_entry": ge, "grid_shape": gs})
            q.append((c, ge, gs, False))
        return q[0][0]

    def reduce_axis(self, op_name, axis, keepdims=False):
        if not (axis is None or isinstance(axis, (int, np.integer))):
            raise NotImplementedError("Only integer axis is currently supported.")
        if axis is not None and axis < 0:
            axis += self.ndim
        partial = np.empty(self.grid.grid_shape, dtype=object)
        for entry in self.grid.get_entry_iterator():
            b = self.blocks[entry]
            oid = self.km.reduce_axis(op_name=op_name, arr=b.oid, axis=axis, keepdims=keepdims,
                                      transposed=b.transposed, syskwargs=b.syskwargs())
            partial[entry] = (oid, b.grid_entry, b.grid_shape, False)
        rshape, rblock = [], []
        for ax in range(self.ndim):
            if axis is None or ax == axis:
                if keepdims:
                    rshape.append(1); rblock.append(1)
                continue
            rshape.append(self.shape[ax]); rblock.append(self.block_shape[ax])
        rdtype = hl.reduce_out_dtype(op_name, self.dtype)
        rgrid = ArrayGrid(tuple(rshape), tuple(rblock), rdtype.name)
        out = BlockArray(rgrid, self.km)
        if axis is None:
            rb = out.blocks[(0,) * len(rshape)]
            rb.oid = self.tree_reduce(op_name, partial.flatten().tolist(), rb.grid_entry, rb.grid_shape)
            return out
        for rentry in rgrid.get_entry_iterator():
            items = []
            for k in range(self.grid.grid_shape[axis]):
                e = list(rentry)
                if keepdims:
                    e[axis] = k
                else:
                    e = e[:axis] + [k] + e[axis:]
                items.append(partial[tuple(e)])
            rb = out.blocks[rentry]
            rb.oid = self.tree_reduce(op_name, items, rb.grid_entry, rb.grid_shape)
        return out

    # ------------------------------------------------------------------ contraction
    def tensordot(self, other, axes=1):
        """blockarray.py:782-855 for axes == 1 and operands of rank <= 2 (`@`)."""
        if axes != 1:
            raise NotImplementedError("tensordot is implemented for axes=1")
        other = self._convert(other)
        if self.ndim < 1 or other.ndim < 1 or self.shape[-1] != other.shape[0]:
            raise ValueError("shape-mismatch for sum")
        if self.block_shape[-1] != other.block_shape[0]:
            raise ValueError("block shapes differ along the contracted axis: %s vs %s"
                             % (self.block_shape, other.block_shape))
        this_axes = self.grid.grid_shape[:-1]
        other_axes = other.grid.grid_shape[1:]
        ksplit = self.grid.grid_shape[-1]
        rdtype = np.result_type(self.dtype, other.dtype)
        rgrid = ArrayGrid(self.shape[:-1] + other.shape[1:], self.block_shape[:-1] + other.block_shape[1:],
                          rdtype.name)
        out = BlockArray(rgrid, self.km)
        for i in itertools.product(*map(range, this_axes)):
            for j in itertools.product(*map(range, other_axes)):
                items = []
                for k in range(ksplit):
                    a, b = self.blocks[i + (k,)], other.blocks[(k,) + j]
                    # schedule on the larger operand (blockarray.py:775-780)
                    big = a if int(np.prod(a.shape)) >= int(np.prod(b.shape)) else b
                    ge, gs = big.true_grid_entry(), big.true_grid_shape()
                    oid = self.km.bop("tensordot", a.oid, b.oid, a.transposed, b.transposed, axes=1,
                                      syskwargs={"grid_entry": ge, "grid_shape": gs})
                    items.append((oid, ge, gs, False))
                rb = out.blocks[i + j]
                rb.oid = self.tree_reduce("sum", items, rb.grid_entry, rb.grid_shape)
        return out

    def __matmul__(self, other):
        if self.ndim > 2:
            raise NotImplementedError("Matrix multiply for tensors of rank > 2 not supported yet.")
        return self.tensordot(other, 1)
