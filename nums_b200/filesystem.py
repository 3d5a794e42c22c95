"""Ingest path on the device: the counterpart of nums/core/filesystem.py for the two formats that feed the hot path
 -- `read_csv` (filesystem.py:227-283, 615-657; the HIGGS example of README.md:159-176) and the
per-block array store `write_fs / read_fs` (filesystem.py:117-175, 386-470).

File I/O stays on the host (there is no GPUDirect Storage on the box); everything after the bytes is on the device:
the text of a block's byte range is uploaded as it is and split / converted by libnumsb200.so (csrc/csv.cu), so no
Python-level float() loop and no host-side array ever exists.  The byte-range rules are the reference's, including
its boundary behaviour: a block starts after the first newline at or beyond `file_start` (unless `file_start` is 0)
and takes every line that starts before `file_end` -- so a line that starts exactly on a block boundary belongs to
no block, as in the reference.
"""
import ctypes as C
import os
import pickle

import numpy as np

from . import _lib
from . import device as dv
from .blockarray import BlockArray
from .grid import ArrayGrid

ARRAY_FILETYPE = "pkl"          # filesystem.py:92
_CSV_ERRORS = {1: "could not convert field", 2: "wrong number of fields", 3: "more than 19 significant digits "
               "with a rounding that needs arbitrary precision", 4: "hexadecimal float (float.fromhex) is not supported",
               5: "integer out of range for int64"}


def get_file_size(filename):
    # storage/utils.py:97-99
    return os.path.getsize(filename)


def get_batches(total_size, num_batches):
    """storage/utils.py:29-61 (Batch.from_num_batches): [start, end) byte ranges."""
    batch_size = (total_size + num_batches - 1) // num_batches
    if total_size < batch_size or batch_size == 0:
        return [[0, total_size]]
    starts = list(range(0, total_size, batch_size))
    count = int(total_size / batch_size)
    batches = [starts[i: i + 2] for i in range(count)]
    if len(batches[-1]) == 1:
        batches[-1].append(total_size)
    if batches[-1][1] != total_size:
        batches.append([batches[-1][1], total_size])
    return batches


def _block_text(filename, file_start, file_end, has_header):
    """The bytes of the lines read_csv_block would parse (filesystem.py:258-279), ending in a newline."""
    size = os.path.getsize(filename)
    with open(filename, "rb") as fh:
        def skip_line(pos):
            fh.seek(pos)
            while True:
                chunk = fh.read(1 << 16)
                if not chunk:
                    return size
                k = chunk.find(b"\n")
                if k >= 0:
                    return pos + k + 1
                pos += len(chunk)

        begin = skip_line(file_start) if file_start != 0 else 0
        if file_start == 0 and has_header and begin < file_end:
            begin = skip_line(0)
        if begin >= file_end or begin >= size:
            return b""
        fh.seek(begin)
        data = fh.read(file_end - begin)
        # the last line taken is the one that starts before file_end: read on to its terminator
        while not data.endswith(b"\n"):
            chunk = fh.read(1 << 16)
            if not chunk:
                break
            k = chunk.find(b"\n")
            if k >= 0:
                data += chunk[: k + 1]
                break
            data += chunk
    if not data.endswith(b"\n"):
        data += b"\n"
    return data


def read_csv_block(filename, file_start, file_end, dtype, delimiter, has_header, device_index=None):
    """filesystem.py:227-283 with the conversion on the device.  Returns (block, shape)."""
    import torch
    dev = torch.cuda.current_device() if device_index is None else int(device_index)
    as_int_of_float = dtype is int                      # the reference's converter for `int` is int(float(x))
    np_dtype = np.dtype(np.float64 if as_int_of_float else dtype)
    if np_dtype not in (np.dtype(np.float64), np.dtype(np.int64)):
        raise NotImplementedError("read_csv on the device converts to float64 or int64, not %s" % np_dtype)
    if len(delimiter.encode()) != 1:
        raise NotImplementedError("read_csv on the device needs a single-byte delimiter")
    data = _block_text(filename, file_start, file_end, has_header)
    out_dtype = np.dtype(np.int64) if as_int_of_float else np_dtype
    if not data:
        empty = dv.empty((0,), out_dtype, dev)
        return empty, (0,)
    first = data[: data.index(b"\n")].strip(b"\r\n")
    cols = first.count(delimiter.encode()) + 1
    lib = _lib.load()
    st = dv.stream_ptr(dev)
    text = dv.upload_bytes(data, dev)
    rows_c = C.c_int64(0)
    sc = dv.scratch(lib.nb_csv_scratch_bytes(len(data), data.count(b"\n")), dev)   # sized by a host memchr pass
    _lib.check(lib.nb_csv_count_rows(dev, text.data_ptr(), len(data), sc.data_ptr(), sc.numel(), C.byref(rows_c), st))
    rows = int(rows_c.value)
    out = dv.empty((rows, cols), np_dtype, dev)
    err = (C.c_int64 * 3)()
    _lib.check(lib.nb_csv_parse(dev, text.data_ptr(), len(data), rows, cols, delimiter.encode()[0], out.code, out.ptr,
                                sc.data_ptr(), sc.numel(), err, st))
    if err[0]:
        raise ValueError("%s: line %d of the block starting at byte %d, field %d: %s"
                         % (filename, err[0], file_start, err[2], _CSV_ERRORS.get(int(err[1]), "error %d" % err[1])))
    if as_int_of_float:
        from .cuda_kernel import KernelCls
        out = KernelCls().astype(out, "int64")
    return out, (rows, cols)


class FileSystem:
    """The array-level operations of nums/core/filesystem.py:286-657 that the ingest path uses."""

    def __init__(self, km):
        self.km = km
        self.km.register("read_csv_block", read_csv_block, {})

    # ---- CSV (filesystem.py:615-657)
    def read_csv(self, filename, dtype=float, delimiter=",", has_header=False, num_workers=4):
        batches = get_batches(get_file_size(filename), num_workers)
        results = []
        for i, (start, end) in enumerate(batches):
            results.append(self.km.call("read_csv_block", filename, start, end, dtype, delimiter, has_header,
                                        syskwargs={"grid_entry": (i,), "grid_shape": (num_workers,),
                                                   "options": {"num_returns": 2}}))
        out_dtype = np.int64 if dtype is int else dtype
        arrays = []
        for block, shape in results:
            if shape[0] == 0:
                continue
            arr = BlockArray(ArrayGrid(shape=shape, block_shape=shape, dtype=np.dtype(out_dtype).name), self.km)
            arr.blocks[(0,) * len(shape)].oid = block
            arrays.append(arr)
        return arrays

    # ---- per-block array store (filesystem.py:95-175, 386-470)
    @staticmethod
    def _entry_path(filename, grid_entry):
        return os.path.join(filename, "_".join(map(str, grid_entry)) + "." + ARRAY_FILETYPE)

    def write_fs(self, ba, filename):
        """application.py:72-75: one pickle per block plus meta.pkl holding the grid."""
        os.makedirs(filename, exist_ok=True)
        for entry in ba.grid.get_entry_iterator():
            blk = ba.blocks[entry]
            host = np.asarray(self.km.get(blk.oid))
            if blk.transposed:
                host = host.T
            with open(self._entry_path(filename, entry), "wb") as fh:
                pickle.dump(np.ascontiguousarray(host), fh)
        with open(os.path.join(filename, "meta.pkl"), "wb") as fh:
            pickle.dump({"grid": ba.grid.to_meta()}, fh)
        return ba

    def read_fs(self, filename):
        """filesystem.py:441-470: blocks are loaded onto the device the grid assigns them to."""
        meta_path = os.path.join(filename, "meta.pkl")
        if not os.path.isfile(meta_path):
            raise FileNotFoundError("Unable to load meta data for file %s" % filename)
        with open(meta_path, "rb") as fh:
            grid = ArrayGrid.from_meta(pickle.load(fh)["grid"])
        ba = BlockArray(grid, self.km)
        for entry in grid.get_entry_iterator():
            with open(self._entry_path(filename, entry), "rb") as fh:
                host = pickle.load(fh)
            ba.blocks[entry].oid = self.km.put(host, syskwargs={"grid_entry": entry, "grid_shape": grid.grid_shape})
        return ba

    def delete_fs(self, filename):
        # filesystem.py:177-190: only a directory that holds meta.pkl is a stored array
        import shutil
        if not os.path.isfile(os.path.join(filename, "meta.pkl")):
            return False
        shutil.rmtree(filename, ignore_errors=True)
        return True
