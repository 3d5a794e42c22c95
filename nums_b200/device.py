"""Device-resident blocks.  PyTorch is used for exactly three things here: allocating HBM
(`torch.empty`), host<->device / peer copies, and stream handles.  All arithmetic goes through
libnumsb200 (see _lib.py); nothing in this package calls a torch compute op on block data.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from . import hostlogic as hl

_NP2TORCH = {
    np.dtype(np.float64): torch.float64,
    np.dtype(np.int64): torch.int64,
    np.dtype(np.bool_): torch.bool,
    np.dtype(np.float32): torch.float32,
    np.dtype(np.int32): torch.int32,
}
_TORCH2NP = {v: k for k, v in _NP2TORCH.items()}


_CUDA_OK = False


def require_cuda():
    # torch.cuda.is_available() costs ~0.3 ms per call (a driver query): ask once, every block allocation comes through here
    global _CUDA_OK
    if _CUDA_OK:
        return
    if not torch.cuda.is_available():
        raise _lib.NumsB200Error("no CUDA device visible: the B200 kernel module has no CPU fallback")
    _CUDA_OK = True


def torch_dtype(dtype):
    dtype = np.dtype(dtype)
    if dtype not in _NP2TORCH:
        raise TypeError("dtype %s is not supported by the B200 kernel module" % dtype)
    return _NP2TORCH[dtype]


class DeviceBlock:
    """One block of a BlockArray in HBM: the object id ("oid") handed back to NumS.

    Wraps a torch tensor used purely as a typed, strided window on device memory.  Like the
    NumPy blocks of the reference it is immutable by convention (numpy_kernel.py:131-135)."""

    __slots__ = ("t", "prep")

    def __init__(self, tensor):
        self.t = tensor
        self.prep = None      # per-block column scales of the int8 Hessian path (fused.glm_prepare), made on first use

    # --- metadata (host side only) ---
    @property
    def shape(self):
        return tuple(self.t.shape)

    @property
    def ndim(self):
        return self.t.dim()

    @property
    def dtype(self):
        return _TORCH2NP[self.t.dtype]

    @property
    def size(self):
        return self.t.numel()

    @property
    def device_index(self):
        return self.t.device.index

    @property
    def strides(self):
        return tuple(self.t.stride())

    @property
    def ptr(self):
        return self.t.data_ptr()

    @property
    def code(self):
        return hl.dtype_code(self.dtype)

    def is_contiguous(self):
        return self.t.is_contiguous()

    @property
    def T(self):
        # deferred transpose: metadata only (reverses all axes like ndarray.T)
        return DeviceBlock(self.t.permute(*reversed(range(self.t.dim()))))

    def __repr__(self):
        return "DeviceBlock(shape=%s, dtype=%s, cuda:%s)" % (self.shape, self.dtype, self.device_index)


def bind_host_to_gpu(device_index):
    """Restricts the calling process to the CPUs next to GPU `device_index` (NVML's ideal CPU set, i.e. the NUMA node of the
    GPU's PCIe root), so that host buffers it allocates and first touches afterwards are local to that GPU.  For the
    one-process-per-GPU layout on multi-socket hosts.  (The 8-GPU box of the round-2 measurements presents one NUMA node
    of 32 CPUs for all GPUs: the call changes nothing there, and its 186 GB/s of aggregate upload are the host's limit.)
    Returns the CPU list, or None when NVML or the topology is not available (nothing is changed then)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(device_index)
        bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cpus = [64 * w + b for w, mask in enumerate(words) for b in range(64) if (int(mask) >> b) & 1]
        allowed = sorted(set(cpus) & os.sched_getaffinity(0))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return None


def stream_ptr(device_index):
    return C.c_void_p(torch.cuda.current_stream(device_index).cuda_stream)


def empty(shape, dtype, device_index):
    require_cuda()
    return DeviceBlock(torch.empty(tuple(int(s) for s in shape), dtype=torch_dtype(dtype),
                                   device=torch.device("cuda", device_index)))


def scratch(nbytes, device_index):
    """Uninitialised scratch of at least nbytes (256-byte aligned by the caching allocator)."""
    return torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=torch.device("cuda", device_index))


_PIN_CHUNK = 1 << 28   # 256 MiB staging chunks for large uploads


def upload_bytes(data, device_index):
    """Raw bytes -> uint8 tensor in HBM (256-byte aligned): the text buffer of the CSV ingest path."""
    require_cuda()
    src = torch.frombuffer(bytearray(data), dtype=torch.uint8)
    out = torch.empty(max(len(data), 8), dtype=torch.uint8, device=torch.device("cuda", device_index))
    for lo in range(0, len(data), _PIN_CHUNK):
        hi = min(len(data), lo + _PIN_CHUNK)
        out[lo:hi].copy_(src[lo:hi], non_blocking=False)
    return out


def from_numpy(arr, device_index):
    """Host ndarray / NumPy scalar -> DeviceBlock: copied straight from the caller's buffer in 256 MiB pieces (at the
    link rate when that buffer is page-locked, through the driver's staging otherwise)."""
    require_cuda()
    arr = np.asarray(arr)
    if arr.dtype == np.dtype(np.int32) or arr.dtype == np.dtype(np.float32):
        pass
    elif arr.dtype.kind in "iu" and arr.dtype != np.dtype(np.int64):
        arr = arr.astype(np.int64)
    elif arr.dtype.kind == "f" and arr.dtype != np.dtype(np.float64):
        arr = arr.astype(np.float64)
    arr = np.ascontiguousarray(arr)
    out = empty(arr.shape, arr.dtype, device_index)
    if arr.size == 0:
        return out
    flat_dst = out.t.view(-1)
    flat_src = torch.from_numpy(arr.reshape(-1))
    if arr.nbytes <= _PIN_CHUNK:
        flat_dst.copy_(flat_src, non_blocking=False)
    else:
        step = _PIN_CHUNK // arr.itemsize
        for lo in range(0, arr.size, step):
            flat_dst[lo:lo + step].copy_(flat_src[lo:lo + step], non_blocking=False)
    return out


def to_numpy(block):
    """DeviceBlock -> host ndarray (synchronises the block's stream)."""
    t = block.t
    if not t.is_contiguous():
        t = contiguous(block).t
    return t.cpu().numpy()


def contiguous(block):
    """C-contiguous copy (through nb_copy_strided) unless the block already is."""
    if block.t.is_contiguous():
        return block
    lib = _lib.load()
    dev = block.device_index
    out = empty(block.shape, block.dtype, dev)
    nd = block.ndim
    if nd > 4:
        raise NotImplementedError("strided copies of more than 4 axes are not supported")
    shape = _lib.i64_array(block.shape)
    ss = _lib.i64_array(block.strides)
    ds = _lib.i64_array(hl.contiguous_strides(block.shape))
    _lib.check(lib.nb_copy_strided(dev, nd, shape, block.t.element_size(), block.ptr, ss, out.ptr, ds,
                                   stream_ptr(dev)))
    return out


def copy_into(dst_view, src_view):
    """dst_view[...] = src_view for two equally shaped strided windows (torch views used as metadata)."""
    lib = _lib.load()
    if tuple(dst_view.shape) != tuple(src_view.shape):
        src_view = src_view.expand(dst_view.shape)   # metadata-only broadcast (stride 0)
    nd = dst_view.dim()
    if nd > 4:
        raise NotImplementedError("strided copies of more than 4 axes are not supported")
    if dst_view.numel() == 0:
        return
    dev = dst_view.device.index
    if src_view.dtype != dst_view.dtype:
        raise TypeError("copy_into: dtype mismatch %s vs %s" % (src_view.dtype, dst_view.dtype))
    shape = _lib.i64_array(dst_view.shape)
    ss = _lib.i64_array(src_view.stride())
    ds = _lib.i64_array(dst_view.stride())
    _lib.check(lib.nb_copy_strided(dev, nd, shape, dst_view.element_size(), src_view.data_ptr(), ss,
                                   dst_view.data_ptr(), ds, stream_ptr(dev)))


def to_device(block, device_index):
    """Moves a block to another GPU of the box (peer copy over NVLink) if it is not already there."""
    if block.device_index == device_index:
        return block
    src = contiguous(block)
    return DeviceBlock(src.t.to(torch.device("cuda", device_index)))
