"""CPU-only checks: the C-ABI library loads and exports what include/nums_b200.h declares, the host
logic (dtype resolution, broadcasting, block shapes, placement) agrees with NumPy / the reference,
and the product refuses to run without a GPU."""
import ast
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, load_golden
from nums_b200 import _lib, build
from nums_b200 import hostlogic as hl
from nums_b200.grid import ArrayGrid, CyclicDeviceGrid, gpu_devices
from nums_b200.kernel_manager import KernelManager

REF = "/root/reference"


def header_functions():
    text = open(os.path.join(ROOT, "include", "nums_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    build.build_library()
    lib = ctypes.CDLL(build.LIB_PATH)
    names = header_functions()
    assert len(names) >= 30
    for name in names:
        assert hasattr(lib, name), "libnumsb200.so does not export %s" % name
    assert sorted(_lib.SIGNATURES) == names
    lib.nb_version.restype = ctypes.c_int
    assert lib.nb_version() == 1


def test_scratch_queries_need_no_gpu():
    lib = _lib.load()
    assert lib.nb_glm_stats_len(256) == 257 * 257 + 257 + 1
    assert lib.nb_where_chunks(10 ** 8) == -(-10 ** 8 // 4096)
    assert lib.nb_reduce_scratch_bytes(1, 10 ** 8, 1) > 0
    assert lib.nb_reduce_scratch_bytes(10 ** 6, 4, 1) == 0
    assert lib.nb_inv_scratch_bytes(257) >= 257 * 257 * 8


def test_product_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from nums_b200.app import make_app
    with pytest.raises(_lib.NumsB200Error):
        make_app()


BOPS = ["add", "sub", "mul", "truediv", "floor_divide", "mod", "pow", "lt", "eq", "bitwise_and", "left_shift",
        "maximum", "fmin", "arctan2", "logaddexp", "logical_or"]


@pytest.mark.parametrize("op", BOPS)
def test_bop_dtype_resolution_matches_numpy(op):
    name = hl.canonical_bop(op)
    for da in (np.float64, np.int64, np.bool_):
        for db in (np.float64, np.int64, np.bool_):
            a, b = np.array(1, dtype=da), np.array(2, dtype=db)
            try:
                want = getattr(np, name)(a, b).dtype   # what the reference does (array/utils.py:40-51)
            except TypeError:
                with pytest.raises(TypeError):
                    hl.resolve_bop(op, da, db)
                continue
            if want not in (np.dtype(np.float64), np.dtype(np.int64), np.dtype(np.bool_)):
                with pytest.raises(NotImplementedError):
                    hl.resolve_bop(op, da, db)
                continue
            _, cdt, odt = hl.resolve_bop(op, da, db)
            assert odt == want
            assert cdt in (hl.NB_F64, hl.NB_I64, hl.NB_BOOL)


def test_uop_dtype_resolution_matches_numpy():
    for op in ["exp", "abs", "negative", "sqrt", "sign", "isnan", "invert", "ceil", "fabs", "square"]:
        for dt in (np.float64, np.int64):
            try:
                want = getattr(np, op)(np.array(1, dtype=dt)).dtype
            except TypeError:
                continue
            assert hl.resolve_uop(op, dt)[2] == want


def test_broadcast_and_collapse():
    shape, sa, sb = hl.broadcast_geometry((6, 1), (1, 1), (6, 4), (4, 1))
    assert shape == (6, 4) and sa == (1, 0) and sb == (4, 1)
    shape, sa, sb = hl.broadcast_geometry((), (), (5,), (1,))
    assert shape == (5,) and sa == (0,) and sb == (1,)
    with pytest.raises(ValueError):
        hl.broadcast_geometry((3,), (1,), (4,), (1,))
    # contiguous same-shape operands collapse to one axis
    assert hl.collapse_dims((2, 3, 4), (12, 4, 1), (12, 4, 1)) == ((24,), (1,), (1,))
    # transposed 2-d operand cannot be merged
    cs = hl.collapse_dims((3, 4), (4, 1), (1, 3))
    assert cs[0] == (3, 4)
    assert hl.reduce_geometry((3, 4, 5), 1) == (3, 4, 5)
    assert hl.reduce_geometry((3, 4, 5), None) == (1, 60, 1)
    assert hl.reduce_geometry((7,), 0) == (1, 7, 1)


def test_reduce_out_dtype():
    for op in ("sum", "min", "max", "all", "any"):
        for dt in (np.float64, np.int64, np.bool_):
            want = getattr(np, op)(np.array([0, 1], dtype=dt)).dtype
            assert hl.reduce_out_dtype(op, dt) == want


def test_block_shape_rule_matches_reference():
    g = load_golden("block_shapes")
    for i in range(len(g["case_ndim"])):
        nd = int(g["case_ndim"][i])
        shape = tuple(int(v) for v in g["case_shape"][i][:nd])
        got = KernelManager.compute_block_shape_static(shape, np.dtype(str(g["case_dtype"][i])),
                                                       tuple(int(v) for v in g["case_cluster"][i]),
                                                       int(g["case_cores"][i]))
        assert got == tuple(int(v) for v in g["block_shape"][i][:nd])
    # the configurations of BASELINE.json
    assert KernelManager.compute_block_shape_static((10 ** 8, 256), np.float64, (8, 1), 8) == (12500000, 256)
    assert KernelManager.compute_block_shape_static((10 ** 6, 2), np.float64, (8, 1), 8) == (10 ** 6, 2)


def test_grid_and_static_placement():
    grid = ArrayGrid((10, 7), (4, 7), "float64")
    assert grid.grid_shape == (3, 1)
    assert grid.get_block_shape((2, 0)) == (2, 7)
    assert grid.get_slice_tuples((1, 0)) == [(4, 8), (0, 7)]
    dg = CyclicDeviceGrid((4, 1), "gpu", gpu_devices(4))
    # row block i of a (G,1)-blocked array lives on GPU i: static placement
    assert [dg.get_device((i, 0), (4, 1)).device for i in range(4)] == [0, 1, 2, 3]
    assert dg.get_device((5, 0), (8, 1)).device == 1
    assert dg.get_device((2,), (4,)).device == 2
    for world in (1, 2, 3, 8):
        n = 1001
        spans = [hl.row_shard(n, world, r) for r in range(world)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        ag = ArrayGrid((n,), (-(-n // world),), "float64")
        assert [tuple(ag.get_slice_tuples((r,))[0]) for r in range(ag.grid_shape[0])] == \
            [s for s in spans if s[1] > s[0]]


def _class_methods(path, cls_name):
    tree = ast.parse(open(path).read())
    for node in tree.body:
        if isinstance(node, ast.ClassDef) and node.name == cls_name:
            out = {}
            for fn in node.body:
                if isinstance(fn, ast.FunctionDef):
                    args = [a.arg for a in fn.args.args + fn.args.kwonlyargs]
                    if fn.args.vararg:
                        args.append(fn.args.vararg.arg)
                    out[fn.name] = args
            return out
    raise KeyError(cls_name)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")
def test_kernel_module_satisfies_reference_interface():
    """What nums.core.backends.utils.check_implementation (:74-87) enforces, checked from the
    reference's source without importing it."""
    import inspect
    from nums_b200 import cuda_kernel
    iface = _class_methods(os.path.join(REF, "nums/core/kernel/kernel_interface.py"), "Kernel")
    assert len(iface) == 39
    for name, args in iface.items():
        assert hasattr(cuda_kernel.KernelCls, name), "KernelCls.%s missing" % name
        have = inspect.getfullargspec(getattr(cuda_kernel.KernelCls, name))
        have_names = set(have.args + have.kwonlyargs + ([have.varargs] if have.varargs else []))
        for a in args:
            if a in ("self", "syskwargs"):
                continue
            assert a in have_names, "KernelCls.%s lacks parameter %s" % (name, a)
    assert hasattr(cuda_kernel, "RNG") and hasattr(cuda_kernel.RNG, "new_block_rng_params")


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference checkout not present")
def test_backend_satisfies_reference_interface():
    from nums_b200.cuda_backend import CudaBackend
    iface = _class_methods(os.path.join(REF, "nums/core/backends/base.py"), "Backend")
    assert len(iface) == 12
    for name in iface:
        assert callable(getattr(CudaBackend, name, None)), "CudaBackend.%s missing" % name


def test_rng_stream_matches_reference_rule():
    from nums_b200 import cuda_kernel
    from oracle import np_kernel as ok
    rng = cuda_kernel.RNG(1337)
    p0, p1 = rng.new_block_rng_params(), rng.new_block_rng_params()
    assert p0 == (1337, 0) and p1 == (1337, 1)
    a = cuda_kernel.block_rng(*p1).normal(0.0, 1.0, 5)
    b = ok.random_block(p1, "normal", (0.0, 1.0, 5), (5,), np.float64)
    np.testing.assert_array_equal(a, b)


def test_parallel_permutation_formulation_equals_numpy():
    """The prefix-sum / pointer-jumping restatement of NumPy's Fisher-Yates that csrc/perm.cu implements, run in NumPy
    (tools/perm_prototype.py): identical to Generator.permutation for sizes around the mask boundaries."""
    from numpy.random import PCG64, Generator
    from tools import perm_prototype as pp
    for n in (1, 2, 3, 4, 5, 8, 9, 255, 256, 257, 1000, 65537, 200_003):
        for jump in (0, 5):
            got, _, rounds, _ = pp.permutation(n, 99, jump)
            assert np.array_equal(got, Generator(PCG64(99).jumped(jump)).permutation(n)), (n, jump)


def test_reference_layers_drive_the_plugin_host_logic():
    """The unmodified reference (oracle/_ref) registers cuda_kernel in its own KernelManager and drives
    CudaBackend / KernelCls through whole programs; the library's compute calls are stubbed
    (tools/dryrun_reference_boundary.py), so only the host logic is exercised.  The values are checked on the GPU by
    tests/test_gpu_reference_boundary.py."""
    import subprocess
    import sys
    from oracle import refharness
    if not refharness.present():
        pytest.skip("oracle/_ref absent (python oracle/make_ref.py)")
    for gpus in ("1", "2"):
        proc = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "dryrun_reference_boundary.py"), gpus],
                              capture_output=True, text=True, timeout=600)
        assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
        assert proc.stdout.count("\nok") + proc.stdout.startswith("ok") >= 14


def test_oracle_ref_is_an_unmodified_copy_of_the_reference():
    """oracle/_ref/nums must be byte-identical to /root/reference/nums (build container only): the boundary tests and the
    timed CPU arm claim to run the reference itself."""
    import hashlib
    import json
    from oracle import make_ref, refharness
    if not refharness.present() or not os.path.isdir(os.path.join(REF, "nums")):
        pytest.skip("needs both oracle/_ref and /root/reference")
    manifest = json.load(open(os.path.join(make_ref.OUT, "MANIFEST.json")))["files"]
    assert len(manifest) > 50
    for rel, digest in manifest.items():
        here = hashlib.sha256(open(os.path.join(make_ref.OUT, rel), "rb").read()).hexdigest()
        there = hashlib.sha256(open(os.path.join(REF, rel), "rb").read()).hexdigest()
        assert here == there == digest, rel


def test_bind_host_to_gpu_changes_nothing_without_a_gpu():
    """device.bind_host_to_gpu (NUMA placement of a rank's host buffers) must be harmless where NVML / CUDA are absent."""
    import os
    import torch
    from nums_b200 import device as dv
    before = os.sched_getaffinity(0)
    got = dv.bind_host_to_gpu(0)
    if not torch.cuda.is_available():
        assert got is None and os.sched_getaffinity(0) == before
    else:
        assert got is None or set(got) <= before
        os.sched_setaffinity(0, before)


def test_int8_flag_selection_thresholds():
    """fused.int8_flags: level 5 is dropped and four planes are chosen only where their error models stay a fifth (a tenth)
    below the 1e-10 bar; NUMS_B200_INT8_PLANES pins the plane count."""
    import os
    from nums_b200 import fused

    class Found:
        def __init__(self, r):
            self.range_log2, self.bad = r, False

    old = os.environ.pop("NUMS_B200_INT8_PLANES", None)
    try:
        small, big = fused.int8_flags(Found(3), 40_000), fused.int8_flags(Found(3), 12_500_000)
        assert small & fused.NB_GLM_INT8_HESSIAN and not small & fused.NB_GLM_INT8_FOUR_PLANES
        assert big & fused.NB_GLM_INT8_FOUR_PLANES and big & fused.NB_GLM_INT8_NO_LEVEL5
        assert not fused.int8_flags(Found(3), 2_000_000) & fused.NB_GLM_INT8_FOUR_PLANES
        assert fused.int8_flags(Found(3), 2_400_000) & fused.NB_GLM_INT8_FOUR_PLANES
        assert not fused.int8_flags(Found(5), 12_500_000) & fused.NB_GLM_INT8_FOUR_PLANES       # heavier tails need more rows
        os.environ["NUMS_B200_INT8_PLANES"] = "5"
        assert not fused.int8_flags(Found(3), 12_500_000) & fused.NB_GLM_INT8_FOUR_PLANES
        os.environ["NUMS_B200_INT8_PLANES"] = "4"
        assert fused.int8_flags(Found(3), 1_000) & fused.NB_GLM_INT8_FOUR_PLANES
    finally:
        os.environ.pop("NUMS_B200_INT8_PLANES", None)
        if old is not None:
            os.environ["NUMS_B200_INT8_PLANES"] = old
