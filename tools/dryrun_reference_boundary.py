"""CPU dry run of the plug-in's HOST logic under the reference's own layers (debug / CI tool, not a product path).

Blocks become zero-filled CPU tensors and every compute entry point of libnumsb200.so is replaced by a stub that
only checks its arguments against the ctypes signature and returns 0.  The unmodified reference (oracle/_ref) then
drives KernelManager -> CudaBackend -> cuda_kernel through the programs of tests/test_gpu_reference_boundary.py:
Python-level faults (argument handling, shapes, multi-returns) surface here without a GPU; values are not compared.
Usage: python tools/dryrun_reference_boundary.py [num_gpus]   -> prints one "ok"/"FAIL" line per program, exit 1 on FAIL.
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import warnings; warnings.simplefilter("ignore")
import contextlib, ctypes
import numpy as np, torch
from nums_b200 import _lib, device as dv
real = _lib.load()
class Fake:
    def __getattr__(self, name):
        if name.endswith("_bytes") or name in ("nb_glm_stats_len","nb_where_chunks","nb_version"):
            return getattr(real, name)
        def f(*a):
            # argument marshalling check through ctypes argtypes
            res, argtypes = _lib.SIGNATURES[name]
            assert len(a)==len(argtypes), (name, len(a), len(argtypes))
            for v,t in zip(a,argtypes):
                if v is None: continue
                if isinstance(v, ctypes._SimpleCData) or isinstance(v, ctypes.Array) or isinstance(v, ctypes._Pointer): continue
                t(v) if t not in (ctypes.c_void_p,) else ctypes.c_void_p(v)
            return 0
        return f
_lib._LIB = Fake()
dv.require_cuda = lambda: None
def empty(shape, dtype, device_index):
    return dv.DeviceBlock(torch.zeros(tuple(int(s) for s in shape), dtype=dv.torch_dtype(dtype)))
dv.empty = empty
dv.scratch = lambda n, d: torch.zeros(max(int(n),256), dtype=torch.uint8)
dv.stream_ptr = lambda d: ctypes.c_void_p(0)
dv.DeviceBlock.device_index = property(lambda self: 0)
dv.to_device = lambda b, d: b
torch.cuda.device = lambda d: contextlib.nullcontext()
torch.cuda.device_count = lambda: int(sys.argv[1]) if len(sys.argv)>1 else 1
torch.cuda.current_device = lambda: 0
class S: 
    cuda_stream=0
    def synchronize(self): pass
torch.cuda.current_stream = lambda d=None: S()
from oracle import refharness as rh
rh.bootstrap()
import test_gpu_reference_boundary as T
G = torch.cuda.device_count()
apps={"cuda":rh.cuda_app(),"serial":rh.serial_app(G),"G":G}
import traceback
failed = []
def run(name,*a):
    try:
        getattr(T,name)(apps,*a); print("ok",name,a)
    except AssertionError as e:
        tb=traceback.extract_tb(e.__traceback__)
        if tb[-1].filename.endswith("test_gpu_reference_boundary.py") and tb[-1].name.startswith("test_"):
            print("ok(values differ as expected)",name,a)
        else:
            traceback.print_exc(); print("FAIL",name,a); failed.append(name)
    except Exception as e:
        traceback.print_exc(); print("FAIL",name,a); failed.append(name)
run("test_cuda_kernel_registers_in_reference_kernel_manager")
run("test_config2_dot_where_argmax_sum")
run("test_block_matmul_and_axis_reductions")
for s,kw in [("lbfgs", dict(penalty="l2", C=2.0, max_iter=8)),("newton", dict(penalty="l2", C=2.0)),("gd", dict(penalty="l2", C=2.0, max_iter=5, lr=1e-4)),("irls", dict(max_iter=5))]:
    run("test_logistic_solvers_of_the_reference",s,kw)
run("test_linear_and_poisson_of_the_reference")
run("test_elementwise_and_integer_semantics_through_nums_numpy")
for d,size,fam in [(2, 10 ** 5, "logistic"), (16, 40000, "logistic"),(8, 8000, "poisson"), (8, 8000, "ridge")]:
    run("test_reference_glm_fit_with_the_fused_newton_routed_in",d,size,fam)
run("test_readme_logistic_regression_unchanged_program")
sys.exit(1 if failed else 0)
