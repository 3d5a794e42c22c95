"""ctypes binding of libnumsb200.so (see include/nums_b200.h).

There is no CPU fallback: if the shared library is missing or a call fails, an exception is
raised.  The library is built in-tree by ``nums_b200.build.build_library`` (nvcc, sm_100a).
"""
import ctypes as C
import os

from . import build as _build

_LIB = None

c_i64 = C.c_int64
c_int = C.c_int
c_vp = C.c_void_p
c_sz = C.c_size_t
c_dbl = C.c_double
P_I64 = C.POINTER(C.c_int64)

# name -> (restype, argtypes); mirrors include/nums_b200.h one to one
SIGNATURES = {
    "nb_last_error": (C.c_char_p, []),
    "nb_version": (c_int, []),
    "nb_device_count": (c_int, []),
    "nb_device_info": (c_int, [c_int, C.POINTER(c_int), C.POINTER(c_int), C.POINTER(c_int)]),
    "nb_bop": (c_int, [c_int, c_int, c_int, P_I64, c_vp, c_int, P_I64, c_vp, c_int, P_I64, c_vp, c_int, c_int, c_vp]),
    "nb_uop": (c_int, [c_int, c_int, c_i64, c_vp, c_int, c_vp, c_int, c_int, c_vp]),
    "nb_where3": (c_int, [c_int, c_i64, c_vp, c_int, c_vp, c_vp, c_vp, c_int, c_vp]),
    "nb_where_chunks": (c_i64, [c_i64]),
    "nb_where_count": (c_int, [c_int, c_i64, c_vp, c_int, c_vp, c_vp, c_vp]),
    "nb_where_fill": (c_int, [c_int, c_i64, c_vp, c_int, c_int, P_I64, P_I64, c_vp, C.POINTER(c_vp), c_vp]),
    "nb_astype": (c_int, [c_int, c_i64, c_vp, c_int, c_vp, c_int, c_vp]),
    "nb_fill": (c_int, [c_int, c_i64, c_vp, c_int, c_dbl, c_i64, c_vp]),
    "nb_eye": (c_int, [c_int, c_i64, c_i64, c_vp, c_int, c_vp]),
    "nb_copy_strided": (c_int, [c_int, c_int, P_I64, c_int, c_vp, P_I64, c_vp, P_I64, c_vp]),
    "nb_pcg64_random": (c_int, [c_int, c_i64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), c_vp, c_vp]),
    "nb_pcg64_permutation_scratch_bytes": (c_i64, [c_i64]),
    "nb_pcg64_permutation": (c_int, [c_int, c_i64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), c_vp, c_vp, c_i64,
                                     c_vp]),
    "nb_csv_scratch_bytes": (c_i64, [c_i64, c_i64]),
    "nb_csv_count_rows": (c_int, [c_int, c_vp, c_i64, c_vp, c_i64, C.POINTER(C.c_int64), c_vp]),
    "nb_csv_parse": (c_int, [c_int, c_vp, c_i64, c_i64, c_int, c_int, c_int, c_vp, c_vp, c_i64, C.POINTER(C.c_int64),
                             c_vp]),
    "nb_pcg64_normal_scratch_bytes": (c_i64, [c_i64]),
    "nb_pcg64_normal": (c_int, [c_int, c_i64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_double, C.c_double,
                                c_int, c_vp, c_vp, c_i64, c_vp]),
    "nb_pcg64_integers_scratch_bytes": (c_i64, [c_i64, C.c_uint64]),
    "nb_pcg64_integers": (c_int, [c_int, c_i64, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), c_i64, C.c_uint64,
                                  c_vp, c_vp, c_i64, c_vp]),
    "nb_reduce_scratch_bytes": (c_sz, [c_i64, c_i64, c_i64]),
    "nb_reduce": (c_int, [c_int, c_int, c_i64, c_i64, c_i64, c_vp, c_int, c_vp, c_int, c_int, c_vp, c_sz, c_vp]),
    "nb_reduce_static": (c_int, [c_int, c_int, c_i64, c_i64, c_i64, c_vp, c_int, c_vp, c_int, c_int, c_vp]),
    "nb_bop_reduce": (c_int, [c_int, c_int, c_i64, c_vp, c_vp, c_vp, c_int, c_vp]),
    "nb_sum_reduce": (c_int, [c_int, c_int, c_i64, C.POINTER(c_vp), c_vp, c_int, c_vp]),
    "nb_argop_scratch_bytes": (c_sz, [c_i64]),
    "nb_argop": (c_int, [c_int, c_int, c_i64, c_vp, c_int, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "nb_matmul_scratch_bytes": (c_sz, [c_int, c_i64, c_i64, c_i64]),
    "nb_matmul_f64": (c_int, [c_int, c_i64, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_vp,
                              c_i64, c_vp, c_sz, c_vp]),
    "nb_matmul_i64": (c_int, [c_int, c_i64, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_i64, c_vp, c_i64, c_vp]),
    "nb_inv_scratch_bytes": (c_sz, [c_i64]),
    "nb_inv_f64": (c_int, [c_int, c_i64, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "nb_cholesky_f64": (c_int, [c_int, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "nb_qr_signs_scratch_bytes": (c_i64, [c_i64]),
    "nb_qr_signs_f64": (c_int, [c_int, c_i64, c_vp, c_i64, c_int, c_vp, c_vp, c_i64, c_vp]),
    "nb_svd_scratch_bytes": (c_i64, [c_i64]),
    "nb_svd_f64": (c_int, [c_int, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_i64, c_vp]),
    "nb_glm_stats_len": (c_i64, [c_i64]),
    "nb_glm_scratch_bytes": (c_sz, [c_int, c_i64, c_i64, c_int]),
    "nb_glm_stats": (c_int, [c_int, c_int, c_vp, c_i64, c_i64, c_i64, c_vp, c_int, c_vp, c_int, c_vp, c_vp, c_sz, c_vp]),
    "nb_glm_prep_bytes": (c_sz, [c_i64]),
    "nb_glm_prepare": (c_int, [c_int, c_vp, c_i64, c_i64, c_i64, c_vp, c_sz, c_vp]),
    "nb_glm_stats_prepared": (c_int, [c_int, c_int, c_vp, c_i64, c_i64, c_i64, c_vp, c_int, c_vp, c_int, c_vp, c_vp,
                                      c_vp, c_sz, c_vp]),
    "nb_glm_newton_scratch_bytes": (c_sz, [c_i64]),
    "nb_glm_newton_update": (c_int, [c_int, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_sz, c_vp]),
    "nb_glm_newton_fit_scratch_bytes": (c_sz, [c_int, c_int, c_i64, c_i64, c_int]),
    "nb_glm_newton_fit": (c_int, [c_int, c_int, c_int, C.POINTER(c_vp), P_I64, c_i64, P_I64, C.POINTER(c_vp), c_int,
                                  C.POINTER(c_vp), c_int, c_vp, c_vp, c_dbl, c_int, C.POINTER(C.c_int32),
                                  C.POINTER(c_dbl), C.POINTER(C.c_int32), c_vp, c_sz, c_vp]),
    "nb_lbfgs_ws_bytes": (c_sz, [c_i64, c_int]),
    "nb_lbfgs_reset": (c_int, [c_int, c_i64, c_int, c_vp, c_sz, c_vp]),
    "nb_lbfgs_direction": (c_int, [c_int, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "nb_lbfgs_trial": (c_int, [c_int, c_i64, c_vp, c_vp, c_dbl, c_vp, c_vp]),
    "nb_lbfgs_judge": (c_int, [c_int, c_i64, c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_dbl, c_dbl, c_int,
                               c_int, c_vp, c_vp]),
    "nb_glm_predict": (c_int, [c_int, c_int, c_vp, c_i64, c_i64, c_i64, c_vp, c_vp, c_vp]),
}


class NumsB200Error(RuntimeError):
    pass


def lib_path():
    return _build.LIB_PATH


def load():
    """Loads libnumsb200.so (once).  Raises if it has not been built."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = lib_path()
    if not os.path.exists(path):
        raise NumsB200Error(
            "libnumsb200.so not found at %s: build it with `python -m nums_b200.build` "
            "(there is no CPU fallback)" % path
        )
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)   # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _LIB = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().nb_last_error()
        raise NumsB200Error(msg.decode("utf-8", "replace") if msg else "libnumsb200 call failed (%d)" % rc)


def i64_array(values):
    values = list(values)
    return (C.c_int64 * max(1, len(values)))(*values)
