"""Generates tests/golden/*.npz by running the REFERENCE ITSELF (nums v0.2.8 at /root/reference) on its
NumPy SerialBackend.  Run in the build container only:  python tests/golden/make_golden.py

The reference does not import on this image as is (missing ray/opt_einsum/boto3/zarr; NumPy-2
removals), so — as SURVEY.md section 8c describes — it is copied to a temp dir (importing
nums.core.settings creates <root>/data), empty stub modules are provided for the absent third-party
packages, and the removed NumPy aliases are restored before `import nums`.  Nothing of the reference
is copied into this repository; only the produced input/output vectors are committed.
"""
import os
import shutil
import sys
import tempfile
import types
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


def bootstrap():
    tmp = tempfile.mkdtemp(prefix="nums_ref_")
    shutil.copytree(os.path.join(REF, "nums"), os.path.join(tmp, "nums"))
    stubs = os.path.join(tmp, "_stubs")
    for name, body in {
        "ray": "def is_initialized():\n    return False\ndef remote(*a, **k):\n    return (lambda f: f)\n",
        "opt_einsum": "import numpy as np\ncontract = np.einsum\n",
        "boto3": "", "botocore": "", "zarr": "",
    }.items():
        os.makedirs(os.path.join(stubs, name))
        with open(os.path.join(stubs, name, "__init__.py"), "w") as fh:
            fh.write(body)
    sys.path.insert(0, stubs)
    sys.path.insert(0, tmp)
    os.environ["NUMS_BACKEND"] = "serial"
    import numpy.ma  # noqa: F401  (must be imported before the aliases below are patched in)
    import scipy.special  # noqa: F401
    import scipy.linalg  # noqa: F401
    np.int = int
    np.object = object
    np.float = float
    np.product = np.prod
    np.NINF = -np.inf
    np.PZERO = 0.0
    np.NZERO = -0.0
    compat = types.ModuleType("numpy.compat")
    compat.asbytes = lambda s: s if isinstance(s, bytes) else str(s).encode("latin1")
    compat.asstr = lambda s: s.decode("latin1") if isinstance(s, bytes) else str(s)
    compat.asunicode = compat.asstr
    sys.modules["numpy.compat"] = compat
    np.compat = compat
    np.lib.stride_tricks.broadcast_to = np.broadcast_to
    return tmp


ONLY = [a.split("=", 1)[1] for a in sys.argv[1:] if a.startswith("--only=")]


def save(name, **arrays):
    if ONLY and name not in ONLY:
        return
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **arrays)
    print("wrote %s (%.1f KB)" % (path, os.path.getsize(path) / 1024))


def main():
    tmp = bootstrap()
    warnings.simplefilter("ignore")
    import nums
    import nums.numpy as nps
    from nums.core import application_manager
    from nums.core.kernel.kernel_manager import KernelManager
    from nums.core.kernel.numpy_kernel import KernelCls
    from nums.models.glms import LogisticRegression, LinearRegression, PoissonRegression, Ridge

    nums.init()
    app = application_manager.instance()
    K = KernelCls()

    # ---------------------------------------------------------------- cfg 1: README example (reduced n)
    nps.random.seed(1337)
    size = 20000
    X1 = nps.random.randn(size // 2, 2)
    y1 = nps.zeros(shape=(size // 2,), dtype=int)
    X2 = nps.random.randn(size // 2, 2) + 2.0
    y2 = nps.ones(shape=(size // 2,), dtype=int)
    X = nps.concatenate([X1, X2], axis=0)
    y = nps.concatenate([y1, y2], axis=0)
    model = LogisticRegression(solver="newton-cg", penalty="l2", C=10)
    model.fit(X, y)
    save("glm_readme", X=X.get(), y=y.get(), row_block=np.int64(X.block_shape[0]),
         beta=model._beta.get(), beta0=model._beta0.get(),
         pred=model.predict(X).get())

    # ---------------------------------------------------------------- logistic d=32, 4 row blocks
    rs = app.random_state(1337)      # block j draws from PCG64(1337).jumped(j)
    n, d, rb = 4096, 32, 1024
    Xb = rs.normal(shape=(n, d), block_shape=(rb, d))
    w_true = np.linspace(-1.0, 1.0, d)
    Xh = Xb.get()
    yh = (Xh @ w_true + 0.5 * np.sin(np.arange(n)) > 0).astype(np.int64)
    yb = app.array(yh, block_shape=(rb,))
    out = {"X": Xh, "y": yh, "row_block": np.int64(rb)}
    for tag, kw in {
        "newton_none": dict(solver="newton", penalty="none", max_iter=4),
        "newton_l2": dict(solver="newton", penalty="l2", C=2.0, max_iter=6),
        "lbfgs_l2": dict(solver="lbfgs", penalty="l2", C=2.0, max_iter=6),
        "gd_l2": dict(solver="gd", penalty="l2", C=2.0, max_iter=5, lr=1e-4),
    }.items():
        m = LogisticRegression(**kw)
        m.fit(Xb, yb)
        out[tag + "_beta"] = m._beta.get()
        out[tag + "_beta0"] = m._beta0.get()
    # first-iteration statistics at beta = 0.1 (g, H, objective) through the reference's own model methods
    m = LogisticRegression(solver="newton", penalty="none")
    m.fit(Xb, yb)  # sets nothing we need, but keeps penalty members consistent
    Xa = app.concatenate([Xb, app.ones(shape=(n, 1), block_shape=(rb, 1), dtype=Xb.dtype)], axis=1,
                         axis_block_size=Xb.block_shape[1])
    beta = app.array(np.full(d + 1, 0.1), block_shape=(d,))
    mu = m.forward(Xa, beta)
    out["stats_beta"] = beta.get()
    out["stats_g"] = m.gradient(Xa, yb, mu, beta=beta).get()
    out["stats_H"] = m.hessian(Xa, yb, mu).get()
    out["stats_obj"] = m.objective(Xa, yb, beta, mu).get()
    save("glm_logistic_d32", **out)

    # ---------------------------------------------------------------- the other solvers / penalties
    from nums.models.glms import Lasso, ElasticNet
    out = {}
    for tag, cls, kw in [
        ("irls", LogisticRegression, dict(solver="irls", penalty="none", max_iter=4)),
        ("irls_l2", LogisticRegression, dict(solver="irls", penalty="l2", C=2.0, max_iter=4)),
        ("block_sgd", LogisticRegression, dict(solver="block_sgd", penalty="l2", C=2.0, max_iter=3, lr=1e-4)),
        ("sgd", LogisticRegression, dict(solver="sgd", penalty="l2", C=2.0, max_iter=25, lr=1e-2, random_state=7)),
        ("newton_l1", LogisticRegression, dict(solver="newton", penalty="l1", C=5.0, max_iter=4)),
        ("newton_en", LogisticRegression, dict(solver="newton", penalty="elasticnet", C=5.0, max_iter=4)),
    ]:
        m = cls(**kw)
        m.fit(Xb, yb)
        out[tag + "_beta"] = m._beta.get()
        out[tag + "_beta0"] = m._beta0.get()
    m = LogisticRegression(solver="newton", penalty="l2", C=2.0, max_iter=6)
    m.fit(Xb, yb)
    out["proba"] = m.predict_proba(Xb).get()
    y_lin32 = Xh @ w_true + 0.1 * np.cos(np.arange(n))
    for tag, cls, kw in [("lasso", Lasso, dict(alpha=0.3, solver="newton", max_iter=3)),
                         ("enet", ElasticNet, dict(alpha=0.3, l1_ratio=0.4, solver="newton", max_iter=3))]:
        m = cls(**kw)
        m.fit(Xb, app.array(y_lin32, block_shape=(rb,)))
        out[tag + "_beta"] = m._beta.get()
        out[tag + "_beta0"] = m._beta0.get()
    out["y_lin32"] = y_lin32
    save("glm_solvers_d32", **out)

    # ---------------------------------------------------------------- linear / ridge / poisson, d=8
    n, d, rb = 3000, 8, 1000
    Xb = rs.normal(shape=(n, d), block_shape=(rb, d))
    Xh = Xb.get()
    coef = np.arange(1, d + 1) / 10.0
    y_lin = Xh @ coef + 0.3 + 0.01 * np.cos(np.arange(n))
    y_poi = np.exp(0.1 * Xh @ coef - 0.1)
    out = {"X": Xh, "y_lin": y_lin, "y_poi": y_poi, "row_block": np.int64(rb)}
    for tag, cls, yv, kw in [
        ("lin_newton", LinearRegression, y_lin, dict(solver="newton", max_iter=3)),
        ("lin_lbfgs", LinearRegression, y_lin, dict(solver="lbfgs", max_iter=8)),
        ("ridge_newton", Ridge, y_lin, dict(solver="newton", alpha=0.5, max_iter=3)),
        ("poi_newton", PoissonRegression, y_poi, dict(solver="newton", max_iter=6)),
        ("poi_lbfgs", PoissonRegression, y_poi, dict(solver="lbfgs", max_iter=6)),
    ]:
        m = cls(**kw)
        m.fit(Xb, app.array(yv, block_shape=(rb,)))
        out[tag + "_beta"] = m._beta.get()
        out[tag + "_beta0"] = m._beta0.get()
    save("glm_linear_poisson_d8", **out)

    # ---------------------------------------------------------------- block kernels, called directly
    i64 = np.int64
    ia = np.array([7, -7, 7, -7, 0, 5, i64(-2) ** 63, 3, 1, -8, 8, 12, -12, 2 ** 40], dtype=i64)
    ib = np.array([2, 2, -2, -2, 3, 0, -1, 41, 64, 64, 64, 18, 18, 3], dtype=i64)
    fa = np.array([1.5, -2.5, 0.0, -0.0, np.inf, -np.inf, np.nan, 7.25, -7.25, 1e-300, 3.0, 2.0])
    fb = np.array([2.0, 2.0, 0.0, 3.0, 1.0, -2.0, 1.0, -2.0, 2.0, 1e300, np.nan, 0.5])
    out = {"ia": ia, "ib": ib, "fa": fa, "fb": fb}
    int_ops = ["add", "sub", "mul", "truediv", "floor_divide", "mod", "fmod", "maximum", "minimum", "lt", "le",
               "gt", "ge", "eq", "ne", "bitwise_and", "bitwise_or", "bitwise_xor", "left_shift", "right_shift",
               "gcd", "lcm", "logical_and", "logical_or", "logical_xor"]
    with np.errstate(all="ignore"):
        for op in int_ops:
            out["i_" + op] = K.bop(op, ia, ib, False, False, {})
        out["i_pow"] = K.bop("pow", np.array([3, -2, 2, 7, 0], dtype=i64), np.array([41, 3, 62, 0, 0], dtype=i64),
                             False, False, {})
        flt_ops = ["add", "sub", "mul", "truediv", "floor_divide", "mod", "fmod", "pow", "maximum", "minimum", "fmax",
                   "fmin", "lt", "le", "gt", "ge", "eq", "ne", "arctan2", "hypot", "copysign", "nextafter",
                   "heaviside", "logaddexp", "logaddexp2", "xlogy", "logical_and", "logical_or", "logical_xor"]
        for op in flt_ops:
            out["f_" + op] = K.bop(op, fa, fb, False, False, {})
        out["mixed_add"] = K.bop("add", fa[:5], ia[:5], False, False, {})
        out["mixed_lt"] = K.bop("lt", ia[:5], fa[:5], False, False, {})
        out["ldexp"] = K.bop("ldexp", fa, np.arange(-6, 6, dtype=i64), False, False, {})
        ua = np.array([-2.5, -1.0, -0.0, 0.0, 0.3, 1.0, 2.5, 1e3, np.inf, -np.inf, np.nan, 0.5])
        out["ua"] = ua
        for op in ["abs", "negative", "positive", "sign", "square", "reciprocal", "sqrt", "cbrt", "exp", "exp2",
                   "expm1", "log", "log2", "log10", "log1p", "sin", "cos", "tan", "arcsin", "arccos", "arctan",
                   "sinh", "cosh", "tanh", "arcsinh", "arccosh", "arctanh", "ceil", "floor", "trunc", "rint",
                   "deg2rad", "rad2deg", "isnan", "isinf", "isfinite", "signbit", "logical_not", "spacing", "fabs",
                   "degrees", "radians"]:
            out["u_" + op] = K.map_uop(op, ua, (), {})
        ui = np.array([-5, -1, 0, 1, 7, i64(-2) ** 63, 2 ** 62], dtype=i64)
        out["ui"] = ui
        for op in ["abs", "negative", "sign", "square", "invert", "sqrt", "exp", "isnan", "logical_not", "ceil",
                   "reciprocal"]:
            out["ui_" + op] = K.map_uop(op, ui, (), {})
    # broadcasting shapes of the Newton loop
    col = np.arange(6, dtype=float).reshape(6, 1) + 1
    mat = np.arange(24, dtype=float).reshape(6, 4) - 7
    out["bc_col_mat"] = K.bop("mul", col, mat, False, False, {})
    out["bc_scalar"] = K.bop("truediv", np.float64(1.0), mat[:, 0] + 8, False, False, {})
    out["bc_T"] = K.bop("add", mat, mat.T.copy(), False, True, {})
    out["col"], out["mat"] = col, mat
    # reductions incl. quirk Q1
    imat = (np.arange(12, dtype=i64).reshape(6, 2) * 3 - 7)
    out["imat"] = imat
    out["red_sum_ax0_int"] = K.reduce_axis("sum", imat, 0, False, False)        # float64 (Q1)
    out["red_sum_ax1_int"] = K.reduce_axis("sum", imat, 1, False, False)
    out["red_sum_all_int"] = K.reduce_axis("sum", imat, None, False, False)
    out["red_sum_ax0_keep"] = K.reduce_axis("sum", imat, 0, True, False)
    out["red_sum_ax0_T"] = K.reduce_axis("sum", imat, 0, False, True)
    out["red_max_ax1"] = K.reduce_axis("max", mat, 1, False, False)
    out["red_min_all"] = K.reduce_axis("min", mat, None, False, False)
    nanv = np.array([1.0, np.nan, -3.0, 2.0])
    out["nanv"] = nanv
    out["red_max_nan"] = K.reduce_axis("max", nanv, None, False, False)
    out["red_nanmax"] = K.reduce_axis("nanmax", nanv, None, False, False)
    out["red_nansum"] = K.reduce_axis("nansum", nanv, None, False, False)
    t3 = np.arange(60, dtype=float).reshape(3, 4, 5)
    out["t3"] = t3
    out["red_t3_ax1"] = K.reduce_axis("sum", t3, 1, False, False)
    out["red_any"] = K.reduce_axis("any", mat > 10, 0, False, False)
    out["red_all"] = K.reduce_axis("all", mat > -8, None, False, False)
    out["bopred_sum"] = K.bop_reduce("sum", mat, mat * 2, False, False)
    out["bopred_max"] = K.bop_reduce("max", nanv, nanv[::-1].copy(), False, False)
    out["sumred"] = K.sum_reduce(mat, mat * 2, mat * 3)
    # arg_op chain, quirk Q2: blocks [5,1] and [5,0] -> 2, not 0
    b0, b1 = np.array([5.0, 1.0]), np.array([5.0, 0.0])
    r0 = K.arg_op("argmax", b0, slice(0, 2))
    r1 = K.arg_op("argmax", b1, slice(2, 4), *r0)
    out["arg_chain"] = np.array([r0[0], r1[0]], dtype=i64)
    out["arg_chain_val"] = np.array([r0[1], r1[1]])
    nanb = np.array([1.0, np.nan, 7.0, np.nan])
    out["arg_nan"] = np.array([K.arg_op("argmax", nanb, slice(0, 4))[0], K.arg_op("argmin", nanb, slice(0, 4))[0]],
                              dtype=i64)
    r0 = K.arg_op("argmin", np.array([4, 2, 2], dtype=i64), slice(10, 13))
    r1 = K.arg_op("argmin", np.array([9, 2, 3], dtype=i64), slice(13, 16), *r0)
    out["arg_chain_min_int"] = np.array([r0[0], r1[0]], dtype=i64)
    # where
    cond = (np.arange(30).reshape(5, 6) % 4 == 1)
    w = K.where(cond, None, None, [(10, 15), (100, 106)])
    out["where_cond"] = cond
    out["where_rows"], out["where_cols"], out["where_count"] = w[0], w[1], np.array(w[2], dtype=i64)
    out["where3"] = K.where(cond, mat.reshape(-1)[:30].reshape(5, 6) if mat.size >= 30 else None,
                            -np.ones((5, 6)), None) if mat.size >= 30 else np.zeros(1)
    xs, ys = np.arange(30, dtype=float).reshape(5, 6), -np.arange(30, dtype=float).reshape(5, 6)
    out["where3"] = K.where(cond, xs, ys, None)
    # tensordot / inv
    A = np.linspace(-1, 1, 7 * 5).reshape(7, 5)
    Bm = np.linspace(2, 3, 5 * 3).reshape(5, 3)
    out["A"], out["Bm"] = A, Bm
    out["mm"] = K.bop("tensordot", A, Bm, False, False, 1)
    out["mm_T"] = K.bop("tensordot", A, A, True, False, 1)
    out["mv"] = K.bop("tensordot", A, Bm[:, 0].copy(), False, False, 1)
    out["vm_T"] = K.bop("tensordot", A, np.arange(7, dtype=float), True, False, 1)
    out["dot"] = K.bop("tensordot", A[:, 0].copy(), A[:, 1].copy(), False, False, 1)
    out["imm"] = K.bop("tensordot", imat.T.copy(), imat, False, False, 1)
    S = A.T @ A + np.eye(5)
    out["S"], out["S_inv"] = S, K.inv(S)
    save("kernels", **out)

    # ---------------------------------------------------------------- BlockArray level (cfg 2 scaled down)
    nps.random.seed(1337)
    n = 100000
    rs2 = app.random_state(1337)
    x = rs2.random(shape=(n,), block_shape=(n // 4,))
    dot = (x.T @ x).get()
    p = rs2.permutation(n, n // 4)
    idx = app.where(p == n // 2)
    s = app.sum(p)
    am, amin = app.argop("argmax", p), app.argop("argmin", x)
    save("blockarray", x=x.get(), p=p.get(), dot=np.float64(dot), where_idx=idx[0].get(),
         sum_p=np.asarray(s.get()), argmax_p=np.asarray(am.get()), argmin_x=np.asarray(amin.get()),
         max_x=np.asarray(app.max(x).get()), mean_x=np.asarray(app.mean(x).get()),
         row_block=np.int64(n // 4))

    # ---------------------------------------------------------------- tall-skinny linear algebra (linalg.py)
    from nums.core import linalg
    rs3 = app.random_state(2024)
    nL, dL, rbL, cbL = 2345, 9, 123, 4                 # the shapes of tests/core/test_linalg.py:155-181
    XL = rs3.normal(shape=(nL, dL), block_shape=(rbL, cbL)) + 0.25
    XLh = XL.get()
    thetaL = np.linspace(0.5, 1.5, dL)
    yLh = XLh @ thetaL + 0.01 * np.cos(np.arange(nL))
    yL = app.array(yLh, block_shape=(rbL,))
    outL = {"X": XLh, "y": yLh, "row_block": np.int64(rbL), "col_block": np.int64(cbL)}
    outL["tsr_R"] = linalg.indirect_tsr(app, XL).get()
    q, r = linalg.indirect_tsqr(app, XL)
    outL["itsqr_Q"], outL["itsqr_R"] = q.get(), r.get()
    q, r = linalg.direct_tsqr(app, XL)
    outL["dtsqr_Q"], outL["dtsqr_R"] = q.get(), r.get()
    u, sv, vt = linalg.svd(app, XL)
    outL["svd_U"], outL["svd_S"], outL["svd_VT"] = u.get(), sv.get(), vt.get()
    outL["cov"] = app.cov(XL, rowvar=False).get()
    outL["pca"] = linalg.pca(app, XL).get()
    outL["lr_theta"] = linalg.linear_regression(app, XL, yL).get()
    outL["fast_lr_theta"] = linalg.fast_linear_regression(app, XL, yL).get()
    outL["rr_theta_0"] = linalg.ridge_regression(app, XL, yL, lamb=0.0).get()
    outL["rr_theta_10"] = linalg.ridge_regression(app, XL, yL, lamb=10.0).get()
    # single column block, ragged last row block (the tall-skinny layout of the GLM path)
    XT = rs3.normal(shape=(2100, 32), block_shape=(512, 32))
    outL["X_tall"] = XT.get()
    q, r = linalg.direct_tsqr(app, XT)
    outL["tall_dtsqr_Q"], outL["tall_dtsqr_R"] = q.get(), r.get()
    outL["tall_tsr_R"] = linalg.indirect_tsr(app, XT).get()
    # SPD matrix: cholesky / inv; upper-triangular blocked inverse (tests/core/test_linalg.py:128-152)
    Ah = XLh.T @ XLh / nL + np.eye(dL)
    A = app.array(Ah, block_shape=(cbL, cbL))
    outL["spd"] = Ah
    outL["chol_L"] = linalg.cholesky(app, A).get()
    outL["inv_spd"] = linalg.inv(app, A).get()
    Rup = np.triu(np.arange(1, 401, dtype=float).reshape(20, 20))
    outL["uppertri"] = Rup
    outL["inv_uppertri"] = linalg.inv_uppertri(app, app.array(Rup, block_shape=(3, 3))).get()
    save("linalg", **outL)

    # ---------------------------------------------------------------- CSV ingest (filesystem.py:227-283, 615-657)
    from nums.core import filesystem as ref_fs
    from nums.core.storage import utils as storage_utils
    rngc = np.random.default_rng(77)
    vals = rngc.standard_normal((257, 5)) * 10.0 ** rngc.integers(-30, 30, size=(257, 5))
    rows = []
    for i, r in enumerate(vals):
        fmt = ["%.18e", "%r", "%.6f", "%g", "%.17g"][i % 5]
        rows.append(",".join((fmt % v) if fmt != "%r" else repr(float(v)) for v in r))
    rows[3] = " 1.5 ,\t-2e3,+.5,7.,1_000.25"                      # whitespace, signs, bare dots, underscores
    rows[4] = "inf,-Infinity,nan,0.0,-0.0"
    rows[5] = "4.9e-324,2.2250738585072011e-308,1.7976931348623157e308,1e309,1e-400"
    rows[6] = "9007199254740993,9007199254740992.5,0.1000000000000000055511151231257827,123456789012345678901234567890,5e-1"
    csv_text = "c0,c1,c2,c3,c4\n" + "\n".join(rows) + "\n"
    csv_path = os.path.join(tmp, "golden.csv")
    with open(csv_path, "w", encoding="utf-8") as fh:
        fh.write(csv_text)
    size = storage_utils.get_file_size(csv_path)
    outC = {"csv_bytes": np.frombuffer(csv_text.encode(), dtype=np.uint8).copy()}
    for nw in (1, 4, 7):
        bts = storage_utils.Batch.from_num_batches(size, nw).batches
        outC["batches_%d" % nw] = np.array(bts, dtype=i64)
        for bi, (s0, e0) in enumerate(bts):
            arr, shp = ref_fs.read_csv_block(csv_path, s0, e0, float, ",", True)
            outC["w%d_b%d" % (nw, bi)] = arr
    # integer file, no header, dtype np.int64 and python int
    int_rows = [",".join(str(int(v)) for v in rngc.integers(-10 ** 15, 10 ** 15, size=3)) for _ in range(40)]
    int_text = "\n".join(int_rows) + "\n"
    int_path = os.path.join(tmp, "golden_int.csv")
    with open(int_path, "w", encoding="utf-8") as fh:
        fh.write(int_text)
    outC["int_bytes"] = np.frombuffer(int_text.encode(), dtype=np.uint8).copy()
    arr, _ = ref_fs.read_csv_block(int_path, 0, len(int_text), np.int64, ",", False)
    outC["int64_all"] = arr
    dec_rows = ["%.3f,%.1f,%d" % (a, b, c) for a, b, c in zip(rngc.standard_normal(30) * 100, rngc.standard_normal(30) * 1e6,
                                                               rngc.integers(-99, 99, size=30))]
    dec_text = "\n".join(dec_rows) + "\n"
    dec_path = os.path.join(tmp, "golden_dec.csv")
    with open(dec_path, "w", encoding="utf-8") as fh:
        fh.write(dec_text)
    outC["dec_bytes"] = np.frombuffer(dec_text.encode(), dtype=np.uint8).copy()
    arr, _ = ref_fs.read_csv_block(dec_path, 0, len(dec_text), int, ",", False)     # converter: int(float(x))
    outC["pyint_all"] = arr
    save("csv", **outC)

    # ---------------------------------------------------------------- block-shape rule
    cases = [((10 ** 8,), "float64", (8, 1), 8), ((10 ** 8, 256), "float64", (8, 1), 8),
             ((10 ** 8, 256), "float64", (2, 1), 2), ((10 ** 4, 10 ** 4), "float64", (4, 1), 4),
             ((10 ** 6, 2), "float64", (1, 1), 1), ((10 ** 8, 1024), "float64", (8, 1), 8),
             ((10 ** 8,), "int64", (4, 1), 4), ((5 * 10 ** 7, 3), "float64", (3, 2), 6)]
    shapes = [KernelManager.compute_block_shape_static(sh, getattr(np, dt), cs, nc) for sh, dt, cs, nc in cases]
    save("block_shapes",
         case_shape=np.array([list(c[0]) + [0] * (2 - len(c[0])) for c in cases], dtype=i64),
         case_ndim=np.array([len(c[0]) for c in cases], dtype=i64),
         case_dtype=np.array([c[1] for c in cases]),
         case_cluster=np.array([list(c[2]) for c in cases], dtype=i64),
         case_cores=np.array([c[3] for c in cases], dtype=i64),
         block_shape=np.array([list(s) + [0] * (2 - len(s)) for s in shapes], dtype=i64))
    shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    main()
