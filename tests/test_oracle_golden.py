"""Pins the oracle (oracle/) against vectors produced by the reference itself (tests/golden)."""
import numpy as np
import pytest

from conftest import load_golden, rel_err
from oracle import glm as oglm
from oracle import np_kernel as ok

# float64 results of the oracle and of the reference come from the same NumPy calls in the same
# order, so they are expected to agree to the last bit; the tolerance only absorbs BLAS threading.
TOL = 1e-12


def test_readme_logistic_newton():
    g = load_golden("glm_readme")
    beta, it = oglm.fit("logistic", g["X"], g["y"], int(g["row_block"]), solver="newton-cg", penalty="l2",
                        lam=1.0 / 10)
    assert rel_err(beta[:-1], g["beta"]) < TOL
    assert abs(beta[-1] - g["beta0"]) < TOL * abs(g["beta0"])
    assert it < 100


@pytest.mark.parametrize("tag,kw", [
    ("newton_none", dict(solver="newton", penalty=None, max_iter=4)),
    ("newton_l2", dict(solver="newton", penalty="l2", lam=0.5, max_iter=6)),
    ("lbfgs_l2", dict(solver="lbfgs", penalty="l2", lam=0.5, max_iter=6)),
    ("gd_l2", dict(solver="gd", penalty="l2", lam=0.5, max_iter=5, lr=1e-4)),
])
def test_logistic_d32(tag, kw):
    g = load_golden("glm_logistic_d32")
    beta, _ = oglm.fit("logistic", g["X"], g["y"], int(g["row_block"]), **kw)
    assert rel_err(beta[:-1], g[tag + "_beta"]) < 1e-10
    assert abs(beta[-1] - g[tag + "_beta0"]) < 1e-10 * max(1.0, abs(g[tag + "_beta0"]))


def test_logistic_stats_d32():
    g = load_golden("glm_logistic_d32")
    rb = int(g["row_block"])
    m = oglm.BlockedGLM("logistic", oglm.split_rows(g["X"], rb), oglm.split_rows(g["y"], rb))
    beta = g["stats_beta"]
    fwd = m.forward(beta)
    assert rel_err(m.gradient(beta, fwd), g["stats_g"]) < TOL
    assert rel_err(m.hessian(beta, fwd), g["stats_H"]) < TOL
    assert rel_err(m.objective(beta, fwd), g["stats_obj"]) < TOL


@pytest.mark.parametrize("tag,family,ykey,kw", [
    ("lin_newton", "linear", "y_lin", dict(solver="newton", max_iter=3)),
    ("lin_lbfgs", "linear", "y_lin", dict(solver="lbfgs", max_iter=8)),
    ("ridge_newton", "linear", "y_lin", dict(solver="newton", penalty="l2", lam=0.5, max_iter=3)),
    ("poi_newton", "poisson", "y_poi", dict(solver="newton", max_iter=6)),
    ("poi_lbfgs", "poisson", "y_poi", dict(solver="lbfgs", max_iter=6)),
])
def test_linear_poisson(tag, family, ykey, kw):
    g = load_golden("glm_linear_poisson_d8")
    beta, _ = oglm.fit(family, g["X"], g[ykey], int(g["row_block"]), **kw)
    assert rel_err(beta[:-1], g[tag + "_beta"]) < 1e-9
    assert abs(beta[-1] - g[tag + "_beta0"]) < 1e-9 * max(1.0, abs(g[tag + "_beta0"]))


INT_OPS = ["add", "sub", "mul", "truediv", "floor_divide", "mod", "fmod", "maximum", "minimum", "lt", "le", "gt",
           "ge", "eq", "ne", "bitwise_and", "bitwise_or", "bitwise_xor", "left_shift", "right_shift", "gcd", "lcm",
           "logical_and", "logical_or", "logical_xor"]


def test_kernels_match_reference():
    g = load_golden("kernels")
    with np.errstate(all="ignore"):
        for op in INT_OPS:
            r = ok.bop(op, g["ia"], g["ib"])
            assert r.dtype == g["i_" + op].dtype, op
            np.testing.assert_array_equal(r, g["i_" + op], err_msg=op)
        for key in [k for k in g if k.startswith("f_")]:
            np.testing.assert_array_equal(ok.bop(key[2:], g["fa"], g["fb"]), g[key], err_msg=key)
        for key in [k for k in g if k.startswith("u_")]:
            np.testing.assert_array_equal(ok.map_uop(key[2:], g["ua"]), g[key], err_msg=key)
    np.testing.assert_array_equal(ok.reduce_axis("sum", g["imat"], 0, False, False), g["red_sum_ax0_int"])
    assert ok.reduce_axis("sum", g["imat"], 0, False, False).dtype == np.float64      # quirk Q1
    assert ok.reduce_axis("sum", g["imat"], None, False, False).dtype == np.int64
    r0 = ok.arg_op("argmax", np.array([5.0, 1.0]), slice(0, 2))
    r1 = ok.arg_op("argmax", np.array([5.0, 0.0]), slice(2, 4), *r0)
    assert [r0[0], r1[0]] == list(g["arg_chain"]) == [0, 2]                             # quirk Q2
    w = ok.where(g["where_cond"], None, None, [(10, 15), (100, 106)])
    np.testing.assert_array_equal(w[0], g["where_rows"])
    np.testing.assert_array_equal(w[1], g["where_cols"])
    np.testing.assert_array_equal(ok.bop("tensordot", g["A"], g["A"], True, False, 1), g["mm_T"])
    np.testing.assert_array_equal(ok.inv(g["S"]), g["S_inv"])


def test_blockarray_level():
    g = load_golden("blockarray")
    rb = int(g["row_block"])
    xb = oglm.split_rows(g["x"], rb)
    dot = oglm.tree_sum([np.matmul(b, b) for b in xb])
    assert rel_err(dot, g["dot"]) < TOL
    pb = oglm.split_rows(g["p"], rb)
    hits = []
    for i, b in enumerate(pb):
        r = ok.where(b == len(g["p"]) // 2, None, None, [(i * rb, (i + 1) * rb)])
        hits.append(r[0])
    np.testing.assert_array_equal(np.concatenate(hits), g["where_idx"])
    red = (None, None)
    for i, b in enumerate(pb):
        red = ok.arg_op("argmax", b, slice(i * rb, (i + 1) * rb), *red)
    assert red[0] == g["argmax_p"]
    assert oglm.tree_sum([np.sum(b) for b in pb]) == g["sum_p"]


@pytest.mark.parametrize("tag,kw", [
    ("irls", dict(solver="irls", penalty=None, max_iter=4)),
    ("irls_l2", dict(solver="irls", penalty="l2", lam=0.5, max_iter=4)),
    ("block_sgd", dict(solver="block_sgd", penalty="l2", lam=0.5, max_iter=3, lr=1e-4)),
    ("sgd", dict(solver="sgd", penalty="l2", lam=0.5, max_iter=25, lr=1e-2, seed=7)),
    ("newton_l1", dict(solver="newton", penalty="l1", lam=0.2, max_iter=4)),
    ("newton_en", dict(solver="newton", penalty="elasticnet", lam=0.2, max_iter=4)),
])
def test_other_solvers_d32(tag, kw):
    g, s = load_golden("glm_logistic_d32"), load_golden("glm_solvers_d32")
    beta, _ = oglm.fit("logistic", g["X"], g["y"], int(g["row_block"]), **kw)
    assert rel_err(beta[:-1], s[tag + "_beta"]) < 1e-9
    assert abs(beta[-1] - s[tag + "_beta0"]) < 1e-9 * max(1.0, abs(s[tag + "_beta0"]))


@pytest.mark.parametrize("tag,kw", [("lasso", dict(penalty="l1", lam=0.3, max_iter=3)),
                                    ("enet", dict(penalty="elasticnet", lam=0.3, l1_ratio=0.4, max_iter=3))])
def test_lasso_elasticnet_d32(tag, kw):
    g, s = load_golden("glm_logistic_d32"), load_golden("glm_solvers_d32")
    beta, _ = oglm.fit("linear", g["X"], s["y_lin32"], int(g["row_block"]), solver="newton", **kw)
    assert rel_err(beta[:-1], s[tag + "_beta"]) < 1e-9


# ---------------------------------------------------------------- tall-skinny linear algebra (nums/core/linalg.py)
def test_linalg_oracle_matches_reference_outputs():
    """oracle/linalg.py against the reference's own outputs (tests/golden/linalg.npz, shapes of the reference's
    tests/core/test_linalg.py:155-181): identical LAPACK calls on identical blocks, so agreement is at rounding
    level -- only the shape a BLAS product is issued in differs."""
    from oracle import linalg as ol
    g = load_golden("linalg")
    X, y, rb, cb = g["X"], g["y"], int(g["row_block"]), int(g["col_block"])
    tol = 1e-12
    assert rel_err(ol.indirect_tsr(X, rb, cb), g["tsr_R"]) <= tol
    for fn, tag in ((ol.indirect_tsqr, "itsqr"), (ol.direct_tsqr, "dtsqr")):
        Q, R = fn(X, rb, cb)
        assert rel_err(Q, g[tag + "_Q"]) <= tol and rel_err(R, g[tag + "_R"]) <= tol
        assert rel_err(Q @ R, X) <= 1e-13
    U, S, VT = ol.svd(X, rb, cb)
    assert rel_err(U, g["svd_U"]) <= tol and rel_err(S, g["svd_S"]) <= tol and rel_err(VT, g["svd_VT"]) <= tol
    assert rel_err(ol.cov(X, rb, cb), g["cov"]) <= tol
    assert rel_err(ol.pca(X, rb, cb), g["pca"]) <= 1e-11
    assert rel_err(ol.linear_regression(X, y, rb, cb), g["lr_theta"]) <= tol
    assert rel_err(ol.fast_linear_regression(X, y, rb, cb), g["fast_lr_theta"]) <= tol
    assert rel_err(ol.ridge_regression(X, y, 0.0, rb, cb), g["rr_theta_0"]) <= tol
    assert rel_err(ol.ridge_regression(X, y, 10.0, rb, cb), g["rr_theta_10"]) <= tol
    XT = g["X_tall"]
    Q, R = ol.direct_tsqr(XT, 512, 32)
    assert rel_err(Q, g["tall_dtsqr_Q"]) <= tol and rel_err(R, g["tall_dtsqr_R"]) <= tol
    assert rel_err(ol.indirect_tsr(XT, 512, 32), g["tall_tsr_R"]) <= tol
    assert rel_err(ol.inv_uppertri(g["uppertri"], 3), g["inv_uppertri"]) <= tol
    assert rel_err(np.linalg.cholesky(g["spd"]), g["chol_L"]) <= tol


# ---------------------------------------------------------------- CSV ingest (nums/core/filesystem.py:227-283)
def test_csv_oracle_matches_reference_blocks():
    """oracle/csvio.py against the reference's own read_csv_block on every byte range of 1, 4 and 7 workers
    (tests/golden/csv.npz): same lines per block, same float()/int() results, bit for bit."""
    from oracle import csvio
    g = load_golden("csv")
    data = g["csv_bytes"].tobytes()
    for nw in (1, 4, 7):
        b = csvio.batches(len(data), nw)
        assert np.array_equal(np.array(b), g["batches_%d" % nw])
        for bi, (s, e) in enumerate(b):
            got = csvio.read_csv_block(data, s, e, float, ",", True)
            want = g["w%d_b%d" % (nw, bi)]
            assert got.shape == want.shape
            if want.size:
                assert np.array_equal(got.view(np.uint64), want.view(np.uint64))
        assert csvio.read_csv(data, float, ",", True, nw).shape == (257, 5)
    assert np.array_equal(csvio.read_csv_block(g["int_bytes"].tobytes(), 0, len(g["int_bytes"]), np.int64), g["int64_all"])
    assert np.array_equal(csvio.read_csv_block(g["dec_bytes"].tobytes(), 0, len(g["dec_bytes"]), int), g["pyint_all"])


def cfg4_inputs(g):
    """The 2*10^5 x 256 slice of BASELINE config 4 as the reference drew it (tests/golden/make_golden_cfg4.py):
    nps.random.seed(1337), row block j from Generator(PCG64(1337).jumped(j)).normal, second block shifted."""
    n, d = int(g["cfg4_n"]), int(g["cfg4_d"])
    half = n // 2
    X = np.empty((n, d))
    for j in range(2):
        rng = np.random.Generator(np.random.PCG64(1337).jumped(j))
        X[j * half:(j + 1) * half] = rng.normal(0.0, 1.0, half * d).reshape(half, d)
    X[half:] += 2.0 * np.sqrt(2.0 / d)
    y = np.concatenate([np.zeros(half, dtype=np.int64), np.ones(half, dtype=np.int64)])
    return X, y


def test_cfg4_slice_d256_fit_iterations_and_first_statistics():
    """The benchmarked shape (d = 256): coefficients, the NUMBER of Newton iterations and g / H / objective at beta = 0
    of the reference itself."""
    g = load_golden("glm_cfg4_d256")
    X, y = cfg4_inputs(g)
    assert abs(float(X.sum()) - float(g["cfg4_x_checksum"])) <= 1e-9 * abs(float(g["cfg4_x_checksum"]))
    rb = int(g["cfg4_row_blocks"][0])
    beta, it = oglm.fit("logistic", X, y, rb, solver="newton-cg", penalty="l2", lam=1.0 / 10)
    assert it == int(g["cfg4_iters"])
    assert np.max(np.abs(beta[:-1] - g["cfg4_beta"])) < 1e-10 and abs(beta[-1] - g["cfg4_beta0"]) < 1e-10
    m = oglm.BlockedGLM("logistic", oglm.split_rows(X, rb), oglm.split_rows(y, rb))
    b0 = np.zeros(X.shape[1] + 1)
    fwd = m.forward(b0)
    assert rel_err(m.gradient(b0, fwd), g["cfg4_g0"]) < TOL
    assert rel_err(m.hessian(b0, fwd), g["cfg4_H0"]) < TOL
    assert rel_err(m.objective(b0, fwd), g["cfg4_obj0"]) < TOL


def test_readme_iteration_counts():
    g = load_golden("glm_cfg4_d256")
    r = load_golden("glm_readme")
    _, it = oglm.fit("logistic", r["X"], r["y"], int(r["row_block"]), solver="newton-cg", penalty="l2", lam=1.0 / 10)
    assert it == int(g["readme_20000_iters"]) == 8
    assert int(g["readme_1000000_iters"]) == 8      # what SURVEY.md 8(d) probed for the README's own size


@pytest.mark.parametrize("nplanes,no_level5,bound", [(5, False, 2.0 ** -36), (5, True, 2.0 ** -30), (4, True, 8.0 * 2.0 ** (3 - 31))])
def test_exact_int8_gram_restatement_against_float64(nplanes, no_level5, bound):
    """oracle/int8_gram.py (the integer restatement of the device's digit pipeline for the linear family) against the
    float64 Gram matrix X^T X of glms.py:318-327: within the error model of each variant, exactly symmetric, an
    all-zero column gives exact zeros, and splitting the rows into chunks changes only the dither (four planes)."""
    from oracle import int8_gram as ig
    rng = np.random.default_rng(3)
    n, d = 2500, 132
    X = rng.standard_normal((n, d)) / 7.0
    X[:, 3] = 0.0
    G = X.T @ X
    s = np.sqrt(np.maximum(np.diag(G), 1e-300))
    s[3] = 1.0
    H = ig.exact_gram_linear(X, nplanes, no_level5)
    scale = bound / np.sqrt(n) if nplanes == 4 or no_level5 else bound
    assert np.max(np.abs(H - G) / np.outer(s, s)) < scale
    assert np.array_equal(H, H.T) and np.all(H[3] == 0.0) and np.all(H[:, 3] == 0.0)
    Hc = ig.exact_gram_linear(X, nplanes, no_level5, chunk_rows=1024)
    if nplanes == 5:
        assert np.array_equal(H, Hc)
    else:
        assert np.max(np.abs(Hc - G) / np.outer(s, s)) < scale
    planes = ig.digit_planes(X, nplanes)
    assert planes.shape == (nplanes, n, d) and planes.min() >= -128 and planes.max() <= 127
