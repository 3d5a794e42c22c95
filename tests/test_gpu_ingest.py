"""GPU parity tests of the ingest path (nums_b200/filesystem.py -> csrc/csv.cu) against the reference's own
read_csv_block outputs (tests/golden/csv.npz), the oracle restatement, and Python's float() / int() bit for bit."""
import os

import numpy as np
import pytest

from conftest import load_golden
from oracle import csvio

pytestmark = pytest.mark.gpu


def bits(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a.view(np.uint64)


def down(b):
    from nums_b200 import device as dv
    return dv.to_numpy(b)


@pytest.fixture(scope="module")
def G():
    return load_golden("csv")


def write(tmp_path, name, data):
    p = os.path.join(str(tmp_path), name)
    with open(p, "wb") as fh:
        fh.write(data)
    return p


def test_read_csv_block_matches_reference_blocks_bit_for_bit(app, G, tmp_path):
    from nums_b200 import filesystem as fs
    path = write(tmp_path, "golden.csv", G["csv_bytes"].tobytes())
    for nw in (1, 4, 7):
        assert np.array_equal(np.array(fs.get_batches(os.path.getsize(path), nw)), G["batches_%d" % nw])
        for bi, (s, e) in enumerate(G["batches_%d" % nw]):
            blk, shape = fs.read_csv_block(path, int(s), int(e), float, ",", True, device_index=0)
            want = G["w%d_b%d" % (nw, bi)]
            assert tuple(shape) == want.shape
            if want.size:
                assert np.array_equal(bits(down(blk)), bits(want)), (nw, bi)


def test_read_csv_end_to_end_and_block_boundaries(app, G, tmp_path):
    data = G["csv_bytes"].tobytes()
    path = write(tmp_path, "golden.csv", data)
    for nw in (1, 2, 4, 7, 16):
        got = app.read_csv(path, dtype=float, delimiter=",", has_header=True, num_workers=nw)
        want = csvio.read_csv(data, float, ",", True, nw)
        assert got.shape == want.shape
        assert np.array_equal(bits(got.get()), bits(want)), nw
    # a line that starts exactly on a block boundary belongs to no block in the reference; the device path agrees
    text = b"".join(b"%d,%d\n" % (i, i * i) for i in range(100, 200))      # 100 lines of 10 bytes
    path2 = write(tmp_path, "aligned.csv", text)
    for nw in (4, 10):
        want = csvio.read_csv(text, float, ",", False, nw)
        got = app.read_csv(path2, num_workers=nw)
        assert want.shape[0] < 100 and got.shape == want.shape and np.array_equal(got.get(), want)
    # CRLF line ends and a file without a final newline
    crlf = b"1.5,2.5\r\n3.5,4.5\r\n5.5,6.5"
    got = app.read_csv(write(tmp_path, "crlf.csv", crlf), num_workers=1).get()
    assert np.array_equal(got, np.array([[1.5, 2.5], [3.5, 4.5], [5.5, 6.5]]))


def test_integer_dtypes(app, G, tmp_path):
    from nums_b200 import filesystem as fs
    p = write(tmp_path, "int.csv", G["int_bytes"].tobytes())
    blk, shape = fs.read_csv_block(p, 0, os.path.getsize(p), np.int64, ",", False, device_index=0)
    got = down(blk)
    assert got.dtype == np.int64 and np.array_equal(got, G["int64_all"])
    p2 = write(tmp_path, "dec.csv", G["dec_bytes"].tobytes())
    blk, _ = fs.read_csv_block(p2, 0, os.path.getsize(p2), int, ",", False, device_index=0)      # int(float(x))
    assert np.array_equal(down(blk), G["pyint_all"])
    edge = b"9223372036854775807,-9223372036854775808, +1_000 \n"
    blk, _ = fs.read_csv_block(write(tmp_path, "edge.csv", edge), 0, len(edge), np.int64, ",", False, device_index=0)
    assert np.array_equal(down(blk), np.array([[2 ** 63 - 1, -2 ** 63, 1000]], dtype=np.int64))
    for bad in (b"9223372036854775808\n", b"1.0\n", b"12a\n", b"1__0\n"):
        with pytest.raises(ValueError):
            fs.read_csv_block(write(tmp_path, "bad.csv", bad), 0, len(bad), np.int64, ",", False, device_index=0)


def test_float_conversion_equals_python_float_on_random_and_hard_decimals(app, tmp_path):
    """Every field is float(text), correctly rounded: shortest repr, %.18e, fixed notation, subnormals, the largest
    finite values, overflow / underflow, exact halfway cases (round half to even) and > 19-digit expansions of doubles."""
    from nums_b200 import filesystem as fs
    rng = np.random.default_rng(2024)
    raw = rng.integers(0, 2 ** 63 - 1, size=60_000, dtype=np.int64).view(np.float64)
    raw = raw[np.isfinite(raw)]
    sub = (rng.integers(1, 2 ** 52, size=2_000, dtype=np.int64)).view(np.float64)            # subnormals
    vals = np.concatenate([raw, -raw[:5000], sub, rng.standard_normal(20_000), rng.standard_normal(5_000) * 1e-310,
                           np.array([0.0, -0.0, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308])])
    strs = []
    for i, v in enumerate(vals):
        k = i % 6
        v = float(v)
        strs.append(repr(v) if k == 0 else ("%.18e" % v) if k == 1 else ("%.16e" % v) if k == 2 else
                    ("%.25e" % v) if k == 3 else ("%.12g" % v) if k == 4 else ("%.3e" % v).upper())
    halfway = ["9007199254740993", "9007199254740995", "18014398509481990", "4503599627370497.5", "1e23", "8.5e22",
               "9.5e21", "2.5e-1", "1125899906842624.125", "0.000000000000000000000000000000000000001e39",
               "1e22", "1e-22", "123456789012345678", "1234567890123456789", "17976931348623157e292", "1e308", "1e-323",
               "2.47e-324", "2.48e-324", "1.7976931348623158e308", "1.7976931348623159e308", "0e999", "0.0e-999",
               "1e400", "-1e400", "1e-400", "00012.5000", ".5", "5.", "+.5e+1", "1_0.2_5e1_0", " 7.25\t", "1E5", "nan",
               "-inf", "Infinity", "iNf"]
    strs += halfway
    while len(strs) % 8:
        strs.append("1")
    want = np.array([float(s) for s in strs]).reshape(-1, 8)
    text = ("\n".join(",".join(strs[i:i + 8]) for i in range(0, len(strs), 8)) + "\n").encode()
    blk, shape = fs.read_csv_block(write(tmp_path, "hard.csv", text), 0, len(text), float, ",", False, device_index=0)
    got = down(blk)
    assert got.shape == want.shape
    neq = bits(got) != bits(want)
    neq &= ~(np.isnan(got) & np.isnan(want))
    assert not neq.any(), [(strs[i], got.reshape(-1)[i], want.reshape(-1)[i]) for i in np.flatnonzero(neq.reshape(-1))[:10]]


def test_malformed_text_raises_like_the_reference(app, tmp_path):
    from nums_b200 import filesystem as fs
    for bad in (b"1,2\n3\n", b"1,2\n3,4,5\n", b"1,abc\n", b"1,\n", b"1,2\n\n3,4\n", b"1e,2\n", b"1,1__0\n",
                b"1,1e5_\n", b"--1,2\n"):
        p = write(tmp_path, "bad.csv", bad)
        with pytest.raises(ValueError):
            fs.read_csv_block(p, 0, len(bad), float, ",", False, device_index=0)
        with pytest.raises(ValueError):
            csvio.read_csv_block(bad, 0, len(bad), float, ",", False)
    blk, shape = fs.read_csv_block(write(tmp_path, "empty.csv", b""), 0, 0, float, ",", False, device_index=0)
    assert shape == (0,)
    # the one accepted input the device path declines: hexadecimal floats (float.fromhex in the reference)
    hexed = b"1,0x1p3\n"
    assert csvio.read_csv_block(hexed, 0, len(hexed))[0, 1] == 8.0
    with pytest.raises(ValueError, match="hexadecimal"):
        fs.read_csv_block(write(tmp_path, "hex.csv", hexed), 0, len(hexed), float, ",", False, device_index=0)


def test_block_store_round_trip_and_fit_from_csv(app, tmp_path):
    """write_fs / read_fs (one file per block + grid meta) and the README's read_csv -> fit flow on a small file."""
    from nums_b200.glms import LogisticRegression
    from oracle import glm as oglm
    rng = np.random.default_rng(5)
    X = rng.standard_normal((3000, 6))
    y = (X @ np.linspace(-1, 1, 6) + 0.3 * rng.standard_normal(3000) > 0).astype(np.float64)
    text = ("\n".join(",".join(repr(float(v)) for v in list(r) + [t]) for r, t in zip(X, y)) + "\n").encode()
    path = write(tmp_path, "data.csv", text)
    data = app.read_csv(path, num_workers=3)
    host = data.get()
    assert np.array_equal(host[:, :6], X) and np.array_equal(host[:, 6], y)
    store = os.path.join(str(tmp_path), "stored_array")
    app.write_fs(data, store)
    back = app.read_fs(store)
    assert back.block_shape == data.block_shape and np.array_equal(back.get(), host)
    assert app.delete_fs(store) and not os.path.exists(store)
    Xb = app.array(host[:, :6], block_shape=(1000, 6))
    yb = app.array(host[:, 6].astype(np.int64), block_shape=(1000,))
    m = LogisticRegression(app=app, solver="newton", penalty="l2", C=1.0, max_iter=20)
    m.fit(Xb, yb)
    want, _ = oglm.fit("logistic", X, y.astype(np.int64), 1000, solver="newton", penalty="l2", lam=1.0, max_iter=20)
    assert np.max(np.abs(m._beta.get() - want[:-1])) < 1e-8
