"""Times the device CSV parse (text already in HBM) and the host float() loop of the reference on the same text."""
import os, sys, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import _lib, device as dv
from nums_b200.app import make_app, set_instance
set_instance(make_app(num_gpus=1))
lib = _lib.load()
rng = np.random.default_rng(0)
cols = 29                                             # HIGGS: label + 28 features
rows_chunk = 20000
vals = rng.standard_normal((rows_chunk, cols))
chunk = ("\n".join(",".join("%.18e" % v for v in r) for r in vals) + "\n").encode()
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
data = chunk * reps
print("text: %.1f MB, %d rows x %d" % (len(data) / 1e6, rows_chunk * reps, cols), flush=True)
text = dv.upload_bytes(data, 0)
st = dv.stream_ptr(0)
rows_c = C.c_int64(0)
sc = dv.scratch(lib.nb_csv_scratch_bytes(len(data), rows_chunk * reps), 0)
out = dv.empty((rows_chunk * reps, cols), np.float64, 0)
err = (C.c_int64 * 3)()
def run():
    _lib.check(lib.nb_csv_count_rows(0, text.data_ptr(), len(data), sc.data_ptr(), sc.numel(), C.byref(rows_c), st))
    _lib.check(lib.nb_csv_parse(0, text.data_ptr(), len(data), rows_c.value, cols, ord(","), out.code, out.ptr,
                                sc.data_ptr(), sc.numel(), err, st))
run(); run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(3): run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 3
assert err[0] == 0 and rows_c.value == rows_chunk * reps
print("device parse: %.2f ms  %.1f GB/s of text, %.2e fields/s" % (ms, len(data) / ms / 1e6, rows_c.value * cols / ms * 1e3))
t0 = time.perf_counter()
lines = chunk.decode().split("\n")[:-1]
arr = np.array([[float(f) for f in line.split(",")] for line in lines])
dt = time.perf_counter() - t0
print("host float() loop (reference's converter), one chunk: %.1f MB/s" % (len(chunk) / dt / 1e6))
got = dv.to_numpy(dv.DeviceBlock(out.t[:rows_chunk]))
assert np.array_equal(got, arr)
