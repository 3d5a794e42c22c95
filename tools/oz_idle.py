"""Does the row pass slow down inside a back-to-back loop because of power/clock capping?  Times k_oz_rows launches one by one
with and without idle gaps between them, and samples SM clock and board power while they run."""
import os, sys, time, threading
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import pynvml
from nums_b200 import device as dv, fused, _lib
from nums_b200.app import make_app, set_instance

set_instance(make_app(num_gpus=1))
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
n, d = 12_500_000, 256
g = torch.Generator(device="cuda:0"); g.manual_seed(1)
X = dv.DeviceBlock(torch.randn((n, d), dtype=torch.float64, device="cuda:0", generator=g))
y = dv.DeviceBlock((torch.rand((n,), dtype=torch.float64, device="cuda:0", generator=g) < 0.5).to(torch.int64))
beta = dv.from_numpy(np.random.default_rng(0).standard_normal(d + 1) * 0.01, 0)
lib = _lib.load()
stats_d = dv.empty((lib.nb_glm_stats_len(d),), np.float64, 0)
prep = fused.glm_prepare(X)
samples = []
stop = False


def sampler():
    while not stop:
        samples.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_MEM),
                        pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0))
        time.sleep(0.002)


for planes in ("5", "4"):
    os.environ["NUMS_B200_INT8_PLANES"] = planes
    flags = 3 | fused.int8_flags(prep, n)
    sc = dv.scratch(lib.nb_glm_scratch_bytes(0, n, d, flags), 0)
    for name, extra in (("rows", 8), ("products", 16), ("both", 0)):
        call = lambda: _lib.check(lib.nb_glm_stats_prepared(0, 1, X.ptr, n, d, d, y.ptr, y.code, beta.ptr, flags | extra,
                                                            prep.buf.data_ptr(), stats_d.ptr, sc.data_ptr(), sc.numel(), dv.stream_ptr(0)))
        for gap in (0.0, 0.05):
            call(); torch.cuda.synchronize()
            samples.clear(); stop = False
            th = threading.Thread(target=sampler); th.start()
            evs = []
            for _ in range(12):
                if gap:
                    torch.cuda.synchronize(); time.sleep(gap)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); call(); e1.record(); evs.append((e0, e1))
            torch.cuda.synchronize()
            stop = True; th.join()
            ms = sorted(a.elapsed_time(b) for a, b in evs[2:])
            sm = sorted(s[0] for s in samples); pw = sorted(s[2] for s in samples)
            print("%s planes %-8s gap %.2fs: median %.3f ms (min %.3f max %.3f)  SM clock median %d min %d MHz, mem %d MHz, power median %.0f max %.0f W"
                  % (planes, name, gap, ms[len(ms) // 2], ms[0], ms[-1], sm[len(sm) // 2], sm[0], samples[0][1], pw[len(pw) // 2], pw[-1]), flush=True)
    del sc
