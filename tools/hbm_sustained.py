"""HBM copy bandwidth as a burst (best single copy after an idle gap) and sustained (back to back for 3 s), with the SM clock and
board power sampled meanwhile: is a kernel inside a long step loop held to less than the burst figure of MEASURED_PEAKS.json?"""
import time, threading
import torch
import pynvml

pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
n = 1 << 30
a = torch.empty(n, dtype=torch.bfloat16, device="cuda:0").normal_()
b = torch.empty_like(a)
bytes_moved = 2 * n * 2


def one():
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); b.copy_(a); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)


for _ in range(3):
    one()
burst = []
for _ in range(10):
    time.sleep(0.1)
    burst.append(one())
print("burst: best %.0f GB/s, median %.0f GB/s" % (bytes_moved / min(burst) / 1e6, bytes_moved / sorted(burst)[5] / 1e6), flush=True)
samples, stop = [], False


def sampler():
    while not stop:
        samples.append((pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM), pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0))
        time.sleep(0.01)


th = threading.Thread(target=sampler); th.start()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
reps = 0
t0 = time.time()
e0.record()
while time.time() - t0 < 3.0:
    for _ in range(50):
        b.copy_(a)
    reps += 50
    torch.cuda.synchronize()
e1.record(); torch.cuda.synchronize()
stop = True; th.join()
ms = e0.elapsed_time(e1) / reps
sm = sorted(s[0] for s in samples); pw = sorted(s[1] for s in samples)
print("sustained (%d copies, 3 s): %.0f GB/s; SM clock median %d MHz (min %d), power median %.0f W (max %.0f), enforced limit %.0f W"
      % (reps, bytes_moved / ms / 1e6, sm[len(sm) // 2], sm[0], pw[len(pw) // 2], pw[-1],
         pynvml.nvmlDeviceGetEnforcedPowerLimit(h) / 1000.0), flush=True)
