"""Two int8 statistics passes per plane count on the benchmark shard, for an ncu launch list:
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum -k regex:k_oz --csv python tools/oz_phase.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from nums_b200 import device as dv, fused
from nums_b200.app import make_app, set_instance

set_instance(make_app(num_gpus=1))
n, d = int(float(sys.argv[1])) if len(sys.argv) > 1 else 12_500_000, 256
g = torch.Generator(device="cuda:0"); g.manual_seed(1)
X = dv.DeviceBlock(torch.randn((n, d), dtype=torch.float64, device="cuda:0", generator=g))
y = dv.DeviceBlock((torch.rand((n,), dtype=torch.float64, device="cuda:0", generator=g) < 0.5).to(torch.int64))
beta = dv.from_numpy(np.random.default_rng(0).standard_normal(d + 1) * 0.01, 0)
ref = dv.to_numpy(fused.glm_stats(X, y, beta, "logistic", True, need_obj=False, hessian="fp64"))
for planes in ("5", "4"):
    os.environ["NUMS_B200_INT8_PLANES"] = planes
    for _ in range(2):
        s = dv.to_numpy(fused.glm_stats(X, y, beta, "logistic", True, need_obj=False, hessian="int8"))
    print("planes %s: max|int8 - fp64| / max|fp64| = %.3e" % (planes, np.max(np.abs(s - ref)) / np.max(np.abs(ref))), flush=True)
