"""Timing experiments on the CTA-pair forward kernel (csrc/mlp_tc2.cu, NERF_EXP bits; build.py --exp N).
Usage on the GPU box:  NERF_B200_LIB=nerf-replication_b200/libnerf_b200_expN.so python profiles/exp_fwd.py
Times the render instantiation (coarse net, 65536 rays x 64 samples = 4.19 M rows) with CUDA events; results of the
experimental flavours are wrong by construction -- only the time is meaningful."""
import json
import os
import sys
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nerf_replication_b200 import config as C            # noqa: E402
from nerf_replication_b200.network import Network       # noqa: E402

R, S = 65536, 64
net = Network(cfg=C.default_cfg(), precision="bf16").cuda()
eng = net.engine()
pts = torch.rand(R, S, 3, device="cuda") * 2 - 1
dirs = torch.nn.functional.normalize(torch.randn(R, 3, device="cuda"), dim=-1)
for _ in range(3):
    eng.mlp_forward(0, pts, dirs)
torch.cuda.synchronize()
ms = []
for _ in range(10):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    eng.mlp_forward(0, pts, dirs)
    e1.record()
    e1.synchronize()
    ms.append(e0.elapsed_time(e1))
ms.sort()
med = ms[len(ms) // 2]
print(json.dumps({"lib": os.environ.get("NERF_B200_LIB", "product"), "ms_median": med, "ms_best": ms[0],
                  "tflops_median": R * S * 1186816 / med / 1e9}))
