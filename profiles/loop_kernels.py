"""Times the 8(f) kernels (ray bank gather, evaluator sums, SSIM) with CUDA events on the launching stream and prints
achieved HBM GB/s against their algorithmic bytes.  python profiles/loop_kernels.py  (B200)"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_replication_b200.engine import Engine  # noqa: E402


def timed(fn, iters=50, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3      # us per call


def main():
    eng = Engine()
    H = W = 800
    V = 8
    imgs = np.random.RandomState(0).randint(0, 256, size=(V, H, W, 3), dtype=np.uint8)
    pose = [[1, 0, 0, 0.5], [0, 1, 0, -1.0], [0, 0, 1, 4.0]]
    out = {}
    for name, u8 in (("u8", True), ("f32", False)):
        b = eng.bank_create(V, H, W, 1111.111, uint8=u8)
        for v in range(V):
            eng.bank_set_view(b, v, imgs[v] if u8 else imgs[v].astype(np.float32), pose)
        px = 3 if u8 else 12
        for n, mode in ((1024, "random"), (H * W, "range"), (H * W, "random")):
            us = timed(lambda: eng.bank_sample(b, 3, n, mode=mode, draw=7))
            # algorithmic bytes: colour gather (3 pixels values) in, rays_o/rays_d/colors out (36 B/ray); allocation of the outputs is inside
            bytes_ = n * (px + 36)
            out["bank_sample_%s_%s_%d" % (name, mode, n)] = {"us": us, "GBps": bytes_ / us / 1e3}
    pred = torch.rand(H * W, 3, device="cuda")
    gt = torch.rand(H * W, 3, device="cuda")
    us = timed(lambda: eng.eval_image(pred, gt))
    out["eval_image_800x800"] = {"us": us, "GBps": 2 * pred.numel() * 4 / us / 1e3}
    us = timed(lambda: eng.eval_ssim(pred, gt, H, W, 7))
    out["eval_ssim_800x800_win7"] = {"us": us, "GBps": 2 * pred.numel() * 4 / us / 1e3, "note": "3 launches: 2 quantise + stencil"}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
