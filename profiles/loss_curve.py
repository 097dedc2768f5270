#!/usr/bin/env python
"""1k-step loss-curve parity (north_star: "training loss curves within 1% over 1k steps").

    python profiles/loss_curve.py [--steps 1000] [--rays 64] > profiles/loss_curve_r1.txt      (on a B200)

Trains three copies from the same init on the same synthetic batches (SURVEY 8d "Training inputs", 100x100 views,
perturb = 0, u = linspace, lr 5e-4 with the reference's per-epoch ExponentialLR applied at step 500):
  oracle  : oracle/nerf_oracle.py Trainer (CPU fp32 torch autograd + torch.optim.Adam) -- the reference's algorithm
  fp32    : engine, fp32 validation kernels
  bf16    : engine, tensor-core path (saving forward + dgrad chain + wgrad), fp32 master weights
and prints the per-step losses and the relative deviations."""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from oracle import nerf_oracle as O  # noqa: E402
from nerf_replication_b200 import config as C  # noqa: E402
from nerf_replication_b200.network import Network  # noqa: E402
from nerf_replication_b200.trainer import FusedTrainer, exponential_lr  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=1000)
ap.add_argument("--rays", type=int, default=64)
ap.add_argument("--weights", default="seed0", choices=["seed0", "fog"])
args = ap.parse_args()

torch.set_num_threads(os.cpu_count() or 8)
cfg = C.default_cfg()
sd = O.init_state_dict(0) if args.weights == "seed0" else O.fog_state_dict(0)
engines = {}
for prec in ("fp32", "bf16"):
    net = Network(cfg=cfg, precision=prec)
    net.load_state_dict(sd)
    engines[prec] = FusedTrainer(net.cuda(), cfg=cfg)
ref = O.Trainer(sd)
worst = {"fp32": 0.0, "bf16": 0.0}
t0 = time.time()
print("# step  loss_oracle  loss_fp32  loss_bf16  rel_fp32  rel_bf16")
for s in range(args.steps):
    lr = exponential_lr(5e-4, 1 if s >= 500 else 0)            # train.py:66: scheduler.step() once per 500-iteration epoch
    ref.set_lr(lr)
    o, d, gt, _, _ = O.synthetic_train_batch(s, 0, args.rays, 100, 100)
    l_ref = ref.step(o, d, gt)
    row = [l_ref]
    for prec, tr in engines.items():
        tr.lr = lr
        l = float(tr.step(o.cuda(), d.cuda(), gt.cuda()))
        worst[prec] = max(worst[prec], abs(l - l_ref) / l_ref)
        row.append(l)
    if s % 10 == 0 or s == args.steps - 1:
        print("%5d  %.8f  %.8f  %.8f  %.3e  %.3e" % (s, row[0], row[1], row[2], abs(row[1] - row[0]) / row[0], abs(row[2] - row[0]) / row[0]))
print("# worst relative deviation over %d steps (%d rays/step, %s weights): fp32 %.3e, bf16 %.3e  (gate 1e-2)  [%.0f s]"
      % (args.steps, args.rays, args.weights, worst["fp32"], worst["bf16"], time.time() - t0))
