"""Fused training step: the composition the reference files define --
render -> MSELoss (trainers/nerf.py:27-33) -> backward -> [DDP mean all-reduce, trainer.py:16-21] ->
clip_grad_value_(40) (trainer.py:59) -> Adam(lr 5e-4, eps 1e-8) (train/optimizer.py:8-28) -- as three engine
calls on one flat gradient buffer.  Data parallel: every rank renders its own rays; the only collective is
one all-reduce of the 1 191 688-float gradient buffer (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""
import torch
import torch.distributed as dist

from .renderer import Renderer


def shard_range(n, rank, world):
    """Contiguous, balanced [begin, end) share of n units for `rank` (image tiles / ray batches)."""
    base, rem = divmod(n, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def allreduce_mean_(flat_grad, world=None, group=None):
    """DDP semantics: gradients are averaged over ranks.  No-op for a single process."""
    if not (dist.is_available() and dist.is_initialized()):
        return flat_grad
    world = dist.get_world_size(group) if world is None else world
    if world > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
        flat_grad.div_(world)
    return flat_grad


def exponential_lr(base_lr, epoch, gamma=0.1, decay_epochs=500):
    """ExponentialLR of src/utils/optimizer/lr_scheduler.py:68-79 (stepped per epoch, train.py:66)."""
    return base_lr * gamma ** (epoch / decay_epochs)


class FusedTrainer:
    def __init__(self, net, cfg=None, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, clip=40.0):
        self.net = net
        self.renderer = Renderer(net, cfg)
        self.lr, self.betas, self.eps, self.clip = lr, betas, eps, clip
        flat = net.flat_params
        self.m, self.v = torch.zeros_like(flat), torch.zeros_like(flat)
        self.step_count = 0

    def forward_backward(self, rays_o, rays_d, rgb_gt, jitter=None, u=None):
        eng = self.renderer._engine()
        return eng.forward_backward(rays_o, rays_d, rgb_gt, self.net.flat_grad, jitter=jitter, u=u)

    def apply(self):
        eng = self.renderer._engine()
        self.step_count += 1
        eng.clip_adam(self.net.flat_params, self.net.flat_grad, self.m, self.v, self.step_count, self.lr,
                      self.betas[0], self.betas[1], self.eps, self.clip)
        self.net.mark_packed()

    def step(self, rays_o, rays_d, rgb_gt, jitter=None, u=None):
        loss = self.forward_backward(rays_o, rays_d, rgb_gt, jitter, u)
        allreduce_mean_(self.net.flat_grad)
        self.apply()
        return loss
