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
    """DDP semantics: gradients are averaged over ranks.  No-op for a single process.  (Stand-alone form: a separate
    division after the sum.  The trainers below fold the 1/world into the backward's grad_scale and call GradReducer.)"""
    if not (dist.is_available() and dist.is_initialized()):
        return flat_grad
    world = dist.get_world_size(group) if world is None else world
    if world > 1:
        dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=group)
        flat_grad.div_(world)
    return flat_grad


def dist_world(group=None):
    return dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1


class GradReducer:
    """The path's one collective: SUM all-reduce of the flat gradient whose 1/world factor is already inside it
    (the backward ran with grad_scale = 1/world, so no division kernel follows) -- DDP's mean, trainer.py:16-21.

    ``overlap`` (GPU / NCCL): the fine net's half of the buffer is final before the coarse net's backward starts
    (``nerf_stream_wait_fine_grads``), so it is reduced on a side stream while ``sample_pdf_bwd`` and the coarse dgrad / wgrad
    kernels run; the coarse half follows on the compute stream, which then joins the side stream before clip + Adam.
    With ``overlap=False`` (and always on CPU tensors / gloo) the whole buffer is reduced in one call after the backward."""

    def __init__(self, n_fine_offset, overlap=True, group=None):
        self.split, self.overlap, self.group = int(n_fine_offset), bool(overlap), group
        self._comm = None

    def __call__(self, flat_grad, engine=None):
        world = dist_world(self.group)
        if world <= 1:
            return flat_grad
        if not (self.overlap and flat_grad.is_cuda and engine is not None):
            dist.all_reduce(flat_grad, op=dist.ReduceOp.SUM, group=self.group)
            return flat_grad
        if self._comm is None:
            self._comm = torch.cuda.Stream(device=flat_grad.device)
        cur = torch.cuda.current_stream(flat_grad.device)
        engine.stream_wait_fine_grads(self._comm)                 # side stream: starts when the fine net's gradients are final
        with torch.cuda.stream(self._comm):
            dist.all_reduce(flat_grad[self.split:], op=dist.ReduceOp.SUM, group=self.group)
        dist.all_reduce(flat_grad[:self.split], op=dist.ReduceOp.SUM, group=self.group)     # after the whole backward
        cur.wait_stream(self._comm)
        return flat_grad


def exponential_lr(base_lr, epoch, gamma=0.1, decay_epochs=500):
    """ExponentialLR of src/utils/optimizer/lr_scheduler.py:68-79 (stepped per epoch, train.py:66)."""
    return base_lr * gamma ** (epoch / decay_epochs)


class FusedTrainer:
    def __init__(self, net, cfg=None, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, clip=40.0, overlap=True):
        self.net = net
        self.renderer = Renderer(net, cfg)
        self.lr, self.betas, self.eps, self.clip = lr, betas, eps, clip
        flat = net.flat_params
        self.m, self.v = torch.zeros_like(flat), torch.zeros_like(flat)
        self.step_count = 0
        self.reduce = GradReducer(flat.numel() // 2, overlap=overlap)
        self._loss = torch.zeros((), device=flat.device)      # persistent: the step's CUDA graph writes the loss here

    def forward_backward(self, rays_o, rays_d, rgb_gt, jitter=None, u=None, grad_scale=None):
        eng = self.renderer._engine()
        scale = 1.0 / dist_world() if grad_scale is None else grad_scale
        return eng.forward_backward(rays_o, rays_d, rgb_gt, self.net.flat_grad, jitter=jitter, u=u, grad_scale=scale, loss=self._loss)

    def apply(self):
        eng = self.renderer._engine()
        self.step_count += 1
        eng.clip_adam(self.net.flat_params, self.net.flat_grad, self.m, self.v, self.step_count, self.lr,
                      self.betas[0], self.betas[1], self.eps, self.clip)
        self.net.mark_packed()

    def step(self, rays_o, rays_d, rgb_gt, jitter=None, u=None):
        """-> this rank's loss (device scalar, overwritten by the next step)."""
        loss = self.forward_backward(rays_o, rays_d, rgb_gt, jitter, u)
        self.reduce(self.net.flat_grad, self.renderer._engine())
        self.apply()
        return loss


# ------------------------------------------------------------------------------------------------------------
# the step driver around the fused step -- train.py:31-87 + trainer.py:10-128 of the reference,
# working (the reference's Trainer unpacks a 4-tuple its NetworkWrapper never returns, trainer.py:53 vs
# trainers/nerf.py:36-42) and with no DataLoader: the images live in the engine's ray bank (dataset.py).
import datetime
import math
import time
from collections import deque


class SmoothedValue:
    """recorder.py:10-39: window median / average and global average of a scalar series."""

    def __init__(self, window_size=20):
        self.deque, self.total, self.count = deque(maxlen=window_size), 0.0, 0

    def update(self, value):
        self.deque.append(value)
        self.count += 1
        self.total += value

    @property
    def median(self):
        return torch.tensor(list(self.deque)).median().item()

    @property
    def avg(self):
        return torch.tensor(list(self.deque)).mean().item()

    @property
    def global_avg(self):
        return self.total / max(self.count, 1)


class Recorder:
    """The part of src/train/recorder.py the loop and the checkpoints use: step / epoch counters, smoothed loss stats,
    batch/data time, ``state_dict() == {"step": ...}`` (:109-119).  TensorBoard output is optional (tensorboardX)."""

    def __init__(self, record_dir=None):
        self.step, self.epoch = 0, 0
        self.loss_stats, self.batch_time, self.data_time = {}, SmoothedValue(), SmoothedValue()
        self.writer = None
        if record_dir is not None:
            try:
                from tensorboardX import SummaryWriter
                self.writer = SummaryWriter(log_dir=record_dir)
            except ImportError:
                self.writer = None

    def update_loss_stats(self, loss_dict):
        for k, v in loss_dict.items():
            self.loss_stats.setdefault(k, SmoothedValue()).update(float(v))

    def record(self, prefix, step=-1, loss_stats=None, image_stats=None):
        if self.writer is None:
            return
        step = step if step >= 0 else self.step
        stats = loss_stats if loss_stats else {k: v.median for k, v in self.loss_stats.items()}
        for k, v in stats.items():
            self.writer.add_scalar("%s/%s" % (prefix, k), float(v), step)

    def state_dict(self):
        return {"step": self.step}

    def load_state_dict(self, d):
        self.step = d["step"]

    def __str__(self):
        stats = "  ".join("%s: %.4f" % (k, v.avg) for k, v in self.loss_stats.items())
        return "  ".join(["epoch: %d" % self.epoch, "step: %d" % self.step, stats,
                          "data: %.4f" % self.data_time.avg, "batch: %.4f" % self.batch_time.avg])


class RayLoader:
    """What train.py's ``make_data_loader`` returns, reduced to what is left when the dataset lives in HBM: an iterable
    over (view index, draw id) per iteration following the reference's sampler stream (view_schedule.py), with
    ``.dataset`` and ``.batch_sampler.sampler.set_epoch`` where train.py:57-61 expects them."""

    def __init__(self, dataset, schedule, rank=0, world=1):
        self.dataset, self.schedule, self.rank, self.world = dataset, schedule, rank, world
        self.batch_sampler = self            # train_loader.batch_sampler.sampler.set_epoch(epoch)
        self.sampler = schedule
        self.drawn = 0

    def __len__(self):
        return len(self.schedule)

    def __iter__(self):
        for views in self.schedule:
            for v in views:
                draw = self.drawn * self.world + self.rank       # distinct pixel draws on every rank and iteration
                self.drawn += 1
                yield v, draw


def make_data_loader(cfg, dataset, is_train=True, is_distributed=False, max_iter=-1, rank=0, world=1, generator=None):
    """src/datasets/make_dataset.py:73-100 ("default" batch sampler, batch_size 1 image per step as lego.yaml:64,76)."""
    from .view_schedule import ViewSchedule
    tr = getattr(cfg, "train", None)
    if is_train:
        batch_size, shuffle = getattr(tr, "batch_size", 1), getattr(tr, "shuffle", True)
    else:
        batch_size, shuffle = getattr(getattr(cfg, "test", None), "batch_size", 1), bool(is_distributed)
    if batch_size != 1:
        raise NotImplementedError("one view per step (cfg.train.batch_size = 1, lego.yaml:64); got %d" % batch_size)
    sched = ViewSchedule(len(dataset), batch_size, shuffle, is_distributed, max_iter, rank, world, generator=generator)
    return RayLoader(dataset, sched, rank, world)


class NetworkWrapper:
    """trainers/nerf.py:9-42 with the return contract the reference's Trainer expects (trainer.py:53):
    ``output, loss, loss_stats, image_stats``; psnr as torchmetrics' PeakSignalNoiseRatio(data_range=1.0)."""

    def __init__(self, net, train_loader=None, cfg=None):
        self.net = net
        self.renderer = Renderer(net, cfg)

    def __call__(self, batch):
        rgb, depth = self.renderer.render(batch)
        gt = batch["colors"].reshape(-1, 3)
        loss = torch.nn.functional.mse_loss(rgb, gt)
        psnr = -10.0 * torch.log10(loss.detach())
        return (rgb, depth), loss, {"loss": loss.detach(), "psnr": psnr}, {}


class Trainer:
    """Mirror of src/train/trainers/trainer.py: ``train(epoch, data_loader, optimizer, recorder)`` and
    ``val(epoch, data_loader, evaluator, recorder)``.  A training iteration is two engine calls --
    ``nerf_bank_train_step`` (pixel draw + ray generation + colour gather + forward + loss + backward) and
    ``optimizer.step()`` (clip 40 + Adam + bf16 weight re-pack) -- with one NCCL all-reduce between them when
    ``world > 1``.  The loss stays on the device; it is read back only every ``log_interval`` iterations."""

    def __init__(self, net, cfg=None, log_interval=10, clip=40.0, rank=0, verbose=True):
        self.net, self.cfg = net, cfg
        self.wrapper = NetworkWrapper(net, cfg=cfg)
        self.renderer = self.wrapper.renderer
        self.log_interval = getattr(cfg, "log_interval", log_interval) if cfg is not None else log_interval
        self.clip, self.local_rank, self.global_step, self.verbose = clip, rank, 0, verbose
        self._loss, self.reduce = None, None

    def _clip_unfused(self, optimizer):
        """trainer.py:59 clip_grad_value_(params, 40).  FusedAdam clips inside its kernel (``optimizer.clip``); for any other
        optimizer (e.g. the torch.optim.Adam the reference's make_optimizer builds) the flat gradient is clamped here, so
        the clip is never silently dropped.  torch.clamp keeps NaN, like clip_grad_value_."""
        from .optimizer import FusedAdam
        if not isinstance(optimizer, FusedAdam) and self.clip:
            self.net.flat_grad.clamp_(-self.clip, self.clip)

    def train(self, epoch, data_loader, optimizer, recorder, jitter=None, u=None):
        ds = data_loader.dataset
        ds.attach(net=self.net)
        optimizer.clip = self.clip                      # clip_grad_value_(…, 40) fused into the Adam kernel (trainer.py:59)
        max_iter, end = len(data_loader), time.time()
        losses = torch.zeros(max_iter, device=self.net.flat_params.device)
        world = dist_world()
        if self._loss is None:
            self._loss = torch.zeros((), device=self.net.flat_params.device)
            self.reduce = GradReducer(self.net.flat_params.numel() // 2)
        for iteration, (view, draw) in enumerate(data_loader):
            data_time = time.time() - end
            eng = self.renderer._engine()
            if eng.cfg.seed != ds.seed:
                eng.configure(seed=ds.seed)             # the pixel draw is keyed by (dataset seed, draw)
            eng.bank_train_step(ds.bank, view, ds.n_rays, self.net.flat_grad, mode="random", draw=draw, jitter=jitter, u=u,
                                grad_scale=1.0 / world, loss=self._loss)
            losses[iteration].copy_(self._loss)         # the step (a replayed CUDA graph) always writes the same scalar
            self.reduce(self.net.flat_grad, eng)
            self._clip_unfused(optimizer)
            optimizer.step()
            iteration += 1
            if self.local_rank > 0:
                continue
            recorder.step += 1
            self.global_step += 1
            recorder.batch_time.update(time.time() - end)   # every iteration (launch-side time; the GPU runs at most one loss read behind)
            recorder.data_time.update(data_time)
            if iteration % self.log_interval == 0 or iteration == max_iter - 1:
                loss = float(losses[iteration - 1])     # the only device->host read of the loop
                recorder.update_loss_stats({"loss": loss, "psnr": -10.0 * math.log10(max(loss, 1e-30))})
                if self.verbose:
                    eta = str(datetime.timedelta(seconds=int(recorder.batch_time.global_avg * (max_iter - iteration))))
                    print("  ".join(["eta: %s" % eta, str(recorder), "lr: %.6f" % optimizer.param_groups[0]["lr"],
                                     "max_mem: %.0f" % (torch.cuda.max_memory_allocated() / 2 ** 20)]))
                recorder.record("train")
            end = time.time()
        return losses

    @torch.no_grad()
    def val(self, epoch, data_loader, evaluator=None, recorder=None):
        ds = data_loader.dataset
        ds.attach(net=self.net)
        stats, n = {}, 0
        for view, _ in data_loader:
            batch = ds.batch(view)
            output, loss, loss_stats, _ = self.wrapper(batch)
            if evaluator is not None:
                evaluator.evaluate(output, batch)
            for k, v in loss_stats.items():
                stats[k] = stats.get(k, 0.0) + float(v)
            n += 1
        stats = {k: v / max(n, 1) for k, v in stats.items()}
        if self.verbose:
            print(["%s: %.4f" % kv for kv in stats.items()])
        if evaluator is not None:
            stats.update(evaluator.summarize())
        if recorder is not None:
            recorder.record("val", epoch, stats, {})
        return stats


def fit(cfg, net, train_set, val_set=None, evaluator=None, rank=0, world=1, resume=True, epochs=None, verbose=True):
    """train.py:31-87: optimizer, scheduler, recorder, resume from cfg.trained_model_dir, then per epoch: ep_iter fused
    steps, scheduler.step(), save every save_ep / save_latest_ep, evaluate every eval_ep (rank 0)."""
    from . import checkpoint as ck
    from .optimizer import make_lr_scheduler, make_optimizer, set_lr_scheduler
    distributed = world > 1
    loader = make_data_loader(cfg, train_set, True, distributed, getattr(cfg, "ep_iter", -1), rank, world)
    val_loader = make_data_loader(cfg, val_set, False) if val_set is not None else None
    trainer = Trainer(net, cfg, rank=rank, verbose=verbose and rank == 0)
    optimizer = make_optimizer(cfg, net, clip=40.0)
    scheduler = make_lr_scheduler(cfg, optimizer)
    recorder = Recorder(getattr(cfg, "record_dir", None) if rank == 0 else None)
    begin = ck.load_model(net, optimizer, scheduler, recorder, cfg.trained_model_dir, resume=resume)
    set_lr_scheduler(cfg, scheduler)
    loader.drawn = begin * len(loader)                  # a resumed run continues the pixel-draw sequence
    last = cfg.train.epoch if epochs is None else min(cfg.train.epoch, begin + epochs)
    history = []
    for epoch in range(begin, last):
        recorder.epoch = epoch
        if distributed:
            loader.batch_sampler.sampler.set_epoch(epoch)
        losses = trainer.train(epoch, loader, optimizer, recorder)
        scheduler.step()
        history.append(losses)
        if rank == 0:
            if (epoch + 1) % cfg.save_ep == 0:
                ck.save_model(net, optimizer, scheduler, recorder, cfg.trained_model_dir, epoch)
            if (epoch + 1) % cfg.save_latest_ep == 0:
                ck.save_model(net, optimizer, scheduler, recorder, cfg.trained_model_dir, epoch, last=True)
            if (epoch + 1) % cfg.eval_ep == 0 and val_loader is not None:
                trainer.val(epoch, val_loader, evaluator, recorder)
    return history, optimizer, scheduler, recorder
