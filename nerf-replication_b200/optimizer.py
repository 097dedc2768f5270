"""Optimizer + lr schedule of the reference (src/train/optimizer.py:8-28, src/train/scheduler.py:9-33,
src/utils/optimizer/lr_scheduler.py:68-79) on the engine's flat buffers.

``FusedAdam`` is a ``torch.optim.Optimizer`` so that everything the reference hangs on its optimizer keeps working --
``optimizer.param_groups[0]["lr"]`` (trainer.py:82), ``torch.optim.lr_scheduler`` subclasses, ``optimizer.state_dict()``
in ``save_model`` / ``load_model`` (net_utils.py:288-343) -- but ``step()`` is one launch of the engine's clip+Adam kernel
over the 1 191 688-float flat parameter / gradient / moment buffers (a-L) instead of 48 small updates.  Its state dict has
torch.optim.Adam's layout (one group per parameter as make_optimizer builds them; ``step / exp_avg / exp_avg_sq`` per
parameter), so checkpoints move between the reference and this engine in both directions.
"""
import torch


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, net, lr=5e-4, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, clip=None):
        if weight_decay != 0.0:
            raise NotImplementedError("FusedAdam implements weight_decay = 0 (configs/nerf/lego.yaml:66); no fallback")
        self.net, self.clip = net, clip
        groups = [{"params": [p], "lr": lr, "weight_decay": weight_decay, "eps": eps}
                  for _, p in net.named_parameters() if p.requires_grad]          # optimizer.py:16-21
        defaults = dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False, foreach=None,
                        capturable=False, differentiable=False, fused=None)
        super().__init__(groups, defaults)
        if len(groups) != len(list(net.parameters())):
            raise NotImplementedError("FusedAdam updates the whole flat buffer: every parameter must require grad")
        flat = net.flat_params
        self.m, self.v = torch.zeros_like(flat), torch.zeros_like(flat)
        self.step_count = 0
        self._bind_state()

    def _bind_state(self):
        """state[p] entries are views into the flat moment buffers (what Optimizer.state_dict() serialises)."""
        o = 0
        for p in self.net.parameters():
            n = p.numel()
            self.state[p] = {"step": torch.tensor(float(self.step_count)), "exp_avg": self.m[o:o + n].view(p.shape),
                             "exp_avg_sq": self.v[o:o + n].view(p.shape)}
            o += n

    def _hyper(self):
        g0 = self.param_groups[0]
        for g in self.param_groups[1:]:
            if (g["lr"], tuple(g["betas"]), g["eps"]) != (g0["lr"], tuple(g0["betas"]), g0["eps"]):
                raise NotImplementedError("FusedAdam needs the same lr / betas / eps in every group (one flat update)")
        return g0["lr"], g0["betas"], g0["eps"]

    @torch.no_grad()
    def step(self, closure=None):
        if closure is not None:
            raise NotImplementedError("FusedAdam.step takes no closure")
        lr, betas, eps = self._hyper()
        self.step_count += 1
        eng = self.net.engine()
        eng.clip_adam(self.net.flat_params, self.net.flat_grad, self.m, self.v, self.step_count, lr, betas[0], betas[1], eps,
                      0.0 if self.clip is None else self.clip)
        self.net.mark_packed()

    def zero_grad(self, set_to_none=False):
        """The fused forward_backward overwrites the flat gradient buffer; keep the views alive instead of dropping them."""
        self.net.flat_grad.zero_()

    def state_dict(self):
        for st in self.state.values():
            st["step"] = torch.tensor(float(self.step_count))
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)          # casts / copies the loaded moments next to their parameters
        steps, o = set(), 0
        for p in self.net.parameters():
            st, n = self.state.get(p, None), p.numel()
            if st:
                self.m[o:o + n].copy_(st["exp_avg"].reshape(-1))
                self.v[o:o + n].copy_(st["exp_avg_sq"].reshape(-1))
                steps.add(int(float(st["step"])))
            else:
                self.m[o:o + n].zero_()
                self.v[o:o + n].zero_()
                steps.add(0)
            o += n
        if len(steps) > 1:
            raise NotImplementedError("FusedAdam: per-parameter step counts differ (%s); one flat update has one step" % sorted(steps))
        self.step_count = steps.pop() if steps else 0
        self._bind_state()


class ExponentialLR(torch.optim.lr_scheduler._LRScheduler):
    """lr = base_lr * gamma ** (epoch / decay_epochs), stepped once per epoch (lr_scheduler.py:68-79, train.py:66)."""

    def __init__(self, optimizer, decay_epochs, gamma=0.1, last_epoch=-1):
        self.decay_epochs, self.gamma = decay_epochs, gamma
        super().__init__(optimizer, last_epoch)

    def get_lr(self):
        return [b * self.gamma ** (self.last_epoch / self.decay_epochs) for b in self.base_lrs]


class MultiStepLR(torch.optim.lr_scheduler.MultiStepLR):
    """scheduler type "multi_step" (scheduler.py:11-14): torch's own MultiStepLR semantics."""


def make_optimizer(cfg, net, clip=None):
    """src/train/optimizer.py:8-28 for cfg.train.optim == "adam" (lego.yaml:68)."""
    tr = cfg.train
    if "adam" not in tr.optim or tr.optim != "adam":
        raise NotImplementedError("nerf_b200 fuses torch.optim.Adam only (cfg.train.optim = %r)" % tr.optim)
    return FusedAdam(net, lr=tr.lr, eps=getattr(tr, "eps", 1e-8), weight_decay=tr.weight_decay, clip=clip)


def make_lr_scheduler(cfg, optimizer):
    """src/train/scheduler.py:9-21."""
    sc = cfg.train.scheduler
    if sc.type == "multi_step":
        return MultiStepLR(optimizer, milestones=list(sc.milestones), gamma=sc.gamma)
    if sc.type == "exponential":
        return ExponentialLR(optimizer, decay_epochs=sc.decay_epochs, gamma=sc.gamma)
    raise NotImplementedError("scheduler type %r" % sc.type)


def set_lr_scheduler(cfg, scheduler):
    """src/train/scheduler.py:24-30: re-apply the config's schedule constants after a checkpoint was loaded."""
    from collections import Counter
    sc = cfg.train.scheduler
    if sc.type == "multi_step":
        scheduler.milestones = Counter(sc.milestones)
    elif sc.type == "exponential":
        scheduler.decay_epochs = sc.decay_epochs
    scheduler.gamma = sc.gamma
