"""Drop-in for ``src.models.nerf.renderer.volume_renderer`` (reference volume_renderer.py:9-432).

``Renderer(net).render(batch)`` returns the reference's 2-tuple ``(rgb [B*N,3], depth [B*N])`` of the final
pass.  Under ``torch.no_grad()`` it is one call into the fused CUDA pipeline; with gradients enabled it goes
through a ``torch.autograd.Function`` whose backward runs the engine's hand-written backward kernels and
hands each ``nn.Parameter`` its gradient, so the reference's ``loss.backward(); clip_grad_value_; optimizer.step``
lines (trainer.py:56-60) work unchanged.  Restriction: the engine keeps the activations of ONE grad-enabled render, so
each such render must be followed by its own backward before the next one (two renders, then one backward over both,
raises).  ``render(batch, return_dict=True)`` (or ``cfg.return_dict``) is the
opt-in extension that also returns acc and the coarse maps.
"""
import torch

from .config import get_cfg, knob


class _RenderFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, renderer, rays_o, rays_d, jitter, u, *params):
        net = renderer.net
        eng = renderer._engine()
        rgb, depth, ctx.forward_id = eng.train_forward(rays_o, rays_d, jitter, u)
        ctx.renderer, ctx.rays = renderer, (rays_o, rays_d)
        return rgb, depth

    @staticmethod
    def backward(ctx, d_rgb, d_depth):
        net = ctx.renderer.net
        eng = ctx.renderer._engine()
        rays_o, rays_d = ctx.rays
        scratch = torch.empty_like(net.flat_params)
        d_rgb = None if d_rgb is None else d_rgb.contiguous()
        d_depth = None if d_depth is None else d_depth.contiguous()
        # one forward, then its backward: the engine keeps the activations of the latest grad-enabled render only and
        # raises (instead of silently using another batch's activations) if a second render ran in between
        eng.train_backward(rays_o, rays_d, d_rgb, d_depth, scratch, zero_grad=True, forward_id=ctx.forward_id)
        grads, o = [], 0
        for p in net.parameters():
            n = p.numel()
            grads.append(scratch[o:o + n].view(p.shape) if p.requires_grad else None)
            o += n
        return (None, None, None, None, None) + tuple(grads)


class Renderer:
    def __init__(self, net, cfg=None):
        cfg = get_cfg() if cfg is None else cfg
        self.net = net
        self.N_samples = knob(cfg, "N_samples", 64)
        self.chunk_size = knob(cfg, "chunk_size", 1024)     # accepted, unused (as in the reference)
        self.white_bkgd = bool(knob(cfg, "white_bkgd", True))
        self.N_importance = knob(cfg, "N_importance", 128)
        self.perturb = bool(knob(cfg, "perturb", True))
        self.sample_size, self.rays_size = 64, 160000
        self.task = getattr(cfg, "task", "test")
        self.perturb = self.perturb if self.task == "train" else False     # volume_renderer.py:21-22
        self.fast_sampling = bool(knob(cfg, "fast_sampling", False))
        self.weights_threshold = knob(cfg, "weights_threshold", 0.25)
        self.return_dict = bool(getattr(cfg, "return_dict", False))
        # fast_sampling (ESS/ERT, volume_renderer.py:116-244,359-374): the engine builds the reference's validity mask in the
        # sample_pdf kernel; masked fine samples get raw = 0 (rendering skips them altogether, training masks them)

    def _engine(self):
        return self.net.engine(n_samples=self.N_samples, n_importance=self.N_importance, white_bkgd=self.white_bkgd,
                               perturb=self.perturb, train_u=(self.task == "train"), fast_sampling=2 if self.fast_sampling else 0,
                               weights_threshold=float(self.weights_threshold))

    def render(self, batch, return_dict=None, jitter=None, u=None):
        rays_o, rays_d = batch["rays_o"], batch["rays_d"]
        self.device = rays_o.device
        B, N_rays, _ = rays_o.shape
        rays_o = rays_o.reshape(B * N_rays, 3).contiguous().float()
        rays_d = rays_d.reshape(B * N_rays, 3).contiguous().float()
        want_dict = self.return_dict if return_dict is None else return_dict
        needs_grad = torch.is_grad_enabled() and any(p.requires_grad for p in self.net.parameters())
        if needs_grad:
            rgb, depth = _RenderFn.apply(self, rays_o, rays_d, jitter, u, *self.net.parameters())
            if not want_dict:
                return rgb, depth
            return {"rgb_map": rgb, "depth_map": depth}
        out = self._engine().render_rays(rays_o, rays_d, jitter=jitter, u=u, full=want_dict)
        if not want_dict:
            return out["rgb"], out["depth"]
        return {"rgb_map": out["rgb"], "depth_map": out["depth"], "acc_map": out["acc"],
                "rgb_map_coarse": out["rgb_coarse"], "depth_map_coarse": out["depth_coarse"],
                "acc_map_coarse": out["acc_coarse"]}
