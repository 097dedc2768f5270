"""Drop-in for ``src.models.nerf.network`` (reference network.py:9-258).

``Network`` keeps the reference's class surface -- an ``nn.Module`` with sub-modules ``model`` /
``model_fine`` (class ``NeRF``) whose parameters have the same names, shapes and registration order, so
``torch.manual_seed(s); Network()`` is bit-identical to the reference and ``latest.pth`` loads
(net_utils.py:346-379) -- but the arithmetic runs in the CUDA engine.  Once the module is on a GPU all 48
parameters become views into ONE flat fp32 buffer (state_dict order) that the engine reads directly; the
gradients likewise live in one flat buffer (``flat_grad``), which is what the all-reduce and the fused Adam
step operate on.
"""
import torch
import torch.nn as nn

from . import _lib as L
from .config import get_cfg
from .engine import Engine


class NeRF(nn.Module):
    """Parameter container with the reference layout (network.py:9-47).  ``forward`` is not used by the
    engine path; it is kept (plain torch) so the module remains a self-contained definition of the math."""

    def __init__(self, D=8, W=256, input_ch=63, input_ch_views=27, skips=(4,), use_viewdirs=True):
        super().__init__()
        if (D, W, input_ch, input_ch_views, list(skips), bool(use_viewdirs)) != (8, 256, 63, 27, [4], True):
            raise NotImplementedError("nerf_b200 implements exactly D=8, W=256, skips=[4], PE 10/4, use_viewdirs=True "
                                      "(configs/nerf/lego.yaml); no fallback for other shapes")
        self.D, self.W, self.input_ch, self.input_ch_views = D, W, input_ch, input_ch_views
        self.skips, self.use_viewdirs = list(skips), True
        self.pts_linears = nn.ModuleList(
            [nn.Linear(input_ch, W)] + [nn.Linear(W + input_ch if i in self.skips else W, W) for i in range(D - 1)])
        self.views_linears = nn.ModuleList([nn.Linear(input_ch_views + W, W // 2)])
        self.feature_linear = nn.Linear(W, W)
        self.alpha_linear = nn.Linear(W, 1)
        self.rgb_linear = nn.Linear(W // 2, 3)

    def forward(self, x):
        pts, views = torch.split(x, [self.input_ch, self.input_ch_views], dim=-1)
        h = pts
        for i, lin in enumerate(self.pts_linears):
            h = torch.relu(lin(h))
            if i in self.skips:
                h = torch.cat([pts, h], -1)
        alpha = self.alpha_linear(h)
        h = torch.cat([self.feature_linear(h), views], -1)
        h = torch.relu(self.views_linears[0](h))
        return torch.cat([self.rgb_linear(h), alpha], -1)


class Network(nn.Module):
    def __init__(self, cfg=None, precision=None):
        super().__init__()
        cfg = get_cfg() if cfg is None else cfg
        ta, nw = cfg.task_arg, cfg.network
        self.N_samples, self.N_importance = ta.N_samples, ta.N_importance
        self.chunk = ta.chunk_size            # accepted and ignored: the fused kernel needs no chunking
        self.batch_size = ta.N_rays
        self.white_bkgd, self.use_viewdirs = ta.white_bkgd, ta.use_viewdirs
        self.device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.sample_size = getattr(cfg, "sample_size", 64)
        self.rays_size = getattr(cfg, "rays_size", 160000)
        for enc, want in ((nw.xyz_encoder, 10), (nw.dir_encoder, 4)):
            if enc.type != "frequency" or enc.input_dim != 3 or enc.freq != want:
                raise NotImplementedError("nerf_b200 implements the frequency encoders 10 (xyz) / 4 (dir) only")
        self.input_ch, self.input_ch_views = 63, 27
        self.embed_fn = lambda x: self.engine().freq_encode(x, 10)
        self.embeddirs_fn = lambda x: self.engine().freq_encode(x, 4)
        kw = dict(D=nw.nerf.D, W=nw.nerf.W, input_ch=63, input_ch_views=27, skips=list(nw.nerf.skips),
                  use_viewdirs=self.use_viewdirs)
        self.model = NeRF(**kw)        # coarse
        self.model_fine = NeRF(**kw)   # fine
        self.precision = precision or getattr(cfg, "precision", "bf16")
        self._engine = None
        self._flat = None
        self._flat_grad = None
        self._packed_version = -1

    # -- flat storage ---------------------------------------------------------------------------
    def _apply(self, fn, *a, **k):
        out = super()._apply(fn, *a, **k)
        self._flat = None
        p0 = next(self.parameters())
        if p0.is_cuda:
            self._flatten()
            self.device = p0.device
        return out

    def _flatten(self):
        params = list(self.parameters())
        dev = params[0].device
        flat = torch.empty(L.NERF_PARAMS_TOTAL, device=dev, dtype=torch.float32)
        o = 0
        for p in params:
            n = p.numel()
            flat[o:o + n].copy_(p.detach().reshape(-1).to(torch.float32))
            p.data = flat[o:o + n].view(p.shape)
            o += n
        assert o == L.NERF_PARAMS_TOTAL
        self._flat, self._flat_grad, self._packed_version = flat, None, -1
        if self._engine is not None and self._engine.device != dev.index:
            self._engine.close()
            self._engine = None

    @property
    def flat_params(self):
        if self._flat is None:
            raise RuntimeError("Network must be moved to a CUDA device first (net.cuda()); nerf_b200 has no CPU path")
        return self._flat

    @property
    def flat_grad(self):
        """One fp32 buffer holding every gradient; each parameter's .grad is a view into it."""
        if self._flat_grad is None:
            self._flat_grad = torch.zeros_like(self.flat_params)
            o = 0
            for p in self.parameters():
                n = p.numel()
                p.grad = self._flat_grad[o:o + n].view(p.shape)
                o += n
        return self._flat_grad

    def engine(self, **cfg_kw):
        flat = self.flat_params
        if self._engine is None:
            self._engine = Engine(precision=self.precision, device=flat.device.index, **cfg_kw)
        elif cfg_kw:
            self._engine.configure(**cfg_kw)
        ver = self._weights_version()
        if self._packed_version != ver or self._engine._flat is not flat:
            self._engine.set_params(flat)          # re-packs the bf16 weight images
            self._packed_version = ver
        return self._engine

    def _weights_version(self):
        # parameters are .data views of the flat buffer: each keeps its own in-place version counter
        return self._flat._version + sum(p._version for p in self.parameters())

    def mark_packed(self):
        """The fused clip+Adam step repacks inside the engine; record that the packed copy is current."""
        self._packed_version = self._weights_version()

    # -- reference interface ----------------------------------------------------------------------
    def batchify(self, fn, chunk):
        return lambda inputs: torch.cat([fn(inputs[i:i + chunk]) for i in range(0, inputs.shape[0], chunk)], 0)

    def forward(self, inputs, viewdirs, valid_mask=None, model=""):
        """inputs [R,S,3], viewdirs [R,3], valid_mask [R,S] bool|None -> raw [R,S,4] (network.py:199-258)."""
        eng = self.engine()
        which = L.NET_FINE if model == "fine" else L.NET_COARSE
        return eng.mlp_forward(which, inputs.contiguous().float(), viewdirs.contiguous().float(), valid_mask)
