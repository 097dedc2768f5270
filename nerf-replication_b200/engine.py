"""Engine: Python face of one ``NerfEngine`` handle (one per process / GPU).

PyTorch is used for device memory and streams only: every method passes raw device pointers and the
current CUDA stream to the C ABI (include/nerf_b200.h).  Nothing here computes on the host and there is
no fallback path -- a missing library or GPU raises.
"""
import ctypes as C

import torch

from . import _lib as L


class EngineError(RuntimeError):
    pass


def _ptr(t, dtype=torch.float32, name="tensor", device=None):
    if t is None:
        return None
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise EngineError("%s must be a CUDA tensor" % name)
    if t.dtype != dtype:
        raise EngineError("%s must be %s, got %s" % (name, dtype, t.dtype))
    if not t.is_contiguous():
        raise EngineError("%s must be contiguous" % name)
    if device is not None and t.device.index != device:
        raise EngineError("%s is on cuda:%s, engine is on cuda:%d" % (name, t.device.index, device))
    return t.data_ptr()


class Engine:
    """Owns a NerfEngine handle.  cfg keys mirror Renderer.__init__ (volume_renderer.py:9-24)."""

    def __init__(self, n_samples=64, n_importance=128, white_bkgd=True, perturb=False, train_u=False,
                 precision="bf16", t_near=2.0, t_far=6.0, seed=0, device=None, fast_sampling=False, weights_threshold=0.25):
        if not torch.cuda.is_available():
            raise EngineError("nerf_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
        self.lib = L.load()
        self.device = torch.cuda.current_device() if device is None else int(device)
        self.cfg = L.NerfConfig()
        self._set_cfg_fields(n_samples, n_importance, white_bkgd, perturb, train_u, precision, t_near, t_far, seed,
                             fast_sampling, weights_threshold)
        h = C.c_void_p()
        rc = self.lib.nerf_create(C.byref(self.cfg), self.device, C.byref(h))
        if rc != 0:
            raise EngineError("nerf_create failed (%d): %s" % (rc, self.lib.nerf_last_error(None).decode()))
        self.h = h
        self._flat = None

    # -- plumbing -------------------------------------------------------------------------------
    def _set_cfg_fields(self, n_samples, n_importance, white_bkgd, perturb, train_u, precision, t_near, t_far, seed,
                        fast_sampling=False, weights_threshold=0.25):
        c = self.cfg
        c.n_samples, c.n_importance = int(n_samples), int(n_importance)
        c.white_bkgd, c.perturb, c.train_u = int(bool(white_bkgd)), int(bool(perturb)), int(bool(train_u))
        c.precision = {"bf16": L.PRECISION_BF16, "fp32": L.PRECISION_FP32}[precision]
        c.t_near, c.t_far, c.seed = float(t_near), float(t_far), int(seed)
        c.fast_sampling, c.weights_threshold = int(fast_sampling), float(weights_threshold)     # 0 off, 1 mask, 2 skip masked rows

    @property
    def precision(self):
        return "fp32" if self.cfg.precision == L.PRECISION_FP32 else "bf16"

    def configure(self, **kw):
        c = self.cfg
        cur = dict(n_samples=c.n_samples, n_importance=c.n_importance, white_bkgd=c.white_bkgd, perturb=c.perturb,
                   train_u=c.train_u, precision=self.precision, t_near=c.t_near, t_far=c.t_far, seed=c.seed,
                   fast_sampling=c.fast_sampling, weights_threshold=c.weights_threshold)
        cur.update(kw)
        self._set_cfg_fields(**cur)
        self._check(self.lib.nerf_set_config(self.h, C.byref(self.cfg)))

    def _check(self, rc):
        if rc != 0:
            raise EngineError("nerf_b200 error %d: %s" % (rc, self.lib.nerf_last_error(self.h).decode()))

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _p(self, t, dtype=torch.float32, name="tensor"):
        return _ptr(t, dtype, name, self.device)

    def close(self):
        if getattr(self, "h", None) is not None and self.h:
            self.lib.nerf_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def launch_count(self):
        return int(self.lib.nerf_launch_count(self.h))

    def profile(self, on=True):
        self._check(self.lib.nerf_profile_enable(self.h, int(on)))

    def profile_read(self):
        """-> (ms spent in fused-MLP launches, launches, MLP rows) since the last read; synchronises."""
        ms, n, rows = C.c_double(), C.c_int64(), C.c_int64()
        self._check(self.lib.nerf_profile_read(self.h, C.byref(ms), C.byref(n), C.byref(rows)))
        return ms.value, n.value, rows.value

    def profile_read_classes(self):
        """-> {"fwd"|"dgrad"|"wgrad": (ms, launches, rows)} since the last read; synchronises."""
        ms, n, rows = (C.c_double * 3)(), (C.c_int64 * 3)(), (C.c_int64 * 3)()
        self._check(self.lib.nerf_profile_read_classes(self.h, ms, n, rows))
        return {k: (ms[i], n[i], rows[i]) for i, k in enumerate(("fwd", "dgrad", "wgrad"))}

    # -- parameters -----------------------------------------------------------------------------
    def set_params(self, flat):
        """flat: fp32 CUDA tensor [NERF_PARAMS_TOTAL] in state_dict order; kept alive by the engine."""
        if flat.numel() != L.NERF_PARAMS_TOTAL:
            raise EngineError("flat parameter buffer must have %d elements" % L.NERF_PARAMS_TOTAL)
        self._check(self.lib.nerf_set_params(self.h, self._p(flat, name="flat"), self._stream()))
        self._flat = flat

    # -- stages ---------------------------------------------------------------------------------
    def get_rays(self, c2w, H, W, focal, cx=None, cy=None, pixel_ids=None, pix0=0, n=None):
        c2w = [float(x) for row in c2w[:3] for x in row[:4]]
        arr = (C.c_double * 12)(*c2w)
        if pixel_ids is not None:
            n = pixel_ids.numel()
        elif n is None:
            n = H * W - pix0
        o = torch.empty(n, 3, device="cuda:%d" % self.device)
        d = torch.empty_like(o)
        self._check(self.lib.nerf_get_rays(self.h, arr, H, W, float(focal), W / 2 if cx is None else cx,
                                           H / 2 if cy is None else cy, self._p(pixel_ids, torch.int64, "pixel_ids"),
                                           pix0, n, o.data_ptr(), d.data_ptr(), self._stream()))
        return o, d

    def stratified_t(self, n, S, jitter=None):
        t = torch.empty(n, S, device="cuda:%d" % self.device)
        self._check(self.lib.nerf_stratified_t(self.h, n, S, self._p(jitter, name="jitter"), t.data_ptr(), self._stream()))
        return t

    def freq_encode(self, x, n_freqs):
        x = x.reshape(-1, 3)
        out = torch.empty(x.shape[0], 3 + 6 * n_freqs, device=x.device)
        self._check(self.lib.nerf_freq_encode(self.h, self._p(x, name="x"), x.shape[0], n_freqs, out.data_ptr(), self._stream()))
        return out

    def mlp_forward(self, which_net, pts, viewdirs, mask=None):
        R, S, _ = pts.shape
        raw = torch.empty(R, S, 4, device=pts.device)
        if mask is not None:
            mask = mask.to(torch.uint8).contiguous()
        self._check(self.lib.nerf_mlp_forward(self.h, which_net, self._p(pts, name="pts"), self._p(viewdirs, name="viewdirs"),
                                              self._p(mask, torch.uint8, "mask"), R, S, raw.data_ptr(), self._stream()))
        return raw

    def composite(self, raw, t, white_bkgd=True, want_weights=False):
        n, S = t.shape
        dev = raw.device
        rgb, depth, acc = torch.empty(n, 3, device=dev), torch.empty(n, device=dev), torch.empty(n, device=dev)
        w = torch.empty(n, S, device=dev) if want_weights else None
        self._check(self.lib.nerf_composite(self.h, self._p(raw, name="raw"), self._p(t, name="t"), n, S, int(white_bkgd),
                                            rgb.data_ptr(), depth.data_ptr(), acc.data_ptr(),
                                            None if w is None else w.data_ptr(), self._stream()))
        return rgb, depth, acc, w

    def sample_pdf(self, raw_c, t_c, n_importance, u=None, cdf_in=None):
        n, Nc = t_c.shape
        Nf, dev = n_importance, t_c.device
        out = dict(t_fine=torch.empty(n, Nf, device=dev),
                   inds=torch.empty(n, Nf, dtype=torch.int64, device=dev),
                   below=torch.empty(n, Nf, dtype=torch.int64, device=dev),
                   above=torch.empty(n, Nf, dtype=torch.int64, device=dev),
                   cdf=torch.empty(n, Nc - 1, device=dev),
                   t_sorted=torch.empty(n, Nc + Nf, device=dev),
                   sort_idx=torch.empty(n, Nc + Nf, dtype=torch.int32, device=dev))
        self._check(self.lib.nerf_sample_pdf(self.h, self._p(raw_c, name="raw_c"), self._p(t_c, name="t_c"), n, Nc, Nf,
                                             self._p(u, name="u"), self._p(cdf_in, name="cdf_in"),
                                             out["t_fine"].data_ptr(), out["inds"].data_ptr(), out["below"].data_ptr(),
                                             out["above"].data_ptr(), out["cdf"].data_ptr(), out["t_sorted"].data_ptr(),
                                             out["sort_idx"].data_ptr(), self._stream()))
        return out

    def sample_pdf_prune(self, raw_c, t_c, n_importance, weights_threshold=0.25, u=None, cdf_in=None):
        """a-M: sample_pdf with the reference's fast_sampling mask -> t_fine, valid (bool [n,Nf]), mask_sorted (bool [n,Nc+Nf]),
        t_sorted, sort_idx."""
        n, Nc = t_c.shape
        Nf, dev = n_importance, t_c.device
        out = dict(t_fine=torch.empty(n, Nf, device=dev), valid=torch.empty(n, Nf, dtype=torch.uint8, device=dev),
                   mask_sorted=torch.empty(n, Nc + Nf, dtype=torch.uint8, device=dev),
                   t_sorted=torch.empty(n, Nc + Nf, device=dev), sort_idx=torch.empty(n, Nc + Nf, dtype=torch.int32, device=dev))
        self._check(self.lib.nerf_sample_pdf_prune(self.h, self._p(raw_c, name="raw_c"), self._p(t_c, name="t_c"), n, Nc, Nf,
                                                   self._p(u, name="u"), self._p(cdf_in, name="cdf_in"), float(weights_threshold),
                                                   out["t_fine"].data_ptr(), out["valid"].data_ptr(), out["mask_sorted"].data_ptr(),
                                                   out["t_sorted"].data_ptr(), out["sort_idx"].data_ptr(), self._stream()))
        out["valid"], out["mask_sorted"] = out["valid"].bool(), out["mask_sorted"].bool()
        return out

    # -- render ---------------------------------------------------------------------------------
    def _render_out(self, n, dev, full):
        t = dict(rgb=torch.empty(n, 3, device=dev), depth=torch.empty(n, device=dev))
        if full:
            t.update(acc=torch.empty(n, device=dev), rgb_coarse=torch.empty(n, 3, device=dev),
                     depth_coarse=torch.empty(n, device=dev), acc_coarse=torch.empty(n, device=dev))
        ro = L.NerfRenderOut()
        for k, v in t.items():
            setattr(ro, k, v.data_ptr())
        return t, ro

    def render_rays(self, rays_o, rays_d, jitter=None, u=None, full=False):
        """rays_o/rays_d [n,3] fp32 CUDA -> dict of maps (rgb, depth[, acc, *_coarse])."""
        n = rays_o.shape[0]
        t, ro = self._render_out(n, rays_o.device, full)
        self._check(self.lib.nerf_render_rays(self.h, self._p(rays_o, name="rays_o"), self._p(rays_d, name="rays_d"), n,
                                              self._p(jitter, name="jitter"), self._p(u, name="u"), C.byref(ro), self._stream()))
        return t

    def render_camera(self, c2w, H, W, focal, cx=None, cy=None, pix0=0, n=None, full=False):
        n = H * W - pix0 if n is None else n
        c2w = [float(x) for row in c2w[:3] for x in row[:4]]
        t, ro = self._render_out(n, torch.device("cuda:%d" % self.device), full)
        self._check(self.lib.nerf_render_camera(self.h, (C.c_double * 12)(*c2w), H, W, float(focal),
                                                W / 2 if cx is None else cx, H / 2 if cy is None else cy, pix0, n,
                                                C.byref(ro), self._stream()))
        return t

    def render_rays_host(self, rays_o, rays_d, rgb_out=None, depth_out=None):
        """Host (CPU) tensors in, host tensors out: the end-to-end path with H2D/D2H inside the call."""
        for x in (rays_o, rays_d):
            if x.is_cuda or x.dtype != torch.float32 or not x.is_contiguous():
                raise EngineError("render_rays_host takes contiguous fp32 CPU tensors")
        n = rays_o.shape[0]
        rgb = torch.empty(n, 3) if rgb_out is None else rgb_out
        depth = torch.empty(n) if depth_out is None else depth_out
        self._check(self.lib.nerf_render_rays_host(self.h, rays_o.data_ptr(), rays_d.data_ptr(), n, rgb.data_ptr(), depth.data_ptr()))
        return rgb, depth

    # -- training -------------------------------------------------------------------------------
    def forward_backward(self, rays_o, rays_d, rgb_gt, flat_grad, jitter=None, u=None, grad_scale=1.0, want_rgb=False, loss=None):
        """One fused step body: flat_grad = grad_scale * d MSE / d params (overwritten), -> loss (device scalar).  Passing the
        same ``loss`` / ``flat_grad`` tensors (and n) on every call lets the engine replay the step as one CUDA graph."""
        n = rays_o.shape[0]
        dev = "cuda:%d" % self.device
        loss = torch.empty((), device=dev) if loss is None else loss
        rgb = torch.empty(n, 3, device=dev) if want_rgb else None

        def inp(t, name):        # rays / colours: CUDA tensors, or (pinned) host tensors that the engine stages itself
            if t.is_cuda:
                return self._p(t, name=name)
            if t.dtype != torch.float32 or not t.is_contiguous():
                raise EngineError("%s must be a contiguous fp32 tensor" % name)
            return t.data_ptr()

        self._check(self.lib.nerf_forward_backward(self.h, inp(rays_o, "rays_o"), inp(rays_d, "rays_d"),
                                                   inp(rgb_gt, "rgb_gt"), n, self._p(jitter, name="jitter"),
                                                   self._p(u, name="u"), float(grad_scale), loss.data_ptr(),
                                                   self._p(flat_grad, name="flat_grad"),
                                                   None if rgb is None else rgb.data_ptr(), self._stream()))
        return (loss, rgb) if want_rgb else loss

    def train_forward(self, rays_o, rays_d, jitter=None, u=None):
        """-> (rgb, depth, forward_id); the handle keeps this forward's activations until its train_backward."""
        n = rays_o.shape[0]
        rgb, depth = torch.empty(n, 3, device=rays_o.device), torch.empty(n, device=rays_o.device)
        fid = C.c_uint64(0)
        self._check(self.lib.nerf_train_forward(self.h, self._p(rays_o, name="rays_o"), self._p(rays_d, name="rays_d"), n,
                                                self._p(jitter, name="jitter"), self._p(u, name="u"), rgb.data_ptr(),
                                                depth.data_ptr(), C.byref(fid), self._stream()))
        return rgb, depth, fid.value

    def train_backward(self, rays_o, rays_d, d_rgb, d_depth, flat_grad, zero_grad=True, forward_id=0):
        """forward_id (from train_forward): raises if a later forward has overwritten that forward's activations."""
        n = rays_o.shape[0]
        self._check(self.lib.nerf_train_backward(self.h, self._p(rays_o, name="rays_o"), self._p(rays_d, name="rays_d"), n,
                                                 self._p(d_rgb, name="d_rgb"), self._p(d_depth, name="d_depth"),
                                                 self._p(flat_grad, name="flat_grad"), int(zero_grad), int(forward_id), self._stream()))

    def stream_wait_fine_grads(self, stream):
        """Makes `stream` (a torch.cuda.Stream) wait until the fine net's half of the last backward's flat_grad is final."""
        self._check(self.lib.nerf_stream_wait_fine_grads(self.h, C.c_void_p(stream.cuda_stream)))

    def set_graph_mode(self, on=True):
        self._check(self.lib.nerf_set_graph_mode(self.h, int(bool(on))))

    def clip_adam(self, flat_param, flat_grad, m, v, step, lr=5e-4, beta1=0.9, beta2=0.999, eps=1e-8, clip=40.0):
        self._check(self.lib.nerf_clip_adam(self.h, self._p(flat_param, name="flat_param"), self._p(flat_grad, name="flat_grad"),
                                            self._p(m, name="m"), self._p(v, name="v"), lr, beta1, beta2, eps, clip, int(step),
                                            self._stream()))

    # -- ray bank + evaluator sums (csrc/bank.cu) ------------------------------------
    def bank_create(self, n_views, H, W, focal, cx=None, cy=None, uint8=False):
        bank = C.c_int(-1)
        self._check(self.lib.nerf_bank_create(self.h, int(n_views), int(H), int(W), float(focal), W / 2 if cx is None else cx,
                                              H / 2 if cy is None else cy, int(bool(uint8)), C.byref(bank)))
        return bank.value

    def bank_destroy(self, bank):
        self._check(self.lib.nerf_bank_destroy(self.h, int(bank)))

    def bank_set_view(self, bank, view, image, c2w):
        """image: contiguous host array [H,W,3], float32 (0..255) or uint8 to match the bank; c2w: 4x4 or 3x4."""
        c2w = [float(x) for row in c2w[:3] for x in row[:4]]
        self._check(self.lib.nerf_bank_set_view(self.h, int(bank), int(view), image.ctypes.data, (C.c_double * 12)(*c2w),
                                                self._stream()))

    _BANK_MODES = {"ids": 0, "range": 1, "random": 2}

    def bank_sample(self, bank, view, n, mode="random", pixel_ids=None, pix0=0, draw=0, want_ids=False):
        """Dataset.__getitem__ on the device -> dict(rays_o, rays_d, colors[, ids])."""
        dev = "cuda:%d" % self.device
        o, d, c = torch.empty(n, 3, device=dev), torch.empty(n, 3, device=dev), torch.empty(n, 3, device=dev)
        ids = torch.empty(n, dtype=torch.int64, device=dev) if want_ids else None
        self._check(self.lib.nerf_bank_sample(self.h, int(bank), int(view), self._BANK_MODES[mode],
                                              self._p(pixel_ids, torch.int64, "pixel_ids"), int(pix0), int(draw), int(n),
                                              o.data_ptr(), d.data_ptr(), c.data_ptr(), None if ids is None else ids.data_ptr(),
                                              self._stream()))
        out = {"rays_o": o, "rays_d": d, "colors": c}
        if want_ids:
            out["ids"] = ids
        return out

    def bank_permutation_host(self, seed, draw, n_pix, n):
        import numpy as np
        ids = np.empty(n, dtype=np.int64)
        rc = self.lib.nerf_bank_permutation_host(int(seed), int(draw), int(n_pix), int(n), ids.ctypes.data)
        if rc != 0:
            raise EngineError("nerf_bank_permutation_host: bad argument")
        return ids

    def bank_train_step(self, bank, view, n, flat_grad, mode="random", pixel_ids=None, pix0=0, draw=0, jitter=None, u=None,
                        grad_scale=1.0, loss=None):
        loss = torch.empty((), device="cuda:%d" % self.device) if loss is None else loss
        self._check(self.lib.nerf_bank_train_step(self.h, int(bank), int(view), self._BANK_MODES[mode],
                                                  self._p(pixel_ids, torch.int64, "pixel_ids"), int(pix0), int(draw), int(n),
                                                  self._p(jitter, name="jitter"), self._p(u, name="u"), float(grad_scale),
                                                  loss.data_ptr(), self._p(flat_grad, name="flat_grad"), None, self._stream()))
        return loss

    def eval_image(self, rgb_pred, rgb_gt):
        """-> device tensors (sse float64 [], sq8 int64 []) of src/evaluators/nerf.py:23-30,96-103 (see nerf_eval_image)."""
        if rgb_pred.numel() != rgb_gt.numel():
            raise EngineError("eval_image: rgb_pred and rgb_gt differ in size")
        dev = "cuda:%d" % self.device
        sse, sq8 = torch.empty((), dtype=torch.float64, device=dev), torch.empty((), dtype=torch.int64, device=dev)
        self._check(self.lib.nerf_eval_image(self.h, self._p(rgb_pred, name="rgb_pred"), self._p(rgb_gt, name="rgb_gt"),
                                             rgb_pred.numel(), sse.data_ptr(), sq8.data_ptr(), self._stream()))
        return sse, sq8

    def eval_image_static(self, rgb_pred, rgb_gt, sse=None, sq8=None):
        """nerf_eval_image_static: the same sums as eval_image, with the reduction scratch in a static device array (no
        cudaMalloc in the call; no counterpart in the reference).  Pass preallocated `sse` (float64 []) and `sq8` (int64 [])
        device tensors to keep this wrapper allocation-free as well."""
        if rgb_pred.numel() != rgb_gt.numel():
            raise EngineError("eval_image_static: rgb_pred and rgb_gt differ in size")
        dev = "cuda:%d" % self.device
        sse = torch.empty((), dtype=torch.float64, device=dev) if sse is None else sse
        sq8 = torch.empty((), dtype=torch.int64, device=dev) if sq8 is None else sq8
        self._check(self.lib.nerf_eval_image_static(self.h, self._p(rgb_pred, name="rgb_pred"), self._p(rgb_gt, name="rgb_gt"),
                                                    rgb_pred.numel(), sse.data_ptr(), sq8.data_ptr(), self._stream()))
        return sse, sq8

    def eval_ssim(self, rgb_pred, rgb_gt, H, W, win_size=7):
        """-> device float64 []: mean SSIM as src/evaluators/nerf.py:71-77 computes it (see nerf_eval_ssim)."""
        if rgb_pred.numel() != H * W * 3 or rgb_gt.numel() != H * W * 3:
            raise EngineError("eval_ssim takes [H,W,3] images")
        s = torch.empty((), dtype=torch.float64, device="cuda:%d" % self.device)
        self._check(self.lib.nerf_eval_ssim(self.h, self._p(rgb_pred, name="rgb_pred"), self._p(rgb_gt, name="rgb_gt"), int(H), int(W),
                                            int(win_size), s.data_ptr(), self._stream()))
        return s / (3 * (H - win_size + 1) * (W - win_size + 1))

    # -- debug ----------------------------------------------------------------------------------
    def umma_probe(self, variant, a, b):
        if variant >= 4:                       # MN-major operands: a = At [K, 128], b = Bt [K, N]
            K, M = a.shape
            N = b.shape[1]
        else:
            M, K = a.shape
            N = b.shape[0]
        d = torch.zeros(M, N, device=a.device)
        rc = L.load_probe().nerf_debug_umma_probe(variant, self._p(a, name="a"), self._p(b, name="b"), d.data_ptr(), M, N, K, self._stream())
        if rc != 0:
            raise EngineError("umma probe launch failed (%d)" % rc)
        return d

    def watchdog(self):
        out = (C.c_uint * 64)()      # [0..8): first record, [8..64): one word per stuck waiter
        self.lib.nerf_debug_tc_watchdog(out)
        return list(out)

    def waitprof(self, on):
        """Diagnostic: start (on=True) / stop the barrier wait profile; returns what was accumulated so far as
        {"fwd"|"bwd": {tag_bucket: (cycles, waits)}} (tag buckets: tc_common.cuh waitprof_add, tag >> 4)."""
        out = (C.c_ulonglong * 1024)()
        self._check(self.lib.nerf_debug_waitprof(int(on), out))
        res = {}
        for k, name in enumerate(("fwd", "bwd")):
            res[name] = {b: (out[512 * k + b], out[512 * k + 256 + b]) for b in range(256) if out[512 * k + 256 + b]}
        return res
