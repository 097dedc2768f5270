"""GPU parity of the training step (trainers/nerf.py:27-33, trainer.py:56-60, train/optimizer.py:8-28):
hand-written backward kernels vs torch autograd on the oracle (the reference's autograd graph, including the
un-detached t_fine path -- SURVEY.md fact 8), the fused clip+Adam step vs torch.optim.Adam, and loss curves."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from gpu_util import flat_from_sd, make_engine, maxabs

pytestmark = pytest.mark.gpu


def _per_tensor(oracle, flat):
    out, o = {}, 0
    sd = oracle.init_state_dict(0)
    for k in oracle.state_dict_keys():
        n = sd[k].numel()
        out[k] = flat[o:o + n]
        o += n
    return out


@pytest.mark.parametrize("weights", ["fog", "fog_seed3", "fog_dense", "sharp", "seed0"])
def test_gradients_match_reference_autograd(oracle, weights):
    """Engine gradients vs torch autograd on the oracle (the reference graph incl. the un-detached t_fine path).
    fine net: 1e-3 of each tensor's max (or the fp32 noise floor where that is larger).
    coarse net: its only signal is d loss / d t_fine through the inverse CDF.  With near-empty bins (seed0 /
    sharp weights) that gradient is noise-dominated in fp32 -- the oracle itself moves by 20%..500% between fp32
    and fp64 (measured: oracle.grads_fp64) -- so there the bound is the oracle's own fp32-vs-fp64 distance;
    the "fog" weights (sigma ~ 1 everywhere, smooth CDF) are well conditioned and held to 2e-2 of max.  Two more such sets
    (another seed and ray batch; a denser fog) verify the un-detached t_fine path on more than one: there the oracle's own
    fp32-vs-fp64 distance is 0.8 % / 1.2 % of max (fog: 0.65 %), so the bound is max(2e-2, 2 x that distance)."""
    sd = {"fog": lambda: oracle.fog_state_dict(0), "fog_seed3": lambda: oracle.fog_state_dict(3),
          "fog_dense": lambda: oracle.fog_state_dict(0, bias=1.5), "sharp": lambda: oracle.init_state_dict(0, sharp=True),
          "seed0": lambda: oracle.init_state_dict(0)}[weights]()
    n = 32
    o, d, gt, _, _ = oracle.synthetic_train_batch({"fog_seed3": 5}.get(weights, 0), 0, n, 100, 100)
    tr = oracle.Trainer(sd)
    loss_ref = tr.forward_backward(o, d, gt)
    g_ref = tr.grads_flat()
    _, g64 = oracle.grads_fp64(sd, o, d, gt)
    eng = make_engine(oracle, sd, "fp32")
    grad = torch.zeros(1191688, device="cuda")
    loss = eng.forward_backward(o.cuda(), d.cuda(), gt.cuda(), grad)
    torch.cuda.synchronize()
    assert abs(float(loss) - float(loss_ref)) <= 1e-6 * float(loss_ref)
    if weights == "sharp":
        g = load_golden("train_sharp_32rays")          # the unmodified reference's own first-step loss
        assert abs(float(loss) - g["losses"][0]) <= 1e-5 * g["losses"][0]
    ref_t, got_t, f64_t = _per_tensor(oracle, g_ref), _per_tensor(oracle, grad.cpu()), _per_tensor(oracle, g64)
    bad = []
    for k in ref_t:
        scale = float(ref_t[k].abs().max())
        if k.startswith("model.") and any(s in k for s in ("views_linears", "feature_linear", "rgb_linear")):
            assert scale == 0.0 and float(got_t[k].abs().max()) == 0.0, k      # exactly-zero grads (fact 8)
            continue
        assert scale > 0, k
        noise = maxabs(ref_t[k], f64_t[k])                 # how far fp32 round-off alone moves this tensor
        coarse = k.startswith("model.")
        fixed = (2e-2 if coarse else 1e-3) * scale
        # ill-conditioned sets: +-1..4 ulp weight perturbations move the fp32 oracle's own coarse gradients by
        # 0.5%..8% of max per tensor (event-driven: single searchsorted flips / tiny denominators), so the coarse
        # bound there is max(noise floor, 15% of max); the fixed bounds apply to fog (coarse) and to every fine tensor
        if coarse and weights == "fog":
            tol = fixed
        elif coarse and weights.startswith("fog"):
            tol = max(fixed, 2.0 * noise)
        else:
            tol = max(fixed, 1.5 * noise, (0.15 * scale) if coarse else 0.0)
        err = maxabs(got_t[k], ref_t[k])
        if err > tol:
            bad.append((k, err, tol, scale, noise))
    assert not bad, bad
    eng.close()


def test_gradients_coarse_only(oracle):
    sd = oracle.init_state_dict(0, sharp=True)
    o, d, gt, _, _ = oracle.synthetic_train_batch(1, 0, 24, 100, 100)
    tr = oracle.Trainer(sd, n_importance=0)
    loss_ref = tr.forward_backward(o, d, gt)
    eng = make_engine(oracle, sd, "fp32", n_importance=0)
    grad = torch.zeros(1191688, device="cuda")
    loss = eng.forward_backward(o.cuda(), d.cuda(), gt.cuda(), grad)
    assert abs(float(loss) - float(loss_ref)) <= 1e-6 * float(loss_ref)
    g_ref = tr.grads_flat()
    assert maxabs(grad[:595844], g_ref[:595844]) <= 1e-3 * float(g_ref.abs().max())
    assert float(grad[595844:].abs().max()) == 0.0
    eng.close()


def test_clip_adam_matches_torch(oracle):
    gen = torch.Generator().manual_seed(4)
    n = 1191688
    p0 = torch.randn(n, generator=gen) * 0.1
    eng = make_engine(oracle, oracle.init_state_dict(0), "bf16")
    p = p0.clone().cuda()
    m, v = torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([ref], lr=5e-4, eps=1e-8)
    for step in range(1, 6):
        g = torch.randn(n, generator=gen) * (100.0 if step == 2 else 1e-3)      # step 2 exercises the +-40 clip
        ref.grad = g.clone()
        torch.nn.utils.clip_grad_value_([ref], 40.0)
        opt.step()
        eng.clip_adam(p, g.cuda(), m, v, step)
        torch.cuda.synchronize()
        assert maxabs(p, ref.detach()) <= 2e-7, step
    eng.close()


def test_autograd_dropin_backward(oracle):
    """loss.backward(); clip_grad_value_; optimizer.step() of trainer.py:56-60 work on the drop-in Renderer."""
    from nerf_replication_b200 import config as C
    from nerf_replication_b200.network import Network
    from nerf_replication_b200.renderer import Renderer
    cfg = C.default_cfg()
    net = Network(cfg=cfg, precision="fp32")
    net.load_state_dict(oracle.init_state_dict(0, sharp=True))
    net = net.cuda()
    r = Renderer(net, cfg=cfg)
    o, d, gt, _, _ = oracle.synthetic_train_batch(0, 0, 32, 100, 100)
    opt = torch.optim.Adam([{"params": [p]} for p in net.parameters()], lr=5e-4, eps=1e-8)
    rgb, depth = r.render({"rays_o": o.cuda()[None], "rays_d": d.cuda()[None]})
    loss = torch.nn.functional.mse_loss(rgb, gt.cuda())
    opt.zero_grad()
    loss.backward()
    tr = oracle.Trainer(oracle.init_state_dict(0, sharp=True))
    loss_ref = tr.forward_backward(o, d, gt)
    assert abs(float(loss) - float(loss_ref)) <= 1e-6 * float(loss_ref)
    g = torch.cat([p.grad.reshape(-1) for p in net.parameters()]).cpu()
    g_ref = tr.grads_flat()
    assert maxabs(g[595844:], g_ref[595844:]) <= 1e-3 * float(g_ref[595844:].abs().max())
    torch.nn.utils.clip_grad_value_(net.parameters(), 40)
    opt.step()
    with torch.no_grad():
        rgb2, _ = r.render({"rays_o": o.cuda()[None], "rays_d": d.cuda()[None]})
    assert maxabs(rgb2, rgb) > 0           # the engine saw the optimizer's in-place update


def test_loss_curve_tracks_oracle(oracle):
    """Loss curves within 1% (north_star) -- 40 steps x 64 rays here; the 1k-step run is profiles/loss_curve_*."""
    from nerf_replication_b200 import config as C
    from nerf_replication_b200.network import Network
    from nerf_replication_b200.trainer import FusedTrainer
    cfg = C.default_cfg()
    sd = oracle.init_state_dict(0)
    net = Network(cfg=cfg, precision="fp32")
    net.load_state_dict(sd)
    net = net.cuda()
    ft = FusedTrainer(net, cfg=cfg)
    tr = oracle.Trainer(sd)
    torch.set_num_threads(8)
    for s in range(40):
        o, d, gt, _, _ = oracle.synthetic_train_batch(s, 0, 64, 100, 100)
        l_ref = tr.step(o, d, gt)
        l = float(ft.step(o.cuda(), d.cuda(), gt.cuda()))
        assert abs(l - l_ref) <= 1e-2 * l_ref, (s, l, l_ref)


def _grad_report(oracle, got, ref):
    rows = []
    ref_t, got_t = _per_tensor(oracle, ref), _per_tensor(oracle, got)
    for k in ref_t:
        scale = float(ref_t[k].abs().max())
        err = maxabs(got_t[k], ref_t[k])
        rows.append((k, err, scale, err / scale if scale > 0 else (0.0 if err == 0 else float("inf"))))
    return rows


@pytest.mark.parametrize("weights", ["fog", "sharp"])
def test_bf16_tensor_core_backward_vs_oracle(oracle, weights):
    """bf16 training path: saving forward (tcgen05) + dgrad chain + wgrad kernels vs torch autograd on the oracle with
    the same bf16-representable weights.  Operands (activations, dZ) are bf16, accumulation fp32; the rounding of the
    gradient tiles compounds through the chain (measured on B200: 0.1% of max at rgb_linear, 1% at layers 5-7, 4% at
    layer 0), so fine-net tensors are held to 6e-2 of each tensor's max and a cosine similarity >= 0.999; the coarse
    net's gradient (amplified through the inverse CDF, see the fp32 test) to 0.15 / cos >= 0.99 on the fog set."""
    sd0 = oracle.fog_state_dict(0) if weights == "fog" else oracle.init_state_dict(0, sharp=True)
    sd = oracle.round_weights_bf16(sd0)
    n = 64
    o, d, gt, _, _ = oracle.synthetic_train_batch(0, 0, n, 100, 100)
    tr = oracle.Trainer(sd)
    loss_ref = tr.forward_backward(o, d, gt)
    g_ref = tr.grads_flat()
    eng = make_engine(oracle, sd, "bf16")
    grad = torch.zeros(1191688, device="cuda")
    loss = eng.forward_backward(o.cuda(), d.cuda(), gt.cuda(), grad)
    torch.cuda.synchronize()
    rep = _grad_report(oracle, grad.cpu(), g_ref)
    for k, err, scale, rel in rep:
        print("%-36s err %.3e  scale %.3e  rel %.3e" % (k, err, scale, rel))
    print("loss", float(loss), float(loss_ref), eng.watchdog())
    assert abs(float(loss) - float(loss_ref)) <= 2e-3 * float(loss_ref)
    bad = []
    for k, err, scale, rel in rep:
        coarse = k.startswith("model.")
        if coarse and any(s in k for s in ("views_linears", "feature_linear", "rgb_linear")):
            if err != 0.0:
                bad.append((k, err, scale))
            continue
        a, b = _per_tensor(oracle, grad.cpu())[k].double(), _per_tensor(oracle, g_ref)[k].double()
        cos = float((a * b).sum() / (a.norm() * b.norm() + 1e-300))
        if coarse and weights == "sharp":
            continue                                     # noise-dominated in the fp32 reference itself (see above)
        # sharp weights (sigma up to ~50, saturating transmittance) amplify the bf16 rounding of the early layers:
        # measured 15% of max / cos 0.98 at fine layer 0; fog: 4% / cos 0.9986
        tol, cmin = (0.15, 0.99) if coarse else ((6e-2, 0.998) if weights == "fog" else (0.25, 0.97))
        if rel > tol or cos < cmin:
            bad.append((k, err, scale, rel, cos))
    assert not bad, bad
    eng.close()


def test_bf16_loss_curve_tracks_oracle(oracle):
    """bf16 tensor-core training (fp32 master weights, fused clip+Adam) vs the fp32 oracle: loss within 1% per step."""
    from nerf_replication_b200 import config as C
    from nerf_replication_b200.network import Network
    from nerf_replication_b200.trainer import FusedTrainer
    cfg = C.default_cfg()
    sd = oracle.init_state_dict(0)
    net = Network(cfg=cfg, precision="bf16")
    net.load_state_dict(sd)
    net = net.cuda()
    ft = FusedTrainer(net, cfg=cfg)
    tr = oracle.Trainer(sd)
    torch.set_num_threads(8)
    worst = 0.0
    for s in range(40):
        o, d, gt, _, _ = oracle.synthetic_train_batch(s, 0, 64, 100, 100)
        l_ref = tr.step(o, d, gt)
        l = float(ft.step(o.cuda(), d.cuda(), gt.cuda()))
        worst = max(worst, abs(l - l_ref) / l_ref)
        assert abs(l - l_ref) <= 1e-2 * l_ref, (s, l, l_ref)
    print("bf16 loss curve: worst relative deviation over 40 steps = %.3e" % worst)


@pytest.mark.parametrize("n_rays,n_importance", [(37, 128), (200, 0), (3, 128)])
def test_bf16_backward_ragged_and_coarse_only(oracle, n_rays, n_importance):
    """Tensor-core training path on row counts that are not multiples of the 128-row tile / 512-row pair unit
    (37 rays -> 2368 + 7104 rows, 3 rays -> 192 + 576 rows) and with N_importance = 0 (the coarse net then takes the
    full backward: rgb head, view branch, PE).  Fog weights, same bounds as the 64-ray test for every tensor whose
    gradient does not pass through the inverse CDF; those (coarse net, hierarchical) are noise-amplified against any
    other arithmetic, so raggedness is checked on ALL tensors by a property instead: the ray-weighted mean of the
    gradients of two sub-batches (different, also ragged, tilings of the same rows) equals the gradient of the batch."""
    sd = oracle.round_weights_bf16(oracle.fog_state_dict(0))
    o, d, gt, _, _ = oracle.synthetic_train_batch(2, 0, n_rays, 100, 100)
    tr = oracle.Trainer(sd, n_importance=n_importance)
    loss_ref = tr.forward_backward(o, d, gt)
    g_ref = tr.grads_flat()
    eng = make_engine(oracle, sd, "bf16", n_importance=n_importance)
    o, d, gt = o.cuda(), d.cuda(), gt.cuda()
    grad = torch.full((1191688,), 7.0, device="cuda")               # zero_grad must clear it
    loss = eng.forward_backward(o, d, gt, grad)
    torch.cuda.synchronize()
    assert abs(float(loss) - float(loss_ref)) <= 2e-3 * float(loss_ref)
    got_t, ref_t = _per_tensor(oracle, grad.cpu()), _per_tensor(oracle, g_ref)
    bad = []
    for k in oracle.state_dict_keys():
        a, b = got_t[k].double(), ref_t[k].double()
        scale = float(b.abs().max())
        if scale == 0.0:
            if float(a.abs().max()) != 0.0:
                bad.append((k, "expected exactly zero", float(a.abs().max())))
            continue
        if k.startswith("model.") and n_importance > 0:
            continue                                                 # through the inverse CDF: property check below
        rel = float((a - b).abs().max()) / scale
        cos = float((a * b).sum() / (a.norm() * b.norm() + 1e-300))
        if rel > 6e-2 or cos < 0.997:           # layer 0 sits at the end of the bf16 chain: measured cos 0.9979 / rel 3.7e-2
            bad.append((k, rel, cos))
    assert not bad, bad
    # grad_scale makes each sub-batch's d rgb equal to the full batch's (2 (rgb - gt) / (3 n)), so that the bf16 gradient
    # tiles round identically and only the fp32 summation order differs
    n_a = n_rays // 2 + 1
    g_a, g_b = torch.zeros_like(grad), torch.zeros_like(grad)
    eng.forward_backward(o[:n_a], d[:n_a], gt[:n_a], g_a, grad_scale=n_a / n_rays)
    eng.forward_backward(o[n_a:], d[n_a:], gt[n_a:], g_b, grad_scale=(n_rays - n_a) / n_rays)
    mix = g_a + g_b
    for k, a in got_t.items():
        b = _per_tensor(oracle, mix.cpu())[k]
        assert maxabs(a, b) <= 1e-3 * float(b.abs().max()) + 1e-12, (k, maxabs(a, b), float(b.abs().max()))
    eng.close()


def test_bf16_full_batch_gradient_is_mean_of_halves(oracle):
    """BASELINE training size (1024 rays x (64 + 128) samples): a size-independent property of the whole
    forward + dgrad + wgrad path -- the loss is a mean over rays, so grad(1024 rays) = (grad(first 512) + grad(last 512)) / 2
    up to fp32 summation order (same rows -> same bf16 operands in both runs)."""
    sd = oracle.round_weights_bf16(oracle.fog_state_dict(0))
    o, d, gt, _, _ = oracle.synthetic_train_batch(5, 0, 1024, 800, 800)
    o, d, gt = o.cuda(), d.cuda(), gt.cuda()
    eng = make_engine(oracle, sd, "bf16")
    u = torch.rand(1024, 128, generator=torch.Generator().manual_seed(3)).cuda()
    g_full, g_a, g_b = (torch.zeros(1191688, device="cuda") for _ in range(3))
    l_full = eng.forward_backward(o, d, gt, g_full, u=u)
    l_a = eng.forward_backward(o[:512], d[:512], gt[:512], g_a, u=u[:512])
    l_b = eng.forward_backward(o[512:], d[512:], gt[512:], g_b, u=u[512:])
    torch.cuda.synchronize()
    assert abs(float(l_full) - 0.5 * (float(l_a) + float(l_b))) <= 1e-6 * float(l_full)
    g_half = 0.5 * (g_a + g_b)
    for k, a in _per_tensor(oracle, g_full.cpu()).items():
        b = _per_tensor(oracle, g_half.cpu())[k]
        scale = float(b.abs().max())
        assert maxabs(a, b) <= 2e-4 * scale + 1e-12, (k, maxabs(a, b), scale)
    eng.close()
