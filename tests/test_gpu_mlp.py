"""GPU parity of the NeRF MLP (Network.forward, network.py:199-258) through the C ABI:
fp32 validation kernels vs the reference goldens (1e-5), the fused tcgen05 bf16 kernel vs the oracle
run with bf16-representable weights, and the tcgen05 descriptor probes."""
import pytest
import torch

from conftest import load_golden
from gpu_util import dev, make_engine, maxabs, quantiles, state_dict_for

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("variant", [0, 1, 2])
@pytest.mark.parametrize("N,K", [(256, 64), (256, 256), (128, 128), (64, 64)])
def test_umma_probe(oracle, variant, N, K):
    """One 128 x N x K bf16 GEMM per operand layout the engine uses (smem SW128 / interleaved / A in TMEM)."""
    eng = make_engine(oracle, oracle.init_state_dict(0), "bf16")
    g = torch.Generator().manual_seed(100 * variant + N + K)
    a = torch.randn(128, K, generator=g).bfloat16().float()
    b = torch.randn(N, K, generator=g).bfloat16().float()
    d = eng.umma_probe(variant, a.cuda(), b.cuda())
    torch.cuda.synchronize()
    ref = a.double() @ b.double().T
    assert maxabs(d, ref) <= 1e-3 * K ** 0.5, eng.watchdog()
    eng.close()


@pytest.mark.parametrize("N,K", [(256, 64), (256, 256), (128, 128), (64, 64)])
def test_umma_probe_cta_pair(oracle, N, K):
    """cta_group::2: one 256 x N x K GEMM over a cluster of two CTAs (layout of the fused kernel mlp_tc2.cu)."""
    eng = make_engine(oracle, oracle.init_state_dict(0), "bf16")
    g = torch.Generator().manual_seed(300 + N + K)
    a = torch.randn(256, K, generator=g).bfloat16().float()
    b = torch.randn(N, K, generator=g).bfloat16().float()
    d = eng.umma_probe(3, a.cuda(), b.cuda())
    torch.cuda.synchronize()
    ref = a.double() @ b.double().T
    assert maxabs(d, ref) <= 1e-3 * K ** 0.5, eng.watchdog()
    eng.close()


@pytest.mark.parametrize("N,K", [(256, 128), (64, 128), (128, 64), (192, 16)])
def test_umma_probe_mn_major(oracle, N, K):
    """Both operands MN-major (wgrad form): D[128 x N] = At^T Bt, reduction over K rows of saved tile images."""
    eng = make_engine(oracle, oracle.init_state_dict(0), "bf16")
    g = torch.Generator().manual_seed(400 + N + K)
    at = torch.randn(K, 128, generator=g).bfloat16().float()
    bt = torch.randn(K, N, generator=g).bfloat16().float()
    d = eng.umma_probe(4, at.cuda(), bt.cuda())
    torch.cuda.synchronize()
    ref = at.double().T @ bt.double()
    assert maxabs(d, ref) <= 1e-3 * K ** 0.5, eng.watchdog()
    eng.close()


@pytest.mark.gpu
@pytest.mark.parametrize("N,K", [(256, 64), (64, 16), (128, 32)])
def test_umma_probe_mn_major_core_matrix(oracle, N, K):
    """MN-major operands in the un-swizzled core-matrix layout of the training arenas (LBO = 128: next 8 rows,
    SBO = 1024: next 8 columns) -- the form the wgrad kernel consumes."""
    eng = make_engine(oracle, oracle.init_state_dict(0), "bf16")
    g = torch.Generator().manual_seed(500 + N + K)
    at = torch.randn(K, 128, generator=g).bfloat16().float()
    bt = torch.randn(K, N, generator=g).bfloat16().float()
    d = eng.umma_probe(5, at.cuda(), bt.cuda())
    torch.cuda.synchronize()
    ref = at.double().T @ bt.double()
    assert maxabs(d, ref) <= 1e-3 * K ** 0.5, eng.watchdog()
    eng.close()


@pytest.mark.parametrize("precision,tol", [("fp32", 1e-5)])
def test_mlp_forward_fp32_golden(oracle, precision, tol):
    g = load_golden("network_forward")
    eng = make_engine(oracle, oracle.init_state_dict(0), precision)
    pts, vd = dev(g["pts"]), dev(g["viewdirs"])
    raw = eng.mlp_forward(1, pts, vd)
    assert raw.shape == (5, 64, 4)
    assert maxabs(raw, g["raw_fine"]) <= tol
    raw_m = eng.mlp_forward(0, pts, vd, mask=dev(g["mask"]))
    assert maxabs(raw_m, g["raw_coarse_masked"]) <= tol
    assert float(raw_m[~dev(g["mask"])].abs().max()) == 0.0          # masked points -> raw = 0 (network.py:235)
    eng.close()


def _bf16_case(oracle, sharp, n_rows_r, S, seed):
    sd = oracle.round_weights_bf16(oracle.init_state_dict(0, sharp=sharp))
    g = torch.Generator().manual_seed(seed)
    pts = (torch.rand(n_rows_r, S, 3, generator=g) - 0.5) * 6.0
    vd = torch.nn.functional.normalize(torch.randn(n_rows_r, 3, generator=g), dim=-1)
    return sd, pts, vd


@pytest.mark.parametrize("sharp", [False, True])
@pytest.mark.parametrize("R,S", [(5, 64), (3, 192), (1, 1), (7, 37), (40, 64)])
def test_mlp_forward_tcgen05_vs_oracle(oracle, sharp, R, S):
    """bf16 operands / fp32 accumulate.  Oracle = fp32 torch with the same bf16-representable weights, so only
    activation rounding (2^-9 relative per layer) is judged: raw within 2e-2 of the pre-activation scale."""
    sd, pts, vd = _bf16_case(oracle, sharp, R, S, 10 * R + S)
    eng = make_engine(oracle, sd, "bf16")
    with torch.no_grad():
        ref_c = oracle.network_forward(sd, pts, vd, None, "")
        ref_f = oracle.network_forward(sd, pts, vd, None, "fine")
    raw_c = eng.mlp_forward(0, pts.cuda(), vd.cuda())
    raw_f = eng.mlp_forward(1, pts.cuda(), vd.cuda())
    torch.cuda.synchronize()
    for raw, ref in ((raw_c, ref_c), (raw_f, ref_f)):
        scale_rgb = float(ref[..., :3].abs().max()) + 1e-6
        scale_sig = float(ref[..., 3].abs().max()) + 1e-6
        assert maxabs(raw[..., :3], ref[..., :3]) <= 2e-2 * max(scale_rgb, 1.0), (quantiles(raw, ref), eng.watchdog())
        assert maxabs(raw[..., 3], ref[..., 3]) <= 2e-2 * max(scale_sig, 1.0), (quantiles(raw, ref), eng.watchdog())
    eng.close()


def test_mlp_biases_enter_the_gemm_at_fp32_accuracy(oracle):
    """The bf16 kernel adds no bias in its epilogue: every layer starts with one K = 16 MMA on a bias slab whose rows hold
    the bias split into three bf16 terms (csrc/mlp_tc.cu pack_bias_slabs_kernel).  With all trunk weights zero the
    activations are the (ReLU'd) biases themselves, and the two CUDA-core heads, which read the fp32 accumulator values,
    return plain sums of them: sigma = sum relu(b_7) + b_alpha, rgb = sum relu(b_views) + b_rgb with unit head weights.
    A bias that reached the accumulator as a single bf16 (2^-9 relative) would miss the float64 sums by ~1e-3 relative."""
    g = torch.Generator().manual_seed(11)
    sd = {k: torch.zeros_like(v) for k, v in oracle.init_state_dict(0).items()}
    for k in sd:
        if k.endswith("bias"):
            sd[k] = torch.randn(sd[k].shape, generator=g) * 1.7 + 0.3          # nothing here is bf16-representable
        if "alpha_linear.weight" in k or "rgb_linear.weight" in k:
            sd[k] = torch.ones_like(sd[k])
    eng = make_engine(oracle, sd, "bf16")
    pts = (torch.rand(9, 37, 3, generator=g) - 0.5) * 6.0
    vd = torch.nn.functional.normalize(torch.randn(9, 3, generator=g), dim=-1)
    for which, prefix in ((0, "model."), (1, "model_fine.")):
        raw = eng.mlp_forward(which, pts.cuda(), vd.cuda()).cpu().double()
        sig = float(sd[prefix + "pts_linears.7.bias"].double().clamp(min=0).sum() + sd[prefix + "alpha_linear.bias"].double())
        hv = float(sd[prefix + "views_linears.0.bias"].double().clamp(min=0).sum())
        rgb = torch.tensor([hv + float(b) for b in sd[prefix + "rgb_linear.bias"].double()], dtype=torch.float64)
        assert float((raw[..., 3] - sig).abs().max()) <= 2e-6 * abs(sig), (float(raw[0, 0, 3]), sig)
        assert float((raw[..., :3] - rgb).abs().max()) <= 2e-6 * float(rgb.abs().max()), (raw[0, 0, :3], rgb)
    eng.close()


def test_mlp_forward_tcgen05_masked_and_multi_tile(oracle):
    """More rows than one wave of 256-row super-tiles per SM (persistent loop), ragged tail, mask."""
    sd, pts, vd = _bf16_case(oracle, False, 1203, 64, 5)      # 76 992 rows = 300.75 super-tiles
    eng = make_engine(oracle, sd, "bf16")
    mask = torch.rand(1203, 64, generator=torch.Generator().manual_seed(1)) > 0.3
    raw = eng.mlp_forward(1, pts.cuda(), vd.cuda(), mask=mask.cuda())
    torch.cuda.synchronize()
    assert float(raw[~mask.cuda()].abs().max()) == 0.0
    sel = torch.arange(0, 1203, 97)
    with torch.no_grad():
        ref = oracle.network_forward(sd, pts[sel], vd[sel], mask[sel], "fine")
    assert maxabs(raw[sel.cuda()], ref) <= 2e-2, eng.watchdog()
    # the fp32 validation kernels and the tcgen05 kernel agree on the same weights
    eng32 = make_engine(oracle, sd, "fp32")
    raw32 = eng32.mlp_forward(1, pts.cuda(), vd.cuda(), mask=mask.cuda())
    assert maxabs(raw, raw32) <= 2e-2
    eng.close(); eng32.close()


def test_mlp_rejects_bad_arguments(oracle):
    from nerf_replication_b200.engine import Engine, EngineError
    eng = Engine(precision="bf16")
    with pytest.raises(EngineError):
        eng.mlp_forward(0, torch.zeros(1, 1, 3).cuda(), torch.zeros(1, 3).cuda())      # before set_params
    with pytest.raises(EngineError):
        eng.set_params(torch.zeros(10).cuda())
    with pytest.raises(EngineError):
        eng.configure(n_samples=2)
    eng.close()
