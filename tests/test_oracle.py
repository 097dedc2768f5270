"""Oracle (oracle/nerf_oracle.py) vs the golden vectors the unmodified reference produced
(oracle/make_golden.py, run in the build container).  CPU only; bit-exact."""
import numpy as np
import pytest
import torch

from conftest import load_golden

RENDER_CASES = ["render_seed0_64rays", "render_sharp_64rays", "render_sharp_bf16w_64rays", "render_seed0_bf16w_64rays",
                "render_sharp_coarse_only", "render_sharp_black_bkgd", "render_sharp_128_256"]


def _sd(oracle, case):
    sd = oracle.init_state_dict(0, sharp="sharp" in case)
    return oracle.round_weights_bf16(sd) if "bf16w" in case else sd


def test_weights_digest(oracle):
    g = load_golden("weights_seed0_digest")
    sd = oracle.init_state_dict(0)
    keys = oracle.state_dict_keys()
    assert len(keys) == 48
    np.testing.assert_array_equal(np.array([sd[k].double().sum().item() for k in keys]), g["sums"])
    flat = torch.cat([sd[k].reshape(-1) for k in keys])
    assert flat.numel() == 1191688 == 2 * oracle.PARAMS_PER_NET
    np.testing.assert_array_equal(flat[:64].numpy(), g["head"])
    np.testing.assert_array_equal(flat[-64:].numpy(), g["tail"])
    assert oracle.FLOP_PER_EVAL == 1186816


def test_get_rays(oracle):
    g = load_golden("get_rays_100")
    o, d = oracle.get_rays_np(g["c2w"], 100, 100, g["pix"])
    np.testing.assert_array_equal(o, g["rays_o"])
    np.testing.assert_array_equal(d, g["rays_d"])
    np.testing.assert_allclose(np.linalg.norm(d, axis=-1), 1.0, atol=1e-6)


def test_network_forward(oracle):
    g = load_golden("network_forward")
    sd = oracle.init_state_dict(0)
    pts, vd = torch.from_numpy(g["pts"]), torch.from_numpy(g["viewdirs"])
    torch.set_num_threads(1)
    with torch.no_grad():
        np.testing.assert_array_equal(oracle.freq_encode(pts.reshape(-1, 3), 10).numpy(), g["pe_pts"])
        np.testing.assert_array_equal(oracle.freq_encode(vd, 4).numpy(), g["pe_dir"])
        np.testing.assert_array_equal(oracle.network_forward(sd, pts, vd, None, "fine").numpy(), g["raw_fine"])
        raw_m = oracle.network_forward(sd, pts, vd, torch.from_numpy(g["mask"]), "").numpy()
    np.testing.assert_array_equal(raw_m, g["raw_coarse_masked"])
    assert (raw_m[~g["mask"]] == 0).all()


@pytest.mark.parametrize("case", RENDER_CASES)
def test_render_golden(oracle, case):
    g = load_golden(case)
    torch.set_num_threads(1)
    taps = {}
    with torch.no_grad():
        rgb, depth = oracle.render(_sd(oracle, case), torch.from_numpy(g["rays_o"]), torch.from_numpy(g["rays_d"]),
                                   int(g["n_samples"]), int(g["n_importance"]), bool(g["white_bkgd"]), taps=taps)
    for k in ("t_coarse", "raw_coarse", "rgb_map", "depth_map", "acc_map", "weights"):
        np.testing.assert_array_equal(taps[k].numpy(), g[k], err_msg=k)
    if int(g["n_importance"]) > 0:
        for k in ("t_fine", "raw_fine", "cdf", "inds", "below", "above", "depth_sorted", "sort_idx"):
            np.testing.assert_array_equal(taps[k].numpy(), g[k], err_msg=k)
        # reference quirk: below/above are clamped at Nc-3
        assert g["below"].max() <= int(g["n_samples"]) - 3 and g["above"].max() <= int(g["n_samples"]) - 3


FAST_CASES = {"fast_sharp_64rays": ("sharp", 0.25), "fast_sharp_thr05_64rays": ("sharp", 0.05), "fast_fog_64rays": ("fog", 0.02),
              "fast_seed0_64rays": ("seed0", 0.25), "fast_thin_64rays": ("thin", 2e-4)}


def fast_state_dict(oracle, kind):
    return {"sharp": lambda: oracle.init_state_dict(0, sharp=True), "fog": lambda: oracle.fog_state_dict(0),
            "seed0": lambda: oracle.init_state_dict(0), "thin": lambda: oracle.fog_state_dict(0, bias=-0.015)}[kind]()


@pytest.mark.parametrize("case", sorted(FAST_CASES))
def test_fast_sampling_golden(oracle, case):
    """a-M: the ESS/ERT mask restatement against outputs of the reference run with fast_sampling=True
    (oracle/make_golden.py --fast-only): mask, masked raw and maps bit for bit."""
    g = load_golden(case)
    kind, thr = FAST_CASES[case]
    assert float(g["weights_threshold"]) == pytest.approx(thr)
    torch.set_num_threads(1)
    taps = {}
    with torch.no_grad():
        oracle.render(fast_state_dict(oracle, kind), torch.from_numpy(g["rays_o"]), torch.from_numpy(g["rays_d"]),
                      taps=taps, fast_sampling=True, weights_threshold=thr)
    for k in ("t_fine", "valid_mask", "valid_sorted", "raw_fine", "rgb_map", "depth_map"):
        np.testing.assert_array_equal(taps[k].numpy(), g[k], err_msg=k)
    raw = torch.from_numpy(g["raw_fine"])
    assert (raw[~torch.from_numpy(g["valid_sorted"])] == 0).all()          # masked points -> raw = 0 (network.py:235)
    empty = torch.from_numpy(g["empty_ray"])
    assert not torch.from_numpy(g["valid_mask"])[empty].any()              # empty rays keep no fine sample (:191)


def test_train_golden(oracle):
    g = load_golden("train_sharp_32rays")
    torch.set_num_threads(1)
    tr = oracle.Trainer(oracle.init_state_dict(0, sharp=True))
    n = int(g["n_rays"])
    o, d, gt, _, _ = oracle.synthetic_train_batch(0, 0, n, 100, 100)
    loss = tr.forward_backward(o, d, gt)
    assert abs(float(loss) - g["losses"][0]) <= 1e-6 * g["losses"][0]
    gf = tr.grads_flat()
    keys = oracle.state_dict_keys()
    offs = np.cumsum([0] + [tr.sd[k].numel() for k in keys])
    gabs = np.array([gf[offs[i]:offs[i + 1]].double().abs().sum().item() for i in range(48)])
    np.testing.assert_allclose(gabs, g["grad0_abs"], rtol=1e-4, atol=1e-12)
    # reference autograd graph: the coarse net gets gradient only through sigma
    for i, k in enumerate(keys):
        if k.startswith("model.") and any(s in k for s in ("views_linears", "feature_linear", "rgb_linear")):
            assert g["grad0_abs"][i] == 0.0 and gabs[i] == 0.0, k
        elif k.startswith("model."):
            assert g["grad0_abs"][i] > 0.0, k


def test_composite_properties(oracle):
    """Size-independent properties of the compositing stage used at full size on the GPU as well."""
    g = torch.Generator().manual_seed(3)
    raw = torch.randn(32, 192, 4, generator=g) * 3
    t = torch.sort(torch.rand(32, 192, generator=g) * 4 + 2, -1).values
    rgb, depth, acc, w = oracle.composite(raw, t, True)
    assert (w >= 0).all() and (acc <= 1 + 1e-5).all()
    assert (rgb >= -1e-6).all() and (rgb <= 1 + 1e-5).all()
    rgb_b, _, _, _ = oracle.composite(raw, t, False)
    torch.testing.assert_close(rgb_b + (1 - acc)[:, None], rgb)


def test_fast_sampling_training_golden(oracle):
    """a-M under autograd: the restated MSE step around the oracle render with fast_sampling=True against the reference's
    own loss and per-tensor gradient sums (oracle/make_golden.py --fast-train-only)."""
    g = load_golden("train_fast_fog_32rays")
    torch.set_num_threads(1)
    o, d, gt, _, _ = oracle.synthetic_train_batch(0, 0, int(g["n_rays"]), 100, 100)
    tr = oracle.Trainer(oracle.fog_state_dict(0), fast_sampling=True, weights_threshold=float(g["weights_threshold"]))
    loss = float(tr.forward_backward(o, d, gt))
    assert abs(loss - float(g["loss"])) <= 1e-6 * loss
    flat, off = tr.grads_flat(), 0
    for i, k in enumerate(oracle.state_dict_keys()):
        n = tr.sd[k].numel()
        seg = flat[off:off + n]
        off += n
        assert abs(float(seg.double().abs().sum()) - g["grad_abs"][i]) <= 1e-4 * max(g["grad_abs"][i], 1e-12), k
        assert np.allclose(seg[:8].numpy(), g["grad_head"][i][:min(8, n)], rtol=1e-3, atol=1e-9 + 1e-4 * float(seg.abs().max())), k
