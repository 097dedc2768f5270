"""GPU parity of the steps either side of the hot path, through the C ABI:
ray bank (Dataset.__getitem__, blender.py:84-170) vs the reference's own items (tests/golden/loop_dataset.npz), evaluator
sums vs Evaluator.psnr_metric goldens, FusedAdam vs torch.optim.Adam incl. checkpoint exchange, and the step driver vs the
oracle training loop."""
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from conftest import load_golden
from gpu_util import dev, make_engine, maxabs
from oracle import loop_oracle as LO

pytestmark = pytest.mark.gpu


def _bank(eng, images, poses, cax, uint8=False):
    V, H, W, _ = images.shape
    focal = W / (2 * np.tan(cax / 2))
    b = eng.bank_create(V, H, W, focal, uint8=uint8)
    for k in range(V):
        eng.bank_set_view(b, k, np.ascontiguousarray(images[k].astype(np.uint8) if uint8 else images[k].astype(np.float32)), poses[k])
    return b


def _ulp_close(a, b):
    """rays_d goes through float64 dot products that numpy (BLAS) and the kernel (FMA) round differently in the last
    float64 bit; after the final rounding to fp32 that can flip one fp32 ulp on a rare element, never more."""
    a, b = a.cpu().numpy(), np.asarray(b)
    diff = np.abs(a.view(np.int32).astype(np.int64) - b.view(np.int32).astype(np.int64))
    return int(diff.max()), float((diff > 0).mean())


def test_bank_reproduces_reference_dataset_items(oracle):
    g = load_golden("loop_dataset")
    cax = float(g["camera_angle_x"])
    eng = make_engine(oracle, oracle.init_state_dict(0), "fp32")
    # val branch: every pixel, fractional (resized) pixel values kept as fp32
    b = _bank(eng, g["val_images"], g["val_poses"], cax)
    for k in (0, 2):
        it = eng.bank_sample(b, k, 24 * 20, mode="range", want_ids=True)
        assert np.array_equal(it["colors"].cpu().numpy(), g["val%d_colors" % k])
        assert np.array_equal(it["rays_o"].cpu().numpy(), g["val%d_rays_o" % k])
        ulp, frac = _ulp_close(it["rays_d"], g["val%d_rays_d" % k])
        assert ulp <= 1 and frac < 0.01, (ulp, frac)
        assert np.array_equal(it["ids"].cpu().numpy(), np.arange(480))
    # train branch with the reference's own np.random.choice ids; uint8 storage gives the same colours as fp32 storage
    bf, bu = _bank(eng, g["train_images"], g["train_poses"], cax), _bank(eng, g["train_images"], g["train_poses"], cax, uint8=True)
    for k in (0, 1):
        ids = dev(g["train%d_ids" % k])
        for bank in (bf, bu):
            it = eng.bank_sample(bank, k, 1024, mode="ids", pixel_ids=ids)
            assert np.array_equal(it["colors"].cpu().numpy(), g["train%d_colors" % k])
            assert np.array_equal(it["rays_o"].cpu().numpy(), g["train%d_rays_o" % k])
            ulp, frac = _ulp_close(it["rays_d"], g["train%d_rays_d" % k])
            assert ulp <= 1 and frac < 0.01, (ulp, frac)
    # lego size: 800x800 view, the reference's 1024 pixels
    img = np.floor(np.random.RandomState(int(g["lego_image_seed"])).uniform(0, 256, size=(1, 800, 800, 3))).astype(np.float32)
    bl = _bank(eng, img, g["lego_pose"][None], cax, uint8=True)
    it = eng.bank_sample(bl, 0, 1024, mode="ids", pixel_ids=dev(g["lego_ids"]))
    assert np.array_equal(it["colors"].cpu().numpy(), g["lego_colors"]) and np.array_equal(it["rays_o"].cpu().numpy(), g["lego_rays_o"])
    ulp, frac = _ulp_close(it["rays_d"], g["lego_rays_d"])
    assert ulp <= 1 and frac < 0.01, (ulp, frac)
    # the bank's rays are the path's own get_rays kernel (a-B), bit for bit
    o2, d2 = eng.get_rays(g["lego_pose"], 800, 800, 800 / (2 * np.tan(cax / 2)), pixel_ids=dev(g["lego_ids"]))
    assert torch.equal(o2, it["rays_o"]) and torch.equal(d2, it["rays_d"])
    eng.close()


def test_bank_device_pixel_draw(oracle):
    """mode "random": the keyed bijection on the device == its integer restatement; distinct pixels; items consistent with
    an explicit-id gather of the same pixels."""
    images, poses, cax = LO.synthetic_scene(2, 800, 800, seed=5)
    eng = make_engine(oracle, oracle.init_state_dict(0), "fp32", seed=1234)
    b = _bank(eng, images, poses, cax, uint8=True)
    seen = set()
    for draw in (0, 1, 77):
        it = eng.bank_sample(b, 1, 1024, mode="random", draw=draw, want_ids=True)
        ids = it["ids"].cpu().numpy()
        assert np.array_equal(ids, LO.bank_permutation(1234, draw, 640000, 1024))
        assert np.array_equal(ids, eng.bank_permutation_host(1234, draw, 640000, 1024))
        assert len(set(ids.tolist())) == 1024 and ids.min() >= 0 and ids.max() < 640000
        seen |= set(ids.tolist())
        ref = LO.dataset_getitem(images[1], poses[1], cax, 800, 800, ids)
        assert np.array_equal(it["colors"].cpu().numpy(), ref["colors"].numpy())
        again = eng.bank_sample(b, 1, 1024, mode="ids", pixel_ids=it["ids"])
        for k in ("rays_o", "rays_d", "colors"):
            assert torch.equal(again[k], it[k])
    assert len(seen) > 3 * 1024 - 32
    # a whole small view drawn at once is a permutation of its pixels
    small = _bank(eng, images[:, :30, :50], poses, cax, uint8=True)
    ids = eng.bank_sample(small, 0, 1500, mode="random", draw=3, want_ids=True)["ids"].cpu().numpy()
    assert sorted(ids.tolist()) == list(range(1500))
    eng.close()


def test_bank_errors(oracle):
    from nerf_replication_b200.engine import EngineError
    images, poses, cax = LO.synthetic_scene(2, 8, 8)
    eng = make_engine(oracle, oracle.init_state_dict(0), "fp32")
    b = eng.bank_create(2, 8, 8, 10.0)
    eng.bank_set_view(b, 0, images[0], poses[0])
    with pytest.raises(EngineError, match="never uploaded"):
        eng.bank_sample(b, 1, 4, mode="range")
    with pytest.raises(EngineError, match="outside the image"):
        eng.bank_sample(b, 0, 65, mode="range")
    with pytest.raises(EngineError, match="without replacement"):
        eng.bank_sample(b, 0, 65, mode="random")
    with pytest.raises(EngineError, match="no such bank"):
        eng.bank_sample(b + 5, 0, 4, mode="range")
    assert eng.bank_sample(b, 0, 64, mode="range")["colors"].shape == (64, 3)
    eng.bank_destroy(b)
    with pytest.raises(EngineError, match="no such bank"):
        eng.bank_sample(b, 0, 4, mode="range")
    eng.close()


def test_evaluator_sums_match_reference_goldens(oracle):
    g = load_golden("loop_evaluator")
    from nerf_replication_b200.evaluator import Evaluator
    eng = make_engine(oracle, oracle.init_state_dict(0), "fp32")
    ev = Evaluator(engine=eng)
    for name in ("noise", "far", "equal", "edges"):
        pred, gt = dev(g[name + "_pred"]), dev(g[name + "_gt"])
        sse, sq8 = eng.eval_image(pred, gt)
        assert int(sq8) == int(g[name + "_sq8"]), name                                   # integer work: bit-exact
        assert ev.psnr_metric(pred, gt) == float(g[name + "_psnr"]), name                # same float64 formula on the same integer
        mse = float(sse) / pred.numel()
        assert abs(mse - float(g[name + "_mse"])) <= 1e-6 * max(float(g[name + "_mse"]), 1e-12), name   # fp32 mean vs float64 sum
    # full image size, against numpy on the same data (size-independent check of the integer sum)
    rng = np.random.RandomState(5)
    gt = rng.uniform(0, 1, size=(640000, 3)).astype(np.float32)
    pred = np.clip(gt + rng.normal(0, 0.2, size=gt.shape), -0.5, 1.5).astype(np.float32)
    sse, sq8 = eng.eval_image(dev(pred), dev(gt))
    assert int(sq8) == LO.uint8_square_sum(pred, gt)
    m_ref, p_ref = LO.evaluate_metrics(pred, gt)
    assert abs(float(sse) / pred.size - float(m_ref)) <= 1e-6 * float(m_ref)
    assert ev.psnr_metric(dev(pred), dev(gt)) == p_ref
    a, b2 = eng.eval_image(dev(pred), dev(gt))
    assert float(a) == float(sse) and int(b2) == int(sq8)                                 # deterministic reduction
    eng.close()


def test_ssim_kernel_matches_skimage_restatement(oracle):
    """SSIM parity is unpinned (scikit-image is not installed anywhere here): the kernel is compared with the restatement
    of skimage's published algorithm on scipy.ndimage.uniform_filter (oracle/loop_oracle.py)."""
    eng = make_engine(oracle, oracle.init_state_dict(0), "fp32")
    rng = np.random.RandomState(9)
    for H, W in ((24, 31), (100, 100), (7, 9)):
        a = rng.uniform(0, 1, size=(H, W, 3)).astype(np.float32)
        b = np.clip(a + rng.normal(0, 0.08, size=a.shape), 0, 1).astype(np.float32)
        got = float(eng.eval_ssim(dev(b), dev(a), H, W, 7))
        assert abs(got - LO.ssim_uint8(b, a)) <= 1e-9, (H, W)
        assert abs(float(eng.eval_ssim(dev(a), dev(a), H, W, 7)) - 1.0) <= 1e-12
    from nerf_replication_b200.engine import EngineError
    with pytest.raises(EngineError, match="win_size"):
        eng.eval_ssim(dev(a), dev(a), 7, 9, 8)
    eng.close()


def _net(oracle, sd, precision, cfg=None):
    from nerf_replication_b200 import config as C
    from nerf_replication_b200.network import Network
    cfg = C.default_cfg() if cfg is None else cfg
    net = Network(cfg=cfg, precision=precision)
    net.load_state_dict(sd)
    return net.cuda(), cfg


def test_bank_train_step_is_sample_plus_forward_backward(oracle):
    images, poses, cax = LO.synthetic_scene(3, 100, 100, seed=6)
    for precision in ("fp32", "bf16"):
        eng = make_engine(oracle, oracle.init_state_dict(0, sharp=True), precision, seed=3)
        b = _bank(eng, images, poses, cax, uint8=True)
        g1, g2 = torch.zeros(1191688, device="cuda"), torch.zeros(1191688, device="cuda")
        l1 = eng.bank_train_step(b, 2, 64, g1, mode="random", draw=5)
        it = eng.bank_sample(b, 2, 64, mode="random", draw=5)
        l2 = eng.forward_backward(it["rays_o"], it["rays_d"], it["colors"], g2)
        assert float(l1) == float(l2)
        assert maxabs(g1, g2) <= 1e-6 * float(g2.abs().max())                    # weight gradients fold with atomics (order varies)
        # and against the oracle on the reference-style item
        ids = LO.bank_permutation(3, 5, 10000, 64)
        ref_it = LO.dataset_getitem(images[2], poses[2], cax, 100, 100, ids)
        tr = oracle.Trainer(oracle.init_state_dict(0, sharp=True))
        loss_ref = float(tr.forward_backward(ref_it["rays_o"], ref_it["rays_d"], ref_it["colors"]))
        assert abs(float(l1) - loss_ref) <= (1e-6 if precision == "fp32" else 1e-2) * loss_ref
        eng.close()


def test_fused_adam_is_torch_adam_and_checkpoints_interchange(oracle, tmp_path):
    from nerf_replication_b200 import checkpoint as ck
    from nerf_replication_b200.optimizer import ExponentialLR, FusedAdam
    from nerf_replication_b200.trainer import FusedTrainer, Recorder
    sd = oracle.init_state_dict(0, sharp=True)
    net, cfg = _net(oracle, sd, "fp32")
    opt = FusedAdam(net, lr=5e-4, eps=1e-8, clip=40.0)
    sch = ExponentialLR(opt, decay_epochs=2, gamma=0.1)
    ref = oracle.Trainer(sd)
    net_b, _ = _net(oracle, sd, "fp32")
    legacy = FusedTrainer(net_b, cfg=cfg)
    for s in range(3):
        o, d, gt, _, _ = oracle.synthetic_train_batch(s, 0, 32, 100, 100)
        loss = net.engine().forward_backward(o.cuda(), d.cuda(), gt.cuda(), net.flat_grad)
        net_b.flat_grad.copy_(net.flat_grad)        # same gradient into both updates (weight gradients fold with atomics,
        opt.step()                                  # so two backward runs differ in the last bits and Adam's first steps
        legacy.apply()                              # are ~ lr * sign(g): compare the update itself)
        loss_ref = ref.step(o, d, gt)
        assert abs(float(loss) - loss_ref) <= 1e-4 * loss_ref                      # multi-step: fp32 round-off compounds through Adam
    assert torch.equal(net.flat_params, net_b.flat_params) and torch.equal(opt.m, legacy.m) and torch.equal(opt.v, legacy.v)
    sch.step()
    assert opt.param_groups[0]["lr"] == LO.exponential_lr(5e-4, 1, 0.1, 2) and opt._hyper()[0] == opt.param_groups[0]["lr"]
    # state dict has torch.optim.Adam's layout: a real torch Adam over the oracle's parameters loads it and continues identically
    st = opt.state_dict()
    assert len(st["param_groups"]) == 48 and sorted(st["state"][0].keys()) == ["exp_avg", "exp_avg_sq", "step"] and float(st["state"][5]["step"]) == 3.0
    off = 0
    for i, k in enumerate(oracle.state_dict_keys()):          # state index i <-> parameter i of the flat layout
        n_i = sd[k].numel()
        assert st["state"][i]["exp_avg"].shape == sd[k].shape
        assert torch.equal(st["state"][i]["exp_avg"].reshape(-1), opt.m[off:off + n_i]) and torch.equal(st["state"][i]["exp_avg_sq"].reshape(-1), opt.v[off:off + n_i])
        off += n_i
    keys = oracle.state_dict_keys()
    tparams = [net.state_dict()[k].detach().cpu().clone().requires_grad_(True) for k in keys]
    topt = torch.optim.Adam([{"params": [p], "lr": 5e-4, "weight_decay": 0.0, "eps": 1e-8} for p in tparams], 5e-4, eps=1e-8)
    topt.load_state_dict({"state": {i: {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in s.items()} for i, s in st["state"].items()},
                          "param_groups": st["param_groups"]})
    o, d, gt, _, _ = oracle.synthetic_train_batch(9, 0, 32, 100, 100)
    net.engine().forward_backward(o.cuda(), d.cuda(), gt.cuda(), net.flat_grad)
    gflat = net.flat_grad.detach().cpu().clone()
    off = 0
    for p in tparams:
        p.grad = gflat[off:off + p.numel()].view(p.shape).clamp(-40, 40).clone()
        off += p.numel()
    topt.step()
    opt.step()
    after = torch.cat([p.detach().reshape(-1) for p in tparams])
    assert maxabs(net.flat_params, after) <= 2.4e-7                                # fp32 round-off only (1-2 ulp of the largest weights)
    assert float(topt.state_dict()["state"][0]["step"]) == 4.0 and opt.step_count == 4
    # save with the engine, resume into a fresh engine: identical state, same continuation
    rec = Recorder()
    rec.step = 4
    d_ck = str(tmp_path / "trained_model")
    ck.save_model(net, opt, sch, rec, d_ck, 0, last=True)
    net2, _ = _net(oracle, oracle.init_state_dict(1), "fp32")
    opt2 = FusedAdam(net2, lr=1.0, clip=40.0)
    sch2 = ExponentialLR(opt2, decay_epochs=2, gamma=0.1)
    rec2 = Recorder()
    assert ck.load_model(net2, opt2, sch2, rec2, d_ck) == 1 and rec2.step == 4 and opt2.step_count == 4
    assert torch.equal(net2.flat_params, net.flat_params) and torch.equal(opt2.m, opt.m) and torch.equal(opt2.v, opt.v)
    assert opt2.param_groups[0]["lr"] == opt.param_groups[0]["lr"]
    o, d, gt, _, _ = oracle.synthetic_train_batch(10, 0, 32, 100, 100)
    net.engine().forward_backward(o.cuda(), d.cuda(), gt.cuda(), net.flat_grad)
    net2.flat_grad.copy_(net.flat_grad)
    opt.step(); opt2.step()
    assert torch.equal(net2.flat_params, net.flat_params) and torch.equal(opt2.m, opt.m)
    # a checkpoint written by a stock torch Adam (what the reference's save_model stores) resumes on the engine
    torch.save({"net": {k: p.detach() for k, p in zip(keys, tparams)}, "optim": topt.state_dict(), "scheduler": sch.state_dict(),
                "recorder": {"step": 4}, "epoch": 6}, os.path.join(d_ck, "latest.pth"))
    net3, _ = _net(oracle, oracle.init_state_dict(2), "fp32")
    opt3 = FusedAdam(net3, lr=1.0)
    assert ck.load_model(net3, opt3, ExponentialLR(opt3, decay_epochs=2, gamma=0.1), Recorder(), d_ck) == 7
    assert maxabs(net3.flat_params, after) == 0.0 and opt3.step_count == 4
    assert maxabs(opt3.m, torch.cat([topt.state[p]["exp_avg"].reshape(-1) for p in tparams])) == 0.0


def _loop_cfg(tmp_path, ep_iter=4, epochs=3):
    from nerf_replication_b200 import config as C
    cfg = C.default_cfg()
    cfg.train = SimpleNamespace(batch_size=1, shuffle=True, lr=5e-4, weight_decay=0.0, eps=1e-8, optim="adam", epoch=epochs,
                                scheduler=SimpleNamespace(type="exponential", gamma=0.1, decay_epochs=2))
    cfg.test = SimpleNamespace(batch_size=1)
    cfg.ep_iter, cfg.save_ep, cfg.save_latest_ep, cfg.eval_ep, cfg.log_interval = ep_iter, 100, 1, 100, 2
    cfg.trained_model_dir = str(tmp_path / "trained_model")
    return cfg


def test_step_driver_follows_the_oracle_loop(oracle, tmp_path):
    """fit(): epochs x ep_iter fused steps on bank-drawn pixels with the per-epoch ExponentialLR, against the oracle
    trainer fed the same views / pixels / lr; then an interrupted + resumed run against the uninterrupted one."""
    from nerf_replication_b200.dataset import Dataset
    from nerf_replication_b200.trainer import fit
    from nerf_replication_b200.view_schedule import ViewSchedule
    images, poses, cax = LO.synthetic_scene(5, 100, 100, seed=8)
    sd = oracle.init_state_dict(0, sharp=True)
    cfg = _loop_cfg(tmp_path)

    def run(epochs, resume, seed_torch=0):
        net, _ = _net(oracle, sd, "fp32", cfg)
        ds = Dataset(images=images, transform_matrices=poses, camera_angle_x=cax, split="train", seed=21, n_rays=48)
        torch.manual_seed(seed_torch)
        hist, opt, sch, rec = fit(cfg, net, ds, resume=resume, epochs=epochs, verbose=False)
        return net, torch.stack(hist).cpu() if hist else None, opt, rec

    torch.manual_seed(0)
    views = []
    for _ in range(3):                                      # what RandomSampler will hand out, replayed from the same global seed
        views.append([b[0] for b in ViewSchedule(5, 1, True, False, 4)])
    net, losses, opt, rec = run(None, resume=False)
    torch.manual_seed(0)
    assert losses.shape == (3, 4) and rec.step == 12 and opt.step_count == 12
    ref = oracle.Trainer(sd)
    draw = 0
    for e in range(3):
        ref.set_lr(LO.exponential_lr(5e-4, e, 0.1, 2))
        for it in range(4):
            ids = LO.bank_permutation(21, draw, 10000, 48)
            item = LO.dataset_getitem(images[views[e][it]], poses[views[e][it]], cax, 100, 100, ids)
            loss_ref = ref.step(item["rays_o"], item["rays_d"], item["colors"])
            assert abs(float(losses[e, it]) - loss_ref) <= 2e-4 * loss_ref, (e, it, float(losses[e, it]), loss_ref)
            draw += 1
    assert opt.param_groups[0]["lr"] == LO.exponential_lr(5e-4, 3, 0.1, 2)
    assert os.path.exists(os.path.join(cfg.trained_model_dir, "latest.pth"))
    # interrupted after 2 epochs, resumed for a third: the checkpoint restores parameters, moments, step, lr schedule and
    # the pixel-draw counter (the view sampler's RNG is not checkpointed, as in the reference: the resumed epoch is replayed
    # from its own seed)
    net_a, l_a, opt_a, _ = run(2, resume=False)
    assert maxabs(l_a, losses[:2]) <= 1e-4 * float(losses.max())     # run to run: atomics order in the weight-gradient fold
    torch.manual_seed(5)
    v_resumed = [b[0] for b in ViewSchedule(5, 1, True, False, 4)]
    net_b, l_b, opt_b, rec_b = run(1, resume=True, seed_torch=5)
    assert rec_b.step == 12 and opt_b.step_count == 12 and l_b.shape == (1, 4)
    assert opt_b.param_groups[0]["lr"] == LO.exponential_lr(5e-4, 3, 0.1, 2)
    ref2 = oracle.Trainer(sd)
    draw = 0
    for e in range(3):
        ref2.set_lr(LO.exponential_lr(5e-4, e, 0.1, 2))
        for it in range(4):
            v = views[e][it] if e < 2 else v_resumed[it]
            item = LO.dataset_getitem(images[v], poses[v], cax, 100, 100, LO.bank_permutation(21, draw, 10000, 48))
            loss_ref = ref2.step(item["rays_o"], item["rays_d"], item["colors"])
            if e == 2:
                assert abs(float(l_b[0, it]) - loss_ref) <= 2e-4 * loss_ref, (it, float(l_b[0, it]), loss_ref)
            draw += 1


def test_validation_pass_scores_rendered_views(oracle, tmp_path):
    """Trainer.val: render every val view through the engine, score it on the device -> the oracle's metrics of the
    oracle-rendered image within the render tolerance."""
    from nerf_replication_b200.dataset import Dataset
    from nerf_replication_b200.evaluator import Evaluator
    from nerf_replication_b200.trainer import Trainer, make_data_loader
    images, poses, cax = LO.synthetic_scene(2, 20, 24, seed=10, integral=False)
    sd = oracle.init_state_dict(0, sharp=True)
    net, cfg = _net(oracle, sd, "fp32")
    val = Dataset(net=net, images=images, transform_matrices=poses, camera_angle_x=cax, split="val")
    ev = Evaluator(net=net, result_dir=str(tmp_path / "result"))
    stats = Trainer(net, cfg, verbose=False).val(0, make_data_loader(cfg, val, is_train=False), ev)
    mses, psnrs, ssims = [], [], []
    for k in range(2):
        item = LO.dataset_getitem(images[k], poses[k], cax, 20, 24)
        with torch.no_grad():
            rgb, _ = oracle.render(sd, item["rays_o"], item["rays_d"])
        m, p = LO.evaluate_metrics(rgb.numpy(), item["colors"].numpy())
        mses.append(float(m)); psnrs.append(float(p))
        ssims.append(LO.ssim_uint8(np.clip(rgb.numpy(), 0, 1).reshape(20, 24, 3), item["colors"].numpy().reshape(20, 24, 3)))
    assert abs(stats["mse"] - np.mean(mses)) <= 1e-4 * np.mean(mses)
    assert abs(stats["psnr"] - np.mean(psnrs)) <= 0.05 and abs(stats["ssim"] - np.mean(ssims)) <= 5e-3
    assert abs(stats["loss"] - np.mean(mses)) <= 1e-4 * np.mean(mses)          # images are in [0,1]: clipping is a no-op for the loss
    assert sorted(os.listdir(tmp_path / "result" / "images")) == ["view000_gt.png", "view000_pred.png", "view001_gt.png", "view001_pred.png"]
