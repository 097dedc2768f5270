"""SURVEY.md 8(f) rows 1-3 without a GPU: the loop oracle (oracle/loop_oracle.py) against the golden vectors produced by
the reference's own Dataset.__getitem__ / samplers / ExponentialLR / Evaluator.psnr_metric / save_model
(oracle/make_golden_loop.py), and the product's host logic (view schedule, lr schedule, checkpoint layout, the keyed
pixel bijection's host copy) against both."""
import json
from types import SimpleNamespace
import os

import numpy as np
import pytest
import torch

from conftest import GOLD, load_golden
from oracle import loop_oracle as LO
from oracle import nerf_oracle as O


def _json(name):
    with open(os.path.join(GOLD, name)) as f:
        return json.load(f)


# ---- oracle vs the reference's outputs -------------------------------------------------------------------------
def test_dataset_getitem_matches_reference_goldens():
    g = load_golden("loop_dataset")
    cax = float(g["camera_angle_x"])
    for k in (0, 2):                                                   # val branch: every pixel, fractional pixel values
        item = LO.dataset_getitem(g["val_images"][k], g["val_poses"][k], cax, 24, 20)
        for key in ("colors", "rays_o", "rays_d"):
            assert np.array_equal(item[key].numpy(), g["val%d_%s" % (k, key)]), (k, key)
    for k in (0, 1):                                                   # train branch with the reference's np.random.choice ids
        item = LO.dataset_getitem(g["train_images"][k], g["train_poses"][k], cax, 40, 48, g["train%d_ids" % k])
        for key in ("colors", "rays_o", "rays_d"):
            assert np.array_equal(item[key].numpy(), g["train%d_%s" % (k, key)]), (k, key)
    img = np.floor(np.random.RandomState(int(g["lego_image_seed"])).uniform(0, 256, size=(800, 800, 3))).astype(np.float32)
    item = LO.dataset_getitem(img, g["lego_pose"], cax, 800, 800, g["lego_ids"])
    for key in ("colors", "rays_o", "rays_d"):
        assert np.array_equal(item[key].numpy(), g["lego_" + key]), key
    # and the path oracle's get_rays agrees with the dataset step (same file:lines, two restatements)
    o, d = O.get_rays_np(g["lego_pose"], 800, 800, g["lego_ids"], cax)
    assert np.array_equal(o, g["lego_rays_o"]) and np.array_equal(d, g["lego_rays_d"])


def test_samplers_match_reference_goldens():
    g = _json("loop_samplers.json")
    for c in g["distributed"]:
        assert LO.distributed_indices(c["n"], c["epoch"], c["rank"], c["world"], c["shuffle"]) == c["indices"]
    for c in g["iteration"]:
        got = LO.iteration_batches(lambda: LO.distributed_indices(c["n"], c["epoch"], c["rank"], c["world"], True),
                                   c["batch_size"], c["iters"])
        assert got == c["batches"]


def test_lr_schedule_matches_reference_golden():
    g = _json("loop_lr.json")
    for e, lr in enumerate(g["lr_by_epoch"]):
        assert LO.exponential_lr(g["base_lr"], e, g["gamma"], g["decay_epochs"]) == lr


def test_evaluator_metrics_match_reference_goldens():
    g = load_golden("loop_evaluator")
    for name in ("noise", "far", "equal", "edges"):
        mse, psnr = LO.evaluate_metrics(g[name + "_pred"], g[name + "_gt"])
        assert mse == g[name + "_mse"] and psnr == g[name + "_psnr"], name
        assert LO.uint8_square_sum(g[name + "_pred"], g[name + "_gt"]) == int(g[name + "_sq8"])
    assert float(g["equal_psnr"]) == 100.0
    # the uint8 wrap-around is real: a plain integer MSE gives a different number on the far-apart case
    p8 = (np.clip(g["far_pred"], 0, 1) * 255).astype(np.uint8).astype(np.int64)
    g8 = (np.clip(g["far_gt"], 0, 1) * 255).astype(np.uint8).astype(np.int64)
    assert int(((p8 - g8) ** 2).sum()) != int(g["far_sq8"])


def test_ssim_restatement_properties():
    rng = np.random.RandomState(0)
    a = rng.uniform(0, 1, size=(24, 31, 3)).astype(np.float32)
    b = np.clip(a + rng.normal(0, 0.1, size=a.shape), 0, 1).astype(np.float32)
    assert LO.ssim_uint8(a, a) == pytest.approx(1.0, abs=1e-12)
    s_ab, s_ba = LO.ssim_uint8(a, b), LO.ssim_uint8(b, a)
    assert 0.0 < s_ab < 1.0 and s_ab == pytest.approx(s_ba, abs=1e-12)


# ---- product host logic ----------------------------------------------------------------------------------------
def test_view_schedule_reproduces_the_reference_sampler_streams():
    from nerf_replication_b200.view_schedule import ViewSchedule, distributed_pass
    g = _json("loop_samplers.json")
    for c in g["distributed"]:
        assert distributed_pass(c["n"], c["epoch"], c["rank"], c["world"], c["shuffle"]) == c["indices"]
        s = ViewSchedule(c["n"], 1, c["shuffle"], True, -1, c["rank"], c["world"])
        s.set_epoch(c["epoch"])
        assert [b[0] for b in s] == c["indices"] and len(s) == len(c["indices"])
    for c in g["iteration"]:
        s = ViewSchedule(c["n"], c["batch_size"], True, True, c["iters"], c["rank"], c["world"])
        s.set_epoch(c["epoch"])
        assert list(s) == c["batches"] and len(s) == c["iters"]
    # single process: torch.utils.data.RandomSampler's stream for a given generator
    gen = torch.Generator().manual_seed(5)
    want = list(torch.utils.data.RandomSampler(range(9), generator=torch.Generator().manual_seed(5)))
    assert [b[0] for b in ViewSchedule(9, 1, True, False, -1, generator=gen)] == want
    assert [b[0] for b in ViewSchedule(4, 1, False)] == [0, 1, 2, 3]


def test_ranks_cover_every_view_once_per_epoch():
    from nerf_replication_b200.view_schedule import distributed_pass
    for n, world in ((100, 8), (100, 3), (5, 8)):
        allv = sum((distributed_pass(n, 4, r, world) for r in range(world)), [])
        assert len(allv) == world * -(-n // world) and set(allv) == set(range(n))


def test_product_exponential_lr_is_the_reference_schedule():
    from nerf_replication_b200.optimizer import ExponentialLR
    g = _json("loop_lr.json")
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([{"params": [p], "lr": g["base_lr"]}], g["base_lr"])
    sch = ExponentialLR(opt, decay_epochs=g["decay_epochs"], gamma=g["gamma"])
    lrs = [opt.param_groups[0]["lr"]]
    for _ in range(600):
        opt.step()
        sch.step()
        lrs.append(opt.param_groups[0]["lr"])
    assert lrs == g["lr_by_epoch"]
    assert sorted(sch.state_dict().keys()) == g["scheduler_state_keys"]


def test_checkpoint_file_has_the_reference_layout(tmp_path):
    """save_model / load_model / load_network on the product Network (CPU, parameters only) + torch.optim.Adam groups as
    make_optimizer builds them: same keys, tensor names, shapes and group structure as the file the reference wrote."""
    from nerf_replication_b200 import checkpoint as ck
    from nerf_replication_b200 import config as C
    from nerf_replication_b200.network import Network
    from nerf_replication_b200.optimizer import ExponentialLR
    from nerf_replication_b200.trainer import Recorder
    g = _json("loop_checkpoint.json")
    torch.manual_seed(0)
    net = Network(cfg=C.default_cfg())
    groups = [{"params": [p], "lr": 5e-4, "weight_decay": 0.0, "eps": 1e-8} for p in net.parameters()]
    opt = torch.optim.Adam(groups, 5e-4, weight_decay=0.0, eps=1e-8)
    sch = ExponentialLR(opt, decay_epochs=500, gamma=0.1)
    for p in net.parameters():
        p.grad = torch.full_like(p, 1e-3)
    opt.step(); opt.step(); sch.step()
    rec = Recorder()
    rec.step = 1000
    d = str(tmp_path / "trained_model")
    assert ck.load_model(net, opt, sch, rec, d) == 0                          # nothing there yet
    ck.save_model(net, opt, sch, rec, d, 1, last=True)
    f = torch.load(os.path.join(d, "latest.pth"), map_location="cpu", weights_only=False)
    assert sorted(f.keys()) == g["keys"] and f["epoch"] == g["epoch"] and f["recorder"] == g["recorder"]
    assert {k: [list(v.shape), str(v.dtype)] for k, v in f["net"].items()} == g["net"]
    assert list(f["net"].keys()) == list(g["net"].keys())                     # registration order = flat layout order
    assert len(f["optim"]["param_groups"]) == g["optim_n_groups"]
    assert [x["params"] for x in f["optim"]["param_groups"]] == g["optim_group_params"]
    assert sorted(f["optim"]["state"][0].keys()) == g["optim_state_keys"] and float(f["optim"]["state"][0]["step"]) == g["optim_state_step"]
    assert sorted(f["scheduler"].keys()) == sorted(g["scheduler"].keys())
    assert f["scheduler"]["last_epoch"] == g["scheduler"]["last_epoch"] and f["scheduler"]["_last_lr"] == g["scheduler"]["_last_lr"]
    # numbered files: at most five kept (net_utils.py:336-343); resume picks latest.pth, else the highest epoch
    for e in range(2, 9):
        ck.save_model(net, opt, sch, rec, d, e)
    kept = sorted(int(x.split(".")[0]) for x in os.listdir(d) if x != "latest.pth")
    assert kept == [4, 5, 6, 7, 8]
    net2 = Network(cfg=C.default_cfg())
    rec2 = Recorder()
    opt2 = torch.optim.Adam([{"params": [p], "lr": 1.0} for p in net2.parameters()], 1.0)
    sch2 = ExponentialLR(opt2, decay_epochs=500, gamma=0.1)
    assert ck.load_model(net2, opt2, sch2, rec2, d) == 2 and rec2.step == 1000                 # latest.pth (epoch 1) wins
    assert all(torch.equal(a, b) for a, b in zip(net.state_dict().values(), net2.state_dict().values()))
    assert opt2.param_groups[0]["lr"] == opt.param_groups[0]["lr"]
    os.remove(os.path.join(d, "latest.pth"))
    assert ck.load_network(Network(cfg=C.default_cfg()), d) == 9                                # highest numbered epoch + 1
    assert ck.load_network(net2, d, epoch=5) == 6
    assert ck.load_model(net2, opt2, sch2, rec2, d, resume=False) == 0 and not os.path.exists(d)  # resume=False wipes (net_utils.py:289-290)


def test_keyed_pixel_bijection_host_copy_equals_the_oracle(built_lib):
    """nerf_bank_permutation_host is plain host C (no CUDA call): the integer restatement must match it bit for bit, and
    the draw must be a set of distinct in-range pixel ids."""
    import ctypes
    from nerf_replication_b200 import _lib
    lib = _lib.load()
    for seed, draw, n_pix, n in ((0, 0, 640000, 1024), (7, 3, 640000, 1024), (2 ** 40 + 5, 2 ** 33, 1920, 1024), (1, 9, 5, 5),
                                 (3, 1, 1, 1), (4, 2, 4096, 4096)):
        ids = np.empty(n, dtype=np.int64)
        assert lib.nerf_bank_permutation_host(seed, draw, n_pix, n, ids.ctypes.data) == 0
        assert np.array_equal(ids, LO.bank_permutation(seed, draw, n_pix, n)), (seed, draw, n_pix, n)
        assert len(set(ids.tolist())) == n and ids.min() >= 0 and ids.max() < n_pix
    full = np.empty(4096, dtype=np.int64)
    lib.nerf_bank_permutation_host(4, 2, 4096, 4096, full.ctypes.data)
    assert sorted(full.tolist()) == list(range(4096))                         # a bijection of [0, n_pix)
    a, b = np.empty(1024, dtype=np.int64), np.empty(1024, dtype=np.int64)
    lib.nerf_bank_permutation_host(0, 0, 640000, 1024, a.ctypes.data)
    lib.nerf_bank_permutation_host(0, 1, 640000, 1024, b.ctypes.data)
    assert len(set(a.tolist()) & set(b.tolist())) < 16                        # different draws are (nearly) disjoint
    assert lib.nerf_bank_permutation_host(0, 0, 10, 11, a.ctypes.data) != 0   # more pixels than the view has
    # uniform over the view: 64 draws x 1024 pixels binned into 16 row bands
    counts = np.zeros(16)
    for dr in range(64):
        lib.nerf_bank_permutation_host(11, dr, 640000, 1024, a.ctypes.data)
        counts += np.bincount(a // 40000, minlength=16)
    exp = 64 * 1024 / 16
    assert ((counts - exp) ** 2 / exp).sum() < 50.0                           # chi^2, 15 dof (p ~ 1e-5 at 50)


def test_dataset_reads_the_blender_layout(tmp_path):
    """Dataset(**cfg.train_dataset) decodes transforms_<split>.json + PNGs like blender.py:26-81 (PIL instead of imageio;
    PNG decoding is lossless) incl. the cv2 bilinear down-sampling for input_ratio != 1."""
    cv2 = pytest.importorskip("cv2")
    from PIL import Image
    from nerf_replication_b200.dataset import Dataset
    images, poses, cax = LO.synthetic_scene(3, 16, 12, seed=4, integral=True)
    root = tmp_path / "nerf_synthetic" / "lego"
    (root / "train").mkdir(parents=True)
    frames = []
    for k in range(3):
        rgba = np.concatenate([images[k].astype(np.uint8), np.full((16, 12, 1), 255, np.uint8)], -1)
        Image.fromarray(rgba, "RGBA").save(root / "train" / ("r_%d.png" % k))
        frames.append({"file_path": "./train/r_%d" % k, "transform_matrix": poses[k].tolist()})
    (root / "transforms_train.json").write_text(json.dumps({"camera_angle_x": cax, "frames": frames}))
    ds = Dataset(data_root=str(tmp_path / "nerf_synthetic"), split="train", input_ratio=1.0, cams=[0, -1, 1], H=16, W=12)
    assert len(ds) == 3 and (ds.H, ds.W) == (16, 12) and ds._host_images.dtype == np.uint8
    assert np.array_equal(ds._host_images, images.astype(np.uint8)) and np.array_equal(ds.transform_matrices, poses)
    assert ds.focal == 12 / (2 * np.tan(cax / 2))
    half = Dataset(data_root=str(tmp_path / "nerf_synthetic"), split="train", input_ratio=0.5, H=16, W=12)
    assert (half.H, half.W) == (8, 6) and half._host_images.dtype == np.float32
    assert np.array_equal(half._host_images[1], cv2.resize(images[1], (6, 8), interpolation=cv2.INTER_LINEAR))
    with pytest.raises(Exception):
        half.bank                                                             # not attached: no CPU path


def test_ray_loader_gives_every_rank_its_own_views_and_pixel_draws():
    """N > 1 without a GPU: two ranks' loaders over one epoch -- views follow the reference's DistributedSampler split, the
    (rank, iteration) -> draw id map is injective, and a resumed loader continues the draw sequence."""
    from nerf_replication_b200.trainer import make_data_loader
    from nerf_replication_b200.view_schedule import distributed_pass
    cfg = SimpleNamespace(train=SimpleNamespace(batch_size=1, shuffle=True), test=SimpleNamespace(batch_size=1))
    ds = list(range(10))                      # the loader only needs len(dataset)
    loaders = [make_data_loader(cfg, ds, True, True, 12, r, 2) for r in range(2)]
    seen = set()
    for r, ld in enumerate(loaders):
        ld.batch_sampler.sampler.set_epoch(3)
        items = list(ld)
        share = distributed_pass(10, 3, r, 2)
        assert [v for v, _ in items] == (share * 3)[:12] and len(ld) == 12
        draws = [d for _, d in items]
        assert draws == [i * 2 + r for i in range(12)]
        seen |= set(draws)
    assert len(seen) == 24
    assert [d for _, d in loaders[1]][:2] == [25, 27]            # next epoch continues where the last one stopped
    with pytest.raises(NotImplementedError):
        make_data_loader(SimpleNamespace(train=SimpleNamespace(batch_size=4, shuffle=True)), ds)
    val = make_data_loader(cfg, ds, is_train=False)
    assert [v for v, _ in val] == list(range(10))
