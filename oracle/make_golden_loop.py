"""TEST INFRASTRUCTURE ONLY -- pins oracle/loop_oracle.py (dataset step, samplers, lr schedule,
evaluator metrics, checkpoint layout) to the unmodified reference and writes tests/golden/loop_*.  Run in the build
container (needs /root/reference):

    python oracle/make_golden_loop.py

The reference modules are executed where they lie.  Environment shims only (no reference edits): the ``imp`` /
``torch.cuda.synchronize`` shims of ref_loader.py plus empty stand-ins for packages that are not installed here and are
not on the code under test: termcolor (net_utils.py:7 colours a message), skimage.metrics / tensorboardX (imported at
module level by evaluators/nerf.py:5 and recorder.py:3; neither is called).  ``Dataset.__init__`` reads PNGs from disk
with imageio (not installed, no dataset here), so the dataset object is created with ``__new__`` and given the
attributes ``__init__`` would set; ``__getitem__`` -- the code being pinned -- then runs unmodified.
"""
import json
import os
import shutil
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
from oracle import loop_oracle as LO      # noqa: E402
from oracle import ref_loader             # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def load_reference_loop_modules():
    ref_loader.load_reference()
    for name in ("termcolor", "skimage", "skimage.metrics", "tensorboardX"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["termcolor"].colored = lambda s, *a, **k: s
    sys.modules["skimage.metrics"].structural_similarity = None
    sys.modules["tensorboardX"].SummaryWriter = object
    cwd = os.getcwd()
    os.chdir(ref_loader.REFERENCE_ROOT)
    try:
        from src.datasets.nerf import blender
        from src.datasets import samplers
        from src.utils.optimizer.lr_scheduler import ExponentialLR
        from src.evaluators.nerf import Evaluator
        from src.utils import net_utils
        # src/train/__init__.py imports the trainer stack (imgaug, tensorboardX ...): load the one file the step driver
        # uses, src/train/optimizer.py, by path
        import imp
        make_optimizer = imp.load_source("ref_train_optimizer", os.path.join("src", "train", "optimizer.py")).make_optimizer
    finally:
        os.chdir(cwd)
    return types.SimpleNamespace(blender=blender, samplers=samplers, ExponentialLR=ExponentialLR, Evaluator=Evaluator,
                                 net_utils=net_utils, make_optimizer=make_optimizer, cfg=ref_loader.load_reference().cfg)


def ref_dataset(ref, images, poses, cax, split):
    ds = ref.blender.Dataset.__new__(ref.blender.Dataset)
    ds.split, ds.H, ds.W, ds.camera_angle_x = split, images.shape[1], images.shape[2], cax
    ds.data = [{"image": images[k], "transform_matrix": poses[k].tolist()} for k in range(len(images))]
    return ds


def dataset_goldens(ref):
    out = {}
    # val branch: every pixel, row-major (blender.py:112-115); fractional pixel values as after cv2.resize
    images, poses, cax = LO.synthetic_scene(3, 24, 20, seed=1, integral=False)
    ds = ref_dataset(ref, images, poses, cax, "val")
    for k in (0, 2):
        item = ds[k]
        mine = LO.dataset_getitem(images[k], poses[k], cax, 24, 20)
        for key in ("colors", "rays_o", "rays_d"):
            assert torch.equal(item[key], mine[key]), ("val", k, key)
            out["val%d_%s" % (k, key)] = item[key].numpy()
        assert int(item["H"]) == 24 and int(item["W"]) == 20 and int(item["id"]) == k
    out.update(val_images=images, val_poses=poses, camera_angle_x=cax)
    # train branch: 1024 pixels by np.random.choice (blender.py:108-110) -- the ids are recovered by replaying the seed
    images, poses, cax = LO.synthetic_scene(2, 40, 48, seed=2, integral=True)
    ds = ref_dataset(ref, images, poses, cax, "train")
    for k in (0, 1):
        np.random.seed(100 + k)
        item = ds[k]
        np.random.seed(100 + k)
        ids = np.random.choice(40 * 48, 1024, replace=False)
        mine = LO.dataset_getitem(images[k], poses[k], cax, 40, 48, ids)
        for key in ("colors", "rays_o", "rays_d"):
            assert torch.equal(item[key], mine[key]), ("train", k, key)
            out["train%d_%s" % (k, key)] = item[key].numpy()
        out["train%d_ids" % k] = ids.astype(np.int64)
    out.update(train_images=images, train_poses=poses)
    # the lego-sized case: one 800x800 view, 1024 reference-drawn pixels (image content irrelevant: colours only gathered)
    rng = np.random.RandomState(3)
    img = np.floor(rng.uniform(0, 256, size=(800, 800, 3))).astype(np.float32)
    pose = LO.synthetic_scene(8, 2, 2)[1][5]
    ds = ref_dataset(ref, img[None], pose[None], cax, "train")
    np.random.seed(7)
    item = ds[0]
    np.random.seed(7)
    ids = np.random.choice(800 * 800, 1024, replace=False)
    mine = LO.dataset_getitem(img, pose, cax, 800, 800, ids)
    for key in ("colors", "rays_o", "rays_d"):
        assert torch.equal(item[key], mine[key]), ("lego", key)
    out.update(lego_ids=ids.astype(np.int64), lego_pose=pose, lego_rays_o=item["rays_o"].numpy(), lego_rays_d=item["rays_d"].numpy(),
               lego_colors=item["colors"].numpy(), lego_image_seed=3)
    np.savez_compressed(os.path.join(GOLD, "loop_dataset.npz"), **out)
    print("golden loop_dataset: val 24x20 x2 views, train 40x48 x2 views, lego 800x800 1024 px -- restatement bit-equal")


def sampler_goldens(ref):
    S = ref.samplers
    out = {"distributed": [], "iteration": []}

    class _DS:                                   # DistributedSampler only calls len(dataset)
        def __init__(self, n):
            self.n = n

        def __len__(self):
            return self.n

    for n, world, shuffle in ((100, 1, True), (100, 8, True), (100, 3, True), (7, 4, False), (5, 8, True)):
        for epoch in (0, 1, 17):
            for rank in range(world):
                s = S.DistributedSampler(_DS(n), num_replicas=world, rank=rank, shuffle=shuffle)
                s.set_epoch(epoch)
                got = list(iter(s))
                assert got == LO.distributed_indices(n, epoch, rank, world, shuffle)
                out["distributed"].append({"n": n, "world": world, "shuffle": shuffle, "epoch": epoch, "rank": rank, "indices": got})
    for n, world, rank, bs, iters in ((100, 2, 1, 1, 500), (10, 1, 0, 1, 25), (10, 1, 0, 4, 7), (3, 4, 2, 1, 5)):
        s = S.DistributedSampler(_DS(n), num_replicas=world, rank=rank, shuffle=True)
        s.set_epoch(3)
        bsamp = torch.utils.data.sampler.BatchSampler(s, bs, False)
        got = [list(b) for b in S.IterationBasedBatchSampler(bsamp, iters)]
        mine = LO.iteration_batches(lambda: LO.distributed_indices(n, 3, rank, world, True), bs, iters)
        assert got == mine and len(got) == iters, (n, world, rank, bs, iters)
        out["iteration"].append({"n": n, "world": world, "rank": rank, "batch_size": bs, "iters": iters, "epoch": 3, "batches": got})
    with open(os.path.join(GOLD, "loop_samplers.json"), "w") as f:
        json.dump(out, f)
    print("golden loop_samplers: %d DistributedSampler passes, %d iteration-based streams" % (len(out["distributed"]), len(out["iteration"])))


def lr_goldens(ref):
    p = torch.nn.Parameter(torch.zeros(1))
    opt = torch.optim.Adam([{"params": [p], "lr": 5e-4}], 5e-4)
    sch = ref.ExponentialLR(opt, decay_epochs=500, gamma=0.1)
    lrs = [opt.param_groups[0]["lr"]]
    for _ in range(600):
        opt.step()
        sch.step()
        lrs.append(opt.param_groups[0]["lr"])
    for e, lr in enumerate(lrs):
        assert lr == LO.exponential_lr(5e-4, e, 0.1, 500), e
    keys = sorted(sch.state_dict().keys())
    with open(os.path.join(GOLD, "loop_lr.json"), "w") as f:
        json.dump({"base_lr": 5e-4, "gamma": 0.1, "decay_epochs": 500, "lr_by_epoch": lrs, "scheduler_state_keys": keys}, f)
    print("golden loop_lr: 601 epochs, lr[500] = %.6g" % lrs[500])


def evaluator_goldens(ref):
    ev = ref.Evaluator()
    rng = np.random.RandomState(11)
    out, cases = {}, {}
    gt = rng.uniform(0, 1, size=(600, 3)).astype(np.float32)
    cases["noise"] = (np.clip(gt + rng.normal(0, 0.05, size=gt.shape), -0.2, 1.2).astype(np.float32), gt)
    cases["far"] = (rng.uniform(-0.3, 1.3, size=gt.shape).astype(np.float32), gt)     # wrap-around of the uint8 difference
    cases["equal"] = (gt.copy(), gt)                                                  # mse < 1e-10 -> 100.0
    cases["edges"] = (np.array([[0.0, 1.0, 0.5], [1.0 / 255, 254.999 / 255, 1.0], [0.99999994, 0.0039215689, 2.0]], np.float32),
                      np.array([[1.0, 0.0, 0.5], [0.0, 1.0, 0.0], [0.0, 1.0, -1.0]], np.float32))
    for name, (pred, g) in cases.items():
        # evaluate(), nerf.py:96-104, up to the psnr call (the rest needs cv2.imwrite into the reference tree + skimage)
        p_c, g_c = np.clip(pred, 0, 1), np.clip(g, 0, 1)
        mse = np.mean((p_c - g_c) ** 2)
        psnr = ev.psnr_metric(p_c, g_c)
        m_mse, m_psnr = LO.evaluate_metrics(pred, g)
        assert mse == m_mse and psnr == m_psnr, name
        n = pred.size
        s8 = LO.uint8_square_sum(pred, g)
        assert (100.0 if s8 / n < 1e-10 else 10 * np.log10(255 ** 2 / (s8 / n))) == psnr, name
        out[name + "_pred"], out[name + "_gt"] = pred, g
        out[name + "_mse"], out[name + "_psnr"], out[name + "_sq8"] = np.float32(mse), np.float64(psnr), np.int64(s8)
    np.savez_compressed(os.path.join(GOLD, "loop_evaluator.npz"), **out)
    print("golden loop_evaluator: %s" % ", ".join("%s psnr %.4f" % (k, out[k + "_psnr"]) for k in cases))


def checkpoint_goldens(ref):
    """save_model / load_model of the reference on its own Network + make_optimizer + ExponentialLR: records the file's
    structure (keys, shapes, dtypes) -- the 14 MB file itself is not committed -- and checks that the reference loads
    a file laid out by the restatement."""
    tmp = os.path.join(ROOT, "gpurun_out", "_ckpt_tmp")
    shutil.rmtree(tmp, ignore_errors=True)
    os.makedirs(tmp)
    rl = ref_loader.load_reference()
    torch.manual_seed(0)
    net = rl.Network()
    cfg = ref.cfg
    opt = ref.make_optimizer(cfg, net)
    sch = ref.ExponentialLR(opt, decay_epochs=500, gamma=0.1)
    for p in net.parameters():
        p.grad = torch.full_like(p, 1e-3)
    opt.step(); opt.step()
    sch.step()
    rec = types.SimpleNamespace(state_dict=lambda: {"step": 1000}, load_state_dict=lambda d: None)
    ref.net_utils.save_model(net, opt, sch, rec, tmp, 1, last=True)
    ck = torch.load(os.path.join(tmp, "latest.pth"), map_location="cpu", weights_only=False)
    mine = LO.checkpoint_dict(net.state_dict(), opt.state_dict(), sch.state_dict(), {"step": 1000}, 1)
    assert sorted(ck.keys()) == sorted(mine.keys())
    desc = {
        "keys": sorted(ck.keys()),
        "net": {k: [list(v.shape), str(v.dtype)] for k, v in ck["net"].items()},
        "optim_param_group_keys": sorted(ck["optim"]["param_groups"][0].keys()),
        "optim_n_groups": len(ck["optim"]["param_groups"]),
        "optim_group_params": [g["params"] for g in ck["optim"]["param_groups"]],
        "optim_state_keys": sorted(ck["optim"]["state"][0].keys()),
        "optim_state_step": float(ck["optim"]["state"][0]["step"]),
        "optim_lr": ck["optim"]["param_groups"][0]["lr"], "optim_eps": ck["optim"]["param_groups"][0]["eps"],
        "scheduler": {k: (v if isinstance(v, (int, float, list, bool)) or v is None else str(v)) for k, v in ck["scheduler"].items()},
        "recorder": ck["recorder"], "epoch": ck["epoch"], "torch": torch.__version__,
    }
    with open(os.path.join(GOLD, "loop_checkpoint.json"), "w") as f:
        json.dump(desc, f, indent=1)
    # the reference resumes from the same layout: epoch + 1, parameters and moments restored
    torch.manual_seed(1)
    net2 = rl.Network()
    opt2 = ref.make_optimizer(cfg, net2)
    sch2 = ref.ExponentialLR(opt2, decay_epochs=500, gamma=0.1)
    with ref_loader.silenced():
        begin = ref.net_utils.load_model(net2, opt2, sch2, rec, tmp, resume=True)
    assert begin == 2 and all(torch.equal(a, b) for a, b in zip(net.state_dict().values(), net2.state_dict().values()))
    shutil.rmtree(tmp, ignore_errors=True)
    print("golden loop_checkpoint: %d net tensors, %d optimizer groups, scheduler keys %s" %
          (len(desc["net"]), desc["optim_n_groups"], sorted(desc["scheduler"].keys())))


if __name__ == "__main__":
    if not ref_loader.available():
        raise SystemExit("reference tree not found at %s" % ref_loader.REFERENCE_ROOT)
    ref = load_reference_loop_modules()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    dataset_goldens(ref)
    sampler_goldens(ref)
    lr_goldens(ref)
    evaluator_goldens(ref)
    checkpoint_goldens(ref)
