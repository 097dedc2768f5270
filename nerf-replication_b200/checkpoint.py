"""Checkpoint files in the reference's layout (src/utils/net_utils.py:288-379): ``{"net", "optim", "scheduler",
"recorder", "epoch"}`` saved as ``<epoch>.pth`` / ``latest.pth`` under ``cfg.trained_model_dir``, at most five numbered
files kept.  ``Network`` has the reference's state_dict keys and ``FusedAdam`` torch.optim.Adam's state layout, so a
``latest.pth`` written by either side resumes on the other.
"""
import os
import shutil

import torch


def _numbered(model_dir):
    return [int(f.split(".")[0]) for f in os.listdir(model_dir) if f != "latest.pth" and f.split(".")[0].isdigit()]


def _pick(model_dir, epoch):
    names = os.listdir(model_dir)
    nums = _numbered(model_dir)
    if not nums and "latest.pth" not in names:
        return None
    if epoch == -1:
        return "latest" if "latest.pth" in names else max(nums)
    return epoch


def load_model(net, optim, scheduler, recorder, model_dir, resume=True, epoch=-1):
    """net_utils.py:288-319 -> the epoch to start from (0 when nothing was loaded)."""
    if not resume and os.path.isdir(model_dir):
        shutil.rmtree(model_dir)
    if not os.path.exists(model_dir):
        return 0
    pth = _pick(model_dir, epoch)
    if pth is None:
        return 0
    path = os.path.join(model_dir, "%s.pth" % pth)
    print("load model: %s" % path)
    ck = torch.load(path, map_location="cpu", weights_only=False)
    net.load_state_dict(ck["net"])
    if "optim" not in ck:
        return 0
    optim.load_state_dict(ck["optim"])
    scheduler.load_state_dict(ck["scheduler"])
    recorder.load_state_dict(ck["recorder"])
    return ck["epoch"] + 1


def save_model(net, optim, scheduler, recorder, model_dir, epoch, last=False):
    """net_utils.py:322-343."""
    os.makedirs(model_dir, exist_ok=True)
    ck = {"net": net.state_dict(), "optim": optim.state_dict(), "scheduler": scheduler.state_dict(),
          "recorder": recorder.state_dict(), "epoch": epoch}
    torch.save(ck, os.path.join(model_dir, "latest.pth" if last else "%d.pth" % epoch))
    nums = _numbered(model_dir)
    if len(nums) > 5:
        os.remove(os.path.join(model_dir, "%d.pth" % min(nums)))


def load_network(net, model_dir, resume=True, epoch=-1, strict=True):
    """net_utils.py:346-379: weights only (evaluation); model_dir may also be a file."""
    if not resume:
        return 0
    if not os.path.exists(model_dir):
        print("pretrained model does not exist")
        return 0
    if os.path.isdir(model_dir):
        pth = _pick(model_dir, epoch)
        if pth is None:
            return 0
        path = os.path.join(model_dir, "%s.pth" % pth)
    else:
        path = model_dir
    print("load model: %s" % path)
    ck = torch.load(path, map_location="cpu", weights_only=False)
    net.load_state_dict(ck["net"], strict=strict)
    return ck["epoch"] + 1 if "epoch" in ck else 0
