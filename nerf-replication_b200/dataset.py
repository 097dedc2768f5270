"""Dataset step in front of the hot path, kept on the device.

Mirror of ``src.datasets.nerf.blender.Dataset`` (reference blender.py:7-184): same constructor kwargs
(``data_root, split, input_ratio, cams, H, W``), same ``__len__`` and the same ``__getitem__`` dictionary
(``colors, rays_o, rays_d, H, W, id``) -- but every view's image and pose is uploaded once into the engine's
ray bank (csrc/bank.cu) and ``__getitem__`` is ONE kernel that writes the batch straight into HBM:
no DataLoader worker, no host->device copy per step.  PNG / JSON decoding stays in Python (PIL, cv2), exactly the
part of the reference that runs once at start-up.

Train split: the reference draws ``np.random.choice(H*W, 1024, replace=False)`` per item (blender.py:108-110).
``__getitem__(i)`` draws the 1024 distinct pixels on the device instead (keyed bijection, seed + draw counter);
``getitem(i, pixel_ids=...)`` takes the caller's ids and then reproduces the reference item bit for bit.
Not carried: the test split's depth / normal maps (blender.py:55-71,133-157) -- nothing on the path reads them.
"""
import json
import math
import os

import numpy as np
import torch

from .engine import EngineError


class Dataset:
    N_TRAIN_PIXELS = 1024                       # blender.py:108

    def __init__(self, net=None, engine=None, images=None, transform_matrices=None, camera_angle_x=None, seed=0,
                 n_rays=None, **kwargs):
        """Either the reference's kwargs (reads ``transforms_<split>.json`` + PNGs under data_root/lego) or arrays:
        images [V,H,W,3] (uint8, or float32 0..255), transform_matrices [V,4,4], camera_angle_x."""
        self.split = kwargs.get("split", "val")                      # blender.py:28
        self.H, self.W = kwargs.get("H", 800), kwargs.get("W", 800)
        self.input_ratio = kwargs.get("input_ratio", 1)
        self.n_rays = self.N_TRAIN_PIXELS if n_rays is None else int(n_rays)
        self.seed, self.draws = int(seed), 0
        if images is None:
            images, transform_matrices, camera_angle_x = self._load_blender(kwargs)
        images = np.asarray(images)
        if images.ndim != 4 or images.shape[-1] != 3:
            raise ValueError("images must be [V,H,W,3]")
        self.H, self.W = int(images.shape[1]), int(images.shape[2])
        self.camera_angle_x = float(camera_angle_x)
        self.transform_matrices = np.asarray(transform_matrices, dtype=np.float64)
        if self.transform_matrices.shape[0] != images.shape[0]:
            raise ValueError("one transform_matrix per image")
        # integral pixel values (a decoded PNG at input_ratio 1) are kept as uint8: same colors, a quarter of the HBM
        if images.dtype != np.uint8:
            images = np.ascontiguousarray(images, dtype=np.float32)
            if np.array_equal(images, np.floor(images)) and images.min() >= 0 and images.max() <= 255:
                images = images.astype(np.uint8)
        self._host_images = np.ascontiguousarray(images)
        self.focal = self.W / (2 * math.tan(self.camera_angle_x / 2))   # blender.py:102
        self.cx, self.cy = self.W / 2, self.H / 2                       # blender.py:103
        self._engine, self._bank = None, None
        if net is not None or engine is not None:
            self.attach(net=net, engine=engine)

    # -- disk (start-up only) ----------------------------------------------------------------------------
    def _load_blender(self, kwargs):
        import cv2
        from PIL import Image
        root = os.path.join(kwargs.get("data_root", "data/nerf_synthetic"), "lego")      # blender.py:26-27
        with open(os.path.join(root, "transforms_%s.json" % self.split)) as f:
            meta = json.load(f)
        H, W = self.H, self.W
        if self.input_ratio != 1.0:                                                      # blender.py:44-45
            H, W = int(H * self.input_ratio), int(W * self.input_ratio)
        images, poses = [], []
        for frame in meta["frames"]:
            img = np.array(Image.open(os.path.join(root, frame["file_path"] + ".png")), dtype=np.float32)[..., :3]
            images.append(cv2.resize(img, (W, H), interpolation=cv2.INTER_LINEAR))      # blender.py:50-52
            poses.append(frame["transform_matrix"])
        return np.stack(images), np.array(poses, dtype=np.float64), meta["camera_angle_x"]

    # -- device residency ---------------------------------------------------------------------------------
    def attach(self, net=None, engine=None):
        """Uploads every view into the engine's ray bank (once).  ``net``: the Network whose engine trains on it."""
        eng = engine if engine is not None else net.engine()
        if self._engine is eng:
            return self
        self.detach()
        self._engine = eng
        self._bank = eng.bank_create(len(self), self.H, self.W, self.focal, self.cx, self.cy,
                                     uint8=self._host_images.dtype == np.uint8)
        for v in range(len(self)):
            eng.bank_set_view(self._bank, v, self._host_images[v], self.transform_matrices[v])
        return self

    def detach(self):
        if self._engine is not None and getattr(self._engine, "h", None):
            self._engine.bank_destroy(self._bank)
        self._engine, self._bank = None, None

    @property
    def bank(self):
        if self._bank is None:
            raise EngineError("Dataset is not attached to an engine yet (Dataset.attach(net)); there is no CPU path")
        return self._bank

    @property
    def engine(self):
        self.bank
        return self._engine

    # -- reference interface ----------------------------------------------------------------------------------
    def __len__(self):
        return int(self._host_images.shape[0])

    def next_draw(self):
        d = self.draws
        self.draws += 1
        return d

    def getitem(self, index, pixel_ids=None, draw=None, want_ids=False):
        eng = self.engine
        if self.split == "train":
            if pixel_ids is not None:
                out = eng.bank_sample(self.bank, index, pixel_ids.numel(), mode="ids", pixel_ids=pixel_ids, want_ids=want_ids)
            else:
                if eng.cfg.seed != self.seed:
                    eng.configure(seed=self.seed)
                out = eng.bank_sample(self.bank, index, self.n_rays, mode="random",
                                      draw=self.next_draw() if draw is None else draw, want_ids=want_ids)
        else:
            out = eng.bank_sample(self.bank, index, self.H * self.W, mode="range", pix0=0, want_ids=want_ids)
        out.update(H=torch.tensor(self.H, dtype=torch.int), W=torch.tensor(self.W, dtype=torch.int),
                   id=torch.tensor(index, dtype=torch.int))              # blender.py:163-165 (host scalars, as collated)
        return out

    def __getitem__(self, index):
        return self.getitem(index)

    def batch(self, index, **kw):
        """The item as the reference's DataLoader would collate it at batch_size 1 (leading dimension 1)."""
        return {k: v[None] for k, v in self.getitem(index, **kw).items()}
