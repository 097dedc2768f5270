"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the steps either side of the hot path:
the dataset step, the view samplers, the lr schedule, the evaluator metrics and the checkpoint layout of
rkin100g/Nerf-Replication.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module; the product path (nerf-replication_b200/) never does.

Pinned by oracle/make_golden_loop.py, which runs the reference's own Dataset.__getitem__, samplers, ExponentialLR,
Evaluator.psnr_metric and save_model/load_model on the same inputs and asserts equality before writing
tests/golden/loop_*.npz|json.  File:line citations are relative to the reference root.
"""
import math

import numpy as np
import torch


# ----------------------------------------------------------------------------------------------------------
# Dataset.__getitem__  (src/datasets/nerf/blender.py:84-170)
def dataset_getitem(image, transform_matrix, camera_angle_x, H, W, pixel_ids=None):
    """image [H,W,3] float32 in 0..255 (blender.py:50-52), transform_matrix 4x4 (nested list or array).
    pixel_ids None = every pixel in row-major order (the val/test branch, :112-115); otherwise the ids the train
    branch drew with np.random.choice (:108-110).  Returns float32 colors, rays_o, rays_d like the returned tensors."""
    transform_matrix = np.array(transform_matrix)                       # :99
    f = W / (2 * math.tan(camera_angle_x / 2))                          # :102
    cx, cy = W / 2, H / 2                                               # :103
    if pixel_ids is None:
        us, vs = np.meshgrid(np.arange(W), np.arange(H))                # :114
        u, v = us.ravel(), vs.ravel()
    else:
        pixel_ids = np.asarray(pixel_ids)
        u, v = pixel_ids % W, pixel_ids // W                            # :109-110
    colors = image[v, u] / 255.0                                        # :117 (float32 / python float -> float32)
    dirs = np.stack([(u - cx) / f, -(v - cy) / f, -np.ones_like(u)], axis=-1)   # :122
    rays_d = (transform_matrix[:3, :3] @ dirs.T).T                      # :125
    rays_d = rays_d / np.linalg.norm(rays_d, axis=-1, keepdims=True)    # :126
    rays_o = np.broadcast_to(transform_matrix[:3, 3], rays_d.shape).copy()      # :127
    return {"colors": torch.from_numpy(colors).float(), "rays_o": torch.from_numpy(rays_o).float(),
            "rays_d": torch.from_numpy(rays_d).float()}                 # :160-162


# ----------------------------------------------------------------------------------------------------------
# The engine's own on-device pixel draw (nerf_bank_sample mode 2): a keyed bijection of [0, n_pix).  The reference
# draws with np.random.choice(H*W, 1024, replace=False) (blender.py:108), which no device kernel can reproduce
# stream for stream; what is kept is the contract -- n distinct pixel ids, uniform over the view.  This is the exact
# integer restatement of csrc/bank.cu (perm_at / make_perm_key) the device ids are compared with.
_M32, _M64 = (1 << 32) - 1, (1 << 64) - 1


def _perm_mix(x):
    x ^= x >> 16
    x = (x * 0x7FEB352D) & _M32
    x ^= x >> 15
    x = (x * 0x846CA68B) & _M32
    x ^= x >> 16
    return x


def bank_permutation(seed, draw, n_pix, n):
    bits = 2
    while bits < 64 and (1 << bits) < n_pix:
        bits += 1
    hb = (bits + 1) // 2
    mask = (1 << hb) - 1
    keys = []
    for r in range(6):
        z = (seed + 0x9E3779B97F4A7C15 * ((draw * 8 + r + 1) & _M64)) & _M64
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _M64
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _M64
        z ^= z >> 31
        keys.append(z & _M32)
    out = np.empty(n, dtype=np.int64)
    for i in range(n):
        x = i
        while True:
            l, r = (x >> hb) & mask, x & mask
            for k in keys:
                l, r = r, l ^ (_perm_mix((r + k) & _M32) & mask)
            x = (l << hb) | r
            if x < n_pix:
                break
        out[i] = x
    return out


# ----------------------------------------------------------------------------------------------------------
# samplers  (src/datasets/samplers.py:50-131, src/datasets/make_dataset.py:37-71)
def distributed_indices(n_items, epoch, rank, world, shuffle=True):
    """DistributedSampler.__iter__ (samplers.py:105-125): epoch-seeded randperm, padded to a multiple of the world
    size, contiguous slice per rank."""
    if shuffle:
        g = torch.Generator()
        g.manual_seed(epoch)
        indices = torch.randperm(n_items, generator=g).tolist()
    else:
        indices = torch.arange(n_items).tolist()
    num_samples = int(math.ceil(n_items * 1.0 / world))
    total = num_samples * world
    indices += indices[: (total - len(indices))]
    return indices[num_samples * rank: num_samples * rank + num_samples]


def iteration_batches(make_pass, batch_size, num_iterations, drop_last=False):
    """BatchSampler (make_dataset.py:55-58) wrapped by IterationBasedBatchSampler (samplers.py:50-72): keep starting
    new passes over the sampler until num_iterations batches have been produced.  make_pass() -> iterable of indices."""
    out = []
    while len(out) < num_iterations:
        batch = []
        for idx in make_pass():
            batch.append(idx)
            if len(batch) == batch_size:
                out.append(batch)
                batch = []
                if len(out) == num_iterations:
                    return out
        if batch and not drop_last:
            out.append(batch)
    return out[:num_iterations]


# ----------------------------------------------------------------------------------------------------------
# lr schedule  (src/utils/optimizer/lr_scheduler.py:68-79, stepped once per epoch in train.py:66)
def exponential_lr(base_lr, epoch, gamma=0.1, decay_epochs=500):
    return base_lr * gamma ** (epoch / decay_epochs)


# ----------------------------------------------------------------------------------------------------------
# Evaluator  (src/evaluators/nerf.py:23-30, 80-103)
def psnr_metric(img_pred, img_gt):
    img_pred = (np.clip(img_pred, 0, 1) * 255).astype(np.uint8)         # :24
    img_gt = (np.clip(img_gt, 0, 1) * 255).astype(np.uint8)             # :25
    mse = np.mean((img_pred - img_gt) ** 2)                             # :26  uint8 arithmetic: both ops wrap mod 256
    if mse < 1e-10:
        return 100.0
    return 10 * np.log10((255 ** 2) / mse)                              # :30


def evaluate_metrics(rgb_pred, rgb_gt):
    """evaluate(), nerf.py:86-103, without the SSIM / image-writing tail: returns (mse float32, psnr float64)."""
    rgb_pred = np.clip(np.asarray(rgb_pred, dtype=np.float32), 0, 1)    # :98
    rgb_gt = np.clip(np.asarray(rgb_gt, dtype=np.float32).reshape(-1, 3), 0, 1)   # :88-89, :99
    mse = np.mean((rgb_pred - rgb_gt) ** 2)                             # :101
    return mse, psnr_metric(rgb_pred, rgb_gt)                           # :104


def uint8_square_sum(rgb_pred, rgb_gt):
    """The exact integer behind psnr_metric's mean: sum(((p8 - g8) mod 256)^2 mod 256)."""
    p8 = (np.clip(np.asarray(rgb_pred, dtype=np.float32), 0, 1) * 255).astype(np.uint8)
    g8 = (np.clip(np.asarray(rgb_gt, dtype=np.float32), 0, 1) * 255).astype(np.uint8)
    return int(((p8 - g8) ** 2).astype(np.uint64).sum())


# ----------------------------------------------------------------------------------------------------------
# structural similarity as the evaluator calls it (nerf.py:71-77): skimage.metrics.structural_similarity(uint8 pred,
# uint8 gt, win_size=7, full=True, channel_axis=2).  scikit-image is not installed here (README pins no version; the
# algorithm below is the published one of skimage >= 0.19: Wang et al. 2004 with a uniform 7x7 window, K1=0.01,
# K2=0.03, sample covariance, data_range 255 for uint8, mean over the border-cropped map, then over channels),
# restated on scipy.ndimage.uniform_filter, which is what skimage itself calls.  PARITY UNPINNED (no skimage to run).
def ssim_uint8(img_pred, img_gt, win_size=7):
    from scipy.ndimage import uniform_filter
    p8 = (np.asarray(img_pred, dtype=np.float32) * 255).astype(np.uint8)      # nerf.py:66-67
    g8 = (np.asarray(img_gt, dtype=np.float32) * 255).astype(np.uint8)
    K1, K2, R = 0.01, 0.03, 255.0
    NP = win_size ** 2
    cov_norm = NP / (NP - 1.0)
    C1, C2 = (K1 * R) ** 2, (K2 * R) ** 2
    pad = (win_size - 1) // 2
    vals = []
    for ch in range(p8.shape[2]):
        x, y = p8[..., ch].astype(np.float64), g8[..., ch].astype(np.float64)
        ux, uy = uniform_filter(x, size=win_size), uniform_filter(y, size=win_size)
        uxx, uyy, uxy = uniform_filter(x * x, size=win_size), uniform_filter(y * y, size=win_size), uniform_filter(x * y, size=win_size)
        vx, vy, vxy = cov_norm * (uxx - ux * ux), cov_norm * (uyy - uy * uy), cov_norm * (uxy - ux * uy)
        S = ((2 * ux * uy + C1) * (2 * vxy + C2)) / ((ux ** 2 + uy ** 2 + C1) * (vx + vy + C2))
        vals.append(S[pad:-pad, pad:-pad].mean(dtype=np.float64))
    return float(np.mean(vals))


# ----------------------------------------------------------------------------------------------------------
# checkpoint layout  (src/utils/net_utils.py:288-343)
def checkpoint_dict(net_state, optim_state, scheduler_state, recorder_state, epoch):
    return {"net": net_state, "optim": optim_state, "scheduler": scheduler_state, "recorder": recorder_state,
            "epoch": epoch}                                             # :324-330


# ----------------------------------------------------------------------------------------------------------
# synthetic scene used by the tests and the bench (no dataset on disk, no network): lego-shaped cameras
#  and a procedural image per view with integral 0..255 values, like a decoded PNG.
def synthetic_scene(n_views, H, W, seed=0, integral=True):
    from oracle import nerf_oracle as O
    rng = np.random.RandomState(seed)
    images = np.empty((n_views, H, W, 3), dtype=np.float32)
    yy, xx = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    for k in range(n_views):
        base = np.stack([127.5 + 127.5 * np.sin(0.07 * xx + k), 127.5 + 127.5 * np.cos(0.05 * yy - k),
                         255.0 * ((xx + yy + 13 * k) % 64) / 63.0], -1)
        img = base + rng.uniform(-8, 8, size=base.shape)
        img = np.clip(img, 0, 255)
        images[k] = np.floor(img) if integral else img
    poses = np.stack([O.lego_pose(k, n_views) for k in range(n_views)])
    return images, poses, O.CAMERA_ANGLE_X
