"""Synthetic lego-shaped workload: there is no dataset on the build or GPU boxes, so the bench and the
examples generate cameras, rays and training batches of the shape of `nerf_synthetic/lego` themselves.  Rays always come
from the engine's own get_rays kernel (a-B) -- nothing here computes on the host beyond a 4x4 pose."""
import math

import numpy as np
import torch

CAMERA_ANGLE_X = 0.6911112070083618      # the public lego transforms_*.json value
ORBIT_RADIUS = 4.031128874


def focal_from_fov(W, camera_angle_x=CAMERA_ANGLE_X):
    return W / (2.0 * math.tan(camera_angle_x / 2.0))          # blender.py:102


def lego_pose(k=0, n_views=8, elevation_deg=30.0, radius=ORBIT_RADIUS):
    """Camera-to-world 4x4 (float64): camera on a sphere looking at the origin, Blender/OpenGL convention (-z forward,
    +y up) as blender.py:122-127 expects; azimuth 2 pi k / n_views."""
    az, el = 2.0 * math.pi * k / n_views, math.radians(elevation_deg)
    eye = radius * np.array([math.cos(el) * math.cos(az), math.cos(el) * math.sin(az), math.sin(el)])
    fwd = -eye / np.linalg.norm(eye)
    right = np.cross(fwd, np.array([0.0, 0.0, 1.0]))
    right /= np.linalg.norm(right)
    up = np.cross(right, fwd)
    c2w = np.eye(4)
    c2w[:3, 0], c2w[:3, 1], c2w[:3, 2], c2w[:3, 3] = right, up, -fwd, eye
    return c2w


def view_rays(engine, k, n_views, H, W, row0=0, row1=None):
    """rays_o, rays_d [rows*W, 3] (device) of rows [row0, row1) of view k."""
    row1 = H if row1 is None else row1
    return engine.get_rays(lego_pose(k, n_views), H, W, focal_from_fov(W), pix0=row0 * W, n=(row1 - row0) * W)


def train_batch(engine, step, rank, n_rays, H=800, W=800, n_views=8):
    """Per step s and rank r: generator seeded with 1000 s + r, pixel ids by torch.randint over view s mod n_views, targets
    torch.rand (SURVEY 8(d) "Training inputs") -> device rays_o, rays_d, rgb_gt."""
    g = torch.Generator().manual_seed(1000 * step + rank)
    pix = torch.randint(0, H * W, (n_rays,), generator=g)
    rgb_gt = torch.rand(n_rays, 3, generator=g)
    dev = "cuda:%d" % engine.device
    o, d = engine.get_rays(lego_pose(step % n_views, n_views), H, W, focal_from_fov(W), pixel_ids=pix.to(dev))
    return o, d, rgb_gt.to(dev)
