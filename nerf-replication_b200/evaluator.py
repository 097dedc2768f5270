"""Evaluator metrics behind the render call, on the device.

Mirror of ``src.evaluators.nerf.Evaluator`` (reference evaluators/nerf.py:14-132): ``evaluate(output, batch)``
accumulates MSE, the uint8-quantised PSNR (including numpy's uint8 wrap-around in ``psnr_metric``, :23-30) and the 7x7
uniform-window SSIM (``skimage.metrics.structural_similarity``, :71-77) per view; ``summarize()`` returns the means.  The
sums are computed by csrc/bank.cu kernels from the rendered rgb still in HBM -- the rendered image is never copied to the
host unless ``result_dir`` asks for the PNGs the reference writes (:56-63).
"""
import os

import numpy as np
import torch


class Evaluator:
    def __init__(self, net=None, engine=None, result_dir=None):
        self.mse, self.psnr, self.ssim, self.imgs = [], [], [], []
        self._net, self._engine = net, engine
        self.result_dir = result_dir

    def _eng(self):
        if self._engine is None:
            if self._net is not None:
                return self._net.engine()
            from .engine import Engine          # Evaluator() as the reference's factory builds it: the sums need no weights
            self._engine = Engine()
        return self._engine

    @staticmethod
    def _psnr_from_sum(sq8, n):
        mse = np.float64(sq8) / n                       # np.mean of the uint8 array (float64; integer sum is exact), nerf.py:26
        if mse < 1e-10:
            return 100.0
        return float(10 * np.log10((255 ** 2) / mse))   # :30 (numpy's log10, as the reference)

    def psnr_metric(self, img_pred, img_gt):
        _, sq8 = self._eng().eval_image(img_pred.contiguous().float(), img_gt.contiguous().float())
        return self._psnr_from_sum(int(sq8), img_pred.numel())

    def ssim_metric(self, img_pred, img_gt, batch=None, id=0, num_imgs=1):
        H, W = img_pred.shape[0], img_pred.shape[1]
        if self.result_dir is not None:                 # nerf.py:53-63: view%03d_pred.png / _gt.png, BGR
            import cv2
            d = os.path.join(self.result_dir, "images")
            os.makedirs(d, exist_ok=True)
            for name, img in (("pred", img_pred), ("gt", img_gt)):
                cv2.imwrite("%s/view%03d_%s.png" % (d, id, name), (img.cpu().numpy()[..., [2, 1, 0]] * 255).astype(np.uint8))
        return float(self._eng().eval_ssim(img_pred.contiguous(), img_gt.contiguous(), H, W, 7))

    def evaluate(self, output, batch):
        rgb_pred, _ = output
        rgb_gt = batch["colors"][..., :3].reshape(-1, 3)                      # nerf.py:88-89
        rgb_pred, rgb_gt = rgb_pred.detach().contiguous().float(), rgb_gt.contiguous().float()
        sse, sq8 = self._eng().eval_image(rgb_pred, rgb_gt)
        H, W, vid = int(batch["H"]), int(batch["W"]), int(batch["id"])
        assert H * W == rgb_pred.shape[0], "H*W=%d vs N_rays=%d" % (H * W, rgb_pred.shape[0])   # nerf.py:111
        ssim = self.ssim_metric(rgb_pred.clamp(0, 1).reshape(H, W, 3), rgb_gt.clamp(0, 1).reshape(H, W, 3), batch, vid, 1)
        n = rgb_pred.numel()
        self.mse.append(float(np.float32(float(sse) / n)))                    # :101-103
        self.psnr.append(self._psnr_from_sum(int(sq8), n))                    # :104-106
        self.ssim.append(ssim)                                                # :120

    def summarize(self):
        avg = lambda xs: float(np.mean(xs)) if xs else 0.0                    # nerf.py:123-132
        return {"mse": avg(self.mse), "psnr": avg(self.psnr), "ssim": avg(self.ssim)}
