"""nerf_b200 -- B200-native (sm_100a) NeRF ray-render / train engine behind the
rkin100g/Nerf-Replication ``Renderer`` / ``Network`` plug-in interfaces.

Layout:
  csrc/         hand-written CUDA (tcgen05 fused MLP, warp-per-ray stages, fp32 validation path) + C ABI
  _lib.py       ctypes binding of libnerf_b200.so (include/nerf_b200.h)
  engine.py     Engine: handle lifecycle + tensor-level wrappers (raw device pointers, current stream)
  network.py    Network  -- drop-in for src.models.nerf.network.Network
  renderer.py   Renderer -- drop-in for src.models.nerf.renderer.volume_renderer.Renderer
  trainer.py    fused train step (forward_backward + all-reduce + clip/Adam), data-parallel over ranks; the step driver
                (Trainer.train / val, fit) of train.py + trainers/trainer.py
  dataset.py    Dataset  -- src.datasets.nerf.blender.Dataset resident in HBM (ray bank, csrc/bank.cu)
  view_schedule.py  the reference sampler stack's index stream (which view each rank trains on)
  optimizer.py  FusedAdam (torch.optim.Adam state layout on the flat buffers), ExponentialLR, make_* factories
  checkpoint.py save_model / load_model / load_network in the reference's file layout
  evaluator.py  Evaluator -- MSE / uint8 PSNR / SSIM from device-side sums
"""
__version__ = "0.1.0"
