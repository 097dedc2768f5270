"""nerf_b200 -- B200-native (sm_100a) NeRF ray-render / train engine behind the
rkin100g/Nerf-Replication ``Renderer`` / ``Network`` plug-in interfaces.

Layout:
  csrc/         hand-written CUDA (tcgen05 fused MLP, warp-per-ray stages, fp32 validation path) + C ABI
  _lib.py       ctypes binding of libnerf_b200.so (include/nerf_b200.h)
  engine.py     Engine: handle lifecycle + tensor-level wrappers (raw device pointers, current stream)
  network.py    Network  -- drop-in for src.models.nerf.network.Network
  renderer.py   Renderer -- drop-in for src.models.nerf.renderer.volume_renderer.Renderer
  trainer.py    fused train step (forward_backward + all-reduce + clip/Adam), data-parallel over ranks
"""
__version__ = "0.1.0"
