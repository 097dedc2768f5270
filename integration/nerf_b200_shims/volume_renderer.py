"""Shim resolved by the reference's factory (src/models/nerf/renderer/make_renderer.py:4-8):
``imp.load_source(cfg.renderer_module, cfg.renderer_path).Renderer(network)``; the training wrapper imports
``Renderer`` by name as well (src/train/trainers/nerf.py:3,13).  YAML: ``renderer_module:
nerf_b200_shims.volume_renderer``."""
import os
import sys

_root = os.environ.get("NERF_B200_ROOT", os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
if _root not in sys.path:
    sys.path.insert(0, _root)

from nerf_replication_b200.renderer import Renderer  # noqa: E402,F401
