"""Shim resolved by the reference's factory (src/models/make_network.py:4-8):
``imp.load_source(cfg.network_module, cfg.network_path).Network()``.
Copy (or symlink) the ``nerf_b200_shims`` directory into the reference's working directory and set
``network_module: nerf_b200_shims.network`` in the YAML; ``NERF_B200_ROOT`` names this repository."""
import os
import sys

_root = os.environ.get("NERF_B200_ROOT", os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
if _root not in sys.path:
    sys.path.insert(0, _root)

from nerf_replication_b200.network import NeRF, Network  # noqa: E402,F401
