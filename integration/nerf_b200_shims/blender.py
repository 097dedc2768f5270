"""Shim resolved by the reference's dataset factory (src/datasets/make_dataset.py:17-31):
``imp.load_source(cfg.train_dataset_module, cfg.train_dataset_path).Dataset(**cfg.train_dataset)``.  YAML:
``train_dataset_module: nerf_b200_shims.blender`` (and ``test_dataset_module``).  The items are CUDA tensors produced by one
kernel from images resident in HBM, so the loader around it must not fork workers or pin memory: use
``nerf_replication_b200.trainer.make_data_loader`` (same argument meaning as src/datasets/make_dataset.py:73-100) instead
of the reference's DataLoader wrapper, or set ``cfg.train.num_workers = 0`` and drop ``pin_memory``."""
import os
import sys

_root = os.environ.get("NERF_B200_ROOT", os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
if _root not in sys.path:
    sys.path.insert(0, _root)

from nerf_replication_b200.dataset import Dataset  # noqa: E402,F401
