"""Shim resolved by the reference's factory (src/evaluators/make_evaluator.py:5-9):
``imp.load_source(cfg.evaluator_module, cfg.evaluator_path).Evaluator()``.  YAML: ``evaluator_module:
nerf_b200_shims.evaluator``.  ``evaluate(output, batch)`` / ``summarize()`` as src/evaluators/nerf.py:80-132; the sums
are computed on the device from the rendered rgb (no device->host copy of the image)."""
import os
import sys

_root = os.environ.get("NERF_B200_ROOT", os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
if _root not in sys.path:
    sys.path.insert(0, _root)

from nerf_replication_b200.evaluator import Evaluator  # noqa: E402,F401
