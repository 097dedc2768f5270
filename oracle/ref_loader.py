"""TEST INFRASTRUCTURE ONLY -- loader for the *unmodified* reference implementation.

Imports rkin100g/Nerf-Replication from ``/root/reference`` (read-only, only present in the
build container, never on the GPU box) so that

  * ``oracle/nerf_oracle.py`` (the CPU restatement that does travel) can be pinned against it, and
  * ``oracle/make_golden.py`` can emit golden vectors under ``tests/golden/``.

Nothing in the product path (``nerf-replication_b200/``), ``bench.py`` or the ``-m gpu`` tests may
import this module.  No reference source is copied: the modules are executed where they lie.

Shims (both are environment shims, not edits of the reference):
  1. a fake ``imp`` module (``imp`` was removed in Python 3.12; used by
     src/models/make_network.py:1 and src/models/nerf/renderer/make_renderer.py:1),
  2. ``torch.cuda.synchronize`` -> no-op on a CPU-only box
     (src/models/nerf/renderer/volume_renderer.py:382,405 call it unconditionally).
The reference parses ``sys.argv`` at import of ``src.config`` (src/config/config.py:199-209) and
resolves paths relative to the cwd (config.py:172-174), so both are set up around the import.
"""
import contextlib
import importlib.util
import io
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("NERF_REFERENCE_ROOT", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "src/models/nerf/renderer/volume_renderer.py"))


def _install_imp_shim():
    if "imp" in sys.modules:
        return
    imp = types.ModuleType("imp")

    def load_source(name, path):
        spec = importlib.util.spec_from_file_location(name, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[name] = mod
        spec.loader.exec_module(mod)
        return mod

    imp.load_source = load_source
    sys.modules["imp"] = imp


_loaded = None


def load_reference(quiet=True):
    """Returns a namespace with the reference's cfg, Network class, Renderer class."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    import torch

    _install_imp_shim()
    if not torch.cuda.is_available():
        torch.cuda.synchronize = lambda *a, **k: None
    old_argv, old_cwd = sys.argv, os.getcwd()
    old_cvd = os.environ.get("CUDA_VISIBLE_DEVICES")
    sys.argv = ["ref_loader", "--cfg_file", "configs/nerf/lego.yaml"]
    os.chdir(REFERENCE_ROOT)
    sys.path.insert(0, REFERENCE_ROOT)
    sink = io.StringIO()
    try:
        with contextlib.redirect_stdout(sink if quiet else sys.stdout), \
             contextlib.redirect_stderr(sink if quiet else sys.stderr):
            from src.config import cfg  # noqa: parses argv + YAML at import
            from src.models.nerf.network import Network
            from src.models.nerf.renderer.volume_renderer import Renderer
    finally:
        sys.argv = old_argv
        os.chdir(old_cwd)
        # parse_cfg sets CUDA_VISIBLE_DEVICES="0" (config.py:139-140); undo it.
        if old_cvd is None:
            os.environ.pop("CUDA_VISIBLE_DEVICES", None)
        else:
            os.environ["CUDA_VISIBLE_DEVICES"] = old_cvd
    ns = types.SimpleNamespace(cfg=cfg, Network=Network, Renderer=Renderer)
    _loaded = ns
    return ns


@contextlib.contextmanager
def silenced():
    """The reference prints progress lines inside render(); silence them in tests."""
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        yield
