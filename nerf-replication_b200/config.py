"""Configuration access for the drop-in classes.

In the reference tree the knobs come from the global YACS singleton ``src.config.cfg``
(src/config/config.py:199-209).  Renderer reads top-level keys with getattr defaults
(volume_renderer.py:14-24); Network reads ``cfg.task_arg.*`` / ``cfg.network.*`` (network.py:129-160).
This module returns that object when it is importable and otherwise a plain namespace restating
configs/nerf/lego.yaml, so the engine also runs standalone (tests, bench).

A renderer knob is read exactly like the reference reads it: top-level ``cfg.<key>``, else the reference's default --
``lego.yaml`` keeps these keys under ``task_arg:``, where the reference's Renderer does NOT look, and a
drop-in must not either (tests/test_integration_shims.py compares the two constructors on the reference's own cfg).
Opt-in extension: a cfg with ``renderer_task_arg_fallback = True`` (the standalone ``default_cfg()`` sets it) also
consults ``cfg.task_arg.<key>`` before the default, so that one place configures Network and Renderer.
"""
from types import SimpleNamespace

_MISSING = object()


def default_cfg():
    """configs/nerf/lego.yaml:1,14-43 restated (only the keys the hot path reads)."""
    return SimpleNamespace(
        task="nerf-replication",
        renderer_task_arg_fallback=True,
        task_arg=SimpleNamespace(N_rays=1024, chunk_size=512, white_bkgd=1, N_samples=64, N_importance=128,
                                 use_viewdirs=True, perturb=1),
        network=SimpleNamespace(
            nerf=SimpleNamespace(W=256, D=8, V_D=1, skips=[4]),
            xyz_encoder=SimpleNamespace(type="frequency", input_dim=3, freq=10),
            dir_encoder=SimpleNamespace(type="frequency", input_dim=3, freq=4)),
    )


_override = None


def set_cfg(cfg):
    """Install a cfg object explicitly (tests / bench); None restores automatic lookup."""
    global _override
    _override = cfg


def get_cfg():
    if _override is not None:
        return _override
    try:
        from src.config import cfg  # the reference's singleton, when running inside its tree
        return cfg
    except Exception:
        return default_cfg()


def _get(node, key):
    if node is None:
        return _MISSING
    if isinstance(node, dict):
        return node.get(key, _MISSING)
    return getattr(node, key, _MISSING)


def knob(cfg, key, default):
    v = _get(cfg, key)
    if v is _MISSING and _get(cfg, "renderer_task_arg_fallback") is True:
        v = _get(_get(cfg, "task_arg") if _get(cfg, "task_arg") is not _MISSING else None, key)
    return default if v is _MISSING else v
