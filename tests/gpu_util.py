"""Helpers shared by the -m gpu parity tests."""
import numpy as np
import torch

from conftest import load_golden


def dev(x, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(x)) if isinstance(x, np.ndarray) else x
    if dtype is not None:
        t = t.to(dtype)
    return t.cuda().contiguous()


def flat_from_sd(oracle, sd):
    return torch.cat([sd[k].reshape(-1) for k in oracle.state_dict_keys()]).float().cuda().contiguous()


def state_dict_for(oracle, case):
    sd = oracle.init_state_dict(0, sharp="sharp" in case)
    return oracle.round_weights_bf16(sd) if "bf16w" in case else sd


def make_engine(oracle, sd, precision, **kw):
    from nerf_replication_b200.engine import Engine
    eng = Engine(precision=precision, **kw)
    flat = flat_from_sd(oracle, sd)
    eng.set_params(flat)
    eng._keep = flat
    return eng


def maxabs(a, b):
    a = a.detach().cpu().double() if isinstance(a, torch.Tensor) else torch.from_numpy(np.asarray(a)).double()
    b = b.detach().cpu().double() if isinstance(b, torch.Tensor) else torch.from_numpy(np.asarray(b)).double()
    return float((a - b).abs().max()) if a.numel() else 0.0


def quantiles(a, b):
    d = (torch.as_tensor(np.asarray(a.detach().cpu() if isinstance(a, torch.Tensor) else a)).double()
         - torch.as_tensor(np.asarray(b.detach().cpu() if isinstance(b, torch.Tensor) else b)).double()).abs().reshape(-1)
    return {q: float(torch.quantile(d, q)) for q in (0.5, 0.9, 0.99)} | {"max": float(d.max())}
