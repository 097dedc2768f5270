"""Stand-alone timing of the HBM-bound stage kernels ("achieved HBM GB/s for the non-MLP stages when run
as standalone kernels"): CUDA events on the launching stream around 20 back-to-back launches on one 65 536-ray chunk
(the engine's own chunk size; inputs/outputs 170-340 MB, larger than the 126 MB L2), algorithmic bytes per ray from
DESIGN.md 5.5.  python profiles/stage_kernels.py  (B200)"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from nerf_replication_b200.engine import Engine  # noqa: E402
from nerf_replication_b200 import _lib as L  # noqa: E402


def hbm_peak():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6555.8, "fallback"


def timed(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3      # us per launch


def main():
    import ctypes as C
    n, Nc, Nf = 65536, 64, 128
    Nt = Nc + Nf
    eng = Engine()
    lib, h = eng.lib, eng.h
    st = lambda: eng._stream()
    dev = "cuda"
    peak, kind = hbm_peak()
    g = torch.Generator(device="cuda").manual_seed(0)
    rays_o = torch.randn(n, 3, device=dev, generator=g)
    rays_d = torch.nn.functional.normalize(torch.randn(n, 3, device=dev, generator=g), dim=-1)
    t_c = torch.empty(n, Nc, device=dev)
    raw_c = torch.randn(n, Nc, 4, device=dev, generator=g)
    raw_f = torch.randn(n, Nt, 4, device=dev, generator=g)
    t_s = torch.empty(n, Nt, device=dev)
    sort_idx = torch.empty(n, Nt, dtype=torch.int32, device=dev)
    rgb, depth, acc = torch.empty(n, 3, device=dev), torch.empty(n, device=dev), torch.empty(n, device=dev)
    pose = (C.c_double * 12)(1, 0, 0, 0.5, 0, 1, 0, -1.0, 0, 0, 1, 4.0)
    p = lambda t: C.c_void_p(t.data_ptr())
    flat = torch.randn(L.NERF_PARAMS_TOTAL, device=dev, generator=g) * 0.05
    grad, m, v = torch.randn_like(flat), torch.zeros_like(flat), torch.zeros_like(flat)
    pts = torch.randn(n * 16, 3, device=dev, generator=g)
    enc = torch.empty(n * 16, 63, device=dev)
    runs = {
        "get_rays_kernel (n rays of one 800x800 view)": (lambda: lib.nerf_get_rays(h, pose, 800, 800, 1111.111, 400.0, 400.0, None, 0, n, p(rays_o), p(rays_d), st()), 24),
        "stratified_t_kernel": (lambda: lib.nerf_stratified_t(h, n, Nc, None, p(t_c), st()), 4 * Nc),
        "freq_encode_kernel (16 points/ray, standalone tap)": (lambda: lib.nerf_freq_encode(h, p(pts), n * 16, 10, p(enc), st()), 16 * (12 + 63 * 4)),
        "composite_kernel (S=192)": (lambda: lib.nerf_composite(h, p(raw_f), p(t_s), n, Nt, 1, p(rgb), p(depth), p(acc), None, st()), Nt * 20 + 20),
        "composite_kernel (S=64)": (lambda: lib.nerf_composite(h, p(raw_c), p(t_c), n, Nc, 1, p(rgb), p(depth), p(acc), None, st()), Nc * 20 + 20),
        "sample_pdf_kernel (+merge; t_sorted only, as in render)": (lambda: lib.nerf_sample_pdf(h, p(raw_c), p(t_c), n, Nc, Nf, None, None, None, None, None, None, None, p(t_s), None, st()), Nc * 20 + Nt * 4),
        "sample_pdf_kernel (+merge; t_sorted + sort_idx, as in training)": (lambda: lib.nerf_sample_pdf(h, p(raw_c), p(t_c), n, Nc, Nf, None, None, None, None, None, None, None, p(t_s), p(sort_idx), st()), Nc * 20 + Nt * 8),
    }
    lib.nerf_stratified_t(h, n, Nc, None, p(t_c), st())
    lib.nerf_sample_pdf(h, p(raw_c), p(t_c), n, Nc, Nf, None, None, None, None, None, None, None, p(t_s), None, st())
    out = {"rays": n, "hbm_peak_gbs": peak, "peak_kind": kind, "kernels": {}}
    for name, (fn, bytes_per_ray) in runs.items():
        us = timed(fn)
        gbs = n * bytes_per_ray / us / 1e3
        out["kernels"][name] = {"us": round(us, 2), "bytes_per_ray": bytes_per_ray, "GBps": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 3)}
    eng.set_params(flat)
    step = [0]

    def adam():
        step[0] += 1
        lib.nerf_clip_adam(h, p(flat), p(grad), p(m), p(v), C.c_float(5e-4), C.c_float(0.9), C.c_float(0.999), C.c_float(1e-8), C.c_float(40.0), step[0], st())
    us = timed(adam)
    bytes_ = L.NERF_PARAMS_TOTAL * 30
    out["kernels"]["clip_adam_kernel + 2 weight re-pack kernels (per step)"] = {"us": round(us, 2), "bytes": bytes_, "GBps": round(bytes_ / us / 1e3, 1),
                                                                              "frac_of_hbm_peak": round(bytes_ / us / 1e3 / peak, 3)}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
