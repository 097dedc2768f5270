#!/usr/bin/env python
"""bench.py -- rays/s of the NeRF render AND train hot path on N B200s, one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--mode all|render|train]

Headline (`value`, `e2e`, `roofline`): BASELINE.json configs[1] ("C2") -- ONE synthetic lego-shaped 800x800 view, N_samples
64 + N_importance 128, bf16 tcgen05 MLP, random-init weights (torch.manual_seed(0); Network()), perturb = 0.  A "step" is one
frame (640 000 rays).  At N > 1 GPUs (one process per GPU, torchrun) the frame is cut into N row bands ("image tiles go to the
GPUs"), rank r renders band r and the bands are gathered (NCCL all-gather of 16 B/ray) so that rank 0 holds the frame:
total work is fixed -> `"scaling": "strong"`, value = 640 000 / frame latency (max over ranks, gather included).

Further records on the same line (same process, after the headline legs):
  weak     (N > 1) N views x N bands, 640 000 rays per GPU and step, no collective -- round 1's figure, kept as a secondary
  train    BASELINE configs[2] ("C3"): 1024 rays / rank / step, fwd + bwd + gradient all-reduce (NCCL) + clip/Adam; device-timed
           value, e2e from pinned host batches, per-kernel-class times and TFLOP/s, all-reduce exposed vs hidden, and the
           checks `param_checksum_equal_across_ranks` and `dp_grad_check` (data-parallel gradient == single-rank gradient of
           the concatenated batch)
  train_c5 BASELINE configs[4]: 8192 rays / rank / step, device-timed
  gpu_baseline  (N = 1) the reference's algorithm as eager PyTorch ON THE SAME B200 (oracle/nerf_oracle.py with tensors on the
           GPU: fp32, TF32 off, 512-row MLP chunks, 64-sample fine blocks = the reference's launch structure), render + train
  cpu_baseline  (N = 1) the same restatement on the host cores (kind "port": the Python reference cannot travel to the box)

value   : device-timed (CUDA events per step, max over ranks), inputs resident in HBM
e2e     : the same frame through the host-buffer path: pinned host rays in, host rgb/depth out, copies inside the timed region
          (N = 1: nerf_render_rays_host, the C-ABI host call; N > 1: pinned rays -> H2D -> Renderer.render -> gather -> D2H on rank 0)
roofline: the fused MLP kernel -- algorithmic FLOPs (1 186 816 per sample evaluation, un-padded) over the CUDA-event duration
          of its launches, vs the measured dense bf16 peak (MEASURED_PEAKS.json)
"""
import argparse
import contextlib
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H = W = 800
N_SAMPLES, N_IMPORTANCE = 64, 128
FLOP_PER_EVAL = 1186816
TRAIN_RAYS, TRAIN_RAYS_C5 = 1024, 8192


def train_flop_per_ray():
    # forward E evals + fine backward (wgrad + dgrad) per fine eval + coarse (sigma-only) backward per coarse eval
    return (2 * N_SAMPLES + N_IMPORTANCE) * FLOP_PER_EVAL + (N_SAMPLES + N_IMPORTANCE) * 2366720 + N_SAMPLES * 1900544


def measured_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        for k in ("hbm_gbs_sustained", "hbm_gbs", "hbm_gbps"):
            if k in d:
                return float(d[k]), "measured"
    return 6500.0, "fallback"


# dram__bytes_read.sum + dram__bytes_write.sum of one launch (the fine-net launch of each class), from the committed round-2
# ncu captures: profiles/stages_render_r2z_metrics.txt (launch list with DRAM bytes; fine-net launches 54.1 MB read + 145.3 MB
# written) and profiles/train_r2z_metrics.txt (ncu --set full of the training step's kernels)
NCU_TRAFFIC_BYTES = {"render": {"fwd": 54.1e6 + 145.3e6},
                     "train": {"fwd": 0.0022e9 + 1.0055e9, "dgrad": 0.0628e9 + 0.9509e9, "wgrad": 1.9327e9 + 0.0056e9}}
# DRAM bytes (read + written) of ALL eight MLP launches of one 1024-ray step in the same capture (forward coarse/fine, dgrad
# fine/coarse, wgrad + heads fine/coarse): 2.721 GB read + 2.475 GB written.  The activation / gradient arenas scale with the
# ray count and the 2.5 MB of packed weights do not matter, so other batch sizes are scaled linearly.
NCU_TRAIN_STEP_DRAM_BYTES_1024 = 2.7211e9 + 2.4751e9


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d.get("bf16_tflops_sustained", d.get("bf16_tflops")), d.get("bf16_tflops"), "measured"
    return 1400.0, 1590.0, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock / throttle reasons of one GPU every 100 ms while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown,
                 "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                 "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                 "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if r & bit:
                        self.reasons.add(k)
            except Exception:
                pass
            self._halt.wait(0.1)

    def finish(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def physical_gpu_index(local_rank):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_rank])
        except Exception:
            return local_rank
    return local_rank


# ------------------------------------------------------------------------------------------------------
def cpu_render_rate(seconds_budget, threads, max_chunks=16, chunk=2048):
    """Times the oracle's render() (the reference's algorithm, CPU fp32 torch) on a bounded sample of the
    C2 workload: successive 2048-ray slices of the 800x800 view until the budget is used."""
    import torch
    from oracle import nerf_oracle as O
    torch.set_num_threads(threads)
    sd = O.init_state_dict(0)
    ro, rd = O.get_rays_np(O.lego_pose(0), H, W)
    ro, rd = torch.from_numpy(ro), torch.from_numpy(rd)
    start = (H // 2) * W          # centre rows: rays that cross the scene volume
    with torch.no_grad():
        O.render(sd, ro[start:start + 256], rd[start:start + 256], N_SAMPLES, N_IMPORTANCE)    # warm-up
        done, t0 = 0, time.perf_counter()
        for c in range(max_chunks):
            a = start + c * chunk
            O.render(sd, ro[a:a + chunk], rd[a:a + chunk], N_SAMPLES, N_IMPORTANCE)
            done += chunk
            if time.perf_counter() - t0 >= seconds_budget:
                break
        dt = time.perf_counter() - t0
    return done / dt, done, dt


def gpu_baseline(render_rays=16384, train_steps=3):
    """The like-for-like "before" figure: the reference's algorithm as eager PyTorch
    on this B200 -- oracle/nerf_oracle.py with every tensor on the GPU: fp32 GEMMs (TF32 off), MLP in chunks of 512 rows
    (network.py:163-171, 230), fine pass in sample blocks of 64 (volume_renderer.py:382-405), ~25 ATen kernels per chunk.
    Render: a bounded slice of the C2 view; train: 1024-ray steps with torch autograd + clip_grad_value_ + torch.optim.Adam.
    The arm is first checked against the CPU oracle on the 64 golden rays so that it is known to run the same algorithm."""
    import numpy as np
    import torch
    from oracle import nerf_oracle as O
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.backends.cudnn.allow_tf32 = False
    dev = torch.device("cuda")
    sd_cpu = O.init_state_dict(0)
    sd = {k: v.to(dev) for k, v in sd_cpu.items()}
    g = np.load(os.path.join(ROOT, "tests", "golden", "render_seed0_64rays.npz"))
    with torch.no_grad():
        rgb_g, dep_g = O.render(sd, torch.from_numpy(g["rays_o"]).to(dev), torch.from_numpy(g["rays_d"]).to(dev), 64, 128)
    check = {"rays": 64, "max_abs_rgb_vs_reference_golden": float((rgb_g.cpu() - torch.from_numpy(g["rgb_map"])).abs().max()),
             "max_abs_depth_vs_reference_golden": float((dep_g.cpu() - torch.from_numpy(g["depth_map"])).abs().max()), "tol_rgb": 1e-4,
             "note": "the golden is the reference on the CPU; cuBLAS accumulates in another order than the CPU GEMM, which flips a few "
                     "searchsorted indices in near-empty bins (SURVEY H1): the reference's own CUDA-vs-CPU distance exceeds 1e-5 on some rays"}
    check["ok"] = check["max_abs_rgb_vs_reference_golden"] <= 1e-4
    ro, rd = O.get_rays_np(O.lego_pose(0), H, W)
    start = (H // 2) * W
    ro, rd = torch.from_numpy(ro[start:start + render_rays]).to(dev), torch.from_numpy(rd[start:start + render_rays]).to(dev)
    with torch.no_grad():
        O.render(sd, ro[:1024], rd[:1024], N_SAMPLES, N_IMPORTANCE)          # warm-up (cuBLAS handles, allocator)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        O.render(sd, ro, rd, N_SAMPLES, N_IMPORTANCE)
        torch.cuda.synchronize()
        dt_r = time.perf_counter() - t0
    tr = O.Trainer(sd)
    batches = [tuple(t.to(dev) for t in O.synthetic_train_batch(s, 0, TRAIN_RAYS, H, W)[:3]) for s in range(train_steps + 1)]
    tr.step(*batches[0])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for b in batches[1:]:
        tr.step(*b)
    torch.cuda.synchronize()
    dt_t = (time.perf_counter() - t0) / train_steps
    return {"kind": "port-cuda",
            "what": "oracle/nerf_oracle.py (reference algorithm, eager PyTorch) on the same B200: fp32, TF32 off, 512-row MLP chunks, 64-sample fine blocks",
            "render": {"value": render_rays / dt_r, "unit": "rays/s", "rays": render_rays, "ms": 1e3 * dt_r,
                       "sample": "%d rays from the centre rows of the C2 view, one pass" % render_rays},
            "train": {"value": TRAIN_RAYS / dt_t, "unit": "rays/s", "rays_per_step": TRAIN_RAYS, "ms_per_step": 1e3 * dt_t, "steps": train_steps,
                      "what": "torch autograd + clip_grad_value_(40) + torch.optim.Adam(5e-4, eps 1e-8), no anomaly mode"},
            "same_algorithm_check": check}


def run_reference(args, rank):
    """--impl reference: the reference's own CPU path (oracle port) on all host threads; rank 0 only."""
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    import torch
    from oracle import nerf_oracle as O
    torch.set_num_threads(threads)
    chunk = 2048
    sd = O.init_state_dict(0)
    ro, rd = O.get_rays_np(O.lego_pose(0), H, W)
    ro, rd = torch.from_numpy(ro), torch.from_numpy(rd)
    start = (H // 2) * W

    def step(i):
        a = start + (i % 64) * chunk
        with torch.no_grad():
            O.render(sd, ro[a:a + chunk], rd[a:a + chunk], N_SAMPLES, N_IMPORTANCE)

    for i in range(args.warmup):
        step(i)
    t0 = time.perf_counter()
    for i in range(args.steps):
        step(args.warmup + i)
    dt = time.perf_counter() - t0
    v = chunk * args.steps / dt
    sample = "%d-ray slices of the 800x800 C2 view per step (oracle/nerf_oracle.py, CPU fp32 torch)" % chunk
    print(json.dumps({
        "impl": "reference", "metric": metric_name(), "value": v, "unit": "rays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, max(args.gpus, 1)),      # the same workload object as the GPU arm's line (one frame)
        "cpu_baseline": {"value": v, "unit": "rays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def metric_name():
    return "rays/sec render (lego 800x800, %d+%d samples); train record alongside" % (N_SAMPLES, N_IMPORTANCE)


def workload_config(args, world):
    extra = ""
    if args.fast_sampling is not None:
        extra = " + fast_sampling (ESS/ERT, weights_threshold %g, masked samples skipped)%s [secondary variant]" % (
            args.fast_sampling, ", fog weights" if args.fog else "")
    return {"workload": ("lego.yaml full 800x800 synthetic-camera render, %d+%d samples, bf16 MLP (BASELINE configs[%d])" % (
                N_SAMPLES, N_IMPORTANCE, 1 if (N_SAMPLES, N_IMPORTANCE) == (64, 128) else 3)) + extra,
            "H": H, "W": W, "n_samples": N_SAMPLES, "n_importance": N_IMPORTANCE, "white_bkgd": True, "perturb": 0,
            "rays_per_step": H * W, "views_per_step": 1, "rays_per_gpu_per_step": H * W // world,
            "parallelism": ("one frame cut into %d row bands, one per GPU; bands all-gathered (NCCL, 16 B/ray) so rank 0 holds the frame" % world)
                           if world > 1 else "1 GPU",
            "l2": "flushed between timed steps (256 MiB write); per-frame workspace also exceeds L2"}


# ------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--mode", default="all", choices=["all", "render", "train"])       # which legs run; the line always has the headline keys
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default=None, choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true")
    # secondary (non-headline) render variant: the reference's fast_sampling (ESS/ERT) knob with this weights_threshold;
    # --fog adds +1 to alpha_linear.bias (sigma ~ 1) so that the mask keeps a realistic share of the fine samples
    ap.add_argument("--fast-sampling", type=float, default=None, metavar="WEIGHTS_THRESHOLD")
    ap.add_argument("--fog", action="store_true")
    # BASELINE.json configs[3] = --samples 128 256
    ap.add_argument("--samples", type=int, nargs=2, default=None, metavar=("N_SAMPLES", "N_IMPORTANCE"))
    ap.add_argument("--train-rays", type=int, default=None)
    ap.add_argument("--no-graph", action="store_true", help="plain stream launches for the training step (per-kernel profiling)")
    ap.add_argument("--no-overlap", action="store_true", help="serialise the gradient all-reduce behind the backward pass")
    args = ap.parse_args()
    global N_SAMPLES, N_IMPORTANCE, TRAIN_RAYS
    if args.samples:
        N_SAMPLES, N_IMPORTANCE = args.samples
    if args.train_rays:
        TRAIN_RAYS = args.train_rays
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200 (there is no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":      # its banner goes to stdout, which carries the JSON line
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    import __graft_entry__
    if rank == 0:
        with contextlib.redirect_stdout(sys.stderr):          # stdout carries exactly one JSON line
            __graft_entry__.build()
    if world > 1:
        dist.barrier()

    from nerf_replication_b200 import build as BUILD
    from nerf_replication_b200 import synthetic as SYN     # lego-shaped cameras / batches; rays from the engine's get_rays kernel
    from nerf_replication_b200 import config as C
    from nerf_replication_b200.network import Network
    from nerf_replication_b200.renderer import Renderer
    from nerf_replication_b200.trainer import FusedTrainer, shard_range

    cfg = C.default_cfg()
    cfg.task_arg.N_samples, cfg.task_arg.N_importance = N_SAMPLES, N_IMPORTANCE
    if args.fast_sampling is not None:
        cfg.fast_sampling, cfg.weights_threshold = True, args.fast_sampling
    precision = args.precision or "bf16"
    torch.manual_seed(0)
    net = Network(cfg=cfg, precision=precision)
    if args.fog:
        with torch.no_grad():
            net.model.alpha_linear.bias += 1.0
            net.model_fine.alpha_linear.bias += 1.0
    net = net.cuda()
    renderer = Renderer(net, cfg=cfg)
    eng = renderer._engine()
    if args.no_graph:
        eng.set_graph_mode(False)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    dev = torch.device("cuda", local_rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps, warmup, mode, after_warmup=None):
        """`warmup` untimed steps, then `steps` timed ones; L2 flushed before every step, outside the brackets.
        mode "events": every step bracketed by its own CUDA event pair, all steps enqueued before the first read-back (the
        flush between two steps keeps the host ahead of the device); "wall": host clock around a step that ends with its own
        synchronisation (host-buffer paths), ranks aligned by a barrier first.  -> per-step ms, max over ranks."""
        for i in range(warmup):
            step_fn(i)
        barrier()
        if after_warmup is not None:
            after_warmup()                       # e.g. reset the launch counter / kernel brackets: only timed steps count
        ms = []
        if mode == "events":
            ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
            for i, (e0, e1) in enumerate(ev):
                flush.fill_(i & 0xFF)
                e0.record()
                step_fn(warmup + i)
                e1.record()
            torch.cuda.synchronize()
            ms = [e0.elapsed_time(e1) for e0, e1 in ev]
        else:
            for i in range(steps):
                flush.fill_(i & 0xFF)
                barrier()
                t0 = time.perf_counter()
                step_fn(warmup + i)
                torch.cuda.synchronize()
                ms.append(1e3 * (time.perf_counter() - t0))
        barrier()
        t = torch.tensor(ms, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.cpu().tolist()

    peak_sus, peak_burst, peak_kind = measured_peaks()
    line = {"metric": metric_name(), "unit": "rays/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": precision, "data": "synthetic",
            "config": workload_config(args, world),
            "build": dict(BUILD.last_build) if rank == 0 else None}

    # ================================================================================================
    # headline: ONE 800x800 frame over `world` GPUs (strong scaling), bands gathered so rank 0 holds the frame
    # ================================================================================================
    r0, r1 = shard_range(H, rank, world)
    n_local = (r1 - r0) * W
    assert H % world == 0, "all-gather of equal bands needs H % n_gpus == 0"
    band_o, band_d = SYN.view_rays(eng, 0, 8, H, W, r0, r1)
    host_o, host_d = band_o.cpu().pin_memory(), band_d.cpu().pin_memory()
    batch = {"rays_o": band_o[None], "rays_d": band_d[None]}
    frame_rgb = torch.empty(H * W, 3, device=dev) if world > 1 else None
    frame_dep = torch.empty(H * W, device=dev) if world > 1 else None
    host_rgb = torch.empty(H * W if rank == 0 else n_local, 3).pin_memory()
    host_dep = torch.empty(H * W if rank == 0 else n_local).pin_memory()

    def gather(rgb, depth):
        dist.all_gather_into_tensor(frame_rgb, rgb.contiguous())
        dist.all_gather_into_tensor(frame_dep, depth.contiguous())

    def render_step_device(i):
        with torch.no_grad():
            rgb, depth = renderer.render(batch)
        if world > 1:
            gather(rgb, depth)

    def render_step_e2e(i):
        if world == 1:
            eng.render_rays_host(host_o, host_d, host_rgb, host_dep)          # the C-ABI host call: chunked H2D / render / D2H
            return
        o, d = host_o.to(dev, non_blocking=True), host_d.to(dev, non_blocking=True)       # run.py:167-169 batch[k].cuda()
        with torch.no_grad():
            rgb, depth = renderer.render({"rays_o": o[None], "rays_d": d[None]})
        gather(rgb, depth)
        if rank == 0:
            host_rgb.copy_(frame_rgb, non_blocking=True)
            host_dep.copy_(frame_dep, non_blocking=True)

    do_render = args.mode in ("all", "render")
    do_train = args.mode in ("all", "train") and args.fast_sampling is None
    value = None
    if do_render:
        sampler = ClockSampler(physical_gpu_index(local_rank))
        sampler.start()
        eng.profile(True)
        mark = {}

        def reset_counters():
            eng.profile_read()
            mark["launches0"] = eng.launch_count

        ms_dev = timed(render_step_device, args.steps, args.warmup, "events", after_warmup=reset_counters)
        launches = eng.launch_count - mark["launches0"]
        mlp_ms, mlp_launches, mlp_rows = eng.profile_read_classes()["fwd"]
        eng.profile(False)
        clocks = sampler.finish()
        ms_e2e = timed(render_step_e2e, args.steps, args.warmup, "wall")

        total_ms = sum(ms_dev)
        value = H * W * args.steps / (total_ms / 1e3)
        e2e_value = H * W * args.steps / (sum(ms_e2e) / 1e3)
        flop_per_ray = (2 * N_SAMPLES + N_IMPORTANCE) * FLOP_PER_EVAL
        roofline = None
        if precision == "bf16" and mlp_launches > 0 and mlp_ms > 0 and args.fast_sampling is None:
            # event brackets around every fused-MLP launch of the timed steps, on the launching stream
            achieved = mlp_rows * FLOP_PER_EVAL / (mlp_ms / 1e3) / 1e12
            roofline = {"bound": "tensor", "achieved": achieved, "peak": peak_sus, "unit": "TFLOP/s", "frac": achieved / peak_sus,
                        "traffic": NCU_TRAFFIC_BYTES["render"]["fwd"], "traffic_unit": "bytes per launch (ncu dram read + write, fine-net launch)",
                        "kernel": "mlp_tc2_forward_kernel (CTA pair, tcgen05 cta_group::2)",
                        "peak_kind": "%s sustained bf16 (burst %.1f)" % (peak_kind, peak_burst),
                        "launches_timed": mlp_launches, "avg_launch_ms": mlp_ms / mlp_launches,
                        "share_of_step": mlp_ms / total_ms if world == 1 else None,
                        "step_tflops_per_gpu": n_local * flop_per_ray / (total_ms / args.steps / 1e3) / 1e12}
        line.update({"value": value, "ms_per_step": total_ms / args.steps,
                     "e2e": {"value": e2e_value, "unit": "rays/s", "h2d_bytes_per_step": H * W * 24, "d2h_bytes_per_step": H * W * 16,
                             "ms_per_step": sum(ms_e2e) / args.steps,
                             "api": "nerf_render_rays_host (C ABI, host buffers)" if world == 1 else
                                    "pinned host rays -> H2D -> Renderer.render (band) -> NCCL all-gather -> D2H of the frame on rank 0"},
                     "gpu_launches": launches, "clocks": clocks, "roofline": roofline})
        if world > 1:
            line["gather"] = {"collective": "2 x ncclAllGather (rgb 12 B/ray, depth 4 B/ray)", "bytes_per_rank_per_frame": n_local * 16}

    # ---- secondary: weak scaling (round 1's figure): N views x N bands, 640 000 rays per GPU, no collective ----------------
    if world > 1 and do_render:
        ro_l, rd_l = [], []
        for v in range(world):
            o, d = SYN.view_rays(eng, v, max(world, 8), H, W, r0, r1)
            ro_l.append(o)
            rd_l.append(d)
        wbatch = {"rays_o": torch.cat(ro_l)[None], "rays_d": torch.cat(rd_l)[None]}

        def weak_step(i):
            with torch.no_grad():
                renderer.render(wbatch)

        ms_w = timed(weak_step, args.steps, args.warmup, "events")
        line["weak"] = {"value": H * W * world * args.steps / (sum(ms_w) / 1e3), "unit": "rays/s", "ms_per_step": sum(ms_w) / args.steps,
                        "scaling": "weak", "rays_per_gpu_per_step": H * W, "views_per_step": world,
                        "what": "%d views x %d row bands, rank r renders band r of every view; no collective, no gather" % (world, world)}
        del wbatch, ro_l, rd_l

    # ================================================================================================
    # train: BASELINE configs[2] (1024 rays / rank / step) and configs[4] (8192)
    # ================================================================================================
    if do_train:
        line["train"] = train_record(args, eng, net, cfg, rank, world, dev, TRAIN_RAYS, timed, barrier, peak_sus, full=True)
        if args.mode == "all" and not args.train_rays:
            line["train_c5"] = train_record(args, eng, net, cfg, rank, world, dev, TRAIN_RAYS_C5, timed, barrier, peak_sus, full=False)

    if not do_render and "train" in line:           # --mode train (profiling runs): the line's headline keys carry the train record
        t = line["train"]
        line.update({"metric": "rays/sec train (lego, %d rays/rank/step)" % TRAIN_RAYS, "value": t["value"], "ms_per_step": t["ms_per_step"],
                     "scaling": "weak", "e2e": t.get("e2e"), "gpu_launches": int(t["gpu_launches_per_step"] * t["steps"]), "roofline": None})
        value = t["value"]

    # ---- baselines on rank 0 of a single-GPU run --------------------------------------------------------------------------
    if rank == 0 and world == 1 and not args.no_gpu_baseline and (N_SAMPLES, N_IMPORTANCE) == (64, 128) and args.fast_sampling is None:
        try:
            gb = gpu_baseline()
            if do_render:
                gb["render"]["speedup_of_this_engine"] = value / gb["render"]["value"]
            if "train" in line:
                gb["train"]["speedup_of_this_engine"] = line["train"]["value"] / gb["train"]["value"]
            line["gpu_baseline"] = gb
        except Exception as e:                                    # a baseline must never take the headline down with it
            line["gpu_baseline"] = {"kind": "port-cuda", "error": repr(e)}
    if rank == 0 and world == 1 and not args.no_cpu_baseline and do_render:
        threads = os.cpu_count() or 1
        v, n_done, dt = cpu_render_rate(12.0, threads)
        line["cpu_baseline"] = {"value": v, "unit": "rays/s", "cores": threads, "kind": "port",
                                "sample": "%d rays (2048-ray slices from the centre rows of the same 800x800 view), %.1f s, oracle/nerf_oracle.py CPU fp32 torch" % (n_done, dt)}
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def train_record(args, eng, net0, cfg, rank, world, dev, n_rays, timed, barrier, peak_sus, full):
    """One training configuration: n_rays / rank / step, synthetic rays + RGB, fwd + bwd + all-reduce + clip/Adam."""
    import torch
    import torch.distributed as dist
    from nerf_replication_b200 import synthetic as SYN
    from nerf_replication_b200.network import Network
    from nerf_replication_b200.trainer import FusedTrainer
    torch.manual_seed(0)
    net = Network(cfg=cfg, precision=net0.precision).cuda()       # every configuration starts from the same seed-0 weights
    trainer = FusedTrainer(net, cfg=cfg, overlap=not args.no_overlap)
    teng = trainer.renderer._engine()
    if args.no_graph:
        teng.set_graph_mode(False)
    steps = args.steps * (10 if n_rays <= 1024 else 2)
    warm = max(args.warmup, 3)
    nb = 8
    host_b, dev_b = [], []
    for s in range(nb):
        o, d, gt = SYN.train_batch(teng, s, rank, n_rays, H, W)
        dev_b.append((o, d, gt))
        host_b.append(tuple(t.cpu().pin_memory() for t in (o, d, gt)))
    loss_host = torch.empty((), pin_memory=True)

    def step_device(i):
        trainer.step(*dev_b[i % nb])

    def step_e2e(i):
        # pinned host batch straight into the step call (the engine stages it: 3 H2D copies), loss read back every step
        loss_host.copy_(trainer.step(*host_b[i % nb]), non_blocking=True)

    flop_step = n_rays * train_flop_per_ray()
    mark = {}
    ms_dev = timed(step_device, steps, warm, "events", after_warmup=lambda: mark.update(l0=teng.launch_count))
    launches_per_step = (teng.launch_count - mark["l0"]) / float(steps)
    ms_step = sum(ms_dev) / steps
    rec = {"workload": "lego.yaml training, %d rays/rank/step, synthetic rays+RGB, fwd+bwd+allreduce+clip/Adam (BASELINE configs[%d])" % (
               n_rays, 2 if n_rays == 1024 else 4),
           "value": n_rays * world / (ms_step / 1e3), "unit": "rays/s", "ms_per_step": ms_step, "steps": steps, "warmup": warm,
           "rays_per_rank": n_rays, "parallelism": "dp%d" % world, "scaling": "weak",
           "step_tflops_per_gpu": flop_step / (ms_step / 1e3) / 1e12, "frac_of_sustained": flop_step / (ms_step / 1e3) / 1e12 / peak_sus,
           "gpu_launches_per_step": launches_per_step, "cuda_graph": not args.no_graph,
           "l2": "flushed between timed steps (256 MiB write)"}
    # the step's second roofline: it moves ~5.2 GB of arenas through HBM per 1024 rays (ncu), next to its 900 GFLOP
    hbm_peak, hbm_kind = measured_hbm()
    step_bytes = NCU_TRAIN_STEP_DRAM_BYTES_1024 * n_rays / 1024.0
    rec["hbm"] = {"dram_bytes_per_step": step_bytes, "achieved_gbs": step_bytes / (ms_step / 1e3) / 1e9, "peak_gbs": hbm_peak,
                  "frac": step_bytes / (ms_step / 1e3) / 1e9 / hbm_peak, "peak_kind": hbm_kind + " copy bandwidth",
                  "source": "ncu dram__bytes_read.sum + dram__bytes_write.sum of the step's 8 MLP launches (profiles/train_r2z_metrics.txt), scaled by rays / 1024"}
    if not full:
        return rec
    ms_e2e = timed(step_e2e, steps, warm, "wall")
    rec["e2e"] = {"value": n_rays * world / (sum(ms_e2e) / steps / 1e3), "unit": "rays/s", "ms_per_step": sum(ms_e2e) / steps,
                  "h2d_bytes_per_step": n_rays * 36 * world, "d2h_bytes_per_step": 4 * world,
                  "api": "FusedTrainer.step on pinned host batches (nerf_forward_backward stages them: H2D inside), loss read back every step"}
    # per-kernel-class times: plain launches with event brackets around the MLP kernels (no graph while profiling)
    teng.profile(True)
    teng.profile_read()
    k_steps = 10
    step_device(0)
    teng.profile_read()
    pe = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    pe[0].record()
    for i in range(k_steps):
        step_device(i)
    pe[1].record()
    torch.cuda.synchronize()
    ms_step_plain = pe[0].elapsed_time(pe[1]) / k_steps        # plain launches with event brackets, no L2 flush in between
    cl = teng.profile_read_classes()
    teng.profile(False)
    n_c, n_f = n_rays * N_SAMPLES, n_rays * (N_SAMPLES + N_IMPORTANCE)
    flops = {"fwd": (n_c + n_f) * FLOP_PER_EVAL, "dgrad": 2 * (n_f * (593408 - 3456) + n_c * 459008), "wgrad": 2 * (n_f * 593408 + n_c * 491264)}
    names = {"fwd": "mlp_tc2_forward_kernel<SAVE>", "dgrad": "mlp_tc2_dgrad_kernel (fused dgrad chain)", "wgrad": "mlp_tc2_wgrad_kernel (+ heads)"}
    kernels = {}
    for k, (ms_k, n_k, _) in cl.items():
        if n_k > 0 and ms_k > 0:
            per_step = ms_k / k_steps
            kernels[k] = {"kernel": names[k], "ms_per_step": per_step, "launches_per_step": n_k / k_steps,
                          "achieved_tflops": flops[k] / (per_step / 1e3) / 1e12, "frac_of_sustained": flops[k] / (per_step / 1e3) / 1e12 / peak_sus}
    if "wgrad" in kernels:
        # wgrad streams every arena block it needs once (tc_arena.cuh: fine net 38 dZ + 36 X + PE + PE(dir) blocks, coarse net
        # 32 + 28 + PE, 16 KB each per 128-row tile): HBM-bound
        hbm_peak, _ = measured_hbm()
        wg_bytes = 16384.0 * (76 * n_f + 61 * n_c) / 128
        kernels["wgrad"]["achieved_gbs"] = wg_bytes / (kernels["wgrad"]["ms_per_step"] / 1e3) / 1e9
        kernels["wgrad"]["frac_of_hbm"] = kernels["wgrad"]["achieved_gbs"] / hbm_peak
    rec["kernels"] = kernels
    rec["mlp_ms_per_step"] = sum(v["ms_per_step"] for v in kernels.values())
    rec["other_ms_per_step"] = ms_step - rec["mlp_ms_per_step"]
    # everything that is not an MLP kernel (stage kernels, memset, clip/Adam/re-pack, launch gaps), measured in the SAME run as
    # the class times: an upper bound of the stages' share (plain launches leave larger gaps than the graph replay)
    rec["non_mlp_ms_per_step_plain_launches"] = ms_step_plain - rec["mlp_ms_per_step"]
    rec["non_mlp_share_plain_launches"] = (ms_step_plain - rec["mlp_ms_per_step"]) / ms_step_plain
    rec["ncu_traffic_bytes_per_launch"] = NCU_TRAFFIC_BYTES["train"]

    # ---- data-parallel checks and the all-reduce's cost (N > 1) -------------------------------------------------------------
    flat = net.flat_params
    chk = torch.stack([flat.double().sum(), flat.view(torch.int32).to(torch.int64).sum().double()])
    if world > 1:
        allc = [torch.empty_like(chk) for _ in range(world)]
        dist.all_gather(allc, chk)
        rec["param_checksum_equal_across_ranks"] = bool(all(torch.equal(a, allc[0]) for a in allc))
        rec["dp_grad_check"] = dp_grad_check(trainer, teng, dev_b[0], world, n_rays)
        g = net.flat_grad
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        for _ in range(5):
            dist.all_reduce(g)
        barrier()
        ev[0].record()
        for _ in range(20):
            dist.all_reduce(g)
        ev[1].record()
        torch.cuda.synchronize()
        ar_alone = ev[0].elapsed_time(ev[1]) / 20
        # the same step with the reduction serialised behind the backward, and with no reduction at all (ranks then diverge:
        # this leg runs last)
        trainer.reduce.overlap = False
        ms_serial = sum(timed(step_device, steps, warm, "events")) / steps
        trainer.reduce = lambda g_, e_=None: g_
        ms_none = sum(timed(step_device, steps, warm, "events")) / steps
        rec["allreduce"] = {"payload_bytes": int(g.numel() * 4), "standalone_ms": ar_alone,
                            "step_ms_overlapped": ms_step if not args.no_overlap else None, "step_ms_serialised": ms_serial,
                            "step_ms_without_allreduce": ms_none,
                            "exposed_ms_overlapped": (ms_step - ms_none) if not args.no_overlap else None,
                            "exposed_ms_serialised": ms_serial - ms_none,
                            "how": "fine-net half reduced on a side stream under the coarse backward, coarse half after it; 1/world folded into grad_scale"}
    else:
        rec["param_checksum_equal_across_ranks"] = True
    return rec


def dp_grad_check(trainer, eng, my_batch, world, n_rays):
    """grad(world ranks x n rays, all-reduced) == grad(1 rank x world*n rays) up to fp32 summation order: every rank gathers all
    batches, runs the single-process step body on the concatenation and compares with the data-parallel result."""
    import torch
    import torch.distributed as dist
    o, d, gt = my_batch
    g_dp = trainer.net.flat_grad
    trainer.forward_backward(o, d, gt)
    trainer.reduce(g_dp, eng)
    torch.cuda.synchronize()
    g_dp = g_dp.clone()
    cat = []
    for t in (o, d, gt):
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t.contiguous())
        cat.append(torch.cat(parts))
    g_one = torch.zeros_like(g_dp)
    eng.forward_backward(cat[0], cat[1], cat[2], g_one, grad_scale=1.0)
    torch.cuda.synchronize()
    half = g_dp.numel() // 2
    out = {"tol_rel": 2e-4, "what": "max |g_dp - g_single| / max |g_single| per net; g_dp = NCCL-reduced gradient of %d ranks x %d rays, "
                             "g_single = one process on the %d concatenated rays" % (world, n_rays, world * n_rays)}
    for name, sl in (("coarse_net", slice(0, half)), ("fine_net", slice(half, None))):
        scale = float(g_one[sl].abs().max())
        out[name + "_max_abs_grad"] = scale
        out[name + "_max_rel_err"] = float((g_dp[sl] - g_one[sl]).abs().max()) / max(scale, 1e-30)
    out["ok"] = bool(out["coarse_net_max_rel_err"] <= 2e-4 and out["fine_net_max_rel_err"] <= 2e-4)
    eng.forward_backward(o, d, gt, trainer.net.flat_grad, grad_scale=1.0 / world)       # leave the engine's buffers at the per-rank size
    return out


if __name__ == "__main__":
    main()
