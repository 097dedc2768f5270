#!/usr/bin/env python
"""bench.py -- rays/s of the NeRF render (default) or train hot path on N B200s.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--mode render|train] [--impl reference]

Workload (BASELINE.json configs[1], "C2"): one synthetic lego-shaped 800x800 view, N_samples 64 +
N_importance 128, bf16 tcgen05 MLP, random-init weights (torch.manual_seed(0); Network()), perturb = 0.
A "step" is one pass of Renderer.render over one frame's worth of rays (640 000 rays / GPU).
Multi-GPU (one process per GPU, launched by torchrun): at N GPUs the job renders N views and every view's
image is cut into N row bands ("image tiles go to the GPUs"), rank r rendering band r of every view -- the
per-GPU work is constant (weak scaling) and there is no data-path collective.
`--mode train` (configs[2], "C3"): 1024 rays / rank / step, fwd + bwd + one gradient all-reduce + clip/Adam.

value   : whole-job rays/s with the rays already resident in HBM (device-timed, CUDA events, max over ranks)
e2e     : same metric through the host-buffer C-ABI call (nerf_render_rays_host): pinned host rays in,
          host rgb/depth out, copies inside the timed region
roofline: the fused MLP kernel -- algorithmic FLOPs (1 186 816 per sample evaluation, un-padded) over the
          CUDA-event duration of its launches, vs the measured dense bf16 peak (MEASURED_PEAKS.json)
cpu_baseline / --impl reference: the reference's algorithm on the host cores (oracle/nerf_oracle.py, the
          CPU restatement pinned bit-exactly to the reference -- the Python reference cannot travel to the box)
"""
import argparse
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
TRAIN_FLOP_PER_RAY = 879869952
TRAIN_RAYS = 1024


def measured_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        for k in ("hbm_gbs_sustained", "hbm_gbs", "hbm_gbps"):
            if k in d:
                return float(d[k]), "measured"
    return 6500.0, "fallback"


# dram__bytes_read.sum + dram__bytes_write.sum of one launch (the fine-net launch of each class), from the committed
# ncu --set full captures: profiles/mlp_r1j_metrics.txt (render) and profiles/train_r1d_metrics.txt (training step)
NCU_TRAFFIC_BYTES = {"render": {"fwd": 56.38e6 + 145.70e6},
                     "train": {"fwd": 0.0021e9 + 1.0045e9, "dgrad": 0.0596e9 + 0.9499e9, "wgrad": 1.9247e9 + 0.0039e9}}


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
        "impl": "reference", "metric": "rays/sec %s (lego 800x800, %d+%d samples)" % (args.mode, N_SAMPLES, N_IMPORTANCE), "value": v, "unit": "rays/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": v, "unit": "rays/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def workload_config(args, world):
    if args.mode == "train":
        return {"workload": "lego.yaml training, %d rays/rank/step, synthetic rays+RGB, fwd+bwd+allreduce+clip/Adam" % TRAIN_RAYS,
                "n_samples": N_SAMPLES, "n_importance": N_IMPORTANCE, "rays_per_rank": TRAIN_RAYS,
                "parallelism": "dp%d" % world, "l2": "flushed between timed steps (256 MiB write)"}
    extra = ""
    if args.fast_sampling is not None:
        extra = " + fast_sampling (ESS/ERT, weights_threshold %g, masked samples skipped)%s [secondary variant]" % (
            args.fast_sampling, ", fog weights" if args.fog else "")
    return {"workload": ("lego.yaml full 800x800 synthetic-camera render, %d+%d samples, bf16 MLP (BASELINE configs[%d])" % (
                N_SAMPLES, N_IMPORTANCE, 1 if (N_SAMPLES, N_IMPORTANCE) == (64, 128) else 3)) + extra,
            "H": H, "W": W, "n_samples": N_SAMPLES, "n_importance": N_IMPORTANCE, "white_bkgd": True, "perturb": 0,
            "rays_per_gpu_per_step": H * W, "views_per_step": world,
            "parallelism": "image row bands over %d GPU(s), no collective" % world,
            "l2": "flushed between timed steps (256 MiB write); per-step workspace (~3 GB) also exceeds L2"}


# ------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--mode", default="render", choices=["render", "train"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--precision", default=None, choices=["bf16", "fp32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    # secondary (non-headline) render variant: the reference's fast_sampling (ESS/ERT) knob with this weights_threshold;
    # --fog adds +1 to alpha_linear.bias (sigma ~ 1) so that the mask keeps a realistic share of the fine samples
    ap.add_argument("--fast-sampling", type=float, default=None, metavar="WEIGHTS_THRESHOLD")
    ap.add_argument("--fog", action="store_true")
    # secondary configurations of BASELINE.json: configs[3] = --samples 128 256, configs[4] = --mode train --train-rays 8192
    ap.add_argument("--samples", type=int, nargs=2, default=None, metavar=("N_SAMPLES", "N_IMPORTANCE"))
    ap.add_argument("--train-rays", type=int, default=None)
    args = ap.parse_args()
    global N_SAMPLES, N_IMPORTANCE, TRAIN_RAYS, TRAIN_FLOP_PER_RAY
    if args.samples:
        N_SAMPLES, N_IMPORTANCE = args.samples
    if args.train_rays:
        TRAIN_RAYS = args.train_rays
    # forward E evals + fine backward (wgrad + dgrad) per fine eval + coarse (sigma-only) backward per coarse eval
    TRAIN_FLOP_PER_RAY = (2 * N_SAMPLES + N_IMPORTANCE) * FLOP_PER_EVAL + (N_SAMPLES + N_IMPORTANCE) * 2366720 + N_SAMPLES * 1900544
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
        __graft_entry__.build()
    if world > 1:
        dist.barrier()

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
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- per-rank inputs -------------------------------------------------------------------------
    if args.mode == "render":
        r0, r1 = shard_range(H, rank, world)          # this rank's row band of every view
        rays_o, rays_d = [], []
        for v in range(world):
            o, d = SYN.view_rays(eng, v, max(world, 8), H, W, r0, r1)
            rays_o.append(o)
            rays_d.append(d)
        host_o = torch.cat(rays_o).cpu().pin_memory()
        host_d = torch.cat(rays_d).cpu().pin_memory()
        dev_o, dev_d = host_o.cuda(), host_d.cuda()
        n_local = dev_o.shape[0]
        host_rgb = torch.empty(n_local, 3).pin_memory()
        host_depth = torch.empty(n_local).pin_memory()
        batch = {"rays_o": dev_o[None], "rays_d": dev_d[None]}

        def step_device(i):
            with torch.no_grad():
                return renderer.render(batch)

        def step_e2e(i):
            eng.render_rays_host(host_o, host_d, host_rgb, host_depth)

        h2d, d2h = n_local * 24, n_local * 16
        rays_per_step_local = n_local
        flop_per_ray = (N_SAMPLES + N_SAMPLES + N_IMPORTANCE) * FLOP_PER_EVAL
    else:
        trainer = FusedTrainer(net, cfg=cfg)
        batches = []
        for s in range(8):
            o, d, gt = SYN.train_batch(eng, s, rank, TRAIN_RAYS, H, W)
            batches.append((o.cpu().pin_memory(), d.cpu().pin_memory(), gt.cpu().pin_memory()))
        dev_batches = [tuple(t.cuda() for t in b) for b in batches]
        loss_host = torch.empty((), pin_memory=True)

        def step_device(i):
            return trainer.step(*dev_batches[i % 8])

        def step_e2e(i):
            o, d, gt = (t.cuda(non_blocking=True) for t in batches[i % 8])
            loss_host.copy_(trainer.step(o, d, gt), non_blocking=True)
            torch.cuda.current_stream().synchronize()

        h2d, d2h = TRAIN_RAYS * 36, 4
        rays_per_step_local = TRAIN_RAYS
        flop_per_ray = TRAIN_FLOP_PER_RAY

    def timed(step_fn, use_events):
        """W warm-ups, then K steps, each bracketed by CUDA events (device) or wall clock (host-buffer path);
        L2 flushed before every step, outside the brackets.  Returns per-step ms (max over ranks)."""
        for i in range(args.warmup):
            step_fn(i)
        barrier()
        eng.profile_read()                       # reset: only the timed steps are bracketed below
        launches0 = eng.launch_count
        ms = []
        for i in range(args.steps):
            flush.fill_(i & 0xFF)
            if use_events:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                step_fn(args.warmup + i)
                e1.record()
                e1.synchronize()
                ms.append(e0.elapsed_time(e1))
            else:
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                step_fn(args.warmup + i)
                torch.cuda.synchronize()
                ms.append(1e3 * (time.perf_counter() - t0))
        barrier()
        stats = (eng.launch_count - launches0, eng.profile_read_classes())
        t = torch.tensor(ms, dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.cpu().tolist(), stats

    sampler = ClockSampler(physical_gpu_index(local_rank))
    sampler.start()
    eng.profile(True)
    ms_dev, (launches, classes) = timed(step_device, True)
    mlp_ms, mlp_launches, mlp_rows = classes["fwd"]
    eng.profile(False)
    clocks = sampler.finish()
    ms_e2e, _ = timed(step_e2e, False)

    bank_loop = None
    if args.mode == "train":
        # the same step driven by Trainer.train from images resident in HBM (ray bank): the pixel draw,
        # ray generation and colour gather run on the device, so a step has no host input at all; one loss read at the end
        import numpy as np
        from nerf_replication_b200.dataset import Dataset
        from nerf_replication_b200.optimizer import FusedAdam
        from nerf_replication_b200.trainer import Recorder, Trainer, make_data_loader
        n_views = 8
        imgs = np.random.RandomState(rank).randint(0, 256, size=(n_views, H, W, 3), dtype=np.uint8)
        poses = np.stack([SYN.lego_pose(v, n_views) for v in range(n_views)])
        ds = Dataset(net=net, images=imgs, transform_matrices=poses, camera_angle_x=SYN.CAMERA_ANGLE_X, split="train", seed=rank,
                     n_rays=TRAIN_RAYS)
        opt = FusedAdam(net, lr=5e-4, eps=1e-8, clip=40.0)
        opt.m, opt.v, opt.step_count = trainer.m, trainer.v, trainer.step_count      # continue the same optimisation state
        opt._bind_state()
        drv = Trainer(net, cfg, log_interval=10 ** 9, verbose=False)
        drv.train(0, make_data_loader(cfg, ds, True, world > 1, args.warmup, rank, world), opt, Recorder())
        barrier()
        t0 = time.perf_counter()
        losses = drv.train(1, make_data_loader(cfg, ds, True, world > 1, args.steps, rank, world), opt, Recorder())
        last_loss = float(losses[-1])
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        bank_loop = {"value": TRAIN_RAYS * world * args.steps / float(dt), "unit": "rays/s", "ms_per_step": 1e3 * float(dt) / args.steps,
                     "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 4.0 / args.steps, "last_loss": last_loss,
                     "api": "Trainer.train: nerf_bank_train_step (device pixel draw + rays + colours + fwd + loss + bwd) + "
                            "FusedAdam.step; %d synthetic %dx%d views resident in HBM; no L2 flush between steps" % (n_views, H, W)}

    total_ms = sum(ms_dev)
    rays_job = rays_per_step_local * world
    value = rays_job * args.steps / (total_ms / 1e3)
    e2e_value = rays_job * args.steps / (sum(ms_e2e) / 1e3)

    peak_sus, peak_burst, peak_kind = measured_peaks()
    roofline = None
    fwd_name = "mlp_tc_forward_kernel" if os.environ.get("NERF_TC_VARIANT") == "1" else "mlp_tc2_forward_kernel (CTA pair, tcgen05 cta_group::2)"
    if precision == "bf16" and mlp_launches > 0 and mlp_ms > 0 and args.fast_sampling is None:   # rows actually evaluated are only known on the device when fast_sampling skips rows
        # algorithmic FLOPs per class: forward 1 186 816 / row; backward per row: fine dgrad 2*(593408-3456),
        # fine wgrad 2*593408, coarse (sigma-only) dgrad 2*459008, coarse wgrad 2*491264
        n_c, n_f = rays_per_step_local * N_SAMPLES, rays_per_step_local * (N_SAMPLES + N_IMPORTANCE)
        flops = {"fwd": mlp_rows * FLOP_PER_EVAL,
                 "dgrad": args.steps * 2 * (n_f * (593408 - 3456) + n_c * 459008),
                 "wgrad": args.steps * 2 * (n_f * 593408 + n_c * 491264)}
        names = {"fwd": fwd_name, "dgrad": "mlp_tc2_dgrad_kernel (fused dgrad chain, CTA pair)", "wgrad": "mlp_tc2_wgrad_kernel (+ heads)"}
        kernels = {}
        for k, (ms_k, n_k, rows_k) in classes.items():
            if n_k > 0 and ms_k > 0:
                kernels[k] = {"kernel": names[k], "ms_total": ms_k, "launches": n_k, "avg_launch_ms": ms_k / n_k,
                              "achieved_tflops": flops[k] / (ms_k / 1e3) / 1e12, "share_of_step": ms_k / total_ms}
        if "wgrad" in kernels:
            # wgrad is HBM-bound: its algorithmic bytes are every arena block it needs, once (tc_arena.cuh: fine net 38 dZ +
            # 36 X + PE + PE(dir) blocks, coarse net 32 + 28 + PE, 16 KB each per 128-row tile)
            wg_bytes = args.steps * 16384.0 * (76 * n_f + 61 * n_c) / 128
            kernels["wgrad"]["achieved_gbs"] = wg_bytes / (kernels["wgrad"]["ms_total"] / 1e3) / 1e9
        top = max(kernels, key=lambda k: kernels[k]["ms_total"])
        traffic = NCU_TRAFFIC_BYTES[args.mode].get(top)
        if top == "wgrad":
            hbm_peak, hbm_kind = measured_hbm()
            achieved = kernels[top]["achieved_gbs"]
            roofline = {"bound": "hbm", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                        "traffic": traffic, "kernel": names[top], "peak_kind": "%s HBM copy bandwidth" % hbm_kind}
        else:
            achieved = kernels[top]["achieved_tflops"]
            roofline = {"bound": "tensor", "achieved": achieved, "peak": peak_sus, "unit": "TFLOP/s", "frac": achieved / peak_sus,
                        "traffic": traffic, "kernel": names[top],
                        "peak_kind": "%s sustained bf16 (burst %.1f)" % (peak_kind, peak_burst)}
        roofline.update({"traffic_unit": "bytes per launch (ncu dram read + write, fine-net launch)",
                         "launches_timed": kernels[top]["launches"], "avg_launch_ms": kernels[top]["avg_launch_ms"],
                         "share_of_step": kernels[top]["share_of_step"], "kernels": kernels,
                         "step_tflops": rays_per_step_local * flop_per_ray / (total_ms / args.steps / 1e3) / 1e12})

    line = {
        "metric": "rays/sec %s (lego 800x800, %d+%d samples)" % (args.mode, N_SAMPLES, N_IMPORTANCE), "value": value, "unit": "rays/s",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": precision, "data": "synthetic",
        "config": workload_config(args, world),
        "e2e": {"value": e2e_value, "unit": "rays/s", "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
                "ms_per_step": sum(ms_e2e) / args.steps, "api": "nerf_render_rays_host (C ABI, host buffers)" if args.mode == "render" else "FusedTrainer.step with pinned host batches"},
        "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
    }
    if rank == 0 and world == 1 and not args.no_cpu_baseline and args.mode == "render":
        threads = os.cpu_count() or 1
        v, n_done, dt = cpu_render_rate(12.0, threads)
        line["cpu_baseline"] = {"value": v, "unit": "rays/s", "cores": threads, "kind": "port",
                                "sample": "%d rays (2048-ray slices from the centre rows of the same 800x800 view), %.1f s, oracle/nerf_oracle.py CPU fp32 torch" % (n_done, dt)}
    if bank_loop is not None:
        line["bank_loop"] = bank_loop
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
