"""Where do the CTA-pair kernels wait?  python profiles/waitprof.py  (GPU box)"""
import sys, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")

from nerf_replication_b200.engine import Engine
from oracle import nerf_oracle as O

NAMES = {"fwd": {0x1: "producer w_empty", 0x5: "forwarder w_full", 0x2: "issuer w_full", 0x3: "issuer act_ready", 0x4: "epi acc_full",
                 0x90: "save leader bulk_wait_read", 0x91: "save before_overwrite total", 0x92: "save named barrier"},
         "bwd": {0x1: "dgrad producer w_empty", 0x5: "dgrad forwarder w_full", 0x2: "dgrad issuer w_full", 0x3: "dgrad issuer act_ready",
                 0x4: "dgrad epi acc_full", 0x81: "wgrad producer empty", 0x85: "wgrad forwarder full", 0x83: "wgrad issuer tmem_free",
                 0x82: "wgrad issuer full", 0x84: "wgrad flush acc_done"}}
for b in range(0x60, 0x70): NAMES["bwd"][b] = "dgrad epi mask full"
for b in range(0x70, 0x80): NAMES["bwd"][b] = "dgrad mask producer m_empty"

def main():
    dev = "cuda:0"
    eng = Engine(device=0, precision="bf16", n_samples=64, n_importance=128, perturb=True, white_bkgd=True)
    sd = O.fog_state_dict(seed=0)
    flat = torch.cat([sd[k].reshape(-1) for k in O.state_dict_keys()]).float().to(dev).contiguous()
    eng.set_params(flat)
    R = 1024
    g = torch.Generator().manual_seed(1)
    o = torch.randn(R, 3, generator=g).mul(0.1).add(torch.tensor([0., 0., 4.])).to(dev)
    d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) * 0.2 + torch.tensor([0., 0., -1.]), dim=-1).to(dev)
    gt = torch.rand(R, 3, generator=g).to(dev)
    grad = torch.zeros_like(flat)
    mode = sys.argv[1] if len(sys.argv) > 1 else "train"
    eng.profile(True)
    for it in range(3):
        if it == 2:
            eng.profile_read_classes(); eng.waitprof(True)
        if mode == "train": eng.forward_backward(o, d, gt, grad)
        else: eng.render_rays(o, d)
    torch.cuda.synchronize()
    res = eng.waitprof(False)
    print("mode", mode, {k: "%.3f ms / %d launches" % (v[0], v[1]) for k, v in eng.profile_read_classes().items()})
    for tu in ("fwd", "bwd"):
        agg = {}
        for b, (cyc, n) in res[tu].items():
            k = NAMES[tu].get(b, hex(b))
            a = agg.setdefault(k, [0, 0]); a[0] += cyc; a[1] += n
        print("== %s (cycles summed over waiting warps, one 1024-ray step)" % tu)
        for k, (cyc, n) in sorted(agg.items(), key=lambda kv: -kv[1][0]):
            print("  %-32s %14d cyc  %9d waits  %8.0f cyc/wait  %8.1f kcyc/SM" % (k, cyc, n, cyc / max(n, 1), cyc / 148e3))

if __name__ == "__main__":
    main()
