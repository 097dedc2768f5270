import sys, torch
sys.path.insert(0, ".")
from nerf_replication_b200.engine import Engine
eng = Engine(precision="bf16")
for variant in (4, 5, 6):
    for N, K in ((64, 16), (256, 64), (128, 32)):
        g = torch.Generator().manual_seed(N + K)
        at = torch.randn(K, 128, generator=g).bfloat16().float()
        bt = torch.randn(K, N, generator=g).bfloat16().float()
        d = eng.umma_probe(variant, at.cuda(), bt.cuda())
        torch.cuda.synchronize()
        ref = at.double().T @ bt.double()
        print("variant", variant, "N", N, "K", K, "maxabs err", float((d.cpu().double() - ref).abs().max()), flush=True)
