"""Pure-read / pure-write / copy DRAM bandwidth of the box (torch kernels, CUDA events, best of 5, 4 GiB buffers >> L2):
what is the roofline of a kernel that only WRITES (the training kernels' arenas)?"""
import torch

def best(fn, n=5):
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)

N = 1 << 30                       # float32 elements: 4 GiB
a = torch.empty(N, dtype=torch.float32, device="cuda")
b = torch.empty(N, dtype=torch.float32, device="cuda")
a.zero_(); b.zero_()
gb = N * 4 / 1e9
print("fill_  (write only)      %.0f GB/s" % (gb / best(lambda: a.fill_(1.0)) * 1e3))
print("zero_  (memset)          %.0f GB/s" % (gb / best(lambda: a.zero_()) * 1e3))
print("sum    (read only)       %.0f GB/s" % (gb / best(lambda: a.sum()) * 1e3))
print("copy_  (read + write)    %.0f GB/s (sum of both directions)" % (2 * gb / best(lambda: b.copy_(a)) * 1e3))
print("add_   (2 reads + write) %.0f GB/s (sum)" % (3 * gb / best(lambda: torch.add(a, b, out=a)) * 1e3))
