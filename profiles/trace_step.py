"""Time line of the saving forward's CTA 0 (diagnostic build: python nerf-replication_b200/build.py --trace;
NERF_B200_LIB=.../libnerf_b200_trace.so python profiles/trace_step.py).  Events: issuer (tile, layer): 1 act_ready seen,
2 first weight stage seen, 3 last MMA + commit issued; epilogue warp 0 of a tile: 1 accumulator seen, 2 epilogue done
(before the hand-over), 3 hand-over signalled."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from nerf_replication_b200.engine import Engine
from oracle import nerf_oracle as O


def main():
    eng = Engine(device=0, precision="bf16")
    sd = O.fog_state_dict(seed=0)
    flat = torch.cat([sd[k].reshape(-1) for k in O.state_dict_keys()]).float().cuda().contiguous()
    eng.set_params(flat)
    eng.set_graph_mode(False)
    o, d, gt, _, _ = O.synthetic_train_batch(0, 0, 1024, 800, 800)
    o, d, gt = o.cuda(), d.cuda(), gt.cuda()
    grad = torch.zeros_like(flat)
    for it in range(3):
        if it == 2:
            eng.waitprof(True)
        eng.forward_backward(o, d, gt, grad)
    torch.cuda.synchronize()
    out = (C.c_ulonglong * 8192)()
    eng._check(eng.lib.nerf_debug_trace(out, 8192))
    ev = sorted((int(v) >> 16, int(v) & 0xFFFF) for v in list(out)[1:8001] if int(v))
    print("records", len(ev))
    if not ev:
        return
    t0 = ev[0][0]
    role = {1: "issuer", 2: "epi", 3: "producer"}
    names = {1: {1: "act_ready seen", 2: "bias slab seen", 3: "tile-layer issued + commit", 4: "stage 0 seen", 5: "stage 1 seen", 6: "stage 2 seen",
                 7: "stage 3 seen", 8: "stage 4 seen"},
             2: {1: "accumulator seen", 2: "epilogue done", 3: "hand-over signalled"},
             3: {4: "slot free for stage 0", 5: "slot free for stage 1", 6: "slot free for stage 2", 7: "slot free for stage 3", 8: "slot free for stage 4",
                 9: "load of stage 0 issued", 10: "load of stage 1 issued", 11: "load of stage 2 issued", 12: "load of stage 3 issued", 13: "load of stage 4 issued"}}
    last = {}
    lo, hi = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (0, 400)
    for t, tag in ev[lo:hi]:
        r, tile, li, e = tag >> 12, (tag >> 8) & 1, (tag >> 4) & 15, tag & 15
        key = (r, tile if r != 3 else 0)
        dt = t - last.get(key, t)
        last[key] = t
        print("%9d  +%6d  %-8s %s layer %d  %s" % (t - t0, dt, role.get(r, "?"), ("tile " + "XY"[tile]) if r != 3 else ("pass %d" % tile), li,
                                                  names.get(r, {}).get(e, str(e))))


if __name__ == "__main__":
    main()
