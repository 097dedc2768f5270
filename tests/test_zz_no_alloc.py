"""nerf_eval_image_static (include/nerf_b200.h): the allocation-free variant of nerf_eval_image -- no counterpart in the
reference.  Its reduction scratch is a file-scope device array, so the call
performs no cudaMalloc.  Checked through the C-ABI (ctypes, preallocated outputs) against nerf_eval_image on two calls in a
row: the uint8 sum bit-exact (integer work), the fp32-term double sum to the evaluator tests' 1e-6 relative tolerance -- and
in fact to the same bits, since both entry points launch the same kernel on the same grid.

Collected last on purpose: a failure here hides nothing.  Runs alone or after the other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ELEMS = 1 << 27          # NERF_EVAL_STATIC_MAX_ELEMS
NERF_E_INVALID = -1


def _images(seed, n):
    rng = np.random.RandomState(seed)
    gt = rng.uniform(0, 1, size=n).astype(np.float32)
    pred = np.clip(gt + rng.normal(0, 0.2, size=n), -0.5, 1.5).astype(np.float32)
    return torch.from_numpy(pred).cuda(), torch.from_numpy(gt).cuda()


def _call(eng, fn, pred, gt, sse, sq8):
    rc = fn(eng.h, C.c_void_p(pred.data_ptr()), C.c_void_p(gt.data_ptr()), pred.numel(), C.c_void_p(sse.data_ptr()),
            C.c_void_p(sq8.data_ptr()), eng._stream())
    assert rc == 0, eng.lib.nerf_last_error(eng.h).decode()


def test_static_scratch_entry_point_matches_eval_image():
    from nerf_replication_b200.engine import Engine
    eng = Engine(precision="fp32")
    lib = eng.lib
    # two calls in a row on different images of different sizes: a full 800x800 frame (grid = 4 blocks per SM), then a
    # ragged one (not a multiple of 4 floats, fewer blocks) that must not see the first call's partials or counter
    cases = [_images(11, 800 * 800 * 3), _images(12, 100 * 100 * 3 + 1), _images(13, 800 * 800 * 3)]
    want = []
    for pred, gt in cases:
        sse, sq8 = torch.zeros((), dtype=torch.float64, device="cuda"), torch.zeros((), dtype=torch.int64, device="cuda")
        _call(eng, lib.nerf_eval_image, pred, gt, sse, sq8)
        torch.cuda.synchronize()
        want.append((float(sse), int(sq8)))
    sse, sq8 = torch.zeros((), dtype=torch.float64, device="cuda"), torch.zeros((), dtype=torch.int64, device="cuda")
    _call(eng, lib.nerf_eval_image_static, *cases[0], sse, sq8)      # first call on this device: resolves the symbol address
    torch.cuda.synchronize()
    res_sse = torch.zeros(2 * len(cases), dtype=torch.float64, device="cuda")        # outputs preallocated: nothing below allocates
    res_sq8 = torch.zeros(2 * len(cases), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for i in range(2 * len(cases)):                                   # back to back on one stream, no sync in between
        pred, gt = cases[i % len(cases)]
        _call(eng, lib.nerf_eval_image_static, pred, gt, res_sse[i], res_sq8[i])
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0                      # no cudaMalloc happened (device free memory did not shrink)
    got = list(zip(res_sse.cpu(), res_sq8.cpu()))
    for i, (a, b) in enumerate(got):
        w_sse, w_sq8 = want[i % len(cases)]
        assert int(b) == w_sq8, (i, int(b), w_sq8)                    # integer work: bit-exact
        assert abs(float(a) - w_sse) <= 1e-6 * w_sse, (i, float(a), w_sse)
        assert float(a) == w_sse, (i, float(a), w_sse)               # same kernel, same grid, same fold order
    # and against numpy's uint8 arithmetic directly (the reference's psnr_metric, src/evaluators/nerf.py:24-26)
    pred, gt = cases[1]
    p8 = (np.clip(pred.cpu().numpy(), 0, 1) * 255).astype(np.uint8)
    g8 = (np.clip(gt.cpu().numpy(), 0, 1) * 255).astype(np.uint8)
    assert want[1][1] == int((((p8 - g8) ** 2).astype(np.uint8)).astype(np.uint64).sum())
    eng.close()


def test_static_scratch_entry_point_rejects_oversize():
    from nerf_replication_b200.engine import Engine
    eng = Engine(precision="fp32")
    x = torch.zeros(16, device="cuda")
    sse, sq8 = torch.zeros((), dtype=torch.float64, device="cuda"), torch.zeros((), dtype=torch.int64, device="cuda")
    args = (eng.h, C.c_void_p(x.data_ptr()), C.c_void_p(x.data_ptr()))
    out = (C.c_void_p(sse.data_ptr()), C.c_void_p(sq8.data_ptr()), eng._stream())
    assert eng.lib.nerf_eval_image_static(*args, MAX_ELEMS + 1, *out) == NERF_E_INVALID      # rejected before any launch
    assert "NERF_EVAL_STATIC_MAX_ELEMS" in eng.lib.nerf_last_error(eng.h).decode()
    assert eng.lib.nerf_eval_image_static(*args, 0, *out) == NERF_E_INVALID
    assert eng.lib.nerf_eval_image_static(*args, 16, *out) == 0                              # and the handle still works
    torch.cuda.synchronize()
    assert float(sse) == 0.0 and int(sq8) == 0
    eng.close()
