"""The C-ABI boundary: libnerf_b200.so builds for sm_100a, loads, and exports exactly the entry points
include/nerf_b200.h declares.  No compute calls (CPU-only container)."""
import ctypes
import os
import re
import subprocess

import pytest

from conftest import ROOT

HEADER = os.path.join(ROOT, "include", "nerf_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nerf_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_hot_path():
    syms = declared_symbols()
    for need in ("nerf_create", "nerf_destroy", "nerf_set_params", "nerf_render_rays", "nerf_render_camera",
                 "nerf_render_rays_host", "nerf_mlp_forward", "nerf_composite", "nerf_sample_pdf", "nerf_get_rays",
                 "nerf_forward_backward", "nerf_clip_adam", "nerf_last_error"):
        assert need in syms


def test_library_exports_every_declared_symbol(built_lib):
    lib = ctypes.CDLL(built_lib)
    for s in declared_symbols():
        assert hasattr(lib, s), "libnerf_b200.so does not export %s" % s


def test_ctypes_signatures_cover_header(built_lib):
    from nerf_replication_b200 import _lib
    assert sorted(_lib.SIGNATURES) == declared_symbols()
    lib = _lib.load()
    assert lib.nerf_version() >= 100


def test_probe_library_is_separate_from_the_product(built_lib):
    """The bring-up probe battery has its own header and library; the product .so does not export it."""
    from nerf_replication_b200 import _lib
    probe_header = open(os.path.join(ROOT, "include", "nerf_b200_probe.h")).read()
    probe_header = re.sub(r"/\*.*?\*/", "", probe_header, flags=re.S)
    assert sorted(set(re.findall(r"\b(nerf_[a-z0-9_]+)\s*\(", probe_header))) == sorted(_lib.PROBE_SIGNATURES)
    probe = _lib.load_probe()
    assert hasattr(probe, "nerf_debug_umma_probe")
    assert not hasattr(ctypes.CDLL(built_lib), "nerf_debug_umma_probe")


def test_header_is_plain_c(tmp_path):
    """No torch / C++ types in the boundary: the header must compile as C."""
    c = tmp_path / "t.c"
    c.write_text('#include "nerf_b200.h"\n#include "nerf_b200_probe.h"\nint main(void){NerfConfig c; (void)c; return NERF_PARAMS_TOTAL == 2*NERF_PARAMS_PER_NET ? 0 : 1;}\n')
    exe = tmp_path / "t"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)])
    assert subprocess.call([str(exe)]) == 0


def test_sass_has_blackwell_tensor_and_tma_instructions(built_lib):
    """The fused MLP must be tcgen05/TMEM + bulk-copy TMA code, not a legacy mma.sync path."""
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    sass = subprocess.run([cuobjdump, "-sass", built_lib], capture_output=True, text=True).stdout
    assert "UTCHMMA" in sass, "no tcgen05.mma (UTCHMMA) in SASS"
    assert "LDTM" in sass, "no tcgen05.ld (LDTM) in SASS"
    assert "UBLKCP" in sass or "UTMALDG" in sass, "no TMA bulk copy in SASS"
    assert "HMMA." not in sass.replace("UTCHMMA", ""), "legacy HMMA found"


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from nerf_replication_b200.engine import Engine, EngineError
    with pytest.raises(EngineError):
        Engine()
    from nerf_replication_b200 import _lib
    lib = _lib.load()
    cfg = _lib.NerfConfig(64, 128, 1, 0, 0, 0, 2.0, 6.0, 0)
    h = ctypes.c_void_p()
    assert lib.nerf_create(ctypes.byref(cfg), 0, ctypes.byref(h)) != 0
    assert b"no CPU fallback" in lib.nerf_last_error(None)
