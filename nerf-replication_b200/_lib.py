"""ctypes binding of libnerf_b200.so -- the thin C-ABI boundary (include/nerf_b200.h).

The library is built in-tree by build.py.  There is no fallback: if it is missing, or there is no
sm_100 GPU, loading / ``nerf_create`` raise."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
# NERF_B200_LIB selects another build of the same library (the -DNERF_TC_WAITPROF diagnostic flavour); never a fallback
LIB_PATH = os.environ.get("NERF_B200_LIB") or os.path.join(HERE, "libnerf_b200.so")

NERF_PARAMS_PER_NET = 595844
NERF_PARAMS_TOTAL = 1191688
PRECISION_BF16, PRECISION_FP32 = 0, 1
NET_COARSE, NET_FINE = 0, 1


class NerfConfig(C.Structure):
    _fields_ = [("n_samples", C.c_int32), ("n_importance", C.c_int32), ("white_bkgd", C.c_int32),
                ("perturb", C.c_int32), ("train_u", C.c_int32), ("precision", C.c_int32),
                ("t_near", C.c_float), ("t_far", C.c_float), ("seed", C.c_uint64),
                ("fast_sampling", C.c_int32), ("weights_threshold", C.c_float)]


class NerfRenderOut(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("rgb", "depth", "acc", "rgb_coarse", "depth_coarse", "acc_coarse")]


_P, _I, _L, _F, _D = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_double
# name -> (restype, argtypes); mirrors include/nerf_b200.h one to one
SIGNATURES = {
    "nerf_version": (_I, []),
    "nerf_create": (_I, [C.POINTER(NerfConfig), _I, C.POINTER(_P)]),
    "nerf_destroy": (_I, [_P]),
    "nerf_last_error": (C.c_char_p, [_P]),
    "nerf_set_config": (_I, [_P, C.POINTER(NerfConfig)]),
    "nerf_set_params": (_I, [_P, _P, _P]),
    "nerf_launch_count": (_L, [_P]),
    "nerf_profile_enable": (_I, [_P, _I]),
    "nerf_profile_read": (_I, [_P, C.POINTER(_D), C.POINTER(_L), C.POINTER(_L)]),
    "nerf_profile_read_classes": (_I, [_P, C.POINTER(_D), C.POINTER(_L), C.POINTER(_L)]),
    "nerf_get_rays": (_I, [_P, C.POINTER(_D), _I, _I, _D, _D, _D, _P, _L, _L, _P, _P, _P]),
    "nerf_stratified_t": (_I, [_P, _L, _I, _P, _P, _P]),
    "nerf_freq_encode": (_I, [_P, _P, _L, _I, _P, _P]),
    "nerf_mlp_forward": (_I, [_P, _I, _P, _P, _P, _L, _I, _P, _P]),
    "nerf_composite": (_I, [_P, _P, _P, _L, _I, _I, _P, _P, _P, _P, _P]),
    "nerf_sample_pdf": (_I, [_P, _P, _P, _L, _I, _I, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "nerf_sample_pdf_prune": (_I, [_P, _P, _P, _L, _I, _I, _P, _P, _F, _P, _P, _P, _P, _P, _P]),
    "nerf_render_rays": (_I, [_P, _P, _P, _L, _P, _P, C.POINTER(NerfRenderOut), _P]),
    "nerf_render_camera": (_I, [_P, C.POINTER(_D), _I, _I, _D, _D, _D, _L, _L, C.POINTER(NerfRenderOut), _P]),
    "nerf_render_rays_host": (_I, [_P, _P, _P, _L, _P, _P]),
    "nerf_forward_backward": (_I, [_P, _P, _P, _P, _L, _P, _P, _F, _P, _P, _P, _P]),
    "nerf_train_forward": (_I, [_P, _P, _P, _L, _P, _P, _P, _P, C.POINTER(C.c_uint64), _P]),
    "nerf_train_backward": (_I, [_P, _P, _P, _L, _P, _P, _P, _I, C.c_uint64, _P]),
    "nerf_stream_wait_fine_grads": (_I, [_P, _P]),
    "nerf_set_graph_mode": (_I, [_P, _I]),
    "nerf_clip_adam": (_I, [_P, _P, _P, _P, _P, _F, _F, _F, _F, _F, _I, _P]),
    "nerf_bank_create": (_I, [_P, _I, _I, _I, _D, _D, _D, _I, C.POINTER(_I)]),
    "nerf_bank_destroy": (_I, [_P, _I]),
    "nerf_bank_set_view": (_I, [_P, _I, _I, _P, C.POINTER(_D), _P]),
    "nerf_bank_sample": (_I, [_P, _I, _I, _I, _P, _L, C.c_uint64, _L, _P, _P, _P, _P, _P]),
    "nerf_bank_permutation_host": (_I, [C.c_uint64, C.c_uint64, _L, _L, _P]),
    "nerf_bank_train_step": (_I, [_P, _I, _I, _I, _P, _L, C.c_uint64, _L, _P, _P, _F, _P, _P, _P, _P]),
    "nerf_eval_image": (_I, [_P, _P, _P, _L, _P, _P, _P]),
    "nerf_eval_image_static": (_I, [_P, _P, _P, _L, _P, _P, _P]),
    "nerf_eval_ssim": (_I, [_P, _P, _P, _I, _I, _I, _P, _P]),
    "nerf_debug_tc_watchdog": (_I, [C.POINTER(C.c_uint)]),
    "nerf_debug_waitprof": (_I, [_I, C.POINTER(C.c_ulonglong)]),
    "nerf_debug_trace": (_I, [C.POINTER(C.c_ulonglong), _I]),
}

# include/nerf_b200_probe.h: the bring-up probe battery, a separate library (not part of the product .so)
PROBE_LIB_PATH = os.path.join(HERE, "libnerf_b200_probe.so")
PROBE_SIGNATURES = {"nerf_debug_umma_probe": (_I, [_I, _P, _P, _P, _I, _I, _I, _P])}

_lib = None
_probe = None


def load():
    """Loads (and prototypes) the shared library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError("libnerf_b200.so is missing (%s): run `python nerf-replication_b200/build.py` or "
                           "`__graft_entry__.build()`; there is no Python/CPU fallback" % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)       # AttributeError if the .so does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    _lib = lib
    return lib


def load_probe():
    """Loads libnerf_b200_probe.so (tests only)."""
    global _probe
    if _probe is None:
        if not os.path.exists(PROBE_LIB_PATH):
            raise RuntimeError("libnerf_b200_probe.so is missing (%s): run `python nerf-replication_b200/build.py`" % PROBE_LIB_PATH)
        lib = C.CDLL(PROBE_LIB_PATH)
        for name, (res, args) in PROBE_SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _probe = lib
    return _probe
