"""Builds libnerf_b200.so in-tree with nvcc for sm_100a (no torch extension machinery: the library is a
plain C-ABI shared object; Python binds it with ctypes).  `python nerf-replication_b200/build.py`."""
import concurrent.futures
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libnerf_b200.so")
OBJ = os.path.join(HERE, "build")
SOURCES = ["api.cu", "stages.cu", "mlp_fp32.cu", "mlp_tc.cu", "mlp_tc2.cu", "mlp_tc2_bwd.cu", "adam.cu", "bank.cu", "probe.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-Wno-deprecated-gpu-targets", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    raise RuntimeError("nvcc not found")


def _digest():
    h = hashlib.sha256()
    for root in (CSRC, os.path.join(os.path.dirname(HERE), "include")):
        for f in sorted(os.listdir(root)):
            with open(os.path.join(root, f), "rb") as fh:
                h.update(f.encode() + fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def up_to_date():
    stamp = OUT + ".stamp"
    return os.path.exists(OUT) and os.path.exists(stamp) and open(stamp).read() == _digest()


def build(force=False, verbose=False, waitprof=False):
    """waitprof=True builds the diagnostic flavour libnerf_b200_waitprof.so (-DNERF_TC_WAITPROF: every barrier wait of the
    CTA-pair kernels is timed); select it with NERF_B200_LIB=<path>.  The product library never carries that code."""
    if waitprof:
        return _build(OUT.replace(".so", "_waitprof.so"), os.path.join(HERE, "build_waitprof"), ["-DNERF_TC_WAITPROF"], verbose, None)
    if not force and up_to_date():
        return OUT
    return _build(OUT, OBJ, [], verbose, OUT + ".stamp")


def _build(out, objdir, extra, verbose, stamp):
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()

    def compile_one(src):
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        cmd = [nvcc] + NVCC_FLAGS + extra + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose and r.stderr.strip():
            print(r.stderr, file=sys.stderr)
        return obj

    with concurrent.futures.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    r = subprocess.run([nvcc, "-shared", "-Wno-deprecated-gpu-targets", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs + ["-lcudart"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    if stamp:
        with open(stamp, "w") as f:
            f.write(_digest())
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True, waitprof="--waitprof" in sys.argv))
