"""Builds the engine in-tree with nvcc for sm_100a (no torch extension machinery: plain C-ABI shared objects; Python binds
them with ctypes).  `python nerf-replication_b200/build.py [--force]`.

  libnerf_b200.so        the product library (include/nerf_b200.h)
  libnerf_b200_probe.so  the tcgen05 descriptor probe battery of the bring-up (include/nerf_b200_probe.h; tests only)

`build()` reports on stdout whether each library was COMPILED by this call or REUSED (sources + flags hash unchanged), so a
driver log shows which of the two happened on a given box.  NERF_B200_REBUILD=1 forces a recompile.
"""
import concurrent.futures
import hashlib
import os
import subprocess
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libnerf_b200.so")
OUT_PROBE = os.path.join(HERE, "libnerf_b200_probe.so")
OBJ = os.path.join(HERE, "build")
SOURCES = ["api.cu", "stages.cu", "mlp_fp32.cu", "pack_weights.cu", "mlp_tc2.cu", "mlp_tc2_bwd.cu", "adam.cu", "bank.cu"]
PROBE_SOURCES = ["probe.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-Wno-deprecated-gpu-targets", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]
last_build = {}      # library file name -> "compiled" | "reused" (what the last build() call did)


def _nvcc():
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if c and (os.path.isabs(c) and os.path.exists(c) or not os.path.isabs(c)):
            return c
    raise RuntimeError("nvcc not found")


def _digest(sources):
    """Hash of everything a library is built from: its sources, every header next to them, the public headers, the flags."""
    h = hashlib.sha256()
    for f in sorted(os.listdir(CSRC)):
        if f in sources or not f.endswith(".cu"):
            with open(os.path.join(CSRC, f), "rb") as fh:
                h.update(f.encode() + fh.read())
    inc = os.path.join(os.path.dirname(HERE), "include")
    for f in sorted(os.listdir(inc)):
        with open(os.path.join(inc, f), "rb") as fh:
            h.update(f.encode() + fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def up_to_date(out=OUT, sources=SOURCES):
    stamp = out + ".stamp"
    return os.path.exists(out) and os.path.exists(stamp) and open(stamp).read() == _digest(sources)


def build_experiment(bits, verbose=False, defines=(), tag=None):
    """Diagnostic flavour libnerf_b200_exp<bits>[_<tag>].so (-DNERF_EXP=bits plus `defines`: timing experiments, some of which
    break results on purpose, see csrc/tc_common.cuh); select it with NERF_B200_LIB=<path>.  Never the product library."""
    name = "exp%d%s" % (bits, "_" + tag if tag else "")
    return _build(OUT.replace(".so", "_%s.so" % name), os.path.join(HERE, "build_" + name), SOURCES,
                  ["-DNERF_EXP=%d" % bits] + ["-D" + d for d in defines], verbose)


def build_trace(verbose=False):
    """Diagnostic flavour libnerf_b200_trace.so (-DNERF_TC_WAITPROF -DNERF_TC_TRACE): CTA 0 of the CTA-pair kernels logs
    time-stamped events (profiles/trace_step.py)."""
    return _build(OUT.replace(".so", "_trace.so"), os.path.join(HERE, "build_trace"), SOURCES, ["-DNERF_TC_WAITPROF", "-DNERF_TC_TRACE"], verbose)


def build(force=False, verbose=False, waitprof=False):
    """waitprof=True builds the diagnostic flavour libnerf_b200_waitprof.so (-DNERF_TC_WAITPROF: every barrier wait of the
    CTA-pair kernels is timed); select it with NERF_B200_LIB=<path>.  The product library never carries that code."""
    if waitprof:
        return _build(OUT.replace(".so", "_waitprof.so"), os.path.join(HERE, "build_waitprof"), SOURCES, ["-DNERF_TC_WAITPROF"], verbose)
    force = force or os.environ.get("NERF_B200_REBUILD") == "1"
    for out, sources in ((OUT, SOURCES), (OUT_PROBE, PROBE_SOURCES)):
        name = os.path.basename(out)
        if not force and up_to_date(out, sources):
            last_build[name] = "reused"
            print("[nerf_b200 build] %s: REUSED (sources + flags hash %s... unchanged)" % (name, _digest(sources)[:12]), flush=True)
            continue
        t0 = time.time()
        _build(out, OBJ, sources, [], verbose)
        with open(out + ".stamp", "w") as f:
            f.write(_digest(sources))
        last_build[name] = "compiled"
        print("[nerf_b200 build] %s: COMPILED %d sources with nvcc for sm_100a in %.1f s" % (name, len(sources), time.time() - t0), flush=True)
    return OUT


def _build(out, objdir, sources, extra, verbose):
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

    with concurrent.futures.ThreadPoolExecutor(max_workers=len(sources)) as ex:
        objs = list(ex.map(compile_one, sources))
    r = subprocess.run([nvcc, "-shared", "-Wno-deprecated-gpu-targets", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs + ["-lcudart"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    return out


if __name__ == "__main__":
    if "--trace" in sys.argv:
        print(build_trace(verbose=True))
        sys.exit(0)
    if "--exp" in sys.argv:
        defs = [sys.argv[i + 1] for i, a in enumerate(sys.argv) if a == "--define"]
        tag = sys.argv[sys.argv.index("--tag") + 1] if "--tag" in sys.argv else None
        print(build_experiment(int(sys.argv[sys.argv.index("--exp") + 1]), verbose=True, defines=defs, tag=tag))
        sys.exit(0)
    print(build(force="--force" in sys.argv, verbose=True, waitprof="--waitprof" in sys.argv))
