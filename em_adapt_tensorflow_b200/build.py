"""Build libweaklabel_cuda.so in-tree with nvcc for sm_100a (B200).

    python -m em_adapt_tensorflow_b200.build [--force] [--verbose]

Outputs (git-ignored, but they travel to the GPU box with the repo snapshot):
    em_adapt_tensorflow_b200/lib/libweaklabel_cuda.so
    em_adapt_tensorflow_b200/lib/dropin/{libweaklabel.so, libestep.so,
        estep/libweaklabel.so, pythonlib/estep/libweaklabel.so}  -> symlinks under the names
        the reference's ctypes callers look for (estep.py:9-12, estep/estep.py:11,
        estep/estep_with_ctype.py:16)
"""
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
LIBDIR = os.path.join(PKG, "lib")
LIB = os.path.join(LIBDIR, "libweaklabel_cuda.so")
SOURCES = ["estep_kernel.cu", "estep_onchip.cu", "softmax_ce.cu", "label_set.cu", "stream_kernels.cu", "api.cu"]
HEADERS = ["estep_kernel.cuh", "estep_device.cuh", "seqsum.cuh", "bulk_tile.cuh", os.path.join(ROOT, "include", "weaklabel_cuda.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "--fmad=false",
              "-shared", "-cudart", "static"]
DROPIN_NAMES = ["libweaklabel.so", "libestep.so", os.path.join("estep", "libweaklabel.so"),
                os.path.join("pythonlib", "estep", "libweaklabel.so")]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libweaklabel_cuda.so cannot be built (there is no CPU fallback)")


def _stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, s) for s in SOURCES] + \
           [h if os.path.isabs(h) else os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def _link(path, target):
    """Create/refresh a symlink atomically; several ranks may do this at the same time."""
    try:
        if os.path.islink(path) and os.readlink(path) == target:
            return
        tmp = "%s.%d.tmp" % (path, os.getpid())
        if os.path.lexists(tmp):
            os.remove(tmp)
        os.symlink(target, tmp)
        os.replace(tmp, path)
    except OSError:
        if not os.path.lexists(path):
            raise


def build(force=False, verbose=False):
    """Compile (if stale) and return the path of libweaklabel_cuda.so.  Safe to call from several
    processes at once (one rank per GPU): the compile is serialised by a lock file."""
    os.makedirs(LIBDIR, exist_ok=True)
    if force or _stale():
        import fcntl
        with open(os.path.join(LIBDIR, ".build.lock"), "w") as lock:
            fcntl.flock(lock, fcntl.LOCK_EX)
            try:
                if force or _stale():               # another process may have built it meanwhile
                    tmp = "%s.%d.tmp" % (LIB, os.getpid())
                    # development knob: extra nvcc flags (e.g. -DWL_SELPROF, -DWL_TRAIL); never set for a product build
                    extra = os.environ.get("WEAKLABEL_NVCC_EXTRA", "").split()
                    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + \
                          ["-o", tmp] + [os.path.join(CSRC, s) for s in SOURCES]
                    if verbose:
                        print(" ".join(cmd))
                    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
                    if verbose or res.returncode != 0:
                        sys.stderr.write(res.stdout)
                    if res.returncode != 0:
                        if os.path.exists(tmp):
                            os.remove(tmp)
                        raise RuntimeError("nvcc failed building libweaklabel_cuda.so")
                    os.replace(tmp, LIB)
            finally:
                fcntl.flock(lock, fcntl.LOCK_UN)
    drop = os.path.join(LIBDIR, "dropin")
    for name in DROPIN_NAMES:
        path = os.path.join(drop, name)
        os.makedirs(os.path.dirname(path), exist_ok=True)
        _link(path, os.path.relpath(LIB, os.path.dirname(path)))
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
