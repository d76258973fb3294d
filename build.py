#!/usr/bin/env python
"""Build the native library in-tree:  python build.py [-j N] [--force]

Compiles every csrc/*.cu for sm_100a (nvcc cross-compiles without a GPU) and links
pytorch_distributed_tutorials_b200/lib/libb200ddp.so.  The .so is git-ignored but travels to the GPU
box with the gpurun snapshot.  No torch headers are needed: the library is a plain C ABI (include/b200ddp.h).
"""
import argparse
import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(ROOT, "csrc")
OBJ = os.path.join(ROOT, "build", "obj")
LIBDIR = os.path.join(ROOT, "pytorch_distributed_tutorials_b200", "lib")
LIB = os.path.join(LIBDIR, "libb200ddp.so")

NVCC = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
# --use_fast_math (flush-to-zero, approximate division / rsqrt / exp / log) only where it is harmless: the tensor-core
# conv / stem kernels, whose arithmetic is bf16 MMAs with fp32 accumulation. BatchNorm statistics and finalize, the
# SGD / all-reduce epilogue and the cross-entropy keep IEEE fp32 semantics (denormal gradients and momentum survive).
FAST_MATH_SOURCES = {"conv_igemm.cu", "stem.cu"}
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
         "-I" + os.path.join(ROOT, "include"), "-I" + CSRC, "--expt-relaxed-constexpr",
         "-diag-suppress", "177"] + os.environ.get("B200DDP_NVCC_FLAGS", "").split()


def _flags_for(src):
    return FLAGS + (["--use_fast_math"] if os.path.basename(src) in FAST_MATH_SOURCES else [])


def _deps_hash(src):
    h = hashlib.sha1()
    for p in [src] + sorted(
            os.path.join(d, f) for d in (CSRC, os.path.join(ROOT, "include")) for f in os.listdir(d)
            if f.endswith((".h", ".cuh"))):
        with open(p, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(ARCH + _flags_for(src)).encode())
    return h.hexdigest()


def _compile(src, force, verbose):
    obj = os.path.join(OBJ, os.path.basename(src) + ".o")
    stamp = obj + ".sha1"
    want = _deps_hash(src)
    if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == want:
        return obj, "", False
    cmd = [NVCC] + ARCH + _flags_for(src) + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    with open(stamp, "w") as fh:
        fh.write(want)
    return obj, r.stderr, True


def build(jobs=None, force=False, verbose=False, quiet=False):
    """Compile and link under an inter-process lock: with `torchrun --nproc_per_node=8` on a fresh tree all ranks
    reach the auto-build at once; the first one builds, the others wait and find everything up to date."""
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
    import fcntl
    with open(os.path.join(ROOT, "build", ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            return _build_locked(jobs, force, verbose, quiet)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)


def _build_locked(jobs, force, verbose, quiet):
    srcs = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cu", ".cpp")))
    jobs = jobs or min(len(srcs), os.cpu_count() or 4)
    objs, rebuilt = [], False
    with cf.ThreadPoolExecutor(jobs) as ex:
        for obj, log, did in ex.map(lambda s: _compile(s, force, verbose), srcs):
            objs.append(obj)
            rebuilt |= did
            if verbose and log:
                print(log)
    if rebuilt or not os.path.exists(LIB):
        tmp = LIB + ".tmp.%d" % os.getpid()
        cmd = [NVCC] + ARCH + ["-shared", "-o", tmp] + objs + ["-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
        os.replace(tmp, LIB)      # never expose a half-written library to a process that is about to dlopen it
    if not quiet:
        print("built %s (%d sources, %s)" % (os.path.relpath(LIB, ROOT), len(srcs),
                                             "rebuilt" if rebuilt else "up to date"))
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("-j", type=int, default=None)
    ap.add_argument("--force", action="store_true")
    ap.add_argument("-v", "--verbose", action="store_true", help="pass -Xptxas -v (registers / spills / smem)")
    a = ap.parse_args()
    try:
        build(a.j, a.force, a.verbose)
    except RuntimeError as e:
        print(e, file=sys.stderr)
        sys.exit(1)
