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
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "--use_fast_math", "-Xcompiler", "-fPIC,-Wall,-Wno-unused-function",
         "-I" + os.path.join(ROOT, "include"), "-I" + CSRC, "--expt-relaxed-constexpr",
         "-diag-suppress", "177"] + os.environ.get("B200DDP_NVCC_FLAGS", "").split()


def _deps_hash(src):
    h = hashlib.sha1()
    for p in [src] + sorted(
            os.path.join(d, f) for d in (CSRC, os.path.join(ROOT, "include")) for f in os.listdir(d)
            if f.endswith((".h", ".cuh"))):
        with open(p, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(ARCH + FLAGS).encode())
    return h.hexdigest()


def _compile(src, force, verbose):
    obj = os.path.join(OBJ, os.path.basename(src) + ".o")
    stamp = obj + ".sha1"
    want = _deps_hash(src)
    if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == want:
        return obj, "", False
    cmd = [NVCC] + ARCH + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    with open(stamp, "w") as fh:
        fh.write(want)
    return obj, r.stderr, True


def build(jobs=None, force=False, verbose=False, quiet=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(LIBDIR, exist_ok=True)
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
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-lpthread", "-ldl", "-lrt"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
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
