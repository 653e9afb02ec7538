"""Build libvcn_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

Replaces the reference's JIT plugin loader (stylegan2_ada/torch_utils/custom_ops.py:46-124): the
library is built ahead of time into ``viscoin-next_b200/csrc/libvcn_b200.so`` so that it travels
with the tree; objects are cached by source hash so rebuilds are incremental.
"""

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
LIB = os.path.join(CSRC, "libvcn_b200.so")
OBJ_DIR = os.path.join(CSRC, "build")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-I", os.path.join(ROOT, "include"),
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libvcn_b200 cannot be built")


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(path):
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    for dep in [path] + sorted(
        os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))
    ) + [os.path.join(ROOT, "include", "vcn_b200.h")]:
        with open(dep, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()[:16]


def _compile(src, verbose):
    os.makedirs(OBJ_DIR, exist_ok=True)
    obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + "." + _digest(src) + ".o")
    if not os.path.exists(obj):
        cmd = [_nvcc()] + NVCC_FLAGS + ["-c", src, "-o", obj]
        proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        log = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".ptxas.log")
        with open(log, "w") as fh:
            fh.write(proc.stdout)
        if proc.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, proc.stdout))
        if verbose:
            print("compiled", os.path.basename(src))
    return obj


def build(verbose=False, force=False):
    """Compile every .cu under csrc/ and link libvcn_b200.so.  Returns the library path."""
    if force and os.path.isdir(OBJ_DIR):
        shutil.rmtree(OBJ_DIR)
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose), srcs))
    stamp = os.path.join(OBJ_DIR, "link.stamp")
    want = " ".join(objs)
    have = open(stamp).read() if os.path.exists(stamp) else ""
    if want != have or not os.path.exists(LIB):
        cmd = [_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart"]
        proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if proc.returncode != 0:
            raise RuntimeError("link failed:\n" + proc.stdout)
        with open(stamp, "w") as fh:
            fh.write(want)
        # drop stale objects
        keep = set(objs)
        for f in os.listdir(OBJ_DIR):
            p = os.path.join(OBJ_DIR, f)
            if f.endswith(".o") and p not in keep:
                os.remove(p)
        if verbose:
            print("linked", LIB)
    return LIB


if __name__ == "__main__":
    print(build(verbose=True, force="--force" in sys.argv))
