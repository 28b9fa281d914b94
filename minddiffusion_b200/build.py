"""Build libmdx_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m minddiffusion_b200.build [--force]

The shared library lands next to this file (git-ignored, but it travels to the GPU box with the snapshot).
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libmdx_b200.so")
SOURCES = ["abi.cu", "gemm_tc.cu", "conv_direct.cu", "norm.cu", "attention.cu", "attention_tc.cu", "elementwise.cu", "rowops.cu", "sampler.cu",
           "glide_ops.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "--use_fast_math" if os.environ.get("MDX_FAST_MATH") else "-DMDX_PRECISE", "-I" + os.path.join(HERE, "..", "include")] + os.environ.get("MDX_NVCC_EXTRA", "").split()


def _digest():
    """Hash of the sources and the compile flags — with the checkout's own path taken out, so that the library built
    in one checkout is recognised as current in a copy of it somewhere else (the GPU box)."""
    h = hashlib.sha256()
    names = sorted(os.listdir(CSRC)) + ["../../include/mdx.h"]
    for name in names:
        path = os.path.join(CSRC, name)
        if os.path.isfile(path):
            h.update(name.encode())
            h.update(open(path, "rb").read())
    h.update(" ".join(FLAGS).replace(HERE, "<pkg>").encode())
    return h.hexdigest()


def build(force=False, verbose=False):
    stamp = os.path.join(OBJ, "stamp")
    digest = _digest()

    def current():
        return os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == digest

    if not force and current():
        return LIB
    if not os.path.exists(NVCC):
        if os.path.exists(LIB):
            return LIB  # GPU box without a toolkit: use the prebuilt library that travelled with the repo
        raise RuntimeError("nvcc not found at %s and no prebuilt %s" % (NVCC, LIB))
    os.makedirs(OBJ, exist_ok=True)
    # one builder at a time: the ranks of a multi-GPU launch all come through here, and a rank must never dlopen a
    # library another rank is still linking.  The lock is advisory (flock); whoever gets it second finds the work done.
    import fcntl
    lock = open(os.path.join(OBJ, "lock"), "w")
    fcntl.flock(lock, fcntl.LOCK_EX)
    try:
        if not force and current():
            return LIB
        return _build_locked(stamp, digest, verbose)
    finally:
        fcntl.flock(lock, fcntl.LOCK_UN)
        lock.close()


def _build_locked(stamp, digest, verbose):

    def compile_one(src):
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    tmp = LIB + ".tmp.%d" % os.getpid()
    cmd = [NVCC, "-shared", "-o", tmp] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    os.replace(tmp, LIB)   # atomic: a reader sees the old library or the new one, never a partial file
    with open(stamp, "w") as f:
        f.write(digest)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
