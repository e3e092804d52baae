"""In-tree build of libcelerite_b200.so (hand-written CUDA for sm_100a behind the C ABI).

    python -m celerite_b200.build          # incremental
    python -m celerite_b200.build --force

nvcc cross-compiles without a GPU; the .so is git-ignored but travels to the GPU box with the
repository snapshot.  No torch, no JIT cache: one object per .cu, compiled in parallel.
"""
import concurrent.futures
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(HERE, "libcelerite_b200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC,-O2", "-I", os.path.join(ROOT, "include")] + os.environ.get("CEL_EXTRA_NVCC_FLAGS", "").split()


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(path):
    h = hashlib.sha256()
    h.update(" ".join(FLAGS).encode())
    with open(path, "rb") as f:
        h.update(f.read())
    for hdr in sorted(os.listdir(CSRC)):
        if hdr.endswith((".h", ".cuh", ".hpp")):
            with open(os.path.join(CSRC, hdr), "rb") as f:
                h.update(f.read())
    with open(os.path.join(ROOT, "include", "celerite_b200.h"), "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def _compile(src, force):
    path = os.path.join(CSRC, src)
    obj = os.path.join(OBJ, src[:-3] + ".o")
    stamp = obj + ".sha"
    dig = _digest(path)
    if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj, False
    cmd = [NVCC] + FLAGS + ["-c", path, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    with open(stamp, "w") as f:
        f.write(dig)
    return obj, True


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    srcs = _sources()
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        res = list(ex.map(lambda s: _compile(s, force), srcs))
    objs = [o for o, _ in res]
    changed = any(c for _, c in res)
    if changed or force or not os.path.exists(LIB):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
        if verbose:
            print("linked", LIB)
    elif verbose:
        print("up to date:", LIB)
    _build_examples(verbose)
    return LIB


def _build_examples(verbose=False):
    """examples/*.cu -> build/bin/ (C++ programs over the C ABI alone; run by the GPU tests)."""
    exdir = os.path.join(ROOT, "examples")
    bindir = os.path.join(ROOT, "build", "bin")
    if not os.path.isdir(exdir):
        return
    os.makedirs(bindir, exist_ok=True)
    for src in sorted(f for f in os.listdir(exdir) if f.endswith(".cu")):
        path = os.path.join(exdir, src)
        out = os.path.join(bindir, src[:-3])
        if os.path.exists(out) and os.path.getmtime(out) >= max(os.path.getmtime(path), os.path.getmtime(LIB)):
            continue
        cmd = [NVCC, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-I", os.path.join(ROOT, "include"),
               path, "-o", out, "-L", HERE, "-lcelerite_b200", "-Xlinker", "-rpath,$ORIGIN/../../celerite_b200"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("example build failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
        if verbose:
            print("built", out)


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
