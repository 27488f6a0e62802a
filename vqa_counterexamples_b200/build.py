"""Builds libvqacx.so in-tree with nvcc for sm_100a (no torch dependency in the library).

Every csrc/*.cu is compiled to an object under build/ (in parallel, only when it or a header changed) and the objects
are linked into vqa_counterexamples_b200/libvqacx.so. `VQACX_DEBUG_BUILD=1` (or build(debug=True)) adds -DVQACX_DEBUG:
the measurement build with the clock64() role counters and timing-experiment options compiled in (include/vqacx.h).
"""
import glob
import hashlib
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libvqacx.so")
OBJ_DIR = os.path.join(HERE, "..", "build", "obj")
NVCC_FLAGS = ["-Xcompiler", "-fPIC", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _headers():
    return sorted(glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(HERE, "..", "include", "*.h")))


def _stamp(flags):
    """Hash of everything an object depends on besides its own source: headers and flags."""
    h = hashlib.sha256(" ".join(flags).encode())
    for p in _headers():
        with open(p, "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def _stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(p) > t for p in sources() + _headers())


def build(force=False, verbose=False, debug=None):
    debug = (os.environ.get("VQACX_DEBUG_BUILD") == "1") if debug is None else debug
    if not force and not _stale() and not debug:
        return OUT
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    flags = NVCC_FLAGS + (["-DVQACX_DEBUG"] if debug else []) + (["-Xptxas", "-v"] if verbose else [])
    os.makedirs(OBJ_DIR, exist_ok=True)
    stamp = _stamp(flags)

    def compile_one(src):
        with open(src, "rb") as f:
            key = hashlib.sha256(f.read() + stamp.encode()).hexdigest()[:16]
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + "." + key + ".o")
        log = ""
        if force or not os.path.exists(obj):
            res = subprocess.run([nvcc] + flags + ["-c", src, "-o", obj + ".tmp"], capture_output=True, text=True)
            if res.returncode != 0:
                raise RuntimeError(f"nvcc failed on {src}:\n" + res.stdout + res.stderr)
            os.replace(obj + ".tmp", obj)
            log = res.stderr
        return obj, log

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as pool:
        results = list(pool.map(compile_one, sources()))
    objs = [o for o, _ in results]
    if verbose:
        print("".join(l for _, l in results))
    res = subprocess.run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a"] + objs + ["-o", OUT + ".tmp"],
                         capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + res.stdout + res.stderr)
    os.replace(OUT + ".tmp", OUT)
    keep = set(objs)
    for p in glob.glob(os.path.join(OBJ_DIR, "*.o")):      # drop objects of older revisions
        if p not in keep:
            os.remove(p)
    return OUT


if __name__ == "__main__":
    print(build(force=True, verbose=True))
