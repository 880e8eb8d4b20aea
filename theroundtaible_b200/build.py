"""Build the sm_100a shared library in-tree (theroundtaible_b200/_lib/librt_engine.so).

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels to the GPU box with
the gpurun snapshot.  Used by __graft_entry__.build() and by `python -m theroundtaible_b200.build`.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "_lib")
LIB = os.path.join(LIB_DIR, "librt_engine.so")
SOURCES = ["engine.cu", "capi.cu", "noalloc.cu", "prefill.cu", "tp.cu", "tokenizer.cpp", "safetensors.cpp"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC,-fvisibility=hidden"]
LINK_FLAGS = ["-shared", "-ldl"]


def _fingerprint() -> str:
    h = hashlib.sha256()
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for fn in sorted(os.listdir(root)):
            if fn.endswith((".cu", ".cuh", ".cpp", ".h")):
                with open(os.path.join(root, fn), "rb") as f:
                    h.update(fn.encode() + b"\0" + f.read())
    h.update((" ".join(NVCC_FLAGS) + os.environ.get("RT_NVCC_EXTRA", "")).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False, trace: bool = False) -> str:
    """``trace``: the profiling variant librt_engine_trace.so (-DRT_TRACE: every kernel records globaltimer stamps, see
    scripts/trace_timeline.py); load it with RT_LIB=<path>.  The product library is never built with it."""
    os.makedirs(LIB_DIR, exist_ok=True)
    tag = "_trace" if trace else ("_" + os.environ["RT_BUILD_TAG"] if os.environ.get("RT_BUILD_TAG") else "")   # experiments: RT_BUILD_TAG + RT_NVCC_EXTRA
    LIB = os.path.join(LIB_DIR, f"librt_engine{tag}.so")
    stamp = os.path.join(LIB_DIR, f"librt_engine{tag}.stamp")
    fp = _fingerprint()
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read().strip() == fp:
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    extra = os.environ.get("RT_NVCC_EXTRA", "").split()      # experiments only (e.g. -DRT_ATTN_STAGES=3)
    if trace:
        extra.append("-DRT_TRACE")
    # one nvcc per translation unit, in parallel, then one link
    from concurrent.futures import ThreadPoolExecutor
    obj_dir = os.path.join(LIB_DIR, "obj" + tag)
    os.makedirs(obj_dir, exist_ok=True)
    vflags = ["-Xptxas", "-v"] if verbose else []

    def compile_one(src):
        obj = os.path.join(obj_dir, os.path.splitext(src)[0] + ".o")
        r = subprocess.run([nvcc, *NVCC_FLAGS, *vflags, *extra, "-c", "-o", obj, os.path.join(CSRC, src)],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n" + r.stdout + r.stderr)
        return obj, r.stderr

    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 4)) as pool:
        results = list(pool.map(compile_one, SOURCES))
    if verbose:
        for _, err in results:
            print(err)
    r = subprocess.run([nvcc, *NVCC_FLAGS, *LINK_FLAGS, "-o", LIB, *[o for o, _ in results]], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + r.stdout + r.stderr)
    with open(stamp, "w") as f:
        f.write(fp)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, trace="--trace" in sys.argv))
