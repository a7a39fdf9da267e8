"""Build the C-ABI CUDA library in-tree (vits_b200/libvitsb200.so) with nvcc for sm_100a.

Each .cu is compiled to an object under vits_b200/build/ (in parallel, re-compiled only when it or a header is
newer) and the objects are linked into the shared library.
"""
from __future__ import annotations

import concurrent.futures
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libvitsb200.so")
SOURCES = ["conv_f32.cu", "aux_kernels.cu", "conv_tc.cu", "rb_fused.cu", "flow_fused.cu", "engine_weights.cu", "engine_run.cu",
           "capi.cu", "noalloc.cu", "p1.cu", "resample.cu"]
NVCC_FLAGS = ["-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-Xcompiler", "-fPIC"]
LINK_FLAGS = ["-shared", "-cudart", "static"]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _headers():
    return [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))] + \
           [os.path.join(HERE, "..", "include", "vitsb200.h")]


def _sources():
    return [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "vitsb200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, lib: str = LIB, extra_flags=()) -> str:
    """lib / extra_flags: a differently built library next to the product one (e.g. libvitsb200_dbg.so with
    -DVITSB_RB_DEBUG: timing-experiment hooks and clock64 traces; select it at run time with VITSB_LIB)."""
    if lib == LIB and not force and not needs_build():
        return LIB
    extra = os.environ.get("VITSB_EXTRA_NVCC", "").split() + list(extra_flags)
    tag = hashlib.sha1(" ".join(NVCC_FLAGS + extra).encode()).hexdigest()[:8]
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdr_t = max(os.path.getmtime(h) for h in _headers())
    jobs = []
    objs = []
    for s in _sources():
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, f"{os.path.splitext(s)[0]}.{tag}.o")
        objs.append(obj)
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(src), hdr_t):
            jobs.append((s, [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]))

    def run(job):
        name, cmd = job
        r = subprocess.run(cmd, capture_output=True, text=True)
        return name, r

    with concurrent.futures.ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as ex:
        for name, r in ex.map(run, jobs):
            if r.returncode != 0:
                raise RuntimeError(f"nvcc failed on {name}:\n" + r.stdout + r.stderr)
            if verbose:
                print(r.stderr)
    r = subprocess.run([nvcc] + NVCC_FLAGS + LINK_FLAGS + ["-o", lib] + objs, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + r.stdout + r.stderr)
    return lib


if __name__ == "__main__":
    if "--debug" in sys.argv:
        print(build(lib=os.path.join(HERE, "libvitsb200_dbg.so"), extra_flags=["-DVITSB_RB_DEBUG"]))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
