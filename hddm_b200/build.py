"""Builds hddm_b200/libwfpt_b200.so in-tree with nvcc for sm_100a (no torch, no JIT cache).

    python -m hddm_b200.build [--force]

Each .cu is compiled to an object under hddm_b200/csrc/_obj/ and linked with `nvcc -shared`.  Files listed in
NO_FMAD are compiled with --fmad=false (the operation-for-operation "faithful" evaluator needs separately rounded
products and sums; see csrc/wfpt_faithful.cuh).  The CUDA runtime is linked statically so that the library loads
without torch.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libwfpt_b200.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

SOURCES = ["api.cu", "cdf.cu", "ktable.cpp", "like_faithful.cu", "like_fast.cu", "like_lean.cu", "noalloc.cu", "sim.cu"]
NO_FMAD = {"like_faithful.cu", "cdf.cu"}
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "-I", INCLUDE, "-I", CSRC]


def _deps(src: str) -> list[str]:
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs += [os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)]
    return [src, os.path.abspath(__file__)] + hdrs


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    srcs = [s for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    jobs = []
    objs = []
    for s in srcs:
        src = os.path.join(CSRC, s)
        obj = os.path.join(OBJ, os.path.splitext(s)[0] + ".o")
        objs.append(obj)
        if force or _stale(obj, _deps(src)):
            cmd = [nvcc] + ARCH + COMMON + (["--fmad=false"] if s in NO_FMAD else []) + ["-c", src, "-o", obj]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
        if verbose:
            print(r.stdout + r.stderr, flush=True)

    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 1)) as ex:
            list(ex.map(run, jobs))
    if jobs or force or _stale(LIB, objs):
        run([nvcc] + ARCH + ["-shared", "-o", LIB] + objs + ["-cudart", "static"])
    # the CPython binding of the hot call (optional: hddm_b200.wfpt falls back to ctypes when it is missing)
    import sysconfig

    ext = os.path.join(HERE, "_wfptcall" + (sysconfig.get_config_var("EXT_SUFFIX") or ".so"))
    src = os.path.join(CSRC, "pywfpt.c")
    if force or _stale(ext, [src, LIB, os.path.join(INCLUDE, "wfpt_b200.h")]):
        run(["gcc", "-O2", "-shared", "-fPIC", "-I", sysconfig.get_paths()["include"], src, "-o", ext, "-L", HERE,
             "-l:libwfpt_b200.so", "-Wl,-rpath,$ORIGIN"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
