"""Build the UNMODIFIED reference Cython sources into oracle/_ref/ (test infrastructure only).

Recipe, not product code.  It compiles, from where they lie under /root/reference/src:

  * wfpt.pyx (+ integrate.pxi, pdf.pxi via its own `include` chain)  -> oracle/_ref/wfpt*.so
  * cddm_data_simulation.pyx                                         -> oracle/_ref/cddm_data_simulation*.so

No reference source is copied into the repository: Cython reads the .pyx in place and writes its generated
C/C++ into oracle/_ref/build/ (git-ignored).  `wfpt.pyx:16` does an unused `import hddm`; the real package cannot
import here (pymc/kabuki/ssms absent), so an EMPTY stand-in package `oracle/_ref/hddm/__init__.py` is generated
next to the built module — the reference source itself stays untouched.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load what this
produces.  On the GPU box /root/reference does not exist; the prebuilt .so files travel with the snapshot.
"""
from __future__ import annotations

import glob
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference/src"
OUT = os.path.join(HERE, "_ref")


CDF_LIB = os.path.join(OUT, "libcdfdif_ref.so")


def build_cdf_ref(force: bool = False) -> bool:
    """The reference's plain-C CDF (src/cdfdif.c) compiled as it lies into oracle/_ref/libcdfdif_ref.so."""
    import subprocess

    if os.path.exists(CDF_LIB) and not force:
        return True
    src = os.path.join(REF_SRC, "cdfdif.c")
    if not os.path.exists(src):
        return False
    os.makedirs(OUT, exist_ok=True)
    subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-o", CDF_LIB, src, "-lm"], check=True)
    return True


def load_cdf_ref():
    """ctypes handle on the reference cdfdif(double t, int x, double* par, double* prob), or ImportError."""
    import ctypes as C

    if not os.path.exists(CDF_LIB):
        raise ImportError("oracle/_ref/libcdfdif_ref.so is not built")
    L = C.CDLL(CDF_LIB)
    L.cdfdif.restype = C.c_double
    L.cdfdif.argtypes = [C.c_double, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    return L


def ref_built() -> bool:
    return bool(glob.glob(os.path.join(OUT, "wfpt*.so"))) and bool(
        glob.glob(os.path.join(OUT, "cddm_data_simulation*.so"))
    )


def build_ref(force: bool = False, verbose: bool = False) -> bool:
    """Returns True when oracle/_ref holds both modules (built now or earlier)."""
    if os.path.isdir(REF_SRC):
        build_cdf_ref(force)
    if ref_built() and not force:
        return True
    if not os.path.isdir(REF_SRC):
        return False
    import numpy as np
    from Cython.Build import cythonize
    from setuptools import Distribution, Extension
    from setuptools.command.build_ext import build_ext

    os.makedirs(OUT, exist_ok=True)
    build_dir = os.path.join(OUT, "build")
    stub = os.path.join(OUT, "hddm")
    os.makedirs(stub, exist_ok=True)
    with open(os.path.join(stub, "__init__.py"), "w") as fh:
        fh.write("# empty stand-in so that the reference's unused `import hddm` (src/wfpt.pyx:16) succeeds\n")

    exts = [
        Extension(
            "wfpt",
            [os.path.join(REF_SRC, "wfpt.pyx")],
            language="c++",
            include_dirs=[np.get_include()],
            extra_compile_args=["-O2", "-w"],
            define_macros=[("NPY_NO_DEPRECATED_API", "NPY_1_7_API_VERSION")],
        ),
        Extension(
            "cddm_data_simulation",
            [os.path.join(REF_SRC, "cddm_data_simulation.pyx")],
            include_dirs=[np.get_include()],
            extra_compile_args=["-O2", "-w"],
            define_macros=[("NPY_NO_DEPRECATED_API", "NPY_1_7_API_VERSION")],
        ),
    ]
    cwd = os.getcwd()
    os.chdir(OUT)
    try:
        ext_modules = cythonize(
            exts,
            build_dir=build_dir,
            compiler_directives={"language_level": "3"},
            quiet=not verbose,
            force=force,
        )
        dist = Distribution({"name": "hddm_ref_oracle", "ext_modules": ext_modules})
        cmd = build_ext(dist)
        cmd.build_lib = OUT
        cmd.build_temp = os.path.join(build_dir, "tmp")
        cmd.inplace = False
        cmd.ensure_finalized()
        cmd.run()
    finally:
        os.chdir(cwd)
    return ref_built()


def load_ref():
    """Import (wfpt_ref, cddm_ref) from oracle/_ref, or raise ImportError."""
    if not ref_built():
        raise ImportError("oracle/_ref is not built (run `python oracle/build_ref.py` where /root/reference exists)")
    import importlib.util

    mods = []
    names = ("wfpt", "cddm_data_simulation", "hddm")
    saved = {n: sys.modules.get(n) for n in names}
    if OUT not in sys.path:
        sys.path.insert(0, OUT)  # only so that `import hddm` finds the empty stand-in
    try:
        for name in ("wfpt", "cddm_data_simulation"):
            path = glob.glob(os.path.join(OUT, name + "*.so"))[0]
            spec = importlib.util.spec_from_file_location(name, path)
            mod = importlib.util.module_from_spec(spec)
            # do NOT register under sys.modules["wfpt"]: the product ships its own top-level `wfpt`
            spec.loader.exec_module(mod)
            mods.append(mod)
    finally:
        try:
            sys.path.remove(OUT)
        except ValueError:
            pass
        # Cython's module init registers itself in sys.modules; take the reference modules (and the empty
        # `hddm` stand-in) out again so that they can never shadow the product's own top-level `wfpt`.
        for name in names:
            if saved[name] is not None:
                sys.modules[name] = saved[name]
            else:
                sys.modules.pop(name, None)
    return tuple(mods)


if __name__ == "__main__":
    ok = build_ref(force="--force" in sys.argv, verbose=True)
    print("oracle/_ref built:", ok)
    sys.exit(0 if ok else 1)
