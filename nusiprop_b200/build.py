"""In-tree build of the CUDA library (sm_100a only) with nvcc.  nvcc cross-compiles, so this
works on a machine without a GPU."""
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
OUT = os.path.join(_HERE, "lib", "libnusiprop_b200.so")
SOURCES = ["nusi_capi.cu", "nusi_host.cpp"]
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    # no implicit contraction: the closed forms must round as written (explicit fma() only)
    "-fmad=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off",
    "-shared",
]


def _stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)]
    deps.append(os.path.join(_HERE, "..", "include", "nusiprop_b200.h"))
    deps.append(os.path.abspath(__file__))
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False):
    if not force and not _stale():
        return OUT
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc] + FLAGS + ["-o", OUT] + [os.path.join(CSRC, s) for s in SOURCES]
    if verbose:
        cmd += ["-Xptxas", "-v"]
        print(" ".join(cmd), file=sys.stderr)
    subprocess.check_call(cmd)
    return OUT


ROOT = os.path.dirname(_HERE)
CONSUMER = os.path.join(ROOT, "tests", "dropin", "_build", "consumer")
CYTHON_DIR = os.path.join(_HERE, "cython")


def build_dropin(force=False):
    """Host-side consumers of the C ABI: the C++ consumer of the drop-in class (tests/dropin) and the
    Cython wrapper nuSIprop.pyprop, both built in-tree so they travel to the GPU box."""
    import glob
    lib = build_library()
    libdir = os.path.dirname(lib)
    src = os.path.join(ROOT, "tests", "dropin", "consumer.cpp")
    hdr = os.path.join(_HERE, "cxx", "nuSIprop.hpp")
    deps = [src, hdr, os.path.join(ROOT, "include", "nusiprop_b200.h")]
    if force or not os.path.exists(CONSUMER) or any(os.path.getmtime(d) > os.path.getmtime(CONSUMER) for d in deps):
        os.makedirs(os.path.dirname(CONSUMER), exist_ok=True)
        subprocess.check_call([os.environ.get("CXX", "g++"), "-std=c++11", "-O2", "-I", os.path.join(_HERE, "cxx"), "-I",
                               os.path.join(ROOT, "include"), src, "-L", libdir, "-lnusiprop_b200", "-Wl,-rpath," + libdir,
                               "-o", CONSUMER])
    ext = glob.glob(os.path.join(CYTHON_DIR, "nuSIprop.*.so"))
    cy_deps = deps[1:] + [os.path.join(CYTHON_DIR, f) for f in ("nuSIprop.pyx", "nuSIprop.pxd", "setup.py")]
    if force or not ext or any(os.path.getmtime(d) > os.path.getmtime(ext[0]) for d in cy_deps):
        subprocess.check_call([sys.executable, "setup.py", "-q", "build_ext", "--inplace"], cwd=CYTHON_DIR,
                              stdout=subprocess.DEVNULL)
    return CONSUMER


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
