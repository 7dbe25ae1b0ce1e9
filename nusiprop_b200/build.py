"""In-tree build of the CUDA library (sm_100a only) with nvcc.  nvcc cross-compiles, so this
works on a machine without a GPU."""
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
OUT = os.path.join(_HERE, "lib", "libnusiprop_b200.so")
# Kernel translation units: they are independent of each other, so they compile in parallel.
SOURCES = ["nusi_capi.cu", "k_rows.cu", "k_alpha.cu", "k_alpha_check.cu", "k_integrate.cu", "k_integrate_split.cu", "k_integrate_cluster.cu", "k_integrate_check.cu", "k_misc.cu", "k_phiphi.cu",
           "nusi_host.cpp"]
OBJ_DIR = os.path.join(_HERE, "lib", "build")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    # no implicit contraction: the closed forms must round as written (explicit fma() only)
    "-fmad=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off",
]


def _deps(path, seen=None):
    """The file and every header it includes (transitively, quoted includes only)."""
    import re
    seen = set() if seen is None else seen
    path = os.path.normpath(path)
    if path in seen or not os.path.exists(path):
        return seen
    seen.add(path)
    for inc in re.findall(r'^\s*#\s*include\s+"([^"]+)"', open(path).read(), re.M):
        _deps(os.path.join(os.path.dirname(path), inc), seen)
    return seen


def build_library(force=False, verbose=False, extra_flags=(), out=None):
    """Compile every translation unit whose object is older than its source or any header, then link.
    extra_flags / out: alternative builds of the same sources (tests: -DNUSI_LOCKED_MATH; tuning experiments)."""
    out = out or OUT
    obj_dir = OBJ_DIR if out == OUT else os.path.join(os.path.dirname(out), "build_" + os.path.basename(out))
    os.makedirs(obj_dir, exist_ok=True)
    nvcc = os.environ.get("NVCC", "nvcc")
    self_t = os.path.getmtime(os.path.abspath(__file__))
    jobs, objs = [], []
    for src in SOURCES:
        sp = os.path.join(CSRC, src)
        ob = os.path.join(obj_dir, os.path.splitext(src)[0] + ".o")
        objs.append(ob)
        if force or not os.path.exists(ob) or os.path.getmtime(ob) < max([self_t] + [os.path.getmtime(d) for d in _deps(sp)]):
            cmd = [nvcc] + FLAGS + list(extra_flags) + ["-c", "-o", ob, sp]
            if verbose:
                cmd += ["-Xptxas", "-v"]
                print(" ".join(cmd), file=sys.stderr)
            jobs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = []
    for src, pr in jobs:
        log = pr.communicate()[0]
        if verbose or pr.returncode != 0:
            sys.stderr.write(log)
        if pr.returncode != 0:
            failed.append(src)
    if failed:
        raise RuntimeError("nvcc failed for: " + ", ".join(failed))
    if jobs or not os.path.exists(out):
        subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out] + objs)
    return out


LOCKED_OUT = os.path.join(_HERE, "lib", "libnusiprop_b200_locked.so")


def build_locked_library(force=False):
    """TEST build of the same sources with -DNUSI_LOCKED_MATH: every elementary function from csrc/nusi_locked_math.h, the
    counterpart of oracle/_ref/libnusi_ref_locked.so (tests/test_locked_parity.py).  Never loaded by the package."""
    return build_library(force=force, extra_flags=["-DNUSI_LOCKED_MATH"], out=LOCKED_OUT)


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
