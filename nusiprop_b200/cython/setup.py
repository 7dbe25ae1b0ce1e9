"""Build the Cython wrapper in place:  python setup.py build_ext --inplace
Replaces the reference's setup.py (hard-coded macOS paths, reference setup.py:7-16): no GSL, no
polylogarithm -- the extension links only libnusiprop_b200.so (built by nusiprop_b200.build)."""
import os

import numpy as np
from Cython.Build import cythonize
from setuptools import Extension, setup

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
LIBDIR = os.path.join(PKG, "lib")

ext = Extension(
    "nuSIprop",
    sources=[os.path.join(HERE, "nuSIprop.pyx")],
    language="c++",
    include_dirs=[os.path.join(PKG, "cxx"), os.path.join(PKG, "..", "include"), np.get_include()],
    library_dirs=[LIBDIR],
    libraries=["nusiprop_b200"],
    runtime_library_dirs=[LIBDIR],
    define_macros=[("NUSIPROP_B200_EXCEPTIONS", "1"), ("NPY_NO_DEPRECATED_API", "NPY_1_7_API_VERSION")],
    extra_compile_args=["-std=c++11", "-O2"],
)
setup(name="nuSIprop", ext_modules=cythonize([ext], compiler_directives={"language_level": 3}, build_dir=os.path.join(HERE, "build")))
