"""ctypes binding of the C ABI in include/nusiprop_b200.h (libnusiprop_b200.so, built in-tree
by nusiprop_b200.build).  Nothing here computes: every function forwards to the CUDA library and
raises if it is missing -- there is no CPU fallback."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# NUSIPROP_B200_LIB: an alternative build of the SAME library (kernel tuning experiments, tools/gpu_variants.sh)
LIB_PATH = os.environ.get("NUSIPROP_B200_LIB") or os.path.join(_HERE, "lib", "libnusiprop_b200.so")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class NusiOpts(C.Structure):
    _fields_ = [
        ("majorana", C.c_int),
        ("non_resonant", C.c_int),
        ("normal_ordering", C.c_int),
        ("N_bins_E", C.c_int),
        ("lEmin", C.c_double),
        ("lEmax", C.c_double),
        ("zmax", C.c_double),
        ("flav", C.c_int),
        ("phiphi", C.c_int),
        ("alphatilde_phiphi_path", C.c_char_p),
        ("alpha_phiphi_path", C.c_char_p),
        ("device", C.c_int),
        ("max_points_per_launch", C.c_int),
        ("elementary_mode", C.c_int),
        ("source_model", C.c_int),
        ("si", C.c_double),
        ("norm", C.c_double),
        # implementation switches (ABI 3, host_copy ABI 4): cross-check kernels and tuning variants, all default 0
        ("integrator", C.c_int),
        ("fine_variant", C.c_int),
        ("builder_per_entry", C.c_int),
        ("group_max", C.c_int),
        ("group_always", C.c_int),
        ("rows_split", C.c_int),
        ("rows_ungrouped", C.c_int),
        ("keep_order", C.c_int),
        ("two_stream", C.c_int),
        ("source_device", C.c_int),
        ("host_copy", C.c_int),
    ]


SWITCHES = ("integrator", "fine_variant", "builder_per_entry", "group_max", "group_always", "rows_split", "rows_ungrouped", "keep_order",
            "two_stream", "source_device", "host_copy")


# name -> (restype, argtypes); also the list tests check against the header
SIGNATURES = {
    "nusi_abi_version": (C.c_int, []),
    "nusi_strerror": (C.c_char_p, [C.c_int]),
    "nusi_last_error": (C.c_char_p, []),
    "nusi_opts_default": (C.c_int, [C.POINTER(NusiOpts)]),
    "nusi_ctx_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(NusiOpts)]),
    "nusi_ctx_destroy": (None, [C.c_void_p]),
    "nusi_ctx_query": (C.c_int, [C.c_void_p, c_int_p, c_int_p, c_int_p, c_double_p]),
    "nusi_ctx_energies": (C.c_int, [C.c_void_p, c_double_p, c_double_p, c_double_p]),
    "nusi_ctx_redshifts": (C.c_int, [C.c_void_p, c_double_p]),
    "nusi_ctx_pmns2": (C.c_int, [C.c_void_p, c_double_p]),
    "nusi_ctx_source": (C.c_int, [C.c_void_p, c_double_p]),
    "nusi_ctx_set_powerlaw": (C.c_int, [C.c_void_p, C.c_double, C.c_double]),
    "nusi_ctx_freestream_energy": (C.c_int, [C.c_void_p, c_double_p]),
    "nusi_energy_integral": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, c_double_p, C.c_void_p]),
    "nusi_energy_integral_static": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, c_double_p, C.c_void_p]),
    "nusi_last_grouping": (C.c_int, [C.c_void_p, c_int_p, c_int_p, c_int_p]),
    "nusi_grid": (C.c_int, [C.c_int, C.c_double, C.c_double, C.c_double, c_int_p, c_double_p, c_double_p, c_double_p, c_double_p]),
    "nusi_freestream_energy": (C.c_int, [C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, c_double_p]),
    "nusi_powerlaw_source": (C.c_int, [C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, c_double_p]),
    "nusi_phiphi_tables_generate": (C.c_int, [C.c_int, C.c_char_p, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, c_double_p]),
    "nusi_masses": (C.c_int, [C.c_double, C.c_int, c_double_p]),
    "nusi_evolve_batch": (C.c_int, [C.c_void_p, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_int_p]),
    "nusi_points_create": (C.c_int, [C.c_void_p, C.c_int, c_double_p, c_double_p, c_double_p, C.POINTER(C.c_void_p)]),
    "nusi_points_destroy": (None, [C.c_void_p]),
    "nusi_points_count": (C.c_int, [C.c_void_p]),
    "nusi_evolve_points": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "nusi_tables_batch": (C.c_int, [C.c_void_p, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_int_p]),
    "nusi_last_timing": (C.c_int, [C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float), c_int_p]),
    "nusi_specfun_eval": (C.c_int, [C.c_int, C.c_int, C.c_int, c_double_p, c_double_p, c_double_p, c_double_p]),
    "nusi_measure_fp64_peak": (C.c_int, [C.c_int, c_double_p]),
}

_lib = None


class NusiError(RuntimeError):
    pass


def load():
    """Load the CUDA library; raise (never fall back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NusiError(
            "CUDA library %s not found: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nusiprop_b200 has no CPU implementation)" % LIB_PATH
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def load_path(path):
    """An ALTERNATIVE build of the same library (tests: the libm-locked build), not cached; pass it to
    Context(_library=...).  The product always uses load()."""
    if not os.path.exists(path):
        raise NusiError("library %s not found" % path)
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def check(rc, what=""):
    if rc != 0:
        lib = load()
        msg = lib.nusi_strerror(rc).decode()
        det = lib.nusi_last_error().decode()
        raise NusiError("%s failed: %s%s" % (what or "nusiprop_b200 call", msg, (" (" + det + ")") if det else ""))
