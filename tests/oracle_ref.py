"""ctypes access to the ORACLE (oracle/_ref/libnusi_ref.so = unmodified reference headers +
refshim).  Test infrastructure only: imported by tests/, __graft_entry__.smoke() and bench.py's
CPU-baseline legs, never by the nusiprop_b200 package."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")
REFERENCE_SRC = os.environ.get("NUSI_REFERENCE_DIR", "/root/reference")

D, I = C.c_double, C.c_int
DP = C.POINTER(C.c_double)


def build_oracle(force=False):
    """(Re)build oracle/_ref from the reference sources where they lie.  Only possible where
    /root/reference exists (this container); the GPU box uses the prebuilt files."""
    have = os.path.exists(os.path.join(REF_DIR, "libnusi_ref.so"))
    if not os.path.isdir(REFERENCE_SRC):
        return have
    if have and not force:
        src_t = max(os.path.getmtime(os.path.join(dp, f)) for dp, _, fs in os.walk(ORACLE_DIR) if "_ref" not in dp for f in fs)
        if os.path.getmtime(os.path.join(REF_DIR, "libnusi_ref.so")) >= src_t:
            return True
    subprocess.check_call(["make", "-C", ORACLE_DIR, "-B", "-j4", "REF=" + REFERENCE_SRC], stdout=subprocess.DEVNULL)
    return True


def available():
    return os.path.exists(os.path.join(REF_DIR, "libnusi_ref.so"))


_libs = {}


def lib(variant=""):
    """variant "" = strict oracle (-O3 -ffp-contract=off); "_fast" = -O2 -mfma contraction on."""
    if variant not in _libs:
        path = os.path.join(REF_DIR, "libnusi_ref%s.so" % variant)
        if not os.path.exists(path):
            raise RuntimeError("oracle library %s missing: run `make -C oracle` where /root/reference exists" % path)
        L = C.CDLL(path)
        L.nusi_ref_dilog.restype = D
        L.nusi_ref_dilog.argtypes = [D]
        for nm in ("nusi_ref_Li2", "nusi_ref_Li3"):
            getattr(L, nm).restype = D
            getattr(L, nm).argtypes = [D]
        for nm in ("nusi_ref_atandiff", "nusi_ref_dilogdiff", "nusi_ref_dilog1mdiff", "nusi_ref_dilog1pdiff", "nusi_ref_dilog1over1mdiff"):
            getattr(L, nm).restype = D
            getattr(L, nm).argtypes = [D, D]
        L.nusi_ref_cdilog.argtypes = [D, D, DP, DP]
        L.nusi_ref_dilogdiff_complex.argtypes = [D, D, D, D, DP, DP]
        L.nusi_ref_evolve.argtypes = [D, D, D, D, D, I, I, I, I, D, D, D, I, I, DP, DP, DP]
        L.nusi_ref_grid.argtypes = [I, D, D, D, C.POINTER(I), DP, DP, DP, DP]
        L.nusi_ref_masses.argtypes = [D, I, DP]
        L.nusi_ref_pmns.argtypes = [I, DP]
        L.nusi_ref_tables.argtypes = [D, D, D, I, I, I, I, D, D, D, I, I, DP, DP, DP]
        L.nusi_ref_source.argtypes = [I, D, D, D, DP]
        L.nusi_ref_check_energy_conservation.restype = D
        L.nusi_ref_check_energy_conservation.argtypes = [D, D, D, D, D, I, I, I, I, D, D, D, I]
        _libs[variant] = L
    return _libs[variant]


def _p(a):
    return a.ctypes.data_as(DP) if a is not None else None


def _silence():
    """The reference floods stderr with 'Negative cross section' text; mute fd 2 around calls."""
    class _Ctx(object):
        def __enter__(self):
            self.saved = os.dup(2)
            self.null = os.open(os.devnull, os.O_WRONLY)
            os.dup2(self.null, 2)
        def __exit__(self, *a):
            os.dup2(self.saved, 2)
            os.close(self.null)
            os.close(self.saved)
    return _Ctx()


def capture_stderr(fn):
    """Run fn() with the process's stderr (fd 2, where the reference prints its warnings) captured; returns (result, text)."""
    import tempfile
    with tempfile.TemporaryFile() as tf:
        saved = os.dup(2)
        os.dup2(tf.fileno(), 2)
        try:
            res = fn()
        finally:
            os.dup2(saved, 2)
            os.close(saved)
        tf.seek(0)
        return res, tf.read().decode(errors="replace")


def evolve(mphi, g, mntot, si=2.0, norm=1.0, majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=300,
           lEmin=12.0, lEmax=17.0, zmax=5.0, flav=2, phiphi=False, variant=""):
    """calculate_flux(...).evolve() of the reference -> (flux_mass[3,N], flux_fla[3,N], energies[N])."""
    N = int(N_bins_E)
    fm, ff, en = np.zeros((3, N)), np.zeros((3, N)), np.zeros(N)
    with _silence():
        lib(variant).nusi_ref_evolve(mphi, g, mntot, si, norm, int(majorana), int(non_resonant), int(normal_ordering), N,
                                     lEmin, lEmax, zmax, int(flav), int(phiphi), _p(fm), _p(ff), _p(en))
    return fm, ff, en


def grid(N_bins_E, lEmin, lEmax, zmax, variant=""):
    N = int(N_bins_E)
    nz = I()
    Emin, Emax, Ec, z = np.zeros(N), np.zeros(N), np.zeros(N), np.zeros(100000)
    lib(variant).nusi_ref_grid(N, lEmin, lEmax, zmax, C.byref(nz), _p(Emin), _p(Emax), _p(Ec), _p(z))
    return nz.value, Emin, Emax, Ec, z[: nz.value].copy()


def masses(mntot, normal_ordering=True):
    mn = np.zeros(3)
    lib().nusi_ref_masses(mntot, int(normal_ordering), _p(mn))
    return mn


def pmns2(normal_ordering=True):
    u = np.zeros(9)
    lib().nusi_ref_pmns(int(normal_ordering), _p(u))
    return u.reshape(3, 3)


def tables(mphi, g, mntot, majorana=True, non_resonant=True, normal_ordering=True, N_bins_E=300, lEmin=12.0, lEmax=17.0,
           zmax=5.0, flav=2, phiphi=False, want_alpha=True, variant=""):
    nz = grid(N_bins_E, lEmin, lEmax, zmax)[0]
    NT = int(N_bins_E) + nz - 2
    G, A = np.zeros(NT), np.zeros(NT)
    T = np.zeros((NT, NT)) if want_alpha else None
    with _silence():
        lib(variant).nusi_ref_tables(mphi, g, mntot, int(majorana), int(non_resonant), int(normal_ordering), int(N_bins_E),
                                     lEmin, lEmax, zmax, int(flav), int(phiphi), _p(G), _p(A), _p(T))
    return G, A, T


def source(N_bins_E, lEmin, lEmax, zmax, variant=""):
    nz = grid(N_bins_E, lEmin, lEmax, zmax)[0]
    out = np.zeros((nz - 1, int(N_bins_E)))
    lib(variant).nusi_ref_source(int(N_bins_E), lEmin, lEmax, zmax, _p(out))
    return out


def relerr(a, b):
    """|a-b|/|b| per element with 0==0 and NaN==NaN counted as exact."""
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    with np.errstate(all="ignore"):
        r = np.abs(a - b) / np.abs(b)
    r[a == b] = 0.0
    r[np.isnan(a) & np.isnan(b)] = 0.0
    r[np.isnan(r)] = np.inf
    return r


def noise_floor(ref, variants, halfwidth=8):
    """Per-bin noise floor of the reference: largest relative spread between the strict oracle `ref`
    [..., N] and its benign variants (FMA-contracted rebuild, +-1-ulp libm), taken over the leading
    axes and widened to a running maximum over +-halfwidth neighbouring energy bins (one ensemble member
    can agree by accident in a bin whose neighbours show the conditioning)."""
    f = np.zeros(ref.shape[-1])
    for v in variants:
        r = relerr(v, ref)
        f = np.maximum(f, r.reshape(-1, r.shape[-1]).max(axis=0))
    return np.array([f[max(0, i - halfwidth): i + halfwidth + 1].max() for i in range(len(f))])


TOLERANCE_CAP = 0.1  # whatever the ensemble's spread: a flux more than 10 % off the oracle fails


def flux_tolerance(ref, variants, k=4.0, strict=1e-9, cap=TOLERANCE_CAP):
    """north_star's 1e-9 wherever the reference is self-consistent to that level, k x its own spread elsewhere,
    never more than `cap` (a bin in which the ensemble scatters by O(1) must not excuse an arbitrary result)."""
    return np.minimum(cap, np.maximum(strict, k * noise_floor(ref, variants)))
