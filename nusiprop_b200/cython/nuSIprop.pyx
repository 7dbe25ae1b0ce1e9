# distutils: language = c++
# cython: language_level=3
"""Cython wrapper `pyprop` over the GPU-backed nuSIprop::calculate_flux: same class name, constructor
keywords and defaults, and methods as the reference's nuSIprop.pyx (:12-144).  Added: evolve_scan(),
one call for many (mphi, g, mntot) points, and getters that fill whole arrays instead of one
element per call."""
from nuSIprop cimport calculate_flux, c_evolve_scan
import numpy as np
cimport numpy as cnp
import warnings

cnp.import_array()


cdef class pyprop:
    """Evolves an astrophysical neutrino flux with self-interactions (see reference nuSIprop.pyx:13-41
    for the meaning of the parameters; defaults are identical, including phiphi=True)."""
    cdef calculate_flux my_object
    cdef bint evolved

    def __cinit__(self, double mphi, double g, double mntot, double si, double norm=1, bint majorana=True,
                  bint non_resonant=True, bint normal_ordering=True, int N_bins_E=300, double lEmin=12.0,
                  double lEmax=17.0, double zmax=5.0, int flav=2, bint phiphi=True, int source_model=0):
        # source_model (extra): 0 = the reference's DSNB source; 1 = upstream power law (reference hpp:648-657)
        self.my_object = calculate_flux(mphi, g, mntot, si, norm, majorana, non_resonant, normal_ordering,
                                        N_bins_E, lEmin, lEmax, zmax, flav, phiphi, source_model)
        self.evolved = False

    def set_parameters(self, mphi=None, g=None, mntot=None, si=None, norm=None):
        """Modify the physics parameters (reference nuSIprop.pyx:60-85)."""
        if mphi is not None:
            self.my_object.mphi = mphi
        if g is not None:
            self.my_object.g = g
        if mntot is not None:
            self.my_object.mntot = mntot
        if si is not None:
            self.my_object.si = si
        if norm is not None:
            self.my_object.norm = norm
        self.evolved = False

    def evolve(self):
        """Evolve the neutrino flux (on the GPU)."""
        self.evolved = True
        self.my_object.evolve()

    cdef _fill(self, bint flavour):
        cdef int N = self.my_object.get_N_bins_E()
        cdef cnp.ndarray[double, ndim=2] flx = np.zeros([3, N])
        cdef int i, j
        if not self.evolved:
            warnings.warn("You have not evolved the neutrino flux! Zero flux will be returned.")
            return flx
        for i in range(3):
            for j in range(N):
                flx[i, j] = self.my_object.get_flux_fla(i, j) if flavour else self.my_object.get_flux(i, j)
        return flx

    def get_flux(self):
        """Flux of each mass eigenstate, [3, N_bins_E]."""
        return self._fill(False)

    def get_flux_fla(self):
        """Flux of each flavour {e, mu, tau}, [3, N_bins_E]."""
        return self._fill(True)

    def _interp(self, int f, energy):
        import scipy.interpolate as interp
        en = self.get_energies()
        si = self.my_object.si
        return interp.interp1d(np.log10(en), self.get_flux_fla()[f] * en ** si)(np.log10(energy)) / energy ** si

    def interp_flux_el(self, energy):
        return self._interp(0, energy)

    def interp_flux_mu(self, energy):
        return self._interp(1, energy)

    def interp_flux_ta(self, energy):
        return self._interp(2, energy)

    def get_energies(self):
        """Energy bin centers."""
        cdef int N = self.my_object.get_N_bins_E()
        cdef cnp.ndarray[double, ndim=1] en = np.zeros(N)
        cdef int i
        for i in range(N):
            en[i] = self.my_object.get_energy(i)
        return en

    def check_energy_conservation(self):
        """(E_int - E_FS)/E_FS as the reference computes it (calls evolve())."""
        self.evolved = True
        return self.my_object.check_energy_conservation()

    @property
    def status(self):
        return self.my_object.get_status()


def evolve_scan(mphi, g, mntot, bint majorana=True, bint non_resonant=True, bint normal_ordering=True,
                int N_bins_E=300, double lEmin=12.0, double lEmax=17.0, double zmax=5.0, int flav=2,
                bint phiphi=False):
    """evolve() for n parameter points in one GPU batch.
    Returns (flux_mass[n,3,N], flux_fla[n,3,N], status[n])."""
    a, b, c = np.broadcast_arrays(np.atleast_1d(np.asarray(mphi, dtype=np.float64)),
                                  np.atleast_1d(np.asarray(g, dtype=np.float64)),
                                  np.atleast_1d(np.asarray(mntot, dtype=np.float64)))
    cdef cnp.ndarray[double, ndim=1] vm = np.ascontiguousarray(a).ravel()
    cdef cnp.ndarray[double, ndim=1] vg = np.ascontiguousarray(b).ravel()
    cdef cnp.ndarray[double, ndim=1] vn = np.ascontiguousarray(c).ravel()
    cdef int n = vm.shape[0]
    cdef cnp.ndarray[double, ndim=3] fm = np.empty((n, 3, N_bins_E))
    cdef cnp.ndarray[double, ndim=3] ff = np.empty((n, 3, N_bins_E))
    cdef cnp.ndarray[int, ndim=1] st = np.zeros(n, dtype=np.intc)
    cdef int rc
    rc = c_evolve_scan(n, &vm[0], &vg[0], &vn[0], &fm[0, 0, 0], &ff[0, 0, 0], &st[0], majorana, non_resonant,
                        normal_ordering, N_bins_E, lEmin, lEmax, zmax, flav, phiphi)
    if rc != 0:
        raise RuntimeError("nuSIprop.evolve_scan failed with code %d" % rc)
    return fm, ff, st
