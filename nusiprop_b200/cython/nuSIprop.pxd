# C++ declarations imported by nuSIprop.pyx (mirrors reference nuSIprop.pxd:2-18).  The header is
# nusiprop_b200/cxx/nuSIprop.hpp: same class, GPU-backed.
cdef extern from "nuSIprop.hpp" namespace "nuSIprop":
    cdef cppclass calculate_flux:
        calculate_flux() except +
        calculate_flux(double, double, double, double,
                       double,
                       bint, bint, bint,
                       int, double, double,
                       double, int, bint, int) except +
        void evolve() except +
        double get_flux(int, int)
        double get_flux_fla(int, int)
        double get_energy(int)
        int get_N_bins_E()
        double check_energy_conservation() except +
        int get_status()

        double mphi, g, mntot, si, norm

    int c_evolve_scan "nuSIprop::evolve_scan"(int n, const double *mphi, const double *g, const double *mntot, double *flux_mass,
                                    double *flux_fla, int *status, bint majorana, bint non_resonant, bint normal_ordering,
                    int N_bins_E, double lEmin, double lEmax, double zmax, int flav, bint phiphi) except +
