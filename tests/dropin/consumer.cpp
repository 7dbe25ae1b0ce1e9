// A consumer of the drop-in class written against the REFERENCE's interface only (constructor,
// public members, evolve(), getters): the same calls as the reference's test.cpp, with the options
// taken from argv and full-precision output so the tests can compare with the oracle.
//   consumer mphi g mntot si norm majorana non_resonant normal_ordering N lEmin lEmax zmax flav phiphi [mode]
// mode: "evolve" (default) | "grid" (no GPU needed) | "copy" (exercise copy/assign semantics) | "conservation"
#include "nuSIprop.hpp"
#include <cstdio>
#include <cstdlib>
#include <cstring>

int main(int argc, char **argv) {
  if (argc < 15) {
    std::fprintf(stderr, "usage: see source\n");
    return 2;
  }
  const char *mode = argc > 15 ? argv[15] : "evolve";
  nuSIprop::calculate_flux evolver(atof(argv[1]), atof(argv[2]), atof(argv[3]), atof(argv[4]), atof(argv[5]), atoi(argv[6]) != 0,
                                   atoi(argv[7]) != 0, atoi(argv[8]) != 0, atoi(argv[9]), atof(argv[10]), atof(argv[11]),
                                   atof(argv[12]), atoi(argv[13]), atoi(argv[14]) != 0, argc > 16 ? atoi(argv[16]) : 0);
  if (!std::strcmp(mode, "grid")) {
    for (int i = 0; i < evolver.get_N_bins_E(); ++i) std::printf("%.17g\n", evolver.get_energy(i));
    std::printf("oob %.17g %.17g\n", evolver.get_energy(-1), evolver.get_flux(3, 0));
    return 0;
  }
  if (!std::strcmp(mode, "conservation")) {
    std::printf("%.17g\n", evolver.check_energy_conservation());
    return 0;
  }
  evolver.evolve();
  if (!std::strcmp(mode, "copy")) {
    nuSIprop::calculate_flux a(evolver);  // copy-ctor keeps the result
    nuSIprop::calculate_flux b;           // default ctor + assignment, the way Cython uses the class
    b = evolver;
    evolver.g *= 10;  // public members are re-read at each evolve()
    evolver.evolve();
    int same = 1, changed = 0;
    for (int i = 0; i < a.get_N_bins_E(); ++i) {
      same &= (a.get_flux_fla(0, i) == b.get_flux_fla(0, i));
      changed |= (a.get_flux_fla(0, i) != evolver.get_flux_fla(0, i));
    }
    std::printf("copy_equal %d original_changed %d\n", same, changed);
    return (same && changed) ? 0 : 1;
  }
  for (int i = 0; i < evolver.get_N_bins_E(); ++i)
    std::printf("%.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", evolver.get_energy(i), evolver.get_flux(0, i), evolver.get_flux(1, i),
                evolver.get_flux(2, i), evolver.get_flux_fla(0, i), evolver.get_flux_fla(1, i), evolver.get_flux_fla(2, i));
  return 0;
}
