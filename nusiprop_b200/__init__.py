"""nusiprop_b200 -- B200-native redshift evolution of astrophysical neutrino fluxes with
self-interactions (the hot path of quarkquartet/nuSIprop), batched over parameter points."""
from .api import Context, Points, pyprop, evolve_scan, masses, specfun, measure_fp64_peak, generate_phiphi_tables  # noqa: F401
from ._lib import NusiError  # noqa: F401

__all__ = ["Context", "Points", "pyprop", "evolve_scan", "masses", "specfun", "measure_fp64_peak", "generate_phiphi_tables", "NusiError"]
