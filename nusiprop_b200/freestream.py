"""Free-streaming diagnostics of the reference (nuSIprop.hpp:666-744): flux_FS_E0, energy_FS,
Lum_times_E.  Cheap host arithmetic (300 pow() calls), kept on the host by design (SURVEY
section 2, row 7).  In this fork they describe the commented-out power-law source, so
check_energy_conservation() compares a DSNB-sourced flux with a power-law energy exactly like
the reference does."""
import math

_W = (5.0 / 9.0, 8.0 / 9.0, 5.0 / 9.0)
_X = (-math.sqrt(3.0 / 5.0), 0.0, math.sqrt(3.0 / 5.0))
E0 = 1e14
N_INTEG_Z = 100


def get_H(z):
    return 1.5e-33 * math.pow(0.692 + 0.308 * (1 + z) * (1 + z) * (1 + z), 0.5)


def get_SFR(z):
    return math.pow(math.pow(1 + z, -3.4 * 10) + math.pow((1 + z) / 5161, 0.3 * 10) + math.pow((1 + z) / 9.06, 3.5 * 10), -1.0 / 10.0)


def _nodes(f, z_min, z_max):
    a = z_min + f * (z_max - z_min) / N_INTEG_Z
    b = z_min + (f + 1.0) * (z_max - z_min) / N_INTEG_Z
    return a, b, [(b - a) / 2.0 * x + (b + a) / 2.0 for x in _X]


def flux_FS_E0(zmax, si):
    res = 0.0
    for f in range(N_INTEG_Z):
        a, b, zs = _nodes(f, 0.0, zmax)
        res += (b - a) / 2.0 * (_W[0] * math.pow(1 + zs[0], -si) * get_SFR(zs[0]) / get_H(zs[0])
                                + _W[1] * math.pow(1 + zs[1], -si) * get_SFR(zs[1]) / get_H(zs[1])
                                + _W[2] * math.pow(1 + zs[2], -si) * get_SFR(zs[2]) / get_H(zs[2]))
    return res


def Lum_times_E(norm_total, z, Em, Ep, si):
    if abs(si - 2) < 1e-5:
        return norm_total * get_SFR(z) * math.pow(E0 / (1 + z), si) * (math.log(Ep / Em) + (2 - si) / 2.0 * (math.log(Ep) ** 2 - math.log(Em) ** 2))
    return norm_total * get_SFR(z) * math.pow(E0 / (1 + z), si) * (math.pow(Ep, 2 - si) - math.pow(Em, 2 - si)) / (2 - si)


def energy_FS(ctx, si, norm):
    zmax = ctx.zmax
    lEmin, lEmax = ctx.opts["lEmin"], ctx.opts["lEmax"]
    norm_total = norm / flux_FS_E0(zmax, si)
    res = 0.0
    Em, Ep = math.pow(10, lEmin), math.pow(10, lEmax)
    for f in range(N_INTEG_Z):
        a, b, zs = _nodes(f, 0.0, zmax)
        res += (b - a) / 2.0 * (_W[0] * Lum_times_E(norm_total, zs[0], Em, Ep, si) / get_H(zs[0])
                                + _W[1] * Lum_times_E(norm_total, zs[1], Em, Ep, si) / get_H(zs[1])
                                + _W[2] * Lum_times_E(norm_total, zs[2], Em, Ep, si) / get_H(zs[2]))
    return res
