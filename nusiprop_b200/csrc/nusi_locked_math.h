// nusi_locked_math.h -- elementary functions written out in IEEE-754 basic operations only (+, -, *, /, sqrt, bit moves;
// no library call, no contraction), so that the SAME source gives the SAME bits under g++ -ffp-contract=off on the host
// and under nvcc -fmad=false on the device.  log / log1p / atan / atan2 / atanh follow the classic argument reductions and
// minimax polynomials of the freely distributable fdlibm (error < 1 ulp each; checked against mpmath in
// tests/test_locked_math.py); hypot is sqrt(x^2 + y^2) without rescaling (the path's arguments are O(1)).
//
// Purpose: the "libm-locked" pair of builds
//     oracle/_ref/libnusi_ref_locked.so        the unmodified reference with log/log1p/atan/atanh/carg/cabs/clog routed here
//     nusiprop_b200/lib/libnusiprop_b200_locked.so   this library with -DNUSI_LOCKED_MATH (nlog, nlog1p, ... route here)
// differ only where the product's arithmetic differs from the reference's -- not by whose libm rounded a logarithm.
// Test infrastructure: the product library proper never includes this header.
#ifndef NUSI_LOCKED_MATH_H
#define NUSI_LOCKED_MATH_H
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define NUSI_LK __host__ __device__ inline
#else
#define NUSI_LK inline
#endif

namespace nusi_locked {

NUSI_LK int hi_word(double x) {
#if defined(__CUDA_ARCH__)
  return __double2hiint(x);
#else
  unsigned long long b;
  memcpy(&b, &x, 8);
  return (int)(b >> 32);
#endif
}
NUSI_LK unsigned lo_word(double x) {
#if defined(__CUDA_ARCH__)
  return (unsigned)__double2loint(x);
#else
  unsigned long long b;
  memcpy(&b, &x, 8);
  return (unsigned)b;
#endif
}
NUSI_LK double with_hi(double x, int hi) {
#if defined(__CUDA_ARCH__)
  return __hiloint2double(hi, __double2loint(x));
#else
  unsigned long long b;
  memcpy(&b, &x, 8);
  b = (b & 0xffffffffull) | ((unsigned long long)(unsigned)hi << 32);
  memcpy(&x, &b, 8);
  return x;
#endif
}

#define NUSI_LK_LN2_HI 6.93147180369123816490e-01
#define NUSI_LK_LN2_LO 1.90821492927058770002e-10
#define NUSI_LK_L1 6.666666666666735130e-01
#define NUSI_LK_L2 3.999999999940941908e-01
#define NUSI_LK_L3 2.857142874366239149e-01
#define NUSI_LK_L4 2.222219843214978396e-01
#define NUSI_LK_L5 1.818357216161805012e-01
#define NUSI_LK_L6 1.531383769920937332e-01
#define NUSI_LK_L7 1.479819860511658591e-01

// log(x): x = 2^k (1+f), sqrt(2)/2 < 1+f < sqrt(2); log(1+f) = 2s + s R(s^2), s = f/(2+f)
NUSI_LK double log(double x) {
  int hx = hi_word(x), k = 0;
  const unsigned lx = lo_word(x);
  if (hx < 0x00100000) {  // x < 2^-1022: zero, negative or subnormal
    if (((hx & 0x7fffffff) | (int)lx) == 0) return -HUGE_VAL;
    if (hx < 0) return (x - x) / 0.0;
    k -= 54;
    x *= 1.80143985094819840000e+16;  // 2^54
    hx = hi_word(x);
  }
  if (hx >= 0x7ff00000) return x + x;
  k += (hx >> 20) - 1023;
  hx &= 0x000fffff;
  int i = (hx + 0x95f64) & 0x100000;
  x = with_hi(x, hx | (i ^ 0x3ff00000));  // x or x/2
  k += (i >> 20);
  const double f = x - 1.0;
  const double dk = (double)k;
  if ((0x000fffff & (2 + hx)) < 3) {  // |f| < 2^-20
    if (f == 0.0) return (k == 0) ? 0.0 : dk * NUSI_LK_LN2_HI + dk * NUSI_LK_LN2_LO;
    const double R = f * f * (0.5 - 0.33333333333333333 * f);
    return (k == 0) ? f - R : dk * NUSI_LK_LN2_HI - ((R - dk * NUSI_LK_LN2_LO) - f);
  }
  const double s = f / (2.0 + f), z = s * s, w = z * z;
  i = hx - 0x6147a;
  const int j = 0x6b851 - hx;
  const double t1 = w * (NUSI_LK_L2 + w * (NUSI_LK_L4 + w * NUSI_LK_L6));
  const double t2 = z * (NUSI_LK_L1 + w * (NUSI_LK_L3 + w * (NUSI_LK_L5 + w * NUSI_LK_L7)));
  i |= j;
  const double R = t2 + t1;
  if (i > 0) {
    const double hfsq = 0.5 * f * f;
    return (k == 0) ? f - (hfsq - s * (hfsq + R)) : dk * NUSI_LK_LN2_HI - ((hfsq - (s * (hfsq + R) + dk * NUSI_LK_LN2_LO)) - f);
  }
  return (k == 0) ? f - s * (f - R) : dk * NUSI_LK_LN2_HI - ((s * (f - R) - dk * NUSI_LK_LN2_LO) - f);
}

// log1p(x): 1+x = 2^k (1+f) with the rounding error of 1+x carried as a correction term c
NUSI_LK double log1p(double x) {
  const int hx = hi_word(x), ax = hx & 0x7fffffff;
  int k = 1, hu = 0;
  double f = 0.0, c = 0.0;
  if (hx < 0x3FDA827A) {  // x < 0.41422
    if (ax >= 0x3ff00000) return (x == -1.0) ? -HUGE_VAL : (x - x) / (x - x);
    if (ax < 0x3e200000) {  // |x| < 2^-29
      if (ax < 0x3c900000) return x;
      return x - x * x * 0.5;
    }
    if (hx > 0 || hx <= (int)0xbfd2bec3) {  // -0.2929 < x < 0.41422
      k = 0;
      f = x;
      hu = 1;
    }
  }
  if (hx >= 0x7ff00000) return x + x;
  if (k != 0) {
    double u;
    if (hx < 0x43400000) {
      u = 1.0 + x;
      hu = hi_word(u);
      k = (hu >> 20) - 1023;
      c = (k > 0) ? 1.0 - (u - x) : x - (u - 1.0);
      c /= u;
    } else {
      u = x;
      hu = hi_word(u);
      k = (hu >> 20) - 1023;
      c = 0.0;
    }
    hu &= 0x000fffff;
    if (hu < 0x6a09e) {
      u = with_hi(u, hu | 0x3ff00000);
    } else {
      k += 1;
      u = with_hi(u, hu | 0x3fe00000);
      hu = (0x00100000 - hu) >> 2;
    }
    f = u - 1.0;
  }
  const double dk = (double)k;
  const double hfsq = 0.5 * f * f;
  if (hu == 0) {  // |f| < 2^-20
    if (f == 0.0) {
      if (k == 0) return 0.0;
      c += dk * NUSI_LK_LN2_LO;
      return dk * NUSI_LK_LN2_HI + c;
    }
    const double R = hfsq * (1.0 - 0.66666666666666666 * f);
    return (k == 0) ? f - R : dk * NUSI_LK_LN2_HI - ((R - (dk * NUSI_LK_LN2_LO + c)) - f);
  }
  const double s = f / (2.0 + f), z = s * s;
  const double R = z * (NUSI_LK_L1 + z * (NUSI_LK_L2 + z * (NUSI_LK_L3 + z * (NUSI_LK_L4 + z * (NUSI_LK_L5 + z * (NUSI_LK_L6 + z * NUSI_LK_L7))))));
  return (k == 0) ? f - (hfsq - s * (hfsq + R)) : dk * NUSI_LK_LN2_HI - ((hfsq - (s * (hfsq + R) + (dk * NUSI_LK_LN2_LO + c))) - f);
}

// atan(x): reduction to [0, 7/16] around 0.5, 1, 1.5, inf; odd minimax polynomial
NUSI_LK double atan(double x) {
  const int hx = hi_word(x), ix = hx & 0x7fffffff;
  if (ix >= 0x44100000) {  // |x| >= 2^66
    if (ix > 0x7ff00000 || (ix == 0x7ff00000 && lo_word(x) != 0)) return x + x;
    return (hx > 0) ? 1.57079632679489655800e+00 + 6.12323399573676603587e-17 : -1.57079632679489655800e+00 - 6.12323399573676603587e-17;
  }
  int id;
  if (ix < 0x3fdc0000) {  // |x| < 0.4375
    if (ix < 0x3e200000) return x;  // |x| < 2^-29
    id = -1;
  } else {
    x = fabs(x);
    if (ix < 0x3ff30000) {    // |x| < 1.1875
      if (ix < 0x3fe60000) {  // 7/16 <= |x| < 11/16
        id = 0;
        x = (2.0 * x - 1.0) / (2.0 + x);
      } else {
        id = 1;
        x = (x - 1.0) / (x + 1.0);
      }
    } else if (ix < 0x40038000) {  // |x| < 2.4375
      id = 2;
      x = (x - 1.5) / (1.0 + 1.5 * x);
    } else {
      id = 3;
      x = -1.0 / x;
    }
  }
  const double z = x * x, w = z * z;
  const double s1 = z * (3.33333333333329318027e-01 + w * (1.42857142725034663711e-01 + w * (9.09088713343650656196e-02 +
                         w * (6.66107313738753120669e-02 + w * (4.97687799461593236017e-02 + w * 1.62858201153657823623e-02)))));
  const double s2 = w * (-1.99999999998764832476e-01 + w * (-1.11111104054623557880e-01 + w * (-7.69187620504482999495e-02 +
                         w * (-5.83357013379057348645e-02 + w * -3.65315727442169155270e-02))));
  if (id < 0) return x - x * (s1 + s2);
  const double hi = (id == 0) ? 4.63647609000806093515e-01 : (id == 1) ? 7.85398163397448278999e-01 : (id == 2) ? 9.82793723247329054082e-01 : 1.57079632679489655800e+00;
  const double lo = (id == 0) ? 2.26987774529616870924e-17 : (id == 1) ? 3.06161699786838301793e-17 : (id == 2) ? 1.39033110312309984516e-17 : 6.12323399573676603587e-17;
  const double r = hi - ((x * (s1 + s2) - lo) - x);
  return (hx < 0) ? -r : r;
}

NUSI_LK double atan2(double y, double x) {
  const double pi = 3.1415926535897931160E+00, pi_lo = 1.2246467991473531772E-16, pi_o_2 = 1.5707963267948965580E+00, pi_o_4 = 7.8539816339744827900E-01;
  const int hx = hi_word(x), ix = hx & 0x7fffffff, hy = hi_word(y), iy = hy & 0x7fffffff;
  const unsigned lx = lo_word(x), ly = lo_word(y);
  if (x != x || y != y) return x + y;
  if (x == 1.0) return nusi_locked::atan(y);
  const int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);  // 2 sign(x) + sign(y)
  if ((iy | (int)ly) == 0) {                          // y = 0
    switch (m) {
      case 0:
      case 1: return y;
      case 2: return pi;
      default: return -pi;
    }
  }
  if ((ix | (int)lx) == 0) return (hy < 0) ? -pi_o_2 : pi_o_2;  // x = 0
  if (ix == 0x7ff00000) {                                        // x = inf
    if (iy == 0x7ff00000) {
      switch (m) {
        case 0: return pi_o_4;
        case 1: return -pi_o_4;
        case 2: return 3.0 * pi_o_4;
        default: return -3.0 * pi_o_4;
      }
    }
    switch (m) {
      case 0: return 0.0;
      case 1: return -0.0;
      case 2: return pi;
      default: return -pi;
    }
  }
  if (iy == 0x7ff00000) return (hy < 0) ? -pi_o_2 : pi_o_2;
  const int k = (iy - ix) >> 20;
  double z;
  if (k > 60)
    z = pi_o_2 + 0.5 * pi_lo;  // |y/x| > 2^60
  else if (hx < 0 && k < -60)
    z = 0.0;  // |y|/x < -2^60
  else
    z = nusi_locked::atan(fabs(y / x));
  switch (m) {
    case 0: return z;
    case 1: return -z;
    case 2: return pi - (z - pi_lo);
    default: return (z - pi_lo) - pi;
  }
}

NUSI_LK double atanh(double x) {
  const int hx = hi_word(x), ix = hx & 0x7fffffff;
  const double ax = fabs(x);
  if (ax > 1.0 || x != x) return (x - x) / (x - x);
  if (ax == 1.0) return x / 0.0;
  if (ix < 0x3e300000) return x;  // |x| < 2^-28
  double t;
  if (ix < 0x3fe00000) {  // |x| < 0.5
    t = ax + ax;
    t = 0.5 * nusi_locked::log1p(t + t * ax / (1.0 - ax));
  } else {
    t = 0.5 * nusi_locked::log1p((ax + ax) / (1.0 - ax));
  }
  return (hx >= 0) ? t : -t;
}

NUSI_LK double hypot(double x, double y) { return sqrt(x * x + y * y); }

}  // namespace nusi_locked
#endif
