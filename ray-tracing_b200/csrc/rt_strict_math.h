// rt_strict_math.h — sin, cos, atan2, asin, log written in plain IEEE double arithmetic (+ - * / sqrt, floor, bit casts) so that
// the SAME source gives the SAME bits on the device (nvcc --fmad=false) and on the host (g++ -ffp-contract=off).
//
// Why: the reference calls libm (GHC primops -> sin/cos/atan2/asin/log/**, Lib.hs:484-513, 1085-1105, 1187-1228, 1071).  glibc's and
// CUDA's implementations differ in the last ulp for some arguments, and in fp32 at scene coordinates of a few hundred units a
// one-ulp change of a scattered direction changes the ROUNDING of the next hit point, which decides whether the next ray
// re-hits its own surface at t_min = 1e-4 (Lib.hs:77, 1309) — paths then separate (tools/parity_classify.py measures it).  Under
// RT_FLAG_STRICT_MATH the kernels evaluate these functions here — fp32 arithmetic through double and rounded once — and the test
// oracle includes this same header, so that same-stream radiance can be compared (nearly) bit for bit.
// Accuracy: the classic fdlibm-style kernels (Cody–Waite reduction by pi/2 in three parts, minimax polynomials), within ~1 ulp
// for |x| < 1e5; larger arguments lose accuracy gracefully but stay deterministic.  Not a libm replacement.
// The product's default arithmetic does NOT use this file (it calls CUDA's libdevice).
#pragma once
#include <stdint.h>
#include <string.h>
#include <math.h>

#if defined(__CUDACC__)
#define RTB_SM_HD __host__ __device__ inline
#else
#define RTB_SM_HD inline
#endif

namespace rtb_strict {

RTB_SM_HD uint64_t bitsOfD(double x) {
#if defined(__CUDA_ARCH__)
  return (uint64_t)__double_as_longlong(x);
#else
  uint64_t b; memcpy(&b, &x, 8); return b;
#endif
}
RTB_SM_HD double dOfBits(uint64_t b) {
#if defined(__CUDA_ARCH__)
  return __longlong_as_double((long long)b);
#else
  double x; memcpy(&x, &b, 8); return x;
#endif
}
RTB_SM_HD bool signBit(double x) { return (bitsOfD(x) >> 63) != 0; }
RTB_SM_HD double absD(double x) { return dOfBits(bitsOfD(x) & 0x7fffffffffffffffULL); }
RTB_SM_HD double nanD() { return dOfBits(0x7ff8000000000000ULL); }
RTB_SM_HD double infD() { return dOfBits(0x7ff0000000000000ULL); }

// ---- sin / cos ---------------------------------------------------------------------------------------------------------------
RTB_SM_HD double kSin(double r) {
  const double S1 = -1.66666666666666324348e-01, S2 = 8.33333333332248946124e-03, S3 = -1.98412698298579493134e-04,
               S4 = 2.75573137070700676789e-06, S5 = -2.50507602534068634195e-08, S6 = 1.58969099521155010221e-10;
  const double z = r * r;
  const double p = S2 + z * (S3 + z * (S4 + z * (S5 + z * S6)));
  return r + (z * r) * (S1 + z * p);
}
RTB_SM_HD double kCos(double r) {
  const double C1 = 4.16666666666666019037e-02, C2 = -1.38888888888741095749e-03, C3 = 2.48015872894767294178e-05,
               C4 = -2.75573143513906633035e-07, C5 = 2.08757232129817482790e-09, C6 = -1.13596475577881948265e-11;
  const double z = r * r;
  const double p = z * (C1 + z * (C2 + z * (C3 + z * (C4 + z * (C5 + z * C6)))));
  return 1.0 - (0.5 * z - z * p);
}
// x = n * pi/2 + r, |r| <= pi/4 (+ rounding); returns n mod 4.  Valid for |x| <= 1e15 (callers guard).
RTB_SM_HD int remPio2(double x, double& r) {
  const double invpio2 = 6.36619772367581382433e-01;
  const double pio2_1 = 1.57079632673412561417e+00;   // first 33 bits of pi/2: n * pio2_1 is exact for |n| < 2^20
  const double pio2_2 = 6.07710050630396597660e-11;   // second 33 bits
  const double pio2_2t = 2.02226624879595063154e-21;  // pi/2 - (pio2_1 + pio2_2)
  const double fn = floor(x * invpio2 + 0.5);
  const double t = x - fn * pio2_1;
  double w = fn * pio2_2;
  const double r1 = t - w;
  w = fn * pio2_2t - ((t - r1) - w);  // the rounding error of t - w, carried into the tail
  r = r1 - w;
  return (int)((long long)fn & 3LL);
}
RTB_SM_HD double sinD(double x) {
  if (!(absD(x) <= 1e15)) return x - x;  // inf, NaN -> NaN; huge finite -> 0 (deterministic; never reached by the renderer)
  double r; const int n = remPio2(x, r);
  switch (n) { case 0: return kSin(r); case 1: return kCos(r); case 2: return -kSin(r); default: return -kCos(r); }
}
RTB_SM_HD double cosD(double x) {
  if (!(absD(x) <= 1e15)) return x - x;
  double r; const int n = remPio2(x, r);
  switch (n) { case 0: return kCos(r); case 1: return -kSin(r); case 2: return -kCos(r); default: return kSin(r); }
}

// ---- atan / atan2 / asin -------------------------------------------------------------------------------------------------------
RTB_SM_HD double atanPos(double x) {  // x >= 0, finite or +inf
  const double hi0 = 4.63647609000806093515e-01, hi1 = 7.85398163397448278999e-01, hi2 = 9.82793723247329054082e-01, hi3 = 1.57079632679489655800e+00;
  const double lo0 = 2.26987774529616870924e-17, lo1 = 3.06161699786838301793e-17, lo2 = 1.39033110312309984516e-17, lo3 = 6.12323399573676603587e-17;
  const double a0 = 3.33333333333329318027e-01, a1 = -1.99999999998764832476e-01, a2 = 1.42857142725034663711e-01, a3 = -1.11111104054623557880e-01,
               a4 = 9.09088713343650656196e-02, a5 = -7.69187620504482999495e-02, a6 = 6.66107313738753120669e-02, a7 = -5.83357013379057348645e-02,
               a8 = 4.97687799461593236017e-02, a9 = -3.65315727442169155270e-02, a10 = 1.62858201153657823623e-02;
  if (x >= 7.378697629483821e19) return hi3 + lo3;  // 2^66 and beyond (also +inf)
  int id; double hi = 0, lo = 0;
  if (x < 0.4375) { if (x < 1.862645149230957e-09) return x; id = -1; }
  else if (x < 1.1875) {
    if (x < 0.6875) { id = 0; hi = hi0; lo = lo0; x = (2.0 * x - 1.0) / (2.0 + x); }
    else { id = 1; hi = hi1; lo = lo1; x = (x - 1.0) / (x + 1.0); }
  } else if (x < 2.4375) { id = 2; hi = hi2; lo = lo2; x = (x - 1.5) / (1.0 + 1.5 * x); }
  else { id = 3; hi = hi3; lo = lo3; x = -1.0 / x; }
  const double z = x * x, w = z * z;
  const double s1 = z * (a0 + w * (a2 + w * (a4 + w * (a6 + w * (a8 + w * a10)))));
  const double s2 = w * (a1 + w * (a3 + w * (a5 + w * (a7 + w * a9))));
  if (id < 0) return x - x * (s1 + s2);
  return hi - ((x * (s1 + s2) - lo) - x);
}
RTB_SM_HD double atan2D(double y, double x) {
  const double pi = 3.1415926535897931160e+00, pi_lo = 1.2246467991473531772e-16, pio2 = 1.5707963267948965580e+00, pio4 = 7.8539816339744827900e-01;
  if (x != x || y != y) return x + y;
  const bool sx = signBit(x), sy = signBit(y);
  if (y == 0.0) return sx ? (sy ? -pi : pi) : y;
  if (x == 0.0) return sy ? -pio2 : pio2;
  const bool ix = absD(x) == infD(), iy = absD(y) == infD();
  if (ix) {
    if (iy) return sx ? (sy ? -3.0 * pio4 : 3.0 * pio4) : (sy ? -pio4 : pio4);
    return sx ? (sy ? -pi : pi) : (sy ? -0.0 : 0.0);
  }
  if (iy) return sy ? -pio2 : pio2;
  double z = atanPos(absD(y / x));
  if (sx) z = pi - (z - pi_lo);
  return sy ? -z : z;
}
RTB_SM_HD double asinD(double x) {  // |x| > 1 -> NaN (sqrt of a negative), as asin
  return atan2D(x, sqrt((1.0 - x) * (1.0 + x)));
}

// ---- log -------------------------------------------------------------------------------------------------------------------------
RTB_SM_HD double logD(double x) {
  const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
  const double L1 = 6.666666666666735130e-01, L2 = 3.999999999940941908e-01, L3 = 2.857142874366239149e-01, L4 = 2.222219843214978396e-01,
               L5 = 1.818357216161805012e-01, L6 = 1.531383769920937332e-01, L7 = 1.479819860511658591e-01;
  if (x != x) return x;
  if (x == 0.0) return -infD();
  if (signBit(x)) return nanD();
  if (x == infD()) return x;
  int k = 0;
  uint64_t b = bitsOfD(x);
  if ((b >> 52) == 0) { x = x * 18014398509481984.0; k = -54; b = bitsOfD(x); }  // subnormal: scale by 2^54
  uint32_t hx = (uint32_t)(b >> 32);
  k += (int)(hx >> 20) - 1023;
  hx &= 0x000fffffu;
  const uint32_t i = (hx + 0x95f64u) & 0x100000u;  // mantissa >= sqrt(2): use x / 2
  b = ((uint64_t)(hx | (i ^ 0x3ff00000u)) << 32) | (b & 0xffffffffULL);
  k += (int)(i >> 20);
  x = dOfBits(b);
  const double f = x - 1.0;
  const double s = f / (2.0 + f), dk = (double)k;
  const double z = s * s, w = z * z;
  const double t1 = w * (L2 + w * (L4 + w * L6));
  const double t2 = z * (L1 + w * (L3 + w * (L5 + w * L7)));
  const double R = t2 + t1, hfsq = 0.5 * f * f;
  return dk * ln2_hi - ((hfsq - (s * (hfsq + R) + dk * ln2_lo)) - f);
}

// ---- T-typed entry points: fp32 arithmetic evaluates through double and rounds once ------------------------------------------------
RTB_SM_HD double sinS(double x) { return sinD(x); }
RTB_SM_HD float sinS(float x) { return (float)sinD((double)x); }
RTB_SM_HD double cosS(double x) { return cosD(x); }
RTB_SM_HD float cosS(float x) { return (float)cosD((double)x); }
RTB_SM_HD double atan2S(double y, double x) { return atan2D(y, x); }
RTB_SM_HD float atan2S(float y, float x) { return (float)atan2D((double)y, (double)x); }
RTB_SM_HD double asinS(double x) { return asinD(x); }
RTB_SM_HD float asinS(float x) { return (float)asinD((double)x); }
RTB_SM_HD double logS(double x) { return logD(x); }
RTB_SM_HD float logS(float x) { return (float)logD((double)x); }

}  // namespace rtb_strict
