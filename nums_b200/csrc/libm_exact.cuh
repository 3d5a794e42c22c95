// Host-libm-exact elementary functions for the block RNG.
//
// NumPy's ziggurat tail (random_standard_normal, distributions.c) calls the C library's log1p on the host; a device
// log1p that is merely accurate differs from it in the last bit on a fraction of the arguments, and a block of
// normal draws would then not be bit-identical to the NumPy backend's.  glibc's log1p is not correctly rounded, so
// "the" result is defined by its algorithm: the fdlibm log1p with glibc's split polynomial (sysdeps/ieee754/dbl-64/
// s_log1p.c), in the variant the dynamic linker selects on every FMA-capable x86-64 (__log1p_fma: the same source
// compiled with -mfma, glibc >= 2.27 multiarch).  The sequence of roundings below — which products are fused with
// which sums — was taken from the disassembly of that function in glibc 2.39 and checked on the host against
// log1p() for 4*10^8 arguments without a mismatch (tools/check_log1p_exact.c).  Every operation is written with
// an explicit rounding intrinsic so that nvcc can neither fuse nor split anything.
#pragma once
#include <cstdint>

namespace nb_libm {

__device__ __forceinline__ int32_t hi_word(double x) { return __double2hiint(x); }
__device__ __forceinline__ double with_hi(double x, int32_t h) { return __hiloint2double(h, __double2loint(x)); }

__device__ __noinline__ double log1p_glibc_fma(double x) {
  const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
  const double Lp1 = 6.666666666666735130e-01, Lp2 = 3.999999999940941908e-01, Lp3 = 2.857142874366239149e-01,
               Lp4 = 2.222219843214978396e-01, Lp5 = 1.818357216161805012e-01, Lp6 = 1.531383769920937332e-01,
               Lp7 = 1.479819860511658591e-01;
  double hfsq, f, c = 0.0, u;
  int32_t k, hu;
  const int32_t hx = hi_word(x), ax = hx & 0x7fffffff;
  bool direct = false;
  if (hx < 0x3FDA827A) {                                   // x < 0.41422
    if (ax >= 0x3ff00000) return (x == -1.0) ? -__longlong_as_double(0x7ff0000000000000ll)
                                             : __longlong_as_double(0x7ff8000000000000ll);
    if (ax < 0x3e200000) {                                 // |x| < 2^-29
      if (ax < 0x3c900000) return x;
      return __fma_rn(-__dmul_rn(x, x), 0.5, x);
    }
    if (hx > 0 || hx <= (int32_t)0xbfd2bec3) direct = true;   // -0.2929 < x < 0.41422
  } else if (hx >= 0x7ff00000) {
    return __dadd_rn(x, x);
  }
  if (direct) {
    k = 0; f = x; hu = 1;
    hfsq = __dmul_rn(__dmul_rn(x, 0.5), x);
  } else {
    if (hx < 0x43400000) {
      u = __dadd_rn(1.0, x);
      hu = hi_word(u);
      k = (hu >> 20) - 1023;
      c = (k > 0) ? __dsub_rn(1.0, __dsub_rn(u, x)) : __dsub_rn(x, __dsub_rn(u, 1.0));
      c = __ddiv_rn(c, u);
    } else {
      u = x; hu = hi_word(u); k = (hu >> 20) - 1023; c = 0.0;
    }
    hu &= 0x000fffff;
    if (hu < 0x6a09e) {
      u = with_hi(u, hu | 0x3ff00000);
    } else {
      k += 1;
      u = with_hi(u, hu | 0x3fe00000);
      hu = (0x00100000 - hu) >> 2;
    }
    f = __dsub_rn(u, 1.0);
    hfsq = __dmul_rn(__dmul_rn(f, 0.5), f);
    if (hu == 0) {                                         // |f| < 2^-20
      const double dk = (double)k;
      if (f == 0.0) return (k == 0) ? 0.0 : __fma_rn(dk, ln2_hi, __fma_rn(dk, ln2_lo, c));
      const double R = __dmul_rn(__fma_rn(-f, 0.66666666666666666, 1.0), hfsq);
      if (k == 0) return __dsub_rn(f, R);
      return __fma_rn(dk, ln2_hi, -__dsub_rn(__dsub_rn(R, __fma_rn(dk, ln2_lo, c)), f));
    }
  }
  const double s = __ddiv_rn(f, __dadd_rn(f, 2.0));
  const double z = __dmul_rn(s, s);
  const double R2 = __fma_rn(z, Lp3, Lp2), R3 = __fma_rn(z, Lp5, Lp4), R4 = __fma_rn(z, Lp7, Lp6);
  const double z2 = __dmul_rn(z, z), z4 = __dmul_rn(z2, z2), z6 = __dmul_rn(z2, z4);
  double R = __dmul_rn(z2, R2);
  R = __fma_rn(z, Lp1, R);
  R = __fma_rn(z4, R3, R);
  R = __fma_rn(z6, R4, R);
  const double t = __dmul_rn(__dadd_rn(R, hfsq), s);
  if (k == 0) return __dsub_rn(f, __dsub_rn(hfsq, t));
  const double dk = (double)k;
  const double c2 = __fma_rn(dk, ln2_lo, c);
  return __fma_rn(dk, ln2_hi, -__dsub_rn(__dsub_rn(hfsq, __dadd_rn(c2, t)), f));
}

}  // namespace nb_libm
