// glibc_libm.cuh — sinf / cosf / expf / logf / powf exactly as the reference computes them.
//
// The reference links glibc 2.39, whose float sin / cos / exp / log / pow are the
// double-precision table + polynomial algorithms of sysdeps/ieee754/flt-32
// (s_sinf.c, s_cosf.c, s_sincosf.h, e_expf.c, e_logf.c, e_powf.c). They are NOT
// correctly rounded, so "evaluate in double and round" does not reproduce them, and
// CUDA's sinf/cosf/... differ from them in the last bit for 0.1-1 % of arguments —
// enough to send a random walk down another path. These restatements do the same
// IEEE double operations in the same order, with a fused multiply-add exactly where
// the library's FMA build has one (x86-64 picks the *_fma variants on every CPU with
// FMA + AVX2; the placement was read off the disassembly, see scripts/gen_glibc_tables.py),
// so the results agree bit for bit (tests/hostsim --selftest-math; tests/test_cpu_abi.py)
// and every shader replays the reference's random walk on the device.
// sincosf is the same arithmetic as (sinf, cosf) lane by lane (checked in the self-test).
#pragma once

#include "vmath.cuh"

// device-only under nvcc (the tables are __device__ data); plain inline for tests/hostsim
#if defined(__CUDACC__)
#define VPT_TABLE static __device__ const
#define VPT_GFN __device__ __forceinline__
#define VPT_GCALL __device__ __noinline__  // the five entry points are real calls: ~10 call sites
#else
#define VPT_TABLE static const
#define VPT_GFN inline
#define VPT_GCALL inline
#endif

namespace vpt {

#include "glibc_tables.inc"

VPT_GFN unsigned g_asuint(float f) { return (unsigned)f2i(f); }
VPT_GFN float    g_asfloat(unsigned u) { return i2f((int)u); }
VPT_GFN unsigned long long g_asuint64(double d) {
#if defined(__CUDA_ARCH__)
  return (unsigned long long)__double_as_longlong(d);
#else
  union { double d; unsigned long long u; } v; v.d = d; return v.u;
#endif
}
VPT_GFN double g_asdouble(unsigned long long u) {
#if defined(__CUDA_ARCH__)
  return __longlong_as_double((long long)u);
#else
  union { double d; unsigned long long u; } v; v.u = u; return v.d;
#endif
}
// tables live in global memory on the device (divergent indices; L1-resident)
VPT_GFN double g_ld(const double* t, int i) {
#if defined(__CUDA_ARCH__)
  return __ldg(t + i);
#else
  return t[i];
#endif
}

// ---- expf (e_expf.c) ---------------------------------------------------------------
VPT_GCALL float g_expf(float x) {
  double   xd     = (double)x;
  unsigned abstop = (g_asuint(x) >> 20) & 0x7ff;
  if (abstop > 0x42a) {  // |x| >= 88 or nan
    if (g_asuint(x) == 0xff800000u) return 0.0f;
    if (abstop > 0x7f7) return x + x;
    if (x > 0x1.62e42ep6f) return i2f(0x7f800000);  // overflow
    if (x < -0x1.9fe368p6f) return 0.0f;            // underflow
    if (x < -0x1.9d1d9ep6f) return 0x1.4p-75f * 0x1.4p-75f;  // __math_may_uflowf
  }
  const double shift = g_expf_c[0], inv_ln2n = g_expf_c[1];
  double z  = fma(inv_ln2n, xd, shift);
  auto   ki = g_asuint64(z);
  double kd = z - shift;
  double r  = fma(inv_ln2n, xd, -kd);
  auto   t  = g_exp2f_tab[ki & 31] + (ki << 47);
  double zz = fma(g_expf_c[2], r, g_expf_c[3]);
  double r2 = r * r;
  double y  = fma(r, g_expf_c[4], 1.0);
  y         = fma(zz, r2, y);
  y         = y * g_asdouble(t);
  return (float)y;
}

// ---- logf (e_logf.c) ---------------------------------------------------------------
VPT_GCALL float g_logf(float x) {
  unsigned ix = g_asuint(x);
  if (ix == 0x3f800000u) return 0.0f;
  if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) {
    if (ix * 2 == 0) return i2f((int)0xff800000);   // log(0) = -inf
    if (ix == 0x7f800000u) return x;                // log(inf) = inf
    if ((ix & 0x80000000u) || ix * 2 >= 0xff000000u) return i2f(0x7fc00000);  // nan
    ix = g_asuint(x * 0x1p23f);                     // subnormal: normalise
    ix -= 23u << 23;
  }
  unsigned tmp = ix - 0x3f330000u;
  int      i   = (tmp >> 19) & 15;
  int      k   = (int)tmp >> 23;
  unsigned iz  = ix - (tmp & 0xff800000u);
  double invc = g_ld(g_logf_tab, 2 * i), logc = g_ld(g_logf_tab, 2 * i + 1);
  double z  = (double)g_asfloat(iz);
  double y0 = fma((double)k, g_logf_c[0], logc);
  double r  = fma(z, invc, -1.0);
  double y  = fma(g_logf_c[2], r, g_logf_c[3]);
  double r2 = r * r;
  double t  = r + y0;
  y         = fma(r2, g_logf_c[1], y);
  y         = fma(r2, y, t);
  return (float)y;
}

// ---- powf (e_powf.c), x >= 0 or any x with finite non-integer-sensitive use -----------
VPT_GFN int g_checkint(unsigned iy) {  // 0: not an integer, 1: odd, 2: even
  int e = (iy >> 23) & 0xff;
  if (e < 0x7f) return 0;
  if (e > 0x7f + 23) return 2;
  if (iy & ((1u << (0x7f + 23 - e)) - 1)) return 0;
  if (iy & (1u << (0x7f + 23 - e))) return 1;
  return 2;
}
VPT_GFN bool g_zeroinfnan(unsigned ix) { return 2 * ix - 1 >= 2u * 0x7f800000u - 1; }
VPT_GCALL float g_powf(float x, float y) {
  unsigned long long sign_bias = 0;
  unsigned ix = g_asuint(x), iy = g_asuint(y);
  if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u || g_zeroinfnan(iy)) {
    if (g_zeroinfnan(iy)) {
      if (2 * iy == 0) return 1.0f;
      if (ix == 0x3f800000u) return 1.0f;
      if (2 * ix > 2u * 0x7f800000u || 2 * iy > 2u * 0x7f800000u) return x + y;
      if (2 * ix == 2 * 0x3f800000u) return 1.0f;
      if ((2 * ix < 2 * 0x3f800000u) == !(iy & 0x80000000u)) return 0.0f;
      return y * y;
    }
    if (g_zeroinfnan(ix)) {
      float x2 = x * x;
      if ((ix & 0x80000000u) && g_checkint(iy) == 1) x2 = -x2;
      return (iy & 0x80000000u) ? 1 / x2 : x2;
    }
    if (ix & 0x80000000u) {  // finite x < 0
      int yint = g_checkint(iy);
      if (yint == 0) return i2f(0x7fc00000);
      if (yint == 1) sign_bias = 1ull << 16;  // SIGN_BIAS = 1 << (EXP2F_TABLE_BITS + 11)
      ix &= 0x7fffffffu;
    }
    if (ix < 0x00800000u) {  // subnormal x
      ix = g_asuint(x * 0x1p23f);
      ix &= 0x7fffffffu;
      ix -= 23u << 23;
    }
  }
  // log2_inline
  unsigned tmp = ix - 0x3f330000u;
  int      i   = (tmp >> 19) & 15;
  unsigned top = tmp & 0xff800000u;
  unsigned iz  = ix - top;
  int      k   = (int)top >> 23;
  double invc = g_ld(g_log2f_tab, 2 * i), logc = g_ld(g_log2f_tab, 2 * i + 1);
  double z  = (double)g_asfloat(iz);
  double r  = fma(z, invc, -1.0);
  double y0 = logc + (double)k;
  double ya = fma(g_log2f_c[0], r, g_log2f_c[1]);
  double p  = fma(g_log2f_c[2], r, g_log2f_c[3]);
  double r2 = r * r;
  double q  = fma(r, g_log2f_c[4], y0);
  double r4 = r2 * r2;
  q         = fma(r2, p, q);
  double logx  = fma(ya, r4, q);
  double ylogx = (double)y * logx;
  if (((g_asuint64(ylogx) >> 47) & 0xffff) > 0x80be) {  // |y log2 x| >= 126
    if (ylogx > 0x1.fffffffd1d571p+6) return sign_bias ? -i2f(0x7f800000) : i2f(0x7f800000);
    if (ylogx <= -150.0) return sign_bias ? -0.0f : 0.0f;
    if (ylogx < -149.0) return sign_bias ? -0x1p-149f : 0x1p-149f;  // __math_may_uflowf
  }
  // exp2_inline
  const double shift = g_exp2f_c[0];
  double kd = ylogx + shift;
  auto   ki = g_asuint64(kd);
  kd        = kd - shift;
  double rr = ylogx - kd;
  auto   t  = g_exp2f_tab[ki & 31] + ((ki + sign_bias) << 47);
  double zz = fma(g_exp2f_c[1], rr, g_exp2f_c[2]);
  double rr2 = rr * rr;
  double yy = fma(rr, g_exp2f_c[3], 1.0);
  yy        = fma(zz, rr2, yy);
  yy        = yy * g_asdouble(t);
  return (float)yy;
}

// ---- sinf / cosf (s_sinf.c, s_cosf.c, s_sincosf.h) -----------------------------------
// table layout per quadrant set p (14 doubles): sign[4], hpi_inv, hpi, c0, c1, s1, c2, s2, c3, s3, c4
VPT_GFN double g_sin_poly(double x, double x2, const double* p) {
  double x3 = x * x2;
  double s1 = fma(x2, g_ld(p, 12), g_ld(p, 10));
  double x7 = x2 * x3;
  double s  = fma(x3, g_ld(p, 8), x);
  return fma(s1, x7, s);
}
VPT_GFN double g_cos_poly(double x2, const double* p) {
  double x4 = x2 * x2;
  double c1 = fma(x2, g_ld(p, 7), g_ld(p, 6));
  double c2 = fma(x2, g_ld(p, 13), g_ld(p, 11));
  double x6 = x2 * x4;
  double c  = fma(x4, g_ld(p, 9), c1);
  return fma(c2, x6, c);
}
// reduce_fast: |x| < 120
VPT_GFN double g_reduce_fast(double x, int& n) {
  double r = x * g_sincosf_tab[4];
  n        = ((int)r + 0x800000) >> 24;
  return fma(-(double)n, g_sincosf_tab[5], x);
}
// reduce_large: 120 <= |x| < inf (integer arithmetic on 4/pi bits)
VPT_GFN double g_reduce_large(unsigned xi, int& n) {
  const unsigned* arr   = g_inv_pio4 + ((xi >> 26) & 15);
  int             shift = (xi >> 23) & 7;
  xi = (xi & 0xffffff) | 0x800000;
  xi <<= shift;
  unsigned long long res0 = (unsigned)(xi * arr[0]);
  unsigned long long res1 = (unsigned long long)xi * arr[4];
  unsigned long long res2 = (unsigned long long)xi * arr[8];
  res0 = (res2 >> 32) | (res0 << 32);
  res0 += res1;
  unsigned long long nn = (res0 + (1ull << 61)) >> 62;
  res0 -= nn << 62;
  n = (int)nn;
  return (double)(long long)res0 * g_pi63[0];
}
// want_cos = 0: sinf, 1: cosf
VPT_GCALL float g_sincosf(float y, int want_cos) {
  double   x      = (double)y;
  unsigned abstop = (g_asuint(y) >> 20) & 0x7ff;
  const double* p = g_sincosf_tab;
  if (abstop <= 0x3f3) {  // |y| < pi/4
    double x2 = x * x;
    if (abstop <= 0x397) return want_cos ? 1.0f : y;  // |y| < 2^-12
    return (float)(want_cos ? g_cos_poly(x2, p) : g_sin_poly(x, x2, p));
  }
  int n, q;
  if (abstop <= 0x42e) {  // |y| < 120
    x = g_reduce_fast(x, n);
    q = n;
  } else if (abstop <= 0x7f7) {
    unsigned xi = g_asuint(y);
    x           = g_reduce_large(xi, n);
    q           = n + (int)(xi >> 31);  // signs and table include the argument's sign, parity does not
  } else {
    return i2f(0x7fc00000);  // inf / nan
  }
  double s = g_ld(g_sincosf_tab, q & 3);
  if (q & 2) p = g_sincosf_tab + 14;
  bool use_sin = ((n ^ want_cos) & 1) == 0;
  return (float)(use_sin ? g_sin_poly(x * s, x * x, p) : g_cos_poly(x * x, p));
}
VPT_GFN float g_sinf(float x) { return g_sincosf(x, 0); }
VPT_GFN float g_cosf(float x) { return g_sincosf(x, 1); }

// ---- the entry points the shaders call -------------------------------------------------
#if defined(__CUDA_ARCH__) && !defined(VPT_LIBM_CUDA)
VPT_FN float vsin(float a) { return g_sinf(a); }
VPT_FN float vcos(float a) { return g_cosf(a); }
VPT_FN float vlog(float a) { return g_logf(a); }
VPT_FN float vexp(float a) { return g_expf(a); }
VPT_FN float vpow(float a, float b) { return g_powf(a, b); }
#else
VPT_FN float vsin(float a) { return sinf(a); }
VPT_FN float vcos(float a) { return cosf(a); }
VPT_FN float vlog(float a) { return logf(a); }
VPT_FN float vexp(float a) { return expf(a); }
VPT_FN float vpow(float a, float b) { return powf(a, b); }
#endif

}  // namespace vpt
