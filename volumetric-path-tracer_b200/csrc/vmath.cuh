// vmath.cuh — float vector / frame arithmetic of the device path.
//
// Semantics follow the reference's yocto_math.h so that results are
// bit-identical when compiled without FMA contraction (-fmad=false):
//   * scalar min/max/clamp/abs are ternaries (yocto_math.h:1354-1359), i.e.
//     NaN propagates from the second operand — NOT fminf/fmaxf;
//   * dot = (ax*bx + ay*by) + az*bz (yocto_math.h:1610);
//   * vec / float is a true division per component (yocto_math.h:1591);
//   * transform_point = ((x*bx + y*by) + z*bz) + o (yocto_math.h:3097).
// The same header compiles for the host (plain g++) so tests/hostsim can
// single-step the device code against the reference; the product library only
// ever runs it on the GPU.
#pragma once

#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define VPT_FN __host__ __device__ __forceinline__
#define VPT_NOINLINE __host__ __device__ __noinline__
// The big shared pieces of the shade stage (texture fetch, BSDF dispatch, light
// sampling, environment) are real calls on the device: force-inlining them made
// k_wf_shade 27k instructions (430 KB) and ncu showed it stalled on instruction
// fetch (stall_no_instruction 14.7 warps per issue, issue slots 13% busy).
#if defined(VPT_INLINE_ALL)
#define VPT_CALL VPT_FN
#else
#define VPT_CALL __host__ __device__ __noinline__
#endif
#else
#define VPT_FN inline
#define VPT_NOINLINE inline
#define VPT_CALL inline
#endif

// Debug-only loop guards (build with -DVPT_GUARD): a loop that exceeds its bound
// records its id and breaks, so a hang can be localised from the host.
#if defined(VPT_GUARD) && defined(__CUDACC__)
__device__ int vpt_guard_flags[16];
#endif
#if defined(VPT_GUARD) && defined(__CUDA_ARCH__)
#define VPT_GUARD_DECL(name) int name = 0
#define VPT_GUARD_STEP(id, name, limit) \
  if (++name > (limit)) {               \
    vpt_guard_flags[id] = name;         \
    break;                              \
  }
#define VPT_COUNT_ADD(id, v) atomicAdd(&vpt_guard_flags[id], (int)(v))
#define VPT_COUNT_MAX(id, v) atomicMax(&vpt_guard_flags[id], (int)(v))
#else
#define VPT_GUARD_DECL(name)
#define VPT_GUARD_STEP(id, name, limit)
#define VPT_COUNT_ADD(id, v)
#define VPT_COUNT_MAX(id, v)
#endif

// Work counters, compiled only into tests/hostsim (VPT_WORK_COUNTERS): they
// measure the algorithmic work per path that the roofline in bench.py uses.
#if defined(VPT_WORK_COUNTERS) && !defined(__CUDACC__)
#include <atomic>
namespace vpt {
struct work_counters {
  std::atomic<long long> box_tests{0}, tri_tests{0}, inst_transforms{0}, scene_queries{0},
      instance_queries{0}, sdf_scene_evals{0}, sdf_fn_evals{0}, sdf_grid_evals{0},
      trilinear{0}, bounces{0}, lightpdf_calls{0}, texel_reads{0}, iq_box_tests{0}, iq_tri_tests{0},
      cull_tests{0}, culled{0}, cull_wrong{0}, node_visits{0}, pops{0}, stale_pops{0}, leaf_pops{0},
      rounds{0}, wide_visits{0}, wide_slot_tests{0}, stale_node_pops{0}, sdf_skips{0},
      sdf_skips_wrong{0};
};
inline work_counters g_work;
}  // namespace vpt
#define VPT_WORK(field, n) (vpt::g_work.field += (n))
#else
#define VPT_WORK(field, n)
#endif

namespace vpt {

constexpr float pif     = 3.14159265358979323846f;
constexpr float flt_max = 3.402823466e+38f;
constexpr float flt_eps = 1.1920928955078125e-07f;
constexpr int   invalidid = -1;

struct f2 { float x, y; };
struct f3 { float x, y, z; };
struct alignas(16) f4 { float x, y, z, w; };
struct alignas(16) i4 { int x, y, z, w; };

// ---- scalars ---------------------------------------------------------------
VPT_FN float vabs(float a) { return a < 0 ? -a : a; }
VPT_FN float vmin(float a, float b) { return (a < b) ? a : b; }
VPT_FN float vmax(float a, float b) { return (a > b) ? a : b; }
VPT_FN int   vmin(int a, int b) { return (a < b) ? a : b; }
VPT_FN int   vmax(int a, int b) { return (a > b) ? a : b; }
VPT_FN float vclamp(float a, float lo, float hi) { return vmin(vmax(a, lo), hi); }
VPT_FN int   vclamp(int a, int lo, int hi) { return vmin(vmax(a, lo), hi); }
VPT_FN float vsqrt(float a) { return sqrtf(a); }
VPT_FN bool  vfinite(float a) { return isfinite(a); }

VPT_FN int   f2i(float f) {
#if defined(__CUDA_ARCH__)
  return __float_as_int(f);
#else
  union { float f; int i; } u; u.f = f; return u.i;
#endif
}
VPT_FN float i2f(int i) {
#if defined(__CUDA_ARCH__)
  return __int_as_float(i);
#else
  union { float f; int i; } u; u.i = i; return u.f;
#endif
}

// libm entry points. On the device sin / cos / log / exp / pow are glibc 2.39's own float
// algorithms restated operation for operation (glibc_libm.cuh, included at the end of this
// header), so that stochastic shaders replay the reference's random walk bit for bit. The host
// build of this header (tests/hostsim) calls the box's libm, i.e. the real thing.
// -DVPT_LIBM_CUDA falls back to CUDA's float functions (last-bit differences, ~2 % faster).
VPT_FN float vsin(float a);
VPT_FN float vcos(float a);
VPT_FN float vlog(float a);
VPT_FN float vexp(float a);
VPT_FN float vpow(float a, float b);
// atanf / atan2f / acosf: the reference links glibc 2.39, whose float versions
// are the fdlibm float algorithms (sysdeps/ieee754/flt-32/s_atanf.c, e_atan2f.c,
// e_acosf.c; not correctly rounded). They use only + - * / sqrt on floats, so
// restating them here keeps environment lookups (atan2, acos) and GGX sampling
// (atan) on the reference's exact bits, on the device and in tests/hostsim.
VPT_FN float vatan(float x) {
  const float atanhi[4] = {4.6364760399e-01f, 7.8539812565e-01f, 9.8279368877e-01f,
      1.5707962513e+00f};
  const float atanlo[4] = {5.0121582440e-09f, 3.7748947079e-08f, 3.4473217170e-08f,
      7.5497894159e-08f};
  const float aT[11] = {3.3333334327e-01f, -2.0000000298e-01f, 1.4285714924e-01f,
      -1.1111110449e-01f, 9.0908870101e-02f, -7.6918758452e-02f, 6.6610731184e-02f,
      -5.8335702866e-02f, 4.9768779427e-02f, -3.6531571299e-02f, 1.6285819933e-02f};
  int hx = f2i(x), ix = hx & 0x7fffffff, id;
  if (ix >= 0x4c000000) {  // |x| >= 2^25 (glibc 2.39 threshold)
    if (ix > 0x7f800000) return x + x;
    return hx > 0 ? atanhi[3] + atanlo[3] : -atanhi[3] - atanlo[3];
  }
  if (ix < 0x3ee00000) {  // |x| < 0.4375
    if (ix < 0x31000000) return x;  // |x| < 2^-29
    id = -1;
  } else {
    x = vabs(x);
    if (ix < 0x3f980000) {  // |x| < 1.1875
      if (ix < 0x3f300000) {
        id = 0;
        x  = (2.0f * x - 1.0f) / (2.0f + x);
      } else {
        id = 1;
        x  = (x - 1.0f) / (x + 1.0f);
      }
    } else {
      if (ix < 0x401c0000) {  // |x| < 2.4375
        id = 2;
        x  = (x - 1.5f) / (1.0f + 1.5f * x);
      } else {
        id = 3;
        x  = -1.0f / x;
      }
    }
  }
  float z  = x * x;
  float w  = z * z;
  float s1 = z * (aT[0] + w * (aT[2] + w * (aT[4] + w * (aT[6] + w * (aT[8] + w * aT[10])))));
  float s2 = w * (aT[1] + w * (aT[3] + w * (aT[5] + w * (aT[7] + w * aT[9]))));
  if (id < 0) return x - x * (s1 + s2);
  float hi = id == 0 ? atanhi[0] : (id == 1 ? atanhi[1] : (id == 2 ? atanhi[2] : atanhi[3]));
  float lo = id == 0 ? atanlo[0] : (id == 1 ? atanlo[1] : (id == 2 ? atanlo[2] : atanlo[3]));
  z        = hi - ((x * (s1 + s2) - lo) - x);
  return (hx < 0) ? -z : z;
}
VPT_FN float vatan2(float y, float x) {
  const float tiny = 1.0e-30f, pi_o_4 = 7.8539818525e-01f, pi_o_2 = 1.5707963705e+00f,
              pi = 3.1415927410e+00f, pi_lo = -8.7422776573e-08f;
  int hx = f2i(x), ix = hx & 0x7fffffff, hy = f2i(y), iy = hy & 0x7fffffff;
  if (ix > 0x7f800000 || iy > 0x7f800000) return x + y;
  if (hx == 0x3f800000) return vatan(y);
  int m = ((hy >> 31) & 1) | ((hx >> 30) & 2);
  if (iy == 0) {
    if (m < 2) return y;
    return m == 2 ? pi + tiny : -pi - tiny;
  }
  if (ix == 0) return (hy < 0) ? -pi_o_2 - tiny : pi_o_2 + tiny;
  if (ix == 0x7f800000) {
    if (iy == 0x7f800000) {
      if (m == 0) return pi_o_4 + tiny;
      if (m == 1) return -pi_o_4 - tiny;
      if (m == 2) return 3.0f * pi_o_4 + tiny;
      return -3.0f * pi_o_4 - tiny;
    }
    if (m == 0) return 0.0f;
    if (m == 1) return -0.0f;
    if (m == 2) return pi + tiny;
    return -pi - tiny;
  }
  if (iy == 0x7f800000) return (hy < 0) ? -pi_o_2 - tiny : pi_o_2 + tiny;
  int   k = (iy - ix) >> 23;
  float z;
  if (k > 60)
    z = pi_o_2 + 0.5f * pi_lo;
  else if (hx < 0 && k < -60)
    z = 0.0f;
  else
    z = vatan(i2f(f2i(y / x) & 0x7fffffff));  // fabsf: clears the sign of -0 too
  if (m == 0) return z;
  if (m == 1) return i2f(f2i(z) ^ (int)0x80000000);
  if (m == 2) return pi - (z - pi_lo);
  return (z - pi_lo) - pi;
}
VPT_FN float vacos(float x) {
  const float pi = 3.1415925026e+00f, pio2_hi = 1.5707962513e+00f, pio2_lo = 7.5497894159e-08f,
              pS0 = 1.6666667163e-01f, pS1 = -3.2556581497e-01f, pS2 = 2.0121252537e-01f,
              pS3 = -4.0055535734e-02f, pS4 = 7.9153501429e-04f, pS5 = 3.4793309169e-05f,
              qS1 = -2.4033949375e+00f, qS2 = 2.0209457874e+00f, qS3 = -6.8828397989e-01f,
              qS4 = 7.7038154006e-02f;
  int hx = f2i(x), ix = hx & 0x7fffffff;
  if (ix == 0x3f800000) return hx > 0 ? 0.0f : pi + 2.0f * pio2_lo;
  if (ix > 0x3f800000) return (x - x) / (x - x);
  if (ix < 0x3f000000) {  // |x| < 0.5
    if (ix <= 0x23000000) return pio2_hi + pio2_lo;
    float z = x * x;
    float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
    float q = 1.0f + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
    float r = p / q;
    return pio2_hi - (x - (pio2_lo - x * r));
  } else if (hx < 0) {  // x < -0.5
    float z = (1.0f + x) * 0.5f;
    float p = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
    float q = 1.0f + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
    float s = vsqrt(z);
    float r = p / q;
    float w = r * s - pio2_lo;
    return pi - 2.0f * (s + w);
  } else {  // x > 0.5
    float z  = (1.0f - x) * 0.5f;
    float s  = vsqrt(z);
    float df = i2f(f2i(s) & (int)0xfffff000);
    float c  = (z - df * df) / (s + df);
    float p  = z * (pS0 + z * (pS1 + z * (pS2 + z * (pS3 + z * (pS4 + z * pS5)))));
    float q  = 1.0f + z * (qS1 + z * (qS2 + z * (qS3 + z * qS4)));
    float r  = p / q;
    float w  = r * s + c;
    return 2.0f * (df + w);
  }
}
VPT_FN float vfmod(float a, float b) { return fmodf(a, b); }

// ---- f2 --------------------------------------------------------------------
VPT_FN f2 operator+(f2 a, f2 b) { return {a.x + b.x, a.y + b.y}; }
VPT_FN f2 operator-(f2 a, f2 b) { return {a.x - b.x, a.y - b.y}; }
VPT_FN f2 operator*(f2 a, float b) { return {a.x * b, a.y * b}; }
VPT_FN f2 operator-(float a, f2 b) { return {a - b.x, a - b.y}; }
VPT_FN float dot(f2 a, f2 b) { return a.x * b.x + a.y * b.y; }
VPT_FN float length(f2 a) { return vsqrt(dot(a, a)); }

// ---- f3 --------------------------------------------------------------------
VPT_FN f3 operator-(f3 a) { return {-a.x, -a.y, -a.z}; }
VPT_FN f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
VPT_FN f3 operator+(f3 a, float b) { return {a.x + b, a.y + b, a.z + b}; }
VPT_FN f3 operator+(float a, f3 b) { return {a + b.x, a + b.y, a + b.z}; }
VPT_FN f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
VPT_FN f3 operator-(f3 a, float b) { return {a.x - b, a.y - b, a.z - b}; }
VPT_FN f3 operator-(float a, f3 b) { return {a - b.x, a - b.y, a - b.z}; }
VPT_FN f3 operator*(f3 a, f3 b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }
VPT_FN f3 operator*(f3 a, float b) { return {a.x * b, a.y * b, a.z * b}; }
VPT_FN f3 operator*(float a, f3 b) { return {a * b.x, a * b.y, a * b.z}; }
VPT_FN f3 operator/(f3 a, f3 b) { return {a.x / b.x, a.y / b.y, a.z / b.z}; }
VPT_FN f3 operator/(f3 a, float b) { return {a.x / b, a.y / b, a.z / b}; }
VPT_FN f3 operator/(float a, f3 b) { return {a / b.x, a / b.y, a / b.z}; }
VPT_FN f3& operator+=(f3& a, f3 b) { return a = a + b; }
VPT_FN f3& operator*=(f3& a, f3 b) { return a = a * b; }
VPT_FN f3& operator*=(f3& a, float b) { return a = a * b; }
VPT_FN bool operator==(f3 a, f3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; }
VPT_FN bool is_zero(f3 a) { return a.x == 0 && a.y == 0 && a.z == 0; }

VPT_FN float dot(f3 a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
VPT_FN f3 cross(f3 a, f3 b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
VPT_FN float length(f3 a) { return vsqrt(dot(a, a)); }
VPT_FN f3 normalize(f3 a) {
  auto l = length(a);
  return (l != 0) ? a / l : a;
}
VPT_FN float distance_squared(f3 a, f3 b) { return dot(a - b, a - b); }
VPT_FN f3 orthonormalize(f3 a, f3 b) { return normalize(a - b * dot(a, b)); }
VPT_FN f3 reflect(f3 w, f3 n) { return -w + 2 * dot(n, w) * n; }
VPT_FN f3 refract(f3 w, f3 n, float inv_eta) {
  auto cosine = dot(n, w);
  auto k      = 1 + inv_eta * inv_eta * (cosine * cosine - 1);
  if (k < 0) return {0, 0, 0};
  return -w * inv_eta + (inv_eta * cosine - vsqrt(k)) * n;
}
VPT_FN f3 vmax(f3 a, float b) { return {vmax(a.x, b), vmax(a.y, b), vmax(a.z, b)}; }
VPT_FN f3 vmin(f3 a, f3 b) { return {vmin(a.x, b.x), vmin(a.y, b.y), vmin(a.z, b.z)}; }
VPT_FN f3 vmax(f3 a, f3 b) { return {vmax(a.x, b.x), vmax(a.y, b.y), vmax(a.z, b.z)}; }
VPT_FN f3 vclamp(f3 a, float lo, float hi) {
  return {vclamp(a.x, lo, hi), vclamp(a.y, lo, hi), vclamp(a.z, lo, hi)};
}
VPT_FN f3 lerp(f3 a, f3 b, float u) { return a * (1 - u) + b * u; }
VPT_FN float max3(f3 a) { return vmax(vmax(a.x, a.y), a.z); }
VPT_FN float min3(f3 a) { return vmin(vmin(a.x, a.y), a.z); }
VPT_FN float sum(f3 a) { return a.x + a.y + a.z; }
VPT_FN float mean(f3 a) { return sum(a) / 3; }
VPT_FN f3 vabs(f3 a) { return {vabs(a.x), vabs(a.y), vabs(a.z)}; }
VPT_FN f3 vsqrt(f3 a) { return {vsqrt(a.x), vsqrt(a.y), vsqrt(a.z)}; }
VPT_FN f3 vexp(f3 a) { return {vexp(a.x), vexp(a.y), vexp(a.z)}; }
VPT_FN f3 vlog(f3 a) { return {vlog(a.x), vlog(a.y), vlog(a.z)}; }
VPT_FN bool vfinite(f3 a) { return vfinite(a.x) && vfinite(a.y) && vfinite(a.z); }
VPT_FN float comp(f3 a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

// ---- f4 --------------------------------------------------------------------
VPT_FN f4 operator+(f4 a, f4 b) { return {a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w}; }
VPT_FN f4 operator*(f4 a, float b) { return {a.x * b, a.y * b, a.z * b, a.w * b}; }
VPT_FN f3 xyz(f4 a) { return {a.x, a.y, a.z}; }

// ---- frames (yocto frame3f: columns x, y, z and origin o) --------------------
struct frame3 { f3 x, y, z, o; };
struct mat3 { f3 x, y, z; };

VPT_FN f3 transform_point(const frame3& a, f3 b) {
  return a.x * b.x + a.y * b.y + a.z * b.z + a.o;
}
VPT_FN f3 transform_vector(const frame3& a, f3 b) {
  return a.x * b.x + a.y * b.y + a.z * b.z;
}
VPT_FN f3 transform_direction(const frame3& a, f3 b) {
  return normalize(transform_vector(a, b));
}
VPT_FN f3 operator*(const mat3& a, f3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
VPT_FN mat3 transpose(const mat3& a) {
  return {{a.x.x, a.y.x, a.z.x}, {a.x.y, a.y.y, a.z.y}, {a.x.z, a.y.z, a.z.z}};
}
VPT_FN f3 transform_direction(const mat3& a, f3 b) { return normalize(a * b); }

// basis_fromz (yocto_math.h:2811-2820)
VPT_FN mat3 basis_fromz(f3 v) {
  auto z    = normalize(v);
  auto sign = copysignf(1.0f, z.z);
  auto a    = -1.0f / (sign + z.z);
  auto b    = z.x * z.y * a;
  auto x    = f3{1.0f + sign * z.x * z.x * a, sign * b, -sign * z.x};
  auto y    = f3{b, sign + z.y * z.y * a, -z.y};
  return {x, y, z};
}

// Frame inverses are evaluated once on the host at upload time with exactly
// these expressions (yocto_math.h:2803-2808, :2948-2956), compiled by the host
// compiler without FMA contraction.
VPT_FN frame3 inverse_frame(const frame3& a, bool non_rigid) {
  mat3 minv;
  if (non_rigid) {
    auto adj = transpose(mat3{cross(a.y, a.z), cross(a.z, a.x), cross(a.x, a.y)});
    auto det = dot(a.x, cross(a.y, a.z));
    auto s   = 1 / det;
    minv     = {adj.x * s, adj.y * s, adj.z * s};
  } else {
    minv = transpose(mat3{a.x, a.y, a.z});
  }
  auto o = -(minv * a.o);
  return {minv.x, minv.y, minv.z, o};
}

}  // namespace vpt

#include "glibc_libm.cuh"
