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
#else
#define VPT_FN inline
#define VPT_NOINLINE inline
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

// libm entry points. glibc's float functions are (near-)correctly rounded;
// rounding a double-precision evaluation to float reproduces them in all but
// vanishingly rare cases, which is what the parity build (VPT_LIBM_F64) uses.
// The throughput build calls CUDA's float functions.
#if defined(__CUDA_ARCH__) && defined(VPT_LIBM_F64)
VPT_FN float vsin(float a) { return (float)sin((double)a); }
VPT_FN float vcos(float a) { return (float)cos((double)a); }
VPT_FN float vlog(float a) { return (float)log((double)a); }
VPT_FN float vexp(float a) { return (float)exp((double)a); }
VPT_FN float vatan(float a) { return (float)atan((double)a); }
VPT_FN float vpow(float a, float b) { return (float)pow((double)a, (double)b); }
#else
VPT_FN float vsin(float a) { return sinf(a); }
VPT_FN float vcos(float a) { return cosf(a); }
VPT_FN float vlog(float a) { return logf(a); }
VPT_FN float vexp(float a) { return expf(a); }
VPT_FN float vatan(float a) { return atanf(a); }
VPT_FN float vpow(float a, float b) { return powf(a, b); }
#endif
// Environment lookups happen once per path; evaluating them in double keeps
// the deterministic shaders' miss pixels on the reference's bits.
#if defined(__CUDA_ARCH__)
VPT_FN float vatan2(float a, float b) { return (float)atan2((double)a, (double)b); }
VPT_FN float vacos(float a) { return (float)acos((double)a); }
#else
VPT_FN float vatan2(float a, float b) { return atan2f(a, b); }
VPT_FN float vacos(float a) { return acosf(a); }
#endif
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
