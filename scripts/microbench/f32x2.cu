// Microbenchmark: issue rate of the packed FP32 instructions of sm_100a (FADD2 / FMUL2 / FFMA2 from PTX
// add / sub / mul / fma.rn.f32x2) against their scalar forms, and a bit-exactness check of the packed
// results against scalar IEEE arithmetic. Measured on B200: un-fused sub + mul
// streams run at 69.6 Tflop/s packed against 36.2 scalar, every half equals the scalar result over
// 2^25 random bit patterns - but ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 whatever
// -fmad says (seen in the product's kernels; an exact packed add has to be written fma(a, 1, b) with
// the unit read from constant memory), and neither the four-slot slab test nor the vec3 operators
// got faster with it: the traversal kernels are not bound by FP issue.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -O3 -o f32x2 f32x2.cu
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void unpk(u64 v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ u64 sub2(u64 a, u64 b) { u64 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ u64 add2_naive(u64 a, u64 b) { u64 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
static __constant__ float units[2] = {1.0f, -1.0f};
__device__ __forceinline__ u64 add2(u64 a, u64 b) { u64 r; float one = units[0]; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(pk(one, one)), "l"(b)); return r; }
__device__ __forceinline__ u64 mul2(u64 a, u64 b) { u64 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }

template <int PACKED>
__global__ void k_rate(float* out, float o, float d, int iters) {
  float a[8], b[8];
  for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 0.001f + i, b[i] = a[i] + 0.5f;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++) {
      if (PACKED) {
        u64 v = mul2(sub2(pk(a[i], b[i]), pk(o, o)), pk(d, d));
        unpk(v, a[i], b[i]);
      } else {
        a[i] = (a[i] - o) * d, b[i] = (b[i] - o) * d;
      }
    }
  }
  float s = 0;
  for (int i = 0; i < 8; i++) s += a[i] + b[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void k_exact(const float* x, const float* y, int n, int* bad) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1 >= n) return;
  float a = x[i], b = x[i + 1], o = y[i], d = y[i + 1];
  float p0, p1;
  unpk(mul2(sub2(pk(a, b), pk(o, o)), pk(d, d)), p0, p1);
  float s0 = __fmul_rn(__fsub_rn(a, o), d), s1 = __fmul_rn(__fsub_rn(b, o), d);
  if (__float_as_uint(p0) != __float_as_uint(s0) && !(p0 != p0 && s0 != s0)) atomicAdd(bad, 1);
  if (__float_as_uint(p1) != __float_as_uint(s1) && !(p1 != p1 && s1 != s1)) atomicAdd(bad, 1);
  // mul then add: ptxas contracts the plain packed pair into FFMA2 (bad[1]); fma(a, 1, b) with the unit
  // from constant memory stays a separate, exact addition (bad[2])
  float q0, q1, r0, r1;
  unpk(add2_naive(mul2(pk(a, b), pk(o, o)), pk(d, d)), q0, q1);
  unpk(add2(mul2(pk(a, b), pk(o, o)), pk(d, d)), r0, r1);
  float t0 = __fadd_rn(__fmul_rn(a, o), d), t1 = __fadd_rn(__fmul_rn(b, o), d);
  if ((__float_as_uint(q0) != __float_as_uint(t0) && !(q0 != q0 && t0 != t0)) ||
      (__float_as_uint(q1) != __float_as_uint(t1) && !(q1 != q1 && t1 != t1))) atomicAdd(bad + 1, 1);
  if ((__float_as_uint(r0) != __float_as_uint(t0) && !(r0 != r0 && t0 != t0)) ||
      (__float_as_uint(r1) != __float_as_uint(t1) && !(r1 != r1 && t1 != t1))) atomicAdd(bad + 2, 1);
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * sizeof(float));
  cudaEvent_t e0, e1; cudaEventCreate(&e0), cudaEventCreate(&e1);
  for (int packed = 0; packed < 2; packed++) {
    for (int rep = 0; rep < 2; rep++) {
      cudaEventRecord(e0);
      if (packed) k_rate<1><<<148 * 8, 256>>>(out, 0.25f, 1.0001f, 20000);
      else k_rate<0><<<148 * 8, 256>>>(out, 0.25f, 1.0001f, 20000);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double flop = 148.0 * 8 * 256 * 20000.0 * 8 * 4;
      if (rep) printf("%s: %.3f ms, %.2f Tflop/s (sub + mul, no fma)\n", packed ? "packed f32x2" : "scalar", ms, flop / ms / 1e9);
    }
  }
  int n = 1 << 24; float *x, *y; int* bad;
  cudaMalloc(&x, n * 4), cudaMalloc(&y, n * 4), cudaMalloc(&bad, 12), cudaMemset(bad, 0, 12);
  unsigned* h = (unsigned*)malloc(n * 4);
  unsigned long long s = 88172645463325252ull;
  for (int pass = 0; pass < 2; pass++) {
    for (int i = 0; i < n; i++) { s ^= s << 13, s ^= s >> 7, s ^= s << 17; h[i] = (unsigned)(s >> 16); }  // any bit pattern
    cudaMemcpy(pass ? y : x, h, n * 4, cudaMemcpyHostToDevice);
  }
  k_exact<<<n / 256, 256>>>(x, y, n, bad);
  int hb[3] = {-1, -1, -1}; cudaMemcpy(hb, bad, 12, cudaMemcpyDeviceToHost);
  printf("packed vs scalar on %d random bit patterns (denormals, infinities, NaNs included): (a - o) * d %d differ; "
         "a * o + d with add.rn.f32x2 %d differ (ptxas may contract this pair); with fma(a, 1, b) %d differ\n", n, hb[0], hb[1], hb[2]);
  return hb[0] != 0 || hb[2] != 0;
}
