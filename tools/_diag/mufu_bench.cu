// Throughput of exp2 variants on sm_100a: MUFU f32, MUFU f16x2, FMA-pipe polynomial (scalar and packed f32x2).
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
__device__ __forceinline__ float ex2f(float x) { float y; asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t ex2h2(uint32_t x) { uint32_t y; asm volatile("ex2.approx.f16x2 %0, %1;" : "=r"(y) : "r"(x)); return y; }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  unsigned long long d;
  asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  unsigned long long d;
  asm volatile("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float poly1(float x) {
  x = fmaxf(x, -125.f);
  const float t = x + 12582912.f;
  const float f = x - (t - 12582912.f);
  float p = fmaf(0.055171654f, f, 0.24261113f);
  p = fmaf(p, f, 0.69326097f);
  p = fmaf(p, f, 0.99992806f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
__device__ __forceinline__ float2 poly2(float2 x) {
  x.x = fmaxf(x.x, -125.f); x.y = fmaxf(x.y, -125.f);
  const float2 mg = make_float2(12582912.f, 12582912.f), nmg = make_float2(-12582912.f, -12582912.f);
  const float2 t = add2(x, mg);
  const float2 n = add2(t, nmg);
  const float2 f = add2(x, make_float2(-n.x, -n.y));
  float2 p = fma2(make_float2(0.055171654f, 0.055171654f), f, make_float2(0.24261113f, 0.24261113f));
  p = fma2(p, f, make_float2(0.69326097f, 0.69326097f));
  p = fma2(p, f, make_float2(0.99992806f, 0.99992806f));
  return make_float2(__int_as_float(__float_as_int(p.x) + (__float_as_int(t.x) << 23)),
                     __int_as_float(__float_as_int(p.y) + (__float_as_int(t.y) << 23)));
}
template <int MODE> __global__ void k(float* out, int iters, float seed) {
  float a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = seed - i * 0.01f - threadIdx.x * 1e-4f;
  for (int it = 0; it < iters; ++it) {
    if (MODE == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = ex2f(a[i]) - 1.5f;
    } else if (MODE == 1) {
#pragma unroll
      for (int i = 0; i < 8; i += 2) {
        __half2 h = __floats2half2_rn(a[i], a[i + 1]);
        uint32_t r = ex2h2(*reinterpret_cast<uint32_t*>(&h));
        float2 f = __half22float2(*reinterpret_cast<__half2*>(&r));
        a[i] = f.x - 1.5f; a[i + 1] = f.y - 1.5f;
      }
    } else if (MODE == 2) {
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = poly1(a[i]) - 1.5f;
    } else if (MODE == 3) {
#pragma unroll
      for (int i = 0; i < 8; i += 2) { float2 r = poly2(make_float2(a[i], a[i + 1])); a[i] = r.x - 1.5f; a[i + 1] = r.y - 1.5f; }
    } else {   // 3 of 8 by polynomial, 5 by MUFU
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = ((i % 8) < 3 ? poly1(a[i]) : ex2f(a[i])) - 1.5f;
    }
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE> void run(const char* name, float* out) {
  const int blocks = 148 * 8, threads = 256, iters = 4096;
  k<MODE><<<blocks, threads>>>(out, 16, -0.3f);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaEventRecord(e0);
  k<MODE><<<blocks, threads>>>(out, iters, -0.3f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double n = double(blocks) * threads * iters * 8;
  printf("%-28s %8.3f ms  %8.1f Gexp/s  (%.2f exp/clk/SM at 1.9 GHz)\n", name, ms, n / ms / 1e6, n / ms / 1e6 / 148 / 1.9);
}
int main() {
  float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  run<0>("MUFU ex2.f32", out); run<1>("MUFU ex2.f16x2 (+cvt)", out); run<2>("poly deg3 scalar", out);
  run<3>("poly deg3 packed f32x2", out); run<4>("mix 3/8 poly + 5/8 MUFU", out);
  return 0;
}
