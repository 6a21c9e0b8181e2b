// Shared device helpers for the MAESTRO-B200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#define MB_DTYPE_F32 0
#define MB_DTYPE_BF16 1

#define MB_CHECK_LAUNCH()                                  \
  do {                                                     \
    cudaError_t e__ = cudaGetLastError();                  \
    if (e__ != cudaSuccess) return static_cast<int>(e__);  \
  } while (0)

typedef __nv_bfloat16 bf16;

static inline int mb_ceil_div(long long a, long long b) { return static_cast<int>((a + b - 1) / b); }

// ---------------------------------------------------------------- programmatic dependent launch
// The step's hot kernels are launched with the programmatic-stream-serialisation attribute (mb_launch): their CTAs may
// become resident while the previous kernel of the stream drains, run their prologue (barrier init, TMEM allocation,
// descriptor prefetch) and then block in pdl_wait() until the previous grid has completed and its writes are
// visible. Rules: no global-memory access before pdl_wait(); pdl_launch() only after the CTA holds every resource it
// may block on (TMEM columns).
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

#include <stdlib.h>
static inline bool mb_pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("MB_PDL");          // MB_PDL=0: plain stream-ordered launches (A/B measurements)
    return !(e && e[0] == '0');
  }();
  return on;
}

// <<<grid, block, smem, st>>> with the PDL attribute; the kernel MUST call pdl_wait() before touching global memory.
template <typename... KArgs, typename... Args>
static inline cudaError_t mb_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                    Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = mb_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---------------------------------------------------------------- vector access
// VecIO<T>: 16-byte global access -> float registers (8 bf16 or 4 fp32 per access).
template <typename T> struct VecIO;

template <> struct VecIO<float> {
  static constexpr int N = 4;
  typedef float4 Raw;   // one 16-byte access, still packed
  __device__ __forceinline__ static Raw load_raw(const float* p) { return *reinterpret_cast<const float4*>(p); }
  __device__ __forceinline__ static Raw zero_raw() { return make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ __forceinline__ static void unpack(const Raw& t, float* v) { v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
  __device__ __forceinline__ static void load(const float* p, float* v) {
    float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ __forceinline__ static void store(float* p, const float* v) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
};

template <> struct VecIO<bf16> {
  static constexpr int N = 8;
  typedef uint4 Raw;
  __device__ __forceinline__ static Raw load_raw(const bf16* p) { return *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ static Raw zero_raw() { return make_uint4(0, 0, 0, 0); }
  __device__ __forceinline__ static void unpack(const Raw& t, float* v) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ __forceinline__ static void load(const bf16* p, float* v) {
    uint4 t = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ __forceinline__ static void store(bf16* p, const float* v) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

__device__ __forceinline__ float to_f32(float x) { return x; }
__device__ __forceinline__ float to_f32(bf16 x) { return __bfloat162float(x); }
template <typename T> __device__ __forceinline__ T from_f32(float x);
template <> __device__ __forceinline__ float from_f32<float>(float x) { return x; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float x) { return __float2bfloat16_rn(x); }

// ---------------------------------------------------------------- packed fp32 (sm_100: FFMA2 / FADD2 / FMUL2)
// Two fp32 operations per issued instruction.  The bandwidth-bound kernels here are bound by issue slots before they
// are bound by HBM (ncu: 55-63 % issue-active at 0.52-0.57 of the copy bandwidth), so halving the FP instruction count
// is what moves them.
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) {
  unsigned long long d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 add2(float2 a, float2 b) {
  unsigned long long d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 mul2(float2 a, float2 b) {
  unsigned long long d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(*reinterpret_cast<unsigned long long*>(&a)),
      "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 splat2(float x) { return make_float2(x, x); }

// 16 bytes -> float2 pairs and back (bf16: 4 pairs per access, fp32: 2 pairs)
template <typename T> struct Vec2IO;
template <> struct Vec2IO<float> {
  static constexpr int NP = 2;
  __device__ __forceinline__ static void unpack(const float4& t, float2* v) { v[0] = make_float2(t.x, t.y); v[1] = make_float2(t.z, t.w); }
  __device__ __forceinline__ static float4 pack(const float2* v) { return make_float4(v[0].x, v[0].y, v[1].x, v[1].y); }
};
template <> struct Vec2IO<bf16> {
  static constexpr int NP = 4;
  __device__ __forceinline__ static void unpack(const uint4& t, float2* v) {
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) v[i] = make_float2(__uint_as_float(w[i] << 16), __uint_as_float(w[i] & 0xffff0000u));
  }
  __device__ __forceinline__ static uint4 pack(const float2* v) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(v[i].x, v[i].y);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// ---------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// both components of a float2 summed over the warp (3 instructions per step instead of 4)
__device__ __forceinline__ float2 warp_sum2(float2 v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1)
    v = add2(v, make_float2(__shfl_xor_sync(0xffffffffu, v.x, o), __shfl_xor_sync(0xffffffffu, v.y, o)));
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---------------------------------------------------------------- math
// 2^x on the MUFU pipe, flush-to-zero form (one instruction, no denormal fix-up code around it).
__device__ __forceinline__ float fast_exp2(float x) {
#ifdef MB_DIAG_NO_MUFU      // timing experiment only (wrong results): is the kernel bound by the MUFU pipe?
  return fmaf(x, 0.001f, 1.0f);
#else
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
#endif
}
__device__ __forceinline__ float gelu_erf(float x) {
  return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float gelu_erf_grad(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * __expf(-0.5f * x * x);
  return cdf + x * pdf;
}

// Fast exact-form GELU for the bf16 tensor-core epilogues: erf by Abramowitz-Stegun 7.1.26 (|err| < 1.5e-7, far
// below bf16 resolution) sharing one exp between erf and the Gaussian term of the derivative.
// q = erfc(|x| / sqrt 2) / 2 (the tail mass beyond |x|) and e = exp(-x^2 / 2); 11 instructions, 2 MUFU (ftz forms:
// no denormal fix-up code).  Coefficients are the A-S ones halved.
__device__ __forceinline__ void gelu_fast_terms(float x, float& q, float& e) {
  const float w = x * (-0.72134752044448170368f * x);      // -x^2 / 2 * log2(e)
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(w));
  const float d = fmaf(fabsf(x), 0.23164189887f, 1.0f);    // 1 + p * |x| / sqrt 2, p = 0.3275911
  float t;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(d));
  float p = fmaf(t, 0.5307027145f, -0.7265760135f);
  p = fmaf(p, t, 0.7107068705f);
  p = fmaf(p, t, -0.142248368f);
  p = fmaf(p, t, 0.127414796f);
  q = (p * t) * e;
}
// gelu(x) = x * Phi(x) = max(x, 0) - |x| * q
__device__ __forceinline__ float gelu_fast(float x) {
  float q, e;
  gelu_fast_terms(x, q, e);
  return fmaxf(x, 0.0f) - fabsf(q * x);
}
// gelu(x) = x * Phi(x) and gelu'(x) = Phi(x) + x * phi(x) from one set of terms (forward epilogue: the derivative is
// what the backward needs saved).  Abramowitz-Stegun 26.2.17 (|error| < 7.5e-8): phi = exp(-x^2 / 2) / sqrt(2 pi) (the
// 1 / sqrt(2 pi) is folded into the exponent), t = 1 / (1 + 0.2316419 |x|), tail mass q = phi * t * poly(t),
// Phi = 1/2 + copysign(1/2 - q, x).  15 FP32 instructions + 2 MUFU (ftz forms) per element.
__device__ __forceinline__ float gelu_fast_both(float x, float& grad) {
  const float w = fmaf(x * -0.72134752044448170368f, x, -1.32574806473616f);   // log2(phi(x))
  float phi, t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(phi) : "f"(w));
  const float d = fmaf(fabsf(x), 0.2316419f, 1.0f);
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(d));
  float p = fmaf(t, 1.330274429f, -1.821255978f);
  p = fmaf(p, t, 1.781477937f);
  p = fmaf(p, t, -0.356563782f);
  p = fmaf(p, t, 0.319381530f);
  const float q = (p * t) * phi;
  const float cdf = 0.5f + copysignf(0.5f - q, x);
  grad = fmaf(x, phi, cdf);
  return x * cdf;
}
// The same for two values at once on the packed fp32 pipe (FFMA2 / FMUL2 / FADD2): identical operations and rounding
// per element as gelu_fast_both, 12.5 instead of 17 issue slots per element (the MUFU, abs and copysign stay scalar).
__device__ __forceinline__ float2 gelu_fast_both2(float2 x, float2& grad) {
  const float2 w = fma2(mul2(x, splat2(-0.72134752044448170368f)), x, splat2(-1.32574806473616f));
  float2 phi, t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(phi.x) : "f"(w.x));
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(phi.y) : "f"(w.y));
  const float2 d = fma2(make_float2(fabsf(x.x), fabsf(x.y)), splat2(0.2316419f), splat2(1.0f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.x) : "f"(d.x));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t.y) : "f"(d.y));
  float2 p = fma2(t, splat2(1.330274429f), splat2(-1.821255978f));
  p = fma2(p, t, splat2(1.781477937f));
  p = fma2(p, t, splat2(-0.356563782f));
  p = fma2(p, t, splat2(0.319381530f));
  const float2 q = mul2(mul2(p, t), phi);
  const float2 hq = add2(splat2(0.5f), make_float2(-q.x, -q.y));
  const float2 cdf = add2(splat2(0.5f), make_float2(copysignf(hq.x, x.x), copysignf(hq.y, x.y)));
  grad = fma2(x, phi, cdf);
  return mul2(x, cdf);
}
// gelu'(x) = Phi(x) + x * phi(x),  Phi(x) = 1/2 + copysign(1/2 - q, x)
__device__ __forceinline__ float gelu_fast_grad(float x) {
  float q, e;
  gelu_fast_terms(x, q, e);
  const float cdf = 0.5f + copysignf(0.5f - q, x);
  return fmaf(x * 0.39894228040143267794f, e, cdf);
}

// Date-encoding channel value: channels [dim-8, dim) of a token are
// [diff, diff, diff, diff, sin_doy, cos_doy, sin_hour, cos_hour] * fac  (reference
// maestro/layers/utils.py:128-173); feat = [diff, sin_doy, cos_doy, sin_hour, cos_hour].
__device__ __forceinline__ float date_channel(const float* feat5, int k /*0..date_dim-1*/, int date_dim, float fac) {
  const int lead = date_dim - 4;
  const float v = (k < lead) ? feat5[0] : feat5[1 + (k - lead)];
  return v * fac;
}

#define MB_DISPATCH_DTYPE(dtype, T, ...)                   \
  if ((dtype) == MB_DTYPE_F32) { typedef float T; __VA_ARGS__; } \
  else if ((dtype) == MB_DTYPE_BF16) { typedef bf16 T; __VA_ARGS__; } \
  else { return -1; }
