// fp32-accumulate CUDA-core GEMM with arbitrary operand strides and a two-level batch.
// This is the arithmetic of the *fp32 check mode* (north star: 1e-4 parity with the reference) and of the
// few contractions too small / oddly shaped for the tcgen05 path.  C[m, n] = alpha * sum_k A(m, k) * B(n, k)
// followed by the shared epilogue (bias, exact GELU with saved derivative, GELU gradient, residual).
#include "common.cuh"
#include "gemm_params.h"

constexpr int TM = 64, TN = 64, TK = 16;

template <typename T, typename TC>
__global__ void __launch_bounds__(256) gemm_simt_kernel(GemmParams P) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int z = blockIdx.z;
  const int z1 = z / P.nb2, z2 = z % P.nb2;
  const T* A = static_cast<const T*>(P.A) + z1 * P.a_s1 + z2 * P.a_s2;
  const T* B = static_cast<const T*>(P.B) + z1 * P.b_s1 + z2 * P.b_s2;
  const long long coff = z1 * P.c_s1 + z2 * P.c_s2;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  // loader mapping: pick the thread->element map that is contiguous along the unit-stride axis
  const bool a_kfast = P.sak == 1, b_kfast = P.sbk == 1;
  for (int k0 = 0; k0 < P.K; k0 += TK) {
#pragma unroll
    for (int it = 0; it < (TM * TK) / 256; ++it) {
      const int e = threadIdx.x + it * 256;
      const int mm = a_kfast ? e / TK : e % TM;
      const int kk = a_kfast ? e % TK : e / TM;
      const int m = m0 + mm, k = k0 + kk;
      As[kk][mm] = (m < P.M && k < P.K) ? to_f32(A[m * P.sam + k * P.sak]) : 0.f;
    }
#pragma unroll
    for (int it = 0; it < (TN * TK) / 256; ++it) {
      const int e = threadIdx.x + it * 256;
      const int nn = b_kfast ? e / TK : e % TN;
      const int kk = b_kfast ? e % TK : e / TN;
      const int n = n0 + nn, k = k0 + kk;
      Bs[kk][nn] = (n < P.N && k < P.K) ? to_f32(B[n * P.sbn + k * P.sbk]) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

  TC* C = static_cast<TC*>(P.C) + coff;
  const T* res = P.residual != nullptr ? static_cast<const T*>(P.residual) + coff : nullptr;
  T* aux = P.aux != nullptr ? static_cast<T*>(P.aux) + coff : nullptr;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= P.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= P.N) continue;
      const size_t o = static_cast<size_t>(m) * P.ldc + n;
      float v = acc[i][j] * P.alpha;
      if (P.bias != nullptr) v += P.bias[n];
      if (P.flags & GEMM_GELU) {
        aux[o] = from_f32<T>(gelu_erf_grad(v));  // saved for the backward: gelu'(pre-activation)
        v = gelu_erf(v);
      }
      if (P.flags & GEMM_DGELU) v *= to_f32(aux[o]);
      if (res != nullptr) v += to_f32(res[o]);
      if (P.flags & GEMM_ACCUM) v += to_f32(C[o]);
      C[o] = from_f32<TC>(v);
    }
  }
}

extern "C" int mb_gemm_simt(const GemmParams* params, int dtype, void* stream) {
  const GemmParams& P = *params;
  if (P.M == 0 || P.N == 0) return 0;
  const int nb = P.nb1 * P.nb2;
  dim3 grid(mb_ceil_div(P.N, TN), mb_ceil_div(P.M, TM), nb);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool of32 = (P.flags & GEMM_OUT_F32) != 0;
  if (dtype == MB_DTYPE_F32) gemm_simt_kernel<float, float><<<grid, 256, 0, st>>>(P);
  else if (dtype == MB_DTYPE_BF16 && of32) gemm_simt_kernel<bf16, float><<<grid, 256, 0, st>>>(P);
  else if (dtype == MB_DTYPE_BF16) gemm_simt_kernel<bf16, bf16><<<grid, 256, 0, st>>>(P);
  else return -1;
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ row softmax (check-mode attention)
// p[r, :] = softmax(scale * s[r, :]) over n columns; s fp32 in, p (T) out.  One warp per row.
template <typename T>
__global__ void __launch_bounds__(256) softmax_rows_kernel(const float* __restrict__ s, T* __restrict__ p,
                                                           long long rows, int n, float scale) {
  const long long row = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* sr = s + row * n;
  float mx = -INFINITY;
  for (int i = lane; i < n; i += 32) mx = fmaxf(mx, sr[i] * scale);
  mx = warp_max(mx);
  float sum = 0.f;
  for (int i = lane; i < n; i += 32) sum += expf(sr[i] * scale - mx);
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
  for (int i = lane; i < n; i += 32) p[row * n + i] = from_f32<T>(expf(sr[i] * scale - mx) * inv);
}

// ds[r, :] = scale * p[r, :] * (dp[r, :] - sum_j dp[r, j] * p[r, j])   (dp fp32 in, overwritten in place by ds)
template <typename T>
__global__ void __launch_bounds__(256) softmax_bwd_rows_kernel(const T* __restrict__ p, float* __restrict__ dp,
                                                               T* __restrict__ ds, long long rows, int n, float scale) {
  const long long row = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float dot = 0.f;
  for (int i = lane; i < n; i += 32) dot += dp[row * n + i] * to_f32(p[row * n + i]);
  dot = warp_sum(dot);
  for (int i = lane; i < n; i += 32) {
    const float pv = to_f32(p[row * n + i]);
    ds[row * n + i] = from_f32<T>(scale * pv * (dp[row * n + i] - dot));
  }
}

extern "C" int mb_softmax_rows(const float* s, void* p, int dtype, long long rows, int n, float scale, void* stream) {
  if (rows == 0) return 0;
  const int blocks = mb_ceil_div(rows * 32, 256);
  MB_DISPATCH_DTYPE(dtype, T,
                    (softmax_rows_kernel<T><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        s, static_cast<T*>(p), rows, n, scale)));
  MB_CHECK_LAUNCH();
  return 0;
}

extern "C" int mb_softmax_bwd_rows(const void* p, float* dp, void* ds, int dtype, long long rows, int n, float scale,
                                   void* stream) {
  if (rows == 0) return 0;
  const int blocks = mb_ceil_div(rows * 32, 256);
  MB_DISPATCH_DTYPE(dtype, T,
                    (softmax_bwd_rows_kernel<T><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(p), dp, static_cast<T*>(ds), rows, n, scale)));
  MB_CHECK_LAUNCH();
  return 0;
}
