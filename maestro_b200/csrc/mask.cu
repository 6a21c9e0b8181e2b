// Mask index math and token gather / scatter of the MAE (reference maestro/ssl/mae.py:228-287).
//   mask_seq:   keys = noise * (1 - struct); the num_masked smallest keys are masked (ties: lower position
//               first = the stable order), both index halves sorted ascending.
//   unmask_seq: decoder input = encoded tokens at kept positions, mask tokens at masked positions,
//               fused here with the decoder positional + date encoding add (ssl/mim.py:254-274).
#include "common.cuh"

// One CTA (1024 threads) per sample row.  rank(i) = #{j : (key_j, j) < (key_i, i)}; masked iff rank < nm.
// Outputs: mask u8 [B, L]; slot i32 [B, L] (index inside the masked / kept list); masked_idx i64 [B, nm] and
// unmasked_idx i64 [B, L - nm], both ascending; seg_counts[seg_id[p]] += 1 for every masked position.  The seg_id table
// has period Lt over the flat token index b * L + i (Lt = L, or a multiple of L when every date of a sample is its own
// sequence: fusion modes "shared" / "monotemp", layers/utils.py:26-33).
__global__ void __launch_bounds__(1024) mask_rank_kernel(const float* __restrict__ noise,
                                                         const unsigned char* __restrict__ structm, int L, int nm,
                                                         unsigned char* __restrict__ mask, int* __restrict__ slot,
                                                         long long* __restrict__ masked_idx,
                                                         long long* __restrict__ unmasked_idx,
                                                         const int* __restrict__ seg_id, int* __restrict__ seg_counts,
                                                         int Lt) {
  extern __shared__ float keys[];  // [L]
  __shared__ int warp_tot[2][32];
  __shared__ int run[2];
  const int b = blockIdx.x;
  const float* nr = noise + static_cast<size_t>(b) * L;
  const unsigned char* sr = structm + static_cast<size_t>(b) * L;
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    // noise * (1 - struct.float()) exactly as the reference forms it
    keys[i] = __fmul_rn(nr[i], 1.0f - static_cast<float>(sr[i] != 0));
  }
  if (threadIdx.x < 2) run[threadIdx.x] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  for (int base = 0; base < L; base += blockDim.x) {
    const int i = base + threadIdx.x;
    int is_m = 0, valid = i < L;
    if (valid) {
      const float ki = keys[i];
      int rank = 0;
      for (int j = 0; j < L; ++j) {
        const float kj = keys[j];
        rank += (kj < ki) || (kj == ki && j < i);
      }
      is_m = rank < nm;
    }
    const unsigned bm = __ballot_sync(0xffffffffu, valid && is_m);
    const unsigned bu = __ballot_sync(0xffffffffu, valid && !is_m);
    const unsigned lt = (1u << lane) - 1u;
    if (lane == 0) { warp_tot[0][w] = __popc(bm); warp_tot[1][w] = __popc(bu); }
    __syncthreads();
    int pm = run[0], pu = run[1];
    for (int k = 0; k < w; ++k) { pm += warp_tot[0][k]; pu += warp_tot[1][k]; }
    pm += __popc(bm & lt);
    pu += __popc(bu & lt);
    if (valid) {
      const size_t o = static_cast<size_t>(b) * L + i;
      mask[o] = static_cast<unsigned char>(is_m);
      if (is_m) {
        slot[o] = pm;
        masked_idx[static_cast<size_t>(b) * nm + pm] = i;
        if (seg_counts != nullptr) atomicAdd(&seg_counts[seg_id[o % Lt]], 1);
      } else {
        slot[o] = pu;
        unmasked_idx[static_cast<size_t>(b) * (L - nm) + pu] = i;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int tm = 0, tu = 0;
      for (int k = 0; k < nw; ++k) { tm += warp_tot[0][k]; tu += warp_tot[1][k]; }
      run[0] += tm;
      run[1] += tu;
    }
    __syncthreads();
  }
}

extern "C" int mb_mask_indices(const float* noise, const unsigned char* structm, int B, int L, int nm,
                               unsigned char* mask, int* slot, long long* masked_idx, long long* unmasked_idx,
                               const int* seg_id, int* seg_counts, int Lt, void* stream) {
  if (B == 0 || L == 0) return 0;
  if (Lt <= 0) Lt = L;
  if (Lt % L != 0) return -2;
  const size_t smem = static_cast<size_t>(L) * sizeof(float);
  if (smem > 200 * 1024) return -2;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(mask_rank_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
  mask_rank_kernel<<<B, 1024, smem, static_cast<cudaStream_t>(stream)>>>(noise, structm, L, nm, mask, slot,
                                                                        masked_idx, unmasked_idx, seg_id, seg_counts, Lt);
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ row gather
// out[b, j, :] = x[b, idx[b, j], :]   (x [B, L, d], idx int64 [B, n]).  One warp per output row.
template <typename T>
__global__ void __launch_bounds__(256) gather_rows_kernel(const T* __restrict__ x, const long long* __restrict__ idx,
                                                          T* __restrict__ out, int B, int L, int n, int d) {
  constexpr int V = VecIO<T>::N;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= B * n) return;
  const int b = row / n;
  const T* src = x + (static_cast<size_t>(b) * L + idx[row]) * d;
  T* dst = out + static_cast<size_t>(row) * d;
  for (int c = lane; c < d / V; c += 32) {
    float v[V];
    VecIO<T>::load(src + c * V, v);
    VecIO<T>::store(dst + c * V, v);
  }
}

extern "C" int mb_gather_rows(const void* x, const long long* idx, void* out, int dtype, int B, int L, int n, int d,
                              void* stream) {
  if (B * n == 0) return 0;
  if (d % 8 != 0) return -2;
  const int blocks = mb_ceil_div(static_cast<long long>(B) * n * 32, 256);
  MB_DISPATCH_DTYPE(dtype, T,
                    (gather_rows_kernel<T><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(x), idx, static_cast<T*>(out), B, L, n, d)));
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ gradient of the keep gather
// dx[b, p, :] = mask[b, p] ? 0 : dkept[b, slot[b, p], :]   (every row of dx is written).
template <typename T>
__global__ void __launch_bounds__(256) scatter_kept_kernel(const T* __restrict__ dkept,
                                                           const unsigned char* __restrict__ mask,
                                                           const int* __restrict__ slot, T* __restrict__ dx, int B,
                                                           int L, int n, int d) {
  constexpr int V = VecIO<T>::N;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= B * L) return;
  const int b = row / L;
  const bool m = mask[row] != 0;
  const T* src = dkept + (static_cast<size_t>(b) * n + (m ? 0 : slot[row])) * d;
  T* dst = dx + static_cast<size_t>(row) * d;
  for (int c = lane; c < d / V; c += 32) {
    float v[V];
    if (m) {
#pragma unroll
      for (int j = 0; j < V; ++j) v[j] = 0.f;
    } else {
      VecIO<T>::load(src + c * V, v);
    }
    VecIO<T>::store(dst + c * V, v);
  }
}

extern "C" int mb_scatter_kept(const void* dkept, const unsigned char* mask, const int* slot, void* dx, int dtype,
                               int B, int L, int n, int d, void* stream) {
  if (B * L == 0) return 0;
  if (d % 8 != 0) return -2;
  const int blocks = mb_ceil_div(static_cast<long long>(B) * L * 32, 256);
  MB_DISPATCH_DTYPE(dtype, T,
                    (scatter_kept_kernel<T><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(dkept), mask, slot, static_cast<T*>(dx), B, L, n, d)));
  MB_CHECK_LAUNCH();
  return 0;
}
