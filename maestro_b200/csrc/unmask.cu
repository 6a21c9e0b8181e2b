// unmask_seq (reference maestro/ssl/mae.py:266-287) fused with dec_add_encodings (ssl/mim.py:254-274).
#include "common.cuh"

// out[b, p, :] = (mask[b,p] ? tok[tok_id[src(b,p)], :] : y[b, slot[b,p], :]) + pos[p, :] + date(b, date_id[p])
//   y       [B, n_kept, d] (T)   encoder->decoder projected tokens
//   tok     [n_kinds, d] fp32    mask-token table, tok_id[p] selects the (modality, band-group) token
//   src_pos [B, L] i32 or null   test hook: position whose tok_id feeds masked position p (reference tie quirk)
//   pos     [L, d] fp32, feat [B, Dtot, 5] fp32, date_id[p] in [0, Dtot)
template <typename T>
__global__ void __launch_bounds__(256) unmask_fwd_kernel(const T* __restrict__ y, const unsigned char* __restrict__ mask,
                                                         const int* __restrict__ slot, const int* __restrict__ tok_id,
                                                         const int* __restrict__ src_pos, const float* __restrict__ tok,
                                                         const float* __restrict__ pos, const float* __restrict__ feat,
                                                         const int* __restrict__ date_id, T* __restrict__ out, int B,
                                                         int L, int n_kept, int d, int Dtot, int date_dim,
                                                         float fac_date) {
  constexpr int V = VecIO<T>::N;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= B * L) return;
  const int b = row / L, p = row % L;
  const bool m = mask[row] != 0;
  const float* f5 = feat + (static_cast<size_t>(b) * Dtot + date_id[p]) * 5;
  const int sp = (m && src_pos != nullptr) ? src_pos[row] : p;
  const float* tk = tok + static_cast<size_t>(tok_id[sp]) * d;
  const T* src = y + (static_cast<size_t>(b) * n_kept + (m ? 0 : slot[row])) * d;
  for (int c = lane; c < d / V; c += 32) {
    float v[V];
    if (m) {
#pragma unroll
      for (int j = 0; j < V; ++j) v[j] = to_f32(from_f32<T>(tk[c * V + j]));  // token cast to the activation dtype first
    } else {
      VecIO<T>::load(src + c * V, v);
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int e = c * V + j;
      float add = pos[static_cast<size_t>(p) * d + e];
      if (e >= d - date_dim) add += date_channel(f5, e - (d - date_dim), date_dim, fac_date);
      v[j] += add;
    }
    VecIO<T>::store(out + static_cast<size_t>(row) * d + c * V, v);
  }
}

extern "C" int mb_unmask_fwd(const void* y, const unsigned char* mask, const int* slot, const int* tok_id,
                             const int* src_pos, const float* tok, const float* pos, const float* feat,
                             const int* date_id, void* out, int dtype, int B, int L, int n_kept, int d, int Dtot,
                             int date_dim, float fac_date, void* stream) {
  if (B * L == 0) return 0;
  if (d % 8 != 0) return -2;
  const int blocks = mb_ceil_div(static_cast<long long>(B) * L * 32, 256);
  MB_DISPATCH_DTYPE(dtype, T,
                    (unmask_fwd_kernel<T><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(y), mask, slot, tok_id, src_pos, tok, pos, feat, date_id,
                        static_cast<T*>(out), B, L, n_kept, d, Dtot, date_dim, fac_date)));
  MB_CHECK_LAUNCH();
  return 0;
}

// Mask-token gradient: dtok[k, :] += sum over masked (b, p) with tok_id[src(b,p)] == k of dx[b, p, :].
// grid (ceil(d / 128), chunks over B*L rows); thread = one channel; fp32 atomics per kind at the end.
constexpr int MAX_TOK_KINDS = 8;
template <typename T>
__global__ void __launch_bounds__(128) masktoken_grad_kernel(const T* __restrict__ dx,
                                                             const unsigned char* __restrict__ mask,
                                                             const int* __restrict__ tok_id,
                                                             const int* __restrict__ src_pos, float* __restrict__ dtok,
                                                             int B, int L, int d, int n_kinds, int rows_per_block) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= d) return;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(B * L, r0 + rows_per_block);
  float acc[MAX_TOK_KINDS];
#pragma unroll
  for (int k = 0; k < MAX_TOK_KINDS; ++k) acc[k] = 0.f;
  for (int r = r0; r < r1; ++r) {
    if (mask[r] == 0) continue;
    const int p = r % L;
    const int sp = src_pos != nullptr ? src_pos[r] : p;
    const int k = tok_id[sp];
    const float v = to_f32(dx[static_cast<size_t>(r) * d + e]);
#pragma unroll
    for (int kk = 0; kk < MAX_TOK_KINDS; ++kk)
      if (kk == k) acc[kk] += v;
  }
#pragma unroll
  for (int k = 0; k < MAX_TOK_KINDS; ++k)
    if (k < n_kinds && acc[k] != 0.f) atomicAdd(dtok + static_cast<size_t>(k) * d + e, acc[k]);
}

extern "C" int mb_masktoken_grad(const void* dx, const unsigned char* mask, const int* tok_id, const int* src_pos,
                                 float* dtok, int dtype, int B, int L, int d, int n_kinds, void* stream) {
  if (B * L == 0) return 0;
  if (n_kinds > MAX_TOK_KINDS) return -2;
  const int bx = mb_ceil_div(d, 128);
  int by = (148 * 8) / bx;
  if (by < 1) by = 1;
  int rows_per_block = mb_ceil_div(static_cast<long long>(B) * L, by);
  if (rows_per_block < 8) rows_per_block = 8;
  by = mb_ceil_div(static_cast<long long>(B) * L, rows_per_block);
  dim3 grid(bx, by);
  MB_DISPATCH_DTYPE(dtype, T,
                    (masktoken_grad_kernel<T><<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(dx), mask, tok_id, src_pos, dtok, B, L, d, n_kinds, rows_per_block)));
  MB_CHECK_LAUNCH();
  return 0;
}
