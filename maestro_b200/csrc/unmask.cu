// unmask_seq (reference maestro/ssl/mae.py:266-287) fused with dec_add_encodings (ssl/mim.py:254-274).
#include "common.cuh"

// out[b, p, :] = (mask[b,p] ? tok[tok_id[src(b,p)], :] : y[b, slot[b,p], :]) + pos[p, :] + date(b, date_id[p])
//   y       [B, n_kept, d] (T)   encoder->decoder projected tokens
//   tok     [n_kinds, d] fp32    mask-token table, tok_id[p] selects the (modality, band-group) token
//   src_pos [B, L] i32 or null   test hook: position whose tok_id feeds masked position p (reference tie quirk)
//   pos     [Lt, d] fp32, feat [B / (Lt / L), Dtot, 5] fp32, date_id[t] in [0, Dtot)
// Tables (tok_id, pos, date_id) have period Lt over the flat token index b * L + p and one feat block per Lt tokens:
// Lt = L for fusion modes "group" / "mod"; for "shared" / "monotemp" every (band group, date) of a sample is its own
// sequence of L = grid^2 tokens (layers/utils.py:26-33) and Lt = G * D * L tokens make up one sample.
template <typename T>
__global__ void __launch_bounds__(256) unmask_fwd_kernel(const T* __restrict__ y, const unsigned char* __restrict__ mask,
                                                         const int* __restrict__ slot, const int* __restrict__ tok_id,
                                                         const int* __restrict__ src_pos, const float* __restrict__ tok,
                                                         const float* __restrict__ pos, const float* __restrict__ feat,
                                                         const int* __restrict__ date_id, T* __restrict__ out, int B,
                                                         int L, int n_kept, int d, int Dtot, int date_dim,
                                                         float fac_date, int Lt) {
  constexpr int V = VecIO<T>::N;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= B * L) return;
  const int b = row / L, p = row % L;
  const int bs = row / Lt, pt = row % Lt;             // sample / index into the per-sample tables
  const bool m = mask[row] != 0;
  const float* f5 = feat + (static_cast<size_t>(bs) * Dtot + date_id[pt]) * 5;
  const int sp = (m && src_pos != nullptr) ? src_pos[row] : p;
  const float* tk = tok + static_cast<size_t>(tok_id[pt - p + sp]) * d;
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
      float add = pos[static_cast<size_t>(pt) * d + e];
      if (e >= d - date_dim) add += date_channel(f5, e - (d - date_dim), date_dim, fac_date);
      v[j] += add;
    }
    VecIO<T>::store(out + static_cast<size_t>(row) * d + c * V, v);
  }
}

extern "C" int mb_unmask_fwd(const void* y, const unsigned char* mask, const int* slot, const int* tok_id,
                             const int* src_pos, const float* tok, const float* pos, const float* feat,
                             const int* date_id, void* out, int dtype, int B, int L, int n_kept, int d, int Dtot,
                             int date_dim, float fac_date, int Lt, void* stream) {
  if (B * L == 0) return 0;
  if (Lt <= 0) Lt = L;
  if (d % 8 != 0 || Lt % L != 0) return -2;
  const int blocks = mb_ceil_div(static_cast<long long>(B) * L * 32, 256);
  MB_DISPATCH_DTYPE(dtype, T,
                    (unmask_fwd_kernel<T><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(y), mask, slot, tok_id, src_pos, tok, pos, feat, date_id,
                        static_cast<T*>(out), B, L, n_kept, d, Dtot, date_dim, fac_date, Lt)));
  MB_CHECK_LAUNCH();
  return 0;
}

// Mask-token gradient: dtok[k, :] += sum over masked (b, p) with tok_id[src(b,p)] == k of dx[b, p, :].
// A thread owns one 16-byte column chunk; the block's threads are laid out as [row lane][column chunk], each row lane
// walks its rows four at a time (mask bytes, then token ids, then dx chunks: three dependent load rounds per four
// rows instead of twelve); fp32 atomics per kind at the end.
constexpr int MAX_TOK_KINDS = 8;
template <typename T, int NK>
__global__ void __launch_bounds__(512) masktoken_grad_kernel(const T* __restrict__ dx,
                                                             const unsigned char* __restrict__ mask,
                                                             const int* __restrict__ tok_id,
                                                             const int* __restrict__ src_pos, float* __restrict__ dtok,
                                                             int B, int L, int d, int n_kinds, int rows_per_block,
                                                             int Lt) {
  constexpr int V = VecIO<T>::N;
  constexpr int U = 4;
  typedef typename VecIO<T>::Raw Raw;
  const int nvec = d / V;
  extern __shared__ float red[];            // [NK][d] partial sums of this block (row lanes combined before the atomics)
  const int cv = threadIdx.x % nvec, rl = threadIdx.x / nvec, nrl = blockDim.x / nvec;
  for (int i = threadIdx.x; i < NK * d; i += blockDim.x) red[i] = 0.f;
  __syncthreads();
  const bool active = rl < nrl;
  const int r0 = blockIdx.x * rows_per_block;
  const int r1 = min(B * L, r0 + rows_per_block);
  float acc[NK][V];
#pragma unroll
  for (int k = 0; k < NK; ++k)
#pragma unroll
    for (int j = 0; j < V; ++j) acc[k][j] = 0.f;
  for (int r = r0 + rl; active && r < r1; r += nrl * U) {
    int kind[U];
    Raw raw[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int ru = r + u * nrl;
      kind[u] = (ru < r1 && mask[ru] != 0) ? ru : -1;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (kind[u] >= 0) {
        const int ru = kind[u];
        raw[u] = VecIO<T>::load_raw(dx + static_cast<size_t>(ru) * d + cv * V);
        const int p = ru % L;
        const int sp = src_pos != nullptr ? src_pos[ru] : p;
        kind[u] = tok_id[ru % Lt - p + sp];
      } else {
        raw[u] = VecIO<T>::zero_raw();
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      float v[V];
      VecIO<T>::unpack(raw[u], v);
#pragma unroll
      for (int kk = 0; kk < NK; ++kk) {
        const float sel = (kk == kind[u]) ? 1.f : 0.f;       // branch-free: the accumulators stay in registers
#pragma unroll
        for (int j = 0; j < V; ++j) acc[kk][j] = fmaf(sel, v[j], acc[kk][j]);
      }
    }
  }
  if (active) {
#pragma unroll
    for (int k = 0; k < NK; ++k)
#pragma unroll
      for (int j = 0; j < V; ++j)   // column (cv * V + j) of kind k lives at [k][j * nvec + cv]: conflict-free across a warp
        if (acc[k][j] != 0.f) atomicAdd(&red[k * d + j * nvec + cv], acc[k][j]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < NK * d; i += blockDim.x) {
    const int k = i / d, c = i - k * d;
    const float v = red[k * d + (c % V) * nvec + c / V];
    if (k < n_kinds && v != 0.f) atomicAdd(dtok + static_cast<size_t>(k) * d + c, v);
  }
}

extern "C" int mb_masktoken_grad(const void* dx, const unsigned char* mask, const int* tok_id, const int* src_pos,
                                 float* dtok, int dtype, int B, int L, int d, int n_kinds, int Lt, void* stream) {
  if (B * L == 0) return 0;
  if (Lt <= 0) Lt = L;
  if (Lt % L != 0) return -2;
  if (n_kinds > MAX_TOK_KINDS || d % 8 != 0 || d / (dtype == MB_DTYPE_F32 ? 4 : 8) > 256) return -2;   // one column chunk per thread
  int blocks = 148;                         // one same-address atomic per column per block at the end
  int rows_per_block = mb_ceil_div(static_cast<long long>(B) * L, blocks);
  if (rows_per_block < 64) rows_per_block = 64;
  blocks = mb_ceil_div(static_cast<long long>(B) * L, rows_per_block);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define MTG(NK_)                                                                                              \
  MB_DISPATCH_DTYPE(dtype, T,                                                                                 \
                    (masktoken_grad_kernel<T, NK_><<<blocks, 512, NK_ * d * sizeof(float), st>>>(static_cast<const T*>(dx), mask, tok_id, \
                                                                           src_pos, dtok, B, L, d, n_kinds,    \
                                                                           rows_per_block, Lt)))
  if (n_kinds <= 1) { MTG(1); }
  else if (n_kinds <= 2) { MTG(2); }
  else if (n_kinds <= 4) { MTG(4); }
  else { MTG(8); }
#undef MTG
  MB_CHECK_LAUNCH();
  return 0;
}
