// Normalisation kernels (bandwidth-bound; warp-shuffle reductions, 128-bit access):
//   * LayerNorm forward / backward with fused residual-gradient add and row remapping (the remap folds the
//     reference's torch.cat / split around encoder_inter, maestro/ssl/mim.py:406-423, into the LN pass);
//   * GroupNorm(1, E) of the patch embedding (maestro/layers/embed.py:55,61): statistics over all E*g*g
//     values of one (b, d) image, apply fused with the positional + date encoding add
//     (maestro/ssl/mim.py:232-252) and the group concat (layers/utils.py:12-47).
#include "common.cuh"
#include "tc_common.cuh"

struct RowMap {
  int n;    // logical rows per sample
  int bs;   // physical rows per sample
  int off;  // physical row offset inside a sample
  __device__ __forceinline__ size_t phys(int r) const { return static_cast<size_t>(r / n) * bs + off + (r % n); }
};


// ------------------------------------------------------------------ LayerNorm forward
// y = (x - mean) * rstd * gamma + beta per row (nn.LayerNorm, eps inside the square root; mean / rstd saved for the
// backward).  HBM-bound: 2 * d * sizeof(T) bytes per row.  Warp w of the grid walks rows w, w + W, w + 2W, ...; a
// lane owns NCH 16-byte column chunks, gamma / beta of those columns stay in registers for the whole kernel, and the
// (still packed) loads of row r + W are in flight while row r is reduced and written.
constexpr int LNF_THREADS = 256;

template <typename T, int NCH>
__global__ void __launch_bounds__(LNF_THREADS)
layernorm_fwd_kernel(const T* __restrict__ x, const float* __restrict__ gamma, const float* __restrict__ beta,
                     T* __restrict__ y, float* __restrict__ mean_out, float* __restrict__ rstd_out, int rows, int d,
                     float eps, RowMap in_map, RowMap out_map) {
  constexpr int V = VecIO<T>::N;
  constexpr int NE = NCH * V;
  typedef typename VecIO<T>::Raw Raw;
  const int lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const int gwarp = blockIdx.x * nwarps + (threadIdx.x >> 5);
  const int total_warps = gridDim.x * nwarps;
  const int nvec = d / V;
  const float inv_d = 1.0f / static_cast<float>(d);

  Raw nx[NCH];
  auto issue_loads = [&](int r) {
    const T* xr = x + in_map.phys(r) * d;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = lane + c * 32;
      nx[c] = idx < nvec ? VecIO<T>::load_raw(xr + idx * V) : VecIO<T>::zero_raw();
    }
  };
  int r = gwarp;
  if (r < rows) issue_loads(r);

  float g[NE], b[NE];
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int idx = lane + c * 32;
#pragma unroll
    for (int j = 0; j < V; j += 4) {
      const bool ok = idx < nvec;
      const float4 gv = ok ? __ldg(reinterpret_cast<const float4*>(gamma + idx * V + j)) : make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 bv = ok ? __ldg(reinterpret_cast<const float4*>(beta + idx * V + j)) : make_float4(0.f, 0.f, 0.f, 0.f);
      g[c * V + j] = gv.x; g[c * V + j + 1] = gv.y; g[c * V + j + 2] = gv.z; g[c * V + j + 3] = gv.w;
      b[c * V + j] = bv.x; b[c * V + j + 1] = bv.y; b[c * V + j + 2] = bv.z; b[c * V + j + 3] = bv.w;
    }
  }

  for (; r < rows; r += total_warps) {
    float v[NE];
#pragma unroll
    for (int c = 0; c < NCH; ++c) VecIO<T>::unpack(nx[c], v + c * V);
    if (r + total_warps < rows) issue_loads(r + total_warps);
    float s = 0.f;
#pragma unroll
    for (int e = 0; e < NE; ++e) s += v[e];                 // padding chunks hold zeros
    const float mean = warp_sum(s) * inv_d;
    float ss = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      if (lane + c * 32 < nvec) {
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float t = v[c * V + j] - mean;
          ss = fmaf(t, t, ss);
        }
      }
    }
    const float rstd = rsqrtf(warp_sum(ss) * inv_d + eps);
    if (lane == 0 && mean_out != nullptr) {
      mean_out[r] = mean;
      rstd_out[r] = rstd;
    }
    const float shift = -mean * rstd;
    T* yr = y + out_map.phys(r) * d;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = lane + c * 32;
      if (idx < nvec) {
        float o[V];
#pragma unroll
        for (int j = 0; j < V; ++j) o[j] = fmaf(fmaf(v[c * V + j], rstd, shift), g[c * V + j], b[c * V + j]);
        VecIO<T>::store(yr + idx * V, o);
      }
    }
  }
}

template <typename T, int NCH>
static void launch_ln_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                          int rows, int d, float eps, RowMap im, RowMap om, cudaStream_t st) {
  const int wpb = LNF_THREADS / 32;
  // at least 2 rows per warp (gamma / beta are loaded once per warp), at most 8 CTAs per SM
  long long blocks = mb_ceil_div(static_cast<long long>(rows), static_cast<long long>(wpb) * 2);
  const long long cap = 148LL * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  layernorm_fwd_kernel<T, NCH><<<static_cast<int>(blocks), LNF_THREADS, 0, st>>>(
      static_cast<const T*>(x), gamma, beta, static_cast<T*>(y), mean, rstd, rows, d, eps, im, om);
}

extern "C" int mb_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean,
                                float* rstd, int dtype, int rows, int d, float eps, int in_n, int in_bs, int in_off,
                                int out_n, int out_bs, int out_off, void* stream) {
  if (rows == 0) return 0;
  if (d > 1024 || d % 8 != 0) return -2;
  RowMap im{in_n, in_bs, in_off}, om{out_n, out_bs, out_off};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define LNF_CASE(n)                                                                                  \
  case n:                                                                                            \
    MB_DISPATCH_DTYPE(dtype, T, (launch_ln_fwd<T, n>(x, gamma, beta, y, mean, rstd, rows, d, eps, im, om, st))); \
    break;
  const int vec = dtype == MB_DTYPE_F32 ? 4 : 8;
  const int nch = mb_ceil_div(d / vec, 32);
  switch (nch) {
    LNF_CASE(1) LNF_CASE(2) LNF_CASE(3) LNF_CASE(4) LNF_CASE(5) LNF_CASE(6) LNF_CASE(7) LNF_CASE(8)
    default: return -2;
  }
#undef LNF_CASE
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ LayerNorm backward
// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)) [+ dres],  g = dy * gamma.
// dgamma += sum_rows dy * xhat, dbeta += sum_rows dy, and optionally dbias += sum_rows dx (the bias gradient of the
// linear layer whose output this LayerNorm consumed: its dY is exactly this dx, so the column sum rides along).
// x / dx / dres rows use in_map, dy rows use out_map (mirrors the forward).
// One CTA per SM, 12 warps, warp w of the grid walks rows w, w + W, w + 2W, ...; lanes own columns, so the column
// sums live in registers for the whole kernel and are combined once per CTA (shared-memory red, then one fp32
// global red per column per CTA: <= 148 per address).
// Every warp owns a ring of `depth` row slots in shared memory that one lane fills with 1D bulk copies
// (cp.async.bulk, mbarrier completion): up to depth x (x, dy, dres) rows per warp are in flight from HBM without
// holding registers, which is what the bandwidth-latency product of the chip needs (~45 KB per SM).
constexpr int LNB_THREADS = 384;

template <typename T, int NCH>
__global__ void __launch_bounds__(LNB_THREADS, 1)
layernorm_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x, const float* __restrict__ mean_in,
                     const float* __restrict__ rstd_in, const float* __restrict__ gamma, const T* __restrict__ dres,
                     T* __restrict__ dx, float* __restrict__ dgamma, float* __restrict__ dbeta,
                     float* __restrict__ dbias, int rows, int d, RowMap in_map, RowMap out_map, int depth) {
  constexpr int V = VecIO<T>::N;
  constexpr int NE = NCH * V;
  typedef typename VecIO<T>::Raw Raw;
  extern __shared__ __align__(128) uint8_t smem_raw[];
  // [3][d] partial dgamma / dbeta / dbias | gamma [d] | barriers [nwarps][depth] | ring [nwarps][depth][3][row]
  float* sm = reinterpret_cast<float*>(smem_raw);
  float* gam = sm + 3 * d;
  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;
  const int nwarps = blockDim.x >> 5;
  uint64_t* bars = reinterpret_cast<uint64_t*>(gam + d) + wid * depth;
  const uint32_t row_bytes = static_cast<uint32_t>(d) * sizeof(T);
  uint8_t* ring = reinterpret_cast<uint8_t*>(reinterpret_cast<uint64_t*>(gam + d) + nwarps * depth) +
                  static_cast<size_t>(wid) * depth * 3 * row_bytes;
  const int gwarp = blockIdx.x * nwarps + wid;
  const int total_warps = gridDim.x * nwarps;
  const int nvec = d / V;
  const bool has_res = dres != nullptr;
  for (int i = threadIdx.x; i < 3 * d; i += blockDim.x) sm[i] = 0.f;
  for (int i = threadIdx.x; i < d; i += blockDim.x) gam[i] = gamma[i];
  if (lane == 0) {
    for (int s = 0; s < depth; ++s) tc::mbar_init(&bars[s], 1);
    tc::fence_barrier_init();
  }
  __syncthreads();

  auto issue_row = [&](int slot, int r) {   // one lane
    const size_t pin = in_map.phys(r), pout = out_map.phys(r);
    uint8_t* dst = ring + static_cast<size_t>(slot) * 3 * row_bytes;
    tc::mbar_expect_tx(&bars[slot], (has_res ? 3u : 2u) * row_bytes);
    tc::bulk_load_1d(dst, x + pin * d, row_bytes, &bars[slot]);
    tc::bulk_load_1d(dst + row_bytes, dy + pout * d, row_bytes, &bars[slot]);
    if (has_res) tc::bulk_load_1d(dst + 2 * row_bytes, dres + pin * d, row_bytes, &bars[slot]);
  };
  if (lane == 0) {
    for (int s = 0; s < depth; ++s) {
      const long long rs = gwarp + static_cast<long long>(s) * total_warps;
      if (rs < rows) issue_row(s, static_cast<int>(rs));
    }
  }

  float ag[NE], ab[NE], ac[NE];
#pragma unroll
  for (int e = 0; e < NE; ++e) { ag[e] = 0.f; ab[e] = 0.f; ac[e] = 0.f; }
  const float inv_d = 1.0f / static_cast<float>(d);

  int slot = 0;
  uint32_t phase = 0;
  for (int r = gwarp; r < rows; r += total_warps) {
    const size_t pin = in_map.phys(r);
    const float mean = mean_in[r], rstd = rstd_in[r];
    float xh[NE], dv[NE];
    Raw cres[NCH];
    tc::mbar_wait(&bars[slot], phase);
    {
      const T* sx = reinterpret_cast<const T*>(ring + static_cast<size_t>(slot) * 3 * row_bytes);
      const T* sdy = reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(sx) + row_bytes);
      const T* sres = reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(sx) + 2 * row_bytes);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const int idx = lane + c * 32;
        const bool ok = idx < nvec;
        const Raw rx = ok ? VecIO<T>::load_raw(sx + idx * V) : VecIO<T>::zero_raw();
        const Raw rdy = ok ? VecIO<T>::load_raw(sdy + idx * V) : VecIO<T>::zero_raw();
        cres[c] = (ok && has_res) ? VecIO<T>::load_raw(sres + idx * V) : VecIO<T>::zero_raw();
        VecIO<T>::unpack(rx, xh + c * V);
        VecIO<T>::unpack(rdy, dv + c * V);
      }
    }
    __syncwarp();   // every lane has read the slot: refill it with the row `depth` rounds ahead
    {
      const long long rn = r + static_cast<long long>(depth) * total_warps;
      if (lane == 0 && rn < rows) issue_row(slot, static_cast<int>(rn));
    }
    if (++slot == depth) { slot = 0; phase ^= 1; }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = lane + c * 32;
      if (idx < nvec) {
        float gv[V];
        VecIO<float>::load(gam + idx * V, gv);
        if (V == 8) VecIO<float>::load(gam + idx * V + 4, gv + 4);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const int e = c * V + j;
          const float h = (xh[e] - mean) * rstd;
          const float gg = dv[e] * gv[j];
          xh[e] = h;
          s1 += gg;
          s2 = fmaf(gg, h, s2);
          ag[e] = fmaf(dv[e], h, ag[e]);
          ab[e] += dv[e];
          dv[e] = gg;
        }
      }
    }
    s1 = warp_sum(s1) * inv_d;
    s2 = warp_sum(s2) * inv_d;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = lane + c * 32;
      if (idx < nvec) {
        float o[V];
        VecIO<T>::unpack(cres[c], o);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const int e = c * V + j;
          o[j] += rstd * (dv[e] - s1 - xh[e] * s2);
          ac[e] += o[j];
        }
        VecIO<T>::store(dx + pin * d + idx * V, o);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int idx = lane + c * 32;
    if (idx < nvec) {
#pragma unroll
      for (int j = 0; j < V; ++j) {
        // column (idx * V + j) lives at [j * nvec + idx]: consecutive lanes -> consecutive banks (conflict-free red)
        atomicAdd(&sm[j * nvec + idx], ag[c * V + j]);
        atomicAdd(&sm[d + j * nvec + idx], ab[c * V + j]);
        if (dbias != nullptr) atomicAdd(&sm[2 * d + j * nvec + idx], ac[c * V + j]);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    const int si = (i % V) * nvec + i / V;
    atomicAdd(dgamma + i, sm[si]);
    atomicAdd(dbeta + i, sm[d + si]);
    if (dbias != nullptr) atomicAdd(dbias + i, sm[2 * d + si]);
  }
}

template <typename T>
static int launch_ln_bwd(const void* dy, const void* x, const float* mean, const float* rstd, const float* gamma,
                         const void* dres, void* dx, float* dgamma, float* dbeta, float* dbias, int rows, int d,
                         RowMap im, RowMap om, cudaStream_t st) {
  constexpr int V = VecIO<T>::N;
  const int nwarps = LNB_THREADS / 32;
  // at least 4 rows per warp before the per-CTA reduction tail pays off; small groups then leave SMs to the
  // other groups' kernels that run concurrently on forked streams
  int blocks = mb_ceil_div(rows, nwarps * 4);
  if (blocks > 148) blocks = 148;
  // shared memory: column partials + gamma + barriers + per-warp row ring (depth slots of x / dy / dres rows)
  const size_t row_bytes = static_cast<size_t>(d) * sizeof(T);
  const size_t fixed = 4 * static_cast<size_t>(d) * sizeof(float);
  int depth = static_cast<int>((200 * 1024 - fixed - 1024) / (nwarps * 3 * row_bytes));
  if (depth > 4) depth = 4;
  if (depth < 1) depth = 1;
  const size_t smem = fixed + static_cast<size_t>(nwarps) * depth * (8 + 3 * row_bytes);
  const int nch = mb_ceil_div(d / V, 32);
#define MB_LNB(N_)                                                                                              \
  case N_:                                                                                                      \
    {                                                                                                           \
      static bool attr_done = false;                                                                            \
      if (!attr_done) {                                                                                         \
        cudaFuncSetAttribute(layernorm_bwd_kernel<T, N_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024); \
        attr_done = true;                                                                                       \
      }                                                                                                         \
    }                                                                                                           \
    layernorm_bwd_kernel<T, N_><<<blocks, LNB_THREADS, smem, st>>>(                                              \
        static_cast<const T*>(dy), static_cast<const T*>(x), mean, rstd, gamma, static_cast<const T*>(dres),    \
        static_cast<T*>(dx), dgamma, dbeta, dbias, rows, d, im, om, depth);                                     \
    break;
  switch (nch) {
    MB_LNB(1) MB_LNB(2) MB_LNB(3) MB_LNB(4) MB_LNB(5) MB_LNB(6) MB_LNB(7) MB_LNB(8)
    default: return -2;
  }
#undef MB_LNB
  return 0;
}

// dbias may be null.  dgamma / dbeta / dbias accumulate (fp32).
extern "C" int mb_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd,
                                const float* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                                float* dbias, int dtype, int rows, int d, int in_n, int in_bs, int in_off, int out_n,
                                int out_bs, int out_off, void* stream) {
  if (rows == 0) return 0;
  if (d > 1024 || d % 8 != 0) return -2;
  RowMap im{in_n, in_bs, in_off}, om{out_n, out_bs, out_off};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc;
  if (dtype == MB_DTYPE_F32) rc = launch_ln_bwd<float>(dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, dbias, rows, d, im, om, st);
  else if (dtype == MB_DTYPE_BF16) rc = launch_ln_bwd<bf16>(dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, dbias, rows, d, im, om, st);
  else return -1;
  if (rc != 0) return rc;
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ GroupNorm(1, E) of the patch embedding
// Phase 1: per-image sum / sum of squares, split over several CTAs per image (fp64 atomics).
template <typename T>
__global__ void __launch_bounds__(256) gn_stats_kernel(const T* __restrict__ y, double* __restrict__ sums,
                                                       size_t elems_per_img, int splits) {
  constexpr int V = VecIO<T>::N;
  const int img = blockIdx.x / splits, part = blockIdx.x % splits;
  const size_t nvec = elems_per_img / V;
  const size_t per = (nvec + splits - 1) / splits;
  const size_t v0 = part * per, v1 = min(nvec, v0 + per);
  const T* base = y + static_cast<size_t>(img) * elems_per_img;
  float s = 0.f, ss = 0.f;
  for (size_t i = v0 + threadIdx.x; i < v1; i += blockDim.x) {
    float v[V];
    VecIO<T>::load(base + i * V, v);
#pragma unroll
    for (int j = 0; j < V; ++j) { s += v[j]; ss += v[j] * v[j]; }
  }
  double ds = warp_sum(static_cast<double>(s)), dss = warp_sum(static_cast<double>(ss));
  __shared__ double red[2][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) { red[0][w] = ds; red[1][w] = dss; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int i = 0; i < (blockDim.x >> 5); ++i) { a += red[0][i]; b += red[1][i]; }
    atomicAdd(&sums[2 * img], a);
    atomicAdd(&sums[2 * img + 1], b);
  }
}

// sums must be zero on entry ([n_img, 2] doubles).
extern "C" int mb_gn_stats(const void* y, int dtype, double* sums, int n_img, long long elems_per_img, void* stream) {
  if (n_img == 0) return 0;
  if (elems_per_img % 8 != 0) return -2;
  int splits = (148 * 4) / n_img;
  if (splits < 1) splits = 1;
  const long long max_splits = elems_per_img / (8 * 256) + 1;
  if (splits > max_splits) splits = static_cast<int>(max_splits);
  MB_DISPATCH_DTYPE(dtype, T,
                    (gn_stats_kernel<T><<<n_img * splits, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(y), sums, static_cast<size_t>(elems_per_img), splits)));
  MB_CHECK_LAUNCH();
  return 0;
}

// Phase 2: normalise + affine + positional + date encoding, written into the group token buffer.
//   y      [n_img * L, E]   conv output (image i = b * D + d)
//   out    [B, out_bs, E]   row = b * out_bs + out_off + d * L + l
//   pos    [L, E] fp32 (already scaled by fac_abs_enc), feat [B, feat_bs, 5] date features (rows feat_off + d).
template <typename T>
__global__ void __launch_bounds__(256) gn_apply_kernel(const T* __restrict__ y, const double* __restrict__ sums,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       const float* __restrict__ pos, const float* __restrict__ feat,
                                                       T* __restrict__ out, float* __restrict__ stats_out, int n_img,
                                                       int D, int L, int E, float eps, int out_bs, int out_off,
                                                       int date_dim, float fac_date, int feat_bs, int feat_off) {
  constexpr int V = VecIO<T>::N;
  extern __shared__ float st_s[];            // [n_img][2]: mean, rstd (the fp64 finalisation runs once per block)
  const double inv_n = 1.0 / (static_cast<double>(L) * E);
  for (int i = threadIdx.x; i < n_img; i += blockDim.x) {
    const double m = sums[2 * i] * inv_n;
    const double var = fmax(sums[2 * i + 1] * inv_n - m * m, 0.0);
    const float mean = static_cast<float>(m);
    const float rstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
    st_s[2 * i] = mean;
    st_s[2 * i + 1] = rstd;
    if (stats_out != nullptr && blockIdx.x == 0) {
      stats_out[2 * i] = mean;
      stats_out[2 * i + 1] = rstd;
    }
  }
  __syncthreads();
  const unsigned nvec_row = E / V;
  const unsigned total = static_cast<unsigned>(n_img) * L * nvec_row;      // host checks < 2^32
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned row = i / nvec_row, cv = i - row * nvec_row;
    const unsigned img = row / L, l = row - img * L;
    const unsigned b = img / D, d = img - b * D;
    const float mean = st_s[2 * img], rstd = st_s[2 * img + 1];
    const float shift = -mean * rstd;
    float v[V], gm[V], bt[V], ps[V];
    VecIO<T>::load(y + static_cast<size_t>(row) * E + cv * V, v);
#pragma unroll
    for (int j = 0; j < V; j += 4) {
      VecIO<float>::load(gamma + cv * V + j, gm + j);
      VecIO<float>::load(beta + cv * V + j, bt + j);
      VecIO<float>::load(pos + static_cast<size_t>(l) * E + cv * V + j, ps + j);
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int e = cv * V + j;
      float o = fmaf(fmaf(v[j], rstd, shift), gm[j], bt[j]) + ps[j];
      if (e >= E - date_dim) o += date_channel(feat + (static_cast<size_t>(b) * feat_bs + feat_off + d) * 5, e - (E - date_dim), date_dim, fac_date);
      v[j] = o;
    }
    const size_t orow = static_cast<size_t>(b) * out_bs + out_off + static_cast<size_t>(d) * L + l;
    VecIO<T>::store(out + orow * E + cv * V, v);
  }
}

extern "C" int mb_gn_apply(const void* y, const double* sums, const float* gamma, const float* beta,
                           const float* pos, const float* feat, void* out, float* stats_out, int dtype, int n_img,
                           int D, int L, int E, float eps, int out_bs, int out_off, int date_dim, float fac_date,
                           int feat_bs, int feat_off, void* stream) {
  if (n_img == 0) return 0;
  if (E % 8 != 0) return -2;
  const long long total = static_cast<long long>(n_img) * L * (E / 4);
  if (total >= 0xffffffffll || n_img > 4096) return -2;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const size_t st_bytes = static_cast<size_t>(n_img) * 2 * sizeof(float);
  MB_DISPATCH_DTYPE(dtype, T,
                    (gn_apply_kernel<T><<<static_cast<int>(blocks), 256, st_bytes, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(y), sums, gamma, beta, pos, feat, static_cast<T*>(out), stats_out,
                        n_img, D, L, E, eps, out_bs, out_off, date_dim, fac_date, feat_bs, feat_off)));
  MB_CHECK_LAUNCH();
  return 0;
}

// Backward phase 1: per image a = sum(dy * gamma), c = sum(dy * gamma * xhat) (fp64 atomics into sums2 [n_img,2]);
// per channel dgamma += sum dy * xhat, dbeta += sum dy (fp32 atomics).  dy is read from the group token buffer.
// grid (n_img, row blocks); the block's threads are laid out as [row lane][16-byte column chunk], every row lane
// keeps 4 rows of loads in flight, and the row lanes are combined in shared memory so that a block issues ONE
// atomic per column (same-address atomics from many blocks are what bounds this kernel).
constexpr int GNR_THREADS = 512;
template <typename T>
__global__ void __launch_bounds__(GNR_THREADS) gn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                    const float* __restrict__ stats,
                                                                    const float* __restrict__ gamma,
                                                                    double* __restrict__ sums2,
                                                                    float* __restrict__ dgamma,
                                                                    float* __restrict__ dbeta, int D, int L, int E,
                                                                    int out_bs, int out_off, int rows_per_block) {
  constexpr int V = VecIO<T>::N;
  constexpr int U = 4;
  typedef typename VecIO<T>::Raw Raw;
  extern __shared__ float cs[];              // [2][E]: dgamma / dbeta partials, column (cv * V + j) at [j * nvec + cv]
  __shared__ double red[2][GNR_THREADS / 32];
  const int img = blockIdx.x;
  const int b = img / D, d = img % D;
  const int l0 = blockIdx.y * rows_per_block, l1 = min(L, l0 + rows_per_block);
  const float mean = stats[2 * img], rstd = stats[2 * img + 1];
  const int nvec = E / V;
  const int cv = threadIdx.x % nvec, rl = threadIdx.x / nvec, nrl = blockDim.x / nvec;
  for (int i = threadIdx.x; i < 2 * E; i += blockDim.x) cs[i] = 0.f;
  __syncthreads();
  float a = 0.f, c = 0.f;
  if (rl < nrl) {
    float ag[V], ab[V], gm[V];
#pragma unroll
    for (int j = 0; j < V; j += 4) VecIO<float>::load(gamma + cv * V + j, gm + j);
#pragma unroll
    for (int j = 0; j < V; ++j) { ag[j] = 0.f; ab[j] = 0.f; }
    for (int l = l0 + rl; l < l1; l += nrl * U) {
      Raw ry[U], rd[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int lu = min(l + u * nrl, l1 - 1);    // clamped rows are loaded again and skipped below
        const size_t yrow = static_cast<size_t>(img) * L + lu;
        const size_t drow = static_cast<size_t>(b) * out_bs + out_off + static_cast<size_t>(d) * L + lu;
        ry[u] = VecIO<T>::load_raw(y + yrow * E + cv * V);
        rd[u] = VecIO<T>::load_raw(dy + drow * E + cv * V);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (l + u * nrl >= l1) break;
        float yv[V], dv[V];
        VecIO<T>::unpack(ry[u], yv);
        VecIO<T>::unpack(rd[u], dv);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float h = (yv[j] - mean) * rstd;
          const float g = dv[j] * gm[j];
          a += g;
          c += g * h;
          ag[j] += dv[j] * h;
          ab[j] += dv[j];
        }
      }
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      atomicAdd(&cs[j * nvec + cv], ag[j]);
      atomicAdd(&cs[E + j * nvec + cv], ab[j]);
    }
  }
  double da = warp_sum(static_cast<double>(a)), dc = warp_sum(static_cast<double>(c));
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) { red[0][w] = da; red[1][w] = dc; }
  __syncthreads();
  for (int i = threadIdx.x; i < E; i += blockDim.x) {
    const int si = (i % V) * nvec + i / V;
    atomicAdd(dgamma + i, cs[si]);
    atomicAdd(dbeta + i, cs[E + si]);
  }
  if (threadIdx.x == 0) {
    double s0 = 0, s1 = 0;
    for (int i = 0; i < ((blockDim.x + 31) >> 5); ++i) { s0 += red[0][i]; s1 += red[1][i]; }
    atomicAdd(&sums2[2 * img], s0);
    atomicAdd(&sums2[2 * img + 1], s1);
  }
}

// Backward phase 2: dconv = rstd * (dy * gamma - a / n - xhat * c / n), rows back in conv order.
template <typename T>
__global__ void __launch_bounds__(256) gn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                           const float* __restrict__ stats,
                                                           const float* __restrict__ gamma,
                                                           const double* __restrict__ sums2, T* __restrict__ dconv,
                                                           int n_img, int D, int L, int E, int out_bs, int out_off) {
  constexpr int V = VecIO<T>::N;
  extern __shared__ float st_s[];            // [n_img][4]: mean, rstd, a / n, c / n (fp64 -> fp32 once per block)
  const double inv_n = 1.0 / (static_cast<double>(L) * E);
  for (int i = threadIdx.x; i < n_img; i += blockDim.x) {
    st_s[4 * i] = stats[2 * i];
    st_s[4 * i + 1] = stats[2 * i + 1];
    st_s[4 * i + 2] = static_cast<float>(sums2[2 * i] * inv_n);
    st_s[4 * i + 3] = static_cast<float>(sums2[2 * i + 1] * inv_n);
  }
  __syncthreads();
  const unsigned nvec_row = E / V;
  const unsigned total = static_cast<unsigned>(n_img) * L * nvec_row;      // host checks < 2^32
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned row = i / nvec_row, cv = i - row * nvec_row;
    const unsigned img = row / L, l = row - img * L;
    const unsigned b = img / D, d = img - b * D;
    const float4 q = *reinterpret_cast<const float4*>(st_s + 4 * img);
    const float mean = q.x, rstd = q.y, ma = q.z, mc = q.w;
    const size_t drow = static_cast<size_t>(b) * out_bs + out_off + static_cast<size_t>(d) * L + l;
    float yv[V], dv[V], gm[V];
    VecIO<T>::load(y + static_cast<size_t>(row) * E + cv * V, yv);
    VecIO<T>::load(dy + drow * E + cv * V, dv);
#pragma unroll
    for (int j = 0; j < V; j += 4) VecIO<float>::load(gamma + cv * V + j, gm + j);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const float h = (yv[j] - mean) * rstd;
      dv[j] = rstd * (dv[j] * gm[j] - ma - h * mc);
    }
    VecIO<T>::store(dconv + static_cast<size_t>(row) * E + cv * V, dv);
  }
}

// sums2 must be zero on entry.
extern "C" int mb_gn_bwd(const void* dy, const void* y, const float* stats, const float* gamma, double* sums2,
                         void* dconv, float* dgamma, float* dbeta, int dtype, int n_img, int D, int L, int E,
                         int out_bs, int out_off, void* stream) {
  if (n_img == 0) return 0;
  if (E % 8 != 0) return -2;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (E / 4 > GNR_THREADS) return -2;        // one column chunk per thread
  int row_blocks = (148 * 2) / n_img;        // ~2 blocks per SM; every block ends in 2 E same-address atomics
  if (row_blocks < 1) row_blocks = 1;
  if (row_blocks > L) row_blocks = L;
  const int rows_per_block = mb_ceil_div(L, row_blocks);
  row_blocks = mb_ceil_div(L, rows_per_block);
  dim3 grid(n_img, row_blocks);
  const long long total = static_cast<long long>(n_img) * L * (E / 4);
  if (total >= 0xffffffffll || n_img > 2048) return -2;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const size_t st_bytes = static_cast<size_t>(n_img) * 4 * sizeof(float);
  MB_DISPATCH_DTYPE(
      dtype, T,
      (gn_bwd_reduce_kernel<T><<<grid, GNR_THREADS, 2 * E * sizeof(float), st>>>(static_cast<const T*>(dy), static_cast<const T*>(y), stats, gamma,
                                                     sums2, dgamma, dbeta, D, L, E, out_bs, out_off, rows_per_block),
       gn_bwd_apply_kernel<T><<<static_cast<int>(blocks), 256, st_bytes, st>>>(static_cast<const T*>(dy),
                                                                        static_cast<const T*>(y), stats, gamma, sums2,
                                                                        static_cast<T*>(dconv), n_img, D, L, E, out_bs,
                                                                        out_off)));
  MB_CHECK_LAUNCH();
  return 0;
}
