// Device-side input pipeline of the pretraining step:
//   * raster_transform: raw raster (u8 / i16 / u16 / f32, as stored on disk) -> fp32 model input with the
//     reference's per-sample dihedral augmentation, log scale and normalisation
//     (maestro/dataset/dataset.py:215-222 astype(float32) -> log(max(x, 1e-10)) -> /= norm_fac;
//      :225-257 transform_rasters: flip axis 2, flip axis 3, swapaxes(2, 3), in that order, same draw for every raster
//      of a sample).  The host then ships the raw dtype over PCIe (aerial u8: 4x fewer bytes than fp32).
//   * struct_mask: MAE.mask_struct (maestro/ssl/mae.py:178-226) with a counter-based Philox4x32-10 generator instead
//     of the host torch.rand loop: one CTA per sample redraws until its group is not entirely masked.
#include "common.cuh"

#define MB_RAW_U8 0
#define MB_RAW_I16 1
#define MB_RAW_U16 2
#define MB_RAW_F32 3

#define MB_AUG_FLIP_H 1
#define MB_AUG_FLIP_W 2
#define MB_AUG_TRANSPOSE 4

// ------------------------------------------------------------------ raster transform
// raw [n_img, C, S_h, S_w] -> out [n_img, C, H, W] fp32; image i belongs to sample i / imgs_per_sample whose
// augmentation bits are aug[sample] (aug == nullptr: identity).  One CTA moves a 32 x 32 output tile of one
// (image, channel) plane through shared memory so that both the reads (along the source row) and the writes
// (along the output row) are coalesced whether or not the sample is transposed.
template <typename Raw>
__global__ void __launch_bounds__(256)
raster_transform_kernel(const Raw* __restrict__ raw, float* __restrict__ out, const unsigned char* __restrict__ aug,
                        int imgs_per_sample, int C, int H, int W, int tiles_x, int tiles_per_plane, float norm_fac,
                        int use_norm, int log_scale) {
  __shared__ float tile[32][33];
  const unsigned plane = blockIdx.x / tiles_per_plane;          // image * C + channel
  const unsigned t = blockIdx.x - plane * tiles_per_plane;
  const int ti = t / tiles_x, tj = t - ti * tiles_x;            // output tile: rows 32 ti.., cols 32 tj..
  const int img = plane / C;
  const int bits = aug ? aug[img / imgs_per_sample] : 0;
  const bool fh = bits & MB_AUG_FLIP_H, fw = bits & MB_AUG_FLIP_W, tr = bits & MB_AUG_TRANSPOSE;
  // source plane is [Hs, Ws]: equal to [H, W], or [W, H] for a transposed sample
  const int Hs = tr ? W : H, Ws = tr ? H : W;
  const Raw* src = raw + static_cast<size_t>(plane) * Hs * Ws;
  float* dst = out + static_cast<size_t>(plane) * H * W;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  // source tile origin before the flips: (row, col) = tr ? (32 tj, 32 ti) : (32 ti, 32 tj)
  const int r0 = tr ? 32 * tj : 32 * ti, c0 = tr ? 32 * ti : 32 * tj;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = r0 + ty + 8 * k, c = c0 + tx;                 // position in flipped(source) coordinates
    if (r < Hs && c < Ws) {
      const int sr = fh ? Hs - 1 - r : r, sc = fw ? Ws - 1 - c : c;
      float v = static_cast<float>(src[static_cast<size_t>(sr) * Ws + sc]);
      if (log_scale) v = logf(v < 1e-10f ? 1e-10f : v);         // NaN stays NaN, as np.maximum / np.log
      if (use_norm) v = __fdiv_rn(v, norm_fac);
      tile[ty + 8 * k][tx] = v;
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int i = 32 * ti + ty + 8 * k, j = 32 * tj + tx;
    if (i < H && j < W) dst[static_cast<size_t>(i) * W + j] = tr ? tile[tx][ty + 8 * k] : tile[ty + 8 * k][tx];
  }
}

// Fast path for rasters whose sides are multiples of 64 (the aerial / DEM tiles): a CTA moves a 64 x 64 tile, every
// thread 4 x (4 consecutive source elements -> one 16-byte store).  Flips only reverse the order of whole 4-element
// chunks and of the elements inside a chunk, so untransposed samples go registers -> HBM directly; transposed samples
// turn the tile through shared memory.
template <typename Raw> struct Raw4;
template <> struct Raw4<unsigned char> {
  __device__ __forceinline__ static void load(const unsigned char* p, float* v) {
    const uchar4 t = *reinterpret_cast<const uchar4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
};
template <> struct Raw4<short> {
  __device__ __forceinline__ static void load(const short* p, float* v) {
    const short4 t = *reinterpret_cast<const short4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
};
template <> struct Raw4<unsigned short> {
  __device__ __forceinline__ static void load(const unsigned short* p, float* v) {
    const ushort4 t = *reinterpret_cast<const ushort4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
};
template <> struct Raw4<float> {
  __device__ __forceinline__ static void load(const float* p, float* v) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
};

template <typename Raw>
__global__ void __launch_bounds__(256)
raster_transform64_kernel(const Raw* __restrict__ raw, float* __restrict__ out, const unsigned char* __restrict__ aug,
                          int imgs_per_sample, int C, int H, int W, int tiles_x, int tiles_per_plane, float norm_fac,
                          int use_norm, int log_scale) {
  __shared__ float tile[64][65];
  const unsigned plane = blockIdx.x / tiles_per_plane;
  const unsigned t = blockIdx.x - plane * tiles_per_plane;
  const int ti = t / tiles_x, tj = t - ti * tiles_x;
  const int img = plane / C;
  const int bits = aug ? aug[img / imgs_per_sample] : 0;
  const bool fh = bits & MB_AUG_FLIP_H, fw = bits & MB_AUG_FLIP_W, tr = bits & MB_AUG_TRANSPOSE;
  const int Hs = tr ? W : H, Ws = tr ? H : W;
  const Raw* src = raw + static_cast<size_t>(plane) * Hs * Ws;
  float* dst = out + static_cast<size_t>(plane) * H * W;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  const int r0 = tr ? 64 * tj : 64 * ti, c0 = tr ? 64 * ti : 64 * tj;
  float v[4][4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int r = r0 + ty + 16 * k, c = c0 + 4 * tx;
    const int sr = fh ? Hs - 1 - r : r, sc = fw ? Ws - 4 - c : c;
    Raw4<Raw>::load(src + static_cast<size_t>(sr) * Ws + sc, v[k]);
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (fw) {
      float a = v[k][0]; v[k][0] = v[k][3]; v[k][3] = a;
      a = v[k][1]; v[k][1] = v[k][2]; v[k][2] = a;
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float x = v[k][e];
      if (log_scale) x = logf(x < 1e-10f ? 1e-10f : x);
      if (use_norm) x = __fdiv_rn(x, norm_fac);
      v[k][e] = x;
    }
  }
  if (!tr) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      *reinterpret_cast<float4*>(dst + static_cast<size_t>(r0 + ty + 16 * k) * W + c0 + 4 * tx) =
          make_float4(v[k][0], v[k][1], v[k][2], v[k][3]);
    return;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k)
#pragma unroll
    for (int e = 0; e < 4; ++e) tile[ty + 16 * k][4 * tx + e] = v[k][e];
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int i = ty + 16 * k;
    *reinterpret_cast<float4*>(dst + static_cast<size_t>(64 * ti + i) * W + 64 * tj + 4 * tx) =
        make_float4(tile[4 * tx][i], tile[4 * tx + 1][i], tile[4 * tx + 2][i], tile[4 * tx + 3][i]);
  }
}

template <typename Raw>
static void launch_raster(const void* raw, float* out, const unsigned char* aug, int n_img, int imgs_per_sample, int C,
                          int H, int W, float norm_fac, int use_norm, int log_scale, cudaStream_t s) {
  const Raw* r = static_cast<const Raw*>(raw);
  const bool fast = H % 64 == 0 && W % 64 == 0 && reinterpret_cast<uintptr_t>(raw) % (4 * sizeof(Raw)) == 0 &&
                    reinterpret_cast<uintptr_t>(out) % 16 == 0;
  const int ts = fast ? 64 : 32;
  const int tiles_x = (W + ts - 1) / ts, tiles_y = (H + ts - 1) / ts;
  const int g = n_img * C * tiles_x * tiles_y, tpp = tiles_x * tiles_y;
  if (fast)
    raster_transform64_kernel<Raw><<<g, 256, 0, s>>>(r, out, aug, imgs_per_sample, C, H, W, tiles_x, tpp, norm_fac,
                                                    use_norm, log_scale);
  else
    raster_transform_kernel<Raw><<<g, 256, 0, s>>>(r, out, aug, imgs_per_sample, C, H, W, tiles_x, tpp, norm_fac,
                                                  use_norm, log_scale);
}

extern "C" int mb_raster_transform(const void* raw, int raw_dtype, float* out, const unsigned char* aug,
                                   int n_img, int imgs_per_sample, int C, int H, int W, float norm_fac,
                                   int use_norm, int log_scale, void* stream) {
  if (n_img <= 0 || C <= 0 || H <= 0 || W <= 0) return 0;
  if (imgs_per_sample <= 0 || raw_dtype < MB_RAW_U8 || raw_dtype > MB_RAW_F32) return static_cast<int>(cudaErrorInvalidValue);
  const long long blocks = static_cast<long long>(n_img) * C * ((W + 31) / 32) * ((H + 31) / 32);
  if (blocks > 0x7fffffffll) return static_cast<int>(cudaErrorInvalidValue);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  switch (raw_dtype) {
    case MB_RAW_U8:
      launch_raster<unsigned char>(raw, out, aug, n_img, imgs_per_sample, C, H, W, norm_fac, use_norm, log_scale, s);
      break;
    case MB_RAW_I16:
      launch_raster<short>(raw, out, aug, n_img, imgs_per_sample, C, H, W, norm_fac, use_norm, log_scale, s);
      break;
    case MB_RAW_U16:
      launch_raster<unsigned short>(raw, out, aug, n_img, imgs_per_sample, C, H, W, norm_fac, use_norm, log_scale, s);
      break;
    default:
      launch_raster<float>(raw, out, aug, n_img, imgs_per_sample, C, H, W, norm_fac, use_norm, log_scale, s);
  }
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ structural mask
struct StructMod {        // one modality of the group, in dataset.groups order
  int G, D, L;            // band groups, dates per band group, locations (grid^2)
  float p_mod, p_bands, p_dates, p_loc;   // <= 0: the draw is skipped (reference: `if self.mask_x[name_mod]`)
};
#define MB_STRUCT_MAX_MODS 8
struct StructParams {
  StructMod mods[MB_STRUCT_MAX_MODS];
  int n_mods;
  int L_group;            // sum of G * D * L
  unsigned long long seed, offset;
};

__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const unsigned hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const unsigned hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u;
    k.y += 0xBB67AE85u;
  }
  return c;
}

// uniform in [0, 1) with 24 random bits; counter = (index, kind | modality << 8, sample, pass + offset)
__device__ __forceinline__ float struct_uniform(const StructParams& P, unsigned sample, unsigned pass, unsigned mod,
                                                unsigned kind, unsigned index) {
  const unsigned long long ctr = P.offset + pass;
  const uint4 c = make_uint4(index, kind | (mod << 8), sample, static_cast<unsigned>(ctr));
  const uint2 k = make_uint2(static_cast<unsigned>(P.seed), static_cast<unsigned>(P.seed >> 32) ^
                                                                static_cast<unsigned>(ctr >> 32));
  return static_cast<float>(philox4x32_10(c, k).x >> 8) * (1.0f / 16777216.0f);
}

// out [B, L_group] u8 (1 = structurally masked); passes [B] (optional): number of redraws of the sample.
__global__ void __launch_bounds__(256) struct_mask_kernel(const StructParams P, unsigned char* __restrict__ out,
                                                          int* __restrict__ passes, int max_pass) {
  const unsigned b = blockIdx.x;
  unsigned char* row = out + static_cast<size_t>(b) * P.L_group;
  int pass = 0;
  for (;; ++pass) {
    int any_open = 0, base = 0;
    for (int m = 0; m < P.n_mods; ++m) {
      const StructMod M = P.mods[m];
      const bool whole = M.p_mod > 0.f && struct_uniform(P, b, pass, m, 0, 0) < M.p_mod;
      const int n = M.G * M.D * M.L;
      for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int l = i % M.L, gd = i / M.L, d = gd % M.D, g = gd / M.D;
        bool mk = whole;
        if (M.p_bands > 0.f) mk |= struct_uniform(P, b, pass, m, 1, g) < M.p_bands;
        if (M.p_dates > 0.f) mk |= struct_uniform(P, b, pass, m, 2, d) < M.p_dates;
        if (M.p_loc > 0.f) mk |= struct_uniform(P, b, pass, m, 3, l) < M.p_loc;
        row[base + i] = mk;
        any_open |= !mk;
      }
      base += n;
    }
    if (__syncthreads_or(any_open) || pass + 1 >= max_pass) break;
  }
  if (passes && threadIdx.x == 0) passes[b] = pass + 1;
}

extern "C" int mb_struct_mask(const StructParams* params, unsigned char* out, int* passes, int B, int max_pass,
                              void* stream) {
  if (B <= 0 || params->L_group <= 0) return 0;
  if (params->n_mods <= 0 || params->n_mods > MB_STRUCT_MAX_MODS) return static_cast<int>(cudaErrorInvalidValue);
  struct_mask_kernel<<<B, 256, 0, static_cast<cudaStream_t>(stream)>>>(*params, out, passes,
                                                                       max_pass > 0 ? max_pass : 1 << 20);
  MB_CHECK_LAUNCH();
  return 0;
}
