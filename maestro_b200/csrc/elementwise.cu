// Bandwidth-bound data-movement kernels of the MAESTRO pretraining step:
//   weight casts (fp32 master -> bf16 shadow, plain and transposed), input resize /
//   elevation rescale (reference maestro/ssl/mim.py:425-437), patch extraction for the
//   strided patch-embed convolution (layers/embed.py:57-60), date features
//   (layers/utils.py:128-173), prediction rows -> image (layers/embed.py:148-160).
#include "common.cuh"

// ------------------------------------------------------------------ casts
__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, bf16* __restrict__ dst, size_t n) {
  size_t i = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x * 8;
  for (; i + 8 <= n; i += stride) {
    float v[8];
    VecIO<float>::load(src + i, v);
    VecIO<float>::load(src + i + 4, v + 4);
    VecIO<bf16>::store(dst + i, v);
  }
  if (i < n) {
    for (size_t j = i; j < n && j < i + 8; ++j) dst[j] = __float2bfloat16_rn(src[j]);
  }
}

extern "C" int mb_cast_f32_to_bf16(const float* src, void* dst, size_t n, void* stream) {
  if (n == 0) return 0;
  const int threads = 256;
  long long blocks = (static_cast<long long>(n) / 8 + threads - 1) / threads;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  cast_f32_bf16_kernel<<<static_cast<int>(blocks), threads, 0, static_cast<cudaStream_t>(stream)>>>(
      src, static_cast<bf16*>(dst), n);
  MB_CHECK_LAUNCH();
  return 0;
}

// dst[c, r] = bf16(src[r, c]); src is [R, C] row-major fp32, dst is [C, R] row-major.
__global__ void cast_transpose_kernel(const float* __restrict__ src, bf16* __restrict__ dst, int R, int C) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (r < R && c < C) ? src[static_cast<size_t>(r) * C + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (c < C && r < R) dst[static_cast<size_t>(c) * R + r] = __float2bfloat16_rn(tile[threadIdx.x][i]);
  }
}

extern "C" int mb_cast_transpose_f32_to_bf16(const float* src, void* dst, int R, int C, void* stream) {
  if (R == 0 || C == 0) return 0;
  dim3 grid(mb_ceil_div(C, 32), mb_ceil_div(R, 32)), block(32, 8);
  cast_transpose_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(src, static_cast<bf16*>(dst), R, C);
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ resize / rescale
// Nearest-neighbour resize as torch F.interpolate(mode="nearest"): src = min(floor(dst * in/out), in-1)
// with the scale held in fp32.  Optionally applies the elevation rescale in the same pass:
// ch>=1 of a rescale_elev modality becomes 30 * (ch0 - ch).
template <typename I>   // index type: unsigned when the element count fits 32 bits (64-bit divisions cost ~10x)
__global__ void resize_nearest_kernel(const float* __restrict__ src, float* __restrict__ dst, int n_img, int C,
                                      int hin, int win, int hout, int wout, int rescale_elev) {
  const I total = static_cast<I>(n_img) * C * hout * wout;
  const float sh = static_cast<float>(hin) / static_cast<float>(hout);
  const float sw = static_cast<float>(win) / static_cast<float>(wout);
  for (I i = static_cast<I>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<I>(gridDim.x) * blockDim.x) {
    const I t0 = i / wout;
    const int x = static_cast<int>(i - t0 * wout);
    const I t1 = t0 / hout;
    const int y = static_cast<int>(t0 - t1 * hout);
    const I img = t1 / C;
    const int c = static_cast<int>(t1 - img * C);
    const int sy = min(static_cast<int>(floorf(y * sh)), hin - 1);
    const int sx = min(static_cast<int>(floorf(x * sw)), win - 1);
    const float* base = src + static_cast<size_t>(img) * C * hin * win;
    float v = base[(static_cast<size_t>(c) * hin + sy) * win + sx];
    if (rescale_elev && c >= 1) {
      const float v0 = base[static_cast<size_t>(sy) * win + sx];
      v = 30.f * (v0 - v);
    }
    dst[i] = v;
  }
}

extern "C" int mb_resize_nearest(const float* src, float* dst, int n_img, int C, int hin, int win, int hout,
                                 int wout, int rescale_elev, void* stream) {
  const size_t total = static_cast<size_t>(n_img) * C * hout * wout;
  if (total == 0) return 0;
  long long blocks = (static_cast<long long>(total) + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  if (total < 0xffffffffull)
    resize_nearest_kernel<unsigned><<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        src, dst, n_img, C, hin, win, hout, wout, rescale_elev);
  else
    resize_nearest_kernel<size_t><<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        src, dst, n_img, C, hin, win, hout, wout, rescale_elev);
  MB_CHECK_LAUNCH();
  return 0;
}

// Bilinear / bicubic resize as torch F.interpolate(mode=..., align_corners=False) (upsample_bilinear2d /
// upsample_bicubic2d, A = -0.75): src = scale * (dst + 0.5) - 0.5 with scale = in / out in fp32; bilinear clamps the
// coordinate at 0 and the second tap at in - 1, bicubic clamps every tap index.  Same optional elevation rescale.
#define MB_RESIZE_NEAREST 0
#define MB_RESIZE_BILINEAR 1
#define MB_RESIZE_BICUBIC 2

__device__ __forceinline__ void cubic_coeffs(float t, float* w) {
  const float A = -0.75f;
  const float x0 = t + 1.f, x1 = t, x2 = 1.f - t, x3 = 2.f - t;
  w[0] = ((A * x0 - 5.f * A) * x0 + 8.f * A) * x0 - 4.f * A;
  w[1] = ((A + 2.f) * x1 - (A + 3.f)) * x1 * x1 + 1.f;
  w[2] = ((A + 2.f) * x2 - (A + 3.f)) * x2 * x2 + 1.f;
  w[3] = ((A * x3 - 5.f * A) * x3 + 8.f * A) * x3 - 4.f * A;
}

template <int MODE>
__device__ __forceinline__ float sample_plane(const float* __restrict__ pl, int hin, int win, float fy, float fx) {
  if (MODE == MB_RESIZE_BILINEAR) {
    fy = fmaxf(fy, 0.f);
    fx = fmaxf(fx, 0.f);
    const int y0 = min(static_cast<int>(fy), hin - 1), x0 = min(static_cast<int>(fx), win - 1);
    const int y1 = y0 + (y0 < hin - 1), x1 = x0 + (x0 < win - 1);
    const float ly = fy - y0, lx = fx - x0, hy = 1.f - ly, hx = 1.f - lx;
    return hy * (hx * pl[y0 * win + x0] + lx * pl[y0 * win + x1]) +
           ly * (hx * pl[y1 * win + x0] + lx * pl[y1 * win + x1]);
  }
  const float yf = floorf(fy), xf = floorf(fx);
  const int iy = static_cast<int>(yf), ix = static_cast<int>(xf);
  float wy[4], wx[4];
  cubic_coeffs(fy - yf, wy);
  cubic_coeffs(fx - xf, wx);
  float acc = 0.f;
#pragma unroll
  for (int a = 0; a < 4; ++a) {
    const int yy = min(max(iy - 1 + a, 0), hin - 1);
    float row = 0.f;
#pragma unroll
    for (int b = 0; b < 4; ++b) row += wx[b] * pl[yy * win + min(max(ix - 1 + b, 0), win - 1)];
    acc += wy[a] * row;
  }
  return acc;
}

template <int MODE>
__global__ void resize_interp_kernel(const float* __restrict__ src, float* __restrict__ dst, int n_img, int C, int hin,
                                     int win, int hout, int wout, int rescale_elev) {
  const size_t total = static_cast<size_t>(n_img) * C * hout * wout;
  const float sh = static_cast<float>(hin) / static_cast<float>(hout);
  const float sw = static_cast<float>(win) / static_cast<float>(wout);
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const size_t t0 = i / wout;
    const int x = static_cast<int>(i - t0 * wout);
    const size_t t1 = t0 / hout;
    const int y = static_cast<int>(t0 - t1 * hout);
    const size_t img = t1 / C;
    const int c = static_cast<int>(t1 - img * C);
    const float fy = sh * (y + 0.5f) - 0.5f, fx = sw * (x + 0.5f) - 0.5f;
    const float* base = src + img * C * hin * win;
    float v = sample_plane<MODE>(base + static_cast<size_t>(c) * hin * win, hin, win, fy, fx);
    if (rescale_elev && c >= 1) v = 30.f * (sample_plane<MODE>(base, hin, win, fy, fx) - v);
    dst[i] = v;
  }
}

extern "C" int mb_resize_nearest(const float* src, float* dst, int n_img, int C, int hin, int win, int hout,
                                 int wout, int rescale_elev, void* stream);

extern "C" int mb_resize(const float* src, float* dst, int n_img, int C, int hin, int win, int hout, int wout, int mode,
                         int rescale_elev, void* stream) {
  if (mode == MB_RESIZE_NEAREST) return mb_resize_nearest(src, dst, n_img, C, hin, win, hout, wout, rescale_elev, stream);
  if (mode != MB_RESIZE_BILINEAR && mode != MB_RESIZE_BICUBIC) return -1;
  const size_t total = static_cast<size_t>(n_img) * C * hout * wout;
  if (total == 0) return 0;
  long long blocks = (static_cast<long long>(total) + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (mode == MB_RESIZE_BILINEAR)
    resize_interp_kernel<MB_RESIZE_BILINEAR><<<static_cast<int>(blocks), 256, 0, s>>>(src, dst, n_img, C, hin, win, hout,
                                                                                      wout, rescale_elev);
  else
    resize_interp_kernel<MB_RESIZE_BICUBIC><<<static_cast<int>(blocks), 256, 0, s>>>(src, dst, n_img, C, hin, win, hout,
                                                                                     wout, rescale_elev);
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ patch extraction
// img [n_img, C, S, S] fp32 -> cols [n_img * g * g, Kpad] (T), column k = (c - c0) * p * p + py * p + px
// (the flattening of Conv2d weight [E, c, p, p]); columns [K, Kpad) are zero filled.
template <typename T>
__global__ void im2col_kernel(const float* __restrict__ img, T* __restrict__ cols, int n_img, int C, int c0, int c,
                              int S, int p, int Kpad) {
  const int g = S / p;
  const int K = c * p * p;
  const size_t total = static_cast<size_t>(n_img) * g * g * Kpad;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int k = static_cast<int>(i % Kpad);
    float v = 0.f;
    if (k < K) {
      size_t row = i / Kpad;
      const int gx = static_cast<int>(row % g);
      row /= g;
      const int gy = static_cast<int>(row % g);
      const size_t im = row / g;
      const int px = k % p;
      const int py = (k / p) % p;
      const int cc = k / (p * p);
      v = img[((im * C + c0 + cc) * S + gy * p + py) * S + gx * p + px];
    }
    cols[i] = from_f32<T>(v);
  }
}

// Vector form for p % 4 == 0: one thread moves 4 consecutive pixels of one patch row (a 16-byte image load, an 8- or
// 16-byte store), 32-bit index arithmetic.
template <typename T>
__global__ void __launch_bounds__(256) im2col_vec4_kernel(const float* __restrict__ img, T* __restrict__ cols,
                                                          unsigned total4, int C, int c0, int c, int S, int p,
                                                          int Kpad) {
  const unsigned g = S / p, K = c * p * p, K4 = Kpad / 4, pp = p * p;
  for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < total4; q += gridDim.x * blockDim.x) {
    const unsigned row = q / K4, k = (q - row * K4) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < K) {
      const unsigned t = row / g, gx = row - t * g, im = t / g, gy = t - im * g;
      const unsigned cc = k / pp, r = k - cc * pp, py = r / p, px = r - py * p;
      v = __ldg(reinterpret_cast<const float4*>(
          img + ((static_cast<size_t>(im) * C + c0 + cc) * S + gy * p + py) * S + gx * p + px));
    }
    T* o = cols + static_cast<size_t>(q) * 4;
    if constexpr (sizeof(T) == 4) {
      *reinterpret_cast<float4*>(o) = v;
    } else {
      __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
      *reinterpret_cast<uint2*>(o) = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
    }
  }
}

extern "C" int mb_im2col(const float* img, void* cols, int dtype, int n_img, int C, int c0, int c, int S, int p,
                         int Kpad, void* stream) {
  const int g = S / p;
  const size_t total = static_cast<size_t>(n_img) * g * g * Kpad;
  if (total == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool vec = p % 4 == 0 && S % 4 == 0 && Kpad % 4 == 0 && total / 4 < 0xffffffffull &&
                   (reinterpret_cast<uintptr_t>(img) & 15) == 0 && (reinterpret_cast<uintptr_t>(cols) & 15) == 0;
  if (vec) {
    const unsigned total4 = static_cast<unsigned>(total / 4);
    long long blocks = (static_cast<long long>(total4) + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    MB_DISPATCH_DTYPE(dtype, T, (im2col_vec4_kernel<T><<<static_cast<int>(blocks), 256, 0, st>>>(
                                    img, static_cast<T*>(cols), total4, C, c0, c, S, p, Kpad)));
    MB_CHECK_LAUNCH();
    return 0;
  }
  long long blocks = (static_cast<long long>(total) + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  MB_DISPATCH_DTYPE(dtype, T,
                    (im2col_kernel<T><<<static_cast<int>(blocks), 256, 0, st>>>(
                        img, static_cast<T*>(cols), n_img, C, c0, c, S, p, Kpad)));
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ prediction rows -> image
// rows [n_img * g * g, c * p * p] (T) with column order (py, px, cc) -> img [n_img, C, S, S] fp32 channels
// [c0, c0 + c)   (PixelifyBands rearrange, layers/embed.py:155-160).  dir = 0: rows -> img; dir = 1: img -> rows
// (the adjoint, used for the gradient of an image-layout loss).
template <typename T>
__global__ void rows_image_kernel(T* __restrict__ rows, float* __restrict__ img, int n_img, int C, int c0, int c,
                                  int S, int p, int dir) {
  const int g = S / p;
  const int K = c * p * p;
  const size_t total = static_cast<size_t>(n_img) * g * g * K;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    // iterate in image order for coalesced image access
    const int x = static_cast<int>(i % S);
    size_t t = i / S;
    const int y = static_cast<int>(t % S);
    t /= S;
    const int cc = static_cast<int>(t % c);
    const size_t im = t / c;
    const int gx = x / p, px = x % p, gy = y / p, py = y % p;
    const size_t r = (im * g + gy) * g + gx;
    const size_t ri = r * K + (static_cast<size_t>(py) * p + px) * c + cc;
    const size_t ii = ((im * C + c0 + cc) * S + y) * S + x;
    if (dir == 0) img[ii] = to_f32(rows[ri]);
    else rows[ri] = from_f32<T>(img[ii]);
  }
}

extern "C" int mb_rows_image(void* rows, int dtype, float* img, int n_img, int C, int c0, int c, int S, int p,
                             int dir, void* stream) {
  const size_t total = static_cast<size_t>(n_img) * S * S * c;
  if (total == 0) return 0;
  long long blocks = (static_cast<long long>(total) + 255) / 256;
  if (blocks > 148 * 32) blocks = 148 * 32;
  MB_DISPATCH_DTYPE(dtype, T,
                    (rows_image_kernel<T><<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<T*>(rows), img, n_img, C, c0, c, S, p, dir)));
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ date features
// dates int16 [B, D, 3] (year, day-of-year, hour), ref int16 [B, 1, 3] -> feat fp32 [B, Dtot, 5] rows doff..doff+D =
// (diff, sin doy, cos doy, sin hour, cos hour) with the reference's fp32 operation order.
__global__ void date_features_kernel(const short* __restrict__ dates, const short* __restrict__ ref,
                                     float* __restrict__ feat_base, int B, int D, int Dtot, int doff) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * D) return;
  const int b = i / D;
  float* feat = feat_base + (static_cast<size_t>(b) * Dtot + doff + (i % D)) * 5 - static_cast<size_t>(i) * 5;
  const float year = static_cast<float>(dates[i * 3 + 0]);
  float doy = __fdiv_rn(static_cast<float>(dates[i * 3 + 1]), 365.25f);
  float hour = __fdiv_rn(static_cast<float>(dates[i * 3 + 2]), 24.0f);
  const float ry = static_cast<float>(ref[b * 3 + 0]);
  const float rd = __fdiv_rn(static_cast<float>(ref[b * 3 + 1]), 365.25f);
  const float diff = __fsub_rn(__fadd_rn(year, doy), __fadd_rn(ry, rd));
  const float two_pi = 6.283185307179586f;
  doy = __fmul_rn(two_pi, doy);
  hour = __fmul_rn(two_pi, hour);
  feat[i * 5 + 0] = diff;
  feat[i * 5 + 1] = sinf(doy);
  feat[i * 5 + 2] = cosf(doy);
  feat[i * 5 + 3] = sinf(hour);
  feat[i * 5 + 4] = cosf(hour);
}

extern "C" int mb_date_features(const short* dates, const short* ref, float* feat, int B, int D, int Dtot, int doff,
                                void* stream) {
  if (B * D == 0) return 0;
  date_features_kernel<<<mb_ceil_div(B * D, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(dates, ref, feat, B, D,
                                                                                              Dtot, doff);
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ column sums (bias gradients)
// out[n] += sum_m dy[m, n]  (fp32).  Each thread owns one 16-byte column vector and strides over the rows of its
// row slab with 4 loads in flight; the CTA's warps cover 8 row phases that are combined in shared memory, then one
// red.add per column per CTA.
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* __restrict__ dy, float* __restrict__ out, int M, int N,
                                                     int rows_per_block) {
  constexpr int V = VecIO<T>::N;
  __shared__ float red[8][32 * V];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + lane) * V;
  const int m0 = blockIdx.y * rows_per_block;
  const int m1 = min(M, m0 + rows_per_block);
  float acc[V];
#pragma unroll
  for (int j = 0; j < V; ++j) acc[j] = 0.f;
  if (col < N) {
    int m = m0 + w;
    for (; m + 24 < m1; m += 32) {
      float v[4][V];
#pragma unroll
      for (int u = 0; u < 4; ++u) VecIO<T>::load(dy + static_cast<size_t>(m + 8 * u) * N + col, v[u]);
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] += v[u][j];
    }
    for (; m < m1; m += 8) {
      float v[V];
      VecIO<T>::load(dy + static_cast<size_t>(m) * N + col, v);
#pragma unroll
      for (int j = 0; j < V; ++j) acc[j] += v[j];
    }
  }
#pragma unroll
  for (int j = 0; j < V; ++j) red[w][lane * V + j] = acc[j];
  __syncthreads();
  for (int i = threadIdx.x; i < 32 * V; i += blockDim.x) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[k][i];
    const int c = blockIdx.x * 32 * V + i;
    if (c < N) atomicAdd(out + c, s);
  }
}

// Any N (e.g. a head's num_classes = 15): lane = column, warps stride over the rows, element-wise loads.
template <typename T>
__global__ void __launch_bounds__(256) colsum_scalar_kernel(const T* __restrict__ dy, float* __restrict__ out, int M,
                                                            int N, int rows_per_block) {
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int col = blockIdx.x * 32 + lane;
  const int m0 = blockIdx.y * rows_per_block;
  const int m1 = min(M, m0 + rows_per_block);
  float acc = 0.f;
  if (col < N)
    for (int m = m0 + w; m < m1; m += 8) acc += to_f32(dy[static_cast<size_t>(m) * N + col]);
  red[w][lane] = acc;
  __syncthreads();
  if (w == 0 && col < N) {
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k) s += red[k][lane];
    atomicAdd(out + col, s);
  }
}

extern "C" int mb_colsum_accum(const void* dy, int dtype, float* out, int M, int N, void* stream) {
  if (M == 0 || N == 0) return 0;
  if (N % 8 != 0) {
    int rows_per_block = mb_ceil_div(M, 148);
    if (rows_per_block < 64) rows_per_block = 64;
    dim3 grid(mb_ceil_div(N, 32), mb_ceil_div(M, rows_per_block));
    MB_DISPATCH_DTYPE(dtype, T,
                      (colsum_scalar_kernel<T><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                          static_cast<const T*>(dy), out, M, N, rows_per_block)));
    MB_CHECK_LAUNCH();
    return 0;
  }
  const int V = dtype == MB_DTYPE_BF16 ? 8 : 4;
  const int bx = mb_ceil_div(N, 32 * V);
  int by = (148 * 4) / bx;
  if (by < 1) by = 1;
  int rows_per_block = mb_ceil_div(M, by);
  if (rows_per_block < 64) rows_per_block = 64;
  by = mb_ceil_div(M, rows_per_block);
  dim3 grid(bx, by);
  MB_DISPATCH_DTYPE(dtype, T,
                    (colsum_kernel<T><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(dy), out, M, N, rows_per_block)));
  MB_CHECK_LAUNCH();
  return 0;
}
