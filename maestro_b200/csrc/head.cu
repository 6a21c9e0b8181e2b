// Probe / finetune side of the MAE and the per-epoch EMA (8f-1):
//   AttentiveReduce          maestro/layers/head.py:28-64    single learned query per head attending over a sequence
//   compute_logits resample  maestro/ssl/mim.py:353-372      bilinear resize of every modality's token grid to the
//                                                             reference grid (F.interpolate on "(b d) c h w")
//   compute_loss_pred        maestro/train/base.py:98-151    cross entropy over (pixels | samples) with a missing value
//   update_ema               maestro/train/base.py:263-274   ema = momentum * ema + (1 - momentum) * param
#include "common.cuh"

// ------------------------------------------------------------------ attentive pooling
// kv [S, L, 2 * C] (k = columns [0, C), v = [C, 2C); head h owns columns h*dh .. h*dh+dh), q [C] fp32 (shared by all
// sequences).  One CTA (4 warps) per (sequence, head): warps stride over the L tokens, lanes over the head dimension.
//   dots[l] = scale * <q_h, k[l]>;  p = softmax_l(dots);  out[s, h*dh + :] = sum_l p[l] * v[l]
// probs [S, H, L] fp32 is saved for the backward.
constexpr int AP_WARPS = 4;
constexpr int AP_MAX_DPL = 4;      // head dim <= 128

template <typename T>
__global__ void __launch_bounds__(AP_WARPS * 32) attn_pool_fwd_kernel(const T* __restrict__ kv, const float* __restrict__ q,
                                                                      T* __restrict__ out, float* __restrict__ probs,
                                                                      int L, int H, int dh, float scale) {
  extern __shared__ float sm[];                 // dots[L] | warp partials
  float* dots = sm;
  __shared__ float red[AP_WARPS];
  __shared__ float accs[AP_WARPS][128];
  const int s = blockIdx.x / H, h = blockIdx.x % H;
  const int C = H * dh, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const T* base = kv + static_cast<size_t>(s) * L * 2 * C + h * dh;
  float qv[AP_MAX_DPL];
#pragma unroll
  for (int j = 0; j < AP_MAX_DPL; ++j) qv[j] = (lane + 32 * j < dh) ? q[h * dh + lane + 32 * j] * scale : 0.f;
  float mx = -INFINITY;
  for (int l = w; l < L; l += AP_WARPS) {
    const T* kr = base + static_cast<size_t>(l) * 2 * C;
    float d = 0.f;
#pragma unroll
    for (int j = 0; j < AP_MAX_DPL; ++j)
      if (lane + 32 * j < dh) d = fmaf(qv[j], to_f32(kr[lane + 32 * j]), d);
    d = warp_sum(d);
    if (lane == 0) dots[l] = d;
    mx = fmaxf(mx, d);
  }
  if (lane == 0) red[w] = mx;
  __syncthreads();
  mx = red[0];
#pragma unroll
  for (int k = 1; k < AP_WARPS; ++k) mx = fmaxf(mx, red[k]);
  __syncthreads();
  float sum = 0.f;
  for (int l = threadIdx.x; l < L; l += blockDim.x) {
    const float e = __expf(dots[l] - mx);
    dots[l] = e;
    sum += e;
  }
  sum = warp_sum(sum);
  if (lane == 0) red[w] = sum;
  __syncthreads();
  sum = 0.f;
#pragma unroll
  for (int k = 0; k < AP_WARPS; ++k) sum += red[k];
  const float inv = 1.0f / sum;
  float acc[AP_MAX_DPL] = {0.f, 0.f, 0.f, 0.f};
  for (int l = w; l < L; l += AP_WARPS) {
    const float p = dots[l] * inv;
    if (lane == 0) probs[(static_cast<size_t>(s) * H + h) * L + l] = p;
    const T* vr = base + static_cast<size_t>(l) * 2 * C + C;
#pragma unroll
    for (int j = 0; j < AP_MAX_DPL; ++j)
      if (lane + 32 * j < dh) acc[j] = fmaf(p, to_f32(vr[lane + 32 * j]), acc[j]);
  }
#pragma unroll
  for (int j = 0; j < AP_MAX_DPL; ++j) accs[w][lane + 32 * j] = acc[j];
  __syncthreads();
  for (int e = threadIdx.x; e < dh; e += blockDim.x) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < AP_WARPS; ++k) v += accs[k][e];
    out[static_cast<size_t>(s) * C + h * dh + e] = from_f32<T>(v);
  }
}

// Backward: dout [S, C] -> dkv [S, L, 2C] (every element written) and dq [C] (fp32, accumulated).
//   dv[l] = p[l] * dout;  dp[l] = <dout, v[l]>;  ds[l] = p[l] * (dp[l] - sum_l p dp);  dk[l] = ds[l] * scale * q
//   dq += scale * sum_l ds[l] * k[l]
template <typename T>
__global__ void __launch_bounds__(AP_WARPS * 32) attn_pool_bwd_kernel(const T* __restrict__ kv, const float* __restrict__ q,
                                                                      const float* __restrict__ probs,
                                                                      const T* __restrict__ dout, T* __restrict__ dkv,
                                                                      float* __restrict__ dq, int L, int H, int dh,
                                                                      float scale) {
  extern __shared__ float sm[];                 // dp[L]
  float* dp = sm;
  __shared__ float red[AP_WARPS];
  __shared__ float accs[AP_WARPS][128];
  const int s = blockIdx.x / H, h = blockIdx.x % H;
  const int C = H * dh, lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const size_t row0 = static_cast<size_t>(s) * L * 2 * C + h * dh;
  const float* pr = probs + (static_cast<size_t>(s) * H + h) * L;
  float go[AP_MAX_DPL], qv[AP_MAX_DPL];
#pragma unroll
  for (int j = 0; j < AP_MAX_DPL; ++j) {
    const bool ok = lane + 32 * j < dh;
    go[j] = ok ? to_f32(dout[static_cast<size_t>(s) * C + h * dh + lane + 32 * j]) : 0.f;
    qv[j] = ok ? q[h * dh + lane + 32 * j] * scale : 0.f;
  }
  float spd = 0.f;
  for (int l = w; l < L; l += AP_WARPS) {
    const T* vr = kv + row0 + static_cast<size_t>(l) * 2 * C + C;
    float d = 0.f;
#pragma unroll
    for (int j = 0; j < AP_MAX_DPL; ++j)
      if (lane + 32 * j < dh) d = fmaf(go[j], to_f32(vr[lane + 32 * j]), d);
    d = warp_sum(d);
    if (lane == 0) dp[l] = d;
    spd = fmaf(pr[l], d, spd);
  }
  if (lane == 0) red[w] = spd;
  __syncthreads();
  spd = 0.f;
#pragma unroll
  for (int k = 0; k < AP_WARPS; ++k) spd += red[k];
  float aq[AP_MAX_DPL] = {0.f, 0.f, 0.f, 0.f};
  for (int l = w; l < L; l += AP_WARPS) {
    const float p = pr[l];
    const float ds = p * (dp[l] - spd);
    const T* kr = kv + row0 + static_cast<size_t>(l) * 2 * C;
    T* dkr = dkv + row0 + static_cast<size_t>(l) * 2 * C;
#pragma unroll
    for (int j = 0; j < AP_MAX_DPL; ++j) {
      if (lane + 32 * j < dh) {
        aq[j] = fmaf(ds, to_f32(kr[lane + 32 * j]), aq[j]);
        dkr[lane + 32 * j] = from_f32<T>(ds * qv[j]);
        dkr[C + lane + 32 * j] = from_f32<T>(p * go[j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < AP_MAX_DPL; ++j) accs[w][lane + 32 * j] = aq[j];
  __syncthreads();
  for (int e = threadIdx.x; e < dh; e += blockDim.x) {
    float v = 0.f;
#pragma unroll
    for (int k = 0; k < AP_WARPS; ++k) v += accs[k][e];
    atomicAdd(dq + h * dh + e, v * scale);
  }
}

extern "C" int mb_attn_pool_fwd(const void* kv, const float* q, void* out, float* probs, int dtype, int S, int L, int H,
                                int dh, float scale, void* stream) {
  if (S == 0 || L == 0) return 0;
  if (dh > 32 * AP_MAX_DPL || static_cast<size_t>(L) * sizeof(float) > 160 * 1024) return -2;
  const size_t smem = static_cast<size_t>(L) * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  MB_DISPATCH_DTYPE(dtype, T, {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(attn_pool_fwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    attn_pool_fwd_kernel<T><<<S * H, AP_WARPS * 32, smem, st>>>(static_cast<const T*>(kv), q, static_cast<T*>(out),
                                                                 probs, L, H, dh, scale);
  });
  MB_CHECK_LAUNCH();
  return 0;
}

extern "C" int mb_attn_pool_bwd(const void* kv, const float* q, const float* probs, const void* dout, void* dkv,
                                float* dq, int dtype, int S, int L, int H, int dh, float scale, void* stream) {
  if (S == 0 || L == 0) return 0;
  if (dh > 32 * AP_MAX_DPL || static_cast<size_t>(L) * sizeof(float) > 160 * 1024) return -2;
  const size_t smem = static_cast<size_t>(L) * sizeof(float);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  MB_DISPATCH_DTYPE(dtype, T, {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(attn_pool_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    attn_pool_bwd_kernel<T><<<S * H, AP_WARPS * 32, smem, st>>>(static_cast<const T*>(kv), q, probs,
                                                                 static_cast<const T*>(dout), static_cast<T*>(dkv), dq, L,
                                                                 H, dh, scale);
  });
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ token-grid resampling
// out[n, oy * wo + ox, :] = sum_{iy, ix} Wy[oy, iy] * Wx[ox, ix] * x[n, iy * wi + ix, :]
// x [n_img, hi * wi, d] channel-last token grids; Wy [ho, hi], Wx [wo, wi] dense fp32 1-D interpolation matrices (the
// host builds them from F.interpolate's own bilinear weights; the backward is the same kernel with the transposes).
// Rows of the matrices hold at most a few non-zeros: zero weights are skipped.  One warp per output token.
template <typename T>
__global__ void __launch_bounds__(256) token_resample_kernel(const T* __restrict__ x, const float* __restrict__ Wy,
                                                             const float* __restrict__ Wx, T* __restrict__ out,
                                                             int n_img, int hi, int wi, int ho, int wo, int d,
                                                             long long in_bs, long long out_bs) {
  constexpr int V = VecIO<T>::N;
  const long long row = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (row >= static_cast<long long>(n_img) * ho * wo) return;
  const int n = static_cast<int>(row / (ho * wo)), o = static_cast<int>(row % (ho * wo));
  const int oy = o / wo, ox = o % wo;
  const T* xin = x + static_cast<size_t>(n) * in_bs;
  T* dst = out + static_cast<size_t>(n) * out_bs + static_cast<size_t>(o) * d;
  for (int c = lane; c < d / V; c += 32) {
    float acc[V];
#pragma unroll
    for (int j = 0; j < V; ++j) acc[j] = 0.f;
    for (int iy = 0; iy < hi; ++iy) {
      const float wy = Wy[oy * hi + iy];
      if (wy == 0.f) continue;
      for (int ix = 0; ix < wi; ++ix) {
        const float wgt = wy * Wx[ox * wi + ix];
        if (wgt == 0.f) continue;
        float v[V];
        VecIO<T>::load(xin + (static_cast<size_t>(iy) * wi + ix) * d + c * V, v);
#pragma unroll
        for (int j = 0; j < V; ++j) acc[j] = fmaf(wgt, v[j], acc[j]);
      }
    }
    VecIO<T>::store(dst + c * V, acc);
  }
}

extern "C" int mb_token_resample(const void* x, const float* Wy, const float* Wx, void* out, int dtype, int n_img,
                                 int hi, int wi, int ho, int wo, int d, long long in_bs, long long out_bs,
                                 void* stream) {
  if (n_img == 0 || ho * wo == 0) return 0;
  if (d % 8 != 0) return -2;
  const int blocks = mb_ceil_div(static_cast<long long>(n_img) * ho * wo * 32, 256);
  MB_DISPATCH_DTYPE(dtype, T,
                    (token_resample_kernel<T><<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(x), Wy, Wx, static_cast<T*>(out), n_img, hi, wi, ho, wo, d, in_bs, out_bs)));
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ cross entropy with a missing value
// logits element (n, c) lives at (n / inner) * outer_stride + c * class_stride + (n % inner): "(b h w) c" rows of a
// [B, 1, C, H, W] segmentation output (inner = H * W, class_stride = H * W, outer_stride = C * H * W) or plain [N, C]
// rows (inner = 1, class_stride = 1, outer_stride = C).  Rows whose target equals `missing` are skipped
// (train/base.py:121-140).  accum[0] += sum of -log softmax[target], accum[1] += rows counted; dlogits (same layout,
// may be null) = softmax - onehot for counted rows, 0 for skipped ones (the caller scales by 1 / count).
__global__ void __launch_bounds__(256) xent_kernel(const float* __restrict__ logits, const long long* __restrict__ target,
                                                   float* __restrict__ dlogits, double* __restrict__ accum,
                                                   long long N, int C, long long inner, long long class_stride,
                                                   long long outer_stride, long long missing) {
  const long long n = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  float loss = 0.f, cnt = 0.f;
  if (n < N) {
    const size_t base = static_cast<size_t>(n / inner) * outer_stride + static_cast<size_t>(n % inner);
    const long long t = target[n];
    if (t != missing) {
      float mx = -INFINITY;
      for (int c = 0; c < C; ++c) mx = fmaxf(mx, logits[base + c * class_stride]);
      float sum = 0.f;
      for (int c = 0; c < C; ++c) sum += __expf(logits[base + c * class_stride] - mx);
      const float lse = mx + __logf(sum);
      loss = lse - logits[base + t * class_stride];
      cnt = 1.f;
      if (dlogits != nullptr)
        for (int c = 0; c < C; ++c)
          dlogits[base + c * class_stride] = __expf(logits[base + c * class_stride] - lse) - (c == t ? 1.f : 0.f);
    } else if (dlogits != nullptr) {
      for (int c = 0; c < C; ++c) dlogits[base + c * class_stride] = 0.f;
    }
  }
  loss = warp_sum(loss);
  cnt = warp_sum(cnt);
  __shared__ float sl[8], sc[8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) { sl[w] = loss; sc[w] = cnt; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0.0, b = 0.0;
    for (int k = 0; k < 8; ++k) { a += sl[k]; b += sc[k]; }
    if (b > 0.0) { atomicAdd(accum, a); atomicAdd(accum + 1, b); }
  }
}

extern "C" int mb_xent_fwd_bwd(const float* logits, const long long* target, float* dlogits, double* accum, long long N,
                               int C, long long inner, long long class_stride, long long outer_stride, long long missing,
                               void* stream) {
  if (N == 0) return 0;
  const int blocks = mb_ceil_div(N, 256);
  xent_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(logits, target, dlogits, accum, N, C, inner,
                                                                      class_stride, outer_stride, missing);
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ EMA of the parameter arena
// ema = momentum * ema + (1 - momentum) * param over the flat fp32 arena (BaseModule.update_ema loops over the
// parameter list with two foreach ops, train/base.py:263-274): one pass, 12 B / parameter.
__global__ void __launch_bounds__(256) ema_kernel(float* __restrict__ ema, const float* __restrict__ param, size_t n,
                                                  float momentum, float om) {
  size_t i = (static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x) * 4;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x * 4;
  for (; i + 4 <= n; i += stride) {
    float4 e = *reinterpret_cast<const float4*>(ema + i);
    const float4 p = *reinterpret_cast<const float4*>(param + i);
    // param_ema.mul_(momentum).add_((1 - momentum) * param): two rounded steps, as the reference's foreach ops
    e.x = __fadd_rn(__fmul_rn(e.x, momentum), __fmul_rn(om, p.x));
    e.y = __fadd_rn(__fmul_rn(e.y, momentum), __fmul_rn(om, p.y));
    e.z = __fadd_rn(__fmul_rn(e.z, momentum), __fmul_rn(om, p.z));
    e.w = __fadd_rn(__fmul_rn(e.w, momentum), __fmul_rn(om, p.w));
    *reinterpret_cast<float4*>(ema + i) = e;
  }
  for (size_t j = i; j < n && j < i + 4; ++j) ema[j] = __fadd_rn(__fmul_rn(ema[j], momentum), __fmul_rn(om, param[j]));
}

// `one_minus` = float(1.0 - momentum) as the host forms it in double precision (not 1.0f - momentum).
extern "C" int mb_ema_update(float* ema, const float* param, size_t n, float momentum, float one_minus, void* stream) {
  if (n == 0) return 0;
  long long blocks = (static_cast<long long>(n) / 4 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  ema_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(ema, param, n, momentum, one_minus);
  MB_CHECK_LAUNCH();
  return 0;
}
