// Patch-group-wise normalised masked reconstruction loss, forward + gradient in one pass
// (reference maestro/train/model.py:195-247, norm_bands / loss_fn select :38-67).
//
// One warp per patch (b, d, l) of one modality.  The target patch (all C channels) is staged in shared
// memory, per norm-band-group mean / unbiased variance are reduced with warp shuffles, and the prediction
// row(s) are streamed once: error accumulated, d(loss)/d(pred) written (zero rows for unmasked patches,
// whose target is never read).  Algorithmic bytes: masked_fraction * target + pred read + dpred write.
#include "common.cuh"

constexpr int LOSS_MAX_C = 16;
constexpr int LOSS_MAX_G = 4;

struct LossParams {
  const float* target;            // [B*D, C, S, S] fp32 (the resized / rescaled batch)
  const void* pred[LOSS_MAX_G];   // per band-group rows [B*D*L, c_g*p*p] (T), column order (py, px, cl)
  void* dpred[LOSS_MAX_G];        // same shape, may be null (forward only)
  const unsigned char* mask;      // group token mask [B, Lg]
  const float* coef;              // device scalar: d(loss)/d(err) weight = w_mod / (sum_w * count_mod) * upstream
  double* accum;                  // [2]: sum of errors over masked elements, number of masked elements
  int B, D, C, S, p, G;
  int Lg, mask_off;               // token index = mask_off + (gi * D + d) * L + l
  int cg[LOSS_MAX_G], c0[LOSS_MAX_G];
  int n_norm;                     // number of norm groups (0 = no target normalisation)
  int norm_of_ch[LOSS_MAX_C];     // channel -> norm group
  int norm_size[LOSS_MAX_C];      // channels per norm group
  int l2;                         // 0: |.|, 1: (.)^2
};

template <typename T>
__global__ void __launch_bounds__(128) group_norm_masked_loss_kernel(LossParams P) {
  extern __shared__ __align__(16) float smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int g = P.S / P.p, L = g * g, pp = P.p * P.p;
  const int npix = pp * P.C;
  float* tg = smem + static_cast<size_t>(w) * npix;
  const long long n_patch = static_cast<long long>(P.B) * P.D * L;
  float err_sum = 0.f, cnt_sum = 0.f;
  const float coef = (P.coef != nullptr) ? *P.coef : 0.f;

  for (long long patch = static_cast<long long>(blockIdx.x) * nw + w; patch < n_patch;
       patch += static_cast<long long>(gridDim.x) * nw) {
    const int l = static_cast<int>(patch % L);
    const long long img = patch / L;  // b * D + d
    const int d = static_cast<int>(img % P.D), b = static_cast<int>(img / P.D);
    unsigned mbits = 0;
    for (int gi = 0; gi < P.G; ++gi) {
      const size_t ti = static_cast<size_t>(b) * P.Lg + P.mask_off + (static_cast<size_t>(gi) * P.D + d) * L + l;
      if (P.mask[ti]) mbits |= 1u << gi;
    }
    if (mbits == 0) {
      // unmasked patch: gradient rows are zero, nothing else to do
      for (int gi = 0; gi < P.G; ++gi) {
        if (P.dpred[gi] == nullptr) continue;
        T* dp = static_cast<T*>(P.dpred[gi]) + static_cast<size_t>(patch) * P.cg[gi] * pp;
        for (int e = lane; e < P.cg[gi] * pp; e += 32) dp[e] = from_f32<T>(0.f);
      }
      continue;
    }
    // stage target patch: global order (ch, py, px) -> smem order (py, px, ch)
    const int gy = l / g, gx = l % g;
    const float* tbase = P.target + static_cast<size_t>(img) * P.C * P.S * P.S;
    for (int i = lane; i < npix; i += 32) {
      const int px = i % P.p, py = (i / P.p) % P.p, ch = i / pp;
      tg[(py * P.p + px) * P.C + ch] = tbase[(static_cast<size_t>(ch) * P.S + gy * P.p + py) * P.S + gx * P.p + px];
    }
    __syncwarp();
    // per norm group statistics
    float mu[LOSS_MAX_C], rs[LOSS_MAX_C];
    if (P.n_norm > 0) {
      int ch0 = 0;
      for (int n = 0; n < P.n_norm; ++n) {
        const int cn = P.norm_size[n];
        const int cnt = pp * cn;
        float s = 0.f;
        for (int i = lane; i < cnt; i += 32) s += tg[(i / cn) * P.C + ch0 + (i % cn)];
        const float mean = warp_sum(s) / cnt;
        float ss = 0.f;
        for (int i = lane; i < cnt; i += 32) {
          const float t = tg[(i / cn) * P.C + ch0 + (i % cn)] - mean;
          ss += t * t;
        }
        const float var = warp_sum(ss) / (cnt - 1);
        mu[n] = mean;
        rs[n] = sqrtf(var + 1.0e-6f);
        ch0 += cn;
      }
    }
    // stream prediction rows
    for (int gi = 0; gi < P.G; ++gi) {
      const int cgi = P.cg[gi], ne = cgi * pp;
      const bool masked = (mbits >> gi) & 1u;
      const T* pr = static_cast<const T*>(P.pred[gi]) + static_cast<size_t>(patch) * ne;
      T* dp = P.dpred[gi] != nullptr ? static_cast<T*>(P.dpred[gi]) + static_cast<size_t>(patch) * ne : nullptr;
      if (!masked) {
        if (dp != nullptr)
          for (int e = lane; e < ne; e += 32) dp[e] = from_f32<T>(0.f);
        continue;
      }
      for (int e = lane; e < ne; e += 32) {
        const int cl = e % cgi, pix = e / cgi;
        const int ch = P.c0[gi] + cl;
        float t = tg[pix * P.C + ch];
        if (P.n_norm > 0) {
          const int n = P.norm_of_ch[ch];
          float m_ = mu[0], r_ = rs[0];
#pragma unroll
          for (int k = 1; k < LOSS_MAX_C; ++k)
            if (k == n) { m_ = mu[k]; r_ = rs[k]; }
          t = (t - m_) / r_;
        }
        const float diff = t - to_f32(pr[e]);  // reference: loss_fn(target - pred)
        float gradv;
        if (P.l2) {
          err_sum += diff * diff;
          gradv = -2.f * diff;
        } else {
          err_sum += fabsf(diff);
          gradv = diff > 0.f ? -1.f : (diff < 0.f ? 1.f : 0.f);
        }
        if (dp != nullptr) dp[e] = from_f32<T>(gradv * coef);
      }
      cnt_sum += static_cast<float>(ne) / 32.f;
    }
    __syncwarp();
  }
  const double es = warp_sum(static_cast<double>(err_sum));
  const double cs = warp_sum(static_cast<double>(cnt_sum));
  if (lane == 0 && (es != 0.0 || cs != 0.0)) {
    atomicAdd(&P.accum[0], es);
    atomicAdd(&P.accum[1], cs);
  }
}

// Vector form of the kernel above for p % 4 == 0 (every modality with real patches; the scalar kernel keeps p = 1, 2).
// The target patch stays CHANNEL-PLANAR in shared memory, the order it has in HBM: float4 rows are fetched eight deep
// per lane (all in flight before the first use) and stored as float4 (conflict-free), the channels of one norm group
// are one contiguous span, so the centred second pass reads float4 as well; per-channel mean / 1 / std go through a
// small shared table and, when a 16-byte prediction chunk holds whole pixels, into registers once per patch.
// Predictions / gradients move as 16-byte chunks with (pixel, channel) tracked incrementally.
// PP / CC > 0: patch side and channel count fixed at compile time (one band group holding every channel), so that the
// index divisions become shifts and the chunk loops unroll -- the generic form is issue-bound on them (~2000 warp
// instructions per aerial patch).  PP = CC = 0: everything from LossParams.
template <typename T, int PP, int CC>
__global__ void __launch_bounds__(128, 5) group_norm_masked_loss_vec_kernel(LossParams P) {
  const int Pp = PP > 0 ? PP : P.p;
  const int PC = CC > 0 ? CC : P.C;
  constexpr int V = VecIO<T>::N;
  constexpr int DEPTH = 8;
  constexpr int PD = 4;     // prediction chunks per lane requested ahead (4 x 16 B x 32 lanes = one aerial patch row set)
  extern __shared__ __align__(16) float smem[];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  const int g = P.S / Pp, L = g * g, pp = Pp * Pp, p4 = Pp / 4, pp4 = pp / 4;
  const int npix = pp * PC;
  float* tg = smem + static_cast<size_t>(w) * (npix + 2 * LOSS_MAX_C);   // [C][p][p]
  float* mu_s = tg + npix;                 // per channel: mean, 1 / sqrt(var + eps) of its norm group
  float* ri_s = mu_s + LOSS_MAX_C;
  const long long n_patch = static_cast<long long>(P.B) * P.D * L;
  const size_t plane = static_cast<size_t>(P.S) * P.S;
  float err_sum = 0.f, cnt_sum = 0.f;
  const float coef = (P.coef != nullptr) ? *P.coef : 0.f;
  float zeros[V];
#pragma unroll
  for (int k = 0; k < V; ++k) zeros[k] = 0.f;

  for (long long patch = static_cast<long long>(blockIdx.x) * nw + w; patch < n_patch;
       patch += static_cast<long long>(gridDim.x) * nw) {
    const int l = static_cast<int>(patch % L);
    const long long img = patch / L;  // b * D + d
    const int d = static_cast<int>(img % P.D), b = static_cast<int>(img / P.D);
    unsigned mbits = 0;
    for (int gi = 0; gi < P.G; ++gi) {
      const size_t ti = static_cast<size_t>(b) * P.Lg + P.mask_off + (static_cast<size_t>(gi) * P.D + d) * L + l;
      if (P.mask[ti]) mbits |= 1u << gi;
    }
    // gradient rows of band groups whose token is not masked are zero
    for (int gi = 0; gi < P.G; ++gi) {
      if (((mbits >> gi) & 1u) || P.dpred[gi] == nullptr) continue;
      T* dp = static_cast<T*>(P.dpred[gi]) + static_cast<size_t>(patch) * P.cg[gi] * pp;
      for (int e = lane * V; e < P.cg[gi] * pp; e += 32 * V) VecIO<T>::store(dp + e, zeros);
    }
    if (mbits == 0) continue;
    // the first PD prediction chunks of band group 0 are requested now, next to the target rows, so that the patch has
    // one memory round trip instead of one per 512-byte chunk row (G = 1 in every shipped configuration)
    const bool use_pre = mbits & 1u;
    typename VecIO<T>::Raw pre[PD];
    if (use_pre) {
      const int ne0 = (CC > 0 ? CC : P.cg[0]) * pp;
      const T* pr0 = static_cast<const T*>(P.pred[0]) + static_cast<size_t>(patch) * ne0;
#pragma unroll
      for (int k = 0; k < PD; ++k) {
        const int e = (lane + 32 * k) * V;
        pre[k] = e < ne0 ? VecIO<T>::load_raw(pr0 + e) : VecIO<T>::zero_raw();
      }
    }
    // stage the target patch, channel-planar as in HBM; group sums on the way
    const int gy = l / g, gx = l % g;
    const float* tbase = P.target + static_cast<size_t>(img) * PC * plane + static_cast<size_t>(gy) * Pp * P.S +
                         gx * Pp;
    int ch0 = 0;
    const int ngroups = P.n_norm > 0 ? P.n_norm : 1;
    for (int n = 0; n < ngroups; ++n) {
      const int cn = P.n_norm > 0 ? P.norm_size[n] : PC;
      const int tot = cn * pp4;              // float4 elements of this norm group's channels
      float4* span = reinterpret_cast<float4*>(tg + ch0 * pp);
      float sacc = 0.f;
      for (int base = lane; base < tot; base += 32 * DEPTH) {
        float4 v[DEPTH];
#pragma unroll
        for (int k = 0; k < DEPTH; ++k) {
          const int idx = base + 32 * k;
          if (idx < tot) {
            const int ch = idx / pp4, i4 = idx - ch * pp4;
            const int py = i4 / p4, px = (i4 - py * p4) * 4;
            v[k] = __ldg(reinterpret_cast<const float4*>(tbase + (ch0 + ch) * plane + static_cast<size_t>(py) * P.S + px));
          }
        }
#pragma unroll
        for (int k = 0; k < DEPTH; ++k) {
          const int idx = base + 32 * k;
          if (idx < tot) {
            span[idx] = v[k];
            sacc += (v[k].x + v[k].y) + (v[k].z + v[k].w);
          }
        }
      }
      if (P.n_norm > 0) {
        const int cnt = pp * cn;
        const float mean = warp_sum(sacc) / cnt;
        __syncwarp();
        float ss = 0.f;
        for (int idx = lane; idx < tot; idx += 32) {
          const float4 t = span[idx];
          const float a0 = t.x - mean, a1 = t.y - mean, a2 = t.z - mean, a3 = t.w - mean;
          ss = fmaf(a0, a0, ss); ss = fmaf(a1, a1, ss); ss = fmaf(a2, a2, ss); ss = fmaf(a3, a3, ss);
        }
        const float var = warp_sum(ss) / (cnt - 1);
        if (lane >= ch0 && lane < ch0 + cn) {
          mu_s[lane] = mean;
          ri_s[lane] = 1.0f / sqrtf(var + 1.0e-6f);
        }
      } else if (lane < PC) {
        mu_s[lane] = 0.f;
        ri_s[lane] = 1.f;
      }
      ch0 += cn;
    }
    __syncwarp();
    // stream prediction rows
    for (int gi = 0; gi < P.G; ++gi) {
      if (!((mbits >> gi) & 1u)) continue;
      const int cgi = CC > 0 ? CC : P.cg[gi], ne = cgi * pp, c0 = CC > 0 ? 0 : P.c0[gi];
      const T* pr = static_cast<const T*>(P.pred[gi]) + static_cast<size_t>(patch) * ne;
      T* dp = P.dpred[gi] != nullptr ? static_cast<T*>(P.dpred[gi]) + static_cast<size_t>(patch) * ne : nullptr;
      const bool whole = (V % cgi) == 0;      // a chunk holds whole pixels: channel of element k is c0 + k % cgi
      float muk[V] = {}, rik[V] = {};
      if (whole) {
#pragma unroll
        for (int k = 0; k < V; ++k) {
          muk[k] = mu_s[c0 + k % cgi];
          rik[k] = ri_s[c0 + k % cgi];
        }
      }
      int it = 0;
      for (int e0 = lane * V; e0 < ne; e0 += 32 * V, ++it) {
        float pv[V], gv[V];
        typename VecIO<T>::Raw r;
        if (gi == 0 && it < PD) {          // use_pre holds: band group 0 is masked here
          r = pre[0];
#pragma unroll
          for (int k = 1; k < PD; ++k)
            if (it == k) r = pre[k];
        } else {
          r = VecIO<T>::load_raw(pr + e0);
        }
        VecIO<T>::unpack(r, pv);
        int pix = e0 / cgi, cl = e0 - pix * cgi;
#pragma unroll
        for (int k = 0; k < V; ++k) {
          const int ch = c0 + cl;
          const float tv = tg[ch * pp + pix];
          const float t = whole ? (tv - muk[k]) * rik[k] : (tv - mu_s[ch]) * ri_s[ch];
          const float diff = t - pv[k];  // reference: loss_fn(target - pred)
          if (P.l2) {
            err_sum = fmaf(diff, diff, err_sum);
            gv[k] = -2.f * diff * coef;
          } else {
            err_sum += fabsf(diff);
            gv[k] = diff > 0.f ? -coef : (diff < 0.f ? coef : 0.f);
          }
          if (++cl == cgi) { cl = 0; ++pix; }
        }
        if (dp != nullptr) VecIO<T>::store(dp + e0, gv);
      }
      cnt_sum += static_cast<float>(ne) / 32.f;
    }
    __syncwarp();
  }
  const double es = warp_sum(static_cast<double>(err_sum));
  const double cs = warp_sum(static_cast<double>(cnt_sum));
  if (lane == 0 && (es != 0.0 || cs != 0.0)) {
    atomicAdd(&P.accum[0], es);
    atomicAdd(&P.accum[1], cs);
  }
}

template <typename T>
static void launch_loss(const LossParams& P, int blocks, int nw, cudaStream_t st) {
  const int pp = P.p * P.p;
  bool vec = P.p % 4 == 0 && P.S % 4 == 0 && (reinterpret_cast<uintptr_t>(P.target) & 15) == 0;
  for (int gi = 0; gi < P.G && vec; ++gi) {
    vec = (P.cg[gi] * pp) % VecIO<T>::N == 0 && (reinterpret_cast<uintptr_t>(P.pred[gi]) & 15) == 0 &&
          (reinterpret_cast<uintptr_t>(P.dpred[gi]) & 15) == 0;
  }
  if (vec) {
    const size_t smem = static_cast<size_t>(nw) * (pp * P.C + 2 * LOSS_MAX_C) * sizeof(float);
    const bool one = P.G == 1 && P.cg[0] == P.C && P.c0[0] == 0;
    void (*kern)(LossParams) = group_norm_masked_loss_vec_kernel<T, 0, 0>;
    if (one && P.p == 16 && P.C == 4) kern = group_norm_masked_loss_vec_kernel<T, 16, 4>;        // aerial / NAIP RGBN
    else if (one && P.p == 32 && P.C == 2) kern = group_norm_masked_loss_vec_kernel<T, 32, 2>;   // DEM + DSM
    else if (one && P.p == 16 && P.C == 3) kern = group_norm_masked_loss_vec_kernel<T, 16, 3>;   // SPOT
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
    // persistent grid: exactly the CTAs that are resident at once (registers / shared memory decide), so no tail wave
    int per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, nw * 32, smem);
    if (per_sm > 0 && blocks > 148 * per_sm) blocks = 148 * per_sm;
    kern<<<blocks, nw * 32, smem, st>>>(P);
  } else {
    const size_t smem = static_cast<size_t>(nw) * pp * P.C * sizeof(float);
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(group_norm_masked_loss_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                           static_cast<int>(smem));
    group_norm_masked_loss_kernel<T><<<blocks, nw * 32, smem, st>>>(P);
  }
}

extern "C" int mb_loss_fwd_bwd(const LossParams* params, int dtype, void* stream) {
  LossParams P = *params;
  if (P.C > LOSS_MAX_C || P.G > LOSS_MAX_G || P.n_norm > LOSS_MAX_C) return -2;
  const int g = P.S / P.p;
  const long long n_patch = static_cast<long long>(P.B) * P.D * g * g;
  if (n_patch == 0) return 0;
  const int nw = 4;
  const size_t smem = static_cast<size_t>(nw) * (P.p * P.p * P.C + 2 * LOSS_MAX_C) * sizeof(float);
  if (smem > 200 * 1024) return -2;
  long long blocks = (n_patch + nw - 1) / nw;
  if (blocks > 148 * 8) blocks = 148 * 8;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == MB_DTYPE_F32) launch_loss<float>(P, static_cast<int>(blocks), nw, st);
  else if (dtype == MB_DTYPE_BF16) launch_loss<bf16>(P, static_cast<int>(blocks), nw, st);
  else return -1;
  MB_CHECK_LAUNCH();
  return 0;
}
