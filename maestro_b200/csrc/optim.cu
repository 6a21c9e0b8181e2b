// Fused AdamW over the flat fp32 parameter arena (reference: torch.optim.AdamW built at
// maestro/train/model.py:135-149).  One pass reads p, g, m, v and writes p, m, v, refreshes the bf16 weight
// shadow used by the tensor-core GEMMs and (optionally) zeroes the gradient for the next step's red.add
// accumulation -- 28 B/param of HBM traffic instead of the several foreach passes of the eager optimiser.
#include "common.cuh"

// hyper (device, fp32[8]): lr, beta1, beta2, eps, weight_decay, 1 - beta1^t, 1 - beta2^t, grad_scale
__global__ void __launch_bounds__(256) adamw_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m,
                                                    float* __restrict__ v, bf16* __restrict__ shadow,
                                                    const float* __restrict__ hyper, size_t n, int zero_grad) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
  const float bc1 = hyper[5], bc2 = hyper[6], gs = hyper[7];
  const float step = lr / bc1;
  const float inv_sqrt_bc2 = rsqrtf(bc2);
  const size_t nvec = n / 4;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec;
       i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    float pv[4], gv[4], mv[4], vv[4];
    VecIO<float>::load(p + i * 4, pv);
    VecIO<float>::load(g + i * 4, gv);
    VecIO<float>::load(m + i * 4, mv);
    VecIO<float>::load(v + i * 4, vv);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gr = gv[j] * gs;
      pv[j] *= (1.f - lr * wd);
      mv[j] = b1 * mv[j] + (1.f - b1) * gr;
      vv[j] = b2 * vv[j] + (1.f - b2) * gr * gr;
      const float denom = sqrtf(vv[j]) * inv_sqrt_bc2 + eps;
      pv[j] -= step * (mv[j] / denom);
      gv[j] = 0.f;
    }
    VecIO<float>::store(p + i * 4, pv);
    VecIO<float>::store(m + i * 4, mv);
    VecIO<float>::store(v + i * 4, vv);
    if (zero_grad) VecIO<float>::store(g + i * 4, gv);
    if (shadow != nullptr) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(pv[0], pv[1]), hi = __floats2bfloat162_rn(pv[2], pv[3]);
      uint2 pk = make_uint2(*reinterpret_cast<uint32_t*>(&lo), *reinterpret_cast<uint32_t*>(&hi));
      *reinterpret_cast<uint2*>(shadow + i * 4) = pk;
    }
  }
}

// n must be a multiple of 4 (the arena pads every parameter to 16 bytes).
extern "C" int mb_adamw_step(float* p, float* g, float* m, float* v, void* shadow, const float* hyper, size_t n,
                             int zero_grad, void* stream) {
  if (n == 0) return 0;
  if (n % 4 != 0) return -2;
  long long blocks = (static_cast<long long>(n / 4) + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  adamw_kernel<<<static_cast<int>(blocks), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      p, g, m, v, static_cast<bf16*>(shadow), hyper, n, zero_grad);
  MB_CHECK_LAUNCH();
  return 0;
}
