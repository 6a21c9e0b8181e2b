// dK / dV kernel of the fused attention backward and the C entry point (included by attn_bwd_tc.cu).
#pragma once

namespace {

template <int DH>
__global__ void __launch_bounds__(128) attn_bwd_dkdv_kernel(const __grid_constant__ CUtensorMap tm_qkv128,
                                                            const __grid_constant__ CUtensorMap tm_qkv64,
                                                            const __grid_constant__ CUtensorMap tm_do64, BwdArgs args) {
  using C_ = BCfg<DH>;
  constexpr int BC = C_::BC;
  constexpr int OFF_K = 0, OFF_V = C_::ROW_TILE, OFF_Q = 2 * C_::ROW_TILE, OFF_DO = OFF_Q + 2 * C_::COL_TILE;
  constexpr int OFF_PT = OFF_DO + 2 * C_::COL_TILE, OFF_DST = OFF_PT + C_::A_TILE, OFF_STAT = OFF_DST + C_::A_TILE;
  constexpr int OFF_BAR = OFF_STAT + 2 * 2 * BC * 4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  float* s_stat = reinterpret_cast<float*>(smem + OFF_STAT);   // [stage][lse2 | delta][64]
  uint64_t* bar_kv = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* bar_q = bar_kv + 1;  // [2]
  uint64_t* bar_s = bar_kv + 3;
  uint64_t* bar_o = bar_kv + 4;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar_kv + 5);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = args.N, H = args.H, inner = H * DH;
  const int r_tiles = (N + BR - 1) / BR;
  const int rt = blockIdx.x % r_tiles, bh = blockIdx.x / r_tiles;
  const int h = bh % H, b = bh / H;
  const int c_tiles = (N + BC - 1) / BC;

  if (tid == 0) {
    tc::mbar_init(bar_kv, 1);
    tc::mbar_init(&bar_q[0], 1);
    tc::mbar_init(&bar_q[1], 1);
    tc::mbar_init(bar_s, 1);
    tc::mbar_init(bar_o, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) {
    tc::tmem_alloc(tmem_ptr, C_::TMEM_DKDV);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tm_st = tmem_base, tm_dpt = tmem_base + BC, tm_dk = tmem_base + 2 * BC, tm_dv = tmem_base + 2 * BC + DH;
  const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
  constexpr uint32_t idesc_s = tc::make_idesc_bf16(BR, BC, 0, 0);    // [128 keys] x [64 queries], K = DH
  constexpr uint32_t idesc_g = tc::make_idesc_bf16(BR, DH, 0, 1);    // [128 keys] x [DH], K = 64 queries (B MN-major)
  const uint32_t sK = tc::smem_u32(smem + OFF_K), sV = tc::smem_u32(smem + OFF_V);
  const uint32_t sPT = tc::smem_u32(smem + OFF_PT), sDST = tc::smem_u32(smem + OFF_DST);

  auto load_q = [&](int i) {
    const int st = i & 1;
    tc::mbar_expect_tx(&bar_q[st], 2 * C_::COL_TILE);
    tc::tma_load_3d(smem + OFF_Q + st * C_::COL_TILE, &tm_qkv64, &bar_q[st], h * DH, i * BC, b);
    tc::tma_load_3d(smem + OFF_DO + st * C_::COL_TILE, &tm_do64, &bar_q[st], h * DH, i * BC, b);
  };
  auto issue_s = [&](int i) {
    const uint32_t sQ = tc::smem_u32(smem + OFF_Q + (i & 1) * C_::COL_TILE);
    const uint32_t sDO = tc::smem_u32(smem + OFF_DO + (i & 1) * C_::COL_TILE);
#pragma unroll
    for (int k = 0; k < DH / 16; ++k)
      tc::umma_bf16(tm_st, tc::make_smem_desc(sK + k * 32, 16, 8 * C_::SW, C_::LT),
                    tc::make_smem_desc(sQ + k * 32, 16, 8 * C_::SW, C_::LT), idesc_s, k > 0);
#pragma unroll
    for (int k = 0; k < DH / 16; ++k)
      tc::umma_bf16(tm_dpt, tc::make_smem_desc(sV + k * 32, 16, 8 * C_::SW, C_::LT),
                    tc::make_smem_desc(sDO + k * 32, 16, 8 * C_::SW, C_::LT), idesc_s, k > 0);
    tc::umma_commit(bar_s);
  };
  auto issue_grads = [&](int i) {
    const uint32_t sQ = tc::smem_u32(smem + OFF_Q + (i & 1) * C_::COL_TILE);
    const uint32_t sDO = tc::smem_u32(smem + OFF_DO + (i & 1) * C_::COL_TILE);
#pragma unroll
    for (int k = 0; k < BC / 16; ++k)
      tc::umma_bf16(tm_dv, tc::make_smem_desc(sPT + k * 32, 16, 8 * C_::A_ROW, C_::A_LT),
                    tc::make_smem_desc(sDO + k * 16 * C_::SW, 16, 8 * C_::SW, C_::LT), idesc_g, (i > 0 || k > 0));
#pragma unroll
    for (int k = 0; k < BC / 16; ++k)
      tc::umma_bf16(tm_dk, tc::make_smem_desc(sDST + k * 32, 16, 8 * C_::A_ROW, C_::A_LT),
                    tc::make_smem_desc(sQ + k * 16 * C_::SW, 16, 8 * C_::SW, C_::LT), idesc_g, (i > 0 || k > 0));
    tc::umma_commit(bar_o);
  };

  if (tid == 0) {
    tc::mbar_expect_tx(bar_kv, 2 * C_::ROW_TILE);
    tc::tma_load_3d(smem + OFF_K, &tm_qkv128, bar_kv, inner + h * DH, rt * BR, b);
    tc::tma_load_3d(smem + OFF_V, &tm_qkv128, bar_kv, 2 * inner + h * DH, rt * BR, b);
    load_q(0);
    if (c_tiles > 1) load_q(1);
    tc::mbar_wait(bar_kv, 0);
    tc::mbar_wait(&bar_q[0], 0);
    tc::tc_fence_after();
    issue_s(0);
  }

  const int row = rt * BR + tid;           // key row owned by this thread
  const bool row_ok = row < N;
  const float c = args.scale_log2, sc = args.scale;
  uint8_t* pt_row = smem + OFF_PT + tid * C_::A_ROW;
  uint8_t* dst_row = smem + OFF_DST + tid * C_::A_ROW;
  const size_t stat_base = (static_cast<size_t>(b) * H + h) * N;

  // per-query statistics (log2-domain lse, delta * scale) of tile i live in s_stat[i & 1]; the global loads of tile
  // i + 1 are issued at the top of iteration i and land in shared memory at its end, so no iteration waits on HBM
  float n_lse = 0.f, n_delta = 0.f;
  auto fetch_stats = [&](int i) {
    const int q = i * BC + tid;
    const bool ok = tid < BC && q < N;
    n_lse = ok ? args.lse[stat_base + q] * 1.4426950408889634f : 0.f;
    n_delta = ok ? args.delta[stat_base + q] * sc : 0.f;      // delta pre-multiplied by the score scale
  };
  auto put_stats = [&](int i) {
    if (tid < BC) {
      float* d = s_stat + (i & 1) * 2 * BC;
      d[tid] = n_lse;
      d[BC + tid] = n_delta;
    }
  };
  fetch_stats(0);
  put_stats(0);
  __syncthreads();

  for (int i = 0; i < c_tiles; ++i) {
    float* st = s_stat + (i & 1) * 2 * BC;
    if (i + 1 < c_tiles) fetch_stats(i + 1);
    tc::mbar_wait(bar_s, i & 1);
    tc::tc_fence_after();
    if (i > 0) {
      tc::mbar_wait(bar_o, (i - 1) & 1);   // gradient MMAs of tile i-1 done: P^T / dS^T buffers, Q/dO stage free
      if (tid == 0 && i + 1 < c_tiles) load_q(i + 1);
    }
    // P^T and dS^T.  No masks: queries past N are zero rows of Q and dO (TMA zero fill), so their columns multiply
    // zero rows in dV += P^T dO and dK += dS^T Q; key rows past N are never stored.  st[] holds finite values there.
#pragma unroll
    for (int ch = 0; ch < BC / 32; ++ch) {
      uint32_t s[32], dp[32];
      tc::tmem_ld_32x32(tm_st + lane_base + ch * 32, s);
      tc::tmem_ld_32x32(tm_dpt + lane_base + ch * 32, dp);
      tc::tmem_ld_wait();
#pragma unroll
      for (int v8 = 0; v8 < 4; ++v8) {
        uint32_t wp[4], wd[4];
        const int q0 = ch * 32 + v8 * 8;
        const float4 l0 = *reinterpret_cast<const float4*>(st + q0), l1 = *reinterpret_cast<const float4*>(st + q0 + 4);
        const float4 e0 = *reinterpret_cast<const float4*>(st + BC + q0);
        const float4 e1 = *reinterpret_cast<const float4*>(st + BC + q0 + 4);
        const float ls[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
        const float ds[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
#pragma unroll
        for (int k2 = 0; k2 < 4; ++k2) {
          float p2[2], d2[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int idx = v8 * 8 + 2 * k2 + e;
            const float p = fast_exp2(fmaf(__uint_as_float(s[idx]), c, -ls[2 * k2 + e]));
            p2[e] = p;
            d2[e] = p * fmaf(__uint_as_float(dp[idx]), sc, -ds[2 * k2 + e]);
          }
          __nv_bfloat162 hp = __floats2bfloat162_rn(p2[0], p2[1]);
          __nv_bfloat162 hd = __floats2bfloat162_rn(d2[0], d2[1]);
          wp[k2] = *reinterpret_cast<uint32_t*>(&hp);
          wd[k2] = *reinterpret_cast<uint32_t*>(&hd);
        }
        store_a_chunk<BC>(pt_row, tid, ch * 4 + v8, wp);
        store_a_chunk<BC>(dst_row, tid, ch * 4 + v8, wd);
      }
    }
    if (i + 1 < c_tiles) put_stats(i + 1);
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc::tc_fence_after();
      issue_grads(i);
      if (i + 1 < c_tiles) {
        tc::mbar_wait(&bar_q[(i + 1) & 1], ((i + 1) >> 1) & 1);
        tc::tc_fence_after();
        issue_s(i + 1);
      }
    }
  }
  tc::mbar_wait(bar_o, (c_tiles - 1) & 1);
  tc::tc_fence_after();
  bf16* out = args.dqkv + (static_cast<size_t>(b) * N + (row_ok ? row : 0)) * 3 * inner + h * DH;
#pragma unroll
  for (int which = 0; which < 2; ++which) {   // 0: dK, 1: dV
#pragma unroll
    for (int ch = 0; ch < (DH + 31) / 32; ++ch) {
      uint32_t r[32];
      tc::tmem_ld_32x32((which == 0 ? tm_dk : tm_dv) + lane_base + ch * 32, r);
      tc::tmem_ld_wait();
#pragma unroll
      for (int v8 = 0; v8 < 4; ++v8) {
        if (ch * 32 + v8 * 8 < DH && row_ok) {
          float v[8];
#pragma unroll
          for (int k2 = 0; k2 < 8; ++k2) v[k2] = __uint_as_float(r[v8 * 8 + k2]);
          VecIO<bf16>::store(out + (which + 1) * inner + ch * 32 + v8 * 8, v);
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, C_::TMEM_DKDV);
}

int tmap3(CUtensorMap* tm, const void* base, int Bn, int N, int cols, int DH, int box_rows) {
  const uint64_t dims[3] = {static_cast<uint64_t>(cols), static_cast<uint64_t>(N), static_cast<uint64_t>(Bn)};
  const uint64_t strides[2] = {static_cast<uint64_t>(cols), static_cast<uint64_t>(N) * cols};
  const uint32_t box[3] = {static_cast<uint32_t>(DH), static_cast<uint32_t>(box_rows), 1};
  return mb_get_tmap(tm, base, 3, dims, strides, box, DH * 2);
}

template <int DH>
int launch_bwd(const void* qkv, const void* o, const void* d_o, const float* lse, void* dqkv, float* delta, int Bn,
               int N, int H, cudaStream_t st) {
  using C_ = BCfg<DH>;
  CUtensorMap q128, q64, do128, do64;
  int rc;
  if ((rc = tmap3(&q128, qkv, Bn, N, 3 * H * DH, DH, BR)) != 0) return rc;
  if ((rc = tmap3(&q64, qkv, Bn, N, 3 * H * DH, DH, C_::BC)) != 0) return rc;
  if ((rc = tmap3(&do128, d_o, Bn, N, H * DH, DH, BR)) != 0) return rc;
  if ((rc = tmap3(&do64, d_o, Bn, N, H * DH, DH, C_::BC)) != 0) return rc;
  BwdArgs a;
  a.o = static_cast<const bf16*>(o);
  a.d_o = static_cast<const bf16*>(d_o);
  a.lse = lse;
  a.delta = delta;
  a.dqkv = static_cast<bf16*>(dqkv);
  a.N = N;
  a.H = H;
  a.scale = 1.0f / sqrtf(static_cast<float>(DH));
  a.scale_log2 = a.scale * 1.4426950408889634f;
  constexpr int SMEM_DQ = 2 * C_::ROW_TILE + 4 * C_::COL_TILE + C_::A_TILE + 128 + 1024;
  constexpr int SMEM_DKDV = 2 * C_::ROW_TILE + 4 * C_::COL_TILE + 2 * C_::A_TILE + 2 * 2 * C_::BC * 4 + 128 + 1024;
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(attn_bwd_dq_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DQ);
    cudaFuncSetAttribute(attn_bwd_dkdv_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DKDV);
    attr = true;
  }
  const int grid = Bn * H * ((N + BR - 1) / BR);
  attn_bwd_dq_kernel<DH><<<grid, 128, SMEM_DQ, st>>>(q128, q64, do128, a);
  attn_bwd_dkdv_kernel<DH><<<grid, 128, SMEM_DKDV, st>>>(q128, q64, do64, a);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : static_cast<int>(e);
}

}  // namespace

// Gradient of mb_attn_fwd: dqkv [Bn*N, 3*H*dh] bf16 from d_o [Bn*N, H*dh]; delta [Bn, H, N] fp32 is scratch.
extern "C" int mb_attn_bwd(const void* qkv, const void* o, const void* d_o, const float* lse, void* dqkv, float* delta,
                           int Bn, int N, int H, int dh, void* stream) {
  if (Bn == 0 || N == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dh) {
    case 16: return launch_bwd<16>(qkv, o, d_o, lse, dqkv, delta, Bn, N, H, st);
    case 32: return launch_bwd<32>(qkv, o, d_o, lse, dqkv, delta, Bn, N, H, st);
    case 64: return launch_bwd<64>(qkv, o, d_o, lse, dqkv, delta, Bn, N, H, st);
    default: return -2;
  }
}
