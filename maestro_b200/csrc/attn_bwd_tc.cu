// Backward of the fused attention on tcgen05 (see attn_tc.cu for layouts).  Two kernels, no atomics:
//   dq kernel    one CTA per (sample, head, 128-query tile), loops over 64-key tiles:
//                S = Q K^T, dP = dO V^T (TMEM) -> dS = P * (dP - delta) * scale (bf16, smem) -> dQ += dS K (TMEM).
//                Also emits delta[row] = sum_d dO * O for the second kernel.
//   dkdv kernel  one CTA per (sample, head, 128-key tile), loops over 64-query tiles:
//                S^T = K Q^T, dP^T = V dO^T (TMEM) -> P^T, dS^T (bf16, smem) -> dV += P^T dO, dK += dS^T Q (TMEM).
// Thread t owns TMEM lane t (a query row in the first kernel, a key row in the second); P is recomputed from
// the saved log-sum-exp, so nothing of size N x N ever touches HBM.
#include "common.cuh"
#include "tc_common.cuh"
#include "tmap.h"

namespace {

constexpr int BR = 128;  // rows owned by the CTA (TMEM lanes)

template <int DH> struct BCfg {
  // columns per inner tile: 32 for dh <= 32 so that every TMEM accumulator of a CTA fits 128 columns and 4 CTAs share
  // an SM (the exp-bound decoder case), 64 for dh = 64
  static constexpr int BC = DH <= 32 ? 32 : 64;
  static constexpr int SW = DH * 2;
  static constexpr int ROW_TILE = BR * DH * 2;   // [128, DH]
  static constexpr int COL_TILE = BC * DH * 2;   // [BC, DH]
  static constexpr int A_ROW = BC * 2;           // row bytes of the [128, BC] bf16 tiles written by the threads
  static constexpr int A_TILE = BR * A_ROW;      // K-major, swizzle span = row bytes (128 B or 64 B)
  static constexpr uint32_t LT = tc::layout_type_for(SW);
  static constexpr uint32_t A_LT = tc::layout_type_for(A_ROW);
  // dK / dV kernel: the inner tile is walked as two SUB-query halves with double-buffered S^T / dP^T accumulators and
  // P^T / dS^T operand buffers, so that the MMAs of one half run under the exp / dS pass of the other
  static constexpr int SUB = BC / 2;
  static constexpr int SUB_TILE = SUB * DH * 2;  // [SUB, DH] Q / dO stage
  static constexpr int NS = DH >= 64 ? 5 : 6;    // Q / dO stages in flight (TMA latency spans ~3 half-tile passes)
  static constexpr int AS_ROW = SUB * 2;         // row bytes of the [128, SUB] bf16 operand halves
  static constexpr int AS_TILE = BR * AS_ROW;
  static constexpr uint32_t AS_LT = tc::layout_type_for(AS_ROW);
  static constexpr int TMEM_DQ = (2 * BC + DH <= 128) ? 128 : 256;
  static constexpr int TMEM_DKDV = (2 * BC + 2 * DH <= 128) ? 128 : 256;
};

struct BwdArgs {
  const bf16* o;
  const bf16* d_o;
  const float* lse;
  float* delta;
  bf16* dqkv;
  int N, H;
  float scale, scale_log2;
};

// 16-byte chunk c16 of row `tid` of a K-major [128, BCc] tile whose swizzle span equals its row (128 B: chunk ^ row % 8;
// 64 B: chunk ^ (row / 2) % 4)
template <int BCc>
__device__ __forceinline__ void store_a_chunk(uint8_t* row_base, int tid, int c16, const uint32_t* w) {
  // swizzle span = row bytes: 128 B: chunk ^ row % 8; 64 B: chunk ^ (row / 2) % 4; 32 B: chunk ^ (row / 4) % 2
  const int x = BCc == 64 ? (tid & 7) : (BCc == 32 ? ((tid >> 1) & 3) : ((tid >> 2) & 1));
  *reinterpret_cast<uint4*>(row_base + ((c16 ^ x) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
}
// the same with a shared-window address (st.shared instead of a generic store)
template <int BCc>
__device__ __forceinline__ void store_a_chunk_s(uint32_t row_base, int tid, int c16, const uint32_t* w) {
  const int x = BCc == 64 ? (tid & 7) : (BCc == 32 ? ((tid >> 1) & 3) : ((tid >> 2) & 1));
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_base + ((c16 ^ x) << 4)), "r"(w[0]), "r"(w[1]),
               "r"(w[2]), "r"(w[3]) : "memory");
}
__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}

// ------------------------------------------------------------------------------------------------ dQ
template <int DH>
__global__ void __launch_bounds__(128) attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tm_qkv128,
                                                          const __grid_constant__ CUtensorMap tm_qkv64,
                                                          const __grid_constant__ CUtensorMap tm_do128, BwdArgs args) {
  using C_ = BCfg<DH>;
  constexpr int BC = C_::BC;
  constexpr int OFF_Q = 0, OFF_DO = C_::ROW_TILE, OFF_K = 2 * C_::ROW_TILE, OFF_V = OFF_K + 2 * C_::COL_TILE;
  constexpr int OFF_DS = OFF_V + 2 * C_::COL_TILE, OFF_BAR = OFF_DS + C_::A_TILE;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bar_q = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* bar_kv = bar_q + 1;  // [2]
  uint64_t* bar_s = bar_q + 3;
  uint64_t* bar_o = bar_q + 4;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar_q + 5);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = args.N, H = args.H, inner = H * DH;
  const int r_tiles = (N + BR - 1) / BR;
  const int rt = blockIdx.x % r_tiles, bh = blockIdx.x / r_tiles;
  const int h = bh % H, b = bh / H;
  const int c_tiles = (N + BC - 1) / BC;

  if (tid == 0) {
    tc::mbar_init(bar_q, 1);
    tc::mbar_init(&bar_kv[0], 1);
    tc::mbar_init(&bar_kv[1], 1);
    tc::mbar_init(bar_s, 1);
    tc::mbar_init(bar_o, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) {
    tc::tmem_alloc(tmem_ptr, C_::TMEM_DQ);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tm_s = tmem_base, tm_dp = tmem_base + BC, tm_dq = tmem_base + 2 * BC;
  const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
  constexpr uint32_t idesc_s = tc::make_idesc_bf16(BR, BC, 0, 0);
  constexpr uint32_t idesc_dq = tc::make_idesc_bf16(BR, DH, 0, 1);
  const uint32_t sQ = tc::smem_u32(smem + OFF_Q), sDO = tc::smem_u32(smem + OFF_DO), sDS = tc::smem_u32(smem + OFF_DS);

  auto load_kv = [&](int j) {
    const int st = j & 1;
    tc::mbar_expect_tx(&bar_kv[st], 2 * C_::COL_TILE);
    tc::tma_load_3d(smem + OFF_K + st * C_::COL_TILE, &tm_qkv64, &bar_kv[st], inner + h * DH, j * BC, b);
    tc::tma_load_3d(smem + OFF_V + st * C_::COL_TILE, &tm_qkv64, &bar_kv[st], 2 * inner + h * DH, j * BC, b);
  };
  auto issue_s_dp = [&](int j) {
    const uint32_t sK = tc::smem_u32(smem + OFF_K + (j & 1) * C_::COL_TILE);
    const uint32_t sV = tc::smem_u32(smem + OFF_V + (j & 1) * C_::COL_TILE);
#pragma unroll
    for (int k = 0; k < DH / 16; ++k)
      tc::umma_bf16(tm_s, tc::make_smem_desc(sQ + k * 32, 16, 8 * C_::SW, C_::LT),
                    tc::make_smem_desc(sK + k * 32, 16, 8 * C_::SW, C_::LT), idesc_s, k > 0);
#pragma unroll
    for (int k = 0; k < DH / 16; ++k)
      tc::umma_bf16(tm_dp, tc::make_smem_desc(sDO + k * 32, 16, 8 * C_::SW, C_::LT),
                    tc::make_smem_desc(sV + k * 32, 16, 8 * C_::SW, C_::LT), idesc_s, k > 0);
    tc::umma_commit(bar_s);
  };
  auto issue_dq = [&](int j) {
    const uint32_t sK = tc::smem_u32(smem + OFF_K + (j & 1) * C_::COL_TILE);
#pragma unroll
    for (int k = 0; k < BC / 16; ++k)
      tc::umma_bf16(tm_dq, tc::make_smem_desc(sDS + k * 32, 16, 8 * C_::A_ROW, C_::A_LT),
                    tc::make_smem_desc(sK + k * 16 * C_::SW, 16, 8 * C_::SW, C_::LT), idesc_dq, (j > 0 || k > 0));
    tc::umma_commit(bar_o);
  };

  if (tid == 0) {
    tc::mbar_expect_tx(bar_q, 2 * C_::ROW_TILE);
    tc::tma_load_3d(smem + OFF_Q, &tm_qkv128, bar_q, h * DH, rt * BR, b);
    tc::tma_load_3d(smem + OFF_DO, &tm_do128, bar_q, h * DH, rt * BR, b);
    load_kv(0);
    if (c_tiles > 1) load_kv(1);
    tc::mbar_wait(bar_q, 0);
    tc::mbar_wait(&bar_kv[0], 0);
    tc::tc_fence_after();
    issue_s_dp(0);
  }

  // per-row constants: lse (log2 domain) and delta = sum_d dO * O
  const int row = rt * BR + tid;
  const bool row_ok = row < N;
  float lse2 = 0.f, delta = 0.f;
  if (row_ok) {
    const size_t ro = (static_cast<size_t>(b) * N + row) * inner + h * DH;
#pragma unroll
    for (int v8 = 0; v8 < DH / 8; ++v8) {
      float a[8], g[8];
      VecIO<bf16>::load(args.o + ro + v8 * 8, a);
      VecIO<bf16>::load(args.d_o + ro + v8 * 8, g);
#pragma unroll
      for (int i = 0; i < 8; ++i) delta = fmaf(a[i], g[i], delta);
    }
    const size_t li = (static_cast<size_t>(b) * H + h) * N + row;
    lse2 = args.lse[li] * 1.4426950408889634f;
    args.delta[li] = delta;
  }
  const float c = args.scale_log2, sc = args.scale;
  const float delta_sc = delta * sc;
  uint8_t* ds_row = smem + OFF_DS + tid * C_::A_ROW;

  for (int j = 0; j < c_tiles; ++j) {
    tc::mbar_wait(bar_s, j & 1);
    tc::tc_fence_after();
    if (j > 0) {
      tc::mbar_wait(bar_o, (j - 1) & 1);           // dQ MMA of tile j-1 done: dS buffer and K/V stage free
      if (tid == 0 && j + 1 < c_tiles) load_kv(j + 1);
    }
    // dS = P * (dP - delta) * scale.  No masks: keys past N are zero rows of K and V, so whatever lands in their dS
    // columns multiplies zero rows of K in dQ += dS K; query rows past N are never stored.
#pragma unroll
    for (int ch = 0; ch < BC / 32; ++ch) {
      uint32_t s[32], dp[32];
      tc::tmem_ld_32x32(tm_s + lane_base + ch * 32, s);
      tc::tmem_ld_32x32(tm_dp + lane_base + ch * 32, dp);
      tc::tmem_ld_wait();
#pragma unroll
      for (int v8 = 0; v8 < 4; ++v8) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          float d2[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int idx = v8 * 8 + 2 * i + e;
            const float p = fast_exp2(fmaf(__uint_as_float(s[idx]), c, -lse2));
            d2[e] = p * fmaf(__uint_as_float(dp[idx]), sc, -delta_sc);
          }
          __nv_bfloat162 hb = __floats2bfloat162_rn(d2[0], d2[1]);
          w[i] = *reinterpret_cast<uint32_t*>(&hb);
        }
        store_a_chunk<BC>(ds_row, tid, ch * 4 + v8, w);
      }
    }
    tc::fence_proxy_async();
    tc::tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc::tc_fence_after();
      issue_dq(j);
      if (j + 1 < c_tiles) {
        tc::mbar_wait(&bar_kv[(j + 1) & 1], ((j + 1) >> 1) & 1);
        tc::tc_fence_after();
        issue_s_dp(j + 1);
      }
    }
  }
  tc::mbar_wait(bar_o, (c_tiles - 1) & 1);
  tc::tc_fence_after();
  {
    bf16* out = args.dqkv + (static_cast<size_t>(b) * N + (row_ok ? row : 0)) * 3 * inner + h * DH;
#pragma unroll
    for (int ch = 0; ch < (DH + 31) / 32; ++ch) {
      uint32_t r[32];
      tc::tmem_ld_32x32(tm_dq + lane_base + ch * 32, r);   // warp-collective: executed by every thread
      tc::tmem_ld_wait();
#pragma unroll
      for (int v8 = 0; v8 < 4; ++v8) {
        if (ch * 32 + v8 * 8 < DH && row_ok) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[v8 * 8 + i]);
          VecIO<bf16>::store(out + ch * 32 + v8 * 8, v);
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, C_::TMEM_DQ);
}

}  // namespace

#include "attn_bwd_dkdv.cuh"
