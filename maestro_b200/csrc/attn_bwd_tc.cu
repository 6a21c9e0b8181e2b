// Backward of the fused attention on tcgen05 (see attn_tc.cu for layouts).  Two kernels, no atomics:
//   dq kernel    one CTA per (sample, head, 128-query tile), loops over key tiles (two half tiles in flight):
//                S = Q K^T, dP = dO V^T (TMEM) -> dS = P * (dP - delta) * scale (bf16, smem) -> dQ += dS K (TMEM).
//                Also emits delta[row] = sum_d dO * O for the second kernel.
//   dkdv kernel  one CTA per (sample, head, 128-key tile), loops over query tiles (two half tiles in flight):
//                S^T = K Q^T, dP^T = V dO^T (TMEM) -> P^T, dS^T (bf16, smem) -> dV += P^T dO, dK += dS^T Q (TMEM).
// Thread t owns TMEM lane t (a query row in the first kernel, a key row in the second); P is recomputed from
// the saved log-sum-exp, so nothing of size N x N ever touches HBM.
#include "common.cuh"
#include "tc_common.cuh"
#include "tmap.h"

struct mb_attn_problem {     // mirror of include/maestro_b200.h
  const void* qkv;
  void* o;
  float* lse;
  const void* d_o;
  void* dqkv;
  float* delta;
  int Bn, N;
};

#ifndef MB_DIAG_TIMING
#define MB_DIAG_TIMING 0
#endif
#if MB_DIAG_TIMING   // diagnostic builds: CTA 0 of the dK / dV kernel stamps clock64() per pass (tools/_diag/attn_timeline.py)
__device__ unsigned long long g_attn_stamps[128 * 8];
#define MB_ASTAMP(pass_i, slot) do { if (blockIdx.x == 0 && (pass_i) < 128) g_attn_stamps[(pass_i) * 8 + (slot)] = clock64(); } while (0)
extern "C" int mb_diag_attn_timing(unsigned long long* host, int n) {
  return static_cast<int>(cudaMemcpyFromSymbol(host, g_attn_stamps, sizeof(unsigned long long) * (n < 1024 ? n : 1024)));
}
#else
#define MB_ASTAMP(pass_i, slot) do { } while (0)
#endif

namespace {

constexpr int BR = 128;  // rows owned by the CTA (TMEM lanes)
#ifndef MB_ATTN_BC_SMALL
#define MB_ATTN_BC_SMALL 32   // inner tile width for dh <= 32 (A/B builds: tools/_diag/build_diag.sh)
#endif

template <int DH> struct BCfg {
  // columns per inner tile: 32 for dh <= 32 so that every TMEM accumulator of a CTA fits 128 columns and 4 CTAs share
  // an SM (the exp-bound decoder case), 64 for dh = 64
  static constexpr int BC = DH <= 32 ? MB_ATTN_BC_SMALL : 64;
  static constexpr int SW = DH * 2;
  static constexpr int ROW_TILE = BR * DH * 2;   // [128, DH]
  static constexpr uint32_t LT = tc::layout_type_for(SW);
  // the inner tile is walked as two SUB-wide halves with double-buffered score accumulators and bf16 operand buffers
  // (K-major, swizzle span = row bytes), so that the MMAs of one half run under the exp / dS pass of the other
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

// Grouped launch (see attn_tc.cu): per problem the four tensor maps of the two kernels, its arguments and CTA range.
constexpr int ATTN_MAX_GROUP = 4;
struct BwdGroup {
  CUtensorMap q128[ATTN_MAX_GROUP], qsub[ATTN_MAX_GROUP], do128[ATTN_MAX_GROUP], dosub[ATTN_MAX_GROUP];
  BwdArgs a[ATTN_MAX_GROUP];
  int cta_end[ATTN_MAX_GROUP];
  int count;
};

// 16-byte chunk c16 of row `tid` of a K-major [128, BCc] tile whose swizzle span equals its row (128 B: chunk ^ row % 8;
// 64 B: chunk ^ (row / 2) % 4; 32 B: chunk ^ (row / 4) % 2)
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
// Same software pipeline as the dK / dV kernel (attn_bwd_dkdv.cuh): SUB-key half tiles, S / dP accumulators and the dS
// operand double-buffered, K / V halves in an NS-stage TMA ring, warps 0-3 = query rows, one lane of warp 4 issues
// every TMA load and MMA.
// NH warps share each TMEM lane quarter (warps w, w + 4, ...) and split the columns of every pass between them.
// NH = 1 (default): one warp per lane quarter.  NH = 2 (8 compute warps per CTA, 32 per SM) was measured SLOWER --
// 5.55 vs 5.10 ms per step over the step's shapes (profiles/r2b_attn_probe_nh2.txt): the fixed cost of a pass
// (barrier waits, fences, TMEM loads) is paid per thread, and halving the columns per thread doubles its share.
#ifndef MB_ATTN_BWD_NH
#define MB_ATTN_BWD_NH 1
#endif
constexpr int NH = MB_ATTN_BWD_NH;
constexpr int BWD_CTRL_WARP = 4 * NH;          // first of three control warps: score MMAs / gradient MMAs / TMA loads
constexpr int DQ_THREADS = 128 * NH + 96;   // + three control warps

template <int DH>
__global__ void __launch_bounds__(DQ_THREADS) attn_bwd_dq_kernel(const __grid_constant__ BwdGroup grp) {
  using C_ = BCfg<DH>;
  int pi = 0, cta0 = 0;
#pragma unroll
  for (int i = 0; i < ATTN_MAX_GROUP - 1; ++i)
    if (i + 1 < grp.count && static_cast<int>(blockIdx.x) >= grp.cta_end[i]) { pi = i + 1; cta0 = grp.cta_end[i]; }
  const CUtensorMap& tm_qkv128 = grp.q128[pi];
  const CUtensorMap& tm_qsub = grp.qsub[pi];
  const CUtensorMap& tm_do128 = grp.do128[pi];
  const BwdArgs& args = grp.a[pi];
  const int bid = static_cast<int>(blockIdx.x) - cta0;
  constexpr int SUB = C_::SUB, NS = C_::NS;
  constexpr int OFF_Q = 0, OFF_DO = C_::ROW_TILE, OFF_K = 2 * C_::ROW_TILE, OFF_V = OFF_K + NS * C_::SUB_TILE;
  constexpr int OFF_DS = OFF_V + NS * C_::SUB_TILE, OFF_BAR = OFF_DS + 2 * C_::AS_TILE;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bar_q = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* bar_kv = bar_q + 1;        // [NS]
  uint64_t* bar_s = bar_kv + NS;       // [2]
  uint64_t* bar_o = bar_s + 2;         // [2]
  uint64_t* bar_p = bar_o + 2;         // [2] dS half written (128 arrivals)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar_p + 2);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = args.N, H = args.H, inner = H * DH;
  const int r_tiles = (N + BR - 1) / BR;
  const int rt = bid % r_tiles, bh = bid / r_tiles;
  const int h = bh % H, b = bh / H;
  const int n_sub = (N + SUB - 1) / SUB;

  if (tid == 0) {
    tc::mbar_init(bar_q, 1);
    for (int i = 0; i < NS; ++i) tc::mbar_init(&bar_kv[i], 1);
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&bar_s[i], 1);
      tc::mbar_init(&bar_o[i], 1);
      tc::mbar_init(&bar_p[i], BR * NH);
    }
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
  pdl_launch();
  pdl_wait();
  // columns: S [2][SUB] | dP [2][SUB] | dQ [DH]
  const uint32_t tm_s = tmem_base, tm_dp = tmem_base + 2 * SUB, tm_dq = tmem_base + 4 * SUB;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const int qrow = tid & 127, half = tid >> 7;      // row of the tile / column half owned by this thread
  constexpr uint32_t idesc_s = tc::make_idesc_bf16(BR, SUB, 0, 0);
  constexpr uint32_t idesc_dq = tc::make_idesc_bf16(BR, DH, 0, 1);
  const uint32_t sQ = tc::smem_u32(smem + OFF_Q), sDO = tc::smem_u32(smem + OFF_DO), sDS = tc::smem_u32(smem + OFF_DS);
  const uint32_t sK0 = tc::smem_u32(smem + OFF_K), sV0 = tc::smem_u32(smem + OFF_V);

  auto load_kv = [&](int j, int stage) {
    tc::mbar_expect_tx(&bar_kv[stage], 2 * C_::SUB_TILE);
    tc::tma_load_3d(smem + OFF_K + stage * C_::SUB_TILE, &tm_qsub, &bar_kv[stage], inner + h * DH, j * SUB, b);
    tc::tma_load_3d(smem + OFF_V + stage * C_::SUB_TILE, &tm_qsub, &bar_kv[stage], 2 * inner + h * DH, j * SUB, b);
  };
  auto issue_s_dp = [&](int j, int stage) {
    const uint32_t sK = sK0 + stage * C_::SUB_TILE, sV = sV0 + stage * C_::SUB_TILE;
    const uint32_t buf = static_cast<uint32_t>(j & 1) * SUB;
#pragma unroll
    for (int k = 0; k < DH / 16; ++k)
      tc::umma_bf16(tm_s + buf, tc::make_smem_desc(sQ + k * 32, 16, 8 * C_::SW, C_::LT),
                    tc::make_smem_desc(sK + k * 32, 16, 8 * C_::SW, C_::LT), idesc_s, k > 0);
#pragma unroll
    for (int k = 0; k < DH / 16; ++k)
      tc::umma_bf16(tm_dp + buf, tc::make_smem_desc(sDO + k * 32, 16, 8 * C_::SW, C_::LT),
                    tc::make_smem_desc(sV + k * 32, 16, 8 * C_::SW, C_::LT), idesc_s, k > 0);
    tc::umma_commit(&bar_s[j & 1]);
  };
  auto issue_dq = [&](int j, int stage) {
    const uint32_t sK = sK0 + stage * C_::SUB_TILE;
    const uint32_t aoff = static_cast<uint32_t>(j & 1) * C_::AS_TILE;
#pragma unroll
    for (int k = 0; k < SUB / 16; ++k)
      tc::umma_bf16(tm_dq, tc::make_smem_desc(sDS + aoff + k * 32, 16, 8 * C_::AS_ROW, C_::AS_LT),
                    tc::make_smem_desc(sK + k * 16 * C_::SW, 16, 8 * C_::SW, C_::LT), idesc_dq, (j > 0 || k > 0));
    tc::umma_commit(&bar_o[j & 1]);
  };

  if (warp >= BWD_CTRL_WARP) {
    // ===================== three control lanes (see attn_bwd_dkdv.cuh): S / dP issuer, dQ issuer, loader =====================
    if (tid == 128 * NH + 64) {
      int l_stage = 0, l_next = 0;
      tc::mbar_expect_tx(bar_q, 2 * C_::ROW_TILE);
      tc::tma_load_3d(smem + OFF_Q, &tm_qkv128, bar_q, h * DH, rt * BR, b);
      tc::tma_load_3d(smem + OFF_DO, &tm_do128, bar_q, h * DH, rt * BR, b);
      for (; l_next < NS && l_next < n_sub; ++l_next) {      // fill the ring
        load_kv(l_next, l_stage);
        if (++l_stage == NS) l_stage = 0;
      }
      for (int j = 1; l_next < n_sub; ++j) {                 // the dQ MMA of half j - 1 frees its K / V stage
        tc::mbar_wait(&bar_o[(j - 1) & 1], ((j - 1) >> 1) & 1);
        load_kv(l_next, l_stage);
        ++l_next;
        if (++l_stage == NS) l_stage = 0;
      }
    } else if (tid == 128 * NH) {
      int s_stage = 0;
      uint32_t s_par = 0;
      tc::mbar_wait(bar_q, 0);
      for (int j = 0; j < n_sub; ++j) {
        if (j >= 2) tc::mbar_wait(&bar_p[j & 1], ((j - 2) >> 1) & 1);   // pass j - 2 has read buffer j & 1
        tc::mbar_wait(&bar_kv[s_stage], s_par);
        tc::tc_fence_after();
        issue_s_dp(j, s_stage);
        if (++s_stage == NS) { s_stage = 0; s_par ^= 1; }
      }
    } else if (tid == 128 * NH + 32) {
      int g_stage = 0;
      for (int j = 0; j < n_sub; ++j) {
        tc::mbar_wait(&bar_p[j & 1], (j >> 1) & 1);          // dS of half j is in shared memory
        tc::tc_fence_after();
        issue_dq(j, g_stage);
        if (++g_stage == NS) g_stage = 0;
      }
    }
  } else {
    // ===================== compute warps: thread = query row =====================
    // per-row constants: lse (log2 domain) and delta = sum_d dO * O (both threads of a row compute them)
    const int row = rt * BR + qrow;
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
      if (half == 0) args.delta[li] = delta;
    }
    const float c = args.scale_log2, sc = args.scale;
    const float delta_sc = delta * sc;
    constexpr int CW = SUB / NH;               // columns of a pass handled by this thread

    for (int j = 0; j < n_sub; ++j) {
      const int bufi = j & 1;
      const uint32_t ds_row = sDS + bufi * C_::AS_TILE + qrow * C_::AS_ROW;
      tc::mbar_wait(&bar_s[bufi], (j >> 1) & 1);
      if (j >= 2) tc::mbar_wait(&bar_o[bufi], ((j - 2) >> 1) & 1);   // the dQ MMA of half j - 2 has read `bufi`
      tc::tc_fence_after();
      // dS = P * (dP - delta) * scale.  No masks: keys past N are zero rows of K and V, so whatever lands in their
      // dS columns multiplies zero rows of K in dQ += dS K; query rows past N are never stored.
      uint32_t s[CW], dp[CW];
      tc::tmem_ld_32xN<CW>(tm_s + lane_base + bufi * SUB + half * CW, s);
      tc::tmem_ld_32xN<CW>(tm_dp + lane_base + bufi * SUB + half * CW, dp);
      tc::tmem_ld_wait();
#pragma unroll
      for (int v8 = 0; v8 < CW / 8; ++v8) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {   // packed fp32 (FFMA2 / FMUL2): 1.5 instead of 3 issue slots of FP32 per element
          const int idx = v8 * 8 + 2 * i;
          const float2 x = fma2(make_float2(__uint_as_float(s[idx]), __uint_as_float(s[idx + 1])), splat2(c), splat2(-lse2));
          const float2 g = fma2(make_float2(__uint_as_float(dp[idx]), __uint_as_float(dp[idx + 1])), splat2(sc), splat2(-delta_sc));
          const float2 d2 = mul2(make_float2(fast_exp2(x.x), fast_exp2(x.y)), g);
          __nv_bfloat162 hb = __floats2bfloat162_rn(d2.x, d2.y);
          w[i] = *reinterpret_cast<uint32_t*>(&hb);
        }
        store_a_chunk_s<SUB>(ds_row, qrow, half * (CW / 8) + v8, w);
      }
      tc::fence_proxy_async();
      tc::tc_fence_before();
      tc::mbar_arrive(&bar_p[bufi]);
    }
    tc::mbar_wait(&bar_o[(n_sub - 1) & 1], ((n_sub - 1) >> 1) & 1);   // a commit covers every MMA issued before it
    tc::tc_fence_after();
    bf16* out = args.dqkv + (static_cast<size_t>(b) * N + (row_ok ? row : 0)) * 3 * inner + h * DH;
    if constexpr (NH == 1) {
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
    } else {
      constexpr int OW = DH / NH;              // output columns written by this thread
      uint32_t r[OW];
      tc::tmem_ld_32xN<OW>(tm_dq + lane_base + half * OW, r);   // warp-collective: executed by every thread
      tc::tmem_ld_wait();
#pragma unroll
      for (int v8 = 0; v8 < OW / 8; ++v8) {
        if (row_ok) {
          float v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[v8 * 8 + i]);
          VecIO<bf16>::store(out + half * OW + v8 * 8, v);
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
