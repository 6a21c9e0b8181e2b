// tcgen05 / TMEM / TMA GEMM for sm_100a (bf16 inputs, fp32 accumulate in tensor memory).
//
//   C[m, n] = epilogue( sum_k A(m, k) * B(n, k) )
//
// Two operand layouts (template MN):
//   MN = false  A is [M, K] and B is [N, K], both K-contiguous       -> forward and dgrad (dgrad reads the
//                                                                        transposed bf16 weight shadow)
//   MN = true   A is [K, M] and B is [K, N], both MN-contiguous      -> wgrad  dW[n, k] = sum_t dY[t, n] X[t, k]
//                                                                        (contraction over tokens, split-K + fp32 red)
// Structure: persistent CTAs (one per SM), warp-specialised: warp 0 = TMA producer, warp 1 = single-thread
// tcgen05.mma issuer, warp 2 = TMEM allocator, warps 4..11 = epilogue (tcgen05.ld -> registers -> fused
// bias / GELU / GELU' / residual -> 16-byte global stores).  The accumulator is double-buffered in TMEM
// (2 x BN columns) so the epilogue of tile i overlaps the main loop of tile i + 1; shared memory holds a
// STAGES-deep ring of 128 x 64 (A) and BN x 64 (B) bf16 tiles in the 128-byte swizzle layout.
#include <mutex>
#include <unordered_map>
#include <string>
#include <cstring>

#include "common.cuh"
#include "gemm_params.h"
#include "tc_common.cuh"
#include "tmap.h"

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int A_BYTES = BM * BK * 2;
constexpr int NUM_THREADS = 384;
constexpr int EPI_WARP0 = 4;

template <int BN> struct Cfg {
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN == 256) ? 4 : 6;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/;
};

struct TcArgs {
  void* C;
  const float* bias;
  const void* residual;
  void* aux;
  int M, N, K;
  int ldc;
  int flags;
  int splits;
  float alpha;
};

template <int BN, bool MN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, TcArgs args) {
  using C_ = Cfg<BN>;
  constexpr int STAGES = C_::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * C_::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int mblocks = (args.M + BM - 1) / BM;
  const int nblocks = (args.N + BN - 1) / BN;
  const int kblocks = (args.K + BK - 1) / BK;
  const int tiles_mn = mblocks * nblocks;
  const int total_tiles = tiles_mn * args.splits;
  const int kb_per_split = (kblocks + args.splits - 1) / args.splits;

  if (warp == 0 && lane == 0) {
    tc::prefetch_tmap(&tmA);
    tc::prefetch_tmap(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&tfull_bar[s], 1);
      tc::mbar_init(&tempty_bar[s], 256);
    }
    tc::fence_barrier_init();
  }
  if (warp == 2) {
    tc::tmem_alloc(tmem_ptr, 2 * BN);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int split = t / tiles_mn, rem = t % tiles_mn;
        const int n_blk = rem / mblocks, m_blk = rem % mblocks;
        const int kb0 = split * kb_per_split;
        const int kb1 = min(kblocks, kb0 + kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sA = smem + stage * C_::STAGE_BYTES;
          uint8_t* sB = sA + A_BYTES;
          tc::mbar_expect_tx(&full_bar[stage], C_::STAGE_BYTES);
          if constexpr (!MN) {
            tc::tma_load_2d(sA, &tmA, &full_bar[stage], kb * BK, m_blk * BM);
            tc::tma_load_2d(sB, &tmB, &full_bar[stage], kb * BK, n_blk * BN);
          } else {
#pragma unroll
            for (int i = 0; i < BM / 64; ++i)
              tc::tma_load_2d(sA + i * 8192, &tmA, &full_bar[stage], m_blk * BM + i * 64, kb * BK);
#pragma unroll
            for (int i = 0; i < BN / 64; ++i)
              tc::tma_load_2d(sB + i * 8192, &tmB, &full_bar[stage], n_blk * BN + i * 64, kb * BK);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc_bf16(BM, BN, MN ? 1 : 0, MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int split = t / tiles_mn;
        const int kb0 = split * kb_per_split;
        const int kb1 = min(kblocks, kb0 + kb_per_split);
        tc::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + stage * C_::STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            uint64_t da, db;
            if constexpr (!MN) {
              // K-major, 128B swizzle: rows are 128 B, 8-row groups 1024 B apart; K advance = 32 B inside the row
              da = tc::make_smem_desc(a_addr + k * UMMA_K * 2, 16, 1024);
              db = tc::make_smem_desc(b_addr + k * UMMA_K * 2, 16, 1024);
            } else {
              // MN-major, 128B swizzle: 64-element MN chunks 8192 B apart (one TMA box each), 8-row K groups
              // 1024 B apart; K advance = 16 rows = 2048 B
              da = tc::make_smem_desc(a_addr + k * UMMA_K * 128, 8192, 1024);
              db = tc::make_smem_desc(b_addr + k * UMMA_K * 128, 8192, 1024);
            }
            tc::umma_bf16(d_tmem, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          tc::umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc::umma_commit(&tfull_bar[acc]);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= EPI_WARP0) {
    // ===================== epilogue (8 warps: 4 TMEM lane groups x 2 column halves) =====================
    const int ew = (warp - EPI_WARP0) & 3;     // == warp % 4 -> TMEM lanes [32 * ew, 32 * ew + 32)
    const int half = (warp - EPI_WARP0) >> 2;  // columns [half * BN / 2, (half + 1) * BN / 2)
    constexpr int CH_PER_HALF = BN / 64;       // 32-column chunks per half
    int acc = 0;
    uint32_t acc_phase = 0;
    const bool out_f32 = (args.flags & GEMM_OUT_F32) != 0;
    const bool f_gelu = (args.flags & GEMM_GELU) != 0, f_dgelu = (args.flags & GEMM_DGELU) != 0;
    const bool has_res = args.residual != nullptr;
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int rem = t % tiles_mn;
      const int n_blk = rem / mblocks, m_blk = rem % mblocks;
      tc::mbar_wait(&tfull_bar[acc], acc_phase);
      tc::tc_fence_after();
      const int row = m_blk * BM + ew * 32 + lane;
      const bool row_ok = row < args.M;
      const size_t row_off = static_cast<size_t>(row_ok ? row : 0) * args.ldc;
#pragma unroll 1
      for (int c2 = 0; c2 < CH_PER_HALF; ++c2) {
        const int ch = half * CH_PER_HALF + c2;
        const int col0 = n_blk * BN + ch * 32;
        if (col0 >= args.N) break;  // warp-uniform
        // issue the global reads of this chunk (GELU' pre-activation / residual) before waiting on TMEM
        uint4 pre[4];
        const bool need_in = f_dgelu || has_res;
        if (need_in) {
          const bf16* src = static_cast<const bf16*>(f_dgelu ? args.aux : args.residual) + row_off + col0;
#pragma unroll
          for (int v8 = 0; v8 < 4; ++v8)
            pre[v8] = (row_ok && col0 + v8 * 8 < args.N) ? __ldg(reinterpret_cast<const uint4*>(src + v8 * 8))
                                                         : make_uint4(0, 0, 0, 0);
        }
        uint32_t r[32];
        tc::tmem_ld_32x32(tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BN + ch * 32, r);
        tc::tmem_ld_wait();
        if (row_ok) {
#pragma unroll
          for (int v8 = 0; v8 < 4; ++v8) {
            const int col = col0 + v8 * 8;
            if (col >= args.N) break;
            float v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[v8 * 8 + j]) * args.alpha;
            if (args.bias != nullptr) {
              const float4 b0 = __ldg(reinterpret_cast<const float4*>(args.bias + col));
              const float4 b1 = __ldg(reinterpret_cast<const float4*>(args.bias + col + 4));
              v[0] += b0.x; v[1] += b0.y; v[2] += b0.z; v[3] += b0.w;
              v[4] += b1.x; v[5] += b1.y; v[6] += b1.z; v[7] += b1.w;
            }
            if (f_gelu) {
              bf16* aux = static_cast<bf16*>(args.aux) + row_off + col;
              VecIO<bf16>::store(aux, v);
#pragma unroll
              for (int j = 0; j < 8; ++j) v[j] = gelu_fast(__bfloat162float(__float2bfloat16_rn(v[j])));
            }
            if (need_in) {
              const uint32_t w[4] = {pre[v8].x, pre[v8].y, pre[v8].z, pre[v8].w};
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float a0 = __uint_as_float(w[j] << 16), a1 = __uint_as_float(w[j] & 0xffff0000u);
                if (f_dgelu) {
                  v[2 * j] *= gelu_fast_grad(a0);
                  v[2 * j + 1] *= gelu_fast_grad(a1);
                } else {
                  v[2 * j] += a0;
                  v[2 * j + 1] += a1;
                }
              }
            }
            if (out_f32) {
              float* c = static_cast<float*>(args.C) + row_off + col;
              if (args.flags & GEMM_ACCUM) {
                atomicAdd(reinterpret_cast<float4*>(c), make_float4(v[0], v[1], v[2], v[3]));
                atomicAdd(reinterpret_cast<float4*>(c + 4), make_float4(v[4], v[5], v[6], v[7]));
              } else {
                VecIO<float>::store(c, v);
                VecIO<float>::store(c + 4, v + 4);
              }
            } else {
              VecIO<bf16>::store(static_cast<bf16*>(args.C) + row_off + col, v);
            }
          }
        }
      }
      tc::tc_fence_before();
      tc::mbar_arrive(&tempty_bar[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, 2 * BN);
}

// ---------------------------------------------------------------- host
// 2D bf16 tensor [d1 rows, d0 cols] with row pitch stride1 (elements), box [b1 rows, b0 cols], 128B swizzle.
int get_tmap(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t stride1, uint32_t b0, uint32_t b1) {
  const uint64_t dims[2] = {d0, d1}, strides[1] = {stride1};
  const uint32_t box[2] = {b0, b1};
  return mb_get_tmap(out, ptr, 2, dims, strides, box, 128);
}

int num_sms() { return mb_num_sms(); }

template <int BN, bool MN>
int launch(const GemmParams& P, cudaStream_t st) {
  using C_ = Cfg<BN>;
  CUtensorMap tmA, tmB;
  int rc;
  if (!MN) {
    if ((rc = get_tmap(&tmA, P.A, P.K, P.M, P.sam, BK, BM)) != 0) return rc;
    if ((rc = get_tmap(&tmB, P.B, P.K, P.N, P.sbn, BK, BN)) != 0) return rc;
  } else {
    if ((rc = get_tmap(&tmA, P.A, P.M, P.K, P.sak, 64, BK)) != 0) return rc;
    if ((rc = get_tmap(&tmB, P.B, P.N, P.K, P.sbk, 64, BK)) != 0) return rc;
  }
  const int mblocks = (P.M + BM - 1) / BM, nblocks = (P.N + BN - 1) / BN, kblocks = (P.K + BK - 1) / BK;
  int splits = 1;
  if (MN && (P.flags & GEMM_ACCUM)) {
    // split the token contraction so that the grid covers the chip (fp32 red.add combines the partials)
    const int tiles = mblocks * nblocks;
    splits = (2 * num_sms() + tiles - 1) / tiles;
    if (splits > kblocks) splits = kblocks;
    if (splits < 1) splits = 1;
    const int per = (kblocks + splits - 1) / splits;
    splits = (kblocks + per - 1) / per;  // no empty split
  }
  TcArgs a;
  a.C = P.C; a.bias = P.bias; a.residual = P.residual; a.aux = P.aux;
  a.M = P.M; a.N = P.N; a.K = P.K; a.ldc = P.ldc; a.flags = P.flags; a.splits = splits; a.alpha = P.alpha;
  const int total = mblocks * nblocks * splits;
  const int grid = total < num_sms() ? total : num_sms();
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(gemm_tc_kernel<BN, MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, C_::SMEM_BYTES);
    attr_set = true;
  }
  gemm_tc_kernel<BN, MN><<<grid, NUM_THREADS, C_::SMEM_BYTES, st>>>(tmA, tmB, a);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : static_cast<int>(e);
}

}  // namespace

// bf16 only.  Supported operand layouts: (sak == 1 && sbk == 1) K-major, (sam == 1 && sbn == 1) MN-major.
// Unbatched (nb1 == nb2 == 1).  N % 8 == 0, ldc % 8 == 0.
extern "C" int mb_gemm_tc(const GemmParams* params, void* stream) {
  const GemmParams& P = *params;
  if (P.M == 0 || P.N == 0) return 0;
  if (P.nb1 != 1 || P.nb2 != 1) return -3;
  if (P.N % 8 != 0 || P.ldc % 8 != 0 || P.K <= 0) return -4;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool kmajor = P.sak == 1 && P.sbk == 1;
  const bool mnmajor = P.sam == 1 && P.sbn == 1;
  bool wide = P.N > 128;
  if (kmajor && wide) {
    // pick the tile width with the smaller quantised cost (waves x tile width; narrow tiles re-read A more)
    const long long mb = (P.M + BM - 1) / BM, sms = num_sms();
    const long long w256 = (mb * ((P.N + 255) / 256) + sms - 1) / sms * 256;
    const long long w128 = (mb * ((P.N + 127) / 128) + sms - 1) / sms * 141;  // 128 * 1.1
    wide = w256 <= w128;
  }
  if (kmajor) return wide ? launch<256, false>(P, st) : launch<128, false>(P, st);
  if (mnmajor) return wide ? launch<256, true>(P, st) : launch<128, true>(P, st);
  return -5;
}
