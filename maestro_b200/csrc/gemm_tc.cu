// tcgen05 / TMEM / TMA GEMM for sm_100a (bf16 inputs, fp32 accumulate in tensor memory), grouped.
//
//   C_p[m, n] = epilogue( sum_k A_p(m, k) * B_p(n, k) )        p = 0 .. count-1 (independent problems, one launch)
//
// Operand layouts (template A_MN / B_MN, any combination the callers need):
//   A K-major  [M, K] (k contiguous), B K-major  [N, K]   -> forward   y = x W^T
//   A K-major  [M, K],                B MN-major [K, N]   -> dgrad     dx = dy W     (W read in place, no transpose)
//   A MN-major [K, M],                B MN-major [K, N]   -> wgrad     dW[n, k] = sum_t dY[t, n] X[t, k]
//                                                            (contraction over tokens, split-K + fp32 red.add)
// Structure: persistent CTAs (one per SM), warp-specialised: warp 0 = TMA producer, warp 1 = single-thread
// tcgen05.mma issuer, warp 2 = TMEM allocator, warps 4..19 = epilogue.  The accumulator is double-buffered in
// TMEM (2 x BN columns) so the epilogue of tile i overlaps the main loop of tile i + 1; shared memory holds a
// STAGES-deep ring of 128 x 64 (A) and BN x 64 (B) bf16 tiles in the 128-byte swizzle layout, plus one 2 KB
// staging buffer per epilogue warp.
// Epilogue (bf16 outputs): tcgen05.ld -> registers -> fused bias / GELU (+ saved gelu') / GELU' /
// residual -> bf16 -> swizzled shared-memory staging -> TMA tile store (full 128-byte lines, clipped at the
// matrix edge by the hardware).  Epilogue inputs (residual, saved gelu') are read with fully coalesced
// 16-byte loads (8 rows x 64 B per instruction, the tile pulled into L2 by a TMA prefetch while the main loop runs)
// and transposed to "thread = row" through the same staging buffer.
// fp32 outputs (wgrad accumulation) use direct 16-byte stores / red.add.
#include <algorithm>
#include <mutex>
#include <unordered_map>
#include <string>
#include <cstdlib>
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
constexpr int NUM_THREADS = 640;
constexpr int EPI_WARP0 = 4;
constexpr int EPI_WARPS = 16;
constexpr int STG_BYTES = 32 * 32 * 2;   // one epilogue warp: 32 rows x 32 bf16 columns
constexpr int MAX_GROUP = 4;

// Diagnostic builds only (tools/_diag/build_diag.sh; results are WRONG, timing only): 1 = the epilogue only reads the
// accumulator and releases it; 2 = everything but the TMA stores; 3 = direct 16-byte global stores instead of staging + TMA.
#ifndef MB_DIAG_EPI
#define MB_DIAG_EPI 0
#endif
// bf16 output path of the epilogue: 1 (default) = one 2 KB TMA tile store per 32 x 32 chunk; 0 = staged chunk ->
// coalesced 16-byte global stores (8 rows x 64 B per warp instruction).  Measured (profiles/r2b_gemm_epilogue.md): equal
// on plain / residual epilogues, TMA 5-15 % ahead where the epilogue is issue-bound (GELU, GELU').  fp32 outputs
// (weight gradients) always go through the staging buffer to coalesced red.add / stores.
#ifndef MB_EPI_TMA_STORE
#define MB_EPI_TMA_STORE 1
#endif
#ifndef MB_EPI_STG_BUFS
#define MB_EPI_STG_BUFS 1   // 2: see Cfg::STG_BUFS (measured: +3 % on the K = 512 GELU / GELU' GEMMs, -2..4 % everywhere else)
#endif

// PAIR: two CTAs of a cluster (one TPC) run ONE 256 x BN tcgen05.mma.cta_group::2 tile; each CTA stages its own 128
// rows of A and HALF of the B tile (BN / 2 rows), so a CTA pulls 32 KB instead of 48 KB per k-block into its SM.
template <int BN, bool PAIR, int EF = -1> struct Cfg {
  static constexpr int B_ROWS = PAIR ? BN / 2 : BN;
  static constexpr int B_BYTES = B_ROWS * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // MB_EPI_STG_BUFS=2 (A/B builds): pair tiles with a compile-time bf16 epilogue alternate between TWO staging buffers
  // per epilogue warp, so a warp only waits for the TMA store before the last one (cp.async.bulk.wait_group.read 1) --
  // paid with one operand stage of the pair's ring (6 -> 5 / 8 -> 6).  Measured (profiles/r2b_summary.md): the operand
  // stage is worth more on every shape except the K = 512 GELU / GELU' GEMMs; off by default.
  static constexpr int STG_BUFS = (PAIR && EF >= 0 && MB_EPI_STG_BUFS == 2) ? 2 : 1;
  static constexpr int STAGES = PAIR ? ((BN == 256) ? (STG_BUFS == 2 ? 5 : 6) : (STG_BUFS == 2 ? 6 : 8)) : ((BN == 256) ? 4 : 6);
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_WARPS * STG_BYTES * STG_BUFS + 1024 /*align slack*/ + 256 /*barriers*/;
};

// MB_DIAG_TIMING (diagnostic builds only): CTA 0 stamps clock64() per tile: [tile][0] MMA warp got the accumulator,
// [1] MMA warp issued the tile's last commit, [2] epilogue warp 4 starts waiting for the accumulator, [3] has it,
// [4] has issued its last store of the tile.  Read back with mb_diag_gemm_timing().
#ifndef MB_DIAG_TIMING
#define MB_DIAG_TIMING 0
#endif
#if MB_DIAG_TIMING
__device__ unsigned long long g_gemm_stamps[64 * 8];
#define MB_STAMP(tile_i, slot) do { if (blockIdx.x == 0 && (tile_i) < 64) g_gemm_stamps[(tile_i) * 8 + (slot)] = clock64(); } while (0)
#else
#define MB_STAMP(tile_i, slot) do { } while (0)
#endif

struct TcProblem {
  void* C;
  const float* bias;
  const void* residual;
  void* aux;
  float* colsum;
  int M, N, K, ldc;
  int splits;
  int mblocks, nblocks;
  int tile_end;          // exclusive prefix sum of mblocks * nblocks * splits over the group
};

struct TcArgs {
  TcProblem p[MAX_GROUP];
  int count;
  int flags;
  float alpha;
  int total_tiles;
  int half_start;        // tiles t >= half_start are HALF-width tiles (two per full tile of the last partial wave)
};

struct TcMaps {
  CUtensorMap a[MAX_GROUP], b[MAX_GROUP], c[MAX_GROUP];
  CUtensorMap x[MAX_GROUP];   // GELU: gelu'(pre-activation) output; GELU': the same tensor as input; else residual input
  CUtensorMap pf[MAX_GROUP];  // the epilogue input again, with a 128-row x 64-column box for L2 prefetch
  CUtensorMap bh[MAX_GROUP];  // K-major B with a half-height box: the B rows of a half-width tail tile
};

struct TileCoord {
  int p, m_blk, n_blk, kb0, kb1;
  int n0, ncols;         // first output column and width of the tile (BN, or BN / 2 for the half tiles of the tail)
};

// With PAIR `t` indexes PAIRS of vertically adjacent m-blocks: P.mblocks then holds the number of pairs and the CTA of
// cluster rank r takes m-block 2 * pair + r.
template <int BN>
__device__ __forceinline__ TileCoord decode_tile(const TcArgs& args, int t, bool n_fastest, int mc_rank = -1) {
  TileCoord tc_;
  int half = -1;
  if (t >= args.half_start) {          // the last partial wave runs as half-width tiles: 0.55 instead of 1 tile time
    const int u = t - args.half_start;
    half = u & 1;
    t = args.half_start + (u >> 1);
  }
  int p = 0, start = 0;
#pragma unroll
  for (int i = 0; i < MAX_GROUP - 1; ++i)
    if (i + 1 < args.count && t >= args.p[i].tile_end) { p = i + 1; start = args.p[i].tile_end; }
  const TcProblem& P = args.p[p];
  const int tl = t - start;
  const int tiles_mn = P.mblocks * P.nblocks;
  const int split = tl / tiles_mn, rem = tl - split * tiles_mn;
  const int kblocks = (P.K + BK - 1) / BK;
  const int per = (kblocks + P.splits - 1) / P.splits;
  tc_.p = p;
  if (n_fastest) {   // activations (A, large) stream from HBM once; the n-blocks of one m-block run side by side
    tc_.m_blk = rem / P.nblocks;
    tc_.n_blk = rem - tc_.m_blk * P.nblocks;
  } else {
    tc_.n_blk = rem / P.mblocks;
    tc_.m_blk = rem - tc_.n_blk * P.mblocks;
  }
  if (mc_rank >= 0) tc_.m_blk = 2 * tc_.m_blk + mc_rank;
  tc_.kb0 = split * per;
  tc_.kb1 = min(kblocks, tc_.kb0 + per);
  tc_.ncols = half < 0 ? BN : BN / 2;
  tc_.n0 = tc_.n_blk * BN + (half > 0 ? BN / 2 : 0);
  return tc_;
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// Compile-time epilogue (EF >= 0): which inputs / outputs exist is known to the compiler, so the per-chunk flag tests,
// pointer tests and their register shuffles disappear (ncu: 158 SASS instructions per 32 x 32 chunk of a PLAIN
// epilogue, ~25 of them essential).  EF < 0: everything decided at run time from args.flags and the pointers.
enum : int { EPI_GENERIC = -1, EPI_PLAIN = 0, EPI_BIAS = 1, EPI_RES = 2, EPI_GELU = 4, EPI_DGELU = 8, EPI_COLSUM = 16 };

template <int BN, bool A_MN, bool B_MN, bool PAIR, int EF>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(const __grid_constant__ TcMaps maps, const __grid_constant__ TcArgs args) {
  using C_ = Cfg<BN, PAIR, EF>;
  constexpr int STAGES = C_::STAGES;
  constexpr int NBUF = C_::STG_BUFS;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint8_t* stg_base = smem + STAGES * C_::STAGE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(stg_base + EPI_WARPS * STG_BYTES * NBUF);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int total_tiles = args.total_tiles;
  // PAIR: the two CTAs of a cluster walk the same (pair) tile sequence; rank 0 (the leader) issues the MMAs for both
  const int mc_rank = PAIR ? static_cast<int>(tc::cluster_ctarank()) : -1;
  const int tile0 = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int tile_step = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < args.count; ++i) {
      tc::prefetch_tmap(&maps.a[i]);
      tc::prefetch_tmap(&maps.b[i]);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);                 // PAIR: only the leader's is used (both CTAs' loads count on it)
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&tfull_bar[s], 1);
      tc::mbar_init(&tempty_bar[s], (PAIR ? 2 : 1) * EPI_WARPS);        // one arrival per epilogue warp; PAIR: the leader's collects both CTAs'
    }
    tc::fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (PAIR) {
      tc::tmem_alloc_pair(tmem_ptr, 2 * BN);
      tc::tmem_relinquish_pair();
    } else {
      tc::tmem_alloc(tmem_ptr, 2 * BN);
      tc::tmem_relinquish();
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) tc::cluster_sync();             // the peer's barriers exist before anything arrives on them
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_launch();          // TMEM is held: dependents may take what the SM has left
  pdl_wait();            // nothing above touches global memory

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = tile0; t < total_tiles; t += tile_step) {
        const TileCoord tc_ = decode_tile<BN>(args, t, !A_MN, mc_rank);
        const CUtensorMap* tmA = &maps.a[tc_.p];
        const CUtensorMap* tmB = &maps.b[tc_.p];
        {
          // pull this tile's epilogue input (residual / saved gelu') into L2 while the main loop runs
          const TcProblem& P = args.p[tc_.p];
          if (MB_DIAG_EPI != 1 && !(args.flags & GEMM_OUT_F32) && ((args.flags & GEMM_DGELU) || P.residual != nullptr)) {
            for (int cc = 0; cc < BN / 64; ++cc) {
              const int c0 = tc_.n0 + cc * 64;
              if (cc * 64 < tc_.ncols && c0 < P.N) tc::tma_prefetch_l2_2d(&maps.pf[tc_.p], c0, tc_.m_blk * BM);
            }
          }
        }
        for (int kb = tc_.kb0; kb < tc_.kb1; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sA = smem + stage * C_::STAGE_BYTES;
          uint8_t* sB = sA + A_BYTES;
          const bool half_tile = tc_.ncols != BN;       // tail tile: half the B rows / columns are loaded (and counted)
          if constexpr (!PAIR) {
            tc::mbar_expect_tx(&full_bar[stage], half_tile ? A_BYTES + C_::B_BYTES / 2 : C_::STAGE_BYTES);
            if constexpr (!A_MN) {
              tc::tma_load_2d(sA, tmA, &full_bar[stage], kb * BK, tc_.m_blk * BM);
            } else {
#pragma unroll
              for (int i = 0; i < BM / 64; ++i)
                tc::tma_load_2d(sA + i * 8192, tmA, &full_bar[stage], tc_.m_blk * BM + i * 64, kb * BK);
            }
            if constexpr (!B_MN) {
              tc::tma_load_2d(sB, half_tile ? &maps.bh[tc_.p] : tmB, &full_bar[stage], kb * BK, tc_.n0);
            } else {
#pragma unroll
              for (int i = 0; i < BN / 64; ++i)
                if (i * 64 < tc_.ncols) tc::tma_load_2d(sB + i * 8192, tmB, &full_bar[stage], tc_.n0 + i * 64, kb * BK);
            }
          } else {
            // both CTAs' bytes are counted on the LEADER's full barrier (the leader alone waits on it and issues)
            const uint32_t fb = tc::mapa_u32(&full_bar[stage], 0);
            if (mc_rank == 0) tc::mbar_expect_tx(&full_bar[stage], half_tile ? 2 * A_BYTES + C_::B_BYTES : 2 * C_::STAGE_BYTES);
            if constexpr (!A_MN) {
              tc::tma_load_2d_pair(sA, tmA, fb, kb * BK, tc_.m_blk * BM);
            } else {
#pragma unroll
              for (int i = 0; i < BM / 64; ++i)
                tc::tma_load_2d_pair(sA + i * 8192, tmA, fb, tc_.m_blk * BM + i * 64, kb * BK);
            }
            const int n0 = tc_.n0 + mc_rank * (tc_.ncols / 2);       // this CTA's half of the B tile (a half tile: the first 64 / 32 rows of the box)
            if constexpr (!B_MN) {
              tc::tma_load_2d_pair(sB, half_tile ? &maps.bh[tc_.p] : tmB, fb, kb * BK, n0);
            } else {
#pragma unroll
              for (int i = 0; i < BN / 128; ++i)
                if (i * 128 < tc_.ncols) tc::tma_load_2d_pair(sB + i * 8192, tmB, fb, n0 + i * 64, kb * BK);
            }
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0 && (!PAIR || mc_rank == 0)) {
      constexpr uint32_t idesc_full = tc::make_idesc_bf16(PAIR ? 2 * BM : BM, BN, A_MN ? 1 : 0, B_MN ? 1 : 0);
      constexpr uint32_t idesc_half = tc::make_idesc_bf16(PAIR ? 2 * BM : BM, BN / 2, A_MN ? 1 : 0, B_MN ? 1 : 0);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int tile_i = 0;
      for (int t = tile0; t < total_tiles; t += tile_step, ++tile_i) {
        const TileCoord tc_ = decode_tile<BN>(args, t, !A_MN, mc_rank);
        tc::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        tc::tc_fence_after();
        MB_STAMP(tile_i, 0);
        const uint32_t d_tmem = tmem_base + acc * BN;
        const uint32_t idesc = tc_.ncols == BN ? idesc_full : idesc_half;
        for (int kb = tc_.kb0; kb < tc_.kb1; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::tc_fence_after();
          const uint32_t a_addr = tc::smem_u32(smem + stage * C_::STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // K-major, 128B swizzle: rows are 128 B, 8-row groups 1024 B apart; K advance = 32 B inside the row.
            // MN-major, 128B swizzle: 64-element MN chunks 8192 B apart (one TMA box each), 8-row K groups 1024 B
            // apart; K advance = 16 rows = 2048 B.
            const uint64_t da = A_MN ? tc::make_smem_desc(a_addr + k * UMMA_K * 128, 8192, 1024)
                                     : tc::make_smem_desc(a_addr + k * UMMA_K * 2, 16, 1024);
            const uint64_t db = B_MN ? tc::make_smem_desc(b_addr + k * UMMA_K * 128, 8192, 1024)
                                     : tc::make_smem_desc(b_addr + k * UMMA_K * 2, 16, 1024);
            if constexpr (PAIR) tc::umma_bf16_pair(d_tmem, da, db, idesc, (kb > tc_.kb0 || k > 0) ? 1u : 0u);
            else tc::umma_bf16(d_tmem, da, db, idesc, (kb > tc_.kb0 || k > 0) ? 1u : 0u);
          }
          if constexpr (PAIR) tc::umma_commit_pair(&empty_bar[stage], 0x3);   // frees the stage in both CTAs
          else tc::umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if constexpr (PAIR) tc::umma_commit_pair(&tfull_bar[acc], 0x3);       // both CTAs' epilogues read their half
        else tc::umma_commit(&tfull_bar[acc]);
        MB_STAMP(tile_i, 1);
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= EPI_WARP0) {
    // ===================== epilogue (16 warps: 4 TMEM lane groups x 4 column quarters) =====================
    // Four warps per scheduler hide the MUFU / shared-memory / TMEM latencies of the fused math; each warp owns a
    // 32-row x (BN / 4)-column slab of the accumulator and walks it in 32-column chunks.
    const int ew = warp & 3;                       // TMEM lanes [32 * ew, 32 * ew + 32) (a warp may only touch its quarter)
    const int quarter = (warp - EPI_WARP0) >> 2;   // columns [quarter * BN / 4, (quarter + 1) * BN / 4)
    constexpr int CH_PER_Q = BN / 128;             // 32-column chunks per warp
    uint8_t* stg = stg_base + (warp - EPI_WARP0) * STG_BYTES * NBUF;
    const uint32_t stg_u32 = tc::smem_u32(stg);
    // staging buffer: 32 rows x 64 B, 64-byte swizzle (16-byte chunk j of row r sits at chunk j ^ ((r >> 1) & 3)),
    // as the TMA store map expects.  "own" = row `lane`; "co" = coalesced view, instruction i covers rows 8i..8i+7.
    const uint32_t own_base = stg_u32 + lane * 64;
    const uint32_t own_x = static_cast<uint32_t>((lane >> 1) & 3);
    const uint32_t co_row = static_cast<uint32_t>(lane >> 2), co_j = static_cast<uint32_t>(lane & 3);
    int acc = 0;
    uint32_t acc_phase = 0;
    uint32_t sboff = 0;       // byte offset of the staging buffer the next store is staged in (NBUF == 2: alternates)
    // hand the accumulator back to the MMA issuer (PAIR: the leader CTA's barrier counts both CTAs' epilogue threads)
    const uint32_t te0 = PAIR ? tc::mapa_u32(&tempty_bar[0], 0) : 0u, te1 = PAIR ? tc::mapa_u32(&tempty_bar[1], 0) : 0u;
    auto release_acc = [&](int a) {   // after tcgen05.wait::ld + fence::before_thread_sync of every lane
      __syncwarp();
      if (lane == 0) {
        if constexpr (PAIR) tc::mbar_arrive_cluster(a == 0 ? te0 : te1);
        else tc::mbar_arrive(&tempty_bar[a]);
      }
    };
    constexpr bool TMA_ST = MB_EPI_TMA_STORE != 0;
    // staged 32 x 32 bf16 chunk (every lane has written its own row) -> global memory, coalesced: instruction i covers
    // rows 8i .. 8i+7, four lanes per 64-byte row segment (whole sectors, no completion to wait for)
    auto flush_chunk = [&](bf16* base, const TcProblem& P, int row0, int col0) {
      __syncwarp();
      const int cc = col0 + static_cast<int>(co_j) * 8;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t rr = i * 8 + co_row;
        const uint32_t a = stg_u32 + rr * 64 + ((co_j ^ ((rr >> 1) & 3)) << 4);
        uint4 v;
        asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
        const int row = row0 + static_cast<int>(rr);
        if (MB_DIAG_EPI != 2 && row < P.M && cc < P.N)
          *reinterpret_cast<uint4*>(base + static_cast<size_t>(row) * P.ldc + cc) = v;
      }
    };
    constexpr bool SPEC = EF >= 0;
    const bool out_f32 = SPEC ? false : (args.flags & GEMM_OUT_F32) != 0;
    const bool f_gelu = SPEC ? (EF & EPI_GELU) != 0 : (args.flags & GEMM_GELU) != 0;
    const bool f_dgelu = SPEC ? (EF & EPI_DGELU) != 0 : (args.flags & GEMM_DGELU) != 0;
    const float alpha = SPEC ? 1.0f : args.alpha;
    int tile_i = -1;
    for (int t = tile0; t < total_tiles; t += tile_step) {
      ++tile_i;
      if (warp == EPI_WARP0 && lane == 0) MB_STAMP(tile_i, 2);
      const TileCoord tc_ = decode_tile<BN>(args, t, !A_MN, mc_rank);
      const TcProblem& P = args.p[tc_.p];
      const int row0 = tc_.m_blk * BM + ew * 32;
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BN;
      int nch = 0;
#pragma unroll
      for (int c2 = 0; c2 < CH_PER_Q; ++c2)
        if ((quarter * CH_PER_Q + c2) * 32 < tc_.ncols && tc_.n0 + (quarter * CH_PER_Q + c2) * 32 < P.N) nch = c2 + 1;
      if (nch == 0) {
        tc::mbar_wait(&tfull_bar[acc], acc_phase);
        tc::tc_fence_before();
        release_acc(acc);
      }
      if (!out_f32) {
        // ---------------- bf16 outputs through shared memory + TMA store ----------------
        const bool has_res = SPEC ? (EF & EPI_RES) != 0 : P.residual != nullptr;
        const bool has_bias = SPEC ? (EF & EPI_BIAS) != 0 : P.bias != nullptr;
        const bool has_colsum = SPEC ? (EF & EPI_COLSUM) != 0 : P.colsum != nullptr;
        const bool need_in = (MB_DIAG_EPI != 1) && (f_dgelu || has_res);
        const bf16* src = static_cast<const bf16*>(f_dgelu ? P.aux : P.residual);
#pragma unroll 1
        for (int c2 = 0; c2 < nch; ++c2) {
          const int ch = quarter * CH_PER_Q + c2;
          const int col0 = tc_.n0 + ch * 32;
          // (1) coalesced prefetch of the epilogue input: instruction i covers rows 8i .. 8i+7, 4 lanes per row
          uint4 pre[4];
          if (need_in) {
            const int cc = col0 + co_j * 8;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const int rr = row0 + i * 8 + co_row;
              pre[i] = (rr < P.M && cc < P.N)
                           ? __ldg(reinterpret_cast<const uint4*>(src + static_cast<size_t>(rr) * P.ldc + cc))
                           : make_uint4(0, 0, 0, 0);
            }
          }
          // (2) accumulator chunk -> registers
          if (c2 == 0) {
            tc::mbar_wait(&tfull_bar[acc], acc_phase);
            tc::tc_fence_after();
            if (warp == EPI_WARP0 && lane == 0) MB_STAMP(tile_i, 3);
          }
          uint32_t r[32];
          tc::tmem_ld_32x32(t_addr + ch * 32, r);
          tc::tmem_ld_wait();
          if (c2 == nch - 1) {   // accumulator slab fully read: hand the TMEM buffer back to the MMA warp
            tc::tc_fence_before();
            release_acc(acc);
          }
#if MB_DIAG_EPI == 1
          continue;
#endif
          // (3) transpose the input chunk to "thread = row" through the staging buffer
          if (need_in) {
            if (TMA_ST && lane == 0) tc::tma_store_wait_read_n<NBUF - 1>();
            __syncwarp();
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              const uint32_t rr = i * 8 + co_row;
              const uint32_t a = stg_u32 + sboff + rr * 64 + ((co_j ^ ((rr >> 1) & 3)) << 4);
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(pre[i].x), "r"(pre[i].y),
                           "r"(pre[i].z), "r"(pre[i].w) : "memory");
            }
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t a = own_base + sboff + ((j ^ own_x) << 4);
              asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(pre[j].x), "=r"(pre[j].y), "=r"(pre[j].z),
                           "=r"(pre[j].w) : "r"(a) : "memory");
            }
            __syncwarp();
          }
          // (4) alpha, bias (packed fp32: one FFMA2 / FADD2 per column pair)
          float2* v2 = reinterpret_cast<float2*>(r);
          if (has_bias) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const float4 b = (col0 + q * 4 < P.N) ? __ldg(reinterpret_cast<const float4*>(P.bias + col0 + q * 4))
                                                    : make_float4(0.f, 0.f, 0.f, 0.f);
              if constexpr (SPEC) {
                v2[2 * q] = add2(v2[2 * q], make_float2(b.x, b.y));
                v2[2 * q + 1] = add2(v2[2 * q + 1], make_float2(b.z, b.w));
              } else {
                v2[2 * q] = fma2(v2[2 * q], splat2(alpha), make_float2(b.x, b.y));
                v2[2 * q + 1] = fma2(v2[2 * q + 1], splat2(alpha), make_float2(b.z, b.w));
              }
            }
          } else if (!SPEC && alpha != 1.0f) {
#pragma unroll
            for (int q = 0; q < 16; ++q) v2[q] = mul2(v2[q], splat2(alpha));
          }
          // (5) GELU: second output = gelu'(pre-activation) (the backward's dgrad then only multiplies by it; the
          // erf / exp terms are shared between the activation and its derivative)
          uint32_t o[16];
          if (f_gelu) {
            if (TMA_ST && lane == 0) tc::tma_store_wait_read_n<NBUF - 1>();
            __syncwarp();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint32_t w[4];
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                float2 d;
                const float2 g = gelu_fast_both2(v2[4 * j + q], d);
                w[q] = pack_bf16x2(d.x, d.y);
                o[4 * j + q] = pack_bf16x2(g.x, g.y);   // packed at once: 16 live registers instead of 32
              }
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(own_base + sboff + ((j ^ own_x) << 4)), "r"(w[0]),
                           "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
            }
            if constexpr (TMA_ST) {
              tc::fence_proxy_async();
              __syncwarp();
              if (lane == 0 && MB_DIAG_EPI != 2) {
                tc::tma_store_2d(&maps.x[tc_.p], stg + sboff, col0, row0);
                tc::tma_store_commit();
              }
              if constexpr (NBUF == 2) sboff ^= STG_BYTES;
            } else {
              flush_chunk(static_cast<bf16*>(P.aux), P, row0, col0);
            }
          }
          // (6) GELU' / residual
          if (need_in) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint32_t w[4] = {pre[j].x, pre[j].y, pre[j].z, pre[j].w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const float2 in = make_float2(__uint_as_float(w[q] << 16), __uint_as_float(w[q] & 0xffff0000u));
                v2[4 * j + q] = f_dgelu ? mul2(v2[4 * j + q], in) : add2(v2[4 * j + q], in);
              }
            }
          }
          // (7) main output
#if MB_DIAG_EPI == 3
          if (!f_gelu && !has_colsum) {
            if (row0 + lane < P.M) {
              bf16* crow = static_cast<bf16*>(P.C) + static_cast<size_t>(row0 + lane) * P.ldc + col0;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (col0 + j * 8 < P.N)
                  *reinterpret_cast<uint4*>(crow + j * 8) =
                      make_uint4(pack_bf16x2(v2[4 * j].x, v2[4 * j].y), pack_bf16x2(v2[4 * j + 1].x, v2[4 * j + 1].y),
                                 pack_bf16x2(v2[4 * j + 2].x, v2[4 * j + 2].y), pack_bf16x2(v2[4 * j + 3].x, v2[4 * j + 3].y));
            }
            continue;
          }
#endif
          if (TMA_ST && lane == 0) tc::tma_store_wait_read_n<NBUF - 1>();
          __syncwarp();
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t w0, w1, w2, w3;
            if (f_gelu) {
              w0 = o[4 * j]; w1 = o[4 * j + 1]; w2 = o[4 * j + 2]; w3 = o[4 * j + 3];
            } else {
              w0 = pack_bf16x2(v2[4 * j].x, v2[4 * j].y); w1 = pack_bf16x2(v2[4 * j + 1].x, v2[4 * j + 1].y);
              w2 = pack_bf16x2(v2[4 * j + 2].x, v2[4 * j + 2].y); w3 = pack_bf16x2(v2[4 * j + 3].x, v2[4 * j + 3].y);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(own_base + sboff + ((j ^ own_x) << 4)), "r"(w0),
                         "r"(w1), "r"(w2), "r"(w3) : "memory");
          }
          if constexpr (TMA_ST) {
            tc::fence_proxy_async();
            __syncwarp();
            if (lane == 0 && MB_DIAG_EPI != 2) {
              tc::tma_store_2d(&maps.c[tc_.p], stg + sboff, col0, row0);
              tc::tma_store_commit();
            }
          } else {
            flush_chunk(static_cast<bf16*>(P.C), P, row0, col0);
          }
          // (8) optional column sums of the rounded outputs (bias gradient of the layer that produced this dY): lane
          // l walks column l of the staged 32 x 32 chunk (one conflict-free 64-byte row segment per step)
          if (has_colsum) {
            const int nrows = min(32, P.M - row0);
            const uint32_t cbase = stg_u32 + sboff + (lane & 7) * 2;
            const uint32_t cj = static_cast<uint32_t>(lane >> 3);
            float cs0 = 0.f, cs1 = 0.f;
#pragma unroll 8
            for (int rr = 0; rr < 32; rr += 2) {
              uint16_t h0, h1;
              const uint32_t a0 = cbase + rr * 64 + ((cj ^ ((rr >> 1) & 3)) << 4);
              asm volatile("ld.shared.u16 %0, [%1];" : "=h"(h0) : "r"(a0) : "memory");
              asm volatile("ld.shared.u16 %0, [%1];" : "=h"(h1) : "r"(a0 + 64) : "memory");
              if (rr < nrows) cs0 += __uint_as_float(static_cast<uint32_t>(h0) << 16);
              if (rr + 1 < nrows) cs1 += __uint_as_float(static_cast<uint32_t>(h1) << 16);
            }
            if (nrows > 0 && col0 + lane < P.N) atomicAdd(P.colsum + col0 + lane, cs0 + cs1);
          }
          if constexpr (TMA_ST && NBUF == 2) sboff ^= STG_BYTES;   // the next chunk (or GELU output) stages into the other buffer
        }
      } else if (nch > 0) {
        // ---------------- fp32 outputs (gradient accumulation): direct 16-byte stores / red.add ----------------
        tc::mbar_wait(&tfull_bar[acc], acc_phase);
        tc::tc_fence_after();
#pragma unroll 1
        for (int c2 = 0; c2 < nch; ++c2) {
          const int ch = quarter * CH_PER_Q + c2;
          const int col0 = tc_.n0 + ch * 32;
          uint32_t r[32];
          tc::tmem_ld_32x32(t_addr + ch * 32, r);
          tc::tmem_ld_wait();
          if (c2 == nch - 1) {
            tc::tc_fence_before();
            release_acc(acc);
          }
          if (MB_DIAG_EPI != 1) {
            // 16 columns at a time through the staging buffer (32 rows x 64 B, same swizzle as the bf16 chunks), then
            // coalesced: one red / store instruction covers 8 rows x 64 B instead of 32 rows x 16 B
            const bool accum = (args.flags & GEMM_ACCUM) != 0;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              __syncwarp();
#pragma unroll
              for (int j = 0; j < 4; ++j)
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(own_base + ((j ^ own_x) << 4)),
                             "r"(r[16 * h + 4 * j]), "r"(r[16 * h + 4 * j + 1]), "r"(r[16 * h + 4 * j + 2]),
                             "r"(r[16 * h + 4 * j + 3]) : "memory");
              __syncwarp();
              const int col = col0 + h * 16 + static_cast<int>(co_j) * 4;
              float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
              if (P.bias != nullptr && col < P.N) b = __ldg(reinterpret_cast<const float4*>(P.bias + col));
#pragma unroll
              for (int i = 0; i < 4; ++i) {
                const uint32_t rr = i * 8 + co_row;
                const uint32_t a = stg_u32 + rr * 64 + ((co_j ^ ((rr >> 1) & 3)) << 4);
                float4 o;
                asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(o.x), "=f"(o.y), "=f"(o.z), "=f"(o.w) : "r"(a) : "memory");
                const int row = row0 + static_cast<int>(rr);
                if (row < P.M && col < P.N) {
                  o.x = fmaf(o.x, alpha, b.x); o.y = fmaf(o.y, alpha, b.y); o.z = fmaf(o.z, alpha, b.z); o.w = fmaf(o.w, alpha, b.w);
                  float4* dst = reinterpret_cast<float4*>(static_cast<float*>(P.C) + static_cast<size_t>(row) * P.ldc + col);
                  if (accum) atomicAdd(dst, o);
                  else *dst = o;
                }
              }
            }
          }
        }
      }
      if (warp == EPI_WARP0 && lane == 0) MB_STAMP(tile_i, 4);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) tc::tma_store_wait_all();   // shared memory must stay valid until the last store has read it
  }

  tc::tc_fence_before();
  __syncthreads();
  if constexpr (PAIR) tc::cluster_sync();             // no CTA leaves while its peer may still read / signal into it
  if (warp == 2) {
    if constexpr (PAIR) tc::tmem_dealloc_pair(tmem_base, 2 * BN);
    else tc::tmem_dealloc(tmem_base, 2 * BN);
  }
}

// ---------------------------------------------------------------- host
// 2D bf16 tensor [d1 rows, d0 cols] with row pitch stride1 (elements), box [b1 rows, b0 cols], 128B swizzle.
int get_tmap(CUtensorMap* out, const void* ptr, uint64_t d0, uint64_t d1, uint64_t stride1, uint32_t b0, uint32_t b1,
             int swizzle = 128) {
  const uint64_t dims[2] = {d0, d1}, strides[1] = {stride1};
  const uint32_t box[2] = {b0, b1};
  return mb_get_tmap(out, ptr, 2, dims, strides, box, swizzle);
}

int num_sms() { return mb_num_sms(); }

// Half-width tiles in the last partial wave: off unless MB_GEMM_HALF_TAIL=1 or the GEMM_HALF_TAIL hint is set.
// Measured (profiles/r2b_summary.md section 7): 0-2 % per launch -- a half tile is not much shorter than a full one
// because the pair main loop is bound by the chip-wide L2 -> SM feed (43 B / cycle / SM = the 6300 B / cycle LTS
// limit), which a partially filled tail wave does not hit anyway.
bool half_tail_on() {
  static const bool on = [] { const char* e = std::getenv("MB_GEMM_HALF_TAIL"); return e && e[0] == '1'; }();
  return on;
}

// PAIR: 2-CTA clusters running cta_group::2 MMAs (see the kernel); tiles are scheduled as vertical pairs of m-blocks.
template <int BN, bool A_MN, bool B_MN, bool MC, int EF = EPI_GENERIC>
int launch(const GemmParams* Ps, int count, cudaStream_t st) {
  using C_ = Cfg<BN, MC, EF>;
  TcMaps maps;
  TcArgs a;
  std::memset(&a, 0, sizeof(a));
  const int flags = Ps[0].flags;
  const bool out_f32 = (flags & GEMM_OUT_F32) != 0;
  int rc;
  long long tiles_mn_total = 0;
  for (int i = 0; i < count; ++i) {
    const GemmParams& P = Ps[i];
    if (!A_MN) rc = get_tmap(&maps.a[i], P.A, P.K, P.M, P.sam, BK, BM);
    else rc = get_tmap(&maps.a[i], P.A, P.M, P.K, P.sak, 64, BK);
    if (rc != 0) return rc;
    if (!B_MN) rc = get_tmap(&maps.b[i], P.B, P.K, P.N, P.sbn, BK, MC ? BN / 2 : BN);
    else rc = get_tmap(&maps.b[i], P.B, P.N, P.K, P.sbk, 64, BK);
    if (rc != 0) return rc;
    maps.bh[i] = maps.b[i];
    if (!B_MN && BN == 256 && (rc = get_tmap(&maps.bh[i], P.B, P.K, P.N, P.sbn, BK, MC ? BN / 4 : BN / 2)) != 0) return rc;
    if (!out_f32) {
      if ((rc = get_tmap(&maps.c[i], P.C, P.N, P.M, P.ldc, 32, 32, 64)) != 0) return rc;
      // x: GELU -> saved gelu'(pre-activation) (second output); GELU' -> the same tensor as epilogue input; else residual
      const void* xp = (flags & (GEMM_GELU | GEMM_DGELU)) ? P.aux : P.residual;
      if ((flags & (GEMM_GELU | GEMM_DGELU)) && xp == nullptr) return -6;
      if (xp != nullptr) {
        if ((reinterpret_cast<uintptr_t>(xp) & 15) != 0) return -11;
        if ((rc = get_tmap(&maps.x[i], xp, P.N, P.M, P.ldc, 32, 32, 64)) != 0) return rc;
        if ((rc = get_tmap(&maps.pf[i], xp, P.N, P.M, P.ldc, 64, BM)) != 0) return rc;
      } else {
        maps.x[i] = maps.c[i];
        maps.pf[i] = maps.c[i];
      }
    } else {
      maps.c[i] = maps.a[i];
      maps.x[i] = maps.a[i];
      maps.pf[i] = maps.a[i];
    }
    TcProblem& q = a.p[i];
    q.C = P.C; q.bias = P.bias; q.residual = P.residual; q.aux = P.aux; q.colsum = P.colsum;
    q.M = P.M; q.N = P.N; q.K = P.K; q.ldc = P.ldc;
    q.mblocks = (P.M + BM - 1) / BM;
    if (MC) q.mblocks = (q.mblocks + 1) / 2;
    q.nblocks = (P.N + BN - 1) / BN;
    tiles_mn_total += static_cast<long long>(q.mblocks) * q.nblocks;
  }
  for (int i = count; i < MAX_GROUP; ++i) {
    maps.a[i] = maps.a[0]; maps.b[i] = maps.b[0]; maps.c[i] = maps.c[0]; maps.x[i] = maps.x[0];
    maps.pf[i] = maps.pf[0]; maps.bh[i] = maps.bh[0];
  }
  // Split-K (weight gradients; fp32 red.add combines the partials): work items of at most `split_target` k-blocks.
  // The target minimises the makespan of the persistent schedule (item t runs on CTA t mod grid; an item costs its
  // k-blocks plus an epilogue) over the whole group -- "cover the chip twice" left e.g. 180 items for 74 CTA pairs
  // (3 waves of 47 k-blocks, the last one 43 % full) where 144 items (2 waves of 59) do the same work 16 % sooner.
  int split_target = 1 << 30;
  if (flags & GEMM_ACCUM) {
    const int slots = MC ? num_sms() / 2 : num_sms();
    int kbs[MAX_GROUP], max_kb = 1;
    for (int i = 0; i < count; ++i) { kbs[i] = (a.p[i].K + BK - 1) / BK; max_kb = std::max(max_kb, kbs[i]); }
    // a 128 x 256 fp32 red epilogue costs about as much as EPI_KB k-blocks (MB_GEMM_SPLIT_EPI: sweeps, tools/gemm_bench.py)
    static const int EPI_KB = [] { const char* e = std::getenv("MB_GEMM_SPLIT_EPI"); return e ? std::atoi(e) : 6; }();
    long long best = -1;
    int prev_sp[MAX_GROUP] = {0, 0, 0, 0};
    int load[256];
    for (int tgt = max_kb; tgt >= std::min(max_kb, 4); --tgt) {
      int sp[MAX_GROUP];
      bool same = true;
      long long items = 0;
      for (int i = 0; i < count; ++i) {
        sp[i] = (kbs[i] + tgt - 1) / tgt;
        same = same && sp[i] == prev_sp[i];
        items += static_cast<long long>(a.p[i].mblocks) * a.p[i].nblocks * sp[i];
      }
      if (same) continue;
      for (int i = 0; i < count; ++i) prev_sp[i] = sp[i];
      if (items > 16 * slots) break;    // finer than 16 waves never wins
      const int grid = static_cast<int>(std::min<long long>(items, std::min(slots, 256)));
      for (int c = 0; c < grid; ++c) load[c] = 0;
      int c = 0;
      for (int i = 0; i < count; ++i) {
        const int len = (kbs[i] + sp[i] - 1) / sp[i] + EPI_KB;
        const int n = a.p[i].mblocks * a.p[i].nblocks * sp[i];
        for (int t = 0; t < n; ++t) { load[c] += len; if (++c == grid) c = 0; }
      }
      int span = 0;
      for (int k = 0; k < grid; ++k) span = std::max(span, load[k]);
      // +items: red traffic grows with the number of partial tiles (tie-break towards fewer splits)
      const long long cost = static_cast<long long>(span) * 4096 + items;
      if (best < 0 || cost < best) { best = cost; split_target = tgt; }
    }
  }
  int total = 0;
  for (int i = 0; i < count; ++i) {
    TcProblem& q = a.p[i];
    const int kblocks = (q.K + BK - 1) / BK;
    int splits = 1;
    if (flags & GEMM_ACCUM) {
      splits = (kblocks + split_target - 1) / split_target;
      const int per = (kblocks + splits - 1) / splits;
      splits = (kblocks + per - 1) / per;  // no empty split
    }
    q.splits = splits;
    total += q.mblocks * q.nblocks * splits;
    q.tile_end = total;
  }
  a.count = count;
  a.flags = flags;
  a.alpha = Ps[0].alpha;
  a.total_tiles = total;
  a.half_start = 1 << 30;
  auto kern = gemm_tc_kernel<BN, A_MN, B_MN, MC, EF>;
  static bool attr_set = false;
  static int max_clusters = 0;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C_::SMEM_BYTES);
    if (MC) {
      cudaLaunchConfig_t qc = {};
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
      qc.gridDim = dim3(num_sms() & ~1); qc.blockDim = dim3(NUM_THREADS); qc.dynamicSmemBytes = C_::SMEM_BYTES;
      qc.attrs = qa; qc.numAttrs = 1;
      if (cudaOccupancyMaxActiveClusters(&max_clusters, kern, &qc) != cudaSuccess || max_clusters < 1) {
        cudaGetLastError();
        max_clusters = num_sms() / 2;
      }
      if (max_clusters > num_sms() / 2) max_clusters = num_sms() / 2;
    }
    attr_set = true;
  }
  // Tail: the R = total mod slots tiles of the last (partial) wave take a whole tile time on R of the slots.  When
  // 2 R <= slots they run as 2 R half-width tiles instead (a 128-wide tile costs 0.55 of a 256-wide one here: same
  // A traffic, half the MMAs and half the epilogue) -- e.g. the N = 768 down-projections: 180 pair tiles on 74
  // clusters = 2 waves + 32 tiles -> 2 waves + 64 half tiles, 2.55 instead of 3 tile times.  bf16 outputs, BN = 256 only.
  if (BN == 256 && (half_tail_on() || (flags & GEMM_HALF_TAIL)) && !(flags & (GEMM_OUT_F32 | GEMM_ACCUM))) {
    const int slots = MC ? max_clusters : num_sms();
    const int R = total % slots;
    if (R > 0 && 2 * R <= slots) {
      a.half_start = total - R;
      total += R;
      a.total_tiles = total;
    }
  }
  if (!MC) {
    const int grid = total < num_sms() ? total : num_sms();
    cudaError_t le = mb_launch(kern, dim3(grid), dim3(NUM_THREADS), C_::SMEM_BYTES, st, maps, a);
    if (le != cudaSuccess) return static_cast<int>(le);
  } else {
    const int clusters = total < max_clusters ? total : max_clusters;
    cudaLaunchConfig_t cfg = {};
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(NUM_THREADS); cfg.dynamicSmemBytes = C_::SMEM_BYTES;
    cfg.stream = st; cfg.attrs = at; cfg.numAttrs = mb_pdl_enabled() ? 2 : 1;
    cudaError_t le = cudaLaunchKernelEx(&cfg, kern, maps, a);
    if (le != cudaSuccess) return static_cast<int>(le);
  }
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : static_cast<int>(e);
}

// MB_GEMM_PAIR = 0: never, 1: always, 2 (default): where it measured faster (tools/gemm_bench.py,
// profiles/r2b_gemm_epilogue.md).  A pair stages 32 KB instead of 48 KB per k-block into each SM, which is what lets
// the epilogue's stores and the operand loads share the L2: every shape of the transformer blocks gains 6-12 % except
// the short-K bias + residual out-projections (equal), which stay on single-CTA tiles.  (Before r2b the pair's
// accumulator hand-back was a cluster-scope release arrive = MEMBAR.ALL.GPU per epilogue thread, and pairs had no
// compile-time epilogues; both made short-K pairs slower than single tiles.)
int pair_mode() {
  static int mode = -1;
  if (mode < 0) {
    const char* e = std::getenv("MB_GEMM_PAIR");
    mode = (e != nullptr && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 2;
  }
  return mode;
}

template <int BN, bool A_MN, bool B_MN, bool MC>
int launch_epi(const GemmParams* Ps, int count, cudaStream_t st, int ef) {
  if constexpr (!A_MN) {
    if constexpr (!B_MN) {      // forward: y = x W^T
      if (ef == EPI_PLAIN) return launch<BN, A_MN, B_MN, MC, EPI_PLAIN>(Ps, count, st);
      if (ef == (EPI_BIAS | EPI_RES)) return launch<BN, A_MN, B_MN, MC, EPI_BIAS | EPI_RES>(Ps, count, st);
      if (ef == (EPI_BIAS | EPI_GELU)) return launch<BN, A_MN, B_MN, MC, EPI_BIAS | EPI_GELU>(Ps, count, st);
    } else {                    // dgrad: dx = dy W
      if (ef == EPI_PLAIN) return launch<BN, A_MN, B_MN, MC, EPI_PLAIN>(Ps, count, st);
      if (ef == EPI_COLSUM) return launch<BN, A_MN, B_MN, MC, EPI_COLSUM>(Ps, count, st);
      if (ef == (EPI_DGELU | EPI_COLSUM)) return launch<BN, A_MN, B_MN, MC, EPI_DGELU | EPI_COLSUM>(Ps, count, st);
      if (ef == EPI_DGELU) return launch<BN, A_MN, B_MN, MC, EPI_DGELU>(Ps, count, st);
    }
  }
  return launch<BN, A_MN, B_MN, MC>(Ps, count, st);
}

template <int BN, bool A_MN, bool B_MN>
int launch_any(const GemmParams* Ps, int count, cudaStream_t st) {
  bool mc = pair_mode() == 1;
  if (pair_mode() == 2) {
    static const int pair_min_k = [] { const char* e = std::getenv("MB_GEMM_PAIR_MINK"); return e ? std::atoi(e) : 4000; }();
    long long singles = 0, pairs = 0;
    int max_k = 0;
    for (int i = 0; i < count; ++i) {
      const long long mb = (Ps[i].M + BM - 1) / BM, nb = (Ps[i].N + BN - 1) / BN;
      singles += mb * nb;
      pairs += (mb + 1) / 2 * nb;
      max_k = Ps[i].K > max_k ? Ps[i].K : max_k;
    }
    const bool pairable = 40 * pairs <= 21 * singles;     // at most 5 % of the pair tiles' rows are padding
    if (Ps[0].flags & GEMM_ACCUM) mc = pairable && max_k >= pair_min_k;
    else if (Ps[0].flags & GEMM_OUT_F32) mc = false;
    else mc = pairable && !(Ps[0].residual != nullptr && max_k <= 1024);
  }
  if (Ps[0].flags & GEMM_PAIR_ON) mc = true;
  if (Ps[0].flags & GEMM_PAIR_OFF) mc = false;
  // compile-time epilogues for the shapes of the transformer blocks (bf16 out, alpha = 1, the same inputs / outputs
  // in every problem of the group); MB_GEMM_EPI=0 keeps the run-time epilogue (A/B measurements)
  static const bool spec_on = [] { const char* e = std::getenv("MB_GEMM_EPI"); return !(e && e[0] == '0'); }();
  int ef = EPI_GENERIC;
  if (spec_on && !(Ps[0].flags & (GEMM_OUT_F32 | GEMM_ACCUM)) && Ps[0].alpha == 1.0f) {
    ef = 0;
    if (Ps[0].flags & GEMM_GELU) ef |= EPI_GELU;
    if (Ps[0].flags & GEMM_DGELU) ef |= EPI_DGELU;
    if (Ps[0].bias != nullptr) ef |= EPI_BIAS;
    if (Ps[0].residual != nullptr) ef |= EPI_RES;
    if (Ps[0].colsum != nullptr) ef |= EPI_COLSUM;
    for (int i = 1; i < count; ++i)
      if ((Ps[i].bias != nullptr) != (Ps[0].bias != nullptr) || (Ps[i].residual != nullptr) != (Ps[0].residual != nullptr) ||
          (Ps[i].colsum != nullptr) != (Ps[0].colsum != nullptr))
        ef = EPI_GENERIC;
  }
  return mc ? launch_epi<BN, A_MN, B_MN, true>(Ps, count, st, ef) : launch_epi<BN, A_MN, B_MN, false>(Ps, count, st, ef);
}

}  // namespace

// bf16 only.  `params` points to `count` (1..4) problems launched as ONE persistent kernel; they must share the
// operand layout, flags and alpha (M, N, K and all pointers may differ).
// Layouts: A K-major (sak == 1) or MN-major (sam == 1); B K-major (sbk == 1) or MN-major (sbn == 1).
// Unbatched (nb1 == nb2 == 1).  N % 8 == 0, ldc % 8 == 0, all base pointers 16-byte aligned.
extern "C" int mb_gemm_tc_grouped(const GemmParams* params, int count, void* stream) {
  if (count < 1 || count > MAX_GROUP) return -2;
  GemmParams live[MAX_GROUP];
  int n = 0;
  for (int i = 0; i < count; ++i) {
    const GemmParams& P = params[i];
    if (P.M == 0 || P.N == 0) continue;
    if (P.nb1 != 1 || P.nb2 != 1) return -3;
    if (P.N % 8 != 0 || P.ldc % 8 != 0 || P.K <= 0) return -4;
    if (P.flags != params[0].flags || P.alpha != params[0].alpha) return -7;
    if ((P.sak == 1) != (params[0].sak == 1) || (P.sbk == 1) != (params[0].sbk == 1)) return -7;
    if ((reinterpret_cast<uintptr_t>(P.C) & 15) != 0) return -11;
    live[n++] = P;
  }
  if (n == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const GemmParams& P0 = live[0];
  const bool a_k = P0.sak == 1, a_mn = !a_k && P0.sam == 1;
  const bool b_k = P0.sbk == 1, b_mn = !b_k && P0.sbn == 1;
  if (!(a_k || a_mn) || !(b_k || b_mn)) return -5;
  int maxN = 0;
  for (int i = 0; i < n; ++i) maxN = live[i].N > maxN ? live[i].N : maxN;
  bool wide = maxN > 128;
  if (wide && !(P0.flags & GEMM_ACCUM)) {
    // pick the tile width with the smaller quantised cost (waves x tile width; narrow tiles re-read A more)
    long long t256 = 0, t128 = 0;
    for (int i = 0; i < n; ++i) {
      const long long mb = (live[i].M + BM - 1) / BM;
      t256 += mb * ((live[i].N + 255) / 256);
      t128 += mb * ((live[i].N + 127) / 128);
    }
    const long long sms = num_sms();
    const long long w256 = (t256 + sms - 1) / sms * 256;
    // measured (profiles/r2b_gemm_epilogue.md): a 128-wide k-block takes 0.76 of a 256-wide one (128 x 128 MMAs read
    // A + B from shared memory at the port's full rate), not 0.5
    const long long w128 = (t128 + sms - 1) / sms * 195;
    wide = w256 <= w128;
  }
  if (a_k && b_k) return wide ? launch_any<256, false, false>(live, n, st) : launch_any<128, false, false>(live, n, st);
  if (a_k && b_mn) return wide ? launch_any<256, false, true>(live, n, st) : launch_any<128, false, true>(live, n, st);
  if (a_mn && b_mn) return wide ? launch_any<256, true, true>(live, n, st) : launch_any<128, true, true>(live, n, st);
  return -5;
}

#if MB_DIAG_TIMING
extern "C" int mb_diag_gemm_timing(unsigned long long* host, int n) {
  return static_cast<int>(cudaMemcpyFromSymbol(host, g_gemm_stamps, sizeof(unsigned long long) * (n < 512 ? n : 512)));
}
#endif

extern "C" int mb_gemm_tc(const GemmParams* params, void* stream) { return mb_gemm_tc_grouped(params, 1, stream); }
