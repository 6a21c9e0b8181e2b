// Fused (flash-style) attention for sm_100a on tcgen05 tensor cores with TMEM accumulators and TMA loads.
// Replaces vit_pytorch Attention.forward's  softmax(q k^T * dh^-0.5) v  (and its autograd), which the reference
// runs unfused, materialising [B, h, N, N] scores (SURVEY 2.1 "Attention").
//
// Layout: qkv [Bn * N, 3 * H * DH] bf16 (row = token; q | k | v column blocks, head h at columns h * DH),
//         o   [Bn * N, H * DH] bf16, lse [Bn, H, N] fp32 (natural log of the softmax denominator, scaled scores).
//
// Forward: one CTA (128 threads) per (sample, head, 128-query tile).  Thread t owns query row t == TMEM lane t, so
// row max / row sum need no shuffles.  Per 64-key tile: S = Q K^T (tcgen05.mma, fp32 in TMEM) -> the row's 64 scores
// in registers (one tcgen05.ld pass; ragged last tiles: two masked passes) -> max, exp2 + sum -> P (bf16) written to shared memory in the K-major 128B-swizzle
// layout -> O_tile = P V (V consumed as an MN-major operand straight from its TMA tile) -> registers accumulate
// O with the online-softmax rescale.  K/V tiles are double-buffered; the next S = Q K^T is issued before the
// current P V result is consumed so the tensor pipe overlaps the exp work.  S (64 columns) + O (<= 64) take 128
// TMEM columns and ~40-64 KB of shared memory, so 3-4 CTAs share an SM and hide each other's MMA / TMEM latencies
// (measured: dh = 32, N = 1024: 303 us with 128-key tiles and 2 CTAs per SM -> 234 us; MUFU.EX2 bound 115 us).
#include "common.cuh"
#include "tc_common.cuh"
#include "tmap.h"

struct mb_attn_problem {     // mirror of include/maestro_b200.h
  const void* qkv;
  void* o;
  float* lse;
  const void* d_o;          // backward only
  void* dqkv;
  float* delta;
  int Bn, N;
};

namespace {

// Full key tiles read their 64 scores from TMEM once (64 registers) instead of once for the max and once for the exp;
// with the register budget pinned to 4 CTAs per SM (3 at dh = 64) the forward is 12 % faster over the step's shapes
// (profiles/r2b_attn_probe.txt: 2.33 -> 2.05 ms; without the pin 2.45 ms -- occupancy matters more than the pass).
#ifndef MB_ATTN_FWD_ONEPASS
#define MB_ATTN_FWD_ONEPASS 1   // A/B builds: tools/_diag/build_diag.sh
#endif
constexpr int BQ = 128;   // query rows per CTA
constexpr int BKV = 64;   // keys per tile: S (64) + O (<= 64) fit 128 TMEM columns -> up to 4 CTAs share an SM

template <int DH> struct ACfg {
  static constexpr int SW = DH * 2;                       // swizzle span == row bytes of a [rows, DH] bf16 tile
  static constexpr int Q_BYTES = BQ * DH * 2;             // Q tile (two 64-row TMA boxes)
  static constexpr int TILE_BYTES = BKV * DH * 2;         // one K / V tile
  static constexpr int P_BYTES = BQ * BKV * 2;            // 16 KB: one [128 x 64] K-major block
  static constexpr int OFF_Q = 0;
  static constexpr int OFF_K = Q_BYTES;                   // 2 stages
  static constexpr int OFF_V = OFF_K + 2 * TILE_BYTES;    // 2 stages
  static constexpr int OFF_P = OFF_V + 2 * TILE_BYTES;
  static constexpr int OFF_BAR = OFF_P + P_BYTES;
  static constexpr int SMEM_BYTES = OFF_BAR + 128 + 1024;  // + barriers + alignment slack
  static constexpr uint32_t LT = tc::layout_type_for(SW);
};

struct AttnArgs {
  bf16* o;
  float* lse;
  int N, H;
  float scale_log2;  // dh^-0.5 * log2(e)
  float scale;
};

// Grouped launch: up to ATTN_MAX_GROUP problems of one head shape (the per-group stacks of a layer: same H and dh,
// different sequence lengths / batch) share one grid; CTA ranges [cta_end[p - 1], cta_end[p]) belong to problem p.
constexpr int ATTN_MAX_GROUP = 4;
struct AttnGroup {
  CUtensorMap tm[ATTN_MAX_GROUP];
  AttnArgs a[ATTN_MAX_GROUP];
  int cta_end[ATTN_MAX_GROUP];
  int count;
};

#ifndef MB_ATTN_FWD_MINB
#define MB_ATTN_FWD_MINB 4
#endif
template <int DH>
__global__ void __launch_bounds__(128, (MB_ATTN_FWD_MINB > 1 && DH == 64) ? MB_ATTN_FWD_MINB - 1 : MB_ATTN_FWD_MINB) attn_fwd_kernel(const __grid_constant__ AttnGroup grp) {
  using C_ = ACfg<DH>;
  int pi = 0, cta0 = 0;
#pragma unroll
  for (int i = 0; i < ATTN_MAX_GROUP - 1; ++i)
    if (i + 1 < grp.count && static_cast<int>(blockIdx.x) >= grp.cta_end[i]) { pi = i + 1; cta0 = grp.cta_end[i]; }
  const CUtensorMap& tm = grp.tm[pi];
  const AttnArgs& args = grp.a[pi];
  const int bid = static_cast<int>(blockIdx.x) - cta0;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  uint64_t* bar_q = reinterpret_cast<uint64_t*>(smem + C_::OFF_BAR);
  uint64_t* bar_kv = bar_q + 1;   // [2]
  uint64_t* bar_s = bar_q + 3;
  uint64_t* bar_o = bar_q + 4;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar_q + 5);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = args.N, H = args.H;
  const int q_tiles = (N + BQ - 1) / BQ;
  const int qt = bid % q_tiles;
  const int bh = bid / q_tiles;
  const int h = bh % H, b = bh / H;
  const int inner = H * DH;
  const int kv_tiles = (N + BKV - 1) / BKV;

  if (tid == 0) {
    tc::prefetch_tmap(&tm);
    tc::mbar_init(bar_q, 1);
    tc::mbar_init(&bar_kv[0], 1);
    tc::mbar_init(&bar_kv[1], 1);
    tc::mbar_init(bar_s, 1);
    tc::mbar_init(bar_o, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) {
    tc::tmem_alloc(tmem_ptr, 128);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_launch();
  pdl_wait();
  const uint32_t tm_s = tmem_base;          // S: columns [0, 64)
  const uint32_t tm_o = tmem_base + BKV;    // O tile: columns [64, 64 + DH)
  const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;

  constexpr uint32_t idesc_s = tc::make_idesc_bf16(BQ, BKV, 0, 0);   // Q (K-major) x K (K-major)
  constexpr uint32_t idesc_o = tc::make_idesc_bf16(BQ, DH, 0, 1);    // P (K-major) x V (MN-major)
  const uint32_t sQ = tc::smem_u32(smem + C_::OFF_Q);
  const uint32_t sP = tc::smem_u32(smem + C_::OFF_P);

  auto load_kv = [&](int j) {
    const int st = j & 1;
    tc::mbar_expect_tx(&bar_kv[st], 2 * C_::TILE_BYTES);
    tc::tma_load_3d(smem + C_::OFF_K + st * C_::TILE_BYTES, &tm, &bar_kv[st], inner + h * DH, j * BKV, b);
    tc::tma_load_3d(smem + C_::OFF_V + st * C_::TILE_BYTES, &tm, &bar_kv[st], 2 * inner + h * DH, j * BKV, b);
  };
  auto issue_s = [&](int j) {
    const uint32_t sK = tc::smem_u32(smem + C_::OFF_K + (j & 1) * C_::TILE_BYTES);
#pragma unroll
    for (int k = 0; k < DH / 16; ++k) {
      const uint64_t da = tc::make_smem_desc(sQ + k * 32, 16, 8 * C_::SW, C_::LT);
      const uint64_t db = tc::make_smem_desc(sK + k * 32, 16, 8 * C_::SW, C_::LT);
      tc::umma_bf16(tm_s, da, db, idesc_s, k > 0 ? 1u : 0u);
    }
    tc::umma_commit(bar_s);
  };
  auto issue_pv = [&](int j) {
    const uint32_t sV = tc::smem_u32(smem + C_::OFF_V + (j & 1) * C_::TILE_BYTES);
#pragma unroll
    for (int k = 0; k < BKV / 16; ++k) {
      const uint64_t da = tc::make_smem_desc(sP + k * 32, 16, 1024, tc::SWIZZLE_128B);
      const uint64_t db = tc::make_smem_desc(sV + k * 16 * C_::SW, 16, 8 * C_::SW, C_::LT);
      tc::umma_bf16(tm_o, da, db, idesc_o, k > 0 ? 1u : 0u);
    }
    tc::umma_commit(bar_o);
  };

  if (tid == 0) {
    tc::mbar_expect_tx(bar_q, C_::Q_BYTES);
    tc::tma_load_3d(smem + C_::OFF_Q, &tm, bar_q, h * DH, qt * BQ, b);
    tc::tma_load_3d(smem + C_::OFF_Q + C_::Q_BYTES / 2, &tm, bar_q, h * DH, qt * BQ + 64, b);
    load_kv(0);
    if (kv_tiles > 1) load_kv(1);
    tc::mbar_wait(bar_q, 0);
    tc::mbar_wait(&bar_kv[0], 0);
    tc::tc_fence_after();
    issue_s(0);
  }

  float o_acc[DH];
#pragma unroll
  for (int i = 0; i < DH; ++i) o_acc[i] = 0.f;
  float m_run = -INFINITY, l_run = 0.f;
  const float c = args.scale_log2;
  uint8_t* p_row = smem + C_::OFF_P + tid * 128;

  for (int j = 0; j < kv_tiles; ++j) {
    const uint32_t ph = j & 1;
    const int kv_valid = min(BKV, N - j * BKV);
    tc::mbar_wait(bar_s, ph);
    tc::tc_fence_after();
    float mx, alpha, mc, lsum;
#if MB_ATTN_FWD_ONEPASS
    if (kv_valid == BKV) {
      // full tile: the 64 scores of the row are read from TMEM ONCE (64 registers), then max -> exp2 -> P
      uint32_t r[64];
      tc::tmem_ld_32x32(tm_s + lane_base, r);
      tc::tmem_ld_32x32(tm_s + lane_base + 32, r + 32);
      tc::tmem_ld_wait();
      float m4[4] = {__uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3])};
#pragma unroll
      for (int i = 4; i < 64; i += 4) {
        m4[0] = fmaxf(m4[0], __uint_as_float(r[i]));
        m4[1] = fmaxf(m4[1], __uint_as_float(r[i + 1]));
        m4[2] = fmaxf(m4[2], __uint_as_float(r[i + 2]));
        m4[3] = fmaxf(m4[3], __uint_as_float(r[i + 3]));
      }
      mx = fmaxf(fmaxf(m_run, fmaxf(m4[0], m4[1])), fmaxf(m4[2], m4[3]));
      alpha = exp2f((m_run - mx) * c);
      mc = mx * c;
      float ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c16 = 0; c16 < 8; ++c16) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {   // scalar on purpose: the packed form (FFMA2 / FADD2) needs aligned register pairs
          // and spills at the 128-register budget that keeps 4 CTAs per SM (measured 2.06 -> 2.20 ms per step)
          const float p0 = fast_exp2(fmaf(__uint_as_float(r[c16 * 8 + 2 * i]), c, -mc));
          const float p1 = fast_exp2(fmaf(__uint_as_float(r[c16 * 8 + 2 * i + 1]), c, -mc));
          ls[i] += p0 + p1;
          __nv_bfloat162 hb = __floats2bfloat162_rn(p0, p1);
          w[i] = *reinterpret_cast<uint32_t*>(&hb);
        }
        *reinterpret_cast<uint4*>(p_row + ((c16 ^ (tid & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
      }
      lsum = (ls[0] + ls[1]) + (ls[2] + ls[3]);
    } else
#endif
    {
    // pass 1: row max over the valid keys (mask-free for chunks that lie entirely inside the sequence)
    mx = m_run;
#pragma unroll 1
    for (int ch = 0; ch < BKV / 32; ++ch) {
      if (ch * 32 >= kv_valid) break;
      uint32_t r[32];
      tc::tmem_ld_32x32(tm_s + lane_base + ch * 32, r);
      tc::tmem_ld_wait();
      if (ch * 32 + 32 <= kv_valid) {
        float m0 = __uint_as_float(r[0]), m1 = __uint_as_float(r[1]);
#pragma unroll
        for (int i = 2; i < 32; i += 2) {
          m0 = fmaxf(m0, __uint_as_float(r[i]));
          m1 = fmaxf(m1, __uint_as_float(r[i + 1]));
        }
        mx = fmaxf(mx, fmaxf(m0, m1));
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i)
          if (ch * 32 + i < kv_valid) mx = fmaxf(mx, __uint_as_float(r[i]));
      }
    }
    alpha = exp2f((m_run - mx) * c);   // m_run = -inf on the first tile -> 0
    mc = mx * c;
    lsum = 0.f;
    // pass 2: p = exp2(s * c - m * c), write bf16 P row (K-major, 128B swizzle), accumulate the row sum
    if (kv_valid == BKV) {
      // full tile: no masks; 4 independent partial sums keep the FADD chain short
      float ls[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 1
      for (int ch = 0; ch < BKV / 32; ++ch) {
        uint32_t r[32];
        tc::tmem_ld_32x32(tm_s + lane_base + ch * 32, r);
        tc::tmem_ld_wait();
#pragma unroll
        for (int v8 = 0; v8 < 4; ++v8) {
          uint32_t w[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float p0 = fast_exp2(fmaf(__uint_as_float(r[v8 * 8 + 2 * i]), c, -mc));
            const float p1 = fast_exp2(fmaf(__uint_as_float(r[v8 * 8 + 2 * i + 1]), c, -mc));
            ls[i] += p0 + p1;
            __nv_bfloat162 hb = __floats2bfloat162_rn(p0, p1);
            w[i] = *reinterpret_cast<uint32_t*>(&hb);
          }
          const int c16 = ch * 4 + v8;                       // 16-byte chunk of the 128-byte P row
          *reinterpret_cast<uint4*>(p_row + ((c16 ^ (tid & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
      lsum = (ls[0] + ls[1]) + (ls[2] + ls[3]);
    } else {
#pragma unroll 1
      for (int ch = 0; ch < BKV / 32; ++ch) {
        uint32_t r[32];
        if (ch * 32 < kv_valid) {
          tc::tmem_ld_32x32(tm_s + lane_base + ch * 32, r);
          tc::tmem_ld_wait();
        }
#pragma unroll
        for (int v8 = 0; v8 < 4; ++v8) {
          uint32_t w[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int col = ch * 32 + v8 * 8 + 2 * i;
            float p0 = 0.f, p1 = 0.f;
            if (col < kv_valid) p0 = fast_exp2(fmaf(__uint_as_float(r[v8 * 8 + 2 * i]), c, -mc));
            if (col + 1 < kv_valid) p1 = fast_exp2(fmaf(__uint_as_float(r[v8 * 8 + 2 * i + 1]), c, -mc));
            lsum += p0 + p1;
            __nv_bfloat162 hb = __floats2bfloat162_rn(p0, p1);
            w[i] = *reinterpret_cast<uint32_t*>(&hb);
          }
          const int c16 = ch * 4 + v8;
          *reinterpret_cast<uint4*>(p_row + ((c16 ^ (tid & 7)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
        }
      }
    }
    }
    l_run = l_run * alpha + lsum;
    m_run = mx;
    tc::fence_proxy_async();     // P (generic-proxy stores) -> visible to the tensor core (async proxy)
    tc::tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      tc::tc_fence_after();
      issue_pv(j);
      if (j + 1 < kv_tiles) {
        tc::mbar_wait(&bar_kv[(j + 1) & 1], ((j + 1) >> 1) & 1);
        tc::tc_fence_after();
        issue_s(j + 1);
      }
    }
    tc::mbar_wait(bar_o, ph);
    tc::tc_fence_after();
#pragma unroll
    for (int ch = 0; ch < (DH + 31) / 32; ++ch) {
      uint32_t r[32];
      tc::tmem_ld_32x32(tm_o + lane_base + ch * 32, r);   // DH = 16 reads 16 junk columns of the allocation
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 32; ++i)
        if (ch * 32 + i < DH) o_acc[ch * 32 + i] = fmaf(o_acc[ch * 32 + i], alpha, __uint_as_float(r[i]));
    }
    // K/V stage (j & 1) and the P buffer are free again: prefetch tile j + 2
    if (tid == 0 && j + 2 < kv_tiles) load_kv(j + 2);
  }

  const int row = qt * BQ + tid;
  if (row < N) {
    const float inv = 1.f / l_run;
    bf16* orow = args.o + (static_cast<size_t>(b) * N + row) * inner + h * DH;
#pragma unroll
    for (int v8 = 0; v8 < DH / 8; ++v8) {
      float v[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = o_acc[v8 * 8 + i] * inv;
      VecIO<bf16>::store(orow + v8 * 8, v);
    }
    args.lse[(static_cast<size_t>(b) * H + h) * N + row] = m_run * args.scale + logf(l_run);
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, 128);
}

int qkv_tmap(CUtensorMap* tm, const void* qkv, int Bn, int N, int cols, int DH, int box_rows) {
  const uint64_t dims[3] = {static_cast<uint64_t>(cols), static_cast<uint64_t>(N), static_cast<uint64_t>(Bn)};
  const uint64_t strides[2] = {static_cast<uint64_t>(cols), static_cast<uint64_t>(N) * cols};
  const uint32_t box[3] = {static_cast<uint32_t>(DH), static_cast<uint32_t>(box_rows), 1};
  return mb_get_tmap(tm, qkv, 3, dims, strides, box, DH * 2);
}

template <int DH>
int launch_fwd(const mb_attn_problem* ps, int count, int H, cudaStream_t st) {
  using C_ = ACfg<DH>;
  AttnGroup g;
  int total = 0;
  for (int i = 0; i < count; ++i) {
    int rc = qkv_tmap(&g.tm[i], ps[i].qkv, ps[i].Bn, ps[i].N, 3 * H * DH, DH, BKV);
    if (rc != 0) return rc;
    AttnArgs& a = g.a[i];
    a.o = static_cast<bf16*>(ps[i].o);
    a.lse = ps[i].lse;
    a.N = ps[i].N;
    a.H = H;
    a.scale = 1.0f / sqrtf(static_cast<float>(DH));
    a.scale_log2 = a.scale * 1.4426950408889634f;
    total += ps[i].Bn * H * ((ps[i].N + BQ - 1) / BQ);
    g.cta_end[i] = total;
  }
  for (int i = count; i < ATTN_MAX_GROUP; ++i) { g.tm[i] = g.tm[0]; g.a[i] = g.a[0]; g.cta_end[i] = total; }
  g.count = count;
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(attn_fwd_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, C_::SMEM_BYTES);
    attr = true;
  }
  mb_launch(attn_fwd_kernel<DH>, dim3(total), dim3(128), C_::SMEM_BYTES, st, g);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : static_cast<int>(e);
}

}  // namespace

// `count` (1..4) attention problems of one head shape as ONE launch (the longest sequences first gives the best tail).
extern "C" int mb_attn_fwd_grouped(const mb_attn_problem* problems, int count, int H, int dh, void* stream) {
  if (count < 1 || count > ATTN_MAX_GROUP) return -2;
  mb_attn_problem live[ATTN_MAX_GROUP];
  int n = 0;
  for (int i = 0; i < count; ++i)
    if (problems[i].Bn > 0 && problems[i].N > 0) live[n++] = problems[i];
  if (n == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dh) {
    case 16: return launch_fwd<16>(live, n, H, st);
    case 32: return launch_fwd<32>(live, n, H, st);
    case 64: return launch_fwd<64>(live, n, H, st);
    default: return -2;
  }
}

// qkv [Bn*N, 3*H*dh] bf16 -> o [Bn*N, H*dh] bf16, lse [Bn, H, N] fp32.  dh in {16, 32, 64}.
extern "C" int mb_attn_fwd(const void* qkv, void* o, float* lse, int Bn, int N, int H, int dh, void* stream) {
  const mb_attn_problem p{qkv, o, lse, nullptr, nullptr, nullptr, Bn, N};
  return mb_attn_fwd_grouped(&p, 1, H, dh, stream);
}
