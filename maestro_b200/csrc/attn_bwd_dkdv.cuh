// dK / dV kernel of the fused attention backward and the C entry point (included by attn_bwd_tc.cu).
#pragma once

namespace {

// Software pipeline over SUB-query half tiles j = 0, 1, ...: half j uses TMEM buffer j & 1 (S^T, dP^T), operand
// buffer j & 1 (P^T, dS^T) and Q / dO stage j % NS.  While the threads run the exp / dS pass of half j, the tensor
// pipe works on S^T / dP^T of half j + 1 (issued one pass earlier) and on the gradient MMAs of half j - 1, so neither
// the MMA latency nor the single-thread issue sequence sits between two passes.  Warps 0-3 (thread = key row = TMEM
// lane) run the passes; one lane each of warps 4 / 5 / 6 issues the score MMAs / the gradient MMAs / the TMA loads,
// driven by mbarriers only.
constexpr int DKDV_THREADS = 128 * NH + 96;   // compute warps (NH per TMEM lane quarter) + three control warps

template <int DH>
__global__ void __launch_bounds__(DKDV_THREADS) attn_bwd_dkdv_kernel(const __grid_constant__ BwdGroup grp) {
  using C_ = BCfg<DH>;
  int pi = 0, cta0 = 0;
#pragma unroll
  for (int i = 0; i < ATTN_MAX_GROUP - 1; ++i)
    if (i + 1 < grp.count && static_cast<int>(blockIdx.x) >= grp.cta_end[i]) { pi = i + 1; cta0 = grp.cta_end[i]; }
  const CUtensorMap& tm_qkv128 = grp.q128[pi];
  const CUtensorMap& tm_qsub = grp.qsub[pi];
  const CUtensorMap& tm_dosub = grp.dosub[pi];
  const BwdArgs& args = grp.a[pi];
  const int bid = static_cast<int>(blockIdx.x) - cta0;
  constexpr int SUB = C_::SUB, NS = C_::NS;
  constexpr int OFF_K = 0, OFF_V = C_::ROW_TILE, OFF_Q = 2 * C_::ROW_TILE, OFF_DO = OFF_Q + NS * C_::SUB_TILE;
  constexpr int OFF_PT = OFF_DO + NS * C_::SUB_TILE, OFF_DST = OFF_PT + 2 * C_::AS_TILE;
  constexpr int OFF_STAT = OFF_DST + 2 * C_::AS_TILE;
  constexpr int OFF_BAR = OFF_STAT + 2 * 2 * SUB * 4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  float* s_stat = reinterpret_cast<float*>(smem + OFF_STAT);   // [j & 1][lse2 | delta][SUB]
  uint64_t* bar_kv = reinterpret_cast<uint64_t*>(smem + OFF_BAR);
  uint64_t* bar_q = bar_kv + 1;        // [NS]
  uint64_t* bar_s = bar_q + NS;        // [2]
  uint64_t* bar_o = bar_s + 2;         // [2]
  uint64_t* bar_p = bar_o + 2;         // [2] operand halves P^T / dS^T written (128 arrivals)
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bar_p + 2);

  const int tid = threadIdx.x, warp = tid >> 5;
  const int N = args.N, H = args.H, inner = H * DH;
  const int r_tiles = (N + BR - 1) / BR;
  const int rt = bid % r_tiles, bh = bid / r_tiles;
  const int h = bh % H, b = bh / H;
  const int n_sub = (N + SUB - 1) / SUB;

  if (tid == 0) {
    tc::mbar_init(bar_kv, 1);
    for (int i = 0; i < NS; ++i) tc::mbar_init(&bar_q[i], 1);
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&bar_s[i], 1);
      tc::mbar_init(&bar_o[i], 1);
      tc::mbar_init(&bar_p[i], BR * NH);
    }
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
  pdl_launch();
  pdl_wait();
  // columns: S^T [2][SUB] | dP^T [2][SUB] | dK [DH] | dV [DH]
  const uint32_t tm_st = tmem_base, tm_dpt = tmem_base + 2 * SUB, tm_dk = tmem_base + 4 * SUB, tm_dv = tm_dk + DH;
  const uint32_t lane_base = static_cast<uint32_t>((warp & 3) * 32) << 16;
  const int krow = tid & 127, half = tid >> 7;       // key row of the tile / column half owned by this thread
  constexpr uint32_t idesc_s = tc::make_idesc_bf16(BR, SUB, 0, 0);   // [128 keys] x [SUB queries], K = DH
  constexpr uint32_t idesc_g = tc::make_idesc_bf16(BR, DH, 0, 1);    // [128 keys] x [DH], K = SUB queries (B MN-major)
  const uint32_t sK = tc::smem_u32(smem + OFF_K), sV = tc::smem_u32(smem + OFF_V);
  const uint32_t sPT = tc::smem_u32(smem + OFF_PT), sDST = tc::smem_u32(smem + OFF_DST);
  const uint32_t sQ0 = tc::smem_u32(smem + OFF_Q), sDO0 = tc::smem_u32(smem + OFF_DO);

  // the lambdas below run on thread 0 only; `stage` = j % NS is tracked by the caller
  auto load_q = [&](int j, int stage) {
    tc::mbar_expect_tx(&bar_q[stage], 2 * C_::SUB_TILE);
    tc::tma_load_3d(smem + OFF_Q + stage * C_::SUB_TILE, &tm_qsub, &bar_q[stage], h * DH, j * SUB, b);
    tc::tma_load_3d(smem + OFF_DO + stage * C_::SUB_TILE, &tm_dosub, &bar_q[stage], h * DH, j * SUB, b);
  };
  auto issue_s = [&](int j, int stage) {
    const uint32_t sQ = sQ0 + stage * C_::SUB_TILE, sDO = sDO0 + stage * C_::SUB_TILE;
    const uint32_t buf = static_cast<uint32_t>(j & 1) * SUB;
#pragma unroll
    for (int k = 0; k < DH / 16; ++k)
      tc::umma_bf16(tm_st + buf, tc::make_smem_desc(sK + k * 32, 16, 8 * C_::SW, C_::LT),
                    tc::make_smem_desc(sQ + k * 32, 16, 8 * C_::SW, C_::LT), idesc_s, k > 0);
#pragma unroll
    for (int k = 0; k < DH / 16; ++k)
      tc::umma_bf16(tm_dpt + buf, tc::make_smem_desc(sV + k * 32, 16, 8 * C_::SW, C_::LT),
                    tc::make_smem_desc(sDO + k * 32, 16, 8 * C_::SW, C_::LT), idesc_s, k > 0);
    tc::umma_commit(&bar_s[j & 1]);
  };
  auto issue_grads = [&](int j, int stage) {
    const uint32_t sQ = sQ0 + stage * C_::SUB_TILE, sDO = sDO0 + stage * C_::SUB_TILE;
    const uint32_t aoff = static_cast<uint32_t>(j & 1) * C_::AS_TILE;
#pragma unroll
    for (int k = 0; k < SUB / 16; ++k)
      tc::umma_bf16(tm_dv, tc::make_smem_desc(sPT + aoff + k * 32, 16, 8 * C_::AS_ROW, C_::AS_LT),
                    tc::make_smem_desc(sDO + k * 16 * C_::SW, 16, 8 * C_::SW, C_::LT), idesc_g, (j > 0 || k > 0));
#pragma unroll
    for (int k = 0; k < SUB / 16; ++k)
      tc::umma_bf16(tm_dk, tc::make_smem_desc(sDST + aoff + k * 32, 16, 8 * C_::AS_ROW, C_::AS_LT),
                    tc::make_smem_desc(sQ + k * 16 * C_::SW, 16, 8 * C_::SW, C_::LT), idesc_g, (j > 0 || k > 0));
    tc::umma_commit(&bar_o[j & 1]);
  };

  if (warp >= BWD_CTRL_WARP) {
    // ===================== three control lanes, one per warp, driven by mbarriers only =====================
    // The per-pass timeline (profiles/r2b_attn_timeline.txt) showed ONE control lane as the bottleneck of the kernel:
    // ~850 cycles to issue the 6 small MMAs of a pass (descriptor arithmetic + tcgen05.mma + commits from a single
    // thread) + ~300 for the loads, 1565 cycles per pass in total, while the compute warps worked ~450 and waited the
    // rest.  The three jobs are independent: warp +0 issues S^T / dP^T, warp +1 the gradient MMAs, warp +2 the TMA loads.
    if (tid == 128 * NH + 64) {
      // ---- loader: K / V once, then the Q / dO ring (a stage is free when the gradient MMAs that read it are done)
      int l_stage = 0, l_next = 0;
      tc::mbar_expect_tx(bar_kv, 2 * C_::ROW_TILE);
      tc::tma_load_3d(smem + OFF_K, &tm_qkv128, bar_kv, inner + h * DH, rt * BR, b);
      tc::tma_load_3d(smem + OFF_V, &tm_qkv128, bar_kv, 2 * inner + h * DH, rt * BR, b);
      for (; l_next < NS && l_next < n_sub; ++l_next) {      // fill the ring
        load_q(l_next, l_stage);
        if (++l_stage == NS) l_stage = 0;
      }
      for (int j = 1; l_next < n_sub; ++j) {                 // stage (j - 1) % NS: read last by the gradient MMAs of half j - 1
        tc::mbar_wait(&bar_o[(j - 1) & 1], ((j - 1) >> 1) & 1);
        load_q(l_next, l_stage);
        ++l_next;
        if (++l_stage == NS) l_stage = 0;
      }
    } else if (tid == 128 * NH) {
      // ---- S^T / dP^T issuer: half j + 2 goes into the TMEM buffer pass j has just drained
      int s_stage = 0;
      uint32_t s_par = 0;
      tc::mbar_wait(bar_kv, 0);
      for (int j = 0; j < n_sub; ++j) {
        if (j >= 2) tc::mbar_wait(&bar_p[j & 1], ((j - 2) >> 1) & 1);   // pass j - 2 has read buffer j & 1
        tc::mbar_wait(&bar_q[s_stage], s_par);
        tc::tc_fence_after();
        MB_ASTAMP(j, 0);
        issue_s(j, s_stage);
        MB_ASTAMP(j, 1);
        if (++s_stage == NS) { s_stage = 0; s_par ^= 1; }
      }
    } else if (tid == 128 * NH + 32) {
      // ---- gradient issuer: dV += P^T dO, dK += dS^T Q of half j once its operands are in shared memory
      int g_stage = 0;
      for (int j = 0; j < n_sub; ++j) {
        tc::mbar_wait(&bar_p[j & 1], (j >> 1) & 1);
        tc::tc_fence_after();
        MB_ASTAMP(j, 2);
        issue_grads(j, g_stage);
        MB_ASTAMP(j, 3);
        if (++g_stage == NS) g_stage = 0;
      }
    }
  } else {
  // ===================== compute warps =====================
  const int row = rt * BR + krow;          // key row owned by this thread (shared by the NH threads of the row)
  const bool row_ok = row < N;
  const float c = args.scale_log2, sc = args.scale;
  const size_t stat_base = (static_cast<size_t>(b) * H + h) * N;

  // per-query statistics (log2-domain lse, delta * scale) of half j live in s_stat[j & 1]; the global loads of half
  // j + 1 are issued at the top of pass j and land in shared memory at its end, so no pass waits on HBM
  float n_lse = 0.f, n_delta = 0.f;
  auto fetch_stats = [&](int j) {
    const int q = j * SUB + tid;
    const bool ok = tid < SUB && q < N;
    n_lse = ok ? args.lse[stat_base + q] : 0.f;               // raw: nothing may wait on these loads before put_stats
    n_delta = ok ? args.delta[stat_base + q] : 0.f;
  };
  auto put_stats = [&](int j) {
    if (tid < SUB) {
      float* d = s_stat + (j & 1) * 2 * SUB;
      d[tid] = n_lse * -1.4426950408889634f;                  // both NEGATED: they are the addends of the packed FMAs
      d[SUB + tid] = n_delta * -sc;                           // delta pre-multiplied by the score scale
    }
  };
  fetch_stats(0);
  put_stats(0);
  asm volatile("bar.sync 1, %0;" ::"n"(128 * NH) : "memory");
  constexpr int CW = SUB / NH;               // columns of a pass handled by this thread
  for (int j = 0; j < n_sub; ++j) {
    const int bufi = j & 1;
    const uint32_t st = tc::smem_u32(s_stat + bufi * 2 * SUB);
    const uint32_t pt_row = sPT + bufi * C_::AS_TILE + krow * C_::AS_ROW;
    const uint32_t dst_row = sDST + bufi * C_::AS_TILE + krow * C_::AS_ROW;
    if (j + 1 < n_sub) fetch_stats(j + 1);
    float4 lsv[CW / 4], dsv[CW / 4];                          // this thread's statistics: in registers before S^T lands
#pragma unroll
    for (int q4 = 0; q4 < CW / 4; ++q4) {
      lsv[q4] = lds_f4(st + (half * CW + q4 * 4) * 4);
      dsv[q4] = lds_f4(st + (SUB + half * CW + q4 * 4) * 4);
    }
    if (tid == 0) MB_ASTAMP(j, 4);
    tc::mbar_wait(&bar_s[bufi], (j >> 1) & 1);
    if (tid == 0) MB_ASTAMP(j, 5);
    if (j >= 2) tc::mbar_wait(&bar_o[bufi], ((j - 2) >> 1) & 1);   // gradient MMAs of half j - 2 have read `bufi`
    tc::tc_fence_after();
    if (tid == 0) MB_ASTAMP(j, 6);
    // P^T and dS^T.  No masks: queries past N are zero rows of Q and dO (TMA zero fill), so their columns multiply
    // zero rows in dV += P^T dO and dK += dS^T Q; key rows past N are never stored.  st[] holds finite values there.
    {
      uint32_t s[CW], dp[CW];
      tc::tmem_ld_32xN<CW>(tm_st + lane_base + bufi * SUB + half * CW, s);
      tc::tmem_ld_32xN<CW>(tm_dpt + lane_base + bufi * SUB + half * CW, dp);
      tc::tmem_ld_wait();
#pragma unroll
      for (int v8 = 0; v8 < CW / 8; ++v8) {
        uint32_t wp[4], wd[4];
        const float4 l0 = lsv[2 * v8], l1 = lsv[2 * v8 + 1], e0 = dsv[2 * v8], e1 = dsv[2 * v8 + 1];
        const float ls[8] = {l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, l1.z, l1.w};
        const float ds[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
#pragma unroll
        for (int k2 = 0; k2 < 4; ++k2) {   // packed fp32 (FFMA2 / FMUL2)
          const int idx = v8 * 8 + 2 * k2;
          const float2 x = fma2(make_float2(__uint_as_float(s[idx]), __uint_as_float(s[idx + 1])), splat2(c),
                                make_float2(ls[2 * k2], ls[2 * k2 + 1]));
          const float2 g = fma2(make_float2(__uint_as_float(dp[idx]), __uint_as_float(dp[idx + 1])), splat2(sc),
                                make_float2(ds[2 * k2], ds[2 * k2 + 1]));
          const float2 p2v = make_float2(fast_exp2(x.x), fast_exp2(x.y));
          const float2 d2v = mul2(p2v, g);
          const float p2[2] = {p2v.x, p2v.y}, d2[2] = {d2v.x, d2v.y};
          __nv_bfloat162 hp = __floats2bfloat162_rn(p2[0], p2[1]);
          __nv_bfloat162 hd = __floats2bfloat162_rn(d2[0], d2[1]);
          wp[k2] = *reinterpret_cast<uint32_t*>(&hp);
          wd[k2] = *reinterpret_cast<uint32_t*>(&hd);
        }
        store_a_chunk_s<SUB>(pt_row, krow, half * (CW / 8) + v8, wp);
        store_a_chunk_s<SUB>(dst_row, krow, half * (CW / 8) + v8, wd);
      }
    }
    if (j + 1 < n_sub) put_stats(j + 1);
    tc::fence_proxy_async();
    tc::tc_fence_before();
    tc::mbar_arrive(&bar_p[bufi]);
    if (tid == 0) MB_ASTAMP(j, 7);
    asm volatile("bar.sync 1, %0;" ::"n"(128 * NH) : "memory");   // next half's statistics are visible to the compute warps
  }
  tc::mbar_wait(&bar_o[(n_sub - 1) & 1], ((n_sub - 1) >> 1) & 1);   // a commit covers every MMA issued before it
  tc::tc_fence_after();
  bf16* out = args.dqkv + (static_cast<size_t>(b) * N + (row_ok ? row : 0)) * 3 * inner + h * DH;
  if constexpr (NH == 1) {
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
  } else {
    constexpr int OW = DH / NH;                // output columns of dK and of dV written by this thread
#pragma unroll
    for (int which = 0; which < 2; ++which) {   // 0: dK, 1: dV
      uint32_t r[OW];
      tc::tmem_ld_32xN<OW>((which == 0 ? tm_dk : tm_dv) + lane_base + half * OW, r);
      tc::tmem_ld_wait();
#pragma unroll
      for (int v8 = 0; v8 < OW / 8; ++v8) {
        if (row_ok) {
          float v[8];
#pragma unroll
          for (int k2 = 0; k2 < 8; ++k2) v[k2] = __uint_as_float(r[v8 * 8 + k2]);
          VecIO<bf16>::store(out + (which + 1) * inner + half * OW + v8 * 8, v);
        }
      }
    }
  }
  }  // compute warps
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
int launch_bwd(const mb_attn_problem* ps, int count, int H, cudaStream_t st) {
  using C_ = BCfg<DH>;
  BwdGroup g;
  int total = 0, rc;
  for (int i = 0; i < count; ++i) {
    const int Bn = ps[i].Bn, N = ps[i].N;
    if ((rc = tmap3(&g.q128[i], ps[i].qkv, Bn, N, 3 * H * DH, DH, BR)) != 0) return rc;
    if ((rc = tmap3(&g.do128[i], ps[i].d_o, Bn, N, H * DH, DH, BR)) != 0) return rc;
    if ((rc = tmap3(&g.qsub[i], ps[i].qkv, Bn, N, 3 * H * DH, DH, C_::SUB)) != 0) return rc;
    if ((rc = tmap3(&g.dosub[i], ps[i].d_o, Bn, N, H * DH, DH, C_::SUB)) != 0) return rc;
    BwdArgs& a = g.a[i];
    a.o = static_cast<const bf16*>(ps[i].o);
    a.d_o = static_cast<const bf16*>(ps[i].d_o);
    a.lse = ps[i].lse;
    a.delta = ps[i].delta;
    a.dqkv = static_cast<bf16*>(ps[i].dqkv);
    a.N = N;
    a.H = H;
    a.scale = 1.0f / sqrtf(static_cast<float>(DH));
    a.scale_log2 = a.scale * 1.4426950408889634f;
    total += Bn * H * ((N + BR - 1) / BR);
    g.cta_end[i] = total;
  }
  for (int i = count; i < ATTN_MAX_GROUP; ++i) {
    g.q128[i] = g.q128[0]; g.do128[i] = g.do128[0]; g.qsub[i] = g.qsub[0]; g.dosub[i] = g.dosub[0];
    g.a[i] = g.a[0]; g.cta_end[i] = total;
  }
  g.count = count;
  constexpr int SMEM_DQ = 2 * C_::ROW_TILE + 2 * C_::NS * C_::SUB_TILE + 2 * C_::AS_TILE + 128 + 1024;
  constexpr int SMEM_DKDV = 2 * C_::ROW_TILE + 2 * C_::NS * C_::SUB_TILE + 4 * C_::AS_TILE + 2 * 2 * C_::SUB * 4 + 128 + 1024;
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(attn_bwd_dq_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DQ);
    cudaFuncSetAttribute(attn_bwd_dkdv_kernel<DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_DKDV);
    attr = true;
  }
  mb_launch(attn_bwd_dq_kernel<DH>, dim3(total), dim3(DQ_THREADS), SMEM_DQ, st, g);
  mb_launch(attn_bwd_dkdv_kernel<DH>, dim3(total), dim3(DKDV_THREADS), SMEM_DKDV, st, g);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? 0 : static_cast<int>(e);
}

}  // namespace

// Gradient of mb_attn_fwd_grouped: per problem dqkv [Bn*N, 3*H*dh] bf16 from d_o [Bn*N, H*dh]; delta [Bn, H, N] fp32 is
// scratch.  Two launches (dQ, then dK / dV) for the whole group.
extern "C" int mb_attn_bwd_grouped(const mb_attn_problem* problems, int count, int H, int dh, void* stream) {
  if (count < 1 || count > ATTN_MAX_GROUP) return -2;
  mb_attn_problem live[ATTN_MAX_GROUP];
  int n = 0;
  for (int i = 0; i < count; ++i)
    if (problems[i].Bn > 0 && problems[i].N > 0) live[n++] = problems[i];
  if (n == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dh) {
    case 16: return launch_bwd<16>(live, n, H, st);
    case 32: return launch_bwd<32>(live, n, H, st);
    case 64: return launch_bwd<64>(live, n, H, st);
    default: return -2;
  }
}

extern "C" int mb_attn_bwd(const void* qkv, const void* o, const void* d_o, const float* lse, void* dqkv, float* delta,
                           int Bn, int N, int H, int dh, void* stream) {
  const mb_attn_problem p{qkv, const_cast<void*>(o), const_cast<float*>(lse), d_o, dqkv, delta, Bn, N};
  return mb_attn_bwd_grouped(&p, 1, H, dh, stream);
}
