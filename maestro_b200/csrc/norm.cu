// Normalisation kernels (bandwidth-bound; warp-shuffle reductions, 128-bit access):
//   * LayerNorm forward / backward with fused residual-gradient add and row remapping (the remap folds the
//     reference's torch.cat / split around encoder_inter, maestro/ssl/mim.py:406-423, into the LN pass);
//   * GroupNorm(1, E) of the patch embedding (maestro/layers/embed.py:55,61): statistics over all E*g*g
//     values of one (b, d) image, apply fused with the positional + date encoding add
//     (maestro/ssl/mim.py:232-252) and the group concat (layers/utils.py:12-47).
#include "common.cuh"
#include "tc_common.cuh"

struct RowMap {
  int n;    // logical rows per sample
  int bs;   // physical rows per sample
  int off;  // physical row offset inside a sample
  __device__ __forceinline__ size_t phys(int r) const { return static_cast<size_t>(r / n) * bs + off + (r % n); }
};


// ------------------------------------------------------------------ LayerNorm forward
// y = (x - mean) * rstd * gamma + beta per row (nn.LayerNorm, eps inside the square root; mean / rstd saved for the
// backward).  HBM-bound: 2 * d * sizeof(T) bytes per row -- but only if the instruction stream lets it be: the r1
// kernel issued 285 warp instructions per 512-wide row (63 % issue-active at 0.52 of the copy bandwidth, 3 CTAs per SM
// at 80 registers).  This one: packed fp32 math (FFMA2 / FADD2 / FMUL2: two elements per instruction), gamma / beta in
// shared memory instead of 2 x NE registers, identity row maps without the integer divisions, the (still packed) loads
// of the next TWO rows of every warp in flight while a row is reduced and written, and a grid of exactly the resident
// CTAs (no partial wave).  Warp w of the grid walks rows w, w + W, w + 2W, ...; a lane owns NCH 16-byte column chunks.
// Grouped: up to LN_MAX_GROUP problems of one width (the per-group stacks of one layer, which step in lockstep) share a
// launch; CTAs are split between the problems in proportion to their rows (`cta_end` prefix), every CTA works on one
// problem exactly as the single-problem kernel would.  One launch of 15 040 rows instead of four of 1 600 - 8 192.
constexpr int LNF_THREADS = 256;
constexpr int LN_MAX_GROUP = 4;

struct LnFwdProblem {
  const void* x;
  const float* gamma;
  const float* beta;
  void* y;
  float* mean;
  float* rstd;
  int rows;
  RowMap in_map, out_map;
  int cta_end;            // exclusive prefix of the CTAs given to problems 0 .. this one
};
struct LnFwdArgs {
  LnFwdProblem p[LN_MAX_GROUP];
  int count, d;
  float eps;
};

// EXACT: see layernorm_bwd_kernel (d == NCH * 32 * V known at compile time, no lane out of range)
template <typename T, int NCH, bool MAPPED, bool EXACT = false>
__global__ void __launch_bounds__(LNF_THREADS, (NCH <= 2 ? 4 : (NCH <= 3 ? 3 : 2)))
layernorm_fwd_kernel(const __grid_constant__ LnFwdArgs args) {
  pdl_launch();
  pdl_wait();
  int pi = 0, cta0 = 0;
#pragma unroll
  for (int i = 0; i < LN_MAX_GROUP - 1; ++i)
    if (i + 1 < args.count && static_cast<int>(blockIdx.x) >= args.p[i].cta_end) { pi = i + 1; cta0 = args.p[i].cta_end; }
  const LnFwdProblem& P = args.p[pi];
  const T* __restrict__ x = static_cast<const T*>(P.x);
  const float* __restrict__ gamma = P.gamma;
  const float* __restrict__ beta = P.beta;
  T* __restrict__ y = static_cast<T*>(P.y);
  float* __restrict__ mean_out = P.mean;
  float* __restrict__ rstd_out = P.rstd;
  constexpr int V = VecIO<T>::N;
  const int rows = P.rows, d = EXACT ? NCH * 32 * V : args.d;
  const float eps = args.eps;
  const RowMap in_map = P.in_map, out_map = P.out_map;
  const int cta = static_cast<int>(blockIdx.x) - cta0, ncta = P.cta_end - cta0;
  constexpr int NP = Vec2IO<T>::NP;          // float2 pairs per 16-byte chunk
  typedef typename VecIO<T>::Raw Raw;
  extern __shared__ __align__(16) float gb[];   // gamma [d] | beta [d]
  const int lane = threadIdx.x & 31;
  const int nwarps = blockDim.x >> 5;
  const int gwarp = cta * nwarps + (threadIdx.x >> 5);
  const int total_warps = ncta * nwarps;
  const int nvec = d / V;
  const float inv_d = 1.0f / static_cast<float>(d);
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    gb[i] = gamma[i];
    gb[d + i] = beta[i];
  }

  Raw n1[NCH], n2[NCH];
  auto issue_loads = [&](int r, Raw* dst) {
    const T* xr = x + (MAPPED ? in_map.phys(r) : static_cast<size_t>(r)) * d;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = lane + c * 32;
      dst[c] = (EXACT || idx < nvec) ? VecIO<T>::load_raw(xr + idx * V) : VecIO<T>::zero_raw();
    }
  };
  int r = gwarp;
  if (r < rows) issue_loads(r, n1);
  if (r + total_warps < rows) issue_loads(r + total_warps, n2);
  __syncthreads();

  for (; r < rows; r += total_warps) {
    float2 v[NCH * NP];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      Vec2IO<T>::unpack(n1[c], v + c * NP);
      n1[c] = n2[c];
    }
    if (r + 2 * total_warps < rows) issue_loads(r + 2 * total_warps, n2);
    float2 s2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int e = 0; e < NCH * NP; ++e) s2 = add2(s2, v[e]);              // padding chunks hold zeros
    const float mean = warp_sum(s2.x + s2.y) * inv_d;
    const float2 nm = splat2(-mean);
    float2 q2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      if (EXACT || lane + c * 32 < nvec) {
#pragma unroll
        for (int j = 0; j < NP; ++j) {
          v[c * NP + j] = add2(v[c * NP + j], nm);                       // centred
          q2 = fma2(v[c * NP + j], v[c * NP + j], q2);
        }
      }
    }
    const float rstd = rsqrtf(warp_sum(q2.x + q2.y) * inv_d + eps);
    if (lane == 0 && mean_out != nullptr) {
      mean_out[r] = mean;
      rstd_out[r] = rstd;
    }
    const float2 rs2 = splat2(rstd);
    T* yr = y + (MAPPED ? out_map.phys(r) : static_cast<size_t>(r)) * d;
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = lane + c * 32;
      if (EXACT || idx < nvec) {
        float2 o[NP];
#pragma unroll
        for (int j = 0; j < NP; j += 2) {
          const float4 g4 = *reinterpret_cast<const float4*>(gb + idx * V + 2 * j);
          const float4 b4 = *reinterpret_cast<const float4*>(gb + d + idx * V + 2 * j);
          o[j] = fma2(mul2(v[c * NP + j], rs2), make_float2(g4.x, g4.y), make_float2(b4.x, b4.y));
          o[j + 1] = fma2(mul2(v[c * NP + j + 1], rs2), make_float2(g4.z, g4.w), make_float2(b4.z, b4.w));
        }
        *reinterpret_cast<Raw*>(yr + idx * V) = Vec2IO<T>::pack(o);
      }
    }
  }
}

struct mb_ln_fwd_problem {      // mirror of include/maestro_b200.h
  const void* x;
  const float* gamma;
  const float* beta;
  void* y;
  float* mean;
  float* rstd;
  int rows;
  int in_n, in_bs, in_off, out_n, out_bs, out_off;
};

// CTAs of a grouped launch: `cap` resident CTAs split in proportion to the rows, at least one per problem and no
// more than a problem can use (rows / rows_per_cta)
static int split_ctas(const int* rows, int count, int cap, int rows_per_cta, int* cta_end) {
  long long total = 0;
  for (int i = 0; i < count; ++i) total += rows[i];
  int used = 0;
  for (int i = 0; i < count; ++i) {
    long long want = (static_cast<long long>(rows[i]) * cap + total - 1) / total;
    const long long mx = mb_ceil_div(rows[i], rows_per_cta);
    if (want > mx) want = mx;
    if (want < 1) want = 1;
    used += static_cast<int>(want);
    cta_end[i] = used;
  }
  return used;
}

template <typename T, int NCH>
static void launch_ln_fwd(const mb_ln_fwd_problem* ps, int count, int d, float eps, cudaStream_t st) {
  const int wpb = LNF_THREADS / 32;
  bool mapped = false;
  for (int i = 0; i < count; ++i)
    mapped = mapped || !(ps[i].in_n == ps[i].rows && ps[i].in_off == 0 && ps[i].out_n == ps[i].rows && ps[i].out_off == 0);
  auto kern = mapped ? layernorm_fwd_kernel<T, NCH, true> : layernorm_fwd_kernel<T, NCH, false>;
  if constexpr (sizeof(T) == 2 && NCH >= 2 && NCH <= 4) {
    if (d == NCH * 32 * VecIO<T>::N)
      kern = mapped ? layernorm_fwd_kernel<T, NCH, true, true> : layernorm_fwd_kernel<T, NCH, false, true>;
  }
  const size_t smem = 2 * static_cast<size_t>(d) * sizeof(float);
  // persistent grid of (at most) exactly the resident CTAs; at least 2 rows per warp (gamma / beta are staged per CTA)
  static int per_sm_tab[4] = {0, 0, 0, 0};
  int* per_sm = per_sm_tab + (d == NCH * 32 * VecIO<T>::N ? 2 : 0);
  if (per_sm[mapped] == 0) {
    int n = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, LNF_THREADS, smem);
    per_sm[mapped] = n > 0 ? n : 1;
  }
  LnFwdArgs a;
  int rows[LN_MAX_GROUP], cta_end[LN_MAX_GROUP];
  for (int i = 0; i < count; ++i) rows[i] = ps[i].rows;
  const int blocks = split_ctas(rows, count, 148 * per_sm[mapped], wpb * 2, cta_end);
  for (int i = 0; i < count; ++i) {
    LnFwdProblem& q = a.p[i];
    q.x = ps[i].x; q.gamma = ps[i].gamma; q.beta = ps[i].beta; q.y = ps[i].y; q.mean = ps[i].mean; q.rstd = ps[i].rstd;
    q.rows = ps[i].rows;
    q.in_map = RowMap{ps[i].in_n, ps[i].in_bs, ps[i].in_off};
    q.out_map = RowMap{ps[i].out_n, ps[i].out_bs, ps[i].out_off};
    q.cta_end = cta_end[i];
  }
  for (int i = count; i < LN_MAX_GROUP; ++i) a.p[i] = a.p[0];
  a.count = count; a.d = d; a.eps = eps;
  mb_launch(kern, dim3(blocks), dim3(LNF_THREADS), smem, st, a);
}

extern "C" int mb_layernorm_fwd_grouped(const mb_ln_fwd_problem* problems, int count, int dtype, int d, float eps,
                                        void* stream) {
  if (count < 1 || count > LN_MAX_GROUP) return -2;
  if (d > 1024 || d % 8 != 0) return -2;
  mb_ln_fwd_problem live[LN_MAX_GROUP];
  int n = 0;
  for (int i = 0; i < count; ++i)
    if (problems[i].rows > 0) live[n++] = problems[i];
  if (n == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define LNF_CASE(k)                                                                  \
  case k:                                                                            \
    MB_DISPATCH_DTYPE(dtype, T, (launch_ln_fwd<T, k>(live, n, d, eps, st)));         \
    break;
  const int vec = dtype == MB_DTYPE_F32 ? 4 : 8;
  const int nch = mb_ceil_div(d / vec, 32);
  switch (nch) {
    LNF_CASE(1) LNF_CASE(2) LNF_CASE(3) LNF_CASE(4) LNF_CASE(5) LNF_CASE(6) LNF_CASE(7) LNF_CASE(8)
    default: return -2;
  }
#undef LNF_CASE
  MB_CHECK_LAUNCH();
  return 0;
}

extern "C" int mb_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean,
                                float* rstd, int dtype, int rows, int d, float eps, int in_n, int in_bs, int in_off,
                                int out_n, int out_bs, int out_off, void* stream) {
  const mb_ln_fwd_problem p{x, gamma, beta, y, mean, rstd, rows, in_n, in_bs, in_off, out_n, out_bs, out_off};
  return mb_layernorm_fwd_grouped(&p, 1, dtype, d, eps, stream);
}

// ------------------------------------------------------------------ LayerNorm backward
// dx = rstd * (g - mean(g) - xhat * mean(g * xhat)) [+ dres],  g = dy * gamma.
// dgamma += sum_rows dy * xhat, dbeta += sum_rows dy, and optionally dbias += sum_rows dx (the bias gradient of the
// linear layer whose output this LayerNorm consumed: its dY is exactly this dx, so the column sum rides along).
// x / dx / dres rows use in_map, dy rows use out_map (mirrors the forward).
// One CTA per SM, 12 warps, warp w of the grid walks rows w, w + W, w + 2W, ...; lanes own columns, so the column
// sums live in registers for the whole kernel and are combined once per CTA (shared-memory red, then one fp32
// global red per column per CTA: <= 148 per address).
// Every warp owns a ring of `depth` row slots in shared memory that one lane fills with 1D bulk copies
// (cp.async.bulk, mbarrier completion): up to depth x (x, dy, dres) rows per warp are in flight from HBM without
// holding registers, which is what the bandwidth-latency product of the chip needs (~45 KB per SM).
// r2: the r1 kernel issued 551 warp instructions per 512-wide row and was issue-bound (ncu: 55 % issue-active with 12
// warps per SM at 0.57 of the copy bandwidth).  Packed fp32 math (two elements per instruction), the two row sums
// reduced together, folded scalars (one FFMA2 chain per output pair) roughly halve the count.
constexpr int LNB_THREADS = 384;

struct LnBwdProblem {
  const void* dy;
  const void* x;
  const float* mean;
  const float* rstd;
  const float* gamma;
  const void* dres;
  void* dx;
  float* dgamma;
  float* dbeta;
  float* dbias;
  int rows;
  RowMap in_map, out_map;
  int cta_end;
};
struct LnBwdArgs {
  LnBwdProblem p[LN_MAX_GROUP];
  int count, d, depth;
};

// EXACT: d == NCH * 32 * V (the 512 / 768 / 1024-wide rows of the shipped models): d is a compile-time constant and
// no lane is ever out of range -- the per-chunk predicates, their divergence bookkeeping (BSSY / BSYNC) and most of
// the address arithmetic disappear from the row loop.
template <typename T, int NCH, bool MAPPED, bool EXACT = false>
__global__ void __launch_bounds__(LNB_THREADS, 1)
layernorm_bwd_kernel(const __grid_constant__ LnBwdArgs args) {
  pdl_launch();
  pdl_wait();
  int pi = 0, cta0 = 0;
#pragma unroll
  for (int i = 0; i < LN_MAX_GROUP - 1; ++i)
    if (i + 1 < args.count && static_cast<int>(blockIdx.x) >= args.p[i].cta_end) { pi = i + 1; cta0 = args.p[i].cta_end; }
  const LnBwdProblem& P = args.p[pi];
  const T* __restrict__ dy = static_cast<const T*>(P.dy);
  const T* __restrict__ x = static_cast<const T*>(P.x);
  const float* __restrict__ mean_in = P.mean;
  const float* __restrict__ rstd_in = P.rstd;
  const float* __restrict__ gamma = P.gamma;
  const T* __restrict__ dres = static_cast<const T*>(P.dres);
  T* __restrict__ dx = static_cast<T*>(P.dx);
  float* __restrict__ dgamma = P.dgamma;
  float* __restrict__ dbeta = P.dbeta;
  float* __restrict__ dbias = P.dbias;
  constexpr int V = VecIO<T>::N;
  const int rows = P.rows, d = EXACT ? NCH * 32 * V : args.d, depth = args.depth;
  const RowMap in_map = P.in_map, out_map = P.out_map;
  const int cta = static_cast<int>(blockIdx.x) - cta0, ncta = P.cta_end - cta0;
  constexpr int NP = Vec2IO<T>::NP;
  constexpr int NE2 = NCH * NP;
  typedef typename VecIO<T>::Raw Raw;
  extern __shared__ __align__(128) uint8_t smem_raw[];
  // [3][d] partial dgamma / dbeta / dbias | gamma [d] | barriers [nwarps][depth] | ring [nwarps][depth][3][row]
  float* sm = reinterpret_cast<float*>(smem_raw);
  float* gam = sm + 3 * d;
  const int lane = threadIdx.x & 31;
  const int wid = threadIdx.x >> 5;
  const int nwarps = blockDim.x >> 5;
  uint64_t* bars = reinterpret_cast<uint64_t*>(gam + d) + wid * depth;
  const uint32_t row_bytes = static_cast<uint32_t>(d) * sizeof(T);
  uint8_t* ring = reinterpret_cast<uint8_t*>(reinterpret_cast<uint64_t*>(gam + d) + nwarps * depth) +
                  static_cast<size_t>(wid) * depth * 3 * row_bytes;
  const int gwarp = cta * nwarps + wid;
  const int total_warps = ncta * nwarps;
  const int nvec = d / V;
  const bool has_res = dres != nullptr;
  const bool has_bias = dbias != nullptr;
  for (int i = threadIdx.x; i < 3 * d; i += blockDim.x) sm[i] = 0.f;
  for (int i = threadIdx.x; i < d; i += blockDim.x) gam[i] = gamma[i];
  if (lane == 0) {
    for (int s = 0; s < depth; ++s) tc::mbar_init(&bars[s], 1);
    tc::fence_barrier_init();
  }
  __syncthreads();

  auto issue_row = [&](int slot, int r) {   // one lane
    const size_t pin = MAPPED ? in_map.phys(r) : static_cast<size_t>(r);
    const size_t pout = MAPPED ? out_map.phys(r) : static_cast<size_t>(r);
    uint8_t* dst = ring + static_cast<size_t>(slot) * 3 * row_bytes;
    tc::mbar_expect_tx(&bars[slot], (has_res ? 3u : 2u) * row_bytes);
    tc::bulk_load_1d(dst, x + pin * d, row_bytes, &bars[slot]);
    tc::bulk_load_1d(dst + row_bytes, dy + pout * d, row_bytes, &bars[slot]);
    if (has_res) tc::bulk_load_1d(dst + 2 * row_bytes, dres + pin * d, row_bytes, &bars[slot]);
  };
  if (lane == 0) {
    for (int s = 0; s < depth; ++s) {
      const long long rs = gwarp + static_cast<long long>(s) * total_warps;
      if (rs < rows) issue_row(s, static_cast<int>(rs));
    }
  }
  float2 ag[NE2], ab[NE2], ac[NE2];
#pragma unroll
  for (int e = 0; e < NE2; ++e) { ag[e] = make_float2(0.f, 0.f); ab[e] = ag[e]; ac[e] = ag[e]; }
  const float inv_d = 1.0f / static_cast<float>(d);

  int slot = 0;
  uint32_t phase = 0;
  for (int r = gwarp; r < rows; r += total_warps) {
    const size_t pin = MAPPED ? in_map.phys(r) : static_cast<size_t>(r);
    const float mean = mean_in[r], rstd = rstd_in[r];
    float2 xh[NE2], dv[NE2];
    Raw cres[NCH];
    tc::mbar_wait(&bars[slot], phase);
    {
      const T* sx = reinterpret_cast<const T*>(ring + static_cast<size_t>(slot) * 3 * row_bytes);
      const T* sdy = reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(sx) + row_bytes);
      const T* sres = reinterpret_cast<const T*>(reinterpret_cast<const uint8_t*>(sx) + 2 * row_bytes);
#pragma unroll
      for (int c = 0; c < NCH; ++c) {
        const int idx = lane + c * 32;
        const bool ok = EXACT || idx < nvec;
        const Raw rx = ok ? VecIO<T>::load_raw(sx + idx * V) : VecIO<T>::zero_raw();
        const Raw rdy = ok ? VecIO<T>::load_raw(sdy + idx * V) : VecIO<T>::zero_raw();
        cres[c] = (ok && has_res) ? VecIO<T>::load_raw(sres + idx * V) : VecIO<T>::zero_raw();
        Vec2IO<T>::unpack(rx, xh + c * NP);
        Vec2IO<T>::unpack(rdy, dv + c * NP);
      }
    }
    __syncwarp();   // every lane has read the slot: refill it with the row `depth` rounds ahead
    {
      const long long rn = r + static_cast<long long>(depth) * total_warps;
      if (lane == 0 && rn < rows) issue_row(slot, static_cast<int>(rn));
    }
    if (++slot == depth) { slot = 0; phase ^= 1; }
    // xhat = x * rstd - mean * rstd; g = dy * gamma; (s1, s2) = (sum g, sum g * xhat); dgamma / dbeta partials
    const float2 rs2 = splat2(rstd), sh2 = splat2(-mean * rstd);
    float2 s1 = make_float2(0.f, 0.f), s2 = make_float2(0.f, 0.f);
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = lane + c * 32;
      if (EXACT || idx < nvec) {
#pragma unroll
        for (int j = 0; j < NP; j += 2) {
          const float4 g4 = *reinterpret_cast<const float4*>(gam + idx * V + 2 * j);
          const float2 gq[2] = {make_float2(g4.x, g4.y), make_float2(g4.z, g4.w)};
#pragma unroll
          for (int k = 0; k < 2; ++k) {
            const int e = c * NP + j + k;
            const float2 h = fma2(xh[e], rs2, sh2);
            ag[e] = fma2(dv[e], h, ag[e]);
            ab[e] = add2(ab[e], dv[e]);
            const float2 gg = mul2(dv[e], gq[k]);
            s1 = add2(s1, gg);
            s2 = fma2(gg, h, s2);
            xh[e] = h;
            dv[e] = gg;
          }
        }
      }
    }
    const float2 tot = warp_sum2(make_float2(s1.x + s1.y, s2.x + s2.y));
    // dx = dres + rstd * g - rstd * mean(g) - xhat * rstd * mean(g * xhat)
    const float2 c1 = splat2(-rstd * tot.x * inv_d), c2 = splat2(-rstd * tot.y * inv_d);
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
      const int idx = lane + c * 32;
      if (EXACT || idx < nvec) {
        float2 o[NP];
        Vec2IO<T>::unpack(cres[c], o);
#pragma unroll
        for (int j = 0; j < NP; ++j) {
          const int e = c * NP + j;
          o[j] = add2(fma2(xh[e], c2, fma2(dv[e], rs2, o[j])), c1);
          if (has_bias) ac[e] = add2(ac[e], o[j]);
        }
        *reinterpret_cast<Raw*>(dx + pin * d + idx * V) = Vec2IO<T>::pack(o);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < NCH; ++c) {
    const int idx = lane + c * 32;
    if (EXACT || idx < nvec) {
#pragma unroll
      for (int j = 0; j < V; ++j) {
        // column (idx * V + j) lives at [j * nvec + idx]: consecutive lanes -> consecutive banks (conflict-free red)
        const int e = c * NP + j / 2;
        atomicAdd(&sm[j * nvec + idx], (j & 1) ? ag[e].y : ag[e].x);
        atomicAdd(&sm[d + j * nvec + idx], (j & 1) ? ab[e].y : ab[e].x);
        if (has_bias) atomicAdd(&sm[2 * d + j * nvec + idx], (j & 1) ? ac[e].y : ac[e].x);
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < d; i += blockDim.x) {
    const int si = (i % V) * nvec + i / V;
    atomicAdd(dgamma + i, sm[si]);
    atomicAdd(dbeta + i, sm[d + si]);
    if (has_bias) atomicAdd(dbias + i, sm[2 * d + si]);
  }
}

struct mb_ln_bwd_problem {      // mirror of include/maestro_b200.h
  const void* dy;
  const void* x;
  const float* mean;
  const float* rstd;
  const float* gamma;
  const void* dres;
  void* dx;
  float* dgamma;
  float* dbeta;
  float* dbias;
  int rows;
  int in_n, in_bs, in_off, out_n, out_bs, out_off;
};

template <typename T>
static int launch_ln_bwd(const mb_ln_bwd_problem* ps, int count, int d, cudaStream_t st) {
  constexpr int V = VecIO<T>::N;
  const int nwarps = LNB_THREADS / 32;
  // shared memory: column partials + gamma + barriers + per-warp row ring (depth slots of x / dy / dres rows)
  const size_t row_bytes = static_cast<size_t>(d) * sizeof(T);
  const size_t fixed = 4 * static_cast<size_t>(d) * sizeof(float);
  int depth = static_cast<int>((200 * 1024 - fixed - 1024) / (nwarps * 3 * row_bytes));
  if (depth > 4) depth = 4;
  if (depth < 1) depth = 1;
  const size_t smem = fixed + static_cast<size_t>(nwarps) * depth * (8 + 3 * row_bytes);
  const int nch = mb_ceil_div(d / V, 32);
  bool mapped = false;
  for (int i = 0; i < count; ++i)
    mapped = mapped || !(ps[i].in_n == ps[i].rows && ps[i].in_off == 0 && ps[i].out_n == ps[i].rows && ps[i].out_off == 0);
  LnBwdArgs a;
  int rows[LN_MAX_GROUP], cta_end[LN_MAX_GROUP];
  for (int i = 0; i < count; ++i) rows[i] = ps[i].rows;
  // one CTA per SM; at least 4 rows per warp before the per-CTA reduction tail pays off
  const int blocks = split_ctas(rows, count, 148, nwarps * 4, cta_end);
  for (int i = 0; i < count; ++i) {
    LnBwdProblem& q = a.p[i];
    q.dy = ps[i].dy; q.x = ps[i].x; q.mean = ps[i].mean; q.rstd = ps[i].rstd; q.gamma = ps[i].gamma; q.dres = ps[i].dres;
    q.dx = ps[i].dx; q.dgamma = ps[i].dgamma; q.dbeta = ps[i].dbeta; q.dbias = ps[i].dbias; q.rows = ps[i].rows;
    q.in_map = RowMap{ps[i].in_n, ps[i].in_bs, ps[i].in_off};
    q.out_map = RowMap{ps[i].out_n, ps[i].out_bs, ps[i].out_off};
    q.cta_end = cta_end[i];
  }
  for (int i = count; i < LN_MAX_GROUP; ++i) a.p[i] = a.p[0];
  a.count = count; a.d = d; a.depth = depth;
#define MB_LNB(N_)                                                                                              \
  case N_:                                                                                                      \
    {                                                                                                           \
      static bool attr_done = false;                                                                            \
      if (!attr_done) {                                                                                         \
        cudaFuncSetAttribute(layernorm_bwd_kernel<T, N_, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024); \
        cudaFuncSetAttribute(layernorm_bwd_kernel<T, N_, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024); \
        attr_done = true;                                                                                       \
      }                                                                                                         \
      auto kern = mapped ? layernorm_bwd_kernel<T, N_, true> : layernorm_bwd_kernel<T, N_, false>;              \
      if constexpr (sizeof(T) == 2 && N_ >= 2 && N_ <= 4) {                                                     \
        if (d == N_ * 32 * V) {                                                                                 \
          static bool attr_x = false;                                                                           \
          if (!attr_x) {                                                                                        \
            cudaFuncSetAttribute(layernorm_bwd_kernel<T, N_, true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024); \
            cudaFuncSetAttribute(layernorm_bwd_kernel<T, N_, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024); \
            attr_x = true;                                                                                      \
          }                                                                                                     \
          kern = mapped ? layernorm_bwd_kernel<T, N_, true, true> : layernorm_bwd_kernel<T, N_, false, true>;   \
        }                                                                                                       \
      }                                                                                                         \
      mb_launch(kern, dim3(blocks), dim3(LNB_THREADS), smem, st, a);                                                            \
    }                                                                                                           \
    break;
  switch (nch) {
    MB_LNB(1) MB_LNB(2) MB_LNB(3) MB_LNB(4) MB_LNB(5) MB_LNB(6) MB_LNB(7) MB_LNB(8)
    default: return -2;
  }
#undef MB_LNB
  return 0;
}

// dbias may be null.  dgamma / dbeta / dbias accumulate (fp32).
extern "C" int mb_layernorm_bwd_grouped(const mb_ln_bwd_problem* problems, int count, int dtype, int d, void* stream) {
  if (count < 1 || count > LN_MAX_GROUP) return -2;
  if (d > 1024 || d % 8 != 0) return -2;
  mb_ln_bwd_problem live[LN_MAX_GROUP];
  int n = 0;
  for (int i = 0; i < count; ++i)
    if (problems[i].rows > 0) live[n++] = problems[i];
  if (n == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc;
  if (dtype == MB_DTYPE_F32) rc = launch_ln_bwd<float>(live, n, d, st);
  else if (dtype == MB_DTYPE_BF16) rc = launch_ln_bwd<bf16>(live, n, d, st);
  else return -1;
  if (rc != 0) return rc;
  MB_CHECK_LAUNCH();
  return 0;
}

extern "C" int mb_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd,
                                const float* gamma, const void* dres, void* dx, float* dgamma, float* dbeta,
                                float* dbias, int dtype, int rows, int d, int in_n, int in_bs, int in_off, int out_n,
                                int out_bs, int out_off, void* stream) {
  const mb_ln_bwd_problem p{dy, x, mean, rstd, gamma, dres, dx, dgamma, dbeta, dbias, rows,
                            in_n, in_bs, in_off, out_n, out_bs, out_off};
  return mb_layernorm_bwd_grouped(&p, 1, dtype, d, stream);
}

// ------------------------------------------------------------------ GroupNorm(1, E) of the patch embedding
// Phase 1: per-image sum / sum of squares, split over several CTAs per image (fp64 atomics).
template <typename T>
__global__ void __launch_bounds__(256) gn_stats_kernel(const T* __restrict__ y, double* __restrict__ sums,
                                                       size_t elems_per_img, int splits) {
  constexpr int V = VecIO<T>::N;
  const int img = blockIdx.x / splits, part = blockIdx.x % splits;
  const size_t nvec = elems_per_img / V;
  const size_t per = (nvec + splits - 1) / splits;
  const size_t v0 = part * per, v1 = min(nvec, v0 + per);
  const T* base = y + static_cast<size_t>(img) * elems_per_img;
  float s = 0.f, ss = 0.f;
  for (size_t i = v0 + threadIdx.x; i < v1; i += blockDim.x) {
    float v[V];
    VecIO<T>::load(base + i * V, v);
#pragma unroll
    for (int j = 0; j < V; ++j) { s += v[j]; ss += v[j] * v[j]; }
  }
  double ds = warp_sum(static_cast<double>(s)), dss = warp_sum(static_cast<double>(ss));
  __shared__ double red[2][8];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) { red[0][w] = ds; red[1][w] = dss; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double a = 0, b = 0;
    for (int i = 0; i < (blockDim.x >> 5); ++i) { a += red[0][i]; b += red[1][i]; }
    atomicAdd(&sums[2 * img], a);
    atomicAdd(&sums[2 * img + 1], b);
  }
}

// sums must be zero on entry ([n_img, 2] doubles).
extern "C" int mb_gn_stats(const void* y, int dtype, double* sums, int n_img, long long elems_per_img, void* stream) {
  if (n_img == 0) return 0;
  if (elems_per_img % 8 != 0) return -2;
  int splits = (148 * 4) / n_img;
  if (splits < 1) splits = 1;
  const long long max_splits = elems_per_img / (8 * 256) + 1;
  if (splits > max_splits) splits = static_cast<int>(max_splits);
  MB_DISPATCH_DTYPE(dtype, T,
                    (gn_stats_kernel<T><<<n_img * splits, 256, 0, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(y), sums, static_cast<size_t>(elems_per_img), splits)));
  MB_CHECK_LAUNCH();
  return 0;
}

// Phase 2: normalise + affine + positional + date encoding, written into the group token buffer.
//   y      [n_img * L, E]   conv output (image i = b * D + d)
//   out    [B, out_bs, E]   row = b * out_bs + out_off + d * L + l
//   pos    [L, E] fp32 (already scaled by fac_abs_enc), feat [B, feat_bs, 5] date features (rows feat_off + d).
template <typename T>
__global__ void __launch_bounds__(256) gn_apply_kernel(const T* __restrict__ y, const double* __restrict__ sums,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       const float* __restrict__ pos, const float* __restrict__ feat,
                                                       T* __restrict__ out, float* __restrict__ stats_out, int n_img,
                                                       int D, int L, int E, float eps, int out_bs, int out_off,
                                                       int date_dim, float fac_date, int feat_bs, int feat_off) {
  constexpr int V = VecIO<T>::N;
  extern __shared__ float st_s[];            // [n_img][2]: mean, rstd (the fp64 finalisation runs once per block)
  const double inv_n = 1.0 / (static_cast<double>(L) * E);
  for (int i = threadIdx.x; i < n_img; i += blockDim.x) {
    const double m = sums[2 * i] * inv_n;
    const double var = fmax(sums[2 * i + 1] * inv_n - m * m, 0.0);
    const float mean = static_cast<float>(m);
    const float rstd = static_cast<float>(1.0 / sqrt(var + static_cast<double>(eps)));
    st_s[2 * i] = mean;
    st_s[2 * i + 1] = rstd;
    if (stats_out != nullptr && blockIdx.x == 0) {
      stats_out[2 * i] = mean;
      stats_out[2 * i + 1] = rstd;
    }
  }
  __syncthreads();
  const unsigned nvec_row = E / V;
  const unsigned total = static_cast<unsigned>(n_img) * L * nvec_row;      // host checks < 2^32
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned row = i / nvec_row, cv = i - row * nvec_row;
    const unsigned img = row / L, l = row - img * L;
    const unsigned b = img / D, d = img - b * D;
    const float mean = st_s[2 * img], rstd = st_s[2 * img + 1];
    const float shift = -mean * rstd;
    float v[V], gm[V], bt[V], ps[V];
    VecIO<T>::load(y + static_cast<size_t>(row) * E + cv * V, v);
#pragma unroll
    for (int j = 0; j < V; j += 4) {
      VecIO<float>::load(gamma + cv * V + j, gm + j);
      VecIO<float>::load(beta + cv * V + j, bt + j);
      VecIO<float>::load(pos + static_cast<size_t>(l) * E + cv * V + j, ps + j);
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const int e = cv * V + j;
      float o = fmaf(fmaf(v[j], rstd, shift), gm[j], bt[j]) + ps[j];
      if (e >= E - date_dim) o += date_channel(feat + (static_cast<size_t>(b) * feat_bs + feat_off + d) * 5, e - (E - date_dim), date_dim, fac_date);
      v[j] = o;
    }
    const size_t orow = static_cast<size_t>(b) * out_bs + out_off + static_cast<size_t>(d) * L + l;
    VecIO<T>::store(out + orow * E + cv * V, v);
  }
}

extern "C" int mb_gn_apply(const void* y, const double* sums, const float* gamma, const float* beta,
                           const float* pos, const float* feat, void* out, float* stats_out, int dtype, int n_img,
                           int D, int L, int E, float eps, int out_bs, int out_off, int date_dim, float fac_date,
                           int feat_bs, int feat_off, void* stream) {
  if (n_img == 0) return 0;
  if (E % 8 != 0) return -2;
  const long long total = static_cast<long long>(n_img) * L * (E / 4);
  if (total >= 0xffffffffll || n_img > 4096) return -2;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const size_t st_bytes = static_cast<size_t>(n_img) * 2 * sizeof(float);
  MB_DISPATCH_DTYPE(dtype, T,
                    (gn_apply_kernel<T><<<static_cast<int>(blocks), 256, st_bytes, static_cast<cudaStream_t>(stream)>>>(
                        static_cast<const T*>(y), sums, gamma, beta, pos, feat, static_cast<T*>(out), stats_out,
                        n_img, D, L, E, eps, out_bs, out_off, date_dim, fac_date, feat_bs, feat_off)));
  MB_CHECK_LAUNCH();
  return 0;
}

// Backward phase 1: per image a = sum(dy * gamma), c = sum(dy * gamma * xhat) (fp64 atomics into sums2 [n_img,2]);
// per channel dgamma += sum dy * xhat, dbeta += sum dy (fp32 atomics).  dy is read from the group token buffer.
// grid (n_img, row blocks); the block's threads are laid out as [row lane][16-byte column chunk], every row lane
// keeps 4 rows of loads in flight, and the row lanes are combined in shared memory so that a block issues ONE
// atomic per column (same-address atomics from many blocks are what bounds this kernel).
constexpr int GNR_THREADS = 512;
template <typename T>
__global__ void __launch_bounds__(GNR_THREADS) gn_bwd_reduce_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                                    const float* __restrict__ stats,
                                                                    const float* __restrict__ gamma,
                                                                    double* __restrict__ sums2,
                                                                    float* __restrict__ dgamma,
                                                                    float* __restrict__ dbeta, int D, int L, int E,
                                                                    int out_bs, int out_off, int rows_per_block) {
  constexpr int V = VecIO<T>::N;
  constexpr int U = 4;
  typedef typename VecIO<T>::Raw Raw;
  extern __shared__ float cs[];              // [2][E]: dgamma / dbeta partials, column (cv * V + j) at [j * nvec + cv]
  __shared__ double red[2][GNR_THREADS / 32];
  const int img = blockIdx.x;
  const int b = img / D, d = img % D;
  const int l0 = blockIdx.y * rows_per_block, l1 = min(L, l0 + rows_per_block);
  const float mean = stats[2 * img], rstd = stats[2 * img + 1];
  const int nvec = E / V;
  const int cv = threadIdx.x % nvec, rl = threadIdx.x / nvec, nrl = blockDim.x / nvec;
  for (int i = threadIdx.x; i < 2 * E; i += blockDim.x) cs[i] = 0.f;
  __syncthreads();
  float a = 0.f, c = 0.f;
  if (rl < nrl) {
    float ag[V], ab[V], gm[V];
#pragma unroll
    for (int j = 0; j < V; j += 4) VecIO<float>::load(gamma + cv * V + j, gm + j);
#pragma unroll
    for (int j = 0; j < V; ++j) { ag[j] = 0.f; ab[j] = 0.f; }
    for (int l = l0 + rl; l < l1; l += nrl * U) {
      Raw ry[U], rd[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int lu = min(l + u * nrl, l1 - 1);    // clamped rows are loaded again and skipped below
        const size_t yrow = static_cast<size_t>(img) * L + lu;
        const size_t drow = static_cast<size_t>(b) * out_bs + out_off + static_cast<size_t>(d) * L + lu;
        ry[u] = VecIO<T>::load_raw(y + yrow * E + cv * V);
        rd[u] = VecIO<T>::load_raw(dy + drow * E + cv * V);
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (l + u * nrl >= l1) break;
        float yv[V], dv[V];
        VecIO<T>::unpack(ry[u], yv);
        VecIO<T>::unpack(rd[u], dv);
#pragma unroll
        for (int j = 0; j < V; ++j) {
          const float h = (yv[j] - mean) * rstd;
          const float g = dv[j] * gm[j];
          a += g;
          c += g * h;
          ag[j] += dv[j] * h;
          ab[j] += dv[j];
        }
      }
    }
#pragma unroll
    for (int j = 0; j < V; ++j) {
      atomicAdd(&cs[j * nvec + cv], ag[j]);
      atomicAdd(&cs[E + j * nvec + cv], ab[j]);
    }
  }
  double da = warp_sum(static_cast<double>(a)), dc = warp_sum(static_cast<double>(c));
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) { red[0][w] = da; red[1][w] = dc; }
  __syncthreads();
  for (int i = threadIdx.x; i < E; i += blockDim.x) {
    const int si = (i % V) * nvec + i / V;
    atomicAdd(dgamma + i, cs[si]);
    atomicAdd(dbeta + i, cs[E + si]);
  }
  if (threadIdx.x == 0) {
    double s0 = 0, s1 = 0;
    for (int i = 0; i < ((blockDim.x + 31) >> 5); ++i) { s0 += red[0][i]; s1 += red[1][i]; }
    atomicAdd(&sums2[2 * img], s0);
    atomicAdd(&sums2[2 * img + 1], s1);
  }
}

// Backward phase 2: dconv = rstd * (dy * gamma - a / n - xhat * c / n), rows back in conv order.
template <typename T>
__global__ void __launch_bounds__(256) gn_bwd_apply_kernel(const T* __restrict__ dy, const T* __restrict__ y,
                                                           const float* __restrict__ stats,
                                                           const float* __restrict__ gamma,
                                                           const double* __restrict__ sums2, T* __restrict__ dconv,
                                                           int n_img, int D, int L, int E, int out_bs, int out_off) {
  constexpr int V = VecIO<T>::N;
  extern __shared__ float st_s[];            // [n_img][4]: mean, rstd, a / n, c / n (fp64 -> fp32 once per block)
  const double inv_n = 1.0 / (static_cast<double>(L) * E);
  for (int i = threadIdx.x; i < n_img; i += blockDim.x) {
    st_s[4 * i] = stats[2 * i];
    st_s[4 * i + 1] = stats[2 * i + 1];
    st_s[4 * i + 2] = static_cast<float>(sums2[2 * i] * inv_n);
    st_s[4 * i + 3] = static_cast<float>(sums2[2 * i + 1] * inv_n);
  }
  __syncthreads();
  const unsigned nvec_row = E / V;
  const unsigned total = static_cast<unsigned>(n_img) * L * nvec_row;      // host checks < 2^32
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const unsigned row = i / nvec_row, cv = i - row * nvec_row;
    const unsigned img = row / L, l = row - img * L;
    const unsigned b = img / D, d = img - b * D;
    const float4 q = *reinterpret_cast<const float4*>(st_s + 4 * img);
    const float mean = q.x, rstd = q.y, ma = q.z, mc = q.w;
    const size_t drow = static_cast<size_t>(b) * out_bs + out_off + static_cast<size_t>(d) * L + l;
    float yv[V], dv[V], gm[V];
    VecIO<T>::load(y + static_cast<size_t>(row) * E + cv * V, yv);
    VecIO<T>::load(dy + drow * E + cv * V, dv);
#pragma unroll
    for (int j = 0; j < V; j += 4) VecIO<float>::load(gamma + cv * V + j, gm + j);
#pragma unroll
    for (int j = 0; j < V; ++j) {
      const float h = (yv[j] - mean) * rstd;
      dv[j] = rstd * (dv[j] * gm[j] - ma - h * mc);
    }
    VecIO<T>::store(dconv + static_cast<size_t>(row) * E + cv * V, dv);
  }
}

// sums2 must be zero on entry.
extern "C" int mb_gn_bwd(const void* dy, const void* y, const float* stats, const float* gamma, double* sums2,
                         void* dconv, float* dgamma, float* dbeta, int dtype, int n_img, int D, int L, int E,
                         int out_bs, int out_off, void* stream) {
  if (n_img == 0) return 0;
  if (E % 8 != 0) return -2;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (E / 4 > GNR_THREADS) return -2;        // one column chunk per thread
  int row_blocks = (148 * 2) / n_img;        // ~2 blocks per SM; every block ends in 2 E same-address atomics
  if (row_blocks < 1) row_blocks = 1;
  if (row_blocks > L) row_blocks = L;
  const int rows_per_block = mb_ceil_div(L, row_blocks);
  row_blocks = mb_ceil_div(L, rows_per_block);
  dim3 grid(n_img, row_blocks);
  const long long total = static_cast<long long>(n_img) * L * (E / 4);
  if (total >= 0xffffffffll || n_img > 2048) return -2;
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const size_t st_bytes = static_cast<size_t>(n_img) * 4 * sizeof(float);
  MB_DISPATCH_DTYPE(
      dtype, T,
      (gn_bwd_reduce_kernel<T><<<grid, GNR_THREADS, 2 * E * sizeof(float), st>>>(static_cast<const T*>(dy), static_cast<const T*>(y), stats, gamma,
                                                     sums2, dgamma, dbeta, D, L, E, out_bs, out_off, rows_per_block),
       gn_bwd_apply_kernel<T><<<static_cast<int>(blocks), 256, st_bytes, st>>>(static_cast<const T*>(dy),
                                                                        static_cast<const T*>(y), stats, gamma, sums2,
                                                                        static_cast<T*>(dconv), n_img, D, L, E, out_bs,
                                                                        out_off)));
  MB_CHECK_LAUNCH();
  return 0;
}

// ------------------------------------------------------------------ GroupNorm forward without any allocation
// mb_gn_fwd_static = mb_gn_stats + mb_gn_apply with the [n_img, 2] fp64 sum scratch in a file-scope device array
// instead of a caller-provided buffer: for callers that cannot allocate (graph capture, serving loops).  No
// counterpart in the reference.  Limit: MB_GN_STATIC_MAX_IMG images per
// call (64 KiB of device memory per GPU that loads the library, held until exit).  Every call shares the array:
// calls that overlap on the GPU (different streams) race.
#define MB_GN_STATIC_MAX_IMG 4096
static __device__ __align__(256) double g_gn_static_sums[2 * MB_GN_STATIC_MAX_IMG];

namespace {
constexpr int GN_STATIC_MAX_DEV = 64;
double* g_gn_static_ptr[GN_STATIC_MAX_DEV] = {};      // device address of the array, per device (host code cannot
                                                      // use the symbol's name as a pointer)
// Once per device, with that device current.
int gn_static_scratch(double** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return static_cast<int>(e);
  if (dev < 0 || dev >= GN_STATIC_MAX_DEV) return -3;
  double* p = __atomic_load_n(&g_gn_static_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_gn_static_sums);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (p == nullptr) return -3;
    __atomic_store_n(&g_gn_static_ptr[dev], p, __ATOMIC_RELEASE);   // same value from every thread that raced here
  }
  *out = p;
  return 0;
}
}  // namespace

extern "C" int mb_gn_static_max_images(void) { return MB_GN_STATIC_MAX_IMG; }

// Returns 0, -2 for unsupported shapes (as mb_gn_stats / mb_gn_apply), -4 when n_img exceeds
// MB_GN_STATIC_MAX_IMG, a positive cudaError_t otherwise.  Calls no cudaMalloc / cudaFree.
extern "C" int mb_gn_fwd_static(const void* y, const float* gamma, const float* beta, const float* pos,
                                const float* feat, void* out, float* stats_out, int dtype, int n_img, int D, int L,
                                int E, float eps, int out_bs, int out_off, int date_dim, float fac_date, int feat_bs,
                                int feat_off, void* stream) {
  if (n_img == 0) return 0;
  if (n_img < 0) return -2;
  if (n_img > MB_GN_STATIC_MAX_IMG) return -4;
  if (E % 8 != 0) return -2;
  double* sums = nullptr;
  int rc = gn_static_scratch(&sums);
  if (rc != 0) return rc;
  // the array keeps the previous call's sums: the statistics kernel accumulates, so start from zero
  cudaError_t e = cudaMemsetAsync(sums, 0, static_cast<size_t>(n_img) * 2 * sizeof(double),
                                  static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return static_cast<int>(e);
  rc = mb_gn_stats(y, dtype, sums, n_img, static_cast<long long>(L) * E, stream);
  if (rc != 0) return rc;
  return mb_gn_apply(y, sums, gamma, beta, pos, feat, out, stats_out, dtype, n_img, D, L, E, eps, out_bs, out_off,
                     date_dim, fac_date, feat_bs, feat_off, stream);
}
