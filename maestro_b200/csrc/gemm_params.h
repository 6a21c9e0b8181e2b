// Parameter block shared by the GEMM entry points (mirrored field-for-field in include/maestro_b200.h
// and in maestro_b200/_lib.py as a ctypes.Structure).
#pragma once

#define GEMM_GELU 1      // out = gelu(acc + bias); aux <- gelu'(acc + bias) (saved for the backward)
#define GEMM_DGELU 2     // out = acc * aux (aux = the derivative the forward GEMM_GELU call saved)
#define GEMM_ACCUM 4     // out += result (fp32 gradient accumulation)
#define GEMM_OUT_F32 8   // C is fp32 regardless of the activation dtype
#define GEMM_PAIR_ON 16  // scheduling hint (tensor-core path): force 2-CTA (cta_group::2) tiles
#define GEMM_PAIR_OFF 32 // scheduling hint (tensor-core path): force single-CTA tiles
#define GEMM_HALF_TAIL 64 // scheduling hint (tensor-core path, bf16 outputs): last partial wave as half-width tiles

struct GemmParams {
  const void* A;         // A(m, k) = A[m * sam + k * sak]
  const void* B;         // B(n, k) = B[n * sbn + k * sbk]
  void* C;               // C[m * ldc + n]
  const float* bias;     // [N] fp32 or null
  const void* residual;  // same layout as C (activation dtype) or null
  void* aux;             // same layout as C (activation dtype) or null
  int M, N, K;
  long long sam, sak, sbn, sbk;
  int ldc;
  int nb1, nb2;          // batch = nb1 * nb2, z = z1 * nb2 + z2
  long long a_s1, a_s2, b_s1, b_s2, c_s1, c_s2;
  float alpha;
  int flags;
  float* colsum;         // optional [N] fp32: colsum[n] += sum_m C[m, n] (the rounded outputs), tensor-core path only
};
