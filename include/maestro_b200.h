/* maestro_b200 -- C ABI of the B200 (sm_100a) kernels behind MAESTRO's MAE pretraining step.
 *
 * The reference (IGNF/MAESTRO) is pure Python on eager PyTorch: it has no FFI of its own.  The drop-in boundary is
 * the Python module interface of maestro/layers + maestro/ssl; this header is what sits underneath
 * it.  Every entry point replaces the torch ops of the reference lines cited beside it.
 *
 * Conventions: plain pointers and sizes, no torch types.  All pointers are DEVICE pointers unless said otherwise.
 * `dtype` selects the activation type of the `void*` tensors: 0 = float32 (fp32 check mode), 1 = bfloat16.
 * `stream` is a cudaStream_t.  Return value: 0 on success, a cudaError_t (> 0) or a negative argument-check code.
 * Kernels never allocate; callers own every buffer.  Functions are re-entrant.
 */
#ifndef MAESTRO_B200_H
#define MAESTRO_B200_H

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MB_DTYPE_F32 0
#define MB_DTYPE_BF16 1

/* Row map: logical row r of a [rows, d] matrix lives at physical row (r / n) * bs + off + (r % n).
 * It folds the reference's torch.cat / split around encoder_inter (maestro/ssl/mim.py:406-423) and the
 * ungroup before Pixelify (maestro/ssl/mim.py:326-341) into the LayerNorm pass.  Identity: n = rows, bs = 0, off = 0. */

/* ---- data movement ------------------------------------------------------------------------------------------ */
/* fp32 master weights -> bf16 shadow (what torch autocast does per op, maestro/conf/trainer.py:13). */
int mb_cast_f32_to_bf16(const float* src, void* dst, size_t n, void* stream);
/* dst[C, R] = bf16(src[R, C]^T): transposed shadow read by the dgrad GEMMs. */
int mb_cast_transpose_f32_to_bf16(const float* src, void* dst, int R, int C, void* stream);
/* BaseMIM.resize_and_rescale, maestro/ssl/mim.py:425-437 (F.interpolate nearest + DEM 30*(ch0-ch)). */
int mb_resize_nearest(const float* src, float* dst, int n_img, int C, int hin, int win, int hout, int wout,
                      int rescale_elev, void* stream);
/* The same with F.interpolate's other modes (align_corners=False): mode 0 nearest, 1 bilinear, 2 bicubic (A = -0.75). */
int mb_resize(const float* src, float* dst, int n_img, int C, int hin, int win, int hout, int wout, int mode,
              int rescale_elev, void* stream);
/* Patch extraction of the strided conv in PatchifyBands, maestro/layers/embed.py:57-60. */
int mb_im2col(const float* img, void* cols, int dtype, int n_img, int C, int c0, int c, int S, int p, int Kpad,
              void* stream);
/* PixelifyBands depth-to-space "(p1 p2 c) h w -> c (h p1) (w p2)", maestro/layers/embed.py:155-160 (dir 0) / adjoint. */
int mb_rows_image(void* rows, int dtype, float* img, int n_img, int C, int c0, int c, int S, int p, int dir,
                  void* stream);
/* encode_dates features, maestro/layers/utils.py:128-160: int16 (year, doy, hour) -> (diff, sin/cos doy, sin/cos hour). */
int mb_date_features(const short* dates, const short* ref, float* feat, int B, int D, int Dtot, int doff, void* stream);
/* Bias gradients: out[n] += sum_m dy[m, n] (autograd of nn.Linear / nn.Conv2d bias). */
int mb_colsum_accum(const void* dy, int dtype, float* out, int M, int N, void* stream);

/* ---- normalisation ------------------------------------------------------------------------------------------ */
/* nn.LayerNorm inside vit_pytorch Attention / FeedForward / Transformer.norm (built at maestro/ssl/mae.py:133-176). */
int mb_layernorm_fwd(const void* x, const float* gamma, const float* beta, void* y, float* mean, float* rstd,
                     int dtype, int rows, int d, float eps, int in_n, int in_bs, int in_off, int out_n, int out_bs,
                     int out_off, void* stream);
/* Its backward, fused with the residual-gradient add (x = f(LN(x)) + x). dgamma / dbeta accumulate.  dbias (may be
 * null) accumulates the column sums of dx: the bias gradient of the nn.Linear that produced this LayerNorm's input
 * stream (to_out / FeedForward down-projection), whose dY is exactly this dx. */
int mb_layernorm_bwd(const void* dy, const void* x, const float* mean, const float* rstd, const float* gamma,
                     const void* dres, void* dx, float* dgamma, float* dbeta, float* dbias, int dtype, int rows, int d,
                     int in_n, int in_bs, int in_off, int out_n, int out_bs, int out_off, void* stream);
/* Grouped forms: `count` (1..4) LayerNorms of one width as ONE launch -- the per-group encoder / decoder stacks of
 * MAE.encode / MAE.decode (maestro/ssl/mae.py:289-306, looped per group by encode_or_decode, maestro/ssl/mim.py:396-
 * 403) step in lockstep, so the four LayerNorms of a layer (1 600 - 8 192 rows each) become one kernel whose CTAs are
 * split between the problems in proportion to their rows.  `problems` is a HOST pointer. */
typedef struct {
  const void* x; const float* gamma; const float* beta; void* y; float* mean; float* rstd;
  int rows;
  int in_n, in_bs, in_off, out_n, out_bs, out_off;       /* row maps (identity: n = rows, bs = 0, off = 0) */
} mb_ln_fwd_problem;
typedef struct {
  const void* dy; const void* x; const float* mean; const float* rstd; const float* gamma; const void* dres;
  void* dx; float* dgamma; float* dbeta; float* dbias;   /* dres / dbias may be NULL */
  int rows;
  int in_n, in_bs, in_off, out_n, out_bs, out_off;
} mb_ln_bwd_problem;
int mb_layernorm_fwd_grouped(const mb_ln_fwd_problem* problems, int count, int dtype, int d, float eps, void* stream);
int mb_layernorm_bwd_grouped(const mb_ln_bwd_problem* problems, int count, int dtype, int d, void* stream);
/* nn.GroupNorm(1, E) of PatchifyBands, maestro/layers/embed.py:55,61: per-image sum / sum-of-squares (fp64). */
int mb_gn_stats(const void* y, int dtype, double* sums, int n_img, long long elems_per_img, void* stream);
/* GroupNorm apply + enc_add_encodings (maestro/ssl/mim.py:232-252) + group_mods concat (layers/utils.py:12-47). */
int mb_gn_apply(const void* y, const double* sums, const float* gamma, const float* beta, const float* pos,
                const float* feat, void* out, float* stats_out, int dtype, int n_img, int D, int L, int E, float eps,
                int out_bs, int out_off, int date_dim, float fac_date, int feat_bs, int feat_off, void* stream);
int mb_gn_bwd(const void* dy, const void* y, const float* stats, const float* gamma, double* sums2, void* dconv,
              float* dgamma, float* dbeta, int dtype, int n_img, int D, int L, int E, int out_bs, int out_off,
              void* stream);
/* mb_gn_stats + mb_gn_apply in one call that allocates nothing and needs no caller scratch: the [n_img, 2] fp64 sums
 * live in a file-scope device array (64 KiB per GPU that loads the library, held until exit).  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc (CUDA graph
 * capture, serving loops).  n_img <= mb_gn_static_max_images() (4096), else -4.  Every call shares the array (cleared
 * on the stream at the start of each call): calls that overlap on the GPU race. */
int mb_gn_static_max_images(void);
int mb_gn_fwd_static(const void* y, const float* gamma, const float* beta, const float* pos, const float* feat,
                     void* out, float* stats_out, int dtype, int n_img, int D, int L, int E, float eps, int out_bs,
                     int out_off, int date_dim, float fac_date, int feat_bs, int feat_off, void* stream);

/* ---- masking ------------------------------------------------------------------------------------------------ */
/* MAE.mask_seq index math, maestro/ssl/mae.py:239-259 (rand*struct -> argsort -> argsort -> gather -> sort x2).
 * B sequences of L tokens.  Lt (0 = L): period of the per-sample tables (here seg_id) over the flat token index
 * b * L + p -- a multiple of L when fusion modes "shared" / "monotemp" flatten the dates of a sample into the batch
 * (maestro/layers/utils.py:26-33) and one sample therefore spans Lt / L sequences. */
int mb_mask_indices(const float* noise, const unsigned char* structm, int B, int L, int nm, unsigned char* mask,
                    int* slot, long long* masked_idx, long long* unmasked_idx, const int* seg_id, int* seg_counts,
                    int Lt, void* stream);
/* x[batch_indices, unmasked_indices], maestro/ssl/mae.py:261. */
int mb_gather_rows(const void* x, const long long* idx, void* out, int dtype, int B, int L, int n, int d, void* stream);
/* Its gradient (index_put of zeros + kept rows). */
int mb_scatter_kept(const void* dkept, const unsigned char* mask, const int* slot, void* dx, int dtype, int B, int L,
                    int n, int d, void* stream);
/* MAE.unmask_seq (maestro/ssl/mae.py:266-287) fused with dec_add_encodings (maestro/ssl/mim.py:254-274). */
int mb_unmask_fwd(const void* y, const unsigned char* mask, const int* slot, const int* tok_id, const int* src_pos,
                  const float* tok, const float* pos, const float* feat, const int* date_id, void* out, int dtype,
                  int B, int L, int n_kept, int d, int Dtot, int date_dim, float fac_date, int Lt, void* stream);
int mb_masktoken_grad(const void* dx, const unsigned char* mask, const int* tok_id, const int* src_pos, float* dtok,
                      int dtype, int B, int L, int d, int n_kinds, int Lt, void* stream);

/* ---- loss --------------------------------------------------------------------------------------------------- */
#define MB_LOSS_MAX_C 16
#define MB_LOSS_MAX_G 4
typedef struct {
  const float* target;              /* [B*D, C, S, S] */
  const void* pred[MB_LOSS_MAX_G];  /* per band-group rows [B*D*L, c_g*p*p], column order (py, px, c) */
  void* dpred[MB_LOSS_MAX_G];       /* gradient rows or NULL */
  const unsigned char* mask;        /* group token mask [B, Lg] */
  const float* coef;                /* device scalar: w_mod / (sum_w * count_mod) * upstream */
  double* accum;                    /* [2]: sum of errors, number of masked elements */
  int B, D, C, S, p, G;
  int Lg, mask_off;
  int cg[MB_LOSS_MAX_G], c0[MB_LOSS_MAX_G];
  int n_norm;
  int norm_of_ch[MB_LOSS_MAX_C];
  int norm_size[MB_LOSS_MAX_C];
  int l2;
} mb_loss_params;                   /* `params` below is a HOST pointer */
/* SSLModule.compute_loss_rec, maestro/train/model.py:195-247 (+ norm_bands / loss_fn select :38-67), fwd + grad. */
int mb_loss_fwd_bwd(const mb_loss_params* params, int dtype, void* stream);

/* ---- contractions ------------------------------------------------------------------------------------------- */
#define MB_GEMM_GELU 1      /* out = gelu(acc + bias) (exact erf form); aux <- gelu'(acc + bias), saved for the backward */
#define MB_GEMM_DGELU 2     /* out = acc * aux (aux = the derivative a MB_GEMM_GELU call saved) */
#define MB_GEMM_ACCUM 4
#define MB_GEMM_OUT_F32 8
#define MB_GEMM_PAIR_ON 16  /* scheduling hint: force 2-CTA (cta_group::2) tiles on the tensor-core path */
#define MB_GEMM_PAIR_OFF 32 /* scheduling hint: force single-CTA tiles */
#define MB_GEMM_HALF_TAIL 64 /* scheduling hint (bf16 outputs): the last partial wave runs as half-width tiles */
typedef struct {
  const void* A;         /* A(m, k) = A[m * sam + k * sak] */
  const void* B;         /* B(n, k) = B[n * sbn + k * sbk] */
  void* C;               /* C[m * ldc + n] */
  const float* bias;
  const void* residual;
  void* aux;             /* [M, ldc] second output (MB_GEMM_GELU) / epilogue input (MB_GEMM_DGELU), activation dtype */
  int M, N, K;
  long long sam, sak, sbn, sbk;
  int ldc;
  int nb1, nb2;
  long long a_s1, a_s2, b_s1, b_s2, c_s1, c_s2;
  float alpha;
  int flags;
  float* colsum;         /* optional [N] fp32, accumulates the column sums of C (bias gradient of the layer whose dY
                            this GEMM produces, e.g. FeedForward up-projection bias from dgrad + GELU'); mb_gemm_tc* only */
} mb_gemm_params;                   /* `params` below is a HOST pointer */
/* nn.Linear / nn.Conv2d(k=p,s=p) / 1x1 conv and their autograd (vit_pytorch to_qkv / to_out / FeedForward,
 * maestro/layers/embed.py:59-60,148-154, maestro/ssl/mae.py:145-154) on the CUDA cores: fp32 check mode. */
int mb_gemm_simt(const mb_gemm_params* params, int dtype, void* stream);
/* The same contractions on tcgen05 tensor cores (bf16 in, fp32 accumulate in TMEM, TMA operand staging). */
int mb_gemm_tc(const mb_gemm_params* params, void* stream);
/* `count` (1..4) independent problems of one layer kind (same operand layouts, flags and alpha; M / N / K and all
 * pointers free) as ONE persistent launch: the per-group encoders / decoders of MAE.encode / MAE.decode
 * (maestro/ssl/mae.py:289-306, looped per group by encode_or_decode, maestro/ssl/mim.py:396-423) step in lockstep.
 * Layouts: A and B each K-major or MN-major (forward: K/K, dgrad: K/MN on the untransposed weight, wgrad: MN/MN). */
int mb_gemm_tc_grouped(const mb_gemm_params* params, int count, void* stream);
/* softmax(q k^T * scale) of vit_pytorch Attention.attend (check-mode path) and its gradient. */
int mb_softmax_rows(const float* s, void* p, int dtype, long long rows, int n, float scale, void* stream);
int mb_softmax_bwd_rows(const void* p, float* dp, void* ds, int dtype, long long rows, int n, float scale,
                        void* stream);

/* Fused flash-style attention on tcgen05 (bf16): softmax(q k^T * dh^-0.5) v of vit_pytorch Attention.forward without
 * materialising the [B, h, N, N] scores.  qkv [Bn*N, 3*H*dh], o [Bn*N, H*dh], lse / delta [Bn, H, N] fp32.
 * dh in {16, 32, 64}. */
int mb_attn_fwd(const void* qkv, void* o, float* lse, int Bn, int N, int H, int dh, void* stream);
int mb_attn_bwd(const void* qkv, const void* o, const void* d_o, const float* lse, void* dqkv, float* delta, int Bn,
                int N, int H, int dh, void* stream);

/* Grouped forms: `count` (1..4) attention problems of one head shape (H, dh) as ONE forward launch / TWO backward
 * launches -- the four per-group stacks of a layer (sequence lengths 50 .. 1024) run as one grid whose CTA ranges
 * belong to the problems in turn.  `problems` is a HOST pointer; d_o / dqkv / delta are read by the backward only. */
typedef struct {
  const void* qkv; void* o; float* lse;
  const void* d_o; void* dqkv; float* delta;
  int Bn, N;
} mb_attn_problem;
int mb_attn_fwd_grouped(const mb_attn_problem* problems, int count, int H, int dh, void* stream);
int mb_attn_bwd_grouped(const mb_attn_problem* problems, int count, int H, int dh, void* stream);

/* ---- device-side input pipeline ----------------------------------------------------------------- */
#define MB_RAW_U8 0
#define MB_RAW_I16 1
#define MB_RAW_U16 2
#define MB_RAW_F32 3
#define MB_AUG_FLIP_H 1     /* np.flip(axis=2) */
#define MB_AUG_FLIP_W 2     /* np.flip(axis=3) */
#define MB_AUG_TRANSPOSE 4  /* np.swapaxes(2, 3), applied after the flips */
/* Raster as stored on disk -> fp32 model input: maestro/dataset/dataset.py:215-222 (astype(float32),
 * log(max(x, 1e-10)) if log_scale, /= norm_fac if use_norm) and transform_rasters :225-257 (the dihedral
 * augmentation; aug[sample] holds the MB_AUG_* bits, NULL = none; image i belongs to sample i / imgs_per_sample).
 * raw [n_img, C, H, W] (or [.., W, H] for a transposed sample), out [n_img, C, H, W]; H == W when transposing. */
int mb_raster_transform(const void* raw, int raw_dtype, float* out, const unsigned char* aug, int n_img,
                        int imgs_per_sample, int C, int H, int W, float norm_fac, int use_norm, int log_scale,
                        void* stream);

#define MB_STRUCT_MAX_MODS 8
typedef struct {
  int G, D, L;                             /* band groups, dates per band group, grid^2 locations */
  float p_mod, p_bands, p_dates, p_loc;    /* <= 0: draw skipped (maestro/ssl/mae.py:199-208) */
} mb_struct_mod;
typedef struct {
  mb_struct_mod mods[MB_STRUCT_MAX_MODS];  /* modalities of the group in dataset.groups order */
  int n_mods;
  int L_group;                             /* sum of G * D * L */
  unsigned long long seed, offset;         /* Philox key / per-step counter offset */
} mb_struct_params;
/* MAE.mask_struct, maestro/ssl/mae.py:178-226, drawn on the device (performance mode; the reference's host
 * torch.rand order stays available on the host side): Philox4x32-10, counter (index, kind | mod << 8, sample,
 * offset + pass), u = (x >> 8) * 2^-24 < p.  One CTA per sample redraws (pass + 1) until its group is not entirely
 * masked.  `params` is a HOST pointer.  out [B, L_group] u8; passes [B] (may be NULL) = draws the sample took. */
int mb_struct_mask(const mb_struct_params* params, unsigned char* out, int* passes, int B, int max_pass,
                   void* stream);

/* ---- optimiser ---------------------------------------------------------------------------------------------- */
/* torch.optim.AdamW built at maestro/train/model.py:135-149, fused over the flat arena (+ bf16 shadow refresh). */
int mb_adamw_step(float* p, float* g, float* m, float* v, void* shadow, const float* hyper, size_t n, int zero_grad,
                  void* stream);

/* BaseModule.update_ema, maestro/train/base.py:263-274, over the flat arena: ema = momentum * ema + one_minus * param
 * (one_minus = float(1.0 - momentum), formed in double on the host as the reference does). */
int mb_ema_update(float* ema, const float* param, size_t n, float momentum, float one_minus, void* stream);

/* ---- probe / finetune heads -------------------------------------------------------------------- */
/* AttentiveReduce, maestro/layers/head.py:28-64: one learned query per head over each of S sequences of L tokens.
 * kv [S, L, 2*H*dh] (to_kv output: k then v), q [H*dh] fp32, out [S, H*dh], probs [S, H, L] fp32 (saved). */
int mb_attn_pool_fwd(const void* kv, const float* q, void* out, float* probs, int dtype, int S, int L, int H, int dh,
                     float scale, void* stream);
/* Its backward: dkv [S, L, 2*H*dh] is fully written, dq [H*dh] fp32 accumulates. */
int mb_attn_pool_bwd(const void* kv, const float* q, const float* probs, const void* dout, void* dkv, float* dq,
                     int dtype, int S, int L, int H, int dh, float scale, void* stream);
/* BaseMIM.compute_logits, maestro/ssl/mim.py:353-372: F.interpolate(bilinear) of channel-last token grids
 * x [n_img, hi*wi, d] -> out [n_img, ho*wo, d] (image n at x + n*in_bs / out + n*out_bs elements) through dense 1-D
 * weight matrices Wy [ho, hi], Wx [wo, wi]; the backward is the same call with the transposed matrices. */
int mb_token_resample(const void* x, const float* Wy, const float* Wx, void* out, int dtype, int n_img, int hi, int wi,
                      int ho, int wo, int d, long long in_bs, long long out_bs, void* stream);
/* BaseModule.compute_loss_pred, maestro/train/base.py:98-151 (F.cross_entropy over rows whose target is not the
 * missing value): logits (n, c) at (n / inner) * outer_stride + c * class_stride + n % inner; accum[0] += sum of row
 * losses, accum[1] += rows counted; dlogits (may be NULL) = softmax - onehot (0 for skipped rows). */
int mb_xent_fwd_bwd(const float* logits, const long long* target, float* dlogits, double* accum, long long N, int C,
                    long long inner, long long class_stride, long long outer_stride, long long missing, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MAESTRO_B200_H */
