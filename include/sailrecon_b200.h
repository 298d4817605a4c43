/* libsailrecon_b200.so — C-ABI of the B200 (sm_100a) hot path of SAIL-Recon.
 *
 * The reference (ShngJZ/self-supervise-sfm) has no FFI of its own: its "plugin API" for this path is
 * the nn.Module surface of sailrecon/layers/{attention,block,rope,mlp,layer_scale}.py,
 * sailrecon/models/aggregator.py and sailrecon/heads/camera_head.py, whose math lands in PyTorch library
 * calls.  Each entry point below replaces one of those library call sites (reference file:line cited);
 * the Python mirror of the module surface (package `sailrecon_b200`) binds them through ctypes.
 *
 * Conventions: plain C types only; all pointers are DEVICE pointers unless said otherwise; every launch is
 * asynchronous on `stream` (a cudaStream_t passed as void*); nothing is allocated, owned or freed by the
 * library; return value 0 = SR_OK, otherwise an SR_ERR_* code with a message in sr_last_error().
 * There is no CPU fallback: on a device that is not compute capability 10.x every compute call returns
 * SR_ERR_ARCH.
 */
#ifndef SAILRECON_B200_H_
#define SAILRECON_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SR_OK 0
#define SR_ERR_SHAPE 1 /* bad shape / alignment / argument                  */
#define SR_ERR_ARCH 2  /* device is not sm_100                              */
#define SR_ERR_CUDA 3  /* a CUDA runtime / driver call failed               */

/* ---- library ---------------------------------------------------------------------------------------- */
int sr_version(void);               /* 10000*major + 100*minor + patch                                       */
const char* sr_last_error(void);    /* thread-local message of the last failing call                          */
int sr_arch_check(void);            /* SR_OK iff current device is compute capability 10.x                    */
int sr_num_kernel_launches(void);   /* kernels launched by this library since load (bench.py gpu_launches)    */

/* ---- LayerNorm: nn.LayerNorm at block.py:50,70 (norm1/norm2), camera_head.py:64,65,77 ------------------
 * x fp32 [rows, C] (row stride ldx), weight/bias fp32 [C] (NULL = no affine, camera_head.py:77 adaln_norm).
 * out: bf16 (out_bf16=1; the cast autocast applies before the next Linear) or fp32, row stride ldo.       */
int sr_layernorm_fwd(const float* x, int64_t ldx, const float* weight, const float* bias, void* out, int out_bf16,
                     int64_t ldo, int64_t rows, int64_t C, float eps, void* stream);
/* Same with a fused row gather: out[r] = LN(x[idx[r]]) (idx int32 [rows], NULL = identity).  Replaces the
 * scene-token gather of aggregator.py:617-625 followed by global_reloc_blocks[i].norm1 (block.py:87).       */
int sr_layernorm_gather_fwd(const float* x, int64_t ldx, const int32_t* idx, const float* weight, const float* bias,
                            void* out, int out_bf16, int64_t ldo, int64_t rows, int64_t C, float eps, void* stream);

/* ---- GEMM + fused epilogue: nn.Linear at attention.py:73 (qkv), :120 (proj), mlp.py:35,38 (fc1, fc2) ----
 * out[M,N] = epilogue(A[M,K] @ W[N,K]^T + bias).  W is nn.Linear.weight as stored ([out,in], row-major). */
#define SR_EPI_BIAS 0          /* y = acc + bias                                 (proj of a plain Linear)   */
#define SR_EPI_BIAS_GELU 1     /* y = gelu_erf(round(acc + bias))                (mlp.py:35-36)             */
#define SR_EPI_LS_RESIDUAL 2   /* out = resid + gamma * round(acc + bias)        (block.py:110-111, fp32)   */
#define SR_EPI_BIAS_SILU 4     /* y = silu(acc + bias): embed_pose feeding poseLN_modulation's SiLU
                                  (camera_head.py:72-74,146-155); fp32 path only                           */
#define SR_EPI_QKV 3           /* bias, then per-head LayerNorm(d) on q,k (attention.py:78) and 2D RoPE
                                  (attention.py:80-82, rope.py:165-207); v = bias only                     */
typedef struct sr_epilogue {
  int32_t kind;          /* SR_EPI_*                                                                        */
  int32_t out_bf16;      /* 1: out is bf16 (tensor-core path rounds here exactly where autocast does)       */
  void* out;             /* [M, N], row stride ldo (elements)                                               */
  int64_t ldo;
  const float* bias;     /* [N] or NULL                                                                     */
  const float* gamma;    /* LS_RESIDUAL: LayerScale gamma [N] (NULL = 1)                                    */
  const float* resid;    /* LS_RESIDUAL: fp32 [M, N], row stride ldr; may alias out                         */
  int64_t ldr;
  /* QKV epilogue */
  int64_t qkv_dim;       /* C: columns [0,C) are q, [C,2C) k, [2C,3C) v of the full qkv Linear              */
  int64_t col_offset;    /* column of the full [3C] output that local column 0 corresponds to               */
  int32_t head_dim;      /* 64                                                                              */
  int32_t qk_norm;       /* apply LayerNorm(head_dim) to q and k                                            */
  const float* qn_w;     /* q_norm.weight/bias, k_norm.weight/bias, fp32 [head_dim]                         */
  const float* qn_b;
  const float* kn_w;
  const float* kn_b;
  float qk_eps;
  int32_t rope;          /* apply 2D RoPE to q and k                                                        */
  const float* rope_cos; /* fp32 [rope_max_pos, head_dim/4] (rope.py:110-123 without the duplication)       */
  const float* rope_sin;
  int32_t rope_max_pos;
  const int32_t* pos;    /* int32 [M, 2] (y, x) positions (rope.py:25-66 + aggregator.py:319-328)           */
  /* LS_RESIDUAL only: 1 = out = resid + gamma * (acc + bias) WITHOUT the bf16 rounding of the Linear output that
   * autocast performs (0, default) — the fp32 accumulate of the split-precision convolution passes (dpt.cu).   */
  int32_t exact_residual;
} sr_epilogue_t;

/* bf16 tensor-core path (tcgen05.mma, TMA-fed): A bf16 [M,K] row stride lda, W bf16 [N,K] row stride ldw.
 * Requirements: K % 64 == 0, N % 128 == 0, lda/ldw multiples of 8, 16-byte aligned pointers.              */
int sr_gemm_bf16(const void* A, int64_t lda, const void* W, int64_t ldw, int64_t M, int64_t N, int64_t K,
                 const sr_epilogue_t* epi, void* stream);
/* The same tensor-core kernel with operands as the Linear GRADIENTS find them in memory (no transposed copies):
 *   out[M, N] = op(A) W   with W bf16 [K, N] (row stride ldw; "MN-major": the contraction index runs over rows) and
 *   a_mn_major = 0: A bf16 [M, K] (row stride lda)      dX[T, K_in]    = dY[T, N_out] W[N_out, K_in]
 *   a_mn_major = 1: A bf16 [K, M] (row stride lda)      dW[N_out, K_in] = dY[T, N_out]^T X[T, K_in]   (K = T, any length)
 * Epilogues: SR_EPI_BIAS (bf16 or fp32 out) and SR_EPI_LS_RESIDUAL (with exact_residual = 1: fp32 accumulate into out).
 * Requirements: N % 128 == 0; M % 64 == 0 (a_mn_major = 1) or K % 64 == 0 (a_mn_major = 0); b_mn_major must be 1.   */
int sr_gemm_bf16_t(const void* A, int64_t lda, int32_t a_mn_major, const void* W, int64_t ldw, int32_t b_mn_major,
                   int64_t M, int64_t N, int64_t K, const sr_epilogue_t* epi, void* stream);
/* fp32 SIMT path (fp32-check mode and the autocast-disabled camera head, sail_recon.py:119): no rounding.  */
int sr_gemm_f32(const float* A, int64_t lda, const float* W, int64_t ldw, int64_t M, int64_t N, int64_t K,
                const sr_epilogue_t* epi, void* stream);

/* ---- Attention: F.scaled_dot_product_attention at attention.py:103-109 --------------------------------
 * Structured instead of dense-masked.  `batch` independent segments; query rows of segment b are
 * [b*q_batch_stride, b*q_batch_stride + q_len) of q, they attend to the shared prefix rows [0, prefix_len)
 * of (pk, pv) followed by the segment's own rows [b*kv_batch_stride, +kv_len) of (k, v).
 *   frame blocks   (aggregator.py:643-670): batch = S, q_len = kv_len = P, no prefix
 *   global blocks  (aggregator.py:743-769): batch = 1, q_len = kv_len = Na*P
 *   reloc blocks   (aggregator.py:672-741 + mask :302-311,:832-851): batch = Nq, own frame + scene prefix
 *   camera trunk   (camera_head.py:163-166 + build_lr_mask :197-228): anchors = 1 segment; queries = Nq
 *                   segments of length 1 with the anchors as prefix
 * Head h of a row lives at columns [h*head_dim, (h+1)*head_dim).  out has the layout the proj Linear
 * consumes (attention.py:119): [rows, num_heads*head_dim].                                               */
typedef struct sr_attn {
  const void* q;  int64_t ldq;
  const void* k;  const void* v;  int64_t ldkv;
  const void* pk; const void* pv; int64_t ldp; int64_t prefix_len;   /* pk == NULL or prefix_len == 0: none */
  void* out;      int64_t ldo;
  int32_t num_heads, head_dim;
  int32_t batch;
  int64_t q_len, q_batch_stride;
  int64_t kv_len, kv_batch_stride;
  float scale;           /* head_dim^-0.5                                                                  */
  /* ring / split-KV carry (global attention sharded over ranks; all NULL for a self-contained call):
   * running row max m and row sum l in the log2 domain, fp32 [num_heads, total_q_rows], and the
   * un-normalised fp32 accumulator acc [total_q_rows, num_heads*head_dim].                               */
  float* carry_m; float* carry_l; float* carry_acc;
  int32_t carry_in;      /* 1: start from the carried state instead of (-inf, 0, 0)                        */
  int32_t carry_out;     /* 1: write the state back and skip the final normalisation + out store           */
  /* training: row-wise log-sum-exp of the scaled scores in the log2 domain, fp32 [num_heads, batch, lse_pad]
   * (lse_pad >= q_len, a multiple of 64; the caller pre-fills the buffer with +inf so padding rows stay +inf).
   * NULL = inference.  Consumed by sr_attn_bwd_bf16.  bf16 tensor-core path only.                           */
  float* lse; int64_t lse_pad;
  /* kernel variant of the tcgen05 path: 0 = default (four-tile kernel, mixed four- / three-tile CTAs, grid order chosen per
   * launch), 3 = three-tile kernel, 4 = uniform four-tile CTAs, 5 / 6 = mixed CTAs with the heavy-first / row-major grid
   * order forced (A/B timing).  Every variant computes the same function; ignored by the fp32 path.          */
  int32_t variant;
  /* optional scratch for split-KV (under-filled launches: few query rows against a long K/V range, e.g. one query frame
   * against a cached 100-anchor scene): sr_attn_workspace_bytes() says how much this call can use (0 = none); NULL or
   * too small = the call runs un-split.  fp32 partial (m, l, O) per K/V chunk, combined by a second small kernel.      */
  void* workspace; int64_t workspace_bytes;
} sr_attn_t;

/* Profiling aid, -DSR_DEBUG builds of the library only (the shipped build never reads the environment and returns an
 * error here): with SR_ATTN_TRACE set, CTA (0,0,0) of sr_attn_fwd_bf16 records clock64() at its pipeline hand-offs for
 * the first 64 K/V tiles; this copies the [3 roles][64 tiles][8 events] table to HOST memory (synchronises the device). */
int sr_debug_trace_read(long long* host_dst, int64_t count);

int sr_attn_fwd_bf16(const sr_attn_t* p, void* stream);  /* tcgen05 flash attention, head_dim 64           */
/* Host-only query (no GPU needed): the work decomposition sr_attn_fwd_bf16 uses for a launch of num_heads x batch
 * (head, segment) rows of q_len query rows on num_sms SMs — n4 CTAs of four 128-row query tiles and n3 CTAs of three per
 * row; all four-tile CTAs are dispatched first, so the lighter ones fill the tail wave.                           */
/* Host-only: bytes of sr_attn_t.workspace the bf16 path would use for this call on a device with num_sms SMs (0 = no split). */
int64_t sr_attn_workspace_bytes(const sr_attn_t* p, int32_t num_sms);
int sr_attn_plan(int64_t q_len, int32_t num_heads, int32_t batch, int32_t num_sms, int32_t* n4, int32_t* n3);
int sr_attn_fwd_f32(const sr_attn_t* p, void* stream);   /* fp32 SIMT, head_dim <= 128 (check mode, camera) */

/* ---- Attention backward (training step, train_imc.py:358-380: autograd through attention.py:103-109) -----
 * Same structured layout as sr_attn.  Inputs: q, k, v (and the prefix pk, pv) as given to the forward, the
 * forward output `out`, its gradient `dout` (bf16, head h at columns [64h, 64h+64)), and the forward's `lse`.
 * `delta` is workspace of the shape of lse (row-wise sum(out * dout) per head, written here first).
 * Outputs (bf16, same row / head-column layout as their inputs; they may be column slices of one [rows, 3C]
 * buffer): dq, dk, dv, and — when dpk != NULL — dpk, dpv [prefix_len rows] (the prefix is shared by all `batch`
 * segments, its gradient is the sum over them).  Two tcgen05 passes: query-stationary CTAs produce dq
 * (S and dP recomputed per K/V tile, dS kept in TMEM as the A operand of dq += dS K), key-stationary CTAs
 * produce dk, dv (S^T, dP^T in TMEM; dv += P^T dO, dk += dS^T Q).  No atomics: results are deterministic.   */
typedef struct sr_attn_bwd {
  const void* q;  int64_t ldq;
  const void* k;  const void* v;  int64_t ldkv;
  const void* pk; const void* pv; int64_t ldp; int64_t prefix_len;
  const void* out;  int64_t ldo;
  const void* dout; int64_t lddo;
  const float* lse; float* delta; int64_t lse_pad;
  void* dq;  int64_t lddq;
  void* dk;  void* dv;  int64_t lddkv;
  void* dpk; void* dpv; int64_t lddp;
  int32_t num_heads, head_dim;
  int32_t batch;
  int64_t q_len, q_batch_stride;
  int64_t kv_len, kv_batch_stride;
  float scale;
} sr_attn_bwd_t;
int sr_attn_bwd_bf16(const sr_attn_bwd_t* p, void* stream);

/* ---- Training-step row kernels (autograd of block.py:86-112; csrc/train_rows.cu) -------------------------
 * The Linear gradients run on sr_gemm_bf16 itself: dX = dY W is sr_gemm_bf16(dY, W^T), dW = dY^T X is
 * sr_gemm_bf16(dY^T, X^T) with fp32 output; sr_transpose_bf16 makes the transposed bf16 operands
 * (dst [C, ldd >= R_pad], columns [R, R_pad) zero-filled so that the contraction length is a multiple of 64).
 * colsum (fp32 [C], NULL = skip): colsum[c] += sum_r src[r, c] in the same pass — the bias gradient of the Linear whose
 * dY is being transposed (needs the 16-byte aligned layout: R_pad % 64 == 0; C, lds, ldd multiples of 8).        */
int sr_transpose_bf16(const void* src, int64_t lds, void* dst, int64_t ldd, int64_t R, int64_t C, int64_t R_pad, float* colsum,
                      void* stream);
/* out[c] (+)= sum_r a[r,c] * b[r,c]   (a fp32 or bf16, b bf16 or NULL = 1): bias and LayerScale gamma gradients */
int sr_colsum(const void* a, int a_bf16, int64_t lda, const void* b, int64_t ldb, float* out, int64_t R, int64_t C,
              int accumulate, void* stream);
/* out = bf16(x * gamma[c]): gradient entering the Linear under a LayerScale (layer_scale.py:22)               */
int sr_scale_cast_bf16(const float* x, int64_t ldx, const float* gamma, void* out, int64_t ldo, int64_t rows, int64_t C,
                       void* stream);
/* out = resid + gamma * z (z bf16 = the Linear's rounded output, kept for d gamma); contiguous [rows, C]       */
int sr_ls_residual_f32(const float* resid, const void* z, const float* gamma, float* out, int64_t rows, int64_t C, void* stream);
/* mlp.py:36 on the saved pre-activation u (bf16): g = gelu_erf(u);  du = dg * gelu'(u)                         */
int sr_gelu_fwd_bf16(const void* u, void* out, int64_t n, void* stream);
int sr_gelu_bwd_bf16(const void* u, const void* dg, void* du, int64_t n, void* stream);
/* Fused forms (one pass over the rows, column sums by shared-memory partials + one atomic per column and block):
 * sr_ls_bwd_bf16: LayerScale backward — dz = bf16(dy * gamma) (the gradient entering the Linear under the LayerScale),
 *   dgamma[c] += sum_r dy * z, dbias[c] += sum_r float(dz) (the Linear's bias gradient; NULL = skip).  dy fp32, z / dz bf16.
 * sr_gelu_bwd_colsum_bf16: du = bf16(dg * gelu'(u)) (du may alias dg) and dbias[c] += sum_r float(du) (fc1's bias gradient). */
int sr_ls_bwd_bf16(const float* dy, int64_t lddy, const void* z, int64_t ldz, const float* gamma, void* dz, int64_t lddz,
                   float* dgamma, float* dbias, int64_t rows, int64_t C, void* stream);
int sr_gelu_bwd_colsum_bf16(const void* u, int64_t ldu, const void* dg, int64_t lddg, void* du, int64_t lddu, float* dbias,
                            int64_t rows, int64_t C, void* stream);
/* LayerNorm gradient (block.py norm1 / norm2, C = 1024): dx = dres + dLN(dy), where dres (fp32) is the gradient arriving
 * over the residual connection (NULL = dx itself, i.e. dx += dLN(dy); dres may alias dx); dweight / dbias (fp32 [C],
 * accumulated with atomics; NULL = skip).  x fp32, dy bf16.                                                         */
int sr_layernorm_bwd(const float* x, int64_t ldx, const float* weight, const void* dy, int64_t lddy, const float* dres,
                     int64_t lddres, float* dx, int64_t lddx, float* dweight, float* dbias, int64_t rows, int64_t C, float eps,
                     void* stream);
/* attention.py:78-82 on a saved raw qkv buffer (bf16 [M, 3C], the qkv Linear's rounded output): per-head
 * LayerNorm(64) + 2D RoPE on q and k, v copied — the un-fused form of SR_EPI_QKV (fields qkv_dim, head_dim,
 * qk_norm, q/k norm parameters, qk_eps, rope tables, pos of `qkv` are read).  _bwd turns the gradient w.r.t. the
 * normalised / rotated q, k (in place in dqkv; the v third is left as is) into the gradient w.r.t. raw and
 * accumulates the q_norm / k_norm parameter gradients (fp32 [64] each, NULL = skip).                             */
int sr_qknorm_rope_fwd(const void* raw, int64_t ldr, void* out, int64_t ldo, const sr_epilogue_t* qkv, int64_t M, void* stream);
int sr_qknorm_rope_bwd(const void* raw, int64_t ldr, void* dqkv, int64_t ldd, const sr_epilogue_t* qkv, float* dqn_w,
                       float* dqn_b, float* dkn_w, float* dkn_b, int64_t M, void* stream);

/* dst[idx[r], :] += src[r, :], idx unique: gradient of the scene-token row gather (aggregator.py:617-625)        */
int sr_scatter_add_rows_f32(float* dst, int64_t ldd, const int32_t* idx, const float* src, int64_t lds, int64_t n, int64_t C,
                            void* stream);

/* torch.optim.Adam(betas, eps) as train/train_imc.py:475 configures it (no weight decay, no amsgrad), one step on a flat
 * fp32 bucket of n parameters (n % 4 == 0, 16-byte aligned): g *= grad_scale; m = b1 m + (1 - b1) g; v = b2 v + (1 - b2) g^2;
 * p -= lr / (1 - b1^step) * m / (sqrt(v) / sqrt(1 - b2^step) + eps).  step counts from 1.                         */
int sr_adam_step_f32(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t n, double lr, double beta1,
                     double beta2, double eps, int64_t step, double grad_scale, void* stream);

/* ---- token re-layout (aggregator.py:355-357,393-399,406-423,603-625) ----------------------------------- */
/* dst[i, :] = src[idx[i], :] for i < n; fp32 rows of C elements; idx int32 (scene-token selection).       */
int sr_gather_rows_f32(const float* src, int64_t lds, const int32_t* idx, float* dst, int64_t ldd, int64_t n,
                       int64_t C, void* stream);
/* dst[r, col_off : col_off + C] = src[r, :] — writes one half of the [.., 2C] concat outputs.             */
int sr_copy_cols_f32(const float* src, int64_t lds, float* dst, int64_t ldd, int64_t col_off, int64_t rows,
                     int64_t C, void* stream);
/* fp32 -> bf16 cast of a [rows, C] matrix (weights, once).                                                */
int sr_cast_f32_to_bf16(const float* src, int64_t lds, void* dst, int64_t ldd, int64_t rows, int64_t C,
                        void* stream);
/* camera_head.py:157-160: out = gate * (LN_noaffine(x) * (1 + scale) + shift) + x, all fp32 [rows, C];
 * mod = [rows, 3C] holding (shift | scale | gate).                                                         */
int sr_adaln_modulate_f32(const float* x, const float* mod, float* out, int64_t rows, int64_t C, float eps,
                          void* stream);


/* LayerScale.forward on its own (layer_scale.py:23): out[r, c] = x[r, c] * gamma[c]; contiguous fp32 (inside a
 * Block the scale is folded into the GEMM epilogue instead).                                                  */
int sr_scale_cols_f32(const float* x, const float* gamma, float* out, int64_t rows, int64_t C, void* stream);

/* ---- 2D RoPE (rope.py:69-207) --------------------------------------------------------------------------- */
/* cos/sin fp32 [max_pos, quarter_dim]: entry (p, f) = cos/sin(p * frequency^(-f/quarter_dim)) (rope.py:107-123
 * before the duplication of :119; the kernels index the duplicate by f mod quarter_dim).                    */
int sr_rope_table_build(float* cos_out, float* sin_out, int32_t max_pos, int32_t quarter_dim, float frequency,
                        void* stream);
/* RotaryPositionEmbedding2D.forward (rope.py:165-207): x, out fp32 [B, H, N, d] contiguous; pos int32 [B, N, 2]. */
int sr_rope2d_f32(const float* x, float* out, const int32_t* pos, const float* rope_cos, const float* rope_sin,
                  int32_t max_pos, int64_t B, int64_t H, int64_t N, int64_t d, void* stream);

/* ---- patch-embed front end (aggregator.py:267-271, patch_embed.py:62-84) --------------------------------- */
/* out[(s*h + py)*w + px, (c*patch + ky)*patch + kx] = (images[s, c, py*patch+ky, px*patch+kx] - mean[c]) / std[c];
 * columns [3*patch*patch, ldo) are zero-filled (K padding for the tensor-core GEMM).  images fp32 [S,3,H,W];
 * out bf16 or fp32 [S*h*w, ldo]; mean3/std3 are HOST pointers to 3 floats.                                    */
int sr_patchify(const float* images, void* out, int out_bf16, int64_t ldo, int64_t S, int64_t H, int64_t W,
                int32_t patch, const float* mean3, const float* std3, void* stream);

/* ---- camera head pose accumulation (camera_head.py:172-184, head_act.py:12-37) -----------------------------
 * pred = first ? delta : pred + delta; activated = [pred[:, :7], relu(pred[:, 7:9])]; all fp32 [rows, 9].     */
int sr_pose_update(float* pred, const float* delta, float* activated, int64_t rows, int32_t first, void* stream);

/* ---- camera geometry after the head ------------------------------------------------------
 * pose_encoding_to_extri_intri (utils/pose_enc.py:68-135, "absT_quaR_FoV") with quat_to_mat (utils/rotation.py:14-44,
 * scalar-last quaternion) fused: pose fp32 [n, 9] = (T[3], quat xyzw[4], fov_h, fov_w) -> extrinsics fp32 [n, 3, 4] =
 * [R | T] and, when intrinsics != NULL, intrinsics fp32 [n, 3, 3] with fy = (H/2) / tan(fov_h/2), fx = (W/2) / tan(fov_w/2),
 * principal point (W/2, H/2).  Called at sail_recon.py:122-124, :235-237.                                      */
int sr_pose_to_extri_intri(const float* pose, float* extrinsics, float* intrinsics, int64_t n, int32_t image_h,
                           int32_t image_w, void* stream);
/* its backward (the training loss differentiates the predicted cameras, train_imc.py:156,175): d pose [n, 9] from
 * d extrinsics [n, 3, 4] and d intrinsics [n, 3, 3] (NULL = none; only the two focal entries carry gradient).   */
int sr_pose_to_extri_intri_bwd(const float* pose, const float* d_extrinsics, const float* d_intrinsics, float* d_pose, int64_t n,
                               int32_t image_h, int32_t image_w, void* stream);
/* unproject_depth_map_to_point_map (utils/geometry.py:19-128; numpy on the host in the reference, sail_recon.py:139-141):
 * depth fp32 [S, H, W], extrinsics [S, 3, 4] (camera from world), intrinsics [S, 3, 3] (zero skew) -> world points
 * fp32 [S, H, W, 3] = R^T ((u - cu) d / fu, (v - cv) d / fv, d) - R^T t.  Everything stays in HBM.              */
int sr_unproject_depth(const float* depth, const float* extrinsics, const float* intrinsics, float* points, int64_t S,
                       int64_t H, int64_t W, void* stream);
/* modulate (heads/camera_head.py:189-194): out = x * (1 + scale) + shift, fp32 [rows, C] with row strides (shift / scale
 * are usually column blocks of the [rows, 3C] modulation tensor).                                              */
int sr_modulate_f32(const float* x, int64_t ldx, const float* shift, int64_t lds, const float* scale, int64_t ldc,
                    float* out, int64_t ldo, int64_t rows, int64_t C, void* stream);

/* ---- DPT dense heads (heads/dpt_head.py:22-344), channels-last fp32 row matrices -------------
 * Convolutions run on sr_gemm_bf16 with split-precision operands (see csrc/dpt.cu): this writes, for x fp32 [N, H, W, C]
 * (pixel stride ldx), the im2col rows out bf16 [N*Ho*Wo, 2*k*k*C] = [hi | lo], hi = bf16(x), lo = bf16(x - hi), column
 * (ky*k + kx)*C + c; k = 1 (stride 1, pad 0: a plain split) or 3; relu = 1 applies max(x, 0) first.                */
int sr_im2col_split_bf16(const float* x, int64_t ldx, void* out, int64_t N, int64_t H, int64_t W, int64_t C, int32_t k,
                         int32_t stride, int32_t pad, int32_t relu, void* stream);
/* 3x3 / stride 1 / padding 1 convolution WITHOUT an im2col matrix ("shifted GEMM"): xs = sr_split_pad_bf16 output, the
 * zero-bordered map as rows [S*(H+2)*(W+2), hi(C) | lo(C)]; output row r (same padded row numbering; border rows hold
 * garbage and are ignored / re-zeroed by the consumers) = sum over taps of xs[r + (ky-1)*(W+2) + (kx-1)] x W_tap.  w bf16
 * [N, 27*C] = (Whi | Whi | Wlo), each [N, (ky, kx, c)]: the three split-precision parts run inside ONE launch, fp32
 * accumulation.  epi: SR_EPI_BIAS with fp32 out, or SR_EPI_LS_RESIDUAL (exact_residual) adding a residual in padded rows.   */
int sr_conv3x3_bf16(const void* xs, int64_t ld_xs, int64_t C, int64_t padded_width, const void* w, int64_t ldw, int64_t rows,
                    int64_t N, const sr_epilogue_t* epi, void* stream);
/* x fp32 channels-last [S, H + 2*in_pad, W + 2*in_pad, ldx] -> out bf16 [S, H+2, W+2, hi(C) | lo(C)], zero border; relu first */
int sr_split_pad_bf16(const float* x, int64_t ldx, int32_t in_pad, void* out, int64_t S, int64_t H, int64_t W, int64_t C,
                      int32_t relu, void* stream);
/* out = act(a + b) between maps [S, H, W, C] that may carry the one-pixel border (pad_* in {0, 1}); b may be NULL; border
 * pixels of a padded output are written as 0; relu = 1: act = ReLU                                                            */
int sr_map_add_f32(const float* a, int32_t pad_a, const float* b, int32_t pad_b, float* out, int32_t pad_out, int64_t S, int64_t H,
                   int64_t W, int64_t C, int32_t relu, void* stream);
/* F.interpolate(mode="bilinear", align_corners=True) (custom_interpolate, dpt_head.py:565-594), x [N,Hi(+2),Wi(+2),C] (in_pad = 1:
 * the source carries the border) -> dense [N,Ho,Wo,C]                                                                           */
int sr_bilinear_nhwc_f32(const float* x, float* out, int64_t N, int64_t Hi, int64_t Wi, int64_t Ho, int64_t Wo, int64_t C,
                         int32_t in_pad, void* stream);
/* ConvTranspose2d(kernel = stride = k) epilogue (dpt_head.py:85-102): g [N*H*W, k*k*C] (GEMM output, columns (ky, kx, c))
 * -> out [N, H*k, W*k, C] + bias[c] (NULL = none)                                                                  */
int sr_pixel_shuffle_f32(const float* g, const float* bias, float* out, int64_t N, int64_t H, int64_t W, int64_t C, int32_t k,
                         void* stream);
/* out[i] = act(a[i] + alpha * b[i mod period]): skip additions (period = n) and the broadcast positional embedding;
 * b may be NULL; relu_out = 1: act = ReLU (the in-place activation of the ResidualConvUnit, dpt_head.py:357,461)     */
int sr_add_bcast_f32(const float* a, const float* b, float* out, int64_t n, int64_t period, float alpha, int32_t relu_out,
                     void* stream);
/* create_uv_grid + position_grid_to_embed (heads/utils.py:11-123) evaluated in double: out fp32 [H, W, C]            */
int sr_uv_posembed_table(float* out, int64_t H, int64_t W, int64_t C, double aspect_ratio, double omega0, void* stream);
/* ReLU -> 1x1 conv (hid -> out_dim) -> activate_head (dpt_head.py:150-160, head_act.py:63-114): h [px, ldh] -> pred
 * [px, out_dim - 1], conf [px].  act: 0 inv_log, 1 exp, 2 linear, 3 relu, 4 sigmoid; conf_act: 0 expp1, 1 expp0, 2 sigmoid */
int sr_dpt_out_head_f32(const float* h, int64_t ldh, const float* w, const float* bias, float* pred, float* conf, int64_t px,
                        int32_t hid, int32_t out_dim, int32_t act, int32_t conf_act, int32_t map_h, int32_t map_w, int32_t pad,
                        void* stream);   /* pad = 1: h is a zero-bordered map [S, map_h + 2, map_w + 2, ldh] */

/* ---- self-supervised training loss (train/train_imc.py:141-256) ----------------------------------
 * All fp32, contiguous.  P = correspondence pairs, K = sampled points per pair, N = frames ("nodes").
 * `rows` = 3 for [.., 3, 4] poses (pad_poses appends (0, 0, 0, 1), train/utils/geometry.py:36-57) or 4 for [.., 4, 4].     */

/* compute_relative_pose (train/utils/geometry.py:252-287): relative [P, 4, 4] = dst @ inverse(src), and its backward.    */
int sr_relative_pose_f32(const float* src_extrinsic, const float* dst_extrinsic, int32_t rows, float* relative, int64_t num_pairs,
                         void* stream);
int sr_relative_pose_bwd_f32(const float* src_extrinsic, const float* dst_extrinsic, int32_t rows, const float* d_relative,
                             float* d_src, float* d_dst, int64_t num_pairs, void* stream);
/* backproject_and_reproject (train/utils/geometry.py:89-158; dst_depth == NULL: perspective division by the projected
 * depth + 1e-6) and backproject_and_reproject_with_approximation (:160-250; dst_depth [P, K]: division by dst_depth *
 * dst_scale + 1e-6).  src_coords [P, K, 2], src_depth [P, K], src_k / dst_k [P, 3, 3], relative [P, rows, 4], src_scale /
 * dst_scale [P] or NULL (= 1) -> out_coords [P, K, 2].  workspace: sr_reproject_workspace_floats(P, K) floats.
 * The backward returns d src_k, d dst_k [P, 3, 3] and d relative [P, rows, 4] from d out_coords [P, K, 2].                */
int64_t sr_reproject_workspace_floats(int64_t num_pairs, int64_t num_points);
int sr_reproject_f32(const float* src_coords, const float* src_depth, const float* dst_depth, const float* src_k,
                     const float* dst_k, const float* relative, int32_t rows, const float* src_scale, const float* dst_scale,
                     float* out_coords, int64_t num_pairs, int64_t num_points, float* workspace, void* stream);
int sr_reproject_bwd_f32(const float* src_coords, const float* src_depth, const float* dst_depth, const float* src_k,
                         const float* dst_k, const float* relative, int32_t rows, const float* src_scale, const float* dst_scale,
                         const float* d_coords, float* d_src_k, float* d_dst_k, float* d_relative, int64_t num_pairs,
                         int64_t num_points, float* workspace, void* stream);
/* No counterpart in the reference (PyTorch allocates whatever it needs).  sr_reproject_f32 for callers that cannot hand
 * over a workspace because they are not allowed to allocate device memory: a plugin inside another engine's closed memory
 * pool, a localisation loop that must not touch cudaMalloc once it is running, a worker that owns nothing but the buffers
 * passed to it.  Same arguments minus `workspace`, same kernels, same result bit for bit.  The per-pair camera records go
 * into a fixed 512 KiB array built into the library (held on every GPU the library is loaded on, until the process ends),
 * so num_pairs <= sr_reproject_static_max_pairs() (4096); more is SR_ERR_SHAPE with a message, never a partial result.
 * The call itself allocates nothing, on either side.  Its first use on a GPU looks the array's address up (do it once at
 * start-up with that GPU current, before capturing any stream).  All calls on one GPU share the array: calls that may run
 * concurrently there corrupt each other, so serialise them on a stream.  Nothing has to be cleared between calls.          */
int64_t sr_reproject_static_max_pairs(void);
int sr_reproject_static_f32(const float* src_coords, const float* src_depth, const float* dst_depth, const float* src_k,
                            const float* dst_k, const float* relative, int32_t rows, const float* src_scale,
                            const float* dst_scale, float* out_coords, int64_t num_pairs, int64_t num_points, void* stream);
/* compute_projective_residual (train/utils/geometry.py:289-303): residual[i] = || predicted[i] - actual[i] ||_2 over n points
 * of 2 coordinates; backward: d predicted = d residual * (predicted - actual) / residual (0 where the residual is 0).      */
int sr_projective_residual_f32(const float* predicted, const float* actual, float* residual, int64_t n, void* stream);
int sr_projective_residual_bwd_f32(const float* predicted, const float* actual, const float* d_residual, float* d_predicted,
                                   int64_t n, void* stream);
/* CDFLossIndexPytorch.compute_weighted_pdf_cdf_vectorized (train/losses/cdf_loss.py:122-186): residuals, weights [P, K]
 * (weights NULL = 1), src_idx / dst_idx int64 [P] -> pmf, cdf, pdf [num_nodes, num_bins].  bin = trunc((r - min_val) *
 * inv_bin_width), inv_bin_width = fp32(1 / bin_width) — torch's GPU division by a Python scalar IS this multiplication
 * (profiles/r02_div_probe.txt); sobel_weight = fp32(1 / (2 bin_width));
 * smooth_kernel: device fp32 [2 * smooth_radius + 1] (reflect padding).  hist_total: scratch, num_nodes * (num_bins + 1).  */
int sr_cdf_tables_f32(const float* residuals, const float* weights, const int64_t* src_idx, const int64_t* dst_idx,
                      int64_t num_pairs, int64_t num_points, int64_t num_nodes, float min_val, float inv_bin_width,
                      int32_t num_bins, float sobel_weight, const float* smooth_kernel, int32_t smooth_radius, float* hist_total,
                      float* pmf, float* cdf, float* pdf, void* stream);
/* compute_weighted_cdf_forward_backward_vectorized (train/losses/cdf_loss.py:188-242): bin = trunc((r - min_val) *
 * inv_bin_width + 0.5); valid = in range and weight > 0; out = cdf[frame, bin] (2.0 when invalid), grad = pdf[frame, bin] *
 * weight (0 when invalid), for the pair's source and destination frame.                                                  */
int sr_cdf_lookup_f32(const float* residuals, const float* weights, const int64_t* src_idx, const int64_t* dst_idx,
                      const float* cdf, const float* pdf, int64_t num_pairs, int64_t num_points, float min_val,
                      float inv_bin_width, int32_t num_bins, float* out_src, float* grad_src, float* out_dst, float* grad_dst,
                      void* stream);
/* compute_loss (train/train_imc.py:141-256) in one call: loss AND its gradient w.r.t. the predicted cameras.             */
typedef struct {
  const float* pred_intrinsics;   /* [N, 3, 3]    cat(pred['intrinsic']), train_imc.py:156                              */
  const float* pred_extrinsics;   /* [N, rows, 4] cat(pred['extrinsic']), :175                                          */
  const float* kprime_to_k;       /* [N, 3, 3] batch['K_prime_to_K'] (:162-165) or NULL = identity                      */
  const int64_t* src_idx;         /* [P] */
  const int64_t* dst_idx;         /* [P] */
  const float* src_coords;        /* [P, K, 2] */
  const float* dst_coords;        /* [P, K, 2] */
  const float* src_depth;         /* [P, K] */
  const float* dst_depth;         /* [P, K] */
  int64_t num_frames, num_pairs, num_points;
  int32_t pose_rows;              /* 3 or 4 */
  int32_t shared_focal;           /* batch['shared_focal'] (:168-172) */
  float min_val, inv_bin_width, sobel_weight;
  int32_t num_bins, smooth_radius;
  const float* smooth_kernel;     /* device fp32 [2 * smooth_radius + 1] */
  float* workspace;               /* sr_reproj_loss_workspace_floats(N, P, K, num_bins) floats */
  int64_t workspace_floats;
  float* loss;                    /* [1] */
  float* d_intrinsics;            /* [N, 3, 3] */
  float* d_extrinsics;            /* [N, rows, 4] */
  float* tables;                  /* optional [3 (pmf, cdf, pdf)][2 (exact, approximated)][N][num_bins], or NULL */
} sr_reproj_loss_t;
int64_t sr_reproj_loss_workspace_floats(int64_t num_frames, int64_t num_pairs, int64_t num_points, int32_t num_bins);
int sr_reproj_cdf_loss_f32(const sr_reproj_loss_t* args, void* stream);

/* ---- camera-head backward, fp32 (sailrecon/heads/camera_head.py:85-186 under autograd; csrc/head_train.cu) -------------
 * The Linear gradients reuse sr_gemm_f32: dX = dY W on a transposed copy of W, dW = dY^T X with the contraction over the T
 * camera-token rows (both operands transposed copies).                                                                    */
int sr_transpose_f32(const float* src, int64_t lds, float* dst, int64_t ldd, int64_t R, int64_t C, void* stream);
/* out[c] (+)= sum_r a[r, c] * b[r, c] (b NULL = 1): bias gradients and d LayerScale gamma = sum_r dy * z (layer_scale.py:22) */
int sr_colsum_f32(const float* a, int64_t lda, const float* b, int64_t ldb, float* out, int64_t R, int64_t C, int32_t accumulate,
                  void* stream);
/* kind 0: nn.GELU (erf form, mlp.py:23,36); kind 1: nn.SiLU (camera_head.py:72).  out = act(u);  du = dg * act'(u)         */
int sr_act_f32(const float* u, float* out, int64_t n, int32_t kind, void* stream);
int sr_act_bwd_f32(const float* u, const float* dg, float* du, int64_t n, int32_t kind, void* stream);
/* out = resid + gamma[c] * z (gamma NULL = 1), all fp32 contiguous [rows, C] (block.py:110-111 un-fused: z is kept)         */
int sr_ls_residual_f32f32(const float* resid, const float* z, const float* gamma, float* out, int64_t rows, int64_t C, void* stream);
/* nn.LayerNorm gradient with an fp32 incoming gradient: dx = (dres ? dres : 0) + dLN(dy); dweight / dbias [C] are ADDED to
 * (NULL = skip; weight NULL = no affine).  x, dy, dres, dx contiguous [rows, C].                                            */
int sr_layernorm_bwd_f32(const float* x, const float* weight, const float* dy, const float* dres, float* dx, float* dweight,
                         float* dbias, int64_t rows, int64_t C, float eps, void* stream);
/* backward of sr_adaln_modulate_f32 (camera_head.py:151-161): d mod [rows, 3C] = [d shift | d scale | d gate] and dx
 * (accumulate_dx = 1: dx += ...).                                                                                          */
int sr_adaln_modulate_bwd_f32(const float* x, const float* mod, const float* dy, float* dx, float* dmod, int64_t rows, int64_t C,
                              float eps, int32_t accumulate_dx, void* stream);
/* attention backward over T <= 128 rows with the build_lr_mask structure (camera_head.py:197-228): row r sees column c iff
 * c < n_anchor or c == r.  q, k, v row views (ld ldqkv), dout (lddo) -> dq, dk, dv (ldd); head h at columns [h * D, +D).     */
int sr_attn_small_bwd_f32(const float* q, const float* k, const float* v, int64_t ldqkv, const float* dout, int64_t lddo, float* dq,
                          float* dk, float* dv, int64_t ldd, int64_t T, int32_t num_heads, int32_t head_dim, int64_t n_anchor,
                          float scale, void* stream);
/* activate_pose backward (head_act.py:12-37 as SAIL-Recon configures it): d pred = d activated, zero where a FoV entry
 * (columns 7, 8 of the 9) was clipped by the relu.  pred is the PRE-activation encoding [rows, 9].                           */
int sr_pose_act_bwd_f32(const float* pred, const float* d_activated, float* d_pred, int64_t rows, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SAILRECON_B200_H_ */
