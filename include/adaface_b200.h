/*
 * adaface_b200 C ABI -- the drop-in boundary of the B200-native AdaPrompt U-Net transformer path.
 *
 * The reference (zyt334/AdaFace) is pure Python/PyTorch and has no FFI layer: its boundary for this
 * path is the nn.Module contract of ldm/modules/attention.py (CrossAttention :147, BasicTransformerBlock
 * :260, SpatialTransformer :287), ldm/modules/embedding_manager.py (AttentionalPooler :100, AdaEmbedding
 * :539) and DDIMSampler.p_sample_ddim (ldm/models/diffusion/ddim.py:222).  The Python host side in
 * adaface_b200/ keeps those signatures; below it, every kernel group is reachable through the
 * stateless extern "C" entry points declared here (plain device pointers and sizes, no torch types),
 * which is what a maintainer binds with ctypes / torch.library.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless stated; the caller owns all buffers and workspaces;
 *     kernels never allocate or free; all launches go to `stream` and are CUDA-graph capturable;
 *   - `bf16` buffers are raw uint16 bfloat16; activations are token-major [rows, C] unless stated;
 *   - return 0 on success; non-zero = error, message via afb_last_error() (thread-local);
 *     2 = unsupported shape/argument, 3 = CUDA runtime error;
 *   - there is no CPU fallback: on a machine without an sm_100 device every entry point returns 3.
 */
#ifndef ADAFACE_B200_H_
#define ADAFACE_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* afb_stream_t;   /* cudaStream_t */
typedef uint16_t afb_bf16;

const char* afb_last_error(void);
int afb_abi_version(void);
/* sha256 (hex) of this header as it was when the library was built: a loader compares it with the header it binds
 * against and refuses a library whose argument structs / signatures may differ (adaface_b200/_lib.py does). */
const char* afb_header_digest(void);
/* Number of kernels launched through this library by the calling process since load (bench.py's gpu_launches). */
int64_t afb_launch_count(void);

/* ---- K14: CFG combine + DDIM update.  Replaces ldm/models/diffusion/ddim.py:252-296 ------------------
 * e = e_u + s (e_c - e_u); pred_x0 = (x - sqrt(1-a_t) e)/sqrt(a_t); x_prev = sqrt(a_prev) pred_x0 +
 * sqrt(1 - a_prev - sigma^2) e + sigma * noise * temperature.  fp32 in/out, n = number of elements of x.
 * `noise` may be NULL when sigma_t == 0. */
int afb_cfg_ddim_step(const float* x, const float* e_cond, const float* e_uncond, const float* noise,
                      float* x_prev, float* pred_x0, int64_t n, float guidance_scale, float a_t, float a_prev,
                      float sigma_t, float sqrt_one_minus_at, float temperature, afb_stream_t stream);

/* ---- K1: GroupNorm(32, C, eps) -> token-major bf16.  Replaces attention.py:71-72,325 (+ the layout change
 * of :327).  x_layout 0: x is fp32 [B,C,HW] (NCHW contiguous); 1: x is bf16 [B,HW,C] (channels-last).
 * stats_ws: fp32 workspace of afb_groupnorm_ws_floats(B, HW, groups) floats.  y: bf16 [B*HW, C].
 * act 0: none; 1: SiLU fused after the affine (the GroupNorm32 -> SiLU pairs of the U-Net's ResBlocks,
 * openaimodel.py:201-205,233-236 - the first "next" row after the transformer path).
 * pre_add: NULL, or fp32 [B, C] with row stride pre_add_ld (x_layout 1 only): a per-(sample, channel) vector added
 * to x before the statistics and the normalisation (a ResBlock's conv bias + time-embedding add, :259-279). */
int64_t afb_groupnorm_ws_floats(int batch, int hw, int groups);
int afb_groupnorm_tokens(const void* x, int x_layout, const float* gamma, const float* beta, afb_bf16* y,
                         float* stats_ws, int batch, int channels, int hw, int groups, float eps, int act,
                         const float* pre_add, int pre_add_ld, afb_stream_t stream);

/* Same with the input given as two token-major bf16 tensors that are (virtually) concatenated along the channel dim:
 * channels [0, channels1) from x [B, HW, channels1], the rest from x2 [B, HW, channels - channels1] - the
 * th.cat([h, hs.pop()], dim=1) in front of every decoder ResBlock (openaimodel.py:1004-1010) is never materialised
 * for its GroupNorm.  x2 == NULL: identical to afb_groupnorm_tokens. */
int afb_groupnorm_tokens_cat(const void* x, int channels1, const afb_bf16* x2, int x_layout, const float* gamma,
                             const float* beta, afb_bf16* y, float* stats_ws, int batch, int channels, int hw,
                             int groups, float eps, int act, const float* pre_add, int pre_add_ld, afb_stream_t stream);

/* ---- allocation-free variant of afb_groupnorm_tokens.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments and results as afb_groupnorm_tokens minus `stats_ws`: the statistics
 * workspace is a file-scope `static __device__` array inside the library (1 MiB = afb_groupnorm_static_ws_max_floats()
 * floats per GPU that loads the library, held until the process exits).  The call allocates nothing (no cudaMalloc /
 * cudaFree / allocating library routine); the caller still owns x and y.  Limit: 2 * batch * groups * 33 floats must
 * fit, i.e. batch * groups <= 3971 (batch <= 124 at 32 groups); larger inputs are rejected with code 2 - use
 * afb_groupnorm_tokens with a caller workspace.  All calls on a device share the one array: calls that overlap on the
 * GPU (different streams, or graph branches without an edge between them) race on it; calls ordered on one stream are
 * safe, and nothing needs to be cleared between them (every slot read in a call is written earlier in that call).
 * The array's device address is resolved once per device on the first call made with that device current. */
int64_t afb_groupnorm_static_ws_max_floats(void);
int afb_groupnorm_tokens_static_ws(const void* x, int x_layout, const float* gamma, const float* beta, afb_bf16* y,
                                   int batch, int channels, int hw, int groups, float eps, int act,
                                   const float* pre_add, int pre_add_ld, afb_stream_t stream);

/* ---- out = a + b + bias (elementwise, bf16 [rows, C] token-major; b and bias optional).  The residual / bias adds
 * around the ResBlock convolutions (openaimodel.py:279) as one vectorised pass instead of two broadcast adds. */
int afb_add_bias_residual(const afb_bf16* a, const afb_bf16* b, const float* bias, afb_bf16* out, int64_t rows,
                          int channels, afb_stream_t stream);

/* ---- K3: LayerNorm over the last dim, bf16 in / bf16 out, fp32 statistics.  attention.py:267-269,277-283.
 * gamma/beta may be NULL (no affine; embedding_manager.py:111-113). */
int afb_layernorm_bf16(const afb_bf16* x, const float* gamma, const float* beta, afb_bf16* y, int64_t rows,
                       int channels, float eps, afb_stream_t stream);

/* ---- K2/K4/K6/K7/K9/K10/K11: Y = epilogue(A[M,K] * W[N,K]^T), bf16 operands, fp32 accumulation in TMEM.
 * Replaces every nn.Linear / 1x1 nn.Conv2d on the path (attention.py:35,55,157-165,302-317). */
enum afb_gemm_epilogue {
  AFB_EPI_STORE = 0,  /* out[m,n] = acc + bias[n] (+ residual[m,n]); row-major, out_dtype selects bf16/fp32 */
  AFB_EPI_GEGLU = 1,  /* W / bias rows interleaved (32 value rows, their 32 gate rows, ...), N % 64 == 0:       */
                      /* out[m,j] = a * gelu_erf(g), out is bf16 [M, N/2] (attention.py:32-39)               */
  AFB_EPI_NCHW = 2    /* out fp32 [B, N, HW] with m = b*HW + t: out = acc + bias + residual (same layout)     */
                      /* (proj_out + 'b (h w) c -> b c h w' + x_in: attention.py:335-341)                    */
};
enum afb_dtype { AFB_BF16 = 0, AFB_F32 = 1 };

typedef struct afb_gemm_args {
  const afb_bf16* a;        /* [M, K] row-major, lda elements (K contiguous); 16-byte aligned, lda % 8 == 0 */
  const afb_bf16* w;        /* [N, K] row-major (nn.Linear weight layout), ldw elements                     */
  void* out;                /* see epilogue                                                                  */
  const float* bias;        /* [N] fp32 or NULL                                                              */
  const void* residual;     /* NULL, or same shape/layout as out with dtype residual_dtype                   */
  int64_t m, n, k;
  int64_t lda, ldw, ldo, ldr;   /* leading dimensions in elements (ldo/ldr unused for AFB_EPI_NCHW)         */
  int epilogue;             /* enum afb_gemm_epilogue */
  int out_dtype;            /* enum afb_dtype (STORE only; GEGLU is bf16, NCHW is fp32)                      */
  int residual_dtype;       /* enum afb_dtype */
  int hw;                   /* tokens per sample, AFB_EPI_NCHW only                                          */
  /* LayerNorm folded around the GEMM (attention.py:269-283: norm1/2/3 feed to_q/k/v and the GEGLU projection).
   * Producer side (STORE epilogue, bf16 out): row_stats_out != NULL makes the epilogue emit, per output row and per
   * (column block, column half), the sum and the sum of squares of the bf16 values it stores:
   * row_stats_out[m][afb_gemm_stats_parts(n)][2] fp32.
   * Consumer side (STORE with bf16 out, or GEGLU): ln_stats != NULL points at such an array for the rows of A
   * (ln_parts partials per row, rows of A have k elements); with W' = W * gamma (given as w),
   * ln_colsum[n] = sum_k W'[n,k] and bias[n] = sum_k W[n,k] * beta[k] (+ the layer's bias), the epilogue computes
   * rstd[m] * (acc - mean[m] * ln_colsum[n]) + bias[n] == LayerNorm(A)[m,:] . W[n,:] + bias. */
  float* row_stats_out;
  const float* ln_stats;
  const float* ln_colsum;
  int ln_parts;
  float ln_eps;
  /* 3x3 convolution (stride 1, zero padding 1) as an implicit GEMM - the ResBlock / Upsample convolutions of
   * openaimodel.py:95-123,203-205,236-238.  conv_h, conv_w > 0: a is the bf16 NHWC image
   * [B, conv_h, conv_w, Cin] with pixel stride lda, m = B*conv_h*conv_w, k = 9*Cin (Cin % 64 == 0), w is
   * [N, 9*Cin] with k = (ky*3 + kx)*Cin + c, i.e. the channels-last memory of the Conv2d weight; out / residual are
   * bf16 [m, N] = NHWC.  STORE epilogue with bf16 out only.  Widths/heights must tile into 128-pixel boxes
   * (powers of two up to 128 or multiples of 128).  0, 0: plain GEMM. */
  int conv_h, conv_w;
  /* conv_fold != 0: nearest-neighbour 2x upsampling folded into the convolution (Upsample.forward,
   * openaimodel.py:95-123: F.interpolate(scale 2, nearest) then the 3x3 conv).  Output pixel (2i+py, 2j+px) only sees
   * 2x2 low-resolution pixels, so each of the four output phases is a 2x2 convolution of the LOW-resolution image
   * a = [B, conv_h, conv_w, Cin] with pre-summed weights: k = 4*Cin, n = 4*Cout, w is [4*Cout, 4*Cin] with row
   * (py*2+px)*Cout + o and column (a*2+c)*Cin + ch = sum of the 3x3 taps (ky, kx) that land on low-res offset
   * (py-1+a, px-1+c); bias is [4*Cout] (the conv bias repeated); out is bf16 [B, 2*conv_h, 2*conv_w, Cout], ldo = Cout.
   * 2.25x fewer FLOPs than convolving the upsampled image, and the upsampled tensor is never written. */
  int conv_fold;
  /* a2 != NULL (plain GEMM, bf16 STORE epilogue): the A operand is the virtual concatenation [a | a2] along K -
   * columns [0, k_split) from a (leading dimension lda), columns [k_split, k) from a2 [M, k - k_split] (leading dimension
   * lda2); k_split % 64 == 0.  The decoder's 1x1 skip convolution over th.cat([h, hs.pop()], 1) in one GEMM. */
  const afb_bf16* a2;
  int64_t k_split, lda2;
  /* conv_stride = 2 (0 and 1 mean stride 1; conv_fold == 0 only): the stride-2 / padding-1 3x3 convolution of
   * Downsample.forward (openaimodel.py:138-165).  conv_h, conv_w are then the OUTPUT image size, a is the input image
   * [B, 2*conv_h, 2*conv_w, Cin], m = B*conv_h*conv_w output pixels; tap (dy, dx) of output pixel (i, j) reads input
   * pixel (2i + dy, 2j + dx), which TMA fetches with an element stride of 2 along both image axes. */
  int conv_stride;
} afb_gemm_args;
int afb_gemm_bf16(const afb_gemm_args* args, afb_stream_t stream);

/* ---- N1: the two convolutions without a 64-channel side (openaimodel.py:575-579 conv_in, :716-720 conv_out).
 * afb_im2col3x3_c4: x bf16 NHWC [B, h, w, 4] -> cols bf16 [B*h*w, 64] with cols[p][(ky*3+kx)*4 + c] = the zero-padded
 * 3x3 neighbourhood of pixel p (columns 36..63 zero); Conv2d(4, Cout, 3, padding=1) is then afb_gemm_bf16 of cols with
 * the weight as [Cout, 64] (same k order, zero-padded).
 * afb_conv3x3_few_out: x bf16 NHWC [B, h, w, channels] (channels % 64 == 0), w8 bf16 [8, 9*channels] with
 * k = (ky*3+kx)*channels + c and rows >= cout zero, bias fp32 [cout] or NULL -> out fp32 NCHW [B, cout, h, w], cout <= 8:
 * 3x3 / stride 1 / padding 1. */
int afb_im2col3x3_c4(const afb_bf16* x, afb_bf16* cols, int batch, int h, int w, afb_stream_t stream);
int afb_conv3x3_few_out(const afb_bf16* x, const afb_bf16* w8, const float* bias, float* out, int batch, int h, int w,
                        int channels, int cout, afb_stream_t stream);
/* number of (sum, sum of squares) partials per row that a STORE GEMM with n output columns emits */
int afb_gemm_stats_parts(int64_t n);

/* ---- K5/K8: fused softmax(Q K^T * scale) V, online softmax, S/O accumulators in TMEM.
 * Replaces attention.py:198-242 (self-attention with optional key mask, and cross-attention over the
 * 77-token context).  q/k/v are bf16 with heads concatenated along the channel dim: element (b, t, h, i)
 * of q is q[(b*q_len + t)*ldq + h*head_dim + i], likewise k/v with kv_len and ldk/ldv; out is bf16
 * [B*q_len, ldo] in the same head-concatenated layout.  head_dim in {40, 80, 160}.
 * key_mask: NULL or uint8 [B, kv_len], 0 = masked key (attention.py:223-232). */
typedef struct afb_attn_args {
  const afb_bf16* q;
  const afb_bf16* k;
  const afb_bf16* v;
  afb_bf16* out;
  const uint8_t* key_mask;
  int batch, heads, head_dim;
  int q_len, kv_len;
  int64_t ldq, ldk, ldv, ldo;
  float scale;
  /* optional (NULL to skip): fp32 [B, heads, q_len], the log-sum-exp of the scaled scores of every query row in the
   * log2 domain, lse = log2(sum_k 2^(scale * log2(e) * s_k)); what afb_attention_bwd needs to rebuild the probabilities */
  float* lse;
} afb_attn_args;
int afb_attention_fwd(const afb_attn_args* args, afb_stream_t stream);

/* ---- K8b: explicit (materialised) 77-token cross-attention for the optional modes of CrossAttention.forward that
 * need the score matrix: save_attn_vars (attention.py:245-255) and the convolutional-attention row replacement
 * (attention.py:208-216 -> ldm/util.py:701-879).  sim / attn are fp32 [B, heads, q_len, kv_len] (the reference's
 * '(b h) i j' with b major); q/k/v/out use the head-concatenated bf16 layout of afb_attention_fwd. */
/* sim[b,h,n,t] = scale * q[b,n,h,:] . k[b,t,h,:]   (attention.py:199) */
int afb_cross_scores(const afb_bf16* q, const afb_bf16* k, float* sim, int batch, int heads, int head_dim, int q_len,
                     int kv_len, int64_t ldq, int64_t ldk, float scale, afb_stream_t stream);
/* replace_rows_by_conv_attn (ldm/util.py:701-879) with conv_attn_mix_weight = 1 and shifted attention maps, in place
 * on sim: for instance i (sample inst_batch[i]) the ks*ks subject tokens tok[i*ks*ks + m] act as a ks x ks filter on
 * the query image [h, w]; column tok[.. + m] of sim receives the conv output shifted by that tap's offset.
 * conv_ws: n_inst * heads * h * w floats.  scale = the attention scale (d^-1/2); ks in {2, 3, 4}. */
int afb_conv_attn_rows(float* sim, const afb_bf16* q, const afb_bf16* k, const int* inst_batch, const int* tok,
                       float* conv_ws, int batch, int heads, int head_dim, int h, int w, int kv_len, int64_t ldq,
                       int64_t ldk, int n_inst, int ks, float scale, afb_stream_t stream);
/* attn = softmax_t(sim) (written when attn != NULL), out[b,n,h,:] = attn . v[b,:,h,:]   (attention.py:238-242) */
int afb_softmax_pv(const float* sim, const afb_bf16* v, float* attn, afb_bf16* out, int batch, int heads, int head_dim,
                   int q_len, int kv_len, int64_t ldv, int64_t ldo, afb_stream_t stream);

/* ---- K12: AttentionalPooler.forward, eval mode.  Replaces embedding_manager.py:146-321.
 * afb_ada_pool_prep (step-invariant, :186-235): fg_q/bg_q = to_k(query embedding) with the layer's attn2.to_k
 *   weight fp32 [C, ctx_dim]; fg_q_emb fp32 [ctx_dim] (shared by the batch), bg_q_emb fp32 [B, ctx_dim];
 *   LayerNorm; grouped 1x1 convs lora_to_fg_q / lora_to_bg_q (fp32 [C/8, C/8], 8 groups); the result is folded
 *   through lora_to_k (fp32 [C/8, C/8]; the grouped conv applied to every token at :227 is linear) into the
 *   pooling queries pool_q fp32 [B, 2, C].  prep_ws: fp32 workspace of afb_ada_pool_prep_ws_floats(B, C) floats.
 * afb_ada_pool: x, q bf16 [B, N, C] (block input x1 and attn2.to_q(norm2(x1))); key_mask NULL or uint8 [B, N]
 *   (0 = masked).  Outputs fp32: fg_out, bg_out [B, C]; attn_fg, attn_bg [B, 8, N] (both NULL to skip).
 *   ws: fp32 workspace of afb_ada_pool_ws_floats(B, N, C) floats. */
int64_t afb_ada_pool_ws_floats(int batch, int n_tokens, int channels);
int64_t afb_ada_pool_prep_ws_floats(int batch, int channels);
/* offset (in floats) inside afb_ada_pool's workspace of the merged softmax statistics fp32 [B, 8, 2] = (max, sum) per
 * (sample, head) that the call leaves behind - the `stats` argument of afb_ada_pool_bwd / afb_ada_pool_bwd_features */
int64_t afb_ada_pool_stats_offset(int batch, int n_tokens, int channels);
int afb_ada_pool_prep(const float* to_k_weight, int ctx_dim, const float* fg_q_emb, const float* bg_q_emb,
                      const float* lora_to_fg_q, const float* lora_to_bg_q, const float* lora_to_k, float* pool_q,
                      float* prep_ws, int batch, int channels, afb_stream_t stream);
/* The same for every (cross-attention layer, subject) of one U-Net call in two launches: `items` is a DEVICE array of
 * n_items descriptors (the caller uploads it once; every pointer inside is a device pointer with the meaning of the
 * afb_ada_pool_prep argument of the same name; ws holds afb_ada_pool_prep_ws_floats(batch, channels) floats).  All
 * items share batch and ctx_dim; max_channels = the largest `channels`. */
typedef struct afb_ada_prep_item {
  const float* to_k_weight;
  const float* fg_q_emb;
  const float* bg_q_emb;
  const float* lora_to_fg_q;
  const float* lora_to_bg_q;
  const float* lora_to_k;
  float* pool_q;
  float* ws;
  int channels;
  int reserved;
} afb_ada_prep_item;
int afb_ada_pool_prep_batched(const afb_ada_prep_item* items, int n_items, int max_channels, int ctx_dim, int batch,
                              afb_stream_t stream);
int afb_ada_pool(const afb_bf16* x, const afb_bf16* q, const float* pool_q, const uint8_t* key_mask, float* fg_out,
                 float* bg_out, float* attn_fg, float* attn_bg, float* ws, int batch, int n_tokens, int channels,
                 float drop_p, uint32_t seed, afb_stream_t stream);
/* drop_p, seed (afb_ada_pool and its two backward entry points): the AttentionalPooler's train-mode Dropout on the
 * attention probabilities (embedding_manager.py:138, 296).  drop_p = 0 is eval mode.  With drop_p > 0 probability
 * (b, head, fg/bg, token) is kept (and scaled by 1 / (1 - drop_p)) when a counter-based hash of (seed, element index)
 * says so; the backward entry points regenerate the same mask from the same (drop_p, seed).  The reference's second
 * dropout (`out_drop` on the pooled features, :308-313) is an elementwise op on a [B, C] tensor and stays with the caller. */

/* ---- K13: AdaEmbedding head.  Replaces embedding_manager.py:847-893.
 * infeat_fg/bg fp32 [B, C]; time_emb fp32 [B, time_ld] (first TD = C/4 used); coeff_w fp32 [K*r, 2C+TD],
 * coeff_b [K*r]; basis fp32 [K, r, D]; bias fp32 [K, D] or NULL; out fp32 [B, K, D].
 * Optional splice of the subject rows into the layer's context: ctx_out bf16 with sample stride
 * ctx_sample_stride elements and rows of D elements, token_pos int32 [K] (device) - vector k of sample b is also
 * written, as bf16, to ctx_out[b * ctx_sample_stride + token_pos[k] * D ...].  Both NULL to skip.
 * basis_sample_stride: 0 for the shared basis [K, r, D]; K*r*D for per-instance zero-shot basis vectors
 * [B, K, r, D] (embedding_manager.py:860-868, no bias in that mode). */
int afb_ada_head(const float* infeat_fg, const float* infeat_bg, const float* time_emb, int time_ld,
                 const float* coeff_w, const float* coeff_b, const float* basis, const float* bias, float* out,
                 afb_bf16* ctx_out, const int* token_pos, int64_t ctx_sample_stride, int64_t basis_sample_stride, int batch, int channels, int td,
                 int num_vectors, int r, int out_dim, afb_stream_t stream);
/* The same head fed straight from the pooling kernel's chunk partials: afb_ada_pool called with fg_out = bg_out = NULL
 * leaves them in its workspace, and this entry point merges them in its prologue (the separate combine launch disappears
 * from the critical path of every cross-attention layer).  pool_ws / n_tokens / batch / channels as given to afb_ada_pool;
 * fg_out, bg_out fp32 [B, C] (optional) receive the pooled features, and the workspace's statistics slot
 * (afb_ada_pool_stats_offset) the merged softmax (max, sum); cached_bg (optional fp32 [B, C]) replaces the pooled bg
 * stream as the head's input (embedding_manager.py:789-792).  Other arguments as afb_ada_head. */
int afb_ada_head_pooled(const float* pool_ws, int n_tokens, float* fg_out, float* bg_out, const float* cached_bg,
                        const float* time_emb, int time_ld, const float* coeff_w, const float* coeff_b, const float* basis,
                        const float* bias, float* out, afb_bf16* ctx_out, const int* token_pos, int64_t ctx_sample_stride,
                        int64_t basis_sample_stride, int batch, int channels, int td, int num_vectors, int r, int out_dim,
                        afb_stream_t stream);

/* ---- N2: StaticLayerwiseEmbedding and the splice of its rows into the layerwise prompt.
 * afb_static_layerwise_embedding replaces StaticLayerwiseEmbedding.forward (not zero-shot),
 *   embedding_manager.py:497-537: out[l, k] = LN(sum_r (rand[l,k,r] + comm[0,k,r]) * basis[k,r]) / sqrt(D) + bias[l,k].
 *   basis_rand_weights fp32 [L, K, r]; basis_comm_weights [1, K, r]; pre_vecs [K, n_pre, D] or NULL (n_pre = 0);
 *   basis_vecs [K, r - n_pre, D]; bias [L, K, D] or NULL; out fp32 [L, K, D].
 * afb_first_token_pos: tokens int64 [B, N] -> first_pos int32 [B], the first position holding `token`, -1 if none
 *   (extract_first_index_in_each_instance, ldm/util.py:2114-2124).
 * afb_splice_static_embedding replaces the inference branch of EmbeddingManager.get_static_embedding for one
 *   placeholder, embedding_manager.py:1341-1352 and 1510-1552: out[b * L + l, n] = static_emb[l, n - first_pos[b]] for
 *   first_pos[b] <= n < first_pos[b] + K, else embedded_text[b, n]; positions past N are dropped.  embedded_text fp32
 *   [B, N, D]; static_emb fp32 [L, K, D]; out_f32 fp32 and/or out_bf16 bf16 [B * L, N, D] (either may be NULL). */
int afb_static_layerwise_embedding(const float* basis_rand_weights, const float* basis_comm_weights, const float* pre_vecs,
                                   int n_pre, const float* basis_vecs, const float* bias, float* out, int num_layers,
                                   int num_vectors, int r, int out_dim, afb_stream_t stream);
int afb_first_token_pos(const int64_t* tokens, int64_t token, int* first_pos, int batch, int n_tokens, afb_stream_t stream);
int afb_splice_static_embedding(const float* embedded_text, const int* first_pos, const float* static_emb, float* out_f32,
                                afb_bf16* out_bf16, int batch, int n_tokens, int num_layers, int num_vectors, int dim,
                                afb_stream_t stream);

/* ---- Backward of the inversion step (BASELINE.json configs[3]: backward through cross-attention, gradients only on the
 * AdaEmbedding parameters).  The reference has no backward code: torch autograd differentiates attention.py:32-59,
 * 195-243, 269-283 and embedding_manager.py:146-321, 847-893 under ddpm.py:599 (manual_backward); these entry points
 * are the analytic derivatives of the forward kernels above.  Products against the frozen weights reuse afb_gemm_bf16
 * (dA = dC W with the transposed weight as `w`).  Gradient buffers marked "+=" are accumulated into (zero them first).
 *
 * afb_geglu_bwd: h bf16 [rows, 2*hidden] = the GEGLU projection in the forward GEMM's interleaved column order (32 value
 *   columns, their 32 gate columns, ...), dout bf16 [rows, hidden] -> dh bf16 [rows, 2*hidden], same order.
 * afb_layernorm_bwd: LayerNorm without affine (the affine part is folded into the consumer's weight): a, g = d(a_hat)
 *   bf16 [rows, C], add = NULL or bf16 [rows, C] (gradient arriving along the residual stream) -> da bf16.
 * afb_attention_bwd_short: kv_len <= 128 (d = 160: <= 80), no key mask - the 77-token cross-attention.  args as in
 *   afb_attention_fwd (`out` unused), dout bf16 [B*q_len, ldo]; dq bf16 [B*q_len, lddq] or NULL; dk, dv fp32
 *   [B*kv_len, lddk] "+=".  Up to 80 keys run on warp-level tensor-core MMAs, longer ones on CUDA cores (see afb_attention_bwd for self-attention).
 * afb_ada_head_bwd: grad_out fp32 [B, K, D] -> d_coeff_w [K*r, 2C+TD], d_coeff_b [K*r], d_basis [K, r, D], d_bias [K, D]
 *   or NULL (all "+="), d_infeat_fg / d_infeat_bg fp32 [B, C] (written); d_feat_ws: fp32 workspace [B, 2C+TD].
 * afb_ada_pool_bwd: gradient of the pooled features w.r.t. the folded pooling queries: stats fp32 [B, 8, 2] = the
 *   forward's (max, sum) per (sample, head) (tail of afb_ada_pool's workspace), fg_out / bg_out the forward outputs,
 *   d_fg / d_bg fp32 [B, C] -> d_pool_q fp32 [B, 2, C] (written).  Gradients towards x and q are not produced.
 * afb_ada_pool_prep_bwd: proj_ws = the forward's workspace (fp32 [(B+1), C]) -> d_lora_to_fg_q, d_lora_to_bg_q,
 *   d_lora_to_k fp32 [C/8, C/8] ("+="). */
int afb_geglu_bwd(const afb_bf16* h, const afb_bf16* dout, afb_bf16* dh, int64_t rows, int hidden, afb_stream_t stream);
int afb_layernorm_bwd(const afb_bf16* a, const afb_bf16* g, const afb_bf16* add, afb_bf16* da, int64_t rows, int channels,
                      float eps, afb_stream_t stream);
int afb_attention_bwd_short(const afb_attn_args* args, const afb_bf16* dout, afb_bf16* dq, float* dk, float* dv, int64_t lddq,
                            int64_t lddk, afb_stream_t stream);
int afb_ada_head_bwd(const float* infeat_fg, const float* infeat_bg, const float* time_emb, int time_ld, const float* coeff_w,
                     const float* coeff_b, const float* basis, const float* grad_out, float* d_coeff_w, float* d_coeff_b,
                     float* d_basis, float* d_bias, float* d_feat_ws, float* d_infeat_fg, float* d_infeat_bg, int batch,
                     int channels, int td, int num_vectors, int r, int out_dim, afb_stream_t stream);
int afb_ada_pool_bwd(const afb_bf16* x, const afb_bf16* q, const float* pool_q, const uint8_t* key_mask, const float* stats,
                     const float* fg_out, const float* bg_out, const float* d_fg, const float* d_bg, float* d_pool_q, int batch,
                     int n_tokens, int channels, float drop_p, uint32_t seed, afb_stream_t stream);
int afb_ada_pool_prep_bwd(const float* proj_ws, const float* lora_to_fg_q, const float* lora_to_bg_q, const float* lora_to_k,
                          const float* d_pool_q, float* d_lora_to_fg_q, float* d_lora_to_bg_q, float* d_lora_to_k, int batch,
                          int channels, afb_stream_t stream);

/* Second stage: what carries a layer's gradient further down the U-Net.
 * afb_attention_bwd: any key length (self-attention), no key mask.  args as in the forward call that produced args->out
 *   and args->lse (both required); dout bf16 with the row pitch of out (ldo); delta_ws fp32 [B * heads * q_len];
 *   dq (zeroed inside), dk, dv fp32 [B * len, ldg].  Warp-level tensor-core MMAs (mma.sync), 128-key tiles, scores computed transposed; a key mask
 *   (args->key_mask) is honoured as in the forward.
 * afb_groupnorm_tokens_bwd: token-major bf16 x / grad_y [B, HW, C] -> dx bf16.  act = 0: plain GroupNorm (the
 *   SpatialTransformer's Normalize, attention.py:71-72); act = 1: GroupNorm -> SiLU (the ResBlocks, openaimodel.py:259-279;
 *   beta required).  pre_add (fp32 [B, C] rows of pitch add_ld, or NULL) is what the forward added to x before the
 *   statistics (a ResBlock's conv bias + time-embedding projection); gamma / beta / pre_add are frozen: no gradient.
 *   ws: fp32 [batch * groups * 4] scratch (zeroed inside) for the streaming three-pass version that large slabs take
 *   (hw >= 2048); NULL keeps the single-kernel (sample, group) slab version for every shape.
 * afb_ada_pool_bwd_features: the pooler's gradient towards x and q (bf16 [B, N, C] each); other arguments as
 *   afb_ada_pool_bwd. */
int afb_attention_bwd(const afb_attn_args* args, const afb_bf16* dout, float* delta_ws, float* dq, float* dk, float* dv,
                      int64_t ldg, afb_stream_t stream);
int afb_groupnorm_tokens_bwd(const afb_bf16* x, const afb_bf16* grad_y, const float* gamma, const float* beta,
                             const float* pre_add, int64_t add_ld, int act, afb_bf16* dx, float* ws, int batch, int channels,
                             int hw, int groups, float eps, afb_stream_t stream);
int afb_ada_pool_bwd_features(const afb_bf16* x, const afb_bf16* q, const float* pool_q, const uint8_t* key_mask,
                              const float* stats, const float* fg_out, const float* bg_out, const float* d_fg, const float* d_bg,
                              afb_bf16* dx, afb_bf16* dq, int batch, int n_tokens, int channels, float drop_p, uint32_t seed,
                              afb_stream_t stream);

/* ---- layout / dtype helpers on the path's edges ---------------------------------------------------------- */
int afb_cast_f32_to_bf16(const float* src, afb_bf16* dst, int64_t n, afb_stream_t stream);
int afb_cast_bf16_to_f32(const afb_bf16* src, float* dst, int64_t n, afb_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* ADAFACE_B200_H_ */
