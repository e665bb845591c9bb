/* tnt_b200.h -- C ABI of the B200-native caption-decoder hot path ("Transform-and-Tell" decoder).
 *
 * The reference (harsh157/news_image_caption) has no FFI of its own: its plugin boundary is the
 * Python nn.Module contract + the AllenNLP registry.  This header is the
 * boundary a replacement sits UNDER: every entry point is extern "C", takes plain device pointers,
 * sizes and a stream, allocates nothing, never throws, and returns 0 or a negative tnt_status.
 * Each entry point cites the reference interface it replaces (paths relative to the reference root).
 *
 * Conventions
 *   - all pointers are DEVICE pointers unless the name ends in _host;
 *   - activations are row-major [rows, cols] with an explicit leading dimension in ELEMENTS;
 *     decoder-layer tensors use the reference's T x B x C order (row = t * B + b);
 *   - dtype: TNT_F32 or TNT_BF16 for activations and GEMM weights; biases, LayerNorm parameters,
 *     softmax statistics, losses and log-probabilities are always fp32; ids are int64 (torch.long);
 *   - `stream` is a cudaStream_t passed as void*; calls are stream-ordered and never synchronise
 *     (the *_host entry points are the exception and say so);
 *   - the caller owns every buffer; workspaces are sized by the matching *_workspace_bytes query.
 */
#ifndef TNT_B200_H_
#define TNT_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* tnt_stream;

typedef enum { TNT_F32 = 0, TNT_BF16 = 1 } tnt_dtype;

typedef enum {
  TNT_OK = 0,
  TNT_ERR_ARG = -1,       /* bad shape / null pointer / unsupported combination */
  TNT_ERR_ALIGN = -2,     /* pointer or leading dimension not aligned for the vector / TMA path */
  TNT_ERR_CUDA = -3,      /* a CUDA runtime / driver call failed; see tnt_last_error() */
  TNT_ERR_WORKSPACE = -4, /* workspace too small */
  TNT_ERR_NODEVICE = -5   /* no sm_100 device */
} tnt_status;

int tnt_version(void);
const char* tnt_last_error(void);          /* thread-local message of the last failing call */
int tnt_device_check(void);                /* TNT_OK iff the current device is sm_100 */

/* ------------------------------------------------------------------------------------------
 * Linear layers.  Replaces GehringLinear / TiedLinear / F.linear call sites:
 *   ttl/tell/modules/linear.py:7-50, multi_head.py:488-526 (in_proj_*), softmax.py:99-132.
 * C[M,N] = epilogue(A[M,K] . W[N,K]^T + bias[N]).  W uses the nn.Linear layout.
 * bf16: hand-written tcgen05 + TMA kernel (fp32 accumulate in TMEM); fp32: SIMT fp32 kernel.
 * ------------------------------------------------------------------------------------------ */
typedef enum {
  TNT_EPI_NONE = 0,      /* + bias (if any) */
  TNT_EPI_RELU = 1,      /* relu(acc + bias)                       decoder_flattened.py:338 */
  TNT_EPI_GLU = 2,       /* W rows pair-interleaved (a0,b0,a1,b1..): out[:, j] = a_j * sigmoid(b_j);
                            N counts the interleaved rows, C has N/2 columns.  decoder_flattened.py:284-286 */
  TNT_EPI_RESIDUAL = 3,  /* acc + bias + residual[M,N]             decoder_flattened.py:288-291 */
  TNT_EPI_RELU_BWD = 5   /* acc * (residual[M,N] > 0): gradient through relu, residual = the saved relu output */
} tnt_epilogue;

/* ------------------------------------------------------------------------------------------
 * Dropout (training mode; F.dropout sites of decoder_flattened.py:109,283,289,308,326,340,342, dynamic.py:305,
 * multi_head.py:463).  Counter based: element `index` of dropout site `site` is KEPT iff
 *     tnt_dropout_bits(seed, site, index) >= (uint32) (p * 2^32)
 * and kept values are scaled by 1 / (1 - p).  Nothing is stored: the backward kernels recompute the mask from
 * (seed, site, index).  p == 0 (a zeroed struct) switches the site off.  tnt_dropout_bits is a host function so
 * that a test / oracle can reproduce every mask bit for bit.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  float p;                /* drop probability in [0, 1) */
  uint32_t site;          /* which dropout call of the model (TNT_DROP_SITE) */
  uint64_t seed;
  int64_t ld;             /* GEMM epilogues: index = group * group_stride + row * ld + column; 0 means ld = N */
  int64_t group_stride;
} tnt_dropout;
uint32_t tnt_dropout_bits(uint64_t seed, uint32_t site, uint64_t index);
/* out[r, c] = dropout(in[r, c]) (+ residual[r, c]); index = r * d.ld + c (d.ld == 0: cols).  In-place allowed. */
int tnt_dropout_apply(const void* in, int64_t ldi, const void* residual, int64_t ldr, void* out, int64_t ldo, int rows,
                      int cols, int dtype, const tnt_dropout* d, tnt_stream stream);
/* dropout sites of the decoder engine: TNT_DROP_SITE(layer, kind) */
enum {
  TNT_DROP_INPUT = 0,      /* layer input before linear1 (input_dropout), index (t*B+b)*d + c */
  TNT_DROP_TAPS = 1,       /* DynamicConv taps after their soft-max (weight_dropout), index ((t*B+b)*H + h)*K + j */
  TNT_DROP_CONV_OUT = 2,   /* linear2 output before the residual (dropout), index (t*B+b)*d + c */
  TNT_DROP_ATTN_OUT = 3,   /* out_proj output of source s before the residual, index (t*B+b)*(n_src*d) + s*d + c */
  TNT_DROP_FFN_OUT = 4,    /* fc2 output before the residual, index (t*B+b)*d + c */
  TNT_DROP_RELU = 5,       /* fc1 output after ReLU (relu_dropout), index (t*B+b)*ffn + c */
  TNT_DROP_ATTN_PROB = 6   /* + source: attention probabilities, index ((b*H + h)*T + t)*(S+2) + key */
};
#define TNT_DROP_SITE(layer, kind) ((uint32_t)((layer) * 16 + (kind)))
#define TNT_DROP_SITE_EMBED 0xFFFFu   /* embedder output, index (t*B+b)*d + c */

/* operand layouts (backward passes reuse the forward weights without transposed copies):
 *   TNT_NT  C[M,N] = A[M,K] . W[N,K]^T     forward        (W = nn.Linear weight [out, in])
 *   TNT_NN  C[M,N] = A[M,K] . W[K,N]       data gradient  dX = dY . W
 *   TNT_TN  C[M,N] = A[K,M]^T . W[K,N]     weight gradient dW = dY^T . X   (K = rows/tokens) */
typedef enum { TNT_NT = 0, TNT_NN = 1, TNT_TN = 2 } tnt_gemm_layout;

typedef struct {
  const void* a;  int64_t lda;        /* [M,K]  dtype_in */
  const void* w;  int64_t ldw;        /* [N,K]  dtype_in */
  const float* bias;                  /* [N] or NULL */
  const void* residual; int64_t ldr;  /* [M,N'] dtype_out, TNT_EPI_RESIDUAL only */
  void* c;        int64_t ldc;        /* [M,N'] dtype_out */
  int M, N, K;
  const int* m_dyn;                   /* optional DEVICE row count (<= M): rows >= *m_dyn are skipped */
  /* grouped problems sharing M,N,K (e.g. one per attention source); element strides, 0 = shared */
  int groups;
  int64_t stride_a, stride_w, stride_bias, stride_res, stride_c;
  int dtype_in, dtype_out, epilogue;
  int layout;                         /* tnt_gemm_layout; with TNT_TN `m_dyn` bounds the reduction length K */
  float alpha;                        /* scales the accumulator before bias / epilogue; 0 means 1 */
  /* optional L2 prefetch hint: `prefetch_bytes` bytes at `prefetch` (16-byte aligned, e.g. the weights of the NEXT
   * GEMM of a latency-bound chain) are requested into L2 (cp.async.bulk.prefetch.L2) by the kernel's prologue */
  const void* prefetch; uint64_t prefetch_bytes;
  tnt_dropout drop;                   /* TNT_EPI_RESIDUAL: dropout(a.w^T + bias) + residual; TNT_EPI_RELU: dropout(relu(.)) */
} tnt_linear_params;
int tnt_linear_fwd(const tnt_linear_params* p, tnt_stream stream);

/* weight_norm fold w = g * v / ||v||_row (linear.py:30-33, nn.utils.weight_norm dim=0), optional
 * GLU pair-interleave of the output rows, cast to `dtype_out`.  g may be NULL (plain cast/copy). */
int tnt_wn_fold(const float* v, const float* g, void* w_out, int rows, int cols, int glu_interleave,
                int dtype_out, tnt_stream stream);
/* bias companion of the GLU interleave: b_out[2j] = b[j], b_out[2j+1] = b[j + n/2] */
int tnt_glu_interleave_bias(const float* b, float* b_out, int n, tnt_stream stream);
int tnt_cast(const void* src, void* dst, int64_t n, int dtype_src, int dtype_dst, tnt_stream stream);
/* the same conversion on at most `max_ctas` CTAs: for casts issued on a side stream while another stream's kernels own the
 * GPU (the bf16 wire buffers of the gradient exchange, trainer.py) -- a full-width grid stalls that stream for its duration */
int tnt_cast_background(const void* src, void* dst, int64_t n, int dtype_src, int dtype_dst, int max_ctas, tnt_stream stream);

/* ------------------------------------------------------------------------------------------
 * DynamicConv1dTBC (causal, weight_softmax, padding_l = K-1, no renorm, no bias).
 *   ttl/tell/modules/convolutions/dynamic.py:85-120,285-336 (forward / _forward_expanded),
 *   :95-99,114-116,338-348 (incremental input buffer).
 * y[t,b,h*R+r] = sum_j softmax_K(logits[t,b,h,:])[j] * x[t-(K-1)+j, b, h*R+r],  x[<0] := history or 0.
 * The softmax runs over ALL K taps; taps reaching before the start are dropped, not renormalised.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* x;  int64_t ldx;        /* [T*B, C] conv input (the GLU output) */
  const void* logits; int64_t ldl;    /* [T*B, H*K] tap logits = weight_linear(x) */
  void* y;        int64_t ldy;        /* [T*B, C] */
  /* incremental decoding: the last `hist_len` (<= K-1) conv inputs, oldest first, [hist_len*B, C];
     NULL for teacher forcing.  If hist_out != NULL the updated window (last min(K-1, hist_len+T)
     inputs) is written there (may alias hist only if hist_len + T <= K-1 is false ... see DESIGN). */
  const void* hist; int hist_len;
  int T, B, C, H, K;
  int dtype;
  tnt_dropout drop;                   /* weight_dropout on the soft-maxed taps (dynamic.py:305) */
} tnt_dynconv_params;
int tnt_dynconv_fwd(const tnt_dynconv_params* p, tnt_stream stream);

typedef struct {
  const void* x;  int64_t ldx;        /* forward input  [T*B, C] */
  const void* logits; int64_t ldl;    /* forward logits [T*B, H*K] */
  const void* dy; int64_t lddy;       /* [T*B, C] */
  void* dx;       int64_t lddx;       /* [T*B, C]   gradient through the value path only */
  void* dlogits;  int64_t lddl;       /* [T*B, H*K] gradient w.r.t. the tap logits */
  int T, B, C, H, K;
  int dtype;
  tnt_dropout drop;                   /* the forward call's tap dropout */
} tnt_dynconv_bwd_params;
int tnt_dynconv_bwd(const tnt_dynconv_bwd_params* p, tnt_stream stream);

/* single-step ring buffer form used by the generation engine (kernel group 5):
 * ring [K-1, B, C] holds conv inputs of steps pos-K+1 .. pos-1 at slot (step mod (K-1)); the new
 * input is consumed and then stored at slot pos mod (K-1). */
typedef struct {
  const void* x;  int64_t ldx;        /* [B, C] */
  const void* logits; int64_t ldl;    /* [B, H*K] */
  void* y;        int64_t ldy;        /* [B, C] */
  void* ring;                          /* [(K-1), B, C] */
  const int* pos;                      /* DEVICE step counter (number of inputs already seen) */
  int B, C, H, K;
  int dtype;
} tnt_dynconv_step_params;
int tnt_dynconv_step(const tnt_dynconv_step_params* p, tnt_stream stream);

/* ------------------------------------------------------------------------------------------
 * LayerNorm over the last dimension of [rows, C] (fp32 statistics).
 *   nn.LayerNorm: decoder_flattened.py:245-268 (variant 0: biased var, eps inside sqrt)
 *   final* LayerNorm: final_roberta2/transformer_v2.py:82-94 (variant 1: unbiased std, eps outside)
 * `groups` > 1 : row r uses gamma/beta block (r % groups) -- the per-source context_attn_lns.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* x; int64_t ldx;
  void* y;       int64_t ldy;
  const float* gamma; const float* beta;   /* [groups, C] */
  int rows, C, groups, variant;
  float eps;
  int dtype;
  /* optional transposed store: output row r = (t*B+b) is written to row (b*T+t); 0 = off */
  int transpose_T, transpose_B;
} tnt_layernorm_params;
int tnt_layernorm_fwd(const tnt_layernorm_params* p, tnt_stream stream);

/* Linear + residual + LayerNorm in ONE launch: c = LayerNorm(a . w^T + bias + residual) * gamma + beta (bf16 in / out, fp32
 * statistics, biased variance, eps inside the sqrt).  Replaces the three GEMM -> add -> nn.LayerNorm sites of a decoder layer
 *   linear2 -> +residual -> conv_layer_norm              ttl/tell/models/decoder_flattened.py:283-291
 *   out_proj[source] -> +residual -> context_attn_lns[source]   decoder_flattened.py:300-330, decoder_faces_objects.py:310-378
 *   fc2 -> +residual -> final_layer_norm                 decoder_flattened.py:336-345
 * `lin` is read as for tnt_linear_fwd with epilogue TNT_EPI_RESIDUAL (or TNT_EPI_NONE: no residual); lin->c receives the
 * NORMALISED rows.  Grouped problems normalise the N columns of every group on their own with gamma / beta block
 * g * stride_gb.  transpose_T/_B > 0 (one group, T * B == M): output row t*B + b is written to row b*T + t.
 * Supported (tnt_linear_ln_supported): TNT_NT, bf16, N = C * BN with C <= 8 CTAs per cluster and BN in {64, 128, 256}
 * (N = 512 ... 2048), no m_dyn, no dropout, 16-byte aligned operands.  A cluster of C CTAs owns a 128-row block; the row
 * statistics cross it through distributed shared memory (csrc/gemm_ln.cu). */
int tnt_linear_ln_supported(const tnt_linear_params* lin, const float* gamma, const float* beta);
int tnt_linear_ln_fwd(const tnt_linear_params* lin, const float* gamma, const float* beta, int64_t stride_gb, float eps,
                      int transpose_T, int transpose_B, tnt_stream stream);

typedef struct {
  const void* x; int64_t ldx;     /* forward input */
  const void* dy; int64_t lddy;
  void* dx; int64_t lddx;
  const float* gamma;
  float* dgamma; float* dbeta;    /* [C] fp32, accumulated with atomics (must be zeroed, or set zero_param_grads) */
  int rows, C;
  float eps;
  int dtype;
  int variant;                    /* 0 = nn.LayerNorm; 1 = final* LayerNorm (unbiased std, eps outside), transformer_v2.py:82-94 */
  int zero_param_grads;           /* 1: the call zeroes dgamma / dbeta itself (no separate memset launches) */
} tnt_layernorm_bwd_params;
int tnt_layernorm_bwd(const tnt_layernorm_bwd_params* p, tnt_stream stream);

/* ------------------------------------------------------------------------------------------
 * Token + position embedding.
 *   AdaptiveEmbedding  ttl/tell/modules/token_embedders/adaptive.py:61-76   (band table . proj * sqrt(d))
 *   Sinusoidal         ttl/tell/modules/token_embedders/positional.py:167-211,231-268
 *   Sum                ttl/tell/modules/token_embedders/sum_text_field_embedder.py:97-118
 * The per-band projection is pre-multiplied at load time (table_i . proj_i^T * sqrt(d)), so the
 * forward is a pure gather: out[t*B+b] = table[ids[b,t]] + pos_table[ids==pad ? pad : t+pad+1+start].
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const int64_t* ids;            /* [B, T] */
  const void* table;             /* [V, d] pre-projected, pre-scaled */
  const float* pos_table;        /* [n_pos, d] */
  void* out; int64_t ldo;        /* [T*B, d] (T x B x C order) */
  const int* start_pos;          /* DEVICE incremental offset or NULL (0) */
  int B, T, d, n_pos, pad_idx;
  int dtype;
  /* id / position guard (nn.Embedding raises IndexError for an id outside [0, vocab), adaptive.py:61-76; the reference
   * grows the sinusoid table on demand, positional.py:231-268).  vocab > 0: a row whose id is out of range is written as
   * NaN and bit TNT_DEVERR_TOKEN_ID is OR-ed into *error_flags; a position beyond the table reads its last row and sets
   * TNT_DEVERR_POSITION.  vocab == 0 / error_flags == NULL: unchecked ids (the caller guarantees the range). */
  int vocab;
  int* error_flags;              /* DEVICE int or NULL */
} tnt_embed_params;
enum { TNT_DEVERR_TOKEN_ID = 1, TNT_DEVERR_POSITION = 2 };
int tnt_embed_fwd(const tnt_embed_params* p, tnt_stream stream);

/* ------------------------------------------------------------------------------------------
 * Context assembly in front of the decoder (the model's _forward, between the encoders and the decoder call).
 *   ttl/tell/models/transformer_faces_objects.py:355-366 : article = sum_l softmax(bert_weight)[l] * hiddens[l]
 *                                                          over the 25 RoBERTa hidden layers, then .transpose(0,1)
 *   ttl/tell/models/transformer_faces_objects.py:371-377 : mask = isnan(x).any(-1); x[mask] = 0  (faces, objects)
 * Outputs are written in the decoder's layout [S, B, E] (the reference's .transpose(0, 1)) and dtype.
 * ------------------------------------------------------------------------------------------ */
#define TNT_MAX_MIX_LAYERS 32
typedef struct {
  const void* layers[TNT_MAX_MIX_LAYERS]; /* n_layers tensors [B, S, E], contiguous, 32-byte aligned */
  const float* weight_logits;             /* [n_layers] fp32 (bert_weight, BEFORE the softmax) */
  void* out;                              /* [S, B, E] */
  int n_layers, B, S, E;
  int dtype_in, dtype_out;
} tnt_context_mix_params;
int tnt_context_mix(const tnt_context_mix_params* p, tnt_stream stream);

typedef struct {
  const void* x;        /* [B, n, d] embeddings, NaN rows = padding */
  void* out;            /* [n, B, d] NaN rows zeroed */
  uint8_t* mask;        /* [B, n] 1 = padding */
  int B, n, d;
  int dtype_in, dtype_out;
} tnt_nan_mask_params;
int tnt_nan_mask(const tnt_nan_mask_params* p, tnt_stream stream);

/* ------------------------------------------------------------------------------------------
 * Multi-head cross attention over one or more memory sources.
 *   ttl/tell/modules/attention/multi_head.py:288-486 (manual path, static_kv)
 *   final_roberta2/transformer_v2.py:213-266        (attention(), MultiHeadedAttention)
 * Inputs are the PROJECTED q / k / v.  Per (b, h):  softmax_fp32(scale * q k^T + mask) v.
 * tell semantics (n_extra = 2): two extra keys after the S memory rows -- the learned
 * bias_k/bias_v row and an all-zero row -- never masked (multi_head.py:355-374,416-427).
 * ------------------------------------------------------------------------------------------ */
typedef enum {
  TNT_MASK_NONE = 0,
  TNT_MASK_KEY_PADDING = 1,   /* mask[b, s] != 0  -> -inf            (multi_head.py:441-449) */
  TNT_MASK_V2 = 2             /* mask[b, t, s] == 0 -> -1e9          (transformer_v2.py:217-218) */
} tnt_mask_mode;

typedef struct {
  const void* q;  int64_t q_stride_t, q_stride_b;    /* element strides; head h at column h*D */
  const void* k;  int64_t k_stride_s, k_stride_b;
  const void* v;  int64_t v_stride_s, v_stride_b;
  void* out;      int64_t o_stride_t, o_stride_b;
  const uint8_t* mask; int64_t m_stride_b, m_stride_t; /* bytes, innermost stride over s is 1 */
  const float* bias_k; const float* bias_v;          /* [H*D] fp32, n_extra == 2 only */
  float* weights;     /* optional [B, T, S + n_extra] fp32 head-AVERAGED probabilities (multi_head.py:478-482) */
  int B, T, S, H, D;
  int n_extra, mask_mode;
  float scale;
  int dtype;
  int mask_S;         /* the mask covers keys [0, mask_S); 0 = all S (trailing k/v rows are then never masked) */
  float* lse;         /* optional [B, H, T] fp32: log-sum-exp of the scaled, masked scores (saved for the backward) */
  int kv_group;       /* beam search: query rows b*kv_group .. +kv_group-1 share the k / v / mask of batch entry b
                         (B counts query rows; 0 or 1 = one k / v entry per query row) */
  tnt_dropout drop;   /* dropout on the probabilities (multi_head.py:463), index ((b*H + h)*T + t)*(S + n_extra) + key;
                         the soft-max normaliser (and `lse`) are those of the un-dropped probabilities; `weights`
                         averages the dropped probabilities, as the reference does */
  const uint32_t* tile_live; /* optional [B] (tnt_xattn_tile_mask): bit i (i < 16) set = keys [128 i, 128 i + 128) of batch row b
                         hold at least one un-masked key.  The tcgen05 kernels skip the other tiles entirely (a masked key
                         contributes exactly 0) and walk the rows longest first; NULL = visit every tile */
} tnt_xattn_params;
int tnt_xattn_fwd(const tnt_xattn_params* p, tnt_stream stream);
/* out[b] = tile bitmask | live-key count << 16 (S <= 2048) of a key-padding mask [B, >= mask_S] (mask_S = 0: S); keys >=
 * mask_S (the extra rows) are never masked */
int tnt_xattn_tile_mask(const uint8_t* mask, int64_t m_stride_b, int B, int S, int mask_S, uint32_t* out, tnt_stream stream);



/* Backward of tnt_xattn_fwd (kernel group 4): given dout (same layout as out) computes dq, dk, dv (same
 * layouts as q, k, v).  `fwd` describes the forward call (out = its output); n_extra must be 0 (callers
 * materialise the bias_k / zero rows as ordinary k / v rows, as the decoder engine does). */
typedef struct {
  tnt_xattn_params fwd;
  const void* dout;
  void* dq; void* dk; void* dv;
  void* scratch; size_t scratch_bytes;   /* tnt_xattn_bwd_scratch_bytes(B, T, H, D) */
  const float* dweights;                 /* optional [B, T, S] fp32: gradient w.r.t. the head-averaged `weights` the forward
                                            returned (multi_head.py:478-482) -- the copy heads of the pointer decoders
                                            differentiate through them; selects the SIMT kernel */
} tnt_xattn_bwd_params;
size_t tnt_xattn_bwd_scratch_bytes(int B, int T, int H, int D);
int tnt_xattn_bwd(const tnt_xattn_bwd_params* p, tnt_stream stream);

/* ------------------------------------------------------------------------------------------
 * AdaptiveSoftmax + AdaptiveLoss.
 *   ttl/tell/modules/softmax.py:144-222, ttl/tell/modules/criteria/adaptive_loss.py:27-73
 *   (= final_roberta2/adaptive_loss.py:25-70);  Generator+CE final_roberta2/transformer_v2.py:54-62 is the
 *   single-cluster case (n_clusters = 1).
 * One cluster = logits z = Xp . W^T (+bias) over `n_cols` classes.  The kernels never write z:
 *   stats[r] = (max, sum exp(z - max), argmax, z[target]) per row.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  const void* x; int64_t ldx;     /* [M, K] */
  const void* w; int64_t ldw;     /* [n_cols, K] */
  const float* bias;              /* [n_cols] or NULL */
  const int* target;              /* [M] cluster-local target column or -1; may be NULL */
  const int* row_map;             /* optional [M]: stats row index for gathered inputs */
  const int* m_dyn;               /* optional DEVICE row count */
  float* stats;                   /* [rows, 4] : max, sum exp(z-max), argmax (int bits), target logit */
  void* workspace; size_t workspace_bytes;
  int M, n_cols, K;
  int dtype;
} tnt_lse_params;
size_t tnt_lse_workspace_bytes(int M, int n_cols, int dtype);
int tnt_lse_fwd(const tnt_lse_params* p, tnt_stream stream);
/* tnt_lse_fwd without a caller workspace.  No counterpart in the reference: added for
 * callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * Same inputs, outputs and results as tnt_lse_fwd; p->workspace / workspace_bytes are ignored -- the partial statistics go
 * to an 8 MiB device array built into the library (held by every GPU that loads it until exit), and the call performs no
 * cudaMalloc / cudaFree or any other allocation.  Limit: M * ceil(n_cols / 32) * 16 bytes <= tnt_lse_fwd_static_max_bytes()
 * (e.g. 256 rows over the 50265-word vocabulary, the decode-step size); larger problems return TNT_ERR_ARG with a message
 * and must use tnt_lse_fwd.  All calls on a GPU share the array: calls that overlap on the GPU (different streams) race;
 * calls ordered on one stream are safe, and nothing needs clearing between them. */
size_t tnt_lse_fwd_static_max_bytes(void);
int tnt_lse_fwd_static(const tnt_lse_params* p, tnt_stream stream);

/* Backward of Generator + CrossEntropyLoss(ignore_index = pad, reduction = 'sum') of the final* scripts
 * (final_roberta2/transformer_v2.py:54-62, custom_v4.py:404-419, 489): from the row statistics tnt_lse_fwd wrote,
 * dz = gscale (softmax(x W^T + b) - onehot(target)) is rebuilt by the logits GEMM's softmax-gradient epilogue (never stored in
 * fp32), then dx [M, K] = dz W (dtype), dW [n_cols, K] = dz^T x (fp32, written), db [n_cols] = column sums (fp32, optional). */
size_t tnt_ce_bwd_workspace_bytes(int M, int n_cols, int dtype);
int tnt_ce_bwd(const void* x, int64_t ldx, const void* w, int64_t ldw, const float* bias, const int* target, const float* stats,
               int M, int n_cols, int K, int pad, float gscale, void* dx, int64_t lddx, float* dW, float* db, int dtype,
               void* workspace, size_t workspace_bytes, tnt_stream stream);
/* nn.Embedding backward: grad[ids[r], :] += scale * dy[r, :] (fp32 atomics) -- Embeddings.lut of the final* decoder */
int tnt_embedding_bwd(const int64_t* ids, int n, const void* dy, int64_t lddy, float scale, float* grad, int64_t ldg, int C,
                      int vocab, int dtype, tnt_stream stream);

/* ------------------------------------------------------------------------------------------
 * Decoder engine: the whole DynamicConvDecoder forward, loss, log-probs and greedy generation.
 *   DynamicConvDecoder.forward      ttl/tell/models/decoder_flattened.py:96-146
 *   DynamicConvDecoderLayer.forward ttl/tell/models/decoder_flattened.py:271-345,
 *                                   ttl/tell/models/decoder_faces_objects.py:283-393
 *   get_normalized_probs            ttl/tell/models/decoder_flattened.py:187-200
 *   _generate (greedy, top-1)       ttl/tell/models/transformer_flattened.py:238-325,
 *                                   ttl/tell/models/transformer_faces_objects.py:399-494
 * Weights are loaded under the reference's own state_dict keys (fp32 device tensors), so a
 * reference checkpoint is the input format.
 * ------------------------------------------------------------------------------------------ */
#define TNT_MAX_SOURCES 4
#define TNT_MAX_LAYERS 16
#define TNT_MAX_CLUSTERS 4

typedef struct {
  int n_layers, d, heads, ffn, vocab;
  int kernels[TNT_MAX_LAYERS];
  int n_sources;
  char source_names[TNT_MAX_SOURCES][16];  /* concat order, e.g. image, article, faces, obj */
  int source_dim[TNT_MAX_SOURCES];
  int n_clusters;                          /* head + tails */
  int cutoffs[TNT_MAX_CLUSTERS];           /* cumulative: [5000, 20000, 50265] */
  int pad_idx;                             /* 1: positions + loss ignore_index */
  int max_positions;                       /* rows of the sinusoid table */
  int dtype;                               /* TNT_BF16 or TNT_F32 (parity mode) */
  int embed_padding_idx;                   /* nn.Embedding(padding_idx) of every adaptive band table: that band-LOCAL row gets
                                              no gradient (adaptive.py:42; config.yaml:39 says 0).  < 0 = none */
} tnt_decoder_config;

typedef struct tnt_decoder tnt_decoder;

tnt_decoder* tnt_decoder_create(const tnt_decoder_config* cfg);
void tnt_decoder_destroy(tnt_decoder* d);
/* bytes of DEVICE memory the prepared (folded / cast / fused) weights need */
size_t tnt_decoder_weight_bytes(const tnt_decoder* d);
int tnt_decoder_bind_weights(tnt_decoder* d, void* weight_mem, size_t bytes);
/* feed one reference state_dict entry (fp32, device, contiguous).  Unknown keys return TNT_ERR_ARG. */
int tnt_decoder_load(tnt_decoder* d, const char* key, const float* data, const int64_t* shape, int ndim,
                     tnt_stream stream);
/* fold weight-norm, fuse per-source projections, pre-project the embedding bands */
int tnt_decoder_finalize(tnt_decoder* d, tnt_stream stream);

typedef struct {
  int B, T;
  int S[TNT_MAX_SOURCES];
  const void* memory[TNT_MAX_SOURCES];      /* [S, B, source_dim] engine dtype (reference layout) */
  const uint8_t* pad_mask[TNT_MAX_SOURCES]; /* [B, S] bool, nonzero = padding */
} tnt_contexts;

size_t tnt_decoder_workspace_bytes(const tnt_decoder* d, int B, int T, const int* S);
/* workspace of tnt_decoder_nll / tnt_decoder_log_prob for n_rows rows */
size_t tnt_decoder_loss_workspace_bytes(const tnt_decoder* d, int n_rows);

/* teacher-forced forward.  ids [B,T] -> X [B,T,d] (engine dtype); attn optional [B,T,S_w+2] fp32: the head-averaged
 * attention of the last layer over source w = the one the reference's layer returns -- "entity" when the decoder has
 * such a source (decoder_entity_pointer.py:326-334 overwrites attn last), else "article" (decoder_flattened.py:325) */
int tnt_decoder_forward(tnt_decoder* d, const int64_t* ids, const tnt_contexts* ctx, void* X_out,
                        float* attn_out, void* workspace, size_t workspace_bytes, tnt_stream stream);
/* Step-wise protocol -- `decoder(prev_target, contexts, incremental_state)` as the reference's own `_generate` loop drives
 * it (transformer_flattened.py:250-275, decoder_flattened.py:96-146, 202-207).  The CALLER owns all state (the mirror keeps
 * it in the incremental_state dict under the reference's "{Class}.{id}.{key}" names, utils/state.py:17-26, so
 * filter_incremental_state / reorder_incremental_state keep working):
 *   kv[s]      projected K / V of source s for all layers, [S+2, B, L*2d] engine dtype (tnt_decoder_kv_cache_bytes),
 *              filled once by tnt_decoder_project_kv (the reference re-projects the memories in every step);
 *   start_pos  tokens already consumed (SinusoidalPositionalEmbedding's `position`, positional.py:170-176);
 *   hist_in[l] the last hist_len[l] (<= K_l - 1) conv inputs of layer l, [hist_len, B, d] oldest first
 *              (`DynamicConv1dTBC.<id>.input_buffer`, dynamic.py:95-99); NULL / 0 before the first step;
 *   hist_out[l] receives the updated buffer: the last min(K_l - 1, hist_len[l] + T) rows (must not alias hist_in[l]).
 * ids [B, T] are the NEW tokens only; X_out [B, T, d]; attn_out optional [B, T, S_w + 2] as in tnt_decoder_forward. */
size_t tnt_decoder_kv_cache_bytes(const tnt_decoder* d, int B, int S);
int tnt_decoder_project_kv(tnt_decoder* d, const tnt_contexts* ctx, void* const* kv_out, tnt_stream stream);
size_t tnt_decoder_incremental_workspace_bytes(const tnt_decoder* d, int B, int T, const int* S);
int tnt_decoder_forward_incremental(tnt_decoder* d, const int64_t* ids, const tnt_contexts* ctx, void* const* kv, int start_pos,
                                    const void* const* hist_in, const int* hist_len, void* const* hist_out, void* X_out,
                                    float* attn_out, void* workspace, size_t workspace_bytes, tnt_stream stream);
/* AdaptiveLoss(reduction='sum'): out[0] = loss sum (nats), out[1] = number of non-pad targets */
int tnt_decoder_nll(tnt_decoder* d, const void* X, const int64_t* target, int n_rows, float* out2,
                    void* workspace, size_t workspace_bytes, tnt_stream stream);
/* AdaptiveSoftmax.get_log_prob: [n_rows, vocab] fp32 */
int tnt_decoder_log_prob(tnt_decoder* d, const void* X, int n_rows, float* out, void* workspace,
                         size_t workspace_bytes, tnt_stream stream);
/* greedy generation with incremental DynamicConv state and cached K/V.
 * seed [B] first token; ids_out [B, gen_len+1] (col 0 = seed, finished rows padded with pad_idx);
 * logp_out [B, gen_len] (0 after the row finished).  Runs all steps; returns in *steps_host (may be
 * NULL) the number of steps after which every row had finished (host sync only if non-NULL). */
/* beam-search generation (models/CaptionModel.py:23-122 scoring / candidate order / finished-beam rule; state fork =
 * reorder_incremental_state, dynamic.py:338-348).  Per caption `beam` (<= 16) hypotheses; the beams share the caption's
 * cached K/V.  Runs exactly gen_len (<= 128) steps like the reference loop.  Outputs, best first:
 * seq_out [B, beam, gen_len] (0 after the end of a hypothesis), logp_out [B, beam, gen_len] (per-word log-probs),
 * p_out [B, beam] (summed log-prob).  ctx->B = number of captions. */
size_t tnt_decoder_beam_workspace_bytes(const tnt_decoder* d, int B, int beam, int gen_len, const int* S);
int tnt_decoder_generate_beam(tnt_decoder* d, const int64_t* seed, const tnt_contexts* ctx, int gen_len, int eos, int beam,
                              int64_t* seq_out, float* logp_out, float* p_out, void* workspace, size_t workspace_bytes,
                              tnt_stream stream);
/* embedder output (adaptive band embedding * sqrt(d) + sinusoidal position, decoder_flattened.py:96-99) for ids [B, T]:
 * out [T*B, d] engine dtype, row t*B + b (the layout the layers run on).  The pointer decoder's `inp_embed`. */
int tnt_decoder_embed(tnt_decoder* d, const int64_t* ids, int B, int T, void* out_tb, tnt_stream stream);
/* Device-side input errors seen by this decoder's kernels since the last call (TNT_DEVERR_* bits, cleared on read):
 * a token id outside [0, vocab) or a position beyond the sinusoid table.  Synchronises `stream`. */
int tnt_decoder_error_flags(tnt_decoder* d, tnt_stream stream);

/* Copy head of DynamicConvDecoderPointer (decoder_pointer.py:138-141, 151-216), in place on the
 * full-vocabulary log-probabilities of tnt_decoder_log_prob:
 *   p_gen  = sigmoid(w_pgen[:d] . inp_embed + w_pgen[d:] . x + b_pgen)          project_p_gens(cat(inp_embed, X))
 *   probs  = p_gen * exp(logprob) + scatter_add(article_ids, (1 - p_gen) * attn[:, :, :S])
 *   logprob <- log(clamp(probs, 1e-10, 1)) */
typedef struct {
  const void* x; int64_t ldx;              /* decoder output [B*T, d] engine dtype, row b*T + t */
  const void* inp_embed; int64_t lde;      /* embedder output [T*B, d], row t*B + b (tnt_decoder_embed) */
  const float* w_pgen; float b_pgen;       /* weight-norm-folded project_p_gens weight [2d] and bias */
  const float* attn; int64_t attn_ld;      /* [B*T, attn_ld >= S] head-averaged article attention of the last layer
                                              (tnt_decoder_forward attn_out: the 2 extra keys are simply not read) */
  const int64_t* article_ids;              /* [B, S] token ids of the article positions */
  float* logprob; int64_t ldp;             /* [B*T, V] fp32, in / out */
  float* p_gen;                            /* optional [B*T] */
  int B, T, S, d, V, dtype;
  int64_t* best_tok; float* best_logp;     /* optional [B*T]: arg-max / max of the mixed row (greedy generation) */
} tnt_pointer_params;
/* inp_embed == NULL selects the entity-pointer head (transformer_entity_pointer.py:198-210): x is then the hidden
 * relu(project_first_p_gens(cat(inp_embed, X))) [B*T, d] and p_gen = sigmoid(w_pgen[:d] . x + b_pgen); `attn` /
 * `article_ids` are the ENTITY attention and the entity token ids. */
int tnt_pointer_mix(const tnt_pointer_params* p, tnt_stream stream);
/* F.nll_loss(logprob, target, ignore_index = pad, reduction = 'sum') (transformer_context_pointer.py:128):
 * out2 = [sum, number of counted rows] */
int tnt_nll_rows(const float* logprob, int64_t ldp, const int64_t* target, int rows, int pad, float* out2, tnt_stream stream);

/* Copy head used inside the generation loop (tnt_decoder_generate_pointer).  Plain pointer head
 * (decoder_pointer.py:138-141): w_hidden == NULL, w_pgen [2d].  Entity-pointer head (transformer_entity_pointer.py:
 * 78-79,198-200): w_hidden = weight-norm-folded project_first_p_gens weight [d_hidden, 2d] in the engine dtype,
 * b_hidden [d_hidden] fp32, w_pgen [d_hidden].  copy_ids [B, S_copy] are the token ids of the positions of the source
 * whose attention is copied (the article, or the entity source when the decoder has one). */
typedef struct {
  const float* w_pgen; float b_pgen;
  const void* w_hidden; const float* b_hidden; int d_hidden;
  const int64_t* copy_ids; int S_copy;
} tnt_pointer_head;
/* Greedy generation through the copy head -- the `_generate` loop of the pointer models with sampling_topk = 1
 * (transformer_entity_pointer.py:375-470): per step the decoder layers (incremental), the full-vocabulary
 * log-probabilities, the copy mix and the arg-max of the mixed row.  Outputs as tnt_decoder_generate. */
size_t tnt_decoder_generate_pointer_workspace_bytes(const tnt_decoder* d, int B, int gen_len, const int* S, int d_hidden);
int tnt_decoder_generate_pointer(tnt_decoder* d, const int64_t* seed, const tnt_contexts* ctx, const tnt_pointer_head* head,
                                 int gen_len, int eos, int64_t* ids_out, float* logp_out, int* steps_host,
                                 void* workspace, size_t workspace_bytes, tnt_stream stream);

size_t tnt_decoder_generate_workspace_bytes(const tnt_decoder* d, int B, int gen_len, const int* S);
int tnt_decoder_generate(tnt_decoder* d, const int64_t* seed, const tnt_contexts* ctx, int gen_len, int eos,
                         int64_t* ids_out, float* logp_out, int* steps_host, void* workspace,
                         size_t workspace_bytes, tnt_stream stream);

/* One training step (kernel group 4 + the forward): teacher-forced forward keeping activations,
 * AdaptiveLoss, and the backward pass through loss, layers, K/V projections and embedding bands.
 *   forward / loss : as tnt_decoder_forward + tnt_decoder_nll, with the dropout of tnt_decoder_set_dropout
 *   gradients      : d(loss_scale * sum-NLL)/d(parameter), fp32, WRITTEN into the buffer registered for each
 *                    reference state_dict key with tnt_decoder_set_grad (same shapes as the parameters;
 *                    weight_norm'ed linears receive weight_g / weight_v gradients, linear.py:30-33).  The tied
 *                    output rows (softmax.py:83-86,111-132) share the embedding bands' buffers:
 *                    adaptive_softmax.head.word_proj.weight / tail.i.2.weight have no buffer of their own.
 *   out2           : [loss sum (nats, unscaled), number of non-pad targets]
 * The parameter pointers given to tnt_decoder_load must still be valid (the backward reads weight_v/_g and
 * the band tables).  Data-parallel training all-reduces the gradient buffers afterwards. */
/* Dropout of the training step (config.yaml:51-69: dropout, input_dropout, weight_dropout, attention_dropout,
 * relu_dropout).  All zero (the default) = the parity configuration.  Step k after this call (k = 0, 1, ...) draws
 * its masks with seed + k.  adaptive_softmax_dropout is 0 in every reference config and is not supported. */
typedef struct {
  float dropout, input_dropout, weight_dropout, attention_dropout, relu_dropout;
  uint64_t seed;
} tnt_dropout_config;
int tnt_decoder_set_dropout(tnt_decoder* d, const tnt_dropout_config* cfg);
size_t tnt_decoder_train_workspace_bytes(const tnt_decoder* d, int B, int T, const int* S);
int tnt_decoder_set_grad(tnt_decoder* d, const char* key, float* grad);
int tnt_decoder_train_step(tnt_decoder* d, const int64_t* ids, const int64_t* target, const tnt_contexts* ctx,
                           float loss_scale, float* out2, void* workspace, size_t workspace_bytes,
                           tnt_stream stream);

/* The same step as two calls, for callers whose forward and loss are separate module calls -- the reference model runs
 * `decoder(...)`, then `criterion(...)`, then the trainer's `loss.backward()` (transformer_faces_objects.py:79-88,
 * callback_apex_trainer.py:229-233); the mirror's autograd Functions sit on these.  train_forward keeps every activation
 * in `workspace` (tnt_decoder_train_workspace_bytes) and copies X [B, T, d] out; train_backward -- same ids / contexts /
 * workspace, untouched in between -- computes the loss and all gradients exactly as tnt_decoder_train_step does. */
int tnt_decoder_train_forward(tnt_decoder* d, const int64_t* ids, const tnt_contexts* ctx, void* X_out, void* workspace,
                              size_t workspace_bytes, tnt_stream stream);
int tnt_decoder_train_backward(tnt_decoder* d, const int64_t* ids, const int64_t* target, const tnt_contexts* ctx,
                               float loss_scale, float* out2, void* workspace, size_t workspace_bytes, tnt_stream stream);

/* Training through a head that sits ON TOP of the decoder -- the copy heads of DynamicConvDecoderPointer and
 * DynamicConvDecoderEntityPointer (decoder_pointer.py:138-216, transformer_entity_pointer.py:123-131, 185-261; ttl/run.sh
 * trains h2_entity_pointer).  The heavy parts stay in the engine, the p_gen / mix arithmetic on [rows] and [rows, S]
 * tensors is left to the host's autograd:
 *   train_forward_ext   = train_forward + attn_out [B, T, S_w + 2] fp32 (last layer, head-averaged, copied source) and
 *                         emb_out [B, T, d] (`inp_embed`: the embedder output before dropout);
 *   target_logprob      lp[r] = log softmax_full(x_r)[target_r] (0 for pad targets): AdaptiveSoftmax.get_log_prob gathered
 *                       at the target without writing the [rows, V] matrix;
 *   train_backward_ext  upstream gradients row_w[r] = dL/d lp_r, dX [B, T, d] (may be NULL), dattn [B, T, S_w + 2] fp32
 *                       (may be NULL), demb [B, T, d] (may be NULL) -> every parameter gradient, as train_backward. */
int tnt_decoder_train_forward_ext(tnt_decoder* d, const int64_t* ids, const tnt_contexts* ctx, void* X_out, float* attn_out,
                                  void* emb_out, void* workspace, size_t workspace_bytes, tnt_stream stream);
int tnt_decoder_target_logprob(tnt_decoder* d, const int64_t* target, const tnt_contexts* ctx, float* lp_out, void* workspace,
                               size_t workspace_bytes, tnt_stream stream);
int tnt_decoder_train_backward_ext(tnt_decoder* d, const int64_t* ids, const int64_t* target, const tnt_contexts* ctx,
                                   const float* row_w, const void* dX, const float* dattn, const void* demb, void* workspace,
                                   size_t workspace_bytes, tnt_stream stream);

/* Overlapping the gradient exchange with the backward pass: the step completes its gradients in stages --
 * stage 0: softmax-only parameters (class_proj, tail projections); stage 1+i: every parameter of layer
 * (n_layers-1-i); last stage: the embedding bands (tied output rows).  After tnt_decoder_train_step has been
 * issued, tnt_decoder_stream_wait_stage makes `stream` wait (device side, no host sync) until that stage's
 * gradient buffers are final, so a communication stream can all-reduce them while earlier layers still run. */
int tnt_decoder_train_stages(const tnt_decoder* d);
int tnt_decoder_stream_wait_stage(tnt_decoder* d, int stage, tnt_stream stream);

/* Data-parallel training, sparse embedding gradient.  The band tables are tied to the softmax output rows: their gradient
 * is a dense softmax part (final at stage 0, so its exchange overlaps the whole backward) plus <= B*T embedding rows that
 * only exist when the backward ends.  With buffers registered here the step does NOT scatter those rows into the dense
 * tables; it packs them -- rows [B*T, d] engine dtype, meta[j] = band << 32 | band-local row, *count = rows written -- so
 * the ranks can all-gather B*T rows (4 MB in bf16) instead of all-reducing the 206 MB tables a second time, and each rank
 * applies every rank's list with tnt_embed_grad_scatter.  rows == NULL switches back to the dense scatter.
 * In this mode the band tables ("...embeddings.i.0.weight") are final at stage 0; the band projections at the last stage. */
int tnt_decoder_set_sparse_embed_grad(tnt_decoder* d, void* rows, int64_t* meta, int* count, int capacity_rows);
int tnt_embed_grad_scatter(const void* rows, int64_t ld_rows, const int64_t* meta, const int* count, int max_rows,
                           float* const* band_grads, const int* band_dims, int n_bands, int skip_row, int dtype,
                           tnt_stream stream);

/* Fused BertAdam step over flat fp32 arenas (the reference's optimizer: config.yaml:126-136, `bert_adam`):
 * g' = grad_scale * (*grad_scale_dev) * g, clipped PER PARAMETER TENSOR to max_grad_norm (segments),
 * m = b1 m + (1-b1) g', v = b2 v + (1-b2) g'^2, p -= lr * (m / (sqrt(v) + eps) + weight_decay * p).
 * Arena segments start at multiples of 64 elements; seg_of_chunk[i] is the segment of elements [64i, 64i+64). */
typedef struct {
  float* params; const float* grads; float* m; float* v;
  int64_t n;                       /* arena length, multiple of 64 */
  const int* seg_of_chunk;         /* [n / 64] or NULL when max_grad_norm <= 0 */
  float* seg_sq; int n_segments;   /* scratch [n_segments] */
  const float* grad_scale_dev;     /* optional device scalar multiplied into the gradients (e.g. 1 / n_tokens) */
  float grad_scale;                /* host scalar, 0 means 1 */
  float max_grad_norm, lr, beta1, beta2, eps, weight_decay;
  int norms_ready;                 /* 1: seg_sq already holds this step's per-tensor sums of squares (tnt_grad_sqnorm) */
  int max_ctas;                    /* > 0: grid cap -- a slice stepped beside the backward pass (trainer, stage-wise) must not
                                      take every SM's thread slots from the latency-bound kernels it overlaps; 0 = fill the GPU */
} tnt_adam_params;
int tnt_bert_adam_step(const tnt_adam_params* p, tnt_stream stream);
/* seg_sq[segment] += sum of squares of grads[first, first + n) (whole 64-element chunks): the norm pass of the step above,
 * callable per arena slice as soon as that slice's gradients are final (the trainer runs it stage by stage on a side
 * stream beside the backward pass; seg_sq must have been zeroed once per step). */
int tnt_grad_sqnorm(const float* grads, const int* seg_of_chunk, float* seg_sq, int64_t first, int64_t n, int max_ctas,
                    tnt_stream stream);   /* max_ctas as in tnt_adam_params */

/* torch.optim.Adam over flat fp32 arenas -- the optimizer of the final* scripts: Adam(betas=(0.9, 0.98), eps=1e-9) driven by
 * the Noam schedule rate(step) = factor * d_model^-0.5 * min(step^-0.5, step * warmup^-1.5)
 * (final_roberta2/custom_v4.py:421-430, 491-498; the schedule itself is host arithmetic).  g' = scale * g + wd * p,
 * bias-corrected moments at 1-based `step`: p -= lr / (1 - b1^step) * m / (sqrt(v) / sqrt(1 - b2^step) + eps). */
int tnt_adam_step(float* params, const float* grads, float* m, float* v, int64_t n, const float* grad_scale_dev,
                  float grad_scale, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                  tnt_stream stream);
/* dst += alpha * (*alpha_dev) * src over n (multiple of 4) floats: how the mirror's autograd Functions add the fused engine
 * backward (computed for an upstream gradient of 1) into `.grad`, scaled by the upstream gradient autograd hands them. */
int tnt_grad_accumulate(float* dst, const float* src, int64_t n, float alpha, const float* alpha_dev, tnt_stream stream);

/* End-to-end convenience used for the e2e benchmark: ids/targets/contexts/masks in PINNED HOST
 * memory, loss read back; performs the H2D copies, forward, loss and D2H on `stream` and
 * synchronises it.  staging = device buffer of tnt_decoder_score_host_staging_bytes(). */
size_t tnt_decoder_score_host_staging_bytes(const tnt_decoder* d, int B, int T, const int* S);
int tnt_decoder_score_host(tnt_decoder* d, const int64_t* ids_host, const int64_t* target_host,
                           const tnt_contexts* ctx_host, float* out2_host, void* staging,
                           size_t staging_bytes, void* workspace, size_t workspace_bytes,
                           tnt_stream stream);

/* Only the LIVE rows of a masked memory source cross PCIe: for caption b the rows s < len_b (len_b = 1 + its last un-masked
 * position in the host mask) are one strided 2-D copy.  A source for which that does not pay (every strided copy costs the
 * DMA engine ~3.5 us on top of its bytes: a source without padding, or a small one) crosses as ONE contiguous copy and its
 * padded tails are cleared on the device.  Either way no kernel ever sees what the host tensor holds in a padded tail.
 * (TNT_H2D_LIVE_ROWS=0 restores the full copy.)  tnt_decoder_last_h2d_bytes = bytes the last score_host / submit call
 * actually copied. */
uint64_t tnt_decoder_last_h2d_bytes(const tnt_decoder* d);

/* The live-row copy on its own (for callers that prefetch their step inputs themselves): dst [S, B, E] device, src [S, B, E]
 * pinned host, mask_host [B, S] bytes (nonzero = padding; NULL = plain copy).  Returns the bytes copied (< 0: tnt_status).
 * Skipped rows of dst keep their previous contents: zero dst once.  When skipping the padding would not pay for the strided
 * copies' fixed cost the whole tensor is copied in one piece instead (the padded rows of dst then hold what src holds). */
int64_t tnt_h2d_live_rows(void* dst, const void* src_host, const uint8_t* mask_host, int S, int B, int E, int dtype,
                          tnt_stream stream);

/* Pipelined form of tnt_decoder_score_host with two slots: submit(slot) issues the H2D copies of that
 * batch on a private copy stream (overlapping the compute of the other slot), then forward + loss + the
 * D2H of (loss, n_tokens) on `stream`; wait(slot) blocks until out2_host of that slot is valid.
 * Each slot needs its own staging buffer and out2_host; the workspace is shared. */
int tnt_decoder_score_host_submit(tnt_decoder* d, int slot, const int64_t* ids_host, const int64_t* target_host,
                                  const tnt_contexts* ctx_host, float* out2_host, void* staging,
                                  size_t staging_bytes, void* workspace, size_t workspace_bytes, tnt_stream stream);
int tnt_decoder_score_host_wait(tnt_decoder* d, int slot);

/* number of kernels this library has launched since load (for bench.py's gpu_launches) */
uint64_t tnt_launch_count(void);
/* Which kernel VARIANT the dispatchers picked since load: launches counted under `name` (-1 = unknown name; NULL resets
 * every counter).  Names: gemm.tc.1cta, gemm.tc.pair (cta_group::2), gemm.splitk, gemm.simt, xattn.tc,
 * xattn.tc.tileskip (all-padding key tiles skipped), xattn.tc.weights (head-averaged weights from the tcgen05 kernel),
 * xattn.simt, xattn.step, xattn_bwd.tc, xattn_bwd.simt, dynconv.{fwd,bwd}.{tc,simt}, dynconv.step.{fast,simt}.
 * Tests use it to pin that the code the benchmark times is the code the parity tests ran. */
int64_t tnt_variant_count(const char* name);

/* Named device-time regions, timed with CUDA events on the launching stream (off by default).
 * Regions: "kv_proj.<source>" (the K/V projection GEMM of tnt_decoder_forward / _generate / _train).
 * tnt_prof_read synchronises the recorded events and returns their summed duration and count. */
void tnt_prof_enable(int on);
void tnt_prof_reset(void);
int tnt_prof_read(const char* name, double* ms_total, int* count);

#ifdef __cplusplus
}
#endif
#endif /* TNT_B200_H_ */
