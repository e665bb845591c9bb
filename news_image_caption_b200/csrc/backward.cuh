// Internal interface of the backward-pass helper kernels (backward.cu) and the attention backward
// (attention_bwd.cu).  Kernel group 4 of the north star: gradients of DynamicConv (dynconv.cu), of the
// multi-source attention and of the AdaptiveSoftmax loss, plus the glue the training step needs.
#pragma once
#include "gemm.cuh"

namespace tnt {

// GLU backward (decoder_flattened.py:284-286): raw [N, 2C] pair-interleaved pre-activations (a0,b0,a1,b1..),
// dh [N, C] -> draw [N, 2C] (same interleaving): da = dh * sigmoid(b), db = dh * a * sigmoid(b) * (1 - sigmoid(b)).
int glu_bwd(const void* raw, long ldraw, const void* dh, long lddh, void* draw, long lddraw, int N, int C, int dtype,
            cudaStream_t st);

// out[n] (+)= sum_m x[m, n] over m < M (device count m_dyn optional).  `out` is fp32 and is zeroed first
// unless accumulate.  deinterleave: column 2j -> out[j], column 2j+1 -> out[j + N/2] (GLU-interleaved bias).
int colsum(const void* x, long ldx, int M, int N, const int* m_dyn, float* out, int dtype, bool deinterleave,
           bool accumulate, cudaStream_t st);

// weight_norm backward (linear.py:30-33): w = g v / |v|_row;  dW [rows, cols] fp32 (row r of the parameter sits
// at row glu_row(r) of dW when glu_interleave) -> dv [rows, cols], dg [rows].  g == NULL: dv = dW (plain copy).
// dst[i] = sum over g < G of parts[g * n + i] (K-split partial weight gradients)
int sum_partials(float* dst, const float* parts, int G, long n, cudaStream_t st);
int wn_bwd(const float* v, const float* g, const float* dW, long lddw, float* dv, float* dg, int rows, int cols,
           bool glu_interleave, cudaStream_t st);

// dst row (t*B + b) = src row (b*T + t)   (to_tb) or the inverse
int permute_rows(const void* src, void* dst, int B, int T, int C, int dtype, bool to_tb, cudaStream_t st);

// dst[rows[j]] += src[j] for j < *count; rows are unique (no atomics).  If map_B > 0 the destination row
// index r = b*T + t is remapped to t*B + b.
int scatter_add_rows(const void* src, long lds, const int* rows, const int* count, int max_rows, void* dst, long ldd,
                     int C, int dtype, int map_B, int map_T, cudaStream_t st);
// out[j] = src[map(rows[j])] for j < *count  (the inverse gather, same optional remap)
int gather_rows(const void* src, long lds, const int* rows, const int* count, int max_rows, void* out, long ldo, int C,
                int dtype, int map_B, int map_T, cudaStream_t st);
// out[j, :] = table[loc[j], :] (fp32 table -> engine dtype) for j < *count
int gather_table_rows(const float* table, long ldt, const int* loc, const int* count, int max_rows, void* out, long ldo,
                      int C, int dtype, cudaStream_t st);
// grad[loc[j], :] += src[j, :] (fp32 atomics; local row `skip_row` = the band's padding_idx is skipped, adaptive.py:42,51)
int scatter_table_grad(const void* src, long lds, const int* loc, const int* count, int max_rows, float* grad, long ldg,
                       int C, int dtype, int skip_row, cudaStream_t st);

// sparse embedding gradient: append band `band`'s rows (src[j], j < counts[band]) to the packed list (see backward.cu)
int pack_band_rows(const void* src, long lds, const int* loc, const int* counts, int band, int n_bands, int max_rows, void* rows_out,
                   long ldo, long long* meta, int* total, int C, int dtype, cudaStream_t st);

// dst[r, 0:C] += sum_{s0 <= s < ns} src[r, s*C : (s+1)*C]   (the residual fan-in of the per-source branches)
int add_col_blocks(void* dst, long ldd, const void* src, long lds, int N, int C, int s0, int ns, int dtype,
                   cudaStream_t st);

// gradient rows of the learned bias_k / bias_v key (multi_head.py:355-358): out[c] = sum_b dkv[(S*B + b)*ld + c]
int kv_extra_grad(const void* dkv, long ld, int S, int B, int ncols, float* out, int dtype, cudaStream_t st);

// per-row (lse, target) for the softmax-gradient GEMM epilogue (EPI_SMGRAD); ignored rows get lse = +inf
int smgrad_rows_head(const float4* head_stats, const float* side, const int* head_t, int N, int n_tails, int cut0,
                     int pad, float* row_lse, int* row_tgt, cudaStream_t st);
int smgrad_rows_tail(const float4* tail_stats, const int* tail_loc, const int* count, int N, int pad, float* row_lse,
                     int* row_tgt, cudaStream_t st);

// attention backward (attention_bwd.cu).  Same tensors as tnt_xattn_params plus the gradients; recomputes the
// softmax statistics.  dq/dk/dv use the q/k/v strides.  dk/dv rows are written for every key row < S (+ extra).
struct XattnBwd {
  tnt_xattn_params f;     // forward description (out = forward output O)
  const void* dout;       // [T,B,E] same strides as out
  void* dq; void* dk; void* dv;
  float* scratch;         // fp32 [B, H, T, D]
  const float* dweights = nullptr;   // optional [B, T, S] fp32 gradient w.r.t. the forward's head-averaged `weights` (SIMT kernel)
};
int xattn_bwd(const XattnBwd* p, cudaStream_t st);

}  // namespace tnt
