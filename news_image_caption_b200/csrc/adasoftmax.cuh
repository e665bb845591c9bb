// Internal interface of the adaptive-softmax kernels (used by the engine and the C ABI).
#pragma once
#include "gemm.cuh"

namespace tnt {

struct AdaCuts {
  int n;                          // clusters (head + tails)
  int cut[TNT_MAX_CLUSTERS];      // cumulative cutoffs, cut[n-1] = vocab
};

struct AdaSoftmax {
  AdaCuts cuts;
  int d, dtype;
  const void* head_w;                       // [cut0 + n_tails, d]  (tied band-0 embedding ; class_proj)
  const void* proj_w[TNT_MAX_CLUSTERS];     // tail i: [d, d]
  const void* out_w[TNT_MAX_CLUSTERS];      // tail i: [cut[i+1]-cut[i], d] (tied band embedding)
  // inference (static weights, bf16): out_w[i] . proj_w[i] folded once at load time, [cut[i+1]-cut[i], d] -- the tail's
  // Sequential(Linear(d, d), Dropout(0), Linear(d, n_i)) (fairseq AdaptiveSoftmax tail, softmax.py:42-60) is ONE linear map,
  // so the per-call projection GEMM of every tail disappears; NULL entries / use_fold == false: the two-GEMM form
  // (training: the factors are updated every step and their gradients are needed separately)
  const void* fold_w[TNT_MAX_CLUSTERS];
  bool use_fold;
};

size_t lse_partials_bytes(int M, int n_cols);
int lse_gemm(const void* x, long ldx, const void* w, long ldw, const float* bias, int M, int n_cols, int K, int dtype,
             const int* m_dyn, const int* target, const int* row_map, int n_lse, float* side, long side_ld,
             float4* partials, float4* stats, cudaStream_t st);

size_t adasoftmax_workspace_bytes(const AdaSoftmax& a, int N, bool logprob);
int adasoftmax_nll(const AdaSoftmax& a, const void* x, long ldx, const int64_t* target, int N, int pad, float* out2,
                   void* ws, size_t ws_bytes, cudaStream_t st);
// gradient of gscale * sum-NLL; `ws_fwd` is the workspace adasoftmax_nll ran in (row statistics are read back).
// dX [N, d] engine dtype (written), d_head_words [cut0, d], d_class_proj [n_tails, d], d_proj[i] [d, d],
// d_out[i] [n_i, d]: fp32, written.
size_t adasoftmax_bwd_workspace_bytes(const AdaSoftmax& a, int N);
// streams lent by the caller: the clusters' chains and weight gradients spread over them.  On return st has every
// contribution to dX; the parameter gradients are complete once the lent streams have drained (the caller joins them).
struct AdaSide { int n; cudaStream_t st[4]; cudaEvent_t fork; cudaEvent_t done[TNT_MAX_CLUSTERS]; };
int adasoftmax_nll_bwd(const AdaSoftmax& a, const void* x, long ldx, int N, int pad, float gscale, void* ws_fwd,
                       void* dX, long lddx, float* d_head_words, float* d_class_proj, float* const* d_proj,
                       float* const* d_out, void* ws, size_t ws_bytes, cudaStream_t st, const float* row_w = nullptr,
                       const AdaSide* sd = nullptr);
// lp[r] = log softmax over the FULL vocabulary of target[r] (0 for pad targets); leaves the row statistics in `ws` for
// adasoftmax_nll_bwd(..., row_w): the gradient of sum_r row_w[r] * (-lp[r])
int adasoftmax_target_logprob(const AdaSoftmax& a, const void* x, long ldx, const int64_t* target, int N, int pad, float* lp,
                              float* out2_scratch, void* ws, size_t ws_bytes, cudaStream_t st);
int band_rows(const int64_t* ids, int N, const int* edges, int n_bands, int* head_scratch, int* rows, int* loc,
              int* counts, cudaStream_t st);
int adasoftmax_greedy(const AdaSoftmax& a, const void* x, long ldx, int N, int64_t* tok, float* logp, void* ws,
                      size_t ws_bytes, cudaStream_t st);
int adasoftmax_logprob(const AdaSoftmax& a, const void* x, long ldx, int N, float* out, void* ws, size_t ws_bytes,
                       cudaStream_t st);

}  // namespace tnt
