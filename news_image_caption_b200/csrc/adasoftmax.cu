// AdaptiveSoftmax head + AdaptiveLoss (kernel group 3): cluster log-sum-exp / NLL / arg-max computed in
// the GEMM epilogue (gemm.cu, EPI_LSE) -- logits are never written -- plus the small finishing kernels.
// Reference: ttl/tell/modules/softmax.py:144-222 (adapt_target, forward, get_log_prob),
//            ttl/tell/modules/criteria/adaptive_loss.py:27-73, final_roberta2/adaptive_loss.py:25-70.
#include "adasoftmax.cuh"
#include "backward.cuh"

#include <algorithm>

namespace tnt {
namespace {

// merge the per-N-tile partials of one row: one warp per row
__global__ void __launch_bounds__(256) lse_finish_kernel(const float4* partials, int n_tiles, int M, const int* m_dyn,
                                                         const int* row_map, float4* stats) {
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int Me = m_dyn ? min(M, *m_dyn) : M;
  if (row >= Me) return;
  LseAcc a;
  a.init();
  for (int t = lane; t < n_tiles; t += 32) {
    const float4 p = partials[static_cast<long>(row) * n_tiles + t];
    a.merge(p.x, p.y, __float_as_int(p.z), p.w);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, a.m, o), s2 = __shfl_xor_sync(0xffffffffu, a.s, o);
    const float t2 = __shfl_xor_sync(0xffffffffu, a.tz, o);
    const int a2 = __shfl_xor_sync(0xffffffffu, a.arg, o);
    a.merge(m2, s2, a2, t2);
  }
  if (lane == 0) stats[row_map ? row_map[row] : row] = make_float4(a.m, a.s, __int_as_float(a.arg), a.tz);
}

// the same merge for SEVERAL clusters in one launch (blockIdx.y = cluster): loss / greedy calls run the clusters' GEMMs back to
// back into per-cluster partial buffers and finish them together (two launches less per call)
struct FinishJob { const float4* partials; int n_tiles; int M; const int* m_dyn; const int* row_map; float4* stats; };
struct FinishJobs { FinishJob j[TNT_MAX_CLUSTERS]; int n; };
__global__ void __launch_bounds__(256) lse_finish_multi_kernel(FinishJobs jobs) {
  pdl_wait();
  const FinishJob& J = jobs.j[blockIdx.y];
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int Me = J.m_dyn ? min(J.M, *J.m_dyn) : J.M;
  if (row >= Me) return;
  LseAcc a;
  a.init();
  for (int t = lane; t < J.n_tiles; t += 32) {
    const float4 p = J.partials[static_cast<long>(row) * J.n_tiles + t];
    a.merge(p.x, p.y, __float_as_int(p.z), p.w);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, a.m, o), s2 = __shfl_xor_sync(0xffffffffu, a.s, o);
    const float t2 = __shfl_xor_sync(0xffffffffu, a.tz, o);
    const int a2 = __shfl_xor_sync(0xffffffffu, a.arg, o);
    a.merge(m2, s2, a2, t2);
  }
  if (lane == 0) J.stats[J.row_map ? J.row_map[row] : row] = make_float4(a.m, a.s, __int_as_float(a.arg), a.tz);
}

// adapt_target (softmax.py:144-167): head targets + ordered per-tail row lists.  Single CTA -> the
// row order inside each tail is the reference's (ascending), so results are run-to-run identical.
__global__ void __launch_bounds__(1024) adapt_target_kernel(const int64_t* target, int N, AdaCuts c, int* head_t,
                                                            int* tail_rows, int* tail_loc, int* counts) {
  __shared__ int warp_cnt[TNT_MAX_CLUSTERS][32];
  __shared__ int base[TNT_MAX_CLUSTERS];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x < TNT_MAX_CLUSTERS) base[threadIdx.x] = 0;
  __syncthreads();
  for (int r0 = 0; r0 < N; r0 += 1024) {
    const int r = r0 + threadIdx.x;
    const long t = r < N ? target[r] : -1;
    int band = -1;
    if (r < N) {
      int ht = static_cast<int>(t);
      for (int i = 1; i < c.n; ++i)
        if (t >= c.cut[i - 1] && t < c.cut[i]) { band = i - 1; ht = c.cut[0] + i - 1; }
      head_t[r] = ht;
    }
    unsigned mybal = 0;
    for (int i = 0; i < c.n - 1; ++i) {
      const unsigned bal = __ballot_sync(0xffffffffu, band == i);
      if (band == i) mybal = bal;
      if (lane == 0) warp_cnt[i][warp] = __popc(bal);
    }
    __syncthreads();
    if (band >= 0) {
      int off = base[band];
      for (int w = 0; w < warp; ++w) off += warp_cnt[band][w];
      off += __popc(mybal & ((1u << lane) - 1));   // rank among the warp's lanes of the same band
      tail_rows[static_cast<long>(band) * N + off] = r;
      tail_loc[static_cast<long>(band) * N + off] = static_cast<int>(t - c.cut[band]);
    }
    __syncthreads();
    if (threadIdx.x < c.n - 1) {
      int s = 0;
      for (int w = 0; w < 32; ++w) s += warp_cnt[threadIdx.x][w];
      base[threadIdx.x] += s;
    }
    __syncthreads();
  }
  if (threadIdx.x < c.n - 1) counts[threadIdx.x] = base[threadIdx.x];
}

template <typename T>
__global__ void __launch_bounds__(256) gather_rows_kernel(const T* x, long ldx, const int* rows, const int* count, int d,
                                                          T* out) {
  const int lane = threadIdx.x & 31;
  const int n = *count;
  for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
    const T* s = x + static_cast<long>(rows[j]) * ldx;
    T* o = out + static_cast<long>(j) * d;
    for (int vc = lane; vc < (d >> 3); vc += 32) {
      float v[8];
      load8(s + vc * 8, v);
      store8(o + vc * 8, v);
    }
  }
}

__device__ __forceinline__ float head_lse(const float4 st, const float* side, int n_tails) {
  float M = st.x;
  for (int i = 0; i < n_tails; ++i) M = fmaxf(M, side[i]);
  float s = st.y * __expf(st.x - M);
  for (int i = 0; i < n_tails; ++i) s += __expf(side[i] - M);
  return M + __logf(s);
}

// AdaptiveLoss(reduction='sum') (adaptive_loss.py:55-65): deterministic single-CTA reduction.
__global__ void __launch_bounds__(1024) nll_reduce_kernel(const int64_t* target, const int* head_t, int N, AdaCuts c,
                                                          int pad, const float4* head_stats, const float* side,
                                                          const float4* tail_stats, const int* tail_loc,
                                                          const int* counts, float* out2) {
  __shared__ float red[32];
  float loss = 0.f, ntok = 0.f;
  const int nt = c.n - 1;
  for (int r = threadIdx.x; r < N; r += 1024) {
    if (target[r] != pad) ntok += 1.f;
    const int ht = head_t[r];
    if (ht != pad) {                                    // ignore_index inside the head cluster
      const float lse = head_lse(head_stats[r], side + static_cast<long>(r) * nt, nt);
      const float zt = ht < c.cut[0] ? head_stats[r].w : side[static_cast<long>(r) * nt + ht - c.cut[0]];
      loss += lse - zt;
    }
  }
  for (int i = 0; i < nt; ++i) {
    const int n = counts[i];
    for (int j = threadIdx.x; j < n; j += 1024) {
      if (tail_loc[static_cast<long>(i) * N + j] != pad) {  // ignore_index inside the tail (local id 1)
        const float4 st = tail_stats[static_cast<long>(i) * N + j];
        loss += st.x + __logf(st.y) - st.w;
      }
    }
  }
  loss = block_sum(loss, red);
  ntok = block_sum(ntok, red);
  if (threadIdx.x == 0) {
    out2[0] = loss;
    out2[1] = ntok;
  }
}

// get_log_prob (softmax.py:193-222) gathered at the target: lp[r] = log softmax_full(x_r)[target_r] -- head term plus, for a
// target in tail i, the tail term (NO in-cluster ignore_index here: this is F.nll_loss over the full-vocabulary
// log-probabilities, as the pointer models compute their loss, transformer_entity_pointer.py:123-131).  pad targets -> 0.
__global__ void __launch_bounds__(1024) target_logprob_kernel(const int64_t* target, const int* head_t, int N, AdaCuts c, int pad,
                                                              const float4* head_stats, const float* side,
                                                              const float4* tail_stats, const int* tail_rows, const int* counts,
                                                              float* lp) {
  const int nt = c.n - 1;
  for (int r = threadIdx.x; r < N; r += 1024) {
    float v = 0.f;
    if (target[r] != pad) {
      const int ht = head_t[r];
      const float lse = head_lse(head_stats[r], side + static_cast<long>(r) * nt, nt);
      const float zt = ht < c.cut[0] ? head_stats[r].w : side[static_cast<long>(r) * nt + ht - c.cut[0]];
      v = zt - lse;
    }
    lp[r] = v;
  }
  __syncthreads();
  for (int i = 0; i < nt; ++i) {
    const int n = counts[i];
    for (int j = threadIdx.x; j < n; j += 1024) {
      const float4 st = tail_stats[static_cast<long>(i) * N + j];
      lp[tail_rows[static_cast<long>(i) * N + j]] += st.w - (st.x + __logf(st.y));
    }
  }
}

// greedy pick: arg-max over the full vocabulary of get_log_prob (softmax.py:193-222) without building it
__global__ void __launch_bounds__(256) greedy_pick_kernel(int N, AdaCuts c, const float4* head_stats, const float* side,
                                                          const float4* tail_stats, int64_t* tok, float* logp) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= N) return;
  const int nt = c.n - 1;
  const float4 hs = head_stats[r];
  const float lse = head_lse(hs, side + static_cast<long>(r) * nt, nt);
  float best = hs.x - lse;
  long arg = __float_as_int(hs.z);
  for (int i = 0; i < nt; ++i) {
    const float4 ts = tail_stats[static_cast<long>(i) * N + r];
    const float lp = side[static_cast<long>(r) * nt + i] - lse - __logf(ts.y);   // prior + (max - lse_tail)
    if (lp > best) {
      best = lp;
      arg = c.cut[i] + __float_as_int(ts.z);
    }
  }
  tok[r] = arg;
  logp[r] = best;
}

// get_log_prob finishing pass over materialised logits: out row = [head words | tail_0 | tail_1 ...]
// `head` [N, cut0 + nt] raw head logits (scratch); tails already sit raw in `out`.
__global__ void __launch_bounds__(256) logprob_norm_kernel(int N, AdaCuts c, const float* head, float* out) {
  __shared__ float red[32];
  __shared__ float bc;
  const int r = blockIdx.x, nt = c.n - 1, nh = c.cut[0] + nt, V = c.cut[c.n - 1];
  const float* h = head + static_cast<long>(r) * nh;
  float* o = out + static_cast<long>(r) * V;
  float m = -INFINITY;
  for (int i = threadIdx.x; i < nh; i += blockDim.x) m = fmaxf(m, h[i]);
  m = warp_max(m);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = red[0];
    for (int w = 1; w < (blockDim.x >> 5); ++w) t = fmaxf(t, red[w]);
    bc = t;
  }
  __syncthreads();
  m = bc;
  float s = 0.f;
  for (int i = threadIdx.x; i < nh; i += blockDim.x) s += __expf(h[i] - m);
  s = block_sum(s, red);
  const float lse = m + logf(s);
  for (int i = threadIdx.x; i < c.cut[0]; i += blockDim.x) o[i] = h[i] - lse;
  for (int k = 0; k < nt; ++k) {
    const float prior = h[c.cut[0] + k] - lse;
    float* z = o + c.cut[k];
    const int n = c.cut[k + 1] - c.cut[k];
    float tm = -INFINITY;
    for (int i = threadIdx.x; i < n; i += blockDim.x) tm = fmaxf(tm, z[i]);
    tm = warp_max(tm);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = tm;
    __syncthreads();
    if (threadIdx.x == 0) {
      float t = red[0];
      for (int w = 1; w < (blockDim.x >> 5); ++w) t = fmaxf(t, red[w]);
      bc = t;
    }
    __syncthreads();
    tm = bc;
    float ts = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) ts += __expf(z[i] - tm);
    ts = block_sum(ts, red);
    const float sub = tm + logf(ts) - prior;
    for (int i = threadIdx.x; i < n; i += blockDim.x) z[i] -= sub;
  }
}

inline size_t al256(size_t x) { return (x + 255) & ~size_t(255); }

}  // namespace

static bool finish_merged() {     // TNT_LSE_FINISH_MERGED=0: one finish launch per cluster, as before
  static const bool on = [] { const char* e = getenv("TNT_LSE_FINISH_MERGED"); return !(e && e[0] == '0'); }();
  return on;
}

static int lse_gemm_deferred(const void* x, long ldx, const void* w, long ldw, const float* bias, int M, int n_cols, int K, int dtype,
                             const int* m_dyn, const int* target, const int* row_map, int n_lse, float* side, long side_ld,
                             float4* partials, float4* stats, FinishJobs* jobs, cudaStream_t st) {
  tnt_linear_params lp = {};
  lp.a = x; lp.lda = ldx; lp.w = w; lp.ldw = ldw; lp.bias = bias;
  lp.M = M; lp.N = n_cols; lp.K = K; lp.m_dyn = m_dyn; lp.groups = 1;
  lp.dtype_in = dtype; lp.dtype_out = TNT_F32; lp.epilogue = EPI_LSE;
  LseExtra ex = {target, partials, n_lse, side, side_ld};
  int n_tiles = 0;
  TNT_TRY(launch_gemm(&lp, EPI_LSE, &ex, &n_tiles, st));
  jobs->j[jobs->n++] = FinishJob{partials, n_tiles, M, m_dyn, row_map, stats};
  return TNT_OK;
}

static int lse_finish_jobs(const FinishJobs& jobs, cudaStream_t st) {
  if (!jobs.n) return TNT_OK;
  int maxM = 0;
  for (int i = 0; i < jobs.n; ++i) maxM = std::max(maxM, jobs.j[i].M);
  TNT_CUDA(launch_pdl(lse_finish_multi_kernel, dim3((maxM + 7) / 8, jobs.n), dim3(256), 0, st, 1, jobs));
  TNT_LAUNCHED();
  return TNT_OK;
}

int lse_gemm(const void* x, long ldx, const void* w, long ldw, const float* bias, int M, int n_cols, int K, int dtype,
             const int* m_dyn, const int* target, const int* row_map, int n_lse, float* side, long side_ld,
             float4* partials, float4* stats, cudaStream_t st) {
  tnt_linear_params lp = {};
  lp.a = x; lp.lda = ldx; lp.w = w; lp.ldw = ldw; lp.bias = bias;
  lp.M = M; lp.N = n_cols; lp.K = K; lp.m_dyn = m_dyn; lp.groups = 1;
  lp.dtype_in = dtype; lp.dtype_out = TNT_F32; lp.epilogue = EPI_LSE;
  LseExtra ex = {target, partials, n_lse, side, side_ld};
  int n_tiles = 0;
  TNT_TRY(launch_gemm(&lp, EPI_LSE, &ex, &n_tiles, st));
  TNT_CUDA(launch_pdl(lse_finish_kernel, dim3((M + 7) / 8), dim3(256), 0, st, 1, partials, n_tiles, M, m_dyn, row_map, stats));
  TNT_LAUNCHED();
  return TNT_OK;
}

size_t lse_partials_bytes(int M, int n_cols) { return al256(static_cast<size_t>(M) * ((n_cols + 31) / 32) * sizeof(float4)); }

// ---------------------------------------------------------------------------- workspace layout
struct AdaWs {
  int* head_t; int* tail_rows; int* tail_loc; int* counts;
  float4* head_stats; float* side; float4* tail_stats; float4* partials;
  float4* partials_c[TNT_MAX_CLUSTERS];     // per cluster (0 = head, 1.. = tails): the merged finish reads them after the last GEMM
  void* xc; void* proj; float* head_raw;
  size_t total;
};

static AdaWs carve(const AdaSoftmax& a, int N, void* ws, bool need_raw) {
  AdaWs w;
  const int nt = a.cuts.n - 1;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += al256(bytes);
    return p;
  };
  int max_cols = a.cuts.cut[0] + nt;
  for (int i = 0; i < nt; ++i) max_cols = std::max(max_cols, a.cuts.cut[i + 1] - a.cuts.cut[i]);
  w.head_t = static_cast<int*>(take(sizeof(int) * N));
  w.tail_rows = static_cast<int*>(take(sizeof(int) * N * std::max(nt, 1)));
  w.tail_loc = static_cast<int*>(take(sizeof(int) * N * std::max(nt, 1)));
  w.counts = static_cast<int*>(take(sizeof(int) * TNT_MAX_CLUSTERS));
  w.head_stats = static_cast<float4*>(take(sizeof(float4) * N));
  w.side = static_cast<float*>(take(sizeof(float) * N * std::max(nt, 1)));
  w.tail_stats = static_cast<float4*>(take(sizeof(float4) * N * std::max(nt, 1)));
  w.partials = static_cast<float4*>(take(lse_partials_bytes(N, max_cols)));
  for (int i = 0; i < TNT_MAX_CLUSTERS; ++i) w.partials_c[i] = nullptr;
  w.partials_c[0] = static_cast<float4*>(take(lse_partials_bytes(N, a.cuts.cut[0] + nt)));
  for (int i = 0; i < nt && i + 1 < TNT_MAX_CLUSTERS; ++i)
    w.partials_c[i + 1] = static_cast<float4*>(take(lse_partials_bytes(N, a.cuts.cut[i + 1] - a.cuts.cut[i])));
  w.xc = take(dtype_size(a.dtype) * N * a.d);
  w.proj = take(dtype_size(a.dtype) * N * a.d);
  w.head_raw = need_raw ? static_cast<float*>(take(sizeof(float) * N * (a.cuts.cut[0] + nt))) : nullptr;
  w.total = off;
  return w;
}

size_t adasoftmax_workspace_bytes(const AdaSoftmax& a, int N, bool logprob) { return carve(a, N, nullptr, logprob).total; }

static int tail_proj(const AdaSoftmax& a, int i, const void* x, long ldx, int M, const int* m_dyn, void* out,
                     cudaStream_t st) {
  tnt_linear_params lp = {};
  lp.a = x; lp.lda = ldx; lp.w = a.proj_w[i]; lp.ldw = a.d; lp.c = out; lp.ldc = a.d;
  lp.M = M; lp.N = a.d; lp.K = a.d; lp.m_dyn = m_dyn; lp.groups = 1;
  lp.dtype_in = a.dtype; lp.dtype_out = a.dtype; lp.epilogue = EPI_NONE;
  return launch_gemm(&lp, -1, nullptr, nullptr, st);
}

int adasoftmax_nll(const AdaSoftmax& a, const void* x, long ldx, const int64_t* target, int N, int pad, float* out2,
                   void* ws, size_t ws_bytes, cudaStream_t st) {
  AdaWs w = carve(a, N, ws, false);
  if (w.total > ws_bytes) {
    set_error("adasoftmax_nll: workspace %zu < %zu", ws_bytes, w.total);
    return TNT_ERR_WORKSPACE;
  }
  const int nt = a.cuts.n - 1;
  adapt_target_kernel<<<1, 1024, 0, st>>>(target, N, a.cuts, w.head_t, w.tail_rows, w.tail_loc, w.counts);
  TNT_LAUNCHED();
  const bool merged = finish_merged() && nt + 1 <= TNT_MAX_CLUSTERS;
  FinishJobs jobs = {};
  if (merged)
    TNT_TRY(lse_gemm_deferred(x, ldx, a.head_w, a.d, nullptr, N, a.cuts.cut[0] + nt, a.d, a.dtype, nullptr, w.head_t, nullptr,
                              a.cuts.cut[0], w.side, nt, w.partials_c[0], w.head_stats, &jobs, st));
  else
    TNT_TRY(lse_gemm(x, ldx, a.head_w, a.d, nullptr, N, a.cuts.cut[0] + nt, a.d, a.dtype, nullptr, w.head_t, nullptr,
                     a.cuts.cut[0], w.side, nt, w.partials, w.head_stats, st));
  for (int i = 0; i < nt; ++i) {
    const int grid = std::min((N + 7) / 8, 4 * num_sms());
    if (a.dtype == TNT_BF16)
      gather_rows_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<const bf16*>(x), ldx, w.tail_rows + static_cast<long>(i) * N,
                                                     w.counts + i, a.d, static_cast<bf16*>(w.xc));
    else
      gather_rows_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(x), ldx,
                                                      w.tail_rows + static_cast<long>(i) * N, w.counts + i, a.d,
                                                      static_cast<float*>(w.xc));
    TNT_LAUNCHED();
    const bool fold = a.use_fold && a.fold_w[i];
    if (!fold) TNT_TRY(tail_proj(a, i, w.xc, a.d, N, w.counts + i, w.proj, st));
    const int ncols = a.cuts.cut[i + 1] - a.cuts.cut[i];
    if (merged)
      TNT_TRY(lse_gemm_deferred(fold ? w.xc : w.proj, a.d, fold ? a.fold_w[i] : a.out_w[i], a.d, nullptr, N, ncols, a.d, a.dtype,
                                w.counts + i, w.tail_loc + static_cast<long>(i) * N, nullptr, ncols, nullptr, 0, w.partials_c[i + 1],
                                w.tail_stats + static_cast<long>(i) * N, &jobs, st));
    else
      TNT_TRY(lse_gemm(fold ? w.xc : w.proj, a.d, fold ? a.fold_w[i] : a.out_w[i], a.d, nullptr, N, ncols, a.d, a.dtype, w.counts + i,
                       w.tail_loc + static_cast<long>(i) * N, nullptr, ncols, nullptr, 0, w.partials,
                       w.tail_stats + static_cast<long>(i) * N, st));
  }
  TNT_TRY(lse_finish_jobs(jobs, st));
  nll_reduce_kernel<<<1, 1024, 0, st>>>(target, w.head_t, N, a.cuts, pad, w.head_stats, w.side, w.tail_stats, w.tail_loc,
                                        w.counts, out2);
  TNT_LAUNCHED();
  return TNT_OK;
}

int adasoftmax_target_logprob(const AdaSoftmax& a, const void* x, long ldx, const int64_t* target, int N, int pad, float* lp,
                              float* out2_scratch, void* ws, size_t ws_bytes, cudaStream_t st) {
  TNT_TRY(adasoftmax_nll(a, x, ldx, target, N, pad, out2_scratch, ws, ws_bytes, st));     // the row statistics of every cluster
  AdaWs w = carve(a, N, ws, false);
  target_logprob_kernel<<<1, 1024, 0, st>>>(target, w.head_t, N, a.cuts, pad, w.head_stats, w.side, w.tail_stats, w.tail_rows,
                                            w.counts, lp);
  TNT_LAUNCHED();
  return TNT_OK;
}

int adasoftmax_greedy(const AdaSoftmax& a, const void* x, long ldx, int N, int64_t* tok, float* logp, void* ws,
                      size_t ws_bytes, cudaStream_t st) {
  AdaWs w = carve(a, N, ws, false);
  if (w.total > ws_bytes) {
    set_error("adasoftmax_greedy: workspace %zu < %zu", ws_bytes, w.total);
    return TNT_ERR_WORKSPACE;
  }
  const int nt = a.cuts.n - 1;
  const bool merged = finish_merged() && nt + 1 <= TNT_MAX_CLUSTERS;
  FinishJobs jobs = {};
  if (merged)
    TNT_TRY(lse_gemm_deferred(x, ldx, a.head_w, a.d, nullptr, N, a.cuts.cut[0] + nt, a.d, a.dtype, nullptr, nullptr, nullptr,
                              a.cuts.cut[0], w.side, nt, w.partials_c[0], w.head_stats, &jobs, st));
  else
    TNT_TRY(lse_gemm(x, ldx, a.head_w, a.d, nullptr, N, a.cuts.cut[0] + nt, a.d, a.dtype, nullptr, nullptr, nullptr,
                     a.cuts.cut[0], w.side, nt, w.partials, w.head_stats, st));
  for (int i = 0; i < nt; ++i) {
    const bool fold = a.use_fold && a.fold_w[i];
    if (!fold) TNT_TRY(tail_proj(a, i, x, ldx, N, nullptr, w.proj, st));
    const int ncols = a.cuts.cut[i + 1] - a.cuts.cut[i];
    if (merged)
      TNT_TRY(lse_gemm_deferred(fold ? x : w.proj, fold ? ldx : a.d, fold ? a.fold_w[i] : a.out_w[i], a.d, nullptr, N, ncols, a.d,
                                a.dtype, nullptr, nullptr, nullptr, ncols, nullptr, 0, w.partials_c[i + 1],
                                w.tail_stats + static_cast<long>(i) * N, &jobs, st));
    else
      TNT_TRY(lse_gemm(fold ? x : w.proj, fold ? ldx : a.d, fold ? a.fold_w[i] : a.out_w[i], a.d, nullptr, N, ncols, a.d, a.dtype,
                       nullptr, nullptr, nullptr, ncols, nullptr, 0, w.partials, w.tail_stats + static_cast<long>(i) * N, st));
  }
  TNT_TRY(lse_finish_jobs(jobs, st));
  greedy_pick_kernel<<<(N + 255) / 256, 256, 0, st>>>(N, a.cuts, w.head_stats, w.side, w.tail_stats, tok, logp);
  TNT_LAUNCHED();
  return TNT_OK;
}

int adasoftmax_logprob(const AdaSoftmax& a, const void* x, long ldx, int N, float* out, void* ws, size_t ws_bytes,
                       cudaStream_t st) {
  AdaWs w = carve(a, N, ws, true);
  if (w.total > ws_bytes) {
    set_error("adasoftmax_logprob: workspace %zu < %zu", ws_bytes, w.total);
    return TNT_ERR_WORKSPACE;
  }
  const int nt = a.cuts.n - 1, V = a.cuts.cut[a.cuts.n - 1], nh = a.cuts.cut[0] + nt;
  tnt_linear_params lp = {};
  lp.a = x; lp.lda = ldx; lp.w = a.head_w; lp.ldw = a.d; lp.c = w.head_raw; lp.ldc = nh;
  lp.M = N; lp.N = nh; lp.K = a.d; lp.groups = 1; lp.dtype_in = a.dtype; lp.dtype_out = TNT_F32; lp.epilogue = EPI_NONE;
  TNT_TRY(launch_gemm(&lp, -1, nullptr, nullptr, st));
  for (int i = 0; i < nt; ++i) {
    const bool fold = a.use_fold && a.fold_w[i];
    if (!fold) TNT_TRY(tail_proj(a, i, x, ldx, N, nullptr, w.proj, st));
    tnt_linear_params tp = {};
    tp.a = fold ? x : w.proj; tp.lda = fold ? ldx : a.d; tp.w = fold ? a.fold_w[i] : a.out_w[i]; tp.ldw = a.d;
    tp.c = out + a.cuts.cut[i]; tp.ldc = V;
    tp.M = N; tp.N = a.cuts.cut[i + 1] - a.cuts.cut[i]; tp.K = a.d; tp.groups = 1;
    tp.dtype_in = a.dtype; tp.dtype_out = TNT_F32; tp.epilogue = EPI_NONE;
    TNT_TRY(launch_gemm(&tp, -1, nullptr, nullptr, st));
  }
  logprob_norm_kernel<<<N, 256, 0, st>>>(N, a.cuts, w.head_raw, out);
  TNT_LAUNCHED();
  return TNT_OK;
}


// ---------------------------------------------------------------------------- backward (kernel group 4)
// Gradient of gscale * AdaptiveLoss(sum) w.r.t. X and the softmax weights.  Needs the row statistics the
// forward (adasoftmax_nll) left in its workspace.  Per cluster: dZ = gscale (softmax(z) - onehot) is produced
// by re-running the logits GEMM with the EPI_SMGRAD epilogue (bf16 dZ, never fp32 logits), then
// dX = dZ . W (NN) and dW = dZ^T . X (TN).  Tail rows are gathered / scattered through the forward's row lists.
struct AdaBwdWs {
  // cluster c = 0: head, c = 1 + i: tail i.  Every cluster has buffers of its own: the clusters' chains (and their
  // weight-gradient GEMMs) run side by side on different streams when the caller lends some (AdaSide)
  float* row_lse[TNT_MAX_CLUSTERS]; int* row_tgt[TNT_MAX_CLUSTERS]; void* dz[TNT_MAX_CLUSTERS]; long ldz[TNT_MAX_CLUSTERS];
  void* xc[TNT_MAX_CLUSTERS]; void* proj[TNT_MAX_CLUSTERS]; void* dact[TNT_MAX_CLUSTERS]; void* dxc[TNT_MAX_CLUSTERS];
  float* row_w[TNT_MAX_CLUSTERS];
  float* dhead;
  size_t total;
};

static AdaBwdWs carve_bwd(const AdaSoftmax& a, int N, void* ws) {
  AdaBwdWs w = {};
  const int nt = a.cuts.n - 1;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += al256(bytes);
    return p;
  };
  const size_t E = dtype_size(a.dtype);
  for (int c = 0; c <= nt; ++c) {
    const int cols = c == 0 ? a.cuts.cut[0] + nt : a.cuts.cut[c] - a.cuts.cut[c - 1];
    w.ldz[c] = (cols + 7) & ~7;
    w.row_lse[c] = static_cast<float*>(take(sizeof(float) * N));
    w.row_tgt[c] = static_cast<int*>(take(sizeof(int) * N));
    w.dz[c] = take(E * N * w.ldz[c]);
    if (c > 0) {
      w.xc[c] = take(E * N * a.d);
      w.proj[c] = take(E * N * a.d);
      w.dact[c] = take(E * N * a.d);
      w.dxc[c] = take(E * N * a.d);
      w.row_w[c] = static_cast<float*>(take(sizeof(float) * N));
    }
  }
  w.dhead = static_cast<float*>(take(sizeof(float) * (a.cuts.cut[0] + nt) * a.d));
  w.total = off;
  return w;
}

// rows of `ids` per vocabulary band [edge[i], edge[i+1]) (ascending row order, as adapt_target builds them for
// the tails): rows / loc are [n_bands, N], counts [n_bands]; head_scratch [N] receives the head targets (unused).
int band_rows(const int64_t* ids, int N, const int* edges, int n_bands, int* head_scratch, int* rows, int* loc,
              int* counts, cudaStream_t st) {
  TNT_CHECK_ARG(n_bands >= 1 && n_bands + 1 <= TNT_MAX_CLUSTERS + 1 && n_bands <= TNT_MAX_CLUSTERS - 1 + 1, "band_rows: bands");
  AdaCuts c;
  c.n = n_bands + 1;                       // cluster 0 = ids below edges[0] (none when edges[0] == 0)
  TNT_CHECK_ARG(c.n <= TNT_MAX_CLUSTERS, "band_rows: at most %d bands", TNT_MAX_CLUSTERS - 1);
  for (int i = 0; i < c.n; ++i) c.cut[i] = edges[i];
  adapt_target_kernel<<<1, 1024, 0, st>>>(ids, N, c, head_scratch, rows, loc, counts);
  TNT_LAUNCHED();
  return TNT_OK;
}

size_t adasoftmax_bwd_workspace_bytes(const AdaSoftmax& a, int N) { return carve_bwd(a, N, nullptr).total; }

static int gemm_call(const void* A, long lda, const void* W, long ldw, void* C, long ldc, int M, int N, int K, int dt_in,
                     int dt_out, int layout, int epi, const int* m_dyn, const void* res, long ldr, const LseExtra* ex,
                     cudaStream_t st) {
  tnt_linear_params lp = {};
  lp.a = A; lp.lda = lda; lp.w = W; lp.ldw = ldw; lp.c = C; lp.ldc = ldc; lp.residual = res; lp.ldr = ldr;
  lp.M = M; lp.N = N; lp.K = K; lp.m_dyn = m_dyn; lp.groups = 1;
  lp.dtype_in = dt_in; lp.dtype_out = dt_out; lp.epilogue = epi == EPI_SMGRAD ? EPI_NONE : epi; lp.layout = layout;
  return launch_gemm(&lp, epi, ex, nullptr, st);
}

int gather_floats(const float* src, const int* rows, const int* count, int N, float* out, cudaStream_t st);   // backward.cu

int adasoftmax_nll_bwd(const AdaSoftmax& a, const void* x, long ldx, int N, int pad, float gscale, void* ws_fwd,
                       void* dX, long lddx, float* d_head_words, float* d_class_proj, float* const* d_proj,
                       float* const* d_out, void* ws, size_t ws_bytes, cudaStream_t st, const float* row_w, const AdaSide* sd) {
  // row_w != NULL: gradient of sum_r row_w[r] * (-lp[r]) with lp = adasoftmax_target_logprob (full log-softmax of the
  // target: the tail term counts for every tail row) instead of the AdaptiveLoss sum with its in-cluster ignore_index
  AdaWs f = carve(a, N, ws_fwd, false);
  AdaBwdWs w = carve_bwd(a, N, ws);
  if (w.total > ws_bytes) {
    set_error("adasoftmax_nll_bwd: workspace %zu < %zu", ws_bytes, w.total);
    return TNT_ERR_WORKSPACE;
  }
  const int nt = a.cuts.n - 1, d = a.d, dt = a.dtype, nh = a.cuts.cut[0] + nt;
  const size_t E = dtype_size(dt);
  // The clusters are independent until their gradients meet in dX, and each is a chain of a few GEMMs with awkward shapes
  // (a tail's data gradient reduces over up to 30265 columns into a handful of output tiles).  With streams lent by the
  // caller (sd) every tail's chain runs on a stream of its own beside the head's (on st), the weight gradients -- nothing
  // waits for them here -- follow on those streams, and st only waits for the tails' dxc before it scatter-adds them.
  const bool par = sd && sd->n > 0;
  auto tail_st = [&](int i) { return par ? sd->st[i % sd->n] : st; };
  cudaStream_t head_wst = par ? sd->st[nt % sd->n] : st;
  if (par) {
    TNT_CUDA(cudaEventRecord(sd->fork, st));
    for (int k = 0; k < sd->n; ++k) TNT_CUDA(cudaStreamWaitEvent(sd->st[k], sd->fork, 0));
  }
  // ---- tails
  for (int i = 0; i < nt; ++i) {
    const int c = i + 1;
    cudaStream_t ts = tail_st(i);
    const int* cnt = f.counts + i;
    const int* rows = f.tail_rows + static_cast<long>(i) * N;
    const int ncols = a.cuts.cut[i + 1] - a.cuts.cut[i];
    // stale rows of the gather / projection buffers must be finite (0 x NaN inside a 64-row reduction block)
    TNT_CUDA(cudaMemsetAsync(w.xc[c], 0, E * N * d, ts));
    TNT_CUDA(cudaMemsetAsync(w.proj[c], 0, E * N * d, ts));
    TNT_TRY(gather_rows(x, ldx, rows, cnt, N, w.xc[c], d, d, dt, 0, 0, ts));
    TNT_TRY(tail_proj(a, i, w.xc[c], d, N, cnt, w.proj[c], ts));
    TNT_TRY(smgrad_rows_tail(f.tail_stats + static_cast<long>(i) * N, f.tail_loc + static_cast<long>(i) * N, cnt, N,
                             row_w ? -1 : pad, w.row_lse[c], w.row_tgt[c], ts));
    LseExtra ex = {w.row_tgt[c], nullptr, 0, nullptr, 0, w.row_lse[c], gscale};
    if (row_w) {
      TNT_TRY(gather_floats(row_w, rows, cnt, N, w.row_w[c], ts));
      ex.row_w = w.row_w[c];
    }
    TNT_TRY(gemm_call(w.proj[c], d, a.out_w[i], d, w.dz[c], w.ldz[c], N, ncols, d, dt, dt, TNT_NT, EPI_SMGRAD, cnt, nullptr, 0, &ex, ts));
    TNT_TRY(gemm_call(w.dz[c], w.ldz[c], a.out_w[i], d, w.dact[c], d, N, d, ncols, dt, dt, TNT_NN, EPI_NONE, cnt, nullptr, 0, nullptr, ts));
    TNT_TRY(gemm_call(w.dact[c], d, a.proj_w[i], d, w.dxc[c], d, N, d, d, dt, dt, TNT_NN, EPI_NONE, cnt, nullptr, 0, nullptr, ts));
    if (par) TNT_CUDA(cudaEventRecord(sd->done[i], ts));
    TNT_TRY(gemm_call(w.dz[c], w.ldz[c], w.proj[c], d, d_out[i], d, ncols, d, N, dt, TNT_F32, TNT_TN, EPI_NONE, cnt, nullptr, 0, nullptr, ts));
    TNT_TRY(gemm_call(w.dact[c], d, w.xc[c], d, d_proj[i], d, d, d, N, dt, TNT_F32, TNT_TN, EPI_NONE, cnt, nullptr, 0, nullptr, ts));
  }
  // ---- head cluster (st); its weight gradient on a lent stream
  TNT_TRY(smgrad_rows_head(f.head_stats, f.side, f.head_t, N, nt, a.cuts.cut[0], pad, w.row_lse[0], w.row_tgt[0], st));
  LseExtra ex = {w.row_tgt[0], nullptr, 0, nullptr, 0, w.row_lse[0], gscale};
  ex.row_w = row_w;
  TNT_TRY(gemm_call(x, ldx, a.head_w, d, w.dz[0], w.ldz[0], N, nh, d, dt, dt, TNT_NT, EPI_SMGRAD, nullptr, nullptr, 0, &ex, st));
  if (par) {
    TNT_CUDA(cudaEventRecord(sd->fork, st));
    TNT_CUDA(cudaStreamWaitEvent(head_wst, sd->fork, 0));
  }
  TNT_TRY(gemm_call(w.dz[0], w.ldz[0], x, ldx, w.dhead, d, nh, d, N, dt, TNT_F32, TNT_TN, EPI_NONE, nullptr, nullptr, 0, nullptr, head_wst));
  TNT_CUDA(cudaMemcpyAsync(d_head_words, w.dhead, sizeof(float) * a.cuts.cut[0] * d, cudaMemcpyDeviceToDevice, head_wst));
  if (nt > 0)
    TNT_CUDA(cudaMemcpyAsync(d_class_proj, w.dhead + static_cast<size_t>(a.cuts.cut[0]) * d, sizeof(float) * nt * d,
                             cudaMemcpyDeviceToDevice, head_wst));
  TNT_TRY(gemm_call(w.dz[0], w.ldz[0], a.head_w, d, dX, lddx, N, d, nh, dt, dt, TNT_NN, EPI_NONE, nullptr, nullptr, 0, nullptr, st));
  // ---- the tails' rows join dX (in cluster order: the sums do not depend on the streams' timing)
  for (int i = 0; i < nt; ++i) {
    if (par) TNT_CUDA(cudaStreamWaitEvent(st, sd->done[i], 0));
    TNT_TRY(scatter_add_rows(w.dxc[i + 1], d, f.tail_rows + static_cast<long>(i) * N, f.counts + i, N, dX, lddx, d, dt, 0, 0, st));
  }
  return TNT_OK;
}

}  // namespace tnt

// ---------------------------------------------------------------------------- Generator + CrossEntropy backward (final*)
namespace tnt {
namespace {
__global__ void __launch_bounds__(256) ce_rows_kernel(const float4* stats, const int* target, int M, int pad, float* row_lse,
                                                      int* row_tgt) {
  const int r = blockIdx.x * 256 + threadIdx.x;
  if (r >= M) return;
  const int t = target[r];
  if (t == pad || t < 0) {                               // CrossEntropyLoss(ignore_index = pad)
    row_lse[r] = INFINITY;
    row_tgt[r] = -1;
    return;
  }
  row_lse[r] = stats[r].x + __logf(stats[r].y);
  row_tgt[r] = t;
}
}  // namespace
}  // namespace tnt

extern "C" size_t tnt_ce_bwd_workspace_bytes(int M, int n_cols, int dtype) {
  const size_t ldz = (static_cast<size_t>(n_cols) + 7) & ~size_t(7);
  return tnt::al256(sizeof(float) * M) + tnt::al256(sizeof(int) * M) + tnt::al256(tnt::dtype_size(dtype) * M * ldz) + 256;
}

// Backward of sum-CrossEntropy over z = x W^T + b (Generator.proj + CrossEntropyLoss(ignore_index=pad, reduction='sum'),
// final_roberta2/transformer_v2.py:54-62, custom_v4.py:404-419, 489): dz = gscale (softmax(z) - onehot(target)) is produced
// by re-running the logits GEMM with the softmax-gradient epilogue from the row statistics tnt_lse_fwd left in `stats`
// (the logits are never stored in fp32); dx = dz W, dW = dz^T x (fp32), db = column sums of dz (fp32).
extern "C" int tnt_ce_bwd(const void* x, int64_t ldx, const void* w, int64_t ldw, const float* bias, const int* target,
                          const float* stats, int M, int n_cols, int K, int pad, float gscale, void* dx, int64_t lddx, float* dW,
                          float* db, int dtype, void* workspace, size_t workspace_bytes, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(x && w && target && stats && dx && dW && workspace && M >= 0 && n_cols > 0 && K > 0, "ce_bwd: bad arguments");
  TNT_CHECK_ARG(workspace_bytes >= tnt_ce_bwd_workspace_bytes(M, n_cols, dtype), "ce_bwd: workspace too small");
  if (M == 0) return TNT_OK;
  cudaStream_t st = as_stream(stream);
  const long ldz = (static_cast<long>(n_cols) + 7) & ~7L;
  char* base = static_cast<char*>(workspace);
  float* row_lse = reinterpret_cast<float*>(base);
  int* row_tgt = reinterpret_cast<int*>(base + al256(sizeof(float) * M));
  void* dz = base + al256(sizeof(float) * M) + al256(sizeof(int) * M);
  ce_rows_kernel<<<(M + 255) / 256, 256, 0, st>>>(reinterpret_cast<const float4*>(stats), target, M, pad, row_lse, row_tgt);
  TNT_LAUNCHED();
  LseExtra ex = {row_tgt, nullptr, 0, nullptr, 0, row_lse, gscale};
  tnt_linear_params lp = {};
  lp.a = x; lp.lda = ldx; lp.w = w; lp.ldw = ldw; lp.bias = bias; lp.c = dz; lp.ldc = ldz;
  lp.M = M; lp.N = n_cols; lp.K = K; lp.groups = 1; lp.dtype_in = dtype; lp.dtype_out = dtype; lp.epilogue = EPI_NONE;
  TNT_TRY(launch_gemm(&lp, EPI_SMGRAD, &ex, nullptr, st));
  TNT_TRY(gemm_call(dz, ldz, w, ldw, dx, lddx, M, K, n_cols, dtype, dtype, TNT_NN, EPI_NONE, nullptr, nullptr, 0, nullptr, st));
  TNT_TRY(gemm_call(dz, ldz, x, ldx, dW, K, n_cols, K, M, dtype, TNT_F32, TNT_TN, EPI_NONE, nullptr, nullptr, 0, nullptr, st));
  if (db) TNT_TRY(colsum(dz, ldz, M, n_cols, nullptr, db, dtype, false, false, st));
  return TNT_OK;
}

extern "C" size_t tnt_lse_workspace_bytes(int M, int n_cols, int dtype) {
  (void)dtype;
  return tnt::lse_partials_bytes(M, n_cols);
}

extern "C" int tnt_lse_fwd(const tnt_lse_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p && p->x && p->w && p->stats && p->workspace, "lse: null pointer");
  TNT_CHECK_ARG(p->workspace_bytes >= lse_partials_bytes(p->M, p->n_cols), "lse: workspace too small");
  return lse_gemm(p->x, p->ldx, p->w, p->ldw, p->bias, p->M, p->n_cols, p->K, p->dtype, p->m_dyn, p->target, p->row_map,
                  p->n_cols, nullptr, 0, static_cast<float4*>(p->workspace), reinterpret_cast<float4*>(p->stats),
                  as_stream(stream));
}

// ------------------------------------------------------------------------------ tnt_lse_fwd without a caller workspace
// The per-(row, column-tile) partial statistics live in a file-scope device array instead of a caller buffer, so the call
// allocates nothing and needs no workspace query: for callers that cannot allocate (CUDA graph capture, serving loops).
// Every GPU that loads the library holds the array until exit.  All calls share it: two calls that overlap on the GPU
// (different streams) race.  Nothing in it has to start at zero -- the GEMM epilogue writes every entry the finish kernel
// reads -- so the last call's contents are simply overwritten.
namespace {
constexpr size_t kLseStaticBytes = size_t(8) << 20;       // 8 MiB: M * ceil(n_cols / 32) <= 524288 partial records
static __device__ __align__(256) unsigned char g_lse_static_scratch[kLseStaticBytes];
}  // namespace

extern "C" size_t tnt_lse_fwd_static_max_bytes(void) { return kLseStaticBytes; }

extern "C" int tnt_lse_fwd_static(const tnt_lse_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p && p->x && p->w && p->stats, "lse_static: null pointer");
  TNT_CHECK_ARG(p->M >= 0 && p->n_cols > 0 && p->K > 0, "lse_static: bad shape");
  const size_t need = lse_partials_bytes(p->M, p->n_cols);
  TNT_CHECK_ARG(need <= kLseStaticBytes,
                "lse_static: M=%d x n_cols=%d needs %zu bytes of scratch, the built-in array holds %zu (M * ceil(n_cols / 32) <= %zu); "
                "use tnt_lse_fwd with a caller workspace",
                p->M, p->n_cols, need, kLseStaticBytes, kLseStaticBytes / sizeof(float4));
  // in host code the array's name is a host-side stand-in: the device address comes from the runtime, once per device,
  // with that device current
  static float4* base[kMaxDevices] = {};
  const int dev = cur_device();
  if (!base[dev]) {
    void* sp = nullptr;
    TNT_CUDA(cudaGetSymbolAddress(&sp, g_lse_static_scratch));
    TNT_CHECK_ARG(sp, "lse_static: no device address for the scratch array");
    base[dev] = static_cast<float4*>(sp);
  }
  if (p->M == 0) return TNT_OK;
  return lse_gemm(p->x, p->ldx, p->w, p->ldw, p->bias, p->M, p->n_cols, p->K, p->dtype, p->m_dyn, p->target, p->row_map,
                  p->n_cols, nullptr, 0, base[dev], reinterpret_cast<float4*>(p->stats), as_stream(stream));
}
