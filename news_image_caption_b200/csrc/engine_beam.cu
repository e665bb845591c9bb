// Decoder engine: beam-search caption generation on top of the incremental step
// (DynamicConv ring buffers + cached K/V).  Scoring rule, candidate order, finished-beam bookkeeping:
//   models/CaptionModel.py:23-122 (beam_search / beam_step), state fork = reorder_incremental_state
//   (ttl/tell/modules/convolutions/dynamic.py:338-348: index_select(1, new_order) of the conv input buffers).
// Mechanism here: the W beams of caption b are rows b*W .. b*W+W-1 of the step batch; they SHARE the caption's
// cached K/V and padding masks (tnt_xattn_params.kv_group), so only the conv ring buffers are forked per step.
// All selection runs on the device (no host round trip per step); steps 1.. replay one CUDA graph.
#include "engine.cuh"

namespace tnt {
namespace {

constexpr int kMaxBeam = 16;
constexpr int kMaxLen = 128;
inline size_t al256(size_t x) { return (x + 255) & ~size_t(255); }

struct BeamWs {
  FwdWs f;
  void* ring[TNT_MAX_LAYERS];
  void* ring_tmp[TNT_MAX_LAYERS];
  int* pos;
  int64_t* cur;                 // [R] token fed to the next step
  float* lp;                    // [R, V] full log-probs of the step
  float* topv; int* topi;       // [R, W] per-row top-W (descending)
  int* parent;                  // [R] source row of each new beam
  int64_t* bseq; float* blogp; float* bsum;     // live beams   [B, W, L], [B, W, L], [B, W]
  int64_t* dseq; float* dlogp; float* dp; int* ndone;   // finished beams, kept sorted by p (descending)
  void* ada; size_t ada_bytes;
  size_t total;
};

BeamWs carve_beam(const tnt_decoder* D, int B, int W, int L, const int* S, void* ws) {
  BeamWs g;
  const int R = B * W;
  g.f = carve_fwd(D, R, 1, S, ws, true, B);
  size_t off = g.f.total;
  auto take = [&](size_t bytes) {
    void* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += al256(bytes);
    return p;
  };
  for (int l = 0; l < D->cfg.n_layers; ++l) {
    const size_t rb = static_cast<size_t>(std::max(D->cfg.kernels[l] - 1, 1)) * R * D->cfg.d * D->esz();
    g.ring[l] = take(rb);
    g.ring_tmp[l] = take(rb);
  }
  g.pos = static_cast<int*>(take(2 * sizeof(int)));
  g.cur = static_cast<int64_t*>(take(sizeof(int64_t) * R));
  g.lp = static_cast<float*>(take(sizeof(float) * R * D->cfg.vocab));
  g.topv = static_cast<float*>(take(sizeof(float) * R * W));
  g.topi = static_cast<int*>(take(sizeof(int) * R * W));
  g.parent = static_cast<int*>(take(sizeof(int) * R));
  g.bseq = static_cast<int64_t*>(take(sizeof(int64_t) * R * L));
  g.blogp = static_cast<float*>(take(sizeof(float) * R * L));
  g.bsum = static_cast<float*>(take(sizeof(float) * R));
  g.dseq = static_cast<int64_t*>(take(sizeof(int64_t) * R * L));
  g.dlogp = static_cast<float*>(take(sizeof(float) * R * L));
  g.dp = static_cast<float*>(take(sizeof(float) * R));
  g.ndone = static_cast<int*>(take(sizeof(int) * B));
  g.ada_bytes = adasoftmax_workspace_bytes(D->ada, R, true);
  g.ada = take(g.ada_bytes);
  g.total = off;
  return g;
}

__global__ void beam_init_kernel(const int64_t* seed, int B, int W, int64_t* cur, float* bsum, int* ndone, int* pos) {
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < B * W; r += gridDim.x * blockDim.x) {
    cur[r] = seed[r / W];
    bsum[r] = 0.f;
    if (r % W == 0) ndone[r / W] = 0;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) pos[0] = 0;
}
__global__ void pos_inc_kernel(int* pos) { pos[0] += 1; }

// (value, index) order of torch.sort(descending): larger value first; equal values by smaller index
__device__ __forceinline__ bool before(float v, int i, float v2, int i2) { return v > v2 || (v == v2 && i < i2); }

// per-row top-W of the log-probs: W block-wide arg-max passes, each restricted to entries ordered strictly after the
// previous pick (ys, ix = torch.sort(logprobsf, 1, True) then [:, :beam] -- CaptionModel.py:41-43)
__global__ void __launch_bounds__(256) topk_rows_kernel(const float* lp, int V, int W, float* topv, int* topi) {
  __shared__ float sv[8];
  __shared__ int si[8];
  const float* row = lp + static_cast<long>(blockIdx.x) * V;
  float pv = INFINITY;
  int pi = -1;
  for (int k = 0; k < W; ++k) {
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < V; i += 256) {
      const float v = row[i];
      if (before(pv, pi, v, i) && before(v, i, bv, bi)) { bv = v; bi = i; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float v2 = __shfl_xor_sync(0xffffffffu, bv, o);
      const int i2 = __shfl_xor_sync(0xffffffffu, bi, o);
      if (before(v2, i2, bv, bi)) { bv = v2; bi = i2; }
    }
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = bv; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    bv = sv[0]; bi = si[0];
    for (int w = 1; w < 8; ++w)
      if (before(sv[w], si[w], bv, bi)) { bv = sv[w]; bi = si[w]; }
    __syncthreads();
    if (threadIdx.x == 0) {
      topv[static_cast<long>(blockIdx.x) * W + k] = bv;
      topi[static_cast<long>(blockIdx.x) * W + k] = bi;
    }
    pv = bv; pi = bi;
  }
}

// One CTA per caption: beam_step + the finished-beam bookkeeping of the caller loop (CaptionModel.py:27-76, 102-112).
//   candidates (c, q) for c < W, q < rows (rows = 1 at t = 0): p = bsum[q] + ys[q, c]; stable sort by -p
//   (creation order c-major, then q); the best W become the new beams (history and conv state forked from q);
//   a beam whose new word is `eos`, or every beam at t = L-1, is appended to the done list with score p and its
//   running sum is set to -1000 so that it is not continued.
__global__ void __launch_bounds__(128) beam_step_kernel(const float* topv, const int* topi, int W, int L, int eos,
                                                        const int* pos, int64_t* cur, int* parent, int64_t* bseq,
                                                        float* blogp, float* bsum, int64_t* dseq, float* dlogp, float* dp,
                                                        int* ndone) {
  __shared__ int64_t s_seq[kMaxBeam * kMaxLen];
  __shared__ float s_lp[kMaxBeam * kMaxLen];
  __shared__ float s_p[kMaxBeam], s_r[kMaxBeam];
  __shared__ int s_q[kMaxBeam], s_c[kMaxBeam];
  const int b = blockIdx.x, t = pos[0];
  if (t >= L) return;
  const int base = b * W;
  for (int i = threadIdx.x; i < W * t; i += blockDim.x) {      // old histories [W][t]
    const int q = i / t, k = i - q * t;
    s_seq[q * kMaxLen + k] = bseq[static_cast<long>(base + q) * L + k];
    s_lp[q * kMaxLen + k] = blogp[static_cast<long>(base + q) * L + k];
  }
  if (threadIdx.x == 0) {
    const int rows = t == 0 ? 1 : W;
    unsigned long long used[4] = {0, 0, 0, 0};                  // W*W <= 256 candidates
    for (int vix = 0; vix < W; ++vix) {
      float bp = 0.f;
      int bc = -1, bq = -1;
      for (int c = 0; c < W; ++c)
        for (int q = 0; q < rows; ++q) {
          const int id = c * W + q;
          if (used[id >> 6] >> (id & 63) & 1ull) continue;
          const float p = bsum[base + q] + topv[static_cast<long>(base + q) * W + c];
          if (bc < 0 || p > bp) { bp = p; bc = c; bq = q; }    // strict >: earlier (c, q) wins ties = stable sort
        }
      used[(bc * W + bq) >> 6] |= 1ull << ((bc * W + bq) & 63);
      s_p[vix] = bp;
      s_q[vix] = bq;
      s_c[vix] = topi[static_cast<long>(base + bq) * W + bc];
      s_r[vix] = topv[static_cast<long>(base + bq) * W + bc];
    }
  }
  __syncthreads();
  // done list: kept sorted by p descending, ties in arrival order, capacity W (the reference sorts ALL finished beams
  // at the end and keeps the best W: same set, same order).  Insertions in beam order; all threads move the rows.
  for (int vix = 0; vix < W; ++vix) {
    const bool fin = s_c[vix] == eos || t == L - 1;
    if (!fin) continue;                                          // uniform: shared memory
    int nd = ndone[b];
    int ins = nd;
    while (ins > 0 && dp[base + ins - 1] < s_p[vix]) --ins;
    __syncthreads();
    if (ins < W) {
      const int last = min(nd, W - 1);                           // entries [ins, last) move down by one
      for (int e = last; e > ins; --e) {
        for (int k = threadIdx.x; k < L; k += blockDim.x) {
          dseq[static_cast<long>(base + e) * L + k] = dseq[static_cast<long>(base + e - 1) * L + k];
          dlogp[static_cast<long>(base + e) * L + k] = dlogp[static_cast<long>(base + e - 1) * L + k];
        }
        if (threadIdx.x == 0) dp[base + e] = dp[base + e - 1];
        __syncthreads();
      }
      const int q = s_q[vix];
      for (int k = threadIdx.x; k < L; k += blockDim.x) {
        dseq[static_cast<long>(base + ins) * L + k] = k < t ? s_seq[q * kMaxLen + k] : (k == t ? static_cast<int64_t>(s_c[vix]) : 0);
        dlogp[static_cast<long>(base + ins) * L + k] = k < t ? s_lp[q * kMaxLen + k] : (k == t ? s_r[vix] : 0.f);
      }
      if (threadIdx.x == 0) {
        dp[base + ins] = s_p[vix];
        ndone[b] = min(nd + 1, W);
      }
    }
    __syncthreads();
  }
  // new live beams
  for (int i = threadIdx.x; i < W * (t + 1); i += blockDim.x) {
    const int vix = i / (t + 1), k = i - vix * (t + 1);
    const int q = s_q[vix];
    bseq[static_cast<long>(base + vix) * L + k] = k < t ? s_seq[q * kMaxLen + k] : static_cast<int64_t>(s_c[vix]);
    blogp[static_cast<long>(base + vix) * L + k] = k < t ? s_lp[q * kMaxLen + k] : s_r[vix];
  }
  if (threadIdx.x < W) {
    const int vix = threadIdx.x;
    const bool fin = s_c[vix] == eos || t == L - 1;
    bsum[base + vix] = fin ? -1000.f : s_p[vix];
    cur[base + vix] = s_c[vix];
    parent[base + vix] = base + s_q[vix];
  }
}

// conv input buffers follow their beam: tmp[s][r] = ring[s][parent[r]] (reorder_incremental_state = index_select(1, order),
// dynamic.py:338-348), for every layer in one launch; copied back by the caller.  16-byte vectors.
struct RingTable {
  const uint4* src[TNT_MAX_LAYERS];
  uint4* dst[TNT_MAX_LAYERS];
  int slots[TNT_MAX_LAYERS];
};
__global__ void __launch_bounds__(256) ring_fork_kernel(RingTable tb, const int* parent, int R, int row_vecs) {
  const int l = blockIdx.y, r = blockIdx.x;
  const int src = parent[r];
  for (int s = 0; s < tb.slots[l]; ++s) {
    const uint4* a = tb.src[l] + (static_cast<long>(s) * R + src) * row_vecs;
    uint4* o = tb.dst[l] + (static_cast<long>(s) * R + r) * row_vecs;
    for (int i = threadIdx.x; i < row_vecs; i += 256) o[i] = a[i];
  }
}

__global__ void beam_finish_kernel(const int64_t* dseq, const float* dlogp, const float* dp, int n, int L, int64_t* seq_out,
                                   float* logp_out, float* p_out) {
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < static_cast<long>(n) * L;
       i += static_cast<long>(gridDim.x) * blockDim.x) {
    seq_out[i] = dseq[i];
    logp_out[i] = dlogp[i];
    if (i % L == 0) p_out[i / L] = dp[i / L];
  }
}

int beam_one_step(tnt_decoder* D, const tnt_contexts* ctx1, const BeamWs& g, int B, int W, int L, int eos, cudaStream_t st) {
  const int R = B * W;
  TNT_TRY(embed(D, g.cur, R, 1, g.f.X0, g.pos, st));
  for (int l = 0; l < D->cfg.n_layers; ++l)
    TNT_TRY(run_layer(D, l, ctx1, g.f, g.f.X0, g.f.X0, R, 1, g.ring[l], g.pos, nullptr, false, st, W));
  TNT_TRY(adasoftmax_logprob(D->ada, g.f.X0, D->cfg.d, R, g.lp, g.ada, g.ada_bytes, st));
  topk_rows_kernel<<<R, 256, 0, st>>>(g.lp, D->cfg.vocab, W, g.topv, g.topi);
  TNT_LAUNCHED();
  beam_step_kernel<<<B, 128, 0, st>>>(g.topv, g.topi, W, L, eos, g.pos, g.cur, g.parent, g.bseq, g.blogp, g.bsum, g.dseq,
                                      g.dlogp, g.dp, g.ndone);
  TNT_LAUNCHED();
  RingTable tb = {};
  const int row_vecs = static_cast<int>(D->cfg.d * D->esz() / 16);
  for (int l = 0; l < D->cfg.n_layers; ++l) {
    tb.src[l] = static_cast<const uint4*>(g.ring[l]);
    tb.dst[l] = static_cast<uint4*>(g.ring_tmp[l]);
    tb.slots[l] = D->cfg.kernels[l] - 1;
  }
  ring_fork_kernel<<<dim3(R, D->cfg.n_layers), 256, 0, st>>>(tb, g.parent, R, row_vecs);
  TNT_LAUNCHED();
  for (int l = 0; l < D->cfg.n_layers; ++l) {
    const size_t rb = static_cast<size_t>(D->cfg.kernels[l] - 1) * R * D->cfg.d * D->esz();
    if (rb) TNT_CUDA(cudaMemcpyAsync(g.ring[l], g.ring_tmp[l], rb, cudaMemcpyDeviceToDevice, st));
  }
  pos_inc_kernel<<<1, 1, 0, st>>>(g.pos);
  TNT_LAUNCHED();
  return TNT_OK;
}

}  // namespace
}  // namespace tnt

using namespace tnt;

extern "C" size_t tnt_decoder_beam_workspace_bytes(const tnt_decoder* d, int B, int beam, int gen_len, const int* S) {
  if (!d || !S || beam < 1 || beam > kMaxBeam || gen_len < 1 || gen_len > kMaxLen) return 0;
  return carve_beam(d, B, beam, gen_len, S, nullptr).total;
}

extern "C" int tnt_decoder_generate_beam(tnt_decoder* D, const int64_t* seed, const tnt_contexts* ctx, int gen_len, int eos,
                                         int beam, int64_t* seq_out, float* logp_out, float* p_out, void* workspace,
                                         size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_generate_beam: decoder not finalized");
  TNT_CHECK_ARG(seed && ctx && seq_out && logp_out && p_out && workspace, "decoder_generate_beam: null pointer");
  TNT_CHECK_ARG(beam >= 1 && beam <= kMaxBeam, "decoder_generate_beam: beam size %d out of [1, %d]", beam, kMaxBeam);
  TNT_CHECK_ARG(gen_len >= 1 && gen_len <= kMaxLen, "decoder_generate_beam: gen_len %d out of [1, %d]", gen_len, kMaxLen);
  TNT_CHECK_ARG(beam <= D->cfg.vocab, "decoder_generate_beam: beam exceeds the vocabulary");
  TNT_CHECK_ARG(gen_len + D->cfg.pad_idx + 1 <= D->cfg.max_positions, "decoder_generate_beam: gen_len exceeds the position table");
  const int B = ctx->B, W = beam, L = gen_len;
  TNT_CHECK_ARG(B > 0, "decoder_generate_beam: empty batch");
  BeamWs g = carve_beam(D, B, W, L, ctx->S, workspace);
  if (g.total > workspace_bytes) {
    set_error("decoder_generate_beam: workspace %zu < %zu", workspace_bytes, g.total);
    return TNT_ERR_WORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  tnt_contexts c1 = *ctx;
  c1.T = 1;
  beam_init_kernel<<<std::min((B * W + 255) / 256, 1024), 256, 0, st>>>(seed, B, W, g.cur, g.bsum, g.ndone, g.pos);
  TNT_LAUNCHED();
  TNT_TRY(project_kv(D, &c1, g.f, st, true));
  TNT_TRY(beam_one_step(D, &c1, g, B, W, L, eos, st));                     // step 0 eagerly
  if (L > 1) {
    cudaStream_t cs;
    TNT_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    int rc = TNT_OK;
    cudaError_t e = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
    const uint64_t n0 = tnt_launch_count();
    if (e == cudaSuccess) {
      rc = beam_one_step(D, &c1, g, B, W, L, eos, cs);
      e = cudaStreamEndCapture(cs, &graph);
    }
    const int per_step = static_cast<int>(tnt_launch_count() - n0);
    if (rc == TNT_OK && e == cudaSuccess) e = cudaGraphInstantiate(&exec, graph, 0);
    if (rc == TNT_OK && e == cudaSuccess) {
      for (int i = 1; i < L && e == cudaSuccess; ++i) e = cudaGraphLaunch(exec, st);
      count_launch((L - 2) * per_step);
    }
    if (exec) cudaGraphExecDestroy(exec);
    if (graph) cudaGraphDestroy(graph);
    cudaStreamDestroy(cs);
    if (rc != TNT_OK) return rc;
    if (e != cudaSuccess) {
      set_error("decoder_generate_beam: CUDA graph replay failed: %s", cudaGetErrorString(e));
      return TNT_ERR_CUDA;
    }
  }
  beam_finish_kernel<<<std::min((B * W * L + 255) / 256, 2048), 256, 0, st>>>(g.dseq, g.dlogp, g.dp, B * W, L, seq_out, logp_out,
                                                                              p_out);
  TNT_LAUNCHED();
  return TNT_OK;
}
