// Decoder engine: greedy generation with incremental DynamicConv ring buffers + cached K/V (kernel
// group 5), and the host-buffer scoring entry used for the end-to-end benchmark.
//   _generate (sampling_topk=1)  ttl/tell/models/transformer_flattened.py:238-325,
//                                ttl/tell/models/transformer_faces_objects.py:399-494
//   incremental conv buffer      ttl/tell/modules/convolutions/dynamic.py:95-99,114-116,338-348
//   incremental positions        ttl/tell/modules/token_embedders/positional.py:170-174,198-199
// Differences in mechanism, not in numbers: finished rows stay in the batch (masked) instead of being
// compacted away every step, K/V projections are computed once instead of once per step and layer,
// and steps 1.. are replayed from one captured CUDA graph (the step counter lives on the device).
#include "engine.cuh"

namespace tnt {

namespace {
inline size_t al256(size_t x) { return (x + 255) & ~size_t(255); }

struct GenWs {
  FwdWs f;
  void* ring[TNT_MAX_LAYERS];
  int* pos; int* done_step; uint8_t* active; int64_t* cur; int64_t* tok; float* logp; void* ada;
  size_t ada_bytes, total;    // `ada` holds tnt_decoder::kMaxSub slots of ada_bytes (one per sub-batch)
};

GenWs carve_gen(const tnt_decoder* D, int B, const int* S, void* ws) {
  GenWs g;
  g.f = carve_fwd(D, B, 1, S, ws, true);
  size_t off = g.f.total;
  auto take = [&](size_t bytes) {
    void* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += al256(bytes);
    return p;
  };
  for (int l = 0; l < D->cfg.n_layers; ++l)
    g.ring[l] = take(static_cast<size_t>(std::max(D->cfg.kernels[l] - 1, 1)) * B * D->cfg.d * D->esz());
  g.pos = static_cast<int*>(take(2 * sizeof(int)));
  g.done_step = g.pos + 1;
  g.active = static_cast<uint8_t*>(take(B));
  g.cur = static_cast<int64_t*>(take(sizeof(int64_t) * B));
  g.tok = static_cast<int64_t*>(take(sizeof(int64_t) * B));
  g.logp = static_cast<float*>(take(sizeof(float) * B));
  g.ada_bytes = al256(adasoftmax_workspace_bytes(D->ada, B, false));
  g.ada = take(g.ada_bytes * tnt_decoder::kMaxSub);
  g.total = off;
  return g;
}

__global__ void gen_init_kernel(const int64_t* seed, int B, int L, int eos, int64_t* cur, uint8_t* active, int64_t* ids_out,
                                int* pos) {
  for (int b = threadIdx.x; b < B; b += blockDim.x) {
    cur[b] = seed[b];
    active[b] = seed[b] != eos;
    ids_out[static_cast<long>(b) * (L + 1)] = seed[b];
  }
  if (threadIdx.x == 0) {
    pos[0] = 0;
    pos[1] = -1;
  }
}

// scatter this step's tokens / log-probs, retire rows that produced EOS (transformer_flattened.py:290-313)
__global__ void __launch_bounds__(1024) gen_update_kernel(const int64_t* tok, const float* logp, int B, int L, int eos,
                                                          int pad, int64_t* cur, uint8_t* active, int64_t* ids_out,
                                                          float* logp_out, int* pos) {
  __shared__ int any_active;
  const int step = pos[0];
  if (threadIdx.x == 0) any_active = 0;
  __syncthreads();
  if (step < L) {
    for (int b = threadIdx.x; b < B; b += blockDim.x) {
      const int64_t t = tok[b];
      if (active[b]) {
        ids_out[static_cast<long>(b) * (L + 1) + step + 1] = t;
        logp_out[static_cast<long>(b) * L + step] = logp[b];
        if (t == eos) active[b] = 0;
        else any_active = 1;
      } else {
        ids_out[static_cast<long>(b) * (L + 1) + step + 1] = pad;
        logp_out[static_cast<long>(b) * L + step] = 0.f;
      }
      cur[b] = t;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    pos[0] = step + 1;
    if (!any_active && pos[1] < 0) pos[1] = step + 1;
  }
}

// one decode step; the captions are cut into sub-batches (engine.cuh) whose chains run on their own streams: while one
// sub-batch streams its K/V cache (HBM-bound), the others run their ~60 latency-bound small launches
int gen_step(tnt_decoder* D, const tnt_contexts* ctx1, const GenWs& g, int B, int L, int eos, int64_t* ids_out,
             float* logp_out, const SubBatches& sb0, cudaStream_t st) {
  SubBatches sb = sb0;
  sb.st[0] = st;                                   // the stream being issued on / captured
  const size_t E = D->esz(), d = D->cfg.d;
  TNT_TRY(sub_fork(D, sb));
  static const bool share_sms = [] { const char* e = getenv("TNT_SUB_SHARE_SMS"); return !(e && e[0] == '0'); }();
  SmBudget budget(sb.n > 1 && share_sms ? num_sms() / sb.n : 0);
  for (int i = 0; i < sb.n; ++i) {
    const FwdWs f = sb.n > 1 ? sub_ws(D, g.f, sb, i, B, 1, ctx1->S) : g.f;
    const tnt_contexts ci = sb.n > 1 ? sub_ctx(*ctx1, sb, i) : *ctx1;
    const int b0 = sb.b0[i], nb = ci.B;
    cudaStream_t si = sb.st[i];
    tnt_embed_params e = {};
    e.ids = g.cur + b0; e.table = D->table; e.pos_table = D->pos_table; e.out = f.X0; e.ldo = D->cfg.d; e.start_pos = g.pos;
    e.B = nb; e.T = 1; e.d = D->cfg.d; e.n_pos = D->cfg.max_positions; e.pad_idx = D->cfg.pad_idx; e.dtype = D->cfg.dtype;
    e.vocab = D->cfg.vocab; e.error_flags = D->err_flags;
    TNT_TRY(tnt_embed_fwd(&e, si));
    for (int l = 0; l < D->cfg.n_layers; ++l) {
      // ring [K-1][nb][d] of this sub-batch: a private block of the layer's ring memory
      void* ring = static_cast<char*>(g.ring[l]) + static_cast<size_t>(std::max(D->cfg.kernels[l] - 1, 1)) * b0 * d * E;
      TNT_TRY(run_layer(D, l, &ci, f, f.X0, f.X0, nb, 1, ring, g.pos, nullptr, false, si));
    }
    TNT_TRY(adasoftmax_greedy(D->ada, f.X0, D->cfg.d, nb, g.tok + b0, g.logp + b0, static_cast<char*>(g.ada) + i * g.ada_bytes,
                              g.ada_bytes, si));
  }
  TNT_TRY(sub_join(D, sb));
  gen_update_kernel<<<1, 1024, 0, st>>>(g.tok, g.logp, B, L, eos, D->cfg.pad_idx, g.cur, g.active, ids_out, logp_out, g.pos);
  TNT_LAUNCHED();
  return TNT_OK;
}

// ---- generation through the copy head (pointer / entity-pointer models)
struct PtrWs {
  GenWs g;
  void* cat;       // [B, 2d]  cat(inp_embed, X) of the step: decoder_pointer.py:139 / decoder_entity_pointer.py:144
  void* hid;       // [B, d_hidden] relu(project_first_p_gens(cat))  (entity-pointer head)
  float* attn;     // [B, S_w + 2] head-averaged attention of the last layer over the copied source
  float* lp;       // [B, V] full-vocabulary log-probabilities, mixed in place
  void* ada; size_t ada_bytes, total;
};

PtrWs carve_ptr(const tnt_decoder* D, int B, const int* S, int d_hidden, void* ws) {
  PtrWs w;
  w.g = carve_gen(D, B, S, ws);
  size_t off = w.g.total;
  auto take = [&](size_t bytes) {
    void* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += al256(bytes);
    return p;
  };
  const int sw = D->article_idx >= 0 ? S[D->article_idx] : 0;
  w.cat = take(static_cast<size_t>(B) * 2 * D->cfg.d * D->esz());
  w.hid = take(static_cast<size_t>(B) * std::max(d_hidden, 8) * D->esz());
  w.attn = static_cast<float*>(take(sizeof(float) * B * (sw + 2)));
  w.lp = static_cast<float*>(take(sizeof(float) * static_cast<size_t>(B) * D->cfg.vocab));
  w.ada_bytes = adasoftmax_workspace_bytes(D->ada, B, true);
  w.ada = take(w.ada_bytes);
  w.g.f.step_scores = static_cast<float*>(take(sizeof(float) * B * D->cfg.heads * (sw + 2)));
  w.g.f.step_lse = static_cast<float*>(take(sizeof(float) * B * D->cfg.heads));
  w.total = off;
  return w;
}

int gen_step_pointer(tnt_decoder* D, const tnt_contexts* ctx1, const PtrWs& w, const tnt_pointer_head* hd, int B, int L,
                     int eos, int64_t* ids_out, float* logp_out, cudaStream_t st) {
  const GenWs& g = w.g;
  const int d = D->cfg.d, dt = D->cfg.dtype;
  const size_t E = D->esz();
  tnt_embed_params e = {};
  e.ids = g.cur; e.table = D->table; e.pos_table = D->pos_table; e.out = g.f.X0; e.ldo = d; e.start_pos = g.pos;
  e.B = B; e.T = 1; e.d = d; e.n_pos = D->cfg.max_positions; e.pad_idx = D->cfg.pad_idx; e.dtype = dt;
  e.vocab = D->cfg.vocab; e.error_flags = D->err_flags;
  TNT_TRY(tnt_embed_fwd(&e, st));
  // inp_embed is kept as the left half of `cat` (the layers overwrite X0)
  TNT_CUDA(cudaMemcpy2DAsync(w.cat, 2 * d * E, g.f.X0, d * E, d * E, B, cudaMemcpyDeviceToDevice, st));
  const int Ls = D->cfg.n_layers;
  for (int l = 0; l < Ls; ++l)
    TNT_TRY(run_layer(D, l, ctx1, g.f, g.f.X0, g.f.X0, B, 1, g.ring[l], g.pos, l == Ls - 1 ? w.attn : nullptr, false, st));
  TNT_TRY(adasoftmax_logprob(D->ada, g.f.X0, d, B, w.lp, w.ada, w.ada_bytes, st));
  tnt_pointer_params pp = {};
  pp.w_pgen = hd->w_pgen; pp.b_pgen = hd->b_pgen;
  if (hd->w_hidden) {
    TNT_CUDA(cudaMemcpy2DAsync(static_cast<char*>(w.cat) + d * E, 2 * d * E, g.f.X0, d * E, d * E, B, cudaMemcpyDeviceToDevice, st));
    tnt_linear_params lp = {};
    lp.a = w.cat; lp.lda = 2 * d; lp.w = hd->w_hidden; lp.ldw = 2 * d; lp.bias = hd->b_hidden; lp.c = w.hid; lp.ldc = hd->d_hidden;
    lp.M = B; lp.N = hd->d_hidden; lp.K = 2 * d; lp.groups = 1; lp.dtype_in = dt; lp.dtype_out = dt; lp.epilogue = TNT_EPI_RELU;
    lp.alpha = 1.f;
    TNT_TRY(tnt_linear_fwd(&lp, st));
    pp.x = w.hid; pp.ldx = hd->d_hidden; pp.inp_embed = nullptr; pp.d = hd->d_hidden;
  } else {
    pp.x = g.f.X0; pp.ldx = d; pp.inp_embed = w.cat; pp.lde = 2 * d; pp.d = d;
  }
  pp.attn = w.attn; pp.attn_ld = hd->S_copy + 2; pp.article_ids = hd->copy_ids;
  pp.logprob = w.lp; pp.ldp = D->cfg.vocab;
  pp.B = B; pp.T = 1; pp.S = hd->S_copy; pp.V = D->cfg.vocab; pp.dtype = dt;
  pp.best_tok = g.tok; pp.best_logp = g.logp;
  TNT_TRY(tnt_pointer_mix(&pp, st));
  gen_update_kernel<<<1, 1024, 0, st>>>(g.tok, g.logp, B, L, eos, D->cfg.pad_idx, g.cur, g.active, ids_out, logp_out, g.pos);
  TNT_LAUNCHED();
  return TNT_OK;
}

// step 0 eagerly, steps 1.. replayed from ONE captured CUDA graph (the step counter lives on the device)
template <typename Step>
int run_steps(int gen_len, cudaStream_t st, Step&& step) {
  TNT_TRY(step(st));
  if (gen_len <= 1) return TNT_OK;
  // capture on a private stream (the caller's may be the legacy stream, which cannot be captured)
  cudaStream_t cs;
  TNT_CUDA(cudaStreamCreateWithFlags(&cs, cudaStreamNonBlocking));
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int rc = TNT_OK;
  cudaError_t e = cudaStreamBeginCapture(cs, cudaStreamCaptureModeThreadLocal);
  const uint64_t n0 = tnt_launch_count();
  if (e == cudaSuccess) {
    rc = step(cs);
    e = cudaStreamEndCapture(cs, &graph);
  }
  const int per_step = static_cast<int>(tnt_launch_count() - n0);   // kernels recorded in the graph
  if (rc == TNT_OK && e == cudaSuccess) e = cudaGraphInstantiate(&exec, graph, 0);
  if (rc == TNT_OK && e == cudaSuccess) {
    for (int i = 1; i < gen_len && e == cudaSuccess; ++i) e = cudaGraphLaunch(exec, st);
    count_launch((gen_len - 2) * per_step);   // capture counted one step; gen_len-1 replays executed
  }
  if (exec) cudaGraphExecDestroy(exec);
  if (graph) cudaGraphDestroy(graph);
  cudaStreamDestroy(cs);
  if (rc != TNT_OK) return rc;
  if (e != cudaSuccess) {
    set_error("decoder_generate: CUDA graph replay failed: %s", cudaGetErrorString(e));
    return TNT_ERR_CUDA;
  }
  return TNT_OK;
}
}  // namespace
}  // namespace tnt

using namespace tnt;

extern "C" size_t tnt_decoder_generate_pointer_workspace_bytes(const tnt_decoder* d, int B, int gen_len, const int* S,
                                                               int d_hidden) {
  (void)gen_len;
  if (!d || !S) return 0;
  return carve_ptr(d, B, S, d_hidden, nullptr).total;
}

extern "C" int tnt_decoder_generate_pointer(tnt_decoder* D, const int64_t* seed, const tnt_contexts* ctx,
                                            const tnt_pointer_head* head, int gen_len, int eos, int64_t* ids_out,
                                            float* logp_out, int* steps_host, void* workspace, size_t workspace_bytes,
                                            tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_generate_pointer: decoder not finalized");
  TNT_CHECK_ARG(seed && ctx && head && ids_out && logp_out && workspace && gen_len > 0, "decoder_generate_pointer: bad arguments");
  TNT_CHECK_ARG(head->w_pgen && head->copy_ids, "decoder_generate_pointer: head weights / copy ids missing");
  TNT_CHECK_ARG(D->article_idx >= 0, "decoder_generate_pointer: the decoder has no article / entity source to copy from");
  TNT_CHECK_ARG(head->S_copy == ctx->S[D->article_idx], "decoder_generate_pointer: copy_ids cover %d positions, source '%s' has %d",
                head->S_copy, D->cfg.source_names[D->article_idx], ctx->S[D->article_idx]);
  TNT_CHECK_ARG(!head->w_hidden || (head->b_hidden && head->d_hidden > 0 && head->d_hidden % 8 == 0),
                "decoder_generate_pointer: hidden layer of the entity-pointer head");
  const int B = ctx->B;
  TNT_CHECK_ARG(gen_len + D->cfg.pad_idx + 1 <= D->cfg.max_positions, "decoder_generate_pointer: gen_len exceeds the position table");
  PtrWs w = carve_ptr(D, B, ctx->S, head->w_hidden ? head->d_hidden : 0, workspace);
  if (w.total > workspace_bytes) {
    set_error("decoder_generate_pointer: workspace %zu < %zu", workspace_bytes, w.total);
    return TNT_ERR_WORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  tnt_contexts c1 = *ctx;
  c1.T = 1;
  gen_init_kernel<<<1, 1024, 0, st>>>(seed, B, gen_len, eos, w.g.cur, w.g.active, ids_out, w.g.pos);
  TNT_LAUNCHED();
  TNT_TRY(project_kv(D, &c1, w.g.f, st, true));
  TNT_TRY(run_steps(gen_len, st, [&](cudaStream_t s) {
    return gen_step_pointer(D, &c1, w, head, B, gen_len, eos, ids_out, logp_out, s);
  }));
  if (steps_host) {
    int two[2];
    TNT_CUDA(cudaMemcpyAsync(two, w.g.pos, sizeof(two), cudaMemcpyDeviceToHost, st));
    TNT_CUDA(cudaStreamSynchronize(st));
    *steps_host = two[1] < 0 ? gen_len : two[1];
  }
  return TNT_OK;
}

extern "C" size_t tnt_decoder_generate_workspace_bytes(const tnt_decoder* d, int B, int gen_len, const int* S) {
  (void)gen_len;
  if (!d || !S) return 0;
  return carve_gen(d, B, S, nullptr).total;
}

extern "C" int tnt_decoder_generate(tnt_decoder* D, const int64_t* seed, const tnt_contexts* ctx, int gen_len, int eos,
                                    int64_t* ids_out, float* logp_out, int* steps_host, void* workspace,
                                    size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_generate: decoder not finalized");
  TNT_CHECK_ARG(seed && ctx && ids_out && logp_out && workspace && gen_len > 0, "decoder_generate: bad arguments");
  const int B = ctx->B;
  TNT_CHECK_ARG(gen_len + D->cfg.pad_idx + 1 <= D->cfg.max_positions, "decoder_generate: gen_len exceeds the position table");
  GenWs g = carve_gen(D, B, ctx->S, workspace);
  if (g.total > workspace_bytes) {
    set_error("decoder_generate: workspace %zu < %zu", workspace_bytes, g.total);
    return TNT_ERR_WORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  tnt_contexts c1 = *ctx;
  c1.T = 1;
  gen_init_kernel<<<1, 1024, 0, st>>>(seed, B, gen_len, eos, g.cur, g.active, ids_out, g.pos);
  TNT_LAUNCHED();
  SubBatches sb;
  TNT_TRY(sub_plan(D, B, 1, 32, st, &sb));
  TNT_TRY(project_kv(D, &c1, g.f, st, true, &sb));
  TNT_TRY(run_steps(gen_len, st, [&](cudaStream_t s) { return gen_step(D, &c1, g, B, gen_len, eos, ids_out, logp_out, sb, s); }));
  if (steps_host) {
    int two[2];
    TNT_CUDA(cudaMemcpyAsync(two, g.pos, sizeof(two), cudaMemcpyDeviceToHost, st));
    TNT_CUDA(cudaStreamSynchronize(st));
    *steps_host = two[1] < 0 ? gen_len : two[1];
  }
  return TNT_OK;
}

// ------------------------------------------------------------------------------ host-buffer scoring
namespace tnt {
namespace {
struct Staging {
  int64_t* ids; int64_t* tgt; void* mem[TNT_MAX_SOURCES]; uint8_t* mask[TNT_MAX_SOURCES]; void* X; float* out2;
  size_t total;
};
Staging carve_staging(const tnt_decoder* D, int B, int T, const int* S, void* base) {
  Staging s;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = base ? static_cast<char*>(base) + off : nullptr;
    off += al256(bytes);
    return p;
  };
  s.ids = static_cast<int64_t*>(take(sizeof(int64_t) * B * T));
  s.tgt = static_cast<int64_t*>(take(sizeof(int64_t) * B * T));
  for (int i = 0; i < D->cfg.n_sources; ++i) {
    s.mem[i] = take(static_cast<size_t>(S[i]) * B * D->cfg.source_dim[i] * D->esz());
    s.mask[i] = static_cast<uint8_t*>(take(static_cast<size_t>(S[i]) * B));
  }
  s.X = take(static_cast<size_t>(B) * T * D->cfg.d * D->esz());
  s.out2 = static_cast<float*>(take(2 * sizeof(float)));
  s.total = off;
  return s;
}

// Host -> device copy of one memory source [S, B, E] (pinned host memory).  With a key-padding mask on the host only the
// LIVE rows travel: caption b's rows s < len_b (len_b = 1 + its last un-masked position) are one strided 2-D copy
// (width = one E-element row, height = len_b, pitch = B rows) -- the padded tail of an article (mean live length ~ 320 of
// 512) is never read by any kernel (a masked key contributes exactly 0), so it need not cross PCIe.  The staging rows that
// are skipped keep stale but FINITE values: the buffer is zeroed the first time this decoder sees it.
// Live-row host -> device copy as ONE kernel that reads the pinned host tensor over PCIe itself (zero copy): CTA = caption b,
// len_b found from the caption's key-padding bytes (also read from the host), rows s < len_b fetched with 16-byte loads (a
// warp per row, 4 chunks per lane in flight) and stored to the device layout.  The DMA-engine form is one strided 2-D copy
// per caption and source: 132 copies per step at B = 64, each with ~3.5 us of fixed cost -- a quarter of the copy engine's
// time (tools/trace_e2e.py), which made the host-buffer path copy-bound on boxes with a slow PCIe link.
__global__ void __launch_bounds__(256) h2d_live_rows_kernel(uint4* dst, const uint4* src, const uint8_t* mask, int S, int B,
                                                            int row16 /* 16-byte chunks per row */) {
  __shared__ int s_len;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    if (threadIdx.x == 0) s_len = 0;
    __syncthreads();
    int last = 0;                                    // 1 + the last un-masked position this thread saw
    for (int i = threadIdx.x; i < S; i += 256)
      if (mask[static_cast<size_t>(b) * S + i] == 0) last = i + 1;
    if (last) atomicMax(&s_len, last);
    __syncthreads();
    const int len = s_len;
    const size_t pitch16 = static_cast<size_t>(B) * row16;
    for (int r = warp; r < len; r += 8) {
      const uint4* sp = src + r * pitch16 + static_cast<size_t>(b) * row16;
      uint4* dp = dst + r * pitch16 + static_cast<size_t>(b) * row16;
      for (int c0 = lane; c0 < row16; c0 += 128) {
        uint4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (c0 + 32 * u < row16) v[u] = sp[c0 + 32 * u];
#pragma unroll
        for (int u = 0; u < 4; ++u)
          if (c0 + 32 * u < row16) dp[c0 + 32 * u] = v[u];
      }
    }
    __syncthreads();
  }
}

// rows s >= len_b of caption b (len_b = 1 + its last un-masked position) of dst [S, B, row16 * 16 bytes] are cleared: the padded
// tail of a source that crossed PCIe as ONE contiguous copy never reaches the kernels as whatever the host tensor held there
__global__ void __launch_bounds__(256) zero_tail_rows_kernel(uint4* dst, const uint8_t* mask, int S, int B, int row16) {
  __shared__ int s_len;
  const int b = blockIdx.x;
  if (threadIdx.x == 0) s_len = 0;
  __syncthreads();
  int last = 0;
  for (int s = threadIdx.x; s < S; s += blockDim.x)
    if (!mask[static_cast<size_t>(b) * S + s]) last = s + 1;
  if (last) atomicMax(&s_len, last);
  __syncthreads();
  const int len = s_len;
  const long n = static_cast<long>(S - len) * row16;
  for (long i = threadIdx.x; i < n; i += blockDim.x) {
    const long s = len + i / row16, c = i % row16;
    dst[(s * B + b) * row16 + c] = make_uint4(0, 0, 0, 0);
  }
}

// device-visible aliases of pinned host pointers (UVA: normally the same address); false = not mapped, use the DMA path
bool host_dev_ptr(const void* host, const void** dev) {
  void* p = nullptr;
  if (cudaHostGetDevicePointer(&p, const_cast<void*>(host), 0) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  *dev = p;
  return true;
}

// mask_dev: the same mask already on the device, uploaded on `st` before this call (or NULL when mask_host is)
int h2d_source(tnt_decoder* D, void* dst, const void* src_host, const uint8_t* mask_host, const uint8_t* mask_dev, int S, int B, int E,
               size_t esz, cudaStream_t st, size_t* bytes_out) {
  const size_t row = static_cast<size_t>(E) * esz, all = static_cast<size_t>(S) * B * row;
  if (!all) return TNT_OK;
  static const bool live_only = [] { const char* e = getenv("TNT_H2D_LIVE_ROWS"); return !(e && e[0] == '0'); }();
  if (!mask_host || !live_only) {
    TNT_CUDA(cudaMemcpyAsync(dst, src_host, all, cudaMemcpyHostToDevice, st));
    if (bytes_out) *bytes_out += all;
    return TNT_OK;
  }
  if (!D->zeroed_staging.count(dst)) {
    TNT_CUDA(cudaMemsetAsync(dst, 0, all, st));
    D->zeroed_staging.insert(dst);
  }
  const size_t pitch = static_cast<size_t>(B) * row;
  {
    // Every strided copy costs the DMA engine ~3.5 us of fixed time on top of its bytes (~40 GB/s beside a running step):
    // a source without padding (the image regions) or with little to skip (faces: 4 rows of 1 KB per caption; objects) is
    // cheaper as ONE contiguous copy of everything than as one copy per caption
    size_t live = 0;
    int n_copies = 0;
    for (int b = 0; b < B; ++b) {
      const uint8_t* m = mask_host + static_cast<size_t>(b) * S;
      int len = S;
      while (len > 0 && m[len - 1]) --len;
      live += row * len;
      n_copies += len > 0;
    }
    const double t_all = static_cast<double>(all) / 40e9, t_live = static_cast<double>(live) / 40e9 + n_copies * 3.5e-6;
    if (t_all <= t_live && (live == all || (mask_dev && row % 16 == 0 && aligned16(dst)))) {
      TNT_CUDA(cudaMemcpyAsync(dst, src_host, all, cudaMemcpyHostToDevice, st));
      if (bytes_out) *bytes_out += all;
      if (live < all) {       // the padded tails came along: cleared on the device, no kernel ever sees the host's contents
        zero_tail_rows_kernel<<<B, 256, 0, st>>>(static_cast<uint4*>(dst), mask_dev, S, B, static_cast<int>(row / 16));
        TNT_LAUNCHED();
      }
      return TNT_OK;
    }
  }
  static const bool zero_copy = [] { const char* e = getenv("TNT_H2D_ZEROCOPY"); return e && e[0] == '1'; }();   // measured slower: off
  const void *src_dev = nullptr, *mask_zc = nullptr;
  if (zero_copy && row % 16 == 0 && aligned16(dst) && aligned16(src_host) && host_dev_ptr(src_host, &src_dev) &&
      host_dev_ptr(mask_host, &mask_zc)) {
    h2d_live_rows_kernel<<<std::min(B, 2 * num_sms()), 256, 0, st>>>(static_cast<uint4*>(dst), static_cast<const uint4*>(src_dev),
                                                                     static_cast<const uint8_t*>(mask_zc), S, B,
                                                                     static_cast<int>(row / 16));
    TNT_LAUNCHED();
    if (bytes_out) {
      for (int b = 0; b < B; ++b) {
        const uint8_t* m = mask_host + static_cast<size_t>(b) * S;
        int len = S;
        while (len > 0 && m[len - 1]) --len;
        *bytes_out += row * len;
      }
    }
    return TNT_OK;
  }
  // The captions' strided copies are dealt round-robin to TNT_H2D_STREAMS streams (default 2).  Beside a running forward
  // pass one queue of 2-D copies drops to ~30 GB/s (tools/trace_e2e.py: 44 MB in 1.46 ms, the rest of the batch exposed);
  // two queues: e2e step 1.78 -> 1.72 ms on a 33 GB/s box, 1.84 (2.9 in one run) -> 1.72 ms on a 55 GB/s box.
  static const int n_q = [] { const char* e = getenv("TNT_H2D_STREAMS"); return e ? std::max(1, std::min(atoi(e), 4)) : 2; }();
  cudaStream_t q[4] = {st, st, st, st};
  const int nq = D ? n_q : 1;
  if (nq > 1) {
    if (!D->ev_cx_fork) {
      TNT_CUDA(cudaEventCreateWithFlags(&D->ev_cx_fork, cudaEventDisableTiming));
      for (int i = 0; i < 3; ++i) {
        TNT_CUDA(cudaStreamCreateWithFlags(&D->copy_extra[i], cudaStreamNonBlocking));
        TNT_CUDA(cudaEventCreateWithFlags(&D->ev_cx_join[i], cudaEventDisableTiming));
      }
    }
    TNT_CUDA(cudaEventRecord(D->ev_cx_fork, st));
    for (int i = 1; i < nq; ++i) {
      q[i] = D->copy_extra[i - 1];
      TNT_CUDA(cudaStreamWaitEvent(q[i], D->ev_cx_fork, 0));
    }
  }
  for (int b = 0; b < B; ++b) {
    const uint8_t* m = mask_host + static_cast<size_t>(b) * S;
    int len = S;
    while (len > 0 && m[len - 1]) --len;
    if (!len) continue;
    TNT_CUDA(cudaMemcpy2DAsync(static_cast<char*>(dst) + b * row, pitch, static_cast<const char*>(src_host) + b * row, pitch, row, len,
                               cudaMemcpyHostToDevice, q[b % nq]));
    if (bytes_out) *bytes_out += row * len;
  }
  for (int i = 1; i < nq; ++i) {
    TNT_CUDA(cudaEventRecord(D->ev_cx_join[i - 1], q[i]));
    TNT_CUDA(cudaStreamWaitEvent(st, D->ev_cx_join[i - 1], 0));
  }
  return TNT_OK;
}
}  // namespace
}  // namespace tnt

extern "C" size_t tnt_decoder_score_host_staging_bytes(const tnt_decoder* d, int B, int T, const int* S) {
  if (!d || !S) return 0;
  return carve_staging(d, B, T, S, nullptr).total;
}

extern "C" size_t tnt_decoder_loss_workspace_bytes(const tnt_decoder* d, int n_rows) {
  return d ? adasoftmax_workspace_bytes(d->ada, n_rows, true) : 0;
}

extern "C" int tnt_decoder_score_host(tnt_decoder* D, const int64_t* ids_host, const int64_t* target_host,
                                      const tnt_contexts* ch, float* out2_host, void* staging, size_t staging_bytes,
                                      void* workspace, size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && ids_host && target_host && ch && out2_host && staging && workspace, "score_host: null pointer");
  const int B = ch->B, T = ch->T;
  Staging s = carve_staging(D, B, T, ch->S, staging);
  if (s.total > staging_bytes) {
    set_error("score_host: staging %zu < %zu", staging_bytes, s.total);
    return TNT_ERR_WORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  TNT_CUDA(cudaMemcpyAsync(s.ids, ids_host, sizeof(int64_t) * B * T, cudaMemcpyHostToDevice, st));
  TNT_CUDA(cudaMemcpyAsync(s.tgt, target_host, sizeof(int64_t) * B * T, cudaMemcpyHostToDevice, st));
  D->last_h2d_bytes = 2 * sizeof(int64_t) * B * T;
  tnt_contexts cd = *ch;
  for (int i = 0; i < D->cfg.n_sources; ++i) {
    if (ch->pad_mask[i] && ch->S[i]) {
      TNT_CUDA(cudaMemcpyAsync(s.mask[i], ch->pad_mask[i], static_cast<size_t>(ch->S[i]) * B, cudaMemcpyHostToDevice, st));
      D->last_h2d_bytes += static_cast<size_t>(ch->S[i]) * B;
      cd.pad_mask[i] = s.mask[i];
    } else {
      cd.pad_mask[i] = nullptr;
    }
    TNT_TRY(h2d_source(D, s.mem[i], ch->memory[i], ch->pad_mask[i], cd.pad_mask[i], ch->S[i], B, D->cfg.source_dim[i], D->esz(), st,
                       &D->last_h2d_bytes));
    cd.memory[i] = s.mem[i];
  }
  const size_t fwd = carve_fwd(D, B, T, ch->S, nullptr, true).total;
  const size_t loss = adasoftmax_workspace_bytes(D->ada, B * T, false);
  if (fwd + loss > workspace_bytes) {
    set_error("score_host: workspace %zu < %zu", workspace_bytes, fwd + loss);
    return TNT_ERR_WORKSPACE;
  }
  TNT_TRY(tnt_decoder_forward(D, s.ids, &cd, s.X, nullptr, workspace, fwd, stream));
  TNT_TRY(tnt_decoder_nll(D, s.X, s.tgt, B * T, s.out2, static_cast<char*>(workspace) + fwd, workspace_bytes - fwd, stream));
  TNT_CUDA(cudaMemcpyAsync(out2_host, s.out2, 2 * sizeof(float), cudaMemcpyDeviceToHost, st));
  TNT_CUDA(cudaStreamSynchronize(st));
  return TNT_OK;
}

// Pipelined form: slot s's H2D copies run on a private copy stream and overlap the compute of the other
// slot; `wait` blocks until slot s's (loss, n_tokens) have landed in out2_host.
extern "C" int tnt_decoder_score_host_submit(tnt_decoder* D, int slot, const int64_t* ids_host,
                                             const int64_t* target_host, const tnt_contexts* ch, float* out2_host,
                                             void* staging, size_t staging_bytes, void* workspace,
                                             size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && ids_host && target_host && ch && out2_host && staging && workspace, "score_host_submit: null pointer");
  TNT_CHECK_ARG(slot == 0 || slot == 1, "score_host_submit: slot must be 0 or 1");
  const int B = ch->B, T = ch->T;
  Staging s = carve_staging(D, B, T, ch->S, staging);
  if (s.total > staging_bytes) {
    set_error("score_host_submit: staging %zu < %zu", staging_bytes, s.total);
    return TNT_ERR_WORKSPACE;
  }
  if (!D->copy_stream) {
    TNT_CUDA(cudaStreamCreateWithFlags(&D->copy_stream, cudaStreamNonBlocking));
    for (int i = 0; i < 2; ++i) {
      TNT_CUDA(cudaEventCreateWithFlags(&D->ev_h2d[i], cudaEventDisableTiming));
      TNT_CUDA(cudaEventCreateWithFlags(&D->ev_free[i], cudaEventDisableTiming));
      TNT_CUDA(cudaEventCreateWithFlags(&D->ev_done[i], cudaEventDisableTiming));
    }
  }
  cudaStream_t st = as_stream(stream), cs = D->copy_stream;
  if (D->slot_used[slot]) TNT_CUDA(cudaStreamWaitEvent(cs, D->ev_free[slot], 0));   // staging of this slot consumed
  TNT_CUDA(cudaMemcpyAsync(s.ids, ids_host, sizeof(int64_t) * B * T, cudaMemcpyHostToDevice, cs));
  TNT_CUDA(cudaMemcpyAsync(s.tgt, target_host, sizeof(int64_t) * B * T, cudaMemcpyHostToDevice, cs));
  D->last_h2d_bytes = 2 * sizeof(int64_t) * B * T;
  tnt_contexts cd = *ch;
  for (int i = 0; i < D->cfg.n_sources; ++i) {
    if (ch->pad_mask[i] && ch->S[i]) {
      TNT_CUDA(cudaMemcpyAsync(s.mask[i], ch->pad_mask[i], static_cast<size_t>(ch->S[i]) * B, cudaMemcpyHostToDevice, cs));
      D->last_h2d_bytes += static_cast<size_t>(ch->S[i]) * B;
      cd.pad_mask[i] = s.mask[i];
    } else {
      cd.pad_mask[i] = nullptr;
    }
    TNT_TRY(h2d_source(D, s.mem[i], ch->memory[i], ch->pad_mask[i], cd.pad_mask[i], ch->S[i], B, D->cfg.source_dim[i], D->esz(), cs,
                       &D->last_h2d_bytes));
    cd.memory[i] = s.mem[i];
  }
  TNT_CUDA(cudaEventRecord(D->ev_h2d[slot], cs));
  TNT_CUDA(cudaStreamWaitEvent(st, D->ev_h2d[slot], 0));
  const size_t fwd = carve_fwd(D, B, T, ch->S, nullptr, true).total;
  const size_t loss = adasoftmax_workspace_bytes(D->ada, B * T, false);
  if (fwd + loss > workspace_bytes) {
    set_error("score_host_submit: workspace %zu < %zu", workspace_bytes, fwd + loss);
    return TNT_ERR_WORKSPACE;
  }
  TNT_TRY(tnt_decoder_forward(D, s.ids, &cd, s.X, nullptr, workspace, fwd, stream));
  TNT_TRY(tnt_decoder_nll(D, s.X, s.tgt, B * T, s.out2, static_cast<char*>(workspace) + fwd, workspace_bytes - fwd, stream));
  TNT_CUDA(cudaMemcpyAsync(out2_host, s.out2, 2 * sizeof(float), cudaMemcpyDeviceToHost, st));
  TNT_CUDA(cudaEventRecord(D->ev_free[slot], st));
  TNT_CUDA(cudaEventRecord(D->ev_done[slot], st));
  D->slot_used[slot] = true;
  return TNT_OK;
}

// The live-row copy on its own, for callers that stage their step inputs themselves (a prefetching data loader: the
// training / generation legs of bench.py): dst [S, B, E] device, src [S, B, E] PINNED host, mask [B, S] host bytes
// (nonzero = padding) or NULL for a plain copy.  The skipped rows of dst are left as they are: zero dst once.
extern "C" int64_t tnt_h2d_live_rows(void* dst, const void* src_host, const uint8_t* mask_host, int S, int B, int E, int dtype,
                                     tnt_stream stream) {
  using namespace tnt;
  if (!dst || !src_host || S < 0 || B <= 0 || E <= 0) {
    set_error("h2d_live_rows: bad arguments");
    return TNT_ERR_ARG;
  }
  const size_t row = static_cast<size_t>(E) * dtype_size(dtype), pitch = static_cast<size_t>(B) * row;
  cudaStream_t st = as_stream(stream);
  int64_t bytes = 0;
  if (!mask_host) {
    if (cudaMemcpyAsync(dst, src_host, pitch * S, cudaMemcpyHostToDevice, st) != cudaSuccess) return TNT_ERR_CUDA;
    return static_cast<int64_t>(pitch * S);
  }
  {
    // one contiguous copy when skipping the padding would not pay for the strided copies' fixed cost (see h2d_source)
    size_t live = 0;
    int n_copies = 0;
    for (int b = 0; b < B; ++b) {
      const uint8_t* m = mask_host + static_cast<size_t>(b) * S;
      int len = S;
      while (len > 0 && m[len - 1]) --len;
      live += row * len;
      n_copies += len > 0;
    }
    if (static_cast<double>(pitch * S) / 40e9 <= static_cast<double>(live) / 40e9 + n_copies * 3.5e-6) {
      if (cudaMemcpyAsync(dst, src_host, pitch * S, cudaMemcpyHostToDevice, st) != cudaSuccess) return TNT_ERR_CUDA;
      return static_cast<int64_t>(pitch * S);
    }
  }
  {
    static const bool zero_copy = [] { const char* e = getenv("TNT_H2D_ZEROCOPY"); return e && e[0] == '1'; }();   // measured slower: off
    const void *src_dev = nullptr, *mask_dev = nullptr;
    if (zero_copy && row % 16 == 0 && aligned16(dst) && aligned16(src_host) && host_dev_ptr(src_host, &src_dev) &&
        host_dev_ptr(mask_host, &mask_dev)) {
      h2d_live_rows_kernel<<<std::min(B, 2 * num_sms()), 256, 0, st>>>(static_cast<uint4*>(dst), static_cast<const uint4*>(src_dev),
                                                                       static_cast<const uint8_t*>(mask_dev), S, B,
                                                                       static_cast<int>(row / 16));
      TNT_LAUNCHED();
      for (int b = 0; b < B; ++b) {
        const uint8_t* m = mask_host + static_cast<size_t>(b) * S;
        int len = S;
        while (len > 0 && m[len - 1]) --len;
        bytes += static_cast<int64_t>(row) * len;
      }
      return bytes;
    }
  }
  for (int b = 0; b < B; ++b) {
    const uint8_t* m = mask_host + static_cast<size_t>(b) * S;
    int len = S;
    while (len > 0 && m[len - 1]) --len;
    if (!len) continue;
    if (cudaMemcpy2DAsync(static_cast<char*>(dst) + b * row, pitch, static_cast<const char*>(src_host) + b * row, pitch, row, len,
                          cudaMemcpyHostToDevice, st) != cudaSuccess) {
      set_error("h2d_live_rows: cudaMemcpy2DAsync failed");
      return TNT_ERR_CUDA;
    }
    bytes += static_cast<int64_t>(row) * len;
  }
  return bytes;
}

// host -> device bytes the last tnt_decoder_score_host / _submit call copied (ids, targets, masks and the LIVE rows of every
// memory source): what bench.py reports as e2e.h2d_bytes_per_step
extern "C" uint64_t tnt_decoder_last_h2d_bytes(const tnt_decoder* D) { return D ? D->last_h2d_bytes : 0; }

extern "C" int tnt_decoder_score_host_wait(tnt_decoder* D, int slot) {
  TNT_CHECK_ARG(D && (slot == 0 || slot == 1) && D->slot_used[slot], "score_host_wait: slot was not submitted");
  TNT_CUDA(cudaEventSynchronize(D->ev_done[slot]));
  return TNT_OK;
}
