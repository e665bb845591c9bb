// Decoder engine: teacher-forced forward, loss / log-probs, greedy generation with incremental state.
//   DynamicConvDecoder.forward       ttl/tell/models/decoder_flattened.py:96-146
//   DynamicConvDecoderLayer.forward  ttl/tell/models/decoder_flattened.py:271-345 (post-LN),
//                                    ttl/tell/models/decoder_faces_objects.py:283-393
//   _generate (top-1)                ttl/tell/models/transformer_flattened.py:238-325
#include "engine.cuh"

namespace tnt {

int xattn_check(const tnt_xattn_params* p);
int xattn_fwd_dispatch(const tnt_xattn_params* p, cudaStream_t st);
bool xattn_step_supported(const tnt_xattn_params* p);
int xattn_fwd_step_multi(const tnt_xattn_params* ps, int n, cudaStream_t st, int* sched = nullptr, const int* const* orders = nullptr,
                         float* scores = nullptr, int scores_src = -1, float* weights_out = nullptr);
int xattn_live_order(const uint32_t* tile_live, int B, int* order, cudaStream_t st);
bool xattn_force_simt();
bool xattn_short_supported(const tnt_xattn_params* p);
int xattn_short_multi(const tnt_xattn_params* ps, int n, cudaStream_t st);
bool linear_ln_enabled();                                                                                    // gemm_ln.cu
bool linear_ln_supported(const tnt_linear_params* p, const float* gamma, const float* beta);
int launch_linear_ln(const tnt_linear_params* p, const float* gamma, const float* beta, long stride_gb, float eps, int tT, int tB,
                     cudaStream_t st);

namespace {
inline size_t al256(size_t x) { return (x + 255) & ~size_t(255); }

int linear(const void* a, long lda, const void* w, long ldw, const float* bias, void* c, long ldc, int M, int N, int K,
           int dtype, int epi, const void* res, long ldr, cudaStream_t st, int groups = 1, long sa = 0, long sw = 0,
           long sb = 0, long sr = 0, long sc = 0, const void* pf = nullptr, size_t pf_bytes = 0,
           const tnt_dropout* drop = nullptr) {
  tnt_linear_params p = {};
  p.a = a; p.lda = lda; p.w = w; p.ldw = ldw; p.bias = bias; p.residual = res; p.ldr = ldr; p.c = c; p.ldc = ldc;
  p.M = M; p.N = N; p.K = K; p.groups = groups;
  p.stride_a = sa; p.stride_w = sw; p.stride_bias = sb; p.stride_res = sr; p.stride_c = sc;
  p.dtype_in = dtype; p.dtype_out = dtype; p.epilogue = epi;
  p.prefetch = pf; p.prefetch_bytes = pf_bytes;
  if (drop) p.drop = *drop;
  return launch_gemm(&p, -1, nullptr, nullptr, st);
}

int layernorm(const void* x, void* y, long ld, const float* g, const float* b, int rows, int C, int groups, int dtype,
              int tT, int tB, cudaStream_t st) {
  tnt_layernorm_params p = {};
  p.x = x; p.ldx = ld; p.y = y; p.ldy = ld; p.gamma = g; p.beta = b;
  p.rows = rows; p.C = C; p.groups = groups; p.variant = 0; p.eps = 1e-5f; p.dtype = dtype;
  p.transpose_T = tT; p.transpose_B = tB;
  return tnt_layernorm_fwd(&p, st);
}
// out = LayerNorm(a . w^T + bias + res): ONE launch (gemm_ln.cu: a cluster of CTAs per 128-row block, row statistics through
// distributed shared memory) when `fuse` and the shape allows it, else the GEMM into `tmp` followed by the LayerNorm kernel.
// Grouped form (per-source out_proj + context_attn_lns): group s reads a + s*sa / w + s*sw / bias + s*sb, writes columns
// [s*sc, s*sc + N) of `out` and normalises them with gamma / beta block s.
int linear_res_ln(const void* a, long lda, const void* w, long ldw, const float* bias, const void* res, long ldr, void* tmp, void* out,
                  long ldo, int M, int N, int K, int dtype, const float* gamma, const float* beta, int groups, long sa, long sw,
                  long sb, long sc, const void* pf, size_t pf_bytes, int tT, int tB, const tnt_dropout* drop, bool fuse,
                  cudaStream_t st) {
  if (fuse) {
    tnt_linear_params p = {};
    p.a = a; p.lda = lda; p.w = w; p.ldw = ldw; p.bias = bias; p.residual = res; p.ldr = ldr; p.c = out; p.ldc = ldo;
    p.M = M; p.N = N; p.K = K; p.groups = groups;
    p.stride_a = sa; p.stride_w = sw; p.stride_bias = sb; p.stride_res = 0; p.stride_c = sc;
    p.dtype_in = dtype; p.dtype_out = dtype; p.epilogue = EPI_RESIDUAL;
    p.prefetch = pf; p.prefetch_bytes = pf_bytes;
    if (linear_ln_supported(&p, gamma, beta)) return launch_linear_ln(&p, gamma, beta, N, 1e-5f, tT, tB, st);
  }
  TNT_TRY(linear(a, lda, w, ldw, bias, tmp, ldo, M, N, K, dtype, EPI_RESIDUAL, res, ldr, st, groups, sa, sw, sb, 0, sc, pf, pf_bytes, drop));
  // the grouped LayerNorm sees [M, groups * N] as M * groups rows of N (block = row % groups)
  return layernorm(tmp, out, groups > 1 ? N : ldo, gamma, beta, M * groups, N, groups, dtype, tT, tB, st);
}
// rows S and S+1 of a K/V buffer: the learned bias_k|bias_v row and the all-zero row that
// add_bias_kv / add_zero_attn append to every source (multi_head.py:355-374, 416-427)
// CTAs beyond `main_ctas` (8 batch rows each, one warp per row) compute the live-tile bitmask of the source's
// key-padding mask for the teacher-forced attention kernels (attention_tc.cu) -- no launch of its own.
template <typename T>
__global__ void __launch_bounds__(256) kv_extra_rows_kernel(T* kv, const float* row, int S, int B, int ld, int main_ctas,
                                                            const uint8_t* mask, uint32_t* tile_live) {
  if (static_cast<int>(blockIdx.x) >= main_ctas) {
    const int b = (blockIdx.x - main_ctas) * 8 + (threadIdx.x >> 5);
    if (b < B) tile_bits_row(mask, S, b, S + 2, S, tile_live, threadIdx.x & 31);
    return;
  }
  const long n = static_cast<long>(B) * ld;
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n; i += static_cast<long>(main_ctas) * blockDim.x) {
    const int c = static_cast<int>(i % ld);
    kv[static_cast<long>(S) * n + i] = from_f<T>(row[c]);
    kv[static_cast<long>(S + 1) * n + i] = from_f<T>(0.f);
  }
}
}  // namespace

FwdWs carve_fwd(const tnt_decoder* D, int B, int T, const int* S, void* ws, bool with_kv, int B_kv) {
  const tnt_decoder_config& c = D->cfg;
  const size_t E = D->esz(), N = static_cast<size_t>(B) * T, d = c.d, ns = c.n_sources;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += al256(bytes);
    return p;
  };
  FwdWs w = {};
  w.X0 = take(N * d * E); w.X1 = take(N * d * E); w.h = take(N * d * E);
  w.logits = take(N * D->max_ld_logits() * E);
  w.cbuf = take(N * d * E); w.ybuf = take(N * d * E);
  w.Q = take(N * ns * d * E); w.AO = take(N * ns * d * E); w.CAT = take(N * ns * d * E);
  w.F = take(N * c.ffn * E);
  w.ybuf2 = w.ybuf;
  w.CATln = w.CAT;
  for (size_t s = 0; s < ns; ++s)
    w.kv[s] = with_kv ? take(static_cast<size_t>(S[s] + 2) * (B_kv >= 0 ? B_kv : B) * c.n_layers * 2 * d * E) : nullptr;
  w.sched = with_kv ? static_cast<int*>(take(2 * tnt_decoder::kMaxSub * sizeof(int))) : nullptr;    // work counters of the decode-step attention (a pair per sub-batch)
  for (size_t s = 0; s < ns; ++s) w.order[s] = with_kv ? static_cast<int*>(take(sizeof(int) * (B_kv >= 0 ? B_kv : B))) : nullptr;
  for (size_t s = 0; s < ns; ++s)
    w.tile_live[s] = with_kv ? static_cast<uint32_t*>(take(sizeof(uint32_t) * (B_kv >= 0 ? B_kv : B))) : nullptr;
  w.total = off;
  return w;
}

static int g_sub_batches = 0;    // test hook (tnt_debug_sub_batches): > 0 overrides TNT_SUB_BATCHES / the default

int sub_plan(tnt_decoder* D, int B, int dflt, int min_rows, cudaStream_t st, SubBatches* sb) {
  static const int env_n = [] { const char* e = getenv("TNT_SUB_BATCHES"); return e ? atoi(e) : -1; }();
  int n = g_sub_batches > 0 ? g_sub_batches : env_n >= 1 ? env_n : dflt;
  if (g_sub_batches > 0) min_rows = 1;
  n = std::max(1, std::min(n, static_cast<int>(tnt_decoder::kMaxSub)));
  while (n > 1 && B / n < min_rows) --n;
  sb->n = n;
  for (int i = 0; i <= n; ++i) sb->b0[i] = static_cast<int>(static_cast<long>(B) * i / n);
  sb->st[0] = st;
  if (n > 1 && !D->sub_fork) {
    TNT_CUDA(cudaEventCreateWithFlags(&D->sub_fork, cudaEventDisableTiming));
    for (int i = 0; i < tnt_decoder::kMaxSub - 1; ++i) {
      TNT_CUDA(cudaStreamCreateWithFlags(&D->sub_streams[i], cudaStreamNonBlocking));
      TNT_CUDA(cudaEventCreateWithFlags(&D->sub_join[i], cudaEventDisableTiming));
    }
  }
  for (int i = 1; i < n; ++i) sb->st[i] = D->sub_streams[i - 1];
  return TNT_OK;
}

int sub_fork(tnt_decoder* D, const SubBatches& sb) {
  if (sb.n <= 1) return TNT_OK;
  TNT_CUDA(cudaEventRecord(D->sub_fork, sb.st[0]));
  for (int i = 1; i < sb.n; ++i) TNT_CUDA(cudaStreamWaitEvent(sb.st[i], D->sub_fork, 0));
  return TNT_OK;
}

int sub_join(tnt_decoder* D, const SubBatches& sb) {
  for (int i = 1; i < sb.n; ++i) {
    TNT_CUDA(cudaEventRecord(D->sub_join[i - 1], sb.st[i]));
    TNT_CUDA(cudaStreamWaitEvent(sb.st[0], D->sub_join[i - 1], 0));
  }
  return TNT_OK;
}

FwdWs sub_ws(const tnt_decoder* D, const FwdWs& w, const SubBatches& sb, int i, int B, int T, const int* S) {
  (void)S;
  const tnt_decoder_config& c = D->cfg;
  const size_t E = D->esz(), r0 = static_cast<size_t>(sb.b0[i]) * T, d = c.d, ns = c.n_sources;
  auto rows = [&](void* p, size_t width) -> void* { return p ? static_cast<char*>(p) + r0 * width * E : nullptr; };
  FwdWs f = w;
  f.X0 = rows(w.X0, d); f.X1 = rows(w.X1, d); f.h = rows(w.h, d);
  f.logits = rows(w.logits, D->max_ld_logits());
  f.cbuf = rows(w.cbuf, d); f.ybuf = rows(w.ybuf, d);
  f.Q = rows(w.Q, ns * d); f.AO = rows(w.AO, ns * d); f.CAT = rows(w.CAT, ns * d);
  f.F = rows(w.F, c.ffn);
  f.ybuf2 = f.ybuf;
  f.CATln = f.CAT;
  f.Xd = nullptr;
  const size_t ldkv = static_cast<size_t>(c.n_layers) * 2 * d;
  for (size_t s = 0; s < ns; ++s) {
    f.kv[s] = w.kv[s] ? static_cast<char*>(w.kv[s]) + static_cast<size_t>(sb.b0[i]) * ldkv * E : nullptr;
    f.lse[s] = nullptr;
    f.tile_live[s] = w.tile_live[s] ? w.tile_live[s] + sb.b0[i] : nullptr;
    f.order[s] = w.order[s] ? w.order[s] + sb.b0[i] : nullptr;
  }
  f.sched = w.sched ? w.sched + 2 * i : nullptr;
  f.kv_B = B;
  return f;
}

tnt_contexts sub_ctx(const tnt_contexts& c, const SubBatches& sb, int i) {
  tnt_contexts s = c;
  s.B = sb.b0[i + 1] - sb.b0[i];
  for (int k = 0; k < TNT_MAX_SOURCES; ++k)
    if (c.pad_mask[k]) s.pad_mask[k] = c.pad_mask[k] + static_cast<size_t>(sb.b0[i]) * c.S[k];
  return s;
}

// K/V of every layer for every source in ONE GEMM per source: [S*B, kdim] x [L*2d, kdim]^T (+bias).
// The reference recomputes these projections in every layer call and every decoding step
// (multi_head.py:347-352 with incremental_state=None, decoder_flattened.py:305,323); same numbers.
int project_kv(tnt_decoder* D, const tnt_contexts* ctx, const FwdWs& w, cudaStream_t st, bool decode_order, const SubBatches* sb) {
  const tnt_decoder_config& c = D->cfg;
  const int ldkv = c.n_layers * 2 * c.d;
  if (w.sched) TNT_CUDA(cudaMemsetAsync(w.sched, 0, 2 * tnt_decoder::kMaxSub * sizeof(int), st));   // every launch leaves them at zero again
  for (int s = 0; s < c.n_sources; ++s) {
    const int grid = std::min<long>((static_cast<long>(ctx->B) * ldkv + 255) / 256, 4L * num_sms());
    const bool scan = w.tile_live[s] && ctx->pad_mask[s] && ctx->S[s] > 0;
    const int extra = scan ? (ctx->B + 7) / 8 : 0;
    if (c.dtype == TNT_BF16)
      kv_extra_rows_kernel<bf16><<<grid + extra, 256, 0, st>>>(static_cast<bf16*>(w.kv[s]), D->kvrow[s], ctx->S[s], ctx->B, ldkv,
                                                               grid, ctx->pad_mask[s], w.tile_live[s]);
    else
      kv_extra_rows_kernel<float><<<grid + extra, 256, 0, st>>>(static_cast<float*>(w.kv[s]), D->kvrow[s], ctx->S[s], ctx->B, ldkv,
                                                                grid, ctx->pad_mask[s], w.tile_live[s]);
    TNT_LAUNCHED();
    // generation: the decode-step attention starts the captions with the most live keys first (once per call)
    if (decode_order && scan && w.order[s] && ctx->B <= 4096) {
      if (sb && sb->n > 1) {                 // an order per sub-batch, relative to its first entry
        for (int i = 0; i < sb->n; ++i)
          TNT_TRY(xattn_live_order(w.tile_live[s] + sb->b0[i], sb->b0[i + 1] - sb->b0[i], w.order[s] + sb->b0[i], st));
      } else {
        TNT_TRY(xattn_live_order(w.tile_live[s], ctx->B, w.order[s], st));
      }
    }
    if (ctx->S[s] == 0) continue;
    TNT_CHECK_ARG(ctx->memory[s], "decoder: context '%s' missing", c.source_names[s]);
    char pname[40];
    snprintf(pname, sizeof(pname), "kv_proj.%s", c.source_names[s]);
    const int pid = prof_begin(pname, st);
    TNT_TRY(linear(ctx->memory[s], c.source_dim[s], D->wkv[s], c.source_dim[s], D->bkv[s], w.kv[s], ldkv,
                   ctx->S[s] * ctx->B, ldkv, c.source_dim[s], c.dtype, EPI_NONE, nullptr, 0, st));
    prof_end(pid, st);
  }
  return TNT_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// K/V projections BESIDE the layer chain (teacher-forced forward) -- MEASURED SLOWER, off unless TNT_KV_BESIDE_SMS / the test
// hook asks for it.  The idea: the projections are throughput-bound (article: 550 GF at B = 64, 0.39 ms on the whole GPU) and
// nothing in a layer needs them before its attention, while the layer chain is latency-bound (14 dependent launches of
// 6-22 us per layer for ~55 us of tensor work).  So the projections run layer by layer (W rows / output columns
// [l*2d, (l+1)*2d) of the fused per-source projection: the same dot products as the one-GEMM form) on a library-owned stream
// with their persistent grids sized for `sms` SMs, the chain keeps the other SMs, and layer l's attention waits for
// kv_ready[l].  Sources with few rows keep the one-GEMM form up front.  Result (B200, tools/sweep_kv_beside.py,
// profiles/r2_kv_beside_sweep.txt): C2 1.41 ms in the plain order, 1.47-1.66 ms at every share from 56 to 132 SMs: the chain's
// one-wave GEMMs (128 tiles) take 2-3 CTA waves on the SMs the projections leave them and stretch 12 -> 30-55 us each.
constexpr int kKvBesideSms = 104; // share of the projections when the mode is asked for without a number (best of the sweep)
static int g_kv_beside = -1;      // test hook (tnt_debug_kv_beside): -1 automatic, 0 off, > 0 on with that many SMs

int kv_beside_sms(const tnt_decoder* D, const tnt_contexts* ctx) {
  static const int env = [] { const char* e = getenv("TNT_KV_BESIDE_SMS"); return e ? atoi(e) : -1; }();
  const int want = g_kv_beside >= 0 ? g_kv_beside : env;
  if (want == 0 || sm_budget() > 0) return 0;          // (a caller that already shares the SMs keeps the plain order)
  long rows = 0;
  for (int s = 0; s < D->cfg.n_sources; ++s) rows += static_cast<long>(ctx->S[s]) * ctx->B;
  if (want < 0 || rows == 0) return 0;      // measured slower than the plain order at every share: off unless asked for
  const int all = num_sms();
  const int sms = want > 0 ? want : kKvBesideSms;
  return std::max(2, std::min(sms, all)) & ~1;
}

// which sources are projected layer by layer (the others keep the one-GEMM form, issued by kv_prologue)
struct KvByLayer { bool on[TNT_MAX_SOURCES] = {}; };

// extra rows + live-tile words of every source, and the one-GEMM projection of the sources that are NOT split by layer
int kv_prologue(tnt_decoder* D, const tnt_contexts* ctx, const FwdWs& w, cudaStream_t ks, long min_rows, KvByLayer* bl) {
  const tnt_decoder_config& c = D->cfg;
  const int ldkv = c.n_layers * 2 * c.d;
  if (w.sched) TNT_CUDA(cudaMemsetAsync(w.sched, 0, 2 * tnt_decoder::kMaxSub * sizeof(int), ks));
  for (int s = 0; s < c.n_sources; ++s) {
    const int grid = std::min<long>((static_cast<long>(ctx->B) * ldkv + 255) / 256, 4L * num_sms());
    const bool scan = w.tile_live[s] && ctx->pad_mask[s] && ctx->S[s] > 0;
    const int extra = scan ? (ctx->B + 7) / 8 : 0;
    if (c.dtype == TNT_BF16)
      kv_extra_rows_kernel<bf16><<<grid + extra, 256, 0, ks>>>(static_cast<bf16*>(w.kv[s]), D->kvrow[s], ctx->S[s], ctx->B, ldkv,
                                                               grid, ctx->pad_mask[s], w.tile_live[s]);
    else
      kv_extra_rows_kernel<float><<<grid + extra, 256, 0, ks>>>(static_cast<float*>(w.kv[s]), D->kvrow[s], ctx->S[s], ctx->B, ldkv,
                                                                grid, ctx->pad_mask[s], w.tile_live[s]);
    TNT_LAUNCHED();
    if (ctx->S[s] == 0) continue;
    TNT_CHECK_ARG(ctx->memory[s], "decoder: context '%s' missing", c.source_names[s]);
    bl->on[s] = static_cast<long>(ctx->S[s]) * ctx->B >= min_rows;
    if (bl->on[s]) continue;
    TNT_TRY(linear(ctx->memory[s], c.source_dim[s], D->wkv[s], c.source_dim[s], D->bkv[s], w.kv[s], ldkv,
                   ctx->S[s] * ctx->B, ldkv, c.source_dim[s], c.dtype, EPI_NONE, nullptr, 0, ks));
  }
  return TNT_OK;
}

// layer l's K/V of the split sources: W rows / output columns [l*2d, (l+1)*2d) of the fused projection
int kv_layer(tnt_decoder* D, const tnt_contexts* ctx, const FwdWs& w, int l, cudaStream_t ks, const KvByLayer& bl) {
  const tnt_decoder_config& c = D->cfg;
  const int ldkv = c.n_layers * 2 * c.d;
  const size_t E = D->esz(), n0 = static_cast<size_t>(l) * 2 * c.d;
  for (int s = 0; s < c.n_sources; ++s) {
    if (!bl.on[s]) continue;
    int pid = -1;
    if (prof_enabled()) {
      char pname[40];
      snprintf(pname, sizeof(pname), "kv_proj.%s", c.source_names[s]);
      pid = prof_begin(pname, ks);
    }
    TNT_TRY(linear(ctx->memory[s], c.source_dim[s], static_cast<const char*>(D->wkv[s]) + n0 * c.source_dim[s] * E, c.source_dim[s],
                   D->bkv[s] ? D->bkv[s] + n0 : nullptr, static_cast<char*>(w.kv[s]) + n0 * E, ldkv, ctx->S[s] * ctx->B, 2 * c.d,
                   c.source_dim[s], c.dtype, EPI_NONE, nullptr, 0, ks));
    prof_end(pid, ks);
  }
  return TNT_OK;
}

int project_kv_beside(tnt_decoder* D, const tnt_contexts* ctx, FwdWs& w, cudaStream_t st, int sms) {
  const tnt_decoder_config& c = D->cfg;
  if (!D->kv_stream) {
    TNT_CUDA(cudaStreamCreateWithFlags(&D->kv_stream, cudaStreamNonBlocking));
    TNT_CUDA(cudaEventCreateWithFlags(&D->kv_fork, cudaEventDisableTiming));
  }
  while (static_cast<int>(D->kv_ready.size()) < c.n_layers) {
    cudaEvent_t e;
    TNT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    D->kv_ready.push_back(e);
  }
  cudaStream_t ks = D->kv_stream;
  // the side stream sees everything issued on `st` so far: the contexts' uploads, and the previous call's attention reads of
  // the K/V buffers this call overwrites
  TNT_CUDA(cudaEventRecord(D->kv_fork, st));
  TNT_CUDA(cudaStreamWaitEvent(ks, D->kv_fork, 0));
  SmBudget budget(sms);
  KvByLayer bl;
  TNT_TRY(kv_prologue(D, ctx, w, ks, 2048, &bl));
  for (int l = 0; l < c.n_layers; ++l) {
    TNT_TRY(kv_layer(D, ctx, w, l, ks, bl));
    TNT_CUDA(cudaEventRecord(D->kv_ready[l], ks));
  }
  w.kv_ready = D->kv_ready.data();
  variant_hit(V_KV_BESIDE);
  return TNT_OK;
}

// K/V projections of the large sources INLINE, layer by layer: layer l's slice is issued on the step's own stream right before
// layer l, so that its attention finds the (live part of the) 135 MB it reads in the 126 MB L2 instead of in HBM.  MEASURED: the
// article attention stays at 30 us per layer (it is not HBM-bound at this size) and the four narrower GEMMs cost 4 x 100 us
// instead of 388 us: C2 1.40 -> 1.42 ms, C4 2.96 -> 3.07 ms (tools/ab_kv_inline.py).  Off unless TNT_KV_INLINE=1 / the test hook.
static int g_kv_inline = -1;      // test hook (tnt_debug_kv_inline): -1 automatic, 0 off, 1 on
bool kv_inline_on(const tnt_decoder* D, const tnt_contexts* ctx) {
  static const int env = [] { const char* e = getenv("TNT_KV_INLINE"); return e ? atoi(e) : -1; }();
  const int want = g_kv_inline >= 0 ? g_kv_inline : env;
  (void)D; (void)ctx;
  return want > 0;
}

int run_layer(tnt_decoder* D, int l, const tnt_contexts* ctx, const FwdWs& w, const void* Xin, void* Xout, int B, int T,
              void* ring, const int* pos, float* attn_out, bool transpose_out, cudaStream_t st, int kv_group,
              const tnt_dropout_config* dr, const ConvHist* ch) {
  const tnt_decoder_config& c = D->cfg;
  const LayerW& L = D->layers[l];
  const int d = c.d, ns = c.n_sources, N = B * T, H = c.heads, dt = c.dtype;
  const int ldl = D->ld_logits(L.K);
  const size_t E = D->esz();
  // conv block: linear1 + GLU -> tap logits -> dynamic conv -> linear2 + residual -> LayerNorm
  // each GEMM of the chain asks for the NEXT GEMM's weights to be staged in L2 while it runs (the step's K/V traffic
  // leaves no weight L2-resident from one step to the next; the small-M GEMMs are otherwise DRAM-latency bound)
  const size_t dd = static_cast<size_t>(d) * d * E;
  const void* next_w1 = l + 1 < c.n_layers ? D->layers[l + 1].w1 : D->ada.head_w;
  const size_t next_w1_bytes = l + 1 < c.n_layers ? 2 * dd : static_cast<size_t>(D->ada.cuts.cut[0] + D->ada.cuts.n - 1) * d * E;
  // dropout sites of the layer (training only; decoder_flattened.py:283,289,308,326,340,342)
  const tnt_dropout d_conv = drop_site(dr, dr ? dr->dropout : 0.f, TNT_DROP_SITE(l, TNT_DROP_CONV_OUT));
  const tnt_dropout d_attn = drop_site(dr, dr ? dr->dropout : 0.f, TNT_DROP_SITE(l, TNT_DROP_ATTN_OUT), static_cast<long>(ns) * d, d);
  const tnt_dropout d_ffn = drop_site(dr, dr ? dr->dropout : 0.f, TNT_DROP_SITE(l, TNT_DROP_FFN_OUT));
  const tnt_dropout d_relu = drop_site(dr, dr ? dr->relu_dropout : 0.f, TNT_DROP_SITE(l, TNT_DROP_RELU));
  const void* X1in = Xin;                 // what linear1 reads: the layer input after input_dropout
  if (dr && dr->input_dropout > 0.f) {
    TNT_CHECK_ARG(w.Xd, "decoder: input_dropout needs the training workspace");
    TNT_TRY(dropout_apply(Xin, d, nullptr, 0, w.Xd, d, N, d, dt, drop_site(dr, dr->input_dropout, TNT_DROP_SITE(l, TNT_DROP_INPUT)), st));
    X1in = w.Xd;
  }
  TNT_TRY(linear(X1in, d, L.w1, d, L.b1, w.h, d, N, 2 * d, d, dt, EPI_GLU, nullptr, 0, st, 1, 0, 0, 0, 0, 0, L.wtap,
                 static_cast<size_t>(H) * L.K * d * E));
  TNT_TRY(linear(w.h, d, L.wtap, d, nullptr, w.logits, ldl, N, H * L.K, d, dt, EPI_NONE, nullptr, 0, st, 1, 0, 0, 0, 0, 0, L.w2, dd));
  if (ring || pos) {
    TNT_CHECK_ARG(T == 1, "decoder: incremental layer call needs T == 1");
    tnt_dynconv_step_params sp = {};
    sp.x = w.h; sp.ldx = d; sp.logits = w.logits; sp.ldl = ldl; sp.y = w.cbuf; sp.ldy = d; sp.ring = ring; sp.pos = pos;
    sp.B = B; sp.C = d; sp.H = H; sp.K = L.K; sp.dtype = dt;
    TNT_TRY(tnt_dynconv_step(&sp, st));
  } else {
    tnt_dynconv_params cp = {};
    cp.x = w.h; cp.ldx = d; cp.logits = w.logits; cp.ldl = ldl; cp.y = w.cbuf; cp.ldy = d;
    cp.T = T; cp.B = B; cp.C = d; cp.H = H; cp.K = L.K; cp.dtype = dt;
    cp.drop = drop_site(dr, dr ? dr->weight_dropout : 0.f, TNT_DROP_SITE(l, TNT_DROP_TAPS));
    if (ch && ch->len > 0) {
      TNT_CHECK_ARG(ch->in && ch->len <= L.K - 1, "decoder: conv history of layer %d has %d rows (K = %d)", l, ch->len, L.K);
      cp.hist = ch->in;
      cp.hist_len = ch->len;
    }
    TNT_TRY(tnt_dynconv_fwd(&cp, st));
    if (ch && ch->out && L.K > 1) {      // new history = last min(K-1, len + T) rows of [history ; this call's conv inputs]
      const int keep = std::min(L.K - 1, ch->len + T), from_x = std::min(T, keep), from_h = keep - from_x;
      const size_t row = static_cast<size_t>(B) * d * E;
      if (from_h > 0)
        TNT_CUDA(cudaMemcpyAsync(ch->out, static_cast<const char*>(ch->in) + static_cast<size_t>(ch->len - from_h) * row, from_h * row,
                                 cudaMemcpyDeviceToDevice, st));
      TNT_CUDA(cudaMemcpyAsync(static_cast<char*>(ch->out) + from_h * row, static_cast<const char*>(w.h) + static_cast<size_t>(T - from_x) * row,
                               from_x * row, cudaMemcpyDeviceToDevice, st));
    }
  }
  // LayerNorm folded into the producing GEMM (inference, bf16): the pre-LayerNorm tensors are not kept
  const bool fuse_ln = !dr && dt == TNT_BF16 && w.ybuf2 == w.ybuf && w.CATln == w.CAT && linear_ln_enabled();
  TNT_TRY(linear_res_ln(w.cbuf, d, L.w2, d, L.b2, Xin, d, w.ybuf, w.X1, d, N, d, d, dt, L.ln_conv_g, L.ln_conv_b, 1, 0, 0, 0, 0,
                        L.wq, ns * dd, 0, 0, &d_conv, fuse_ln, st));
  // multi-source cross attention: one stacked query projection, per-source attention, grouped out-proj
  TNT_TRY(linear(w.X1, d, L.wq, d, L.bq, w.Q, static_cast<long>(ns) * d, N, ns * d, d, dt, EPI_NONE, nullptr, 0, st, 1, 0, 0, 0, 0, 0,
                 L.wo, ns * dd));
  if (w.kv_ready) TNT_CUDA(cudaStreamWaitEvent(st, w.kv_ready[l], 0));   // K/V of this layer come from the side stream (project_kv_beside)
  const long ldkv = static_cast<long>(c.n_layers) * 2 * d;
  tnt_xattn_params aps[TNT_MAX_SOURCES];
  bool all_step = T == 1;
  for (int s = 0; s < ns; ++s) {
    tnt_xattn_params& ap = aps[s];
    ap = tnt_xattn_params{};
    ap.q = static_cast<const char*>(w.Q) + static_cast<size_t>(s) * d * E;
    ap.q_stride_t = static_cast<long>(B) * ns * d; ap.q_stride_b = static_cast<long>(ns) * d;
    ap.k = static_cast<const char*>(w.kv[s]) + static_cast<size_t>(l) * 2 * d * E;
    ap.k_stride_s = static_cast<long>(w.kv_B ? w.kv_B : B / (kv_group > 1 ? kv_group : 1)) * ldkv; ap.k_stride_b = ldkv;   // cache rows = context entries
    ap.v = static_cast<const char*>(ap.k) + static_cast<size_t>(d) * E;
    ap.v_stride_s = ap.k_stride_s; ap.v_stride_b = ap.k_stride_b;
    ap.out = static_cast<char*>(w.AO) + static_cast<size_t>(s) * d * E;
    ap.o_stride_t = ap.q_stride_t; ap.o_stride_b = ap.q_stride_b;
    ap.mask = ctx->pad_mask[s]; ap.m_stride_b = ctx->S[s]; ap.m_stride_t = 0;
    ap.mask_mode = (ctx->pad_mask[s] && ctx->S[s] > 0) ? TNT_MASK_KEY_PADDING : TNT_MASK_NONE;
    // decode step with a score scratch (generation through the copy head): the step kernel dumps its scores and a small
    // kernel head-averages them, so the layer stays on the one-launch step path
    const bool step_weights = T == 1 && w.step_scores && attn_out && s == D->article_idx;
    ap.weights = (attn_out && s == D->article_idx && !step_weights) ? attn_out : nullptr;
    ap.lse = step_weights ? w.step_lse : w.lse[s];
    // the K/V buffer carries the 2 extra rows (project_kv), so they are plain keys the mask does not cover
    ap.B = B; ap.T = T; ap.S = ctx->S[s] + 2; ap.H = H; ap.D = d / H; ap.n_extra = 0;
    ap.mask_S = ctx->S[s];
    ap.kv_group = kv_group;
    ap.drop = drop_site(dr, dr ? dr->attention_dropout : 0.f, TNT_DROP_SITE(l, TNT_DROP_ATTN_PROB + s));
    static const bool tile_skip = [] { const char* e = getenv("TNT_TILE_SKIP"); return !(e && e[0] == '0'); }();
    ap.tile_live = (tile_skip && kv_group <= 1 && ap.mask_mode == TNT_MASK_KEY_PADDING) ? w.tile_live[s] : nullptr;
    if (ctx->S[s] == 0) ap.mask_mode = TNT_MASK_NONE;
    ap.scale = 1.0f / sqrtf(static_cast<float>(d / H));
    ap.dtype = dt;
    TNT_TRY(xattn_check(&ap));
    all_step = all_step && !xattn_force_simt() && xattn_step_supported(&ap);
  }
  if (!all_step && attn_out && D->article_idx >= 0) aps[D->article_idx].weights = attn_out;   // general kernels return the weights themselves
  if (all_step) {
    const int pid = prof_enabled() ? prof_begin("xattn_step.merged", st) : -1;   // -1 while a graph is being captured
    const int* orders[TNT_MAX_SOURCES];
    for (int s = 0; s < ns; ++s)
      orders[s] = (aps[s].mask_mode == TNT_MASK_KEY_PADDING && ctx->B <= 4096 && ring) ? w.order[s] : nullptr;
    const bool sw = w.step_scores && attn_out && D->article_idx >= 0;
    TNT_TRY(xattn_fwd_step_multi(aps, ns, st, w.sched, orders, sw ? w.step_scores : nullptr, sw ? D->article_idx : -1,
                                 sw ? attn_out : nullptr));   // decode step: every source of the layer in one launch
    prof_end(pid, st);
  } else {
    // the short sources of the layer (image regions, faces, objects: <= 128 keys) share ONE launch of the warp-level kernel
    tnt_xattn_params shorts[TNT_MAX_SOURCES];
    bool is_short[TNT_MAX_SOURCES] = {};
    int n_short = 0;
    for (int s = 0; s < ns; ++s) {
      is_short[s] = !xattn_force_simt() && xattn_short_supported(&aps[s]);
      if (is_short[s]) shorts[n_short++] = aps[s];
    }
    if (n_short) {
      const int pid = prof_enabled() ? prof_begin("xattn.short", st) : -1;
      TNT_TRY(xattn_short_multi(shorts, n_short, st));
      prof_end(pid, st);
    }
    for (int s = 0; s < ns; ++s) {
      if (is_short[s]) continue;
      const tnt_xattn_params& ap = aps[s];
      int pid = -1;
      if (prof_enabled()) {
        char pname[48];
        snprintf(pname, sizeof(pname), T == 1 ? "xattn_step.%s" : "xattn.%s", c.source_names[s]);
        pid = prof_begin(pname, st);
      }
      TNT_TRY(xattn_fwd_dispatch(&ap, st));
      prof_end(pid, st);
    }
  }
  if (fuse_ln) {
    TNT_TRY(linear_res_ln(w.AO, static_cast<long>(ns) * d, L.wo, d, L.bo, w.X1, d, w.CAT, w.CATln, static_cast<long>(ns) * d, N, d, d,
                          dt, L.ln_src_g, L.ln_src_b, ns, d, static_cast<long>(d) * d, d, d, L.wc, ns * dd, 0, 0, &d_attn, true, st));
  } else {
    TNT_TRY(linear(w.AO, static_cast<long>(ns) * d, L.wo, d, L.bo, w.CAT, static_cast<long>(ns) * d, N, d, d, dt, EPI_RESIDUAL,
                   w.X1, d, st, ns, d, static_cast<long>(d) * d, d, 0, d, L.wc, ns * dd, &d_attn));
    TNT_TRY(layernorm(w.CAT, w.CATln, d, L.ln_src_g, L.ln_src_b, N * ns, d, ns, dt, 0, 0, st));
  }
  TNT_TRY(linear(w.CATln, static_cast<long>(ns) * d, L.wc, static_cast<long>(ns) * d, L.bc, w.X0, d, N, d, ns * d, dt, EPI_NONE,
                 nullptr, 0, st, 1, 0, 0, 0, 0, 0, L.wf1, static_cast<size_t>(c.ffn) * d * E));
  // feed-forward
  TNT_TRY(linear(w.X0, d, L.wf1, d, L.bf1, w.F, c.ffn, N, c.ffn, d, dt, EPI_RELU, nullptr, 0, st, 1, 0, 0, 0, 0, 0, L.wf2,
                 static_cast<size_t>(c.ffn) * d * E, &d_relu));
  // (fc2 has K = ffn = 4096: a cluster CTA owns a 128 x 256 tile and walks 64 k-blocks, 31 us at B*T = 2048 where the
  // persistent GEMM's 128 narrower tiles + the LayerNorm kernel take 27 us; at a decode step's few rows the split-K form wins)
  TNT_TRY(linear_res_ln(w.F, c.ffn, L.wf2, c.ffn, L.bf2, w.X0, d, w.ybuf2, Xout, d, N, d, c.ffn, dt, L.ln_f_g, L.ln_f_b, 1, 0, 0, 0, 0,
                        next_w1, next_w1_bytes, transpose_out ? T : 0, transpose_out ? B : 0, &d_ffn,
                        fuse_ln && c.ffn <= 2048, st));
  return TNT_OK;
}

int embed(tnt_decoder* D, const int64_t* ids, int B, int T, void* out, const int* start_pos, cudaStream_t st) {
  tnt_embed_params e = {};
  e.ids = ids; e.table = D->table; e.pos_table = D->pos_table; e.out = out; e.ldo = D->cfg.d; e.start_pos = start_pos;
  e.B = B; e.T = T; e.d = D->cfg.d; e.n_pos = D->cfg.max_positions; e.pad_idx = D->cfg.pad_idx; e.dtype = D->cfg.dtype;
  e.vocab = D->cfg.vocab; e.error_flags = D->err_flags;
  return tnt_embed_fwd(&e, st);
}

}  // namespace tnt

using namespace tnt;

extern "C" void tnt_debug_sub_batches(int n) { tnt::g_sub_batches = n > 0 ? n : 0; }
extern "C" void tnt_debug_kv_beside(int sms) { tnt::g_kv_beside = sms; }
extern "C" void tnt_debug_kv_inline(int on) { tnt::g_kv_inline = on; }
// SMs the K/V projections' grids are sized for when the forward runs them beside the layer chain (0: that mode is off)
extern "C" int tnt_debug_kv_beside_get(void) {
  const char* e = getenv("TNT_KV_BESIDE_SMS");
  const int want = tnt::g_kv_beside >= 0 ? tnt::g_kv_beside : e ? atoi(e) : -1;
  if (want == 0) return 0;
  return std::max(2, std::min(want > 0 ? want : tnt::kKvBesideSms, tnt::num_sms())) & ~1;
}

extern "C" size_t tnt_decoder_workspace_bytes(const tnt_decoder* d, int B, int T, const int* S) {
  if (!d || !S) return 0;
  return carve_fwd(d, B, T, S, nullptr, true).total;
}

extern "C" int tnt_decoder_forward(tnt_decoder* D, const int64_t* ids, const tnt_contexts* ctx, void* X_out,
                                   float* attn_out, void* workspace, size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_forward: decoder not finalized");
  TNT_CHECK_ARG(ids && ctx && X_out && workspace, "decoder_forward: null pointer");
  const int B = ctx->B, T = ctx->T;
  TNT_CHECK_ARG(B > 0 && T > 0, "decoder_forward: empty batch");
  TNT_CHECK_ARG(T + D->cfg.pad_idx + 1 <= D->cfg.max_positions, "decoder_forward: T=%d exceeds the position table", T);
  TNT_CHECK_ALIGN((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "decoder_forward: workspace alignment");
  FwdWs w = carve_fwd(D, B, T, ctx->S, workspace, true);
  if (w.total > workspace_bytes) {
    set_error("decoder_forward: workspace %zu < %zu", workspace_bytes, w.total);
    return TNT_ERR_WORKSPACE;
  }
  cudaStream_t st = as_stream(stream);
  // sub-batches (engine.cuh): the layer chains of a few row ranges side by side, each on its own stream
  SubBatches sb;
  TNT_TRY(sub_plan(D, B, 1, 16, st, &sb));
  const int kv_sms = sb.n > 1 ? 0 : kv_beside_sms(D, ctx);
  const bool kv_inline = !kv_sms && sb.n == 1 && kv_inline_on(D, ctx);
  KvByLayer kv_bl;
  if (kv_sms) TNT_TRY(project_kv_beside(D, ctx, w, st, kv_sms));
  else if (kv_inline) TNT_TRY(kv_prologue(D, ctx, w, st, 16384, &kv_bl));
  else TNT_TRY(project_kv(D, ctx, w, st));
  const int Ls = D->cfg.n_layers;
  const size_t E = D->esz();
  TNT_TRY(sub_fork(D, sb));
  static const bool share_sms = [] { const char* e = getenv("TNT_SUB_SHARE_SMS"); return !(e && e[0] == '0'); }();
  SmBudget budget(sb.n > 1 && share_sms ? num_sms() / sb.n : 0);
  for (int i = 0; i < sb.n; ++i) {
    const FwdWs wi = sb.n > 1 ? sub_ws(D, w, sb, i, B, T, ctx->S) : w;
    const tnt_contexts ci = sb.n > 1 ? sub_ctx(*ctx, sb, i) : *ctx;
    const int b0 = sb.b0[i], nb = ci.B;
    TNT_TRY(embed(D, ids + static_cast<size_t>(b0) * T, nb, T, wi.X0, nullptr, sb.st[i]));
    // layer l reads `cur` and writes `out`; Xin == X0 is safe: X0 is only overwritten by context_fc after its last use
    // as residual.  The last layer writes batch-major rows straight into the caller's X (and attention weights).
    void* cur = wi.X0;
    for (int l = 0; l < Ls; ++l) {
      const bool last = l == Ls - 1;
      void* out = last ? static_cast<char*>(X_out) + static_cast<size_t>(b0) * T * D->cfg.d * E : wi.X0;
      float* ao = nullptr;
      if (last && attn_out && D->article_idx >= 0)
        ao = attn_out + static_cast<size_t>(b0) * T * (ctx->S[D->article_idx] + 2);
      if (kv_inline) TNT_TRY(kv_layer(D, ctx, w, l, st, kv_bl));
      TNT_TRY(run_layer(D, l, &ci, wi, cur, out, nb, T, nullptr, nullptr, ao, last, sb.st[i]));
      cur = out;
    }
  }
  TNT_TRY(sub_join(D, sb));
  return TNT_OK;
}

// ---------------------------------------------------------------------------------------------------------------------
// Step-wise protocol of the reference (`forward(prev, contexts, incremental_state)`, decoder_flattened.py:96-146 driven by
// `_generate`, transformer_flattened.py:250-275): the caller owns every piece of state -- the per-layer conv input buffers
// (dynamic.py:95-99), the position offset (positional.py:170-176) and, so that the memory projections are not redone every
// step as the reference does, the projected K/V of the contexts.
namespace tnt_inc {
__global__ void set_int_kernel(int* p, int v) { *p = v; }
}  // namespace tnt_inc

extern "C" size_t tnt_decoder_kv_cache_bytes(const tnt_decoder* d, int B, int S) {
  if (!d || B < 0 || S < 0) return 0;
  return static_cast<size_t>(S + 2) * B * d->cfg.n_layers * 2 * d->cfg.d * d->esz();
}

extern "C" int tnt_decoder_project_kv(tnt_decoder* D, const tnt_contexts* ctx, void* const* kv_out, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized && ctx && kv_out && ctx->B > 0, "decoder_project_kv: bad arguments");
  FwdWs w = {};
  for (int s = 0; s < D->cfg.n_sources; ++s) {
    TNT_CHECK_ARG(kv_out[s] && aligned16(kv_out[s]), "decoder_project_kv: K/V buffer of source %d missing / unaligned", s);
    w.kv[s] = kv_out[s];
  }
  return project_kv(D, ctx, w, as_stream(stream));
}

extern "C" size_t tnt_decoder_incremental_workspace_bytes(const tnt_decoder* d, int B, int T, const int* S) {
  if (!d || !S) return 0;
  return carve_fwd(d, B, T, S, nullptr, false).total + 256;
}

extern "C" int tnt_decoder_forward_incremental(tnt_decoder* D, const int64_t* ids, const tnt_contexts* ctx, void* const* kv,
                                               int start_pos, const void* const* hist_in, const int* hist_len,
                                               void* const* hist_out, void* X_out, float* attn_out, void* workspace,
                                               size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_forward_incremental: decoder not finalized");
  TNT_CHECK_ARG(ids && ctx && kv && X_out && workspace && hist_len, "decoder_forward_incremental: null pointer");
  const int B = ctx->B, T = ctx->T, Ls = D->cfg.n_layers;
  TNT_CHECK_ARG(B > 0 && T > 0 && start_pos >= 0, "decoder_forward_incremental: empty batch");
  TNT_CHECK_ARG(start_pos + T + D->cfg.pad_idx + 1 <= D->cfg.max_positions,
                "decoder_forward_incremental: position %d exceeds the position table", start_pos + T);
  TNT_CHECK_ALIGN((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "decoder_forward_incremental: workspace alignment");
  FwdWs w = carve_fwd(D, B, T, ctx->S, workspace, false);
  if (w.total + 256 > workspace_bytes) {
    set_error("decoder_forward_incremental: workspace %zu < %zu", workspace_bytes, w.total + 256);
    return TNT_ERR_WORKSPACE;
  }
  for (int s = 0; s < D->cfg.n_sources; ++s) {
    TNT_CHECK_ARG(kv[s], "decoder_forward_incremental: K/V cache of source %d missing (tnt_decoder_project_kv)", s);
    w.kv[s] = const_cast<void*>(kv[s]);
  }
  cudaStream_t st = as_stream(stream);
  int* pos_dev = reinterpret_cast<int*>(static_cast<char*>(workspace) + w.total);
  tnt_inc::set_int_kernel<<<1, 1, 0, st>>>(pos_dev, start_pos);
  TNT_LAUNCHED();
  TNT_TRY(embed(D, ids, B, T, w.X0, pos_dev, st));
  void* cur = w.X0;
  for (int l = 0; l < Ls; ++l) {
    const bool last = l == Ls - 1;
    void* out = last ? X_out : w.X0;
    ConvHist ch = {hist_in ? hist_in[l] : nullptr, hist_len[l], hist_out ? hist_out[l] : nullptr};
    TNT_TRY(run_layer(D, l, ctx, w, cur, out, B, T, nullptr, nullptr, last ? attn_out : nullptr, last, st, 1, nullptr, &ch));
    cur = out;
  }
  return TNT_OK;
}

extern "C" int tnt_decoder_embed(tnt_decoder* D, const int64_t* ids, int B, int T, void* out_tb, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized && ids && out_tb && B > 0 && T > 0, "decoder_embed: bad arguments");
  TNT_CHECK_ARG(T + D->cfg.pad_idx + 1 <= D->cfg.max_positions, "decoder_embed: T=%d exceeds the position table", T);
  return embed(D, ids, B, T, out_tb, nullptr, as_stream(stream));
}

extern "C" int tnt_decoder_error_flags(tnt_decoder* D, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->err_flags, "decoder_error_flags: decoder has no weight memory bound");
  int flags = 0;
  cudaStream_t st = as_stream(stream);
  TNT_CUDA(cudaMemcpyAsync(&flags, D->err_flags, sizeof(int), cudaMemcpyDeviceToHost, st));
  TNT_CUDA(cudaStreamSynchronize(st));
  if (flags) TNT_CUDA(cudaMemsetAsync(D->err_flags, 0, sizeof(int), st));
  return flags;
}

extern "C" int tnt_decoder_nll(tnt_decoder* D, const void* X, const int64_t* target, int n_rows, float* out2,
                               void* workspace, size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized && X && target && out2 && workspace, "decoder_nll: bad arguments");
  return adasoftmax_nll(D->ada, X, D->cfg.d, target, n_rows, D->cfg.pad_idx, out2, workspace, workspace_bytes,
                        as_stream(stream));
}

extern "C" int tnt_decoder_log_prob(tnt_decoder* D, const void* X, int n_rows, float* out, void* workspace,
                                    size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized && X && out && workspace, "decoder_log_prob: bad arguments");
  return adasoftmax_logprob(D->ada, X, D->cfg.d, n_rows, out, workspace, workspace_bytes, as_stream(stream));
}
