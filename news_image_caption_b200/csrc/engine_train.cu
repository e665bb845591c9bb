// Decoder engine: one training step = teacher-forced forward (activations kept) + AdaptiveLoss + backward
// through the loss, every decoder layer, the K/V projections and the embedding -- gradients land in the
// caller's fp32 buffers, one per reference state_dict key (tnt_decoder_set_grad).
// Dropout (tnt_decoder_set_dropout) is counter based: the backward recomputes every mask from (seed, site, index).
// What is differentiated:
//   DynamicConvDecoderLayer.forward  ttl/tell/models/decoder_flattened.py:271-345, decoder_faces_objects.py:283-393
//   MultiHeadAttention               ttl/tell/modules/attention/multi_head.py:288-486
//   GehringLinear weight_norm        ttl/tell/modules/linear.py:7-33
//   AdaptiveSoftmax / AdaptiveLoss   ttl/tell/modules/softmax.py:144-191, criteria/adaptive_loss.py:27-73
//   AdaptiveEmbedding                ttl/tell/modules/token_embedders/adaptive.py:61-76
// The exchange step of data-parallel training (gradient all-reduce) happens outside, over the flat gradient
// arena these buffers live in.
#include "backward.cuh"
#include "engine.cuh"

namespace tnt {

int xattn_check(const tnt_xattn_params* p);

namespace {
inline size_t al256(size_t x) { return (x + 255) & ~size_t(255); }

struct LayerAct {
  void *h, *logits, *cbuf, *yc, *X1, *Q, *AO, *CATpre, *CATln, *X0, *F, *yf, *Xout, *Xd;
  float* lse[TNT_MAX_SOURCES];
};

// Gradient tensors that a weight-gradient GEMM / bias column sum reads.  Those run on the parameter-gradient stream, beside
// the data-gradient chain of the SAME and the NEXT layer, so each tensor has a buffer of its own (never rewritten inside a
// layer) and the layers alternate between two sets: layer l's set is rewritten by layer l-2, which first waits for layer
// l's parameter-gradient work.
struct PGradBufs {
  void *gffn;                                  // [N, d]    dL/d(fc2 output), dropout mask applied
  void *dF;                                    // [N, ffn]  dL/d(fc1 output)
  void *dX0;                                   // [N, d]    dL/d(context_fc output)
  void *gop;                                   // [N, ns*d] dL/d(out_proj outputs), dropout mask applied
  void *dQ;                                    // [N, ns*d]
  void *gl2;                                   // [N, d]    dL/d(linear2 output), dropout mask applied
  void *dlogits;                               // [N, ldl]
  void *draw;                                  // [N, 2d]   dL/d(linear1 output)
};

struct TrainWs {
  void* emb;                                   // layer-0 input [N, d]
  LayerAct act[TNT_MAX_LAYERS];
  void* kv[TNT_MAX_SOURCES];
  void* dkv[TNT_MAX_SOURCES];
  uint32_t* tile_live[TNT_MAX_SOURCES];        // [B] live 128-key tiles per context entry
  void *Xbt, *dXbt;                            // decoder output / its gradient in B x T row order
  void *gA, *gB, *gC, *gD;                     // [N, d] gradient ping-pong buffers (data-gradient chain only)
  void *dAO, *dCAT, *dCATp;                    // [N, ns*d]
  void *raw;                                   // [N, 2d]
  PGradBufs pg[2];                             // what the parameter-gradient stream reads, by layer parity
  float* dW[tnt_decoder::kSideStreams];        // fp32 scratch for weight-normed matrices / K-V grads, one per side stream
  size_t dW_floats;
  float* attn_scr;                             // [B, H, T, D] fp32
  float* vecs[tnt_decoder::kSideStreams];      // [L*2d] fp32 scratch (K/V bias rows), one per side stream
  int *band_rows, *band_loc, *band_cnt, *band_head;
  void *G, *Hrows, *dH;                        // embedding backward [N, d]
  void* loss_fwd; size_t loss_fwd_bytes;
  void* loss_bwd; size_t loss_bwd_bytes;
  size_t total;
};

TrainWs carve_train(const tnt_decoder* D, int B, int T, const int* S, void* ws) {
  const tnt_decoder_config& c = D->cfg;
  const size_t E = D->esz(), N = static_cast<size_t>(B) * T, d = c.d, ns = c.n_sources, L = c.n_layers;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    void* p = ws ? static_cast<char*>(ws) + off : nullptr;
    off += al256(bytes);
    return p;
  };
  TrainWs w;
  w.emb = take(N * d * E);
  for (size_t l = 0; l < L; ++l) {
    LayerAct& a = w.act[l];
    a.h = take(N * d * E); a.logits = take(N * D->ld_logits(c.kernels[l]) * E); a.cbuf = take(N * d * E);
    a.yc = take(N * d * E); a.X1 = take(N * d * E); a.Q = take(N * ns * d * E); a.AO = take(N * ns * d * E);
    a.CATpre = take(N * ns * d * E); a.CATln = take(N * ns * d * E); a.X0 = take(N * d * E);
    a.F = take(N * c.ffn * E); a.yf = take(N * d * E); a.Xout = take(N * d * E); a.Xd = take(N * d * E);
    for (size_t s = 0; s < TNT_MAX_SOURCES; ++s)
      a.lse[s] = s < ns ? static_cast<float*>(take(N * c.heads * sizeof(float))) : nullptr;
  }
  size_t max_kdim = d;
  for (size_t s = 0; s < ns; ++s) {
    const size_t bytes = static_cast<size_t>(S[s] + 2) * B * L * 2 * d * E;
    w.kv[s] = take(bytes);
    w.dkv[s] = take(bytes);
    w.tile_live[s] = static_cast<uint32_t*>(take(sizeof(uint32_t) * B));
    max_kdim = std::max<size_t>(max_kdim, c.source_dim[s]);
  }
  w.Xbt = take(N * d * E); w.dXbt = take(N * d * E);
  w.gA = take(N * d * E); w.gB = take(N * d * E); w.gC = take(N * d * E); w.gD = take(N * d * E);
  w.dAO = take(N * ns * d * E); w.dCAT = take(N * ns * d * E); w.dCATp = take(N * ns * d * E);
  w.raw = take(N * 2 * d * E);
  for (PGradBufs& g : w.pg) {
    g.gffn = take(N * d * E); g.dF = take(N * c.ffn * E); g.dX0 = take(N * d * E);
    g.gop = take(N * ns * d * E); g.dQ = take(N * ns * d * E); g.gl2 = take(N * d * E);
    g.dlogits = take(N * D->max_ld_logits() * E); g.draw = take(N * 2 * d * E);
  }
  size_t dw = std::max<size_t>(2 * d * d, static_cast<size_t>(c.ffn) * d);
  dw = std::max(dw, d * ns * d);
  dw = std::max(dw, L * 2 * d * max_kdim);
  for (float*& p : w.dW) p = static_cast<float*>(take(dw * sizeof(float)));
  w.dW_floats = dw;
  w.attn_scr = static_cast<float*>(take(N * d * sizeof(float)));
  for (float*& p : w.vecs) p = static_cast<float*>(take(L * 2 * d * sizeof(float)));
  w.band_rows = static_cast<int*>(take(sizeof(int) * N * TNT_MAX_CLUSTERS));
  w.band_loc = static_cast<int*>(take(sizeof(int) * N * TNT_MAX_CLUSTERS));
  w.band_cnt = static_cast<int*>(take(sizeof(int) * TNT_MAX_CLUSTERS));
  w.band_head = static_cast<int*>(take(sizeof(int) * N));
  w.G = take(N * d * E); w.Hrows = take(N * d * E); w.dH = take(N * d * E);
  w.loss_fwd_bytes = adasoftmax_workspace_bytes(D->ada, static_cast<int>(N), false);
  w.loss_fwd = take(w.loss_fwd_bytes);
  w.loss_bwd_bytes = adasoftmax_bwd_workspace_bytes(D->ada, static_cast<int>(N));
  w.loss_bwd = take(w.loss_bwd_bytes);
  w.total = off;
  return w;
}

// ------------------------------------------------------------------------------ small call helpers
struct Ctx {
  tnt_decoder* D;
  cudaStream_t st;
  int dt;
  const tnt_dropout_config* dr;     // this step's dropout (seed already advanced) or NULL
  int n_side = 0;                   // parameter-gradient streams in use (0: everything on `st`)
  int job = 0;                      // side stream (and scratch slot) of the parameter-gradient job being issued
  bool side_pending = false;        // work issued on a side stream since the last join
  float* grad(const std::string& key) {
    auto it = D->grads.find(key);
    return it == D->grads.end() ? nullptr : it->second;
  }
  const float* raw(const std::string& key) {
    auto it = D->raw.find(key);
    return it == D->raw.end() ? nullptr : it->second.data;
  }
};

int gemm_on(cudaStream_t st, int dt_in, int layout, int epi, const void* A, long lda, const void* W, long ldw, void* C, long ldc,
            int M, int N, int K, int dt_out, const void* res = nullptr, long ldr = 0, const int* m_dyn = nullptr, float alpha = 0.f,
            const float* bias = nullptr, int groups = 1, long sa = 0, long sw = 0, long sr = 0, long sc = 0) {
  // (dropout never rides on a backward GEMM: masked gradients are materialised with dropout_apply)
  tnt_linear_params p = {};
  p.a = A; p.lda = lda; p.w = W; p.ldw = ldw; p.c = C; p.ldc = ldc; p.residual = res; p.ldr = ldr; p.bias = bias;
  p.M = M; p.N = N; p.K = K; p.m_dyn = m_dyn; p.groups = groups;
  p.stride_a = sa; p.stride_w = sw; p.stride_res = sr; p.stride_c = sc;
  p.dtype_in = dt_in; p.dtype_out = dt_out; p.epilogue = epi; p.layout = layout; p.alpha = alpha;
  return launch_gemm(&p, -1, nullptr, nullptr, st);
}
// data-gradient chain (the step's stream)
template <typename... Args>
int gemm(Ctx& c, Args... args) { return gemm_on(c.st, c.dt, args...); }

int ln_bwd(Ctx& c, const void* x, long ldx, const void* dy, long lddy, void* dx, long lddx, const float* gamma,
           float* dgamma, float* dbeta, int rows, int C) {
  TNT_CHECK_ARG(dgamma && dbeta, "train: LayerNorm gradient buffer missing");
  tnt_layernorm_bwd_params p = {};
  p.zero_param_grads = 1;
  p.x = x; p.ldx = ldx; p.dy = dy; p.lddy = lddy; p.dx = dx; p.lddx = lddx; p.gamma = gamma; p.dgamma = dgamma;
  p.dbeta = dbeta; p.rows = rows; p.C = C; p.eps = 1e-5f; p.dtype = c.dt;
  return tnt_layernorm_bwd(&p, c.st);
}

// Parameter-gradient streams.  The backward's critical path is the DATA-gradient chain (dX flows from kernel to kernel,
// every launch latency-bound at B*T = 2048 rows); the weight-gradient GEMMs, the weight-norm backward and the bias column
// sums only consume what that chain produces.  Each parameter is one JOB (weight-gradient GEMM -> weight-norm backward,
// bias column sum) issued on a side stream, round-robin over a few of them -- the jobs are as latency-bound as the chain,
// one stream alone fell 0.75 ms behind it:
//   side_fork  -- starts a job: picks the next side stream, which then sees everything issued on `st` so far (call it right
//                 after the producer of the tensor the job reads, before the next chain kernel is issued);
//   side_join  -- `st` waits for all side work issued so far.
// What the jobs read is never rewritten inside a layer (PGradBufs); the cross-layer reuse is ordered by
// tnt_decoder::side_layer_done (layer_bwd).  Scratch (w.dW / w.vecs) is per side stream, so it is ordered in-stream.
int side_fork(Ctx& c) {
  if (!c.n_side) return TNT_OK;
  c.job = (c.job + 1) % c.n_side;
  TNT_CUDA(cudaEventRecord(c.D->side_fork, c.st));
  TNT_CUDA(cudaStreamWaitEvent(c.D->side_streams[c.job], c.D->side_fork, 0));
  c.side_pending = true;
  return TNT_OK;
}
int side_join(Ctx& c) {
  if (!c.n_side || !c.side_pending) return TNT_OK;
  for (int k = 0; k < c.n_side; ++k) {
    TNT_CUDA(cudaEventRecord(c.D->side_join[k], c.D->side_streams[k]));
    TNT_CUDA(cudaStreamWaitEvent(c.st, c.D->side_join[k], 0));
  }
  c.side_pending = false;
  return TNT_OK;
}
inline cudaStream_t pstream(Ctx& c) { return c.n_side ? c.D->side_streams[c.job] : c.st; }
// weight-gradient GEMM (TN: reduction over the tokens) on the parameter-gradient stream
// The grids (and tile shapes) of the side-stream GEMMs are sized for 64 SMs, so that a chain GEMM of 128 CTAs finds most
// of its SMs free; the dynamic tile scheduler lets either kernel absorb what the other leaves.  Measured at B = 64:
// all SMs 6.56 ms per step, 96: 6.51, 64: 6.38, 32: 6.50 (TNT_SIDE_SM_BUDGET = n, 0 = all)
inline int side_sm_budget() {
  static const int n = [] { const char* e = getenv("TNT_SIDE_SM_BUDGET"); return e ? atoi(e) : 64; }();
  return n;
}
template <typename... Args>
int wgemm(Ctx& c, Args... args) {
  SmBudget budget(c.n_side ? side_sm_budget() : 0);
  return gemm_on(pstream(c), c.dt, TNT_TN, EPI_NONE, args...);
}

// gradient of a GehringLinear: dWp [rows(prepared order), cols] fp32 -> weight_v / weight_g (or plain weight)
int linear_param_grads(Ctx& c, const std::string& prefix, const float* dWp, int rows, int cols, bool glu) {
  const float* g = c.raw(prefix + "weight_g");
  if (g) {
    const float* v = c.raw(prefix + "weight_v");
    float *dv = c.grad(prefix + "weight_v"), *dg = c.grad(prefix + "weight_g");
    TNT_CHECK_ARG(v && dv && dg, "train: '%sweight_v/_g' parameter or gradient buffer missing", prefix.c_str());
    return wn_bwd(v, g, dWp, cols, dv, dg, rows, cols, glu, pstream(c));
  }
  const float* v = c.raw(prefix + "weight");
  float* dv = c.grad(prefix + "weight");
  TNT_CHECK_ARG(v && dv, "train: '%sweight' parameter or gradient buffer missing", prefix.c_str());
  return wn_bwd(v, nullptr, dWp, cols, dv, nullptr, rows, cols, glu, pstream(c));
}

int bias_grad(Ctx& c, const std::string& key, const void* dy, long ld, int M, int N, bool deinterleave, long offset = 0) {
  float* g = c.grad(key);
  TNT_CHECK_ARG(g, "train: gradient buffer for '%s' missing", key.c_str());
  return colsum(dy, ld, M, N, nullptr, g + offset, c.dt, deinterleave, false, pstream(c));
}

inline const char* off(const void* p, size_t elems, size_t E) { return static_cast<const char*>(p) + elems * E; }
inline char* off(void* p, size_t elems, size_t E) { return static_cast<char*>(p) + elems * E; }


// ------------------------------------------------------------------------------ forward (activations kept)
FwdWs layer_ws(const TrainWs& w, int l, int ns) {
  FwdWs f = {};
  const LayerAct& a = w.act[l];
  f.X0 = a.X0; f.X1 = a.X1; f.h = a.h; f.logits = a.logits; f.cbuf = a.cbuf; f.ybuf = a.yc; f.ybuf2 = a.yf;
  f.Q = a.Q; f.AO = a.AO; f.CAT = a.CATpre; f.CATln = a.CATln; f.F = a.F; f.Xd = a.Xd;
  for (int s = 0; s < ns; ++s) {
    f.kv[s] = w.kv[s];
    f.lse[s] = a.lse[s];
    f.tile_live[s] = w.tile_live[s];
  }
  return f;
}

tnt_xattn_params attn_params(const tnt_decoder* D, const tnt_contexts* ctx, const TrainWs& w, int l, int s,
                             const tnt_dropout_config* dr) {
  const tnt_decoder_config& c = D->cfg;
  const int d = c.d, ns = c.n_sources, B = ctx->B, T = ctx->T;
  const size_t E = D->esz();
  const long ldkv = static_cast<long>(c.n_layers) * 2 * d;
  const LayerAct& a = w.act[l];
  tnt_xattn_params ap = {};
  ap.q = off(a.Q, static_cast<size_t>(s) * d, E);
  ap.q_stride_t = static_cast<long>(B) * ns * d; ap.q_stride_b = static_cast<long>(ns) * d;
  ap.k = off(w.kv[s], static_cast<size_t>(l) * 2 * d, E);
  ap.k_stride_s = static_cast<long>(B) * ldkv; ap.k_stride_b = ldkv;
  ap.v = off(ap.k, d, E);
  ap.v_stride_s = ap.k_stride_s; ap.v_stride_b = ap.k_stride_b;
  ap.out = off(a.AO, static_cast<size_t>(s) * d, E);
  ap.o_stride_t = ap.q_stride_t; ap.o_stride_b = ap.q_stride_b;
  ap.mask = ctx->pad_mask[s]; ap.m_stride_b = ctx->S[s]; ap.m_stride_t = 0;
  ap.mask_mode = (ctx->pad_mask[s] && ctx->S[s] > 0) ? TNT_MASK_KEY_PADDING : TNT_MASK_NONE;
  ap.B = B; ap.T = T; ap.S = ctx->S[s] + 2; ap.H = c.heads; ap.D = d / c.heads; ap.n_extra = 0;
  ap.mask_S = ctx->S[s];
  ap.scale = 1.0f / sqrtf(static_cast<float>(d / c.heads));
  ap.dtype = c.dtype;
  ap.lse = a.lse[s];
  ap.drop = drop_site(dr, dr ? dr->attention_dropout : 0.f, TNT_DROP_SITE(l, TNT_DROP_ATTN_PROB + s));
  ap.tile_live = (ap.mask_mode == TNT_MASK_KEY_PADDING) ? w.tile_live[s] : nullptr;
  return ap;
}

int kv_param_grads(Ctx& c, const tnt_contexts* ctx, TrainWs& w, int l, int s);

// ------------------------------------------------------------------------------ one layer, backward
// in: gA = dL/dXout [N, d] (T x B order);  out: gA = dL/dXin
// `dweights` (last layer only, copy heads): gradient w.r.t. the head-averaged attention weights the forward returned
// Data-gradient chain on c.st; after each tensor a parameter gradient needs is produced, side_fork() and that parameter's
// weight-gradient GEMM (+ weight-norm backward) and bias column sum go to the parameter-gradient stream.
int layer_bwd(Ctx& c, const tnt_contexts* ctx, TrainWs& w, int l, const void* Xin, const float* dweights = nullptr) {
  tnt_decoder* D = c.D;
  const tnt_decoder_config& cf = D->cfg;
  const LayerW& L = D->layers[l];
  const LayerAct& a = w.act[l];
  const PGradBufs& G = w.pg[l & 1];
  const int d = cf.d, ns = cf.n_sources, B = ctx->B, T = ctx->T, N = B * T, H = cf.heads, dt = c.dt, ffn = cf.ffn;
  const int HK = H * L.K, ldl = D->ld_logits(L.K);
  const long nsd = static_cast<long>(ns) * d;
  const size_t E = D->esz();
  const std::string p = "layers." + std::to_string(l) + ".";
  const tnt_dropout_config* dr = c.dr;
  const float p_out = dr ? dr->dropout : 0.f;
  // this layer rewrites the PGradBufs set layer l+2 used: its parameter-gradient work must have read them
  if (l + 2 < cf.n_layers)
    for (int k = 0; k < c.n_side; ++k)
      TNT_CUDA(cudaStreamWaitEvent(c.st, D->side_layer_done[(l + 2) * tnt_decoder::kSideStreams + k], 0));
  // ---- final LayerNorm and feed-forward: Xout = LN(yf), yf = drop(fc2(drop_relu(relu(fc1(X0))))) + X0
  void* dyf = p_out > 0.f ? w.gB : G.gffn;        // gradient w.r.t. yf; G.gffn = the same with the fc2 dropout mask applied
  TNT_TRY(ln_bwd(c, a.yf, d, w.gA, d, dyf, d, L.ln_f_g, c.grad(p + "final_layer_norm.weight"),
                 c.grad(p + "final_layer_norm.bias"), N, d));
  if (p_out > 0.f)
    TNT_TRY(dropout_apply(dyf, d, nullptr, 0, G.gffn, d, N, d, dt, drop_site(dr, p_out, TNT_DROP_SITE(l, TNT_DROP_FFN_OUT)), c.st));
  TNT_TRY(side_fork(c));
  TNT_TRY(wgemm(c, G.gffn, d, a.F, ffn, w.dW[c.job], ffn, d, ffn, N, TNT_F32));
  TNT_TRY(linear_param_grads(c, p + "fc2.", w.dW[c.job], d, ffn, false));
  TNT_TRY(bias_grad(c, p + "fc2.bias", G.gffn, d, N, d, false));
  // a.F holds the dropped-out ReLU output: F > 0 iff kept and active, and the kept elements carry 1 / (1 - p_relu)
  const float relu_scale = (dr && dr->relu_dropout > 0.f) ? 1.f / (1.f - dr->relu_dropout) : 0.f;
  TNT_TRY(gemm(c, TNT_NN, EPI_RELU_BWD, G.gffn, d, L.wf2, ffn, G.dF, ffn, N, ffn, d, dt, a.F, ffn, nullptr, relu_scale));
  TNT_TRY(side_fork(c));
  TNT_TRY(wgemm(c, G.dF, ffn, a.X0, d, w.dW[c.job], d, ffn, d, N, TNT_F32));
  TNT_TRY(linear_param_grads(c, p + "fc1.", w.dW[c.job], ffn, d, false));
  TNT_TRY(bias_grad(c, p + "fc1.bias", G.dF, ffn, N, ffn, false));
  TNT_TRY(gemm(c, TNT_NN, EPI_RESIDUAL, G.dF, ffn, L.wf1, d, G.dX0, d, N, d, ffn, dt, dyf, d));       // dX0
  // ---- context_fc over the concatenated, layer-normed per-source branches
  TNT_TRY(side_fork(c));
  TNT_TRY(wgemm(c, G.dX0, d, a.CATln, nsd, w.dW[c.job], nsd, d, ns * d, N, TNT_F32));
  TNT_TRY(linear_param_grads(c, p + "context_fc.", w.dW[c.job], d, ns * d, false));
  TNT_TRY(bias_grad(c, p + "context_fc.bias", G.dX0, d, N, d, false));
  TNT_TRY(gemm(c, TNT_NN, EPI_NONE, G.dX0, d, L.wc, nsd, w.dCAT, nsd, N, ns * d, d, dt));
  void* dcatp = p_out > 0.f ? w.dCATp : G.gop;    // gradient w.r.t. CATpre; G.gop = the same with the out_proj dropout mask
  for (int s = 0; s < ns; ++s) {
    const std::string ln = p + "context_attn_lns." + cf.source_names[s] + ".";
    TNT_TRY(ln_bwd(c, off(a.CATpre, static_cast<size_t>(s) * d, E), nsd, off(w.dCAT, static_cast<size_t>(s) * d, E), nsd,
                   off(dcatp, static_cast<size_t>(s) * d, E), nsd, L.ln_src_g + s * d, c.grad(ln + "weight"),
                   c.grad(ln + "bias"), N, d));
  }
  // ---- per-source attention branch: CATpre_s = drop(out_proj_s(attn_s)) + X1
  if (p_out > 0.f)
    TNT_TRY(dropout_apply(dcatp, nsd, nullptr, 0, G.gop, nsd, N, ns * d, dt,
                          drop_site(dr, p_out, TNT_DROP_SITE(l, TNT_DROP_ATTN_OUT), nsd), c.st));
  for (int s = 0; s < ns; ++s) {
    const std::string at = p + "context_attns." + cf.source_names[s] + ".";
    float* dwo = c.grad(at + "out_proj.weight");
    TNT_CHECK_ARG(dwo, "train: gradient buffer for '%sout_proj.weight' missing", at.c_str());
    TNT_TRY(side_fork(c));
    TNT_TRY(wgemm(c, off(G.gop, static_cast<size_t>(s) * d, E), nsd, off(a.AO, static_cast<size_t>(s) * d, E), nsd, dwo, d, d, d, N,
                  TNT_F32));
    TNT_TRY(bias_grad(c, at + "out_proj.bias", off(G.gop, static_cast<size_t>(s) * d, E), nsd, N, d, false));
  }
  TNT_TRY(gemm(c, TNT_NN, EPI_NONE, G.gop, nsd, L.wo, d, w.dAO, nsd, N, d, d, dt, nullptr, 0, nullptr, 0.f, nullptr, ns,
               d, static_cast<long>(d) * d, 0, d));
  for (int s = 0; s < ns; ++s) {
    const std::string at = p + "context_attns." + cf.source_names[s] + ".";
    XattnBwd xb;
    xb.f = attn_params(D, ctx, w, l, s, dr);
    xb.dout = off(w.dAO, static_cast<size_t>(s) * d, E);
    xb.dq = off(G.dQ, static_cast<size_t>(s) * d, E);
    xb.dk = off(w.dkv[s], static_cast<size_t>(l) * 2 * d, E);
    xb.dv = off(xb.dk, d, E);
    xb.scratch = w.attn_scr;
    xb.dweights = (dweights && s == D->article_idx) ? dweights : nullptr;
    TNT_TRY(xattn_bwd(&xb, c.st));
    // query projection parameters of this source: q_proj_weight or rows [0, d) of in_proj_weight
    float* dwq = c.grad(at + "in_proj_weight");
    if (!dwq) dwq = c.grad(at + "q_proj_weight");
    TNT_CHECK_ARG(dwq, "train: gradient buffer for the query projection of '%s' missing", at.c_str());
    TNT_TRY(side_fork(c));
    TNT_TRY(wgemm(c, off(G.dQ, static_cast<size_t>(s) * d, E), nsd, a.X1, d, dwq, d, d, d, N, TNT_F32));
    TNT_TRY(bias_grad(c, at + "in_proj_bias", off(G.dQ, static_cast<size_t>(s) * d, E), nsd, N, d, false));
    // K/V projection parameters of this layer and source: dkv's columns of layer l are final
    TNT_TRY(kv_param_grads(c, ctx, w, l, s));
  }
  // dX1 = dQ . Wq + sum_s dCATpre_s
  TNT_TRY(gemm(c, TNT_NN, EPI_RESIDUAL, G.dQ, nsd, L.wq, d, w.gD, d, N, d, ns * d, dt, dcatp, nsd));
  TNT_TRY(add_col_blocks(w.gD, d, dcatp, nsd, N, d, 1, ns, dt, c.st));
  // ---- conv block: X1 = LN(yc), yc = linear2(conv(h)) + Xin, h = GLU(linear1(Xin))
  void* dyc = p_out > 0.f ? w.gB : G.gl2;         // gradient w.r.t. yc; G.gl2 = the same with the linear2 dropout mask
  TNT_TRY(ln_bwd(c, a.yc, d, w.gD, d, dyc, d, L.ln_conv_g, c.grad(p + "conv_layer_norm.weight"),
                 c.grad(p + "conv_layer_norm.bias"), N, d));
  if (p_out > 0.f)
    TNT_TRY(dropout_apply(dyc, d, nullptr, 0, G.gl2, d, N, d, dt, drop_site(dr, p_out, TNT_DROP_SITE(l, TNT_DROP_CONV_OUT)), c.st));
  TNT_TRY(side_fork(c));
  TNT_TRY(wgemm(c, G.gl2, d, a.cbuf, d, w.dW[c.job], d, d, d, N, TNT_F32));
  TNT_TRY(linear_param_grads(c, p + "linear2.", w.dW[c.job], d, d, false));
  TNT_TRY(bias_grad(c, p + "linear2.bias", G.gl2, d, N, d, false));
  TNT_TRY(gemm(c, TNT_NN, EPI_NONE, G.gl2, d, L.w2, d, w.gC, d, N, d, d, dt));                        // gC = dcbuf
  {
    tnt_dynconv_bwd_params bp = {};
    bp.x = a.h; bp.ldx = d; bp.logits = a.logits; bp.ldl = ldl; bp.dy = w.gC; bp.lddy = d; bp.dx = w.gD; bp.lddx = d;
    bp.dlogits = G.dlogits; bp.lddl = ldl; bp.T = T; bp.B = B; bp.C = d; bp.H = H; bp.K = L.K; bp.dtype = dt;
    bp.drop = drop_site(dr, dr ? dr->weight_dropout : 0.f, TNT_DROP_SITE(l, TNT_DROP_TAPS));
    TNT_TRY(tnt_dynconv_bwd(&bp, c.st));                                                             // gD = dh (value path)
  }
  {
    float* dwt = c.grad(p + "conv.weight_linear.weight");
    TNT_CHECK_ARG(dwt, "train: gradient buffer for '%sconv.weight_linear.weight' missing", p.c_str());
    TNT_TRY(side_fork(c));
    TNT_TRY(wgemm(c, G.dlogits, ldl, a.h, d, dwt, d, HK, d, N, TNT_F32));
  }
  TNT_TRY(gemm(c, TNT_NN, EPI_RESIDUAL, G.dlogits, ldl, L.wtap, d, w.gC, d, N, d, HK, dt, w.gD, d));   // gC = dh
  const bool in_drop = dr && dr->input_dropout > 0.f;
  const void* X1in = in_drop ? a.Xd : Xin;        // what linear1 read in the forward
  TNT_TRY(gemm(c, TNT_NT, EPI_NONE, X1in, d, L.w1, d, w.raw, 2 * d, N, 2 * d, d, dt, nullptr, 0, nullptr, 0.f, L.b1));
  TNT_TRY(glu_bwd(w.raw, 2 * d, w.gC, d, G.draw, 2 * d, N, d, dt, c.st));
  TNT_TRY(side_fork(c));
  TNT_TRY(wgemm(c, G.draw, 2 * d, X1in, d, w.dW[c.job], d, 2 * d, d, N, TNT_F32));
  TNT_TRY(linear_param_grads(c, p + "linear1.", w.dW[c.job], 2 * d, d, true));
  TNT_TRY(bias_grad(c, p + "linear1.bias", G.draw, 2 * d, N, 2 * d, true));
  if (in_drop) {                                  // gA = dXin = drop(draw . W1) + dyc
    TNT_TRY(gemm(c, TNT_NN, EPI_NONE, G.draw, 2 * d, L.w1, d, w.gC, d, N, d, 2 * d, dt));
    TNT_TRY(dropout_apply(w.gC, d, dyc, d, w.gA, d, N, d, dt, drop_site(dr, dr->input_dropout, TNT_DROP_SITE(l, TNT_DROP_INPUT)), c.st));
  } else {
    TNT_TRY(gemm(c, TNT_NN, EPI_RESIDUAL, G.draw, 2 * d, L.w1, d, w.gA, d, N, d, 2 * d, dt, dyc, d));    // gA = dXin
  }
  return TNT_OK;
}


// ------------------------------------------------------------------------------ K/V projection parameters
// of layer l, source s (k/v_proj_weight or in_proj_weight rows [d, 3d), in_proj_bias[d:], bias_k, bias_v): issued on the
// parameter-gradient stream as soon as the layer's attention backward has written dkv's columns of layer l.
int kv_param_grads(Ctx& c, const tnt_contexts* ctx, TrainWs& w, int l, int s) {
  tnt_decoder* D = c.D;
  const tnt_decoder_config& cf = D->cfg;
  const int d = cf.d, kd = cf.source_dim[s], B = ctx->B, S = ctx->S[s];
  const int ldkv = cf.n_layers * 2 * d;
  const size_t E = D->esz();
  const std::string at = "layers." + std::to_string(l) + ".context_attns." + cf.source_names[s] + ".";
  const void* dkv = off(w.dkv[s], static_cast<size_t>(l) * 2 * d, E);     // columns [K | V] of this layer
  float *gk = c.grad(at + "bias_k"), *gv = c.grad(at + "bias_v"), *gb = c.grad(at + "in_proj_bias");
  TNT_CHECK_ARG(gk && gv && gb, "train: gradient buffer for '%sbias_k / bias_v / in_proj_bias' missing", at.c_str());
  TNT_TRY(side_fork(c));
  cudaStream_t ss = pstream(c);
  TNT_TRY(kv_extra_grad(dkv, ldkv, S, B, 2 * d, w.vecs[c.job], c.dt, ss));        // the learned extra key row (index S)
  TNT_CUDA(cudaMemcpyAsync(gk, w.vecs[c.job], sizeof(float) * d, cudaMemcpyDeviceToDevice, ss));
  TNT_CUDA(cudaMemcpyAsync(gv, w.vecs[c.job] + d, sizeof(float) * d, cudaMemcpyDeviceToDevice, ss));
  TNT_TRY(colsum(dkv, ldkv, S * B, 2 * d, nullptr, gb + d, c.dt, false, false, ss));
  float* gi = c.grad(at + "in_proj_weight");
  float* dst = gi ? gi + static_cast<size_t>(d) * d : w.dW[c.job];
  // A long reduction (the article: S*B = 32768 rows) over only (2d/128) x (kd/256) output tiles leaves most SMs idle, and
  // narrower tiles are L2->SM bound (128 x 128 tiles: 213 us = 46 % of the tensor peak).  The rows are split into G
  // groups of one grouped launch -- wide tiles on every SM -- whose fp32 partials are summed afterwards.
  int G = 1;
  const int Kr = S * B;
  if (Kr >= 8192 && c.dt == TNT_BF16) {
    const long tiles = static_cast<long>((2 * d + 127) / 128) * ((kd + 255) / 256);
    while (G < 8 && tiles * G * 4 < 3L * num_sms() && Kr % (2 * G) == 0 && Kr / (2 * G) >= 2048) G *= 2;
    const size_t need = (static_cast<size_t>(G) + (gi ? 0 : 1)) * 2 * d * kd;
    if (need > w.dW_floats || (static_cast<size_t>(2) * d * kd) % 4) G = 1;
  }
  if (G > 1) {
    float* parts = w.dW[c.job] + (gi ? 0 : static_cast<size_t>(2) * d * kd);
    TNT_TRY(wgemm(c, dkv, ldkv, ctx->memory[s], kd, parts, kd, 2 * d, kd, Kr / G, TNT_F32, nullptr, 0, nullptr, 0.f,
                  nullptr, G, static_cast<long>(Kr / G) * ldkv, static_cast<long>(Kr / G) * kd, 0, static_cast<long>(2) * d * kd));
    TNT_TRY(sum_partials(dst, parts, G, static_cast<long>(2) * d * kd, ss));
  } else if (S * B > 0) {
    TNT_TRY(wgemm(c, dkv, ldkv, ctx->memory[s], kd, dst, kd, 2 * d, kd, S * B, TNT_F32));
  } else {
    TNT_CUDA(cudaMemsetAsync(dst, 0, sizeof(float) * 2 * d * kd, ss));
  }
  if (!gi) {
    float *gwk = c.grad(at + "k_proj_weight"), *gwv = c.grad(at + "v_proj_weight");
    TNT_CHECK_ARG(gwk && gwv, "train: gradient buffer for '%sk/v_proj_weight' missing", at.c_str());
    TNT_CUDA(cudaMemcpyAsync(gwk, w.dW[c.job], sizeof(float) * d * kd, cudaMemcpyDeviceToDevice, ss));
    TNT_CUDA(cudaMemcpyAsync(gwv, w.dW[c.job] + static_cast<size_t>(d) * kd, sizeof(float) * d * kd, cudaMemcpyDeviceToDevice, ss));
  }
  return TNT_OK;
}

// ------------------------------------------------------------------------------ embedding bands
// out = table_i[id - lo_i] . proj_i^T * sqrt(d) + position (adaptive.py:61-76); gE = dL/d(out) [N, d] in T x B order.
// Band tables are tied to the softmax output rows: their gradient buffers already hold the softmax part.
int embed_bwd(Ctx& c, const int64_t* ids, int B, int T, TrainWs& w, const void* gE) {
  tnt_decoder* D = c.D;
  const tnt_decoder_config& cf = D->cfg;
  const int d = cf.d, N = B * T, nb = cf.n_clusters;
  const size_t E = D->esz();
  int edges[TNT_MAX_CLUSTERS + 1];
  edges[0] = 0;
  for (int i = 0; i < nb; ++i) edges[i + 1] = cf.cutoffs[i];
  TNT_CHECK_ARG(nb + 1 <= TNT_MAX_CLUSTERS, "train: embedding backward supports at most %d vocabulary bands", TNT_MAX_CLUSTERS - 1);
  TNT_TRY(band_rows(ids, N, edges, nb, w.band_head, w.band_rows, w.band_loc, w.band_cnt, c.st));
  const float sq = sqrtf(static_cast<float>(d));
  for (int i = 0; i < nb; ++i) {
    const std::string e = "embedder.token_embedder_adaptive.embeddings." + std::to_string(i);
    const int di = D->emb_dim[i];
    const float* table = c.raw(e + ".0.weight");
    float *gt = c.grad(e + ".0.weight"), *gp = c.grad(e + ".1.weight");
    TNT_CHECK_ARG(table && gt && gp, "train: embedding band %d parameter or gradient buffer missing", i);
    const int* rows = w.band_rows + static_cast<long>(i) * N;
    const int* loc = w.band_loc + static_cast<long>(i) * N;
    const int* cnt = w.band_cnt + i;
    TNT_CUDA(cudaMemsetAsync(w.G, 0, E * N * d, c.st));       // pad rows of the 64-row reduction blocks must be 0
    TNT_CUDA(cudaMemsetAsync(w.Hrows, 0, E * N * d, c.st));
    TNT_TRY(gather_rows(gE, d, rows, cnt, N, w.G, d, d, c.dt, B, T, c.st));
    TNT_TRY(gather_table_rows(table, di, loc, cnt, N, w.Hrows, di, di, c.dt, c.st));
    // d proj_i [d, dim_i] = sqrt(d) G^T . H
    TNT_TRY(gemm(c, TNT_TN, EPI_NONE, w.G, d, w.Hrows, di, gp, di, d, di, N, TNT_F32, nullptr, 0, cnt, sq));
    // d H [rows, dim_i] = sqrt(d) G . proj_i, scattered into the band table (padding row 0 skipped)
    TNT_TRY(gemm(c, TNT_NN, EPI_NONE, w.G, d, D->emb_proj[i], di, w.dH, di, N, di, d, c.dt, nullptr, 0, cnt, sq));
    if (D->sparse_rows) {
      // data-parallel training: the rows join the packed list the ranks all-gather (tnt_decoder_set_sparse_embed_grad); the
      // dense band-table gradient then holds the softmax part only and is final at stage 0
      TNT_CHECK_ARG(di == d && N <= D->sparse_capacity, "train: sparse embedding gradient needs band dim == d and capacity >= B*T");
      TNT_TRY(pack_band_rows(w.dH, di, loc, w.band_cnt, i, nb, N, D->sparse_rows, d, D->sparse_meta, D->sparse_count, di, c.dt, c.st));
      continue;
    }
    // every band is an nn.Embedding(band_size, dim, padding_idx): its LOCAL row padding_idx gets no gradient (adaptive.py:42)
    const int skip = D->cfg.embed_padding_idx >= 0 ? D->cfg.embed_padding_idx : -1;
    TNT_TRY(scatter_table_grad(w.dH, di, loc, cnt, N, gt, di, di, c.dt, skip, c.st));
  }
  return TNT_OK;
}

}  // namespace
}  // namespace tnt

using namespace tnt;

extern "C" size_t tnt_decoder_train_workspace_bytes(const tnt_decoder* d, int B, int T, const int* S) {
  if (!d || !S) return 0;
  return carve_train(d, B, T, S, nullptr).total;
}

extern "C" int tnt_decoder_set_dropout(tnt_decoder* d, const tnt_dropout_config* cfg) {
  TNT_CHECK_ARG(d, "decoder_set_dropout: null decoder");
  tnt_dropout_config z = {};
  const tnt_dropout_config& c = cfg ? *cfg : z;
  const float ps[5] = {c.dropout, c.input_dropout, c.weight_dropout, c.attention_dropout, c.relu_dropout};
  for (float p : ps) TNT_CHECK_ARG(p >= 0.f && p < 1.f, "decoder_set_dropout: probability %f outside [0, 1)", p);
  d->drop_cfg = c;
  d->drop_step = 0;
  return TNT_OK;
}

extern "C" int tnt_decoder_set_sparse_embed_grad(tnt_decoder* d, void* rows, int64_t* meta, int* count, int capacity_rows) {
  TNT_CHECK_ARG(d, "decoder_set_sparse_embed_grad: null decoder");
  TNT_CHECK_ARG(!rows || (meta && count && capacity_rows > 0 && aligned16(rows)), "decoder_set_sparse_embed_grad: bad buffers");
  d->sparse_rows = rows;
  d->sparse_meta = reinterpret_cast<long long*>(meta);
  d->sparse_count = count;
  d->sparse_capacity = rows ? capacity_rows : 0;
  return TNT_OK;
}

extern "C" int tnt_decoder_set_grad(tnt_decoder* d, const char* key, float* grad) {
  TNT_CHECK_ARG(d && key, "decoder_set_grad: bad arguments");
  if (grad) d->grads[key] = grad;
  else d->grads.erase(key);
  return TNT_OK;
}

static int g_train_side = -1;   // test hook (tnt_debug_train_side_streams): >= 0 overrides TNT_TRAIN_SIDE_STREAMS for decoders
                                // that have not stepped yet

namespace {
struct StepSetup {
  TrainWs w;
  tnt_dropout_config drc;
  const tnt_dropout_config* dr;
};

int step_setup(tnt_decoder* D, const tnt_contexts* ctx, void* workspace, size_t workspace_bytes, const char* who, StepSetup* s) {
  const tnt_decoder_config& cf = D->cfg;
  const int B = ctx->B, T = ctx->T, Ls = cf.n_layers;
  TNT_CHECK_ARG(B > 0 && T > 0, "%s: empty batch", who);
  TNT_CHECK_ARG(T + cf.pad_idx + 1 <= cf.max_positions, "%s: T=%d exceeds the position table", who, T);
  TNT_CHECK_ALIGN((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "%s: workspace alignment", who);
  s->w = carve_train(D, B, T, ctx->S, workspace);
  if (s->w.total > workspace_bytes) {
    set_error("%s: workspace %zu < %zu", who, workspace_bytes, s->w.total);
    return TNT_ERR_WORKSPACE;
  }
  // TNT_TRAIN_SIDE_STREAMS = 0..4 parameter-gradient streams (default 3; TNT_TRAIN_SIDE_STREAM=0 also switches them off)
  static const int want_side = [] {
    const char* off1 = getenv("TNT_TRAIN_SIDE_STREAM");
    if (off1 && atoi(off1) == 0) return 0;
    const char* n = getenv("TNT_TRAIN_SIDE_STREAMS");
    return n ? std::max(0, std::min(atoi(n), static_cast<int>(tnt_decoder::kSideStreams))) : 3;
  }();
  const int want = g_train_side >= 0 ? std::min(g_train_side, static_cast<int>(tnt_decoder::kSideStreams)) : want_side;
  if (want && !D->side_fork) {
    int lo = 0, hi = 0;
    TNT_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    // measured (B = 64, one GPU): with ONE side stream the highest priority was better (7.05 vs 7.12 ms per step), with three
    // the lowest is (6.57 vs 6.62 ms): the chain is the critical path, freed SMs go to its next kernel first
    const char* pe = getenv("TNT_SIDE_PRIO");
    for (int k = 0; k < want; ++k)
      TNT_CUDA(cudaStreamCreateWithPriority(&D->side_streams[k], cudaStreamNonBlocking, (pe && pe[0] == 'h') ? hi : lo));
    // (all kSideStreams join events: the loss backward also uses them as per-tail "chain done" events, whatever `want`)
    for (int k = 0; k < tnt_decoder::kSideStreams; ++k) TNT_CUDA(cudaEventCreateWithFlags(&D->side_join[k], cudaEventDisableTiming));
    TNT_CUDA(cudaStreamCreateWithFlags(&D->stage_stream, cudaStreamNonBlocking));
    TNT_CUDA(cudaEventCreateWithFlags(&D->side_fork, cudaEventDisableTiming));
    D->n_side = want;
  }
  while (static_cast<int>(D->stage_events.size()) < Ls + 2) {
    cudaEvent_t e;
    TNT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    D->stage_events.push_back(e);
  }
  while (static_cast<int>(D->side_layer_done.size()) < (Ls + 1) * tnt_decoder::kSideStreams) {
    cudaEvent_t e;
    TNT_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    D->side_layer_done.push_back(e);
  }
  return TNT_OK;
}

// teacher-forced forward keeping every activation the backward needs; ends with X in B x T row order in w.Xbt
int train_forward(tnt_decoder* D, const int64_t* ids, const tnt_contexts* ctx, TrainWs& w, const tnt_dropout_config* dr,
                  cudaStream_t st, float* attn_out = nullptr, void* emb_out = nullptr) {
  const tnt_decoder_config& cf = D->cfg;
  const int B = ctx->B, T = ctx->T, N = B * T, d = cf.d, ns = cf.n_sources, Ls = cf.n_layers;
  {
    FwdWs f0 = layer_ws(w, 0, ns);
    TNT_TRY(project_kv(D, ctx, f0, st));
  }
  TNT_TRY(embed(D, ids, B, T, w.emb, nullptr, st));
  // `inp_embed` of the pointer decoders = the embedder output BEFORE dropout, batch first (decoder_pointer.py:100-101)
  if (emb_out) TNT_TRY(permute_rows(w.emb, emb_out, B, T, d, cf.dtype, false, st));
  const tnt_dropout d_emb = drop_site(dr, dr ? dr->dropout : 0.f, TNT_DROP_SITE_EMBED);     // decoder_flattened.py:109
  if (d_emb.p > 0.f) TNT_TRY(dropout_apply(w.emb, d, nullptr, 0, w.emb, d, N, d, cf.dtype, d_emb, st));
  for (int l = 0; l < Ls; ++l) {
    FwdWs f = layer_ws(w, l, ns);
    const void* xin = l == 0 ? w.emb : w.act[l - 1].Xout;
    TNT_TRY(run_layer(D, l, ctx, f, xin, w.act[l].Xout, B, T, nullptr, nullptr, l == Ls - 1 ? attn_out : nullptr, false, st, 1, dr));
  }
  return permute_rows(w.act[Ls - 1].Xout, w.Xbt, B, T, d, cf.dtype, false, st);     // T x B -> B x T rows
}

// AdaptiveLoss on w.Xbt, then the backward pass through loss, layers, K/V projections and embedding bands
// upstream gradients of a head that sits on top of the decoder (the copy heads of the pointer decoders, driven by autograd):
// row_w[r] = dL/d lp_r with lp = adasoftmax_target_logprob (already run on w.Xbt: the statistics are in w.loss_fwd),
// dX = dL/dX [B, T, d] beyond the softmax path, dattn = dL/d(returned attention weights) [B, T, S_w + 2],
// demb = dL/d inp_embed [B, T, d]
struct ExtGrads { const float* row_w; const void* dX; const float* dattn; const void* demb; };

int train_backward(tnt_decoder* D, const int64_t* ids, const int64_t* target, const tnt_contexts* ctx, float loss_scale,
                   float* out2, TrainWs& w, const tnt_dropout_config* dr, cudaStream_t st, const ExtGrads* ext = nullptr) {
  const tnt_decoder_config& cf = D->cfg;
  const int B = ctx->B, T = ctx->T, N = B * T, d = cf.d, ns = cf.n_sources, Ls = cf.n_layers;
  Ctx c{D, st, cf.dtype, dr};
  c.n_side = D->n_side;
  if (!ext) TNT_TRY(adasoftmax_nll(D->ada, w.Xbt, d, target, N, cf.pad_idx, out2, w.loss_fwd, w.loss_fwd_bytes, st));
  // ---------------- backward: loss
  const int nt = cf.n_clusters - 1;
  float* d_proj[TNT_MAX_CLUSTERS] = {};
  float* d_out[TNT_MAX_CLUSTERS] = {};
  float* d_words = c.grad("embedder.token_embedder_adaptive.embeddings.0.0.weight");
  float* d_class = nt > 0 ? c.grad("adaptive_softmax.head.class_proj.weight") : nullptr;
  TNT_CHECK_ARG(d_words && (nt == 0 || d_class), "decoder_train_step: gradient buffers of the softmax head missing");
  for (int i = 0; i < nt; ++i) {
    d_proj[i] = c.grad("adaptive_softmax.tail." + std::to_string(i) + ".0.weight");
    d_out[i] = c.grad("embedder.token_embedder_adaptive.embeddings." + std::to_string(i + 1) + ".0.weight");
    TNT_CHECK_ARG(d_proj[i] && d_out[i], "decoder_train_step: gradient buffers of softmax tail %d missing", i);
  }
  // stage `stage` is final when the chain AND the parameter-gradient work issued so far are: the event is recorded on a
  // kernel-less stream that waits for both (neither the chain nor the side streams wait for each other here); `slot`
  // = 0..Ls-1 after layer `slot`'s backward (those events also order the reuse of its PGradBufs), Ls after the loss
  auto record_stage = [&](int stage, int slot) -> int {
    if (!c.n_side) {
      TNT_CUDA(cudaEventRecord(D->stage_events[stage], st));
      return TNT_OK;
    }
    cudaStream_t s0 = D->stage_stream;
    for (int k = 0; k < c.n_side; ++k) {
      cudaEvent_t done = D->side_layer_done[slot * tnt_decoder::kSideStreams + k];
      TNT_CUDA(cudaEventRecord(done, D->side_streams[k]));
      TNT_CUDA(cudaStreamWaitEvent(s0, done, 0));
    }
    TNT_CUDA(cudaEventRecord(D->side_fork, st));
    TNT_CUDA(cudaStreamWaitEvent(s0, D->side_fork, 0));
    TNT_CUDA(cudaEventRecord(D->stage_events[stage], s0));
    return TNT_OK;
  };
  // the clusters of the loss backward spread over the parameter-gradient streams (adasoftmax.cu); ext: dz = row_w (onehot - p)
  AdaSide sd = {};
  static const bool loss_par = !(getenv("TNT_LOSS_BWD_STREAMS") && atoi(getenv("TNT_LOSS_BWD_STREAMS")) == 0);
  if (c.n_side && loss_par) {
    sd.n = std::min(c.n_side, 4);
    for (int k = 0; k < sd.n; ++k) sd.st[k] = D->side_streams[k];
    sd.fork = D->side_fork;
    for (int i = 0; i < nt && i < tnt_decoder::kSideStreams; ++i) sd.done[i] = D->side_join[i];
    if (nt > tnt_decoder::kSideStreams) sd.n = 0;
    c.side_pending = sd.n > 0;
  }
  TNT_TRY(adasoftmax_nll_bwd(D->ada, w.Xbt, d, N, cf.pad_idx, ext ? -1.f : loss_scale, w.loss_fwd, w.dXbt, d, d_words, d_class,
                             d_proj, d_out, w.loss_bwd, w.loss_bwd_bytes, st, ext ? ext->row_w : nullptr, sd.n ? &sd : nullptr));
  if (ext && ext->dX) TNT_TRY(add_col_blocks(w.dXbt, d, ext->dX, d, N, d, 0, 1, cf.dtype, st));
  TNT_TRY(permute_rows(w.dXbt, w.gA, B, T, d, cf.dtype, true, st));                    // B x T -> T x B rows
  // ---------------- backward: layers, K/V projections, embedding
  TNT_TRY(record_stage(0, Ls));
  for (int l = Ls - 1; l >= 0; --l) {
    const void* xin = l == 0 ? w.emb : w.act[l - 1].Xout;
    TNT_TRY(layer_bwd(c, ctx, w, l, xin, (ext && l == Ls - 1) ? ext->dattn : nullptr));
    TNT_TRY(record_stage(Ls - l, l));
  }
  const tnt_dropout d_emb = drop_site(dr, dr ? dr->dropout : 0.f, TNT_DROP_SITE_EMBED);
  if (d_emb.p > 0.f) TNT_TRY(dropout_apply(w.gA, d, nullptr, 0, w.gA, d, N, d, cf.dtype, d_emb, st));
  if (ext && ext->demb) {                        // inp_embed is the PRE-dropout embedder output
    TNT_TRY(permute_rows(ext->demb, w.gB, B, T, d, cf.dtype, true, st));
    TNT_TRY(add_col_blocks(w.gA, d, w.gB, d, N, d, 0, 1, cf.dtype, st));
  }
  // the band tables are tied to the softmax output rows: embed_bwd ADDS its rows to what the loss backward's weight-gradient
  // GEMMs (side streams) wrote -- stage 0 is long final by now, the wait only states the order
  if (c.n_side) TNT_CUDA(cudaStreamWaitEvent(st, D->stage_events[0], 0));
  TNT_TRY(embed_bwd(c, ids, B, T, w, w.gA));
  TNT_TRY(side_join(c));                         // the caller's stream has every gradient from here on
  TNT_CUDA(cudaEventRecord(D->stage_events[Ls + 1], st));
  return TNT_OK;
}

// the dropout configuration of the step being issued: step k after tnt_decoder_set_dropout draws its masks with seed + k
const tnt_dropout_config* next_dropout(tnt_decoder* D, tnt_dropout_config* drc) {
  *drc = D->drop_cfg;
  const bool on = drc->dropout > 0.f || drc->input_dropout > 0.f || drc->weight_dropout > 0.f || drc->attention_dropout > 0.f ||
                  drc->relu_dropout > 0.f;
  drc->seed += D->drop_step++;
  return on ? drc : nullptr;
}
}  // namespace

extern "C" int tnt_decoder_train_step(tnt_decoder* D, const int64_t* ids, const int64_t* target, const tnt_contexts* ctx,
                                      float loss_scale, float* out2, void* workspace, size_t workspace_bytes,
                                      tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_train_step: decoder not finalized");
  TNT_CHECK_ARG(ids && target && ctx && out2 && workspace, "decoder_train_step: null pointer");
  StepSetup s;
  TNT_TRY(step_setup(D, ctx, workspace, workspace_bytes, "decoder_train_step", &s));
  s.dr = next_dropout(D, &s.drc);
  cudaStream_t st = as_stream(stream);
  TNT_TRY(train_forward(D, ids, ctx, s.w, s.dr, st));
  return train_backward(D, ids, target, ctx, loss_scale, out2, s.w, s.dr, st);
}

// The same step in two calls, for callers whose forward and loss are separate module calls (the reference model computes
// `decoder(...)`, then `criterion(...)`, then `loss.backward()`: transformer_faces_objects.py:79-88,
// callback_apex_trainer.py:229-233).  train_forward keeps the activations in `workspace`; train_backward must be given
// the SAME ids / contexts / workspace, untouched in between.
extern "C" int tnt_decoder_train_forward(tnt_decoder* D, const int64_t* ids, const tnt_contexts* ctx, void* X_out, void* workspace,
                                         size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_train_forward: decoder not finalized");
  TNT_CHECK_ARG(ids && ctx && X_out && workspace, "decoder_train_forward: null pointer");
  StepSetup s;
  TNT_TRY(step_setup(D, ctx, workspace, workspace_bytes, "decoder_train_forward", &s));
  D->pending_dr_on = next_dropout(D, &D->pending_drc) != nullptr;
  D->pending_ws = workspace;
  cudaStream_t st = as_stream(stream);
  TNT_TRY(train_forward(D, ids, ctx, s.w, D->pending_dr_on ? &D->pending_drc : nullptr, st));
  TNT_CUDA(cudaMemcpyAsync(X_out, s.w.Xbt, static_cast<size_t>(ctx->B) * ctx->T * D->cfg.d * D->esz(), cudaMemcpyDeviceToDevice, st));
  return TNT_OK;
}

// Variants for a head on top of the decoder (the copy heads of DynamicConvDecoderPointer / ...EntityPointer, whose small
// p_gen / mix arithmetic the host side leaves to autograd): the forward also returns the last layer's head-averaged
// attention over the copied source and `inp_embed`; tnt_decoder_target_logprob gives log softmax_full(x_r)[target_r] per
// row; the backward takes the upstream gradients w.r.t. those four quantities.
extern "C" int tnt_decoder_train_forward_ext(tnt_decoder* D, const int64_t* ids, const tnt_contexts* ctx, void* X_out, float* attn_out,
                                             void* emb_out, void* workspace, size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_train_forward_ext: decoder not finalized");
  TNT_CHECK_ARG(ids && ctx && X_out && workspace, "decoder_train_forward_ext: null pointer");
  StepSetup s;
  TNT_TRY(step_setup(D, ctx, workspace, workspace_bytes, "decoder_train_forward_ext", &s));
  D->pending_dr_on = next_dropout(D, &D->pending_drc) != nullptr;
  D->pending_ws = workspace;
  cudaStream_t st = as_stream(stream);
  TNT_TRY(train_forward(D, ids, ctx, s.w, D->pending_dr_on ? &D->pending_drc : nullptr, st, attn_out, emb_out));
  TNT_CUDA(cudaMemcpyAsync(X_out, s.w.Xbt, static_cast<size_t>(ctx->B) * ctx->T * D->cfg.d * D->esz(), cudaMemcpyDeviceToDevice, st));
  return TNT_OK;
}

extern "C" int tnt_decoder_target_logprob(tnt_decoder* D, const int64_t* target, const tnt_contexts* ctx, float* lp_out,
                                          void* workspace, size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized && target && ctx && lp_out && workspace, "decoder_target_logprob: bad arguments");
  TNT_CHECK_ARG(D->pending_ws == workspace, "decoder_target_logprob: no matching tnt_decoder_train_forward_ext on this workspace");
  StepSetup s;
  TNT_TRY(step_setup(D, ctx, workspace, workspace_bytes, "decoder_target_logprob", &s));
  float* scratch2 = s.w.vecs[0];                   // the AdaptiveLoss sum of the statistics pass is not used here
  return adasoftmax_target_logprob(D->ada, s.w.Xbt, D->cfg.d, target, ctx->B * ctx->T, D->cfg.pad_idx, lp_out, scratch2,
                                   s.w.loss_fwd, s.w.loss_fwd_bytes, as_stream(stream));
}

extern "C" int tnt_decoder_train_backward_ext(tnt_decoder* D, const int64_t* ids, const int64_t* target, const tnt_contexts* ctx,
                                              const float* row_w, const void* dX, const float* dattn, const void* demb,
                                              void* workspace, size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_train_backward_ext: decoder not finalized");
  TNT_CHECK_ARG(ids && target && ctx && row_w && workspace, "decoder_train_backward_ext: null pointer");
  TNT_CHECK_ARG(D->pending_ws == workspace, "decoder_train_backward_ext: no matching forward on this workspace");
  TNT_CHECK_ARG(!dattn || D->article_idx >= 0, "decoder_train_backward_ext: the decoder returns no attention weights");
  StepSetup s;
  TNT_TRY(step_setup(D, ctx, workspace, workspace_bytes, "decoder_train_backward_ext", &s));
  D->pending_ws = nullptr;
  ExtGrads ext = {row_w, dX, dattn, demb};
  return train_backward(D, ids, target, ctx, 1.f, nullptr, s.w, D->pending_dr_on ? &D->pending_drc : nullptr, as_stream(stream), &ext);
}

extern "C" int tnt_decoder_train_backward(tnt_decoder* D, const int64_t* ids, const int64_t* target, const tnt_contexts* ctx,
                                          float loss_scale, float* out2, void* workspace, size_t workspace_bytes, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->finalized, "decoder_train_backward: decoder not finalized");
  TNT_CHECK_ARG(ids && target && ctx && out2 && workspace, "decoder_train_backward: null pointer");
  TNT_CHECK_ARG(D->pending_ws == workspace, "decoder_train_backward: no matching tnt_decoder_train_forward on this workspace");
  StepSetup s;
  TNT_TRY(step_setup(D, ctx, workspace, workspace_bytes, "decoder_train_backward", &s));
  D->pending_ws = nullptr;
  return train_backward(D, ids, target, ctx, loss_scale, out2, s.w, D->pending_dr_on ? &D->pending_drc : nullptr, as_stream(stream));
}

extern "C" void tnt_debug_train_side_streams(int n) { g_train_side = n; }

extern "C" int tnt_decoder_train_stages(const tnt_decoder* d) { return d ? d->cfg.n_layers + 2 : 0; }

extern "C" int tnt_decoder_stream_wait_stage(tnt_decoder* d, int stage, tnt_stream stream) {
  TNT_CHECK_ARG(d && stage >= 0 && stage < static_cast<int>(d->stage_events.size()),
                "stream_wait_stage: no training step has been issued, or bad stage");
  TNT_CUDA(cudaStreamWaitEvent(as_stream(stream), d->stage_events[stage], 0));
  return TNT_OK;
}
