// Decoder engine: construction, loading a reference state_dict, weight preparation.
// State-dict keys follow the reference modules:
//   ttl/tell/models/decoder_flattened.py:213-269, decoder_faces_objects.py:211-281 (layer members),
//   ttl/tell/modules/linear.py:7-33 (weight_g / weight_v), multi_head.py:234-262 (in_proj / *_proj_weight),
//   token_embedders/adaptive.py:36-59, softmax.py:83-132 (tied head / tails).
#include <string.h>

#include "engine.cuh"

namespace tnt {
namespace {

__global__ void scale_kernel(const float* s, float* d, long n, float a) {
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long>(gridDim.x) * blockDim.x)
    d[i] = s[i] * a;
}

inline size_t al256(size_t x) { return (x + 255) & ~size_t(255); }

struct Arena {
  char* base;
  size_t off = 0;
  explicit Arena(char* b) : base(b) {}
  void* take(size_t bytes) {
    void* p = base ? base + off : nullptr;
    off += al256(bytes);
    return p;
  }
  float* takef(size_t n) { return static_cast<float*>(take(n * 4)); }
};

// lays out every prepared tensor; base == nullptr only measures
constexpr int kMaxFoldJobs = 1024;

size_t plan(tnt_decoder* D, char* base) {
  const tnt_decoder_config& c = D->cfg;
  const size_t E = D->esz(), d = c.d, ns = c.n_sources, L = c.n_layers;
  Arena a(base);
  D->layers.resize(L);
  for (size_t l = 0; l < L; ++l) {
    LayerW& w = D->layers[l];
    w.K = c.kernels[l];
    w.w1 = a.take(2 * d * d * E); w.b1 = a.takef(2 * d);
    w.wtap = a.take(static_cast<size_t>(c.heads) * w.K * d * E);
    w.w2 = a.take(d * d * E); w.b2 = a.takef(d);
    w.ln_conv_g = a.takef(d); w.ln_conv_b = a.takef(d);
    w.wq = a.take(ns * d * d * E); w.bq = a.takef(ns * d);
    w.wo = a.take(ns * d * d * E); w.bo = a.takef(ns * d);
    w.bias_k = a.takef(ns * d); w.bias_v = a.takef(ns * d);
    w.ln_src_g = a.takef(ns * d); w.ln_src_b = a.takef(ns * d);
    w.wc = a.take(d * ns * d * E); w.bc = a.takef(d);
    w.wf1 = a.take(static_cast<size_t>(c.ffn) * d * E); w.bf1 = a.takef(c.ffn);
    w.wf2 = a.take(static_cast<size_t>(c.ffn) * d * E); w.bf2 = a.takef(d);
    w.ln_f_g = a.takef(d); w.ln_f_b = a.takef(d);
  }
  for (size_t s = 0; s < ns; ++s) {
    D->wkv[s] = a.take(L * 2 * d * c.source_dim[s] * E);
    D->bkv[s] = a.takef(L * 2 * d);
    D->kvrow[s] = a.takef(L * 2 * d);
  }
  D->table = a.take(static_cast<size_t>(c.vocab) * d * E);
  D->pos_table = a.takef(static_cast<size_t>(c.max_positions) * d);
  const int nt = c.n_clusters - 1;
  D->ada.cuts.n = c.n_clusters;
  for (int i = 0; i < c.n_clusters; ++i) D->ada.cuts.cut[i] = c.cutoffs[i];
  D->ada.d = c.d;
  D->ada.dtype = c.dtype;
  D->ada.head_w = a.take(static_cast<size_t>(c.cutoffs[0] + nt) * d * E);
  for (int i = 0; i < nt; ++i) {
    D->ada.proj_w[i] = a.take(d * d * E);
    D->ada.out_w[i] = a.take(static_cast<size_t>(c.cutoffs[i + 1] - c.cutoffs[i]) * d * E);
    D->ada.fold_w[i] = a.take(static_cast<size_t>(c.cutoffs[i + 1] - c.cutoffs[i]) * d * E);
  }
  D->scratch_dd = a.take(d * d * 4);
  for (int i = 0; i < c.n_clusters; ++i) D->emb_proj[i] = a.take(d * d * E);
  {
    size_t max_band = 0;
    int lo = 0;
    for (int i = 0; i < c.n_clusters; ++i) {
      max_band = std::max<size_t>(max_band, c.cutoffs[i] - lo);
      lo = c.cutoffs[i];
    }
    D->emb_scratch = a.take(max_band * d * E);
  }
  D->fold_dev = static_cast<FoldJob*>(a.take(sizeof(FoldJob) * kMaxFoldJobs));
  D->err_flags = static_cast<int*>(a.take(256));
  return a.off;
}

struct Loader {
  tnt_decoder* D;
  cudaStream_t st;
  int err = TNT_OK;
  // row-wise copies (matrix folds / casts and fp32 vectors) are queued and run as ONE launch (flush): the ~70 fold launches
  // and ~130 small device copies of a refresh were 0.4 ms of every training step
  std::vector<FoldJob> jobs;
  int total_rows = 0;
  void push(const float* v, const float* g, void* out, int rows, int cols, bool glu, bool out_f32) {
    if (jobs.size() >= static_cast<size_t>(kMaxFoldJobs)) {        // cannot happen for <= 4 sources x 8 layers
      chk(tnt_wn_fold(v, g, out, rows, cols, glu, out_f32 ? TNT_F32 : D->cfg.dtype, st));
      return;
    }
    jobs.push_back(FoldJob{v, g, out, rows, cols, glu ? 1 : 0, out_f32 ? 1 : 0, total_rows});
    total_rows += rows;
  }
  void flush() {
    if (jobs.empty() || err != TNT_OK) return;
    if (!(jobs == D->fold_host)) {         // first refresh (or re-bound pointers): upload the list
      if (cudaMemcpyAsync(D->fold_dev, jobs.data(), sizeof(FoldJob) * jobs.size(), cudaMemcpyHostToDevice, st) != cudaSuccess ||
          cudaStreamSynchronize(st) != cudaSuccess) {
        set_error("decoder_finalize: uploading the refresh job list failed");
        err = TNT_ERR_CUDA;
        return;
      }
      D->fold_host = jobs;
    }
    chk(wn_fold_batch(D->fold_dev, static_cast<int>(jobs.size()), total_rows, D->cfg.dtype, st));
    jobs.clear();
    total_rows = 0;
  }

  const RawTensor* find(const std::string& k) {
    auto it = D->raw.find(k);
    return it == D->raw.end() ? nullptr : &it->second;
  }
  const RawTensor* need(const std::string& k, int64_t numel) {
    const RawTensor* t = find(k);
    if (!t) {
      if (err == TNT_OK) set_error("decoder_finalize: missing state_dict entry '%s'", k.c_str());
      err = TNT_ERR_ARG;
      return nullptr;
    }
    if (numel >= 0 && t->numel() != numel) {
      if (err == TNT_OK)
        set_error("decoder_finalize: '%s' has %lld elements, expected %lld", k.c_str(), (long long)t->numel(),
                  (long long)numel);
      err = TNT_ERR_ARG;
      return nullptr;
    }
    return t;
  }
  void chk(int s) {
    if (s != TNT_OK && err == TNT_OK) err = s;
  }
  // GehringLinear weight (weight_g/weight_v or plain weight) -> prepared matrix
  void linear(const std::string& p, void* w, float* b, int rows, int cols, bool glu) {
    const RawTensor* g = find(p + "weight_g");
    if (g) {
      const RawTensor* v = need(p + "weight_v", static_cast<int64_t>(rows) * cols);
      if (v && need(p + "weight_g", rows)) push(v->data, g->data, w, rows, cols, glu, false);
    } else {
      const RawTensor* v = need(p + "weight", static_cast<int64_t>(rows) * cols);
      if (v) push(v->data, nullptr, w, rows, cols, glu, false);
    }
    if (b) {
      const RawTensor* bb = need(p + "bias", rows);
      if (bb) {
        if (glu) chk(tnt_glu_interleave_bias(bb->data, b, rows, st));
        else vec(bb->data, b, rows);
      }
    }
  }
  void matrix(const float* src, void* dst, int rows, int cols) {   // plain cast of a weight slice
    push(src, nullptr, dst, rows, cols, false, false);
  }
  void matrix_now(const float* src, void* dst, int rows, int cols) {   // same, launched immediately (dst is reused right away)
    chk(tnt_wn_fold(src, nullptr, dst, rows, cols, 0, D->cfg.dtype, st));
  }
  void vec(const float* src, float* dst, size_t n) {               // fp32 vector copy; long ones as rows of 1024
    if (n > 4096 && n % 1024 == 0) push(src, nullptr, dst, static_cast<int>(n / 1024), 1024, false, true);
    else push(src, nullptr, dst, 1, static_cast<int>(n), false, true);
  }
  void vec_key(const std::string& k, float* dst, size_t n) {
    const RawTensor* t = need(k, n);
    if (t) vec(t->data, dst, n);
  }
};

}  // namespace
}  // namespace tnt

using namespace tnt;

extern "C" tnt_decoder* tnt_decoder_create(const tnt_decoder_config* cfg) {
  if (!cfg || cfg->n_layers < 1 || cfg->n_layers > TNT_MAX_LAYERS || cfg->d % 8 || cfg->heads < 1 ||
      cfg->d % cfg->heads || (cfg->d / cfg->heads) % 8 || cfg->n_sources < 1 || cfg->n_sources > TNT_MAX_SOURCES ||
      cfg->n_clusters < 1 || cfg->n_clusters > TNT_MAX_CLUSTERS || cfg->ffn % 8 ||
      (cfg->dtype != TNT_F32 && cfg->dtype != TNT_BF16)) {
    set_error("decoder_create: unsupported configuration (d and d/heads must be multiples of 8, ...)");
    return nullptr;
  }
  for (int s = 0; s < cfg->n_sources; ++s)
    if (cfg->source_dim[s] % 8) {
      set_error("decoder_create: source_dim[%d] = %d must be a multiple of 8", s, cfg->source_dim[s]);
      return nullptr;
    }
  tnt_decoder* D = new tnt_decoder();
  D->cfg = *cfg;
  for (int s = 0; s < cfg->n_sources; ++s)
    if (!strcmp(cfg->source_names[s], "article") && D->article_idx < 0) D->article_idx = s;
  for (int s = 0; s < cfg->n_sources; ++s)      // the entity-pointer layer returns the ENTITY attention (decoder_entity_pointer.py:326-334)
    if (!strcmp(cfg->source_names[s], "entity")) D->article_idx = s;
  D->wbytes = plan(D, nullptr);
  return D;
}

extern "C" void tnt_decoder_destroy(tnt_decoder* d) {
  if (!d) return;
  // library-owned streams and events (cudaStreamDestroy lets queued work finish; the handles are released afterwards)
  auto ev = [](cudaEvent_t& e) { if (e) cudaEventDestroy(e); e = nullptr; };
  auto st = [](cudaStream_t& s) { if (s) cudaStreamDestroy(s); s = nullptr; };
  for (cudaEvent_t& e : d->stage_events) ev(e);
  for (cudaEvent_t& e : d->side_layer_done) ev(e);
  for (cudaStream_t& s : d->side_streams) st(s);
  for (cudaEvent_t& e : d->side_join) ev(e);
  ev(d->side_fork);
  st(d->stage_stream);
  for (cudaStream_t& s : d->sub_streams) st(s);
  for (cudaEvent_t& e : d->sub_join) ev(e);
  ev(d->sub_fork);
  st(d->kv_stream);
  ev(d->kv_fork);
  for (cudaEvent_t& e : d->kv_ready) ev(e);
  st(d->copy_stream);
  for (cudaStream_t& s : d->copy_extra) st(s);
  for (cudaEvent_t& e : d->ev_cx_join) ev(e);
  ev(d->ev_cx_fork);
  for (int i = 0; i < 2; ++i) { ev(d->ev_h2d[i]); ev(d->ev_free[i]); ev(d->ev_done[i]); }
  delete d;
}

extern "C" size_t tnt_decoder_weight_bytes(const tnt_decoder* d) { return d ? d->wbytes : 0; }

extern "C" int tnt_decoder_bind_weights(tnt_decoder* d, void* weight_mem, size_t bytes) {
  TNT_CHECK_ARG(d && weight_mem, "decoder_bind_weights: null");
  TNT_CHECK_ARG(bytes >= d->wbytes, "decoder_bind_weights: %zu bytes given, %zu needed", bytes, d->wbytes);
  TNT_CHECK_ALIGN((reinterpret_cast<uintptr_t>(weight_mem) & 255) == 0, "decoder_bind_weights: 256-byte alignment");
  d->wmem = static_cast<char*>(weight_mem);
  plan(d, d->wmem);
  d->finalized = false;
  return TNT_OK;
}

extern "C" int tnt_decoder_load(tnt_decoder* d, const char* key, const float* data, const int64_t* shape, int ndim,
                                tnt_stream) {
  TNT_CHECK_ARG(d && key && data && ndim >= 0 && ndim <= 4, "decoder_load: bad arguments");
  RawTensor t;
  t.data = data;
  t.shape.assign(shape, shape + ndim);
  d->raw[key] = t;
  d->finalized = false;
  return TNT_OK;
}

extern "C" int tnt_decoder_finalize(tnt_decoder* D, tnt_stream stream) {
  TNT_CHECK_ARG(D && D->wmem, "decoder_finalize: bind the weight memory first");
  const tnt_decoder_config& c = D->cfg;
  if (!D->finalized) TNT_CUDA(cudaMemsetAsync(D->err_flags, 0, sizeof(int), as_stream(stream)));   // first finalize after bind
  Loader L{D, as_stream(stream)};
  const int d = c.d, ns = c.n_sources;
  for (int l = 0; l < c.n_layers; ++l) {
    LayerW& w = D->layers[l];
    const std::string p = "layers." + std::to_string(l) + ".";
    L.linear(p + "linear1.", w.w1, w.b1, 2 * d, d, true);
    {
      const RawTensor* t = L.need(p + "conv.weight_linear.weight", static_cast<int64_t>(c.heads) * w.K * d);
      if (t) L.matrix(t->data, w.wtap, c.heads * w.K, d);
    }
    L.linear(p + "linear2.", w.w2, w.b2, d, d, false);
    L.vec_key(p + "conv_layer_norm.weight", w.ln_conv_g, d);
    L.vec_key(p + "conv_layer_norm.bias", w.ln_conv_b, d);
    for (int s = 0; s < ns; ++s) {
      const std::string a = p + "context_attns." + c.source_names[s] + ".";
      const int kd = c.source_dim[s];
      const size_t E = D->esz();
      char* wq = static_cast<char*>(w.wq) + static_cast<size_t>(s) * d * d * E;
      char* wk = static_cast<char*>(D->wkv[s]) + static_cast<size_t>(l) * 2 * d * kd * E;
      char* wv = wk + static_cast<size_t>(d) * kd * E;
      if (const RawTensor* ip = L.find(a + "in_proj_weight")) {
        if (kd != d || ip->numel() != 3LL * d * d) {
          set_error("decoder_finalize: '%sin_proj_weight' needs source_dim == d", a.c_str());
          return TNT_ERR_ARG;
        }
        L.matrix(ip->data, wq, d, d);
        L.matrix(ip->data + static_cast<size_t>(d) * d, wk, d, d);
        L.matrix(ip->data + 2 * static_cast<size_t>(d) * d, wv, d, d);
      } else {
        const RawTensor* q = L.need(a + "q_proj_weight", static_cast<int64_t>(d) * d);
        const RawTensor* k = L.need(a + "k_proj_weight", static_cast<int64_t>(d) * kd);
        const RawTensor* v = L.need(a + "v_proj_weight", static_cast<int64_t>(d) * kd);
        if (q && k && v) {
          L.matrix(q->data, wq, d, d);
          L.matrix(k->data, wk, d, kd);
          L.matrix(v->data, wv, d, kd);
        }
      }
      if (const RawTensor* b = L.need(a + "in_proj_bias", 3 * d)) {
        L.vec(b->data, w.bq + s * d, d);
        L.vec(b->data + d, D->bkv[s] + static_cast<size_t>(l) * 2 * d, d);
        L.vec(b->data + 2 * d, D->bkv[s] + static_cast<size_t>(l) * 2 * d + d, d);
      }
      L.vec_key(a + "bias_k", w.bias_k + s * d, d);
      L.vec_key(a + "bias_v", w.bias_v + s * d, d);
      L.vec_key(a + "bias_k", D->kvrow[s] + static_cast<size_t>(l) * 2 * d, d);
      L.vec_key(a + "bias_v", D->kvrow[s] + static_cast<size_t>(l) * 2 * d + d, d);
      if (const RawTensor* o = L.need(a + "out_proj.weight", static_cast<int64_t>(d) * d))
        L.matrix(o->data, static_cast<char*>(w.wo) + static_cast<size_t>(s) * d * d * E, d, d);
      L.vec_key(a + "out_proj.bias", w.bo + s * d, d);
      const std::string ln = p + "context_attn_lns." + c.source_names[s] + ".";
      L.vec_key(ln + "weight", w.ln_src_g + s * d, d);
      L.vec_key(ln + "bias", w.ln_src_b + s * d, d);
    }
    L.linear(p + "context_fc.", w.wc, w.bc, d, ns * d, false);
    L.linear(p + "fc1.", w.wf1, w.bf1, c.ffn, d, false);
    L.linear(p + "fc2.", w.wf2, w.bf2, d, c.ffn, false);
    L.vec_key(p + "final_layer_norm.weight", w.ln_f_g, d);
    L.vec_key(p + "final_layer_norm.bias", w.ln_f_b, d);
  }
  // embedding bands: table_i . proj_i^T * sqrt(d), pre-multiplied once (adaptive.py:61-76)
  {
    int lo = 0;
    for (int i = 0; i < c.n_clusters; ++i) {
      const std::string e = "embedder.token_embedder_adaptive.embeddings." + std::to_string(i);
      const int n_i = c.cutoffs[i] - lo;
      const RawTensor* tab = L.need(e + ".0.weight", -1);
      const RawTensor* prj = L.need(e + ".1.weight", -1);
      if (tab && prj) {
        const int dim_i = static_cast<int>(tab->numel() / n_i);
        if (tab->numel() != static_cast<int64_t>(n_i) * dim_i || prj->numel() != static_cast<int64_t>(d) * dim_i ||
            dim_i > d) {
          set_error("decoder_finalize: embedding band %d has an unexpected shape", i);
          return TNT_ERR_ARG;
        }
        D->emb_dim[i] = dim_i;
        L.matrix_now(prj->data, D->emb_proj[i], d, dim_i);
        // table_i (engine dtype copy) . proj_i^T * sqrt(d): tcgen05 GEMM for the bf16 engine, fp32 SIMT for parity mode
        L.matrix_now(tab->data, D->emb_scratch, n_i, dim_i);
        tnt_linear_params lp = {};
        lp.a = D->emb_scratch; lp.lda = dim_i; lp.w = D->emb_proj[i]; lp.ldw = dim_i;
        lp.c = static_cast<char*>(D->table) + static_cast<size_t>(lo) * d * D->esz(); lp.ldc = d;
        lp.M = n_i; lp.N = d; lp.K = dim_i; lp.groups = 1; lp.alpha = sqrtf(static_cast<float>(d));
        lp.dtype_in = c.dtype; lp.dtype_out = c.dtype; lp.epilogue = TNT_EPI_NONE;
        L.chk(tnt_linear_fwd(&lp, stream));
      }
      lo = c.cutoffs[i];
    }
  }
  if (const RawTensor* pt = L.need("embedder.token_embedder_position.weights", -1)) {
    const int64_t rows = pt->numel() / d;
    if (rows < c.max_positions) {
      set_error("decoder_finalize: position table has %lld rows, config says %d", (long long)rows, c.max_positions);
      return TNT_ERR_ARG;
    }
    L.vec(pt->data, D->pos_table, static_cast<size_t>(c.max_positions) * d);
  }
  // adaptive softmax: head = [tied band-0 words ; class_proj] (softmax.py:11-40), tails proj + tied out
  {
    const int nt = c.n_clusters - 1;
    const size_t E = D->esz();
    if (const RawTensor* wp = L.need("adaptive_softmax.head.word_proj.weight", static_cast<int64_t>(c.cutoffs[0]) * d))
      L.matrix(wp->data, const_cast<void*>(D->ada.head_w), c.cutoffs[0], d);
    if (nt > 0) {
      if (const RawTensor* cp = L.need("adaptive_softmax.head.class_proj.weight", static_cast<int64_t>(nt) * d))
        L.matrix(cp->data, static_cast<char*>(const_cast<void*>(D->ada.head_w)) + static_cast<size_t>(c.cutoffs[0]) * d * E,
                 nt, d);
    }
    for (int i = 0; i < nt; ++i) {
      const std::string t = "adaptive_softmax.tail." + std::to_string(i);
      if (const RawTensor* p0 = L.need(t + ".0.weight", static_cast<int64_t>(d) * d))
        L.matrix(p0->data, const_cast<void*>(D->ada.proj_w[i]), d, d);
      const int n_i = c.cutoffs[i + 1] - c.cutoffs[i];
      if (const RawTensor* p2 = L.need(t + ".2.weight", static_cast<int64_t>(n_i) * d))
        L.matrix(p2->data, const_cast<void*>(D->ada.out_w[i]), n_i, d);
    }
  }
  L.flush();
  if (L.err != TNT_OK) return L.err;
  // inference, bf16: fold every tail's two linear maps into one (adasoftmax.cuh); a decoder with registered gradient
  // buffers is being trained -- its factors change every step and are differentiated separately -- and keeps the two GEMMs
  {
    static const bool fold_on = [] { const char* e = getenv("TNT_FOLD_TAILS"); return !(e && e[0] == '0'); }();
    const int nt = c.n_clusters - 1;
    D->ada.use_fold = fold_on && c.dtype == TNT_BF16 && D->grads.empty() && nt > 0;
    for (int i = 0; i < nt && D->ada.use_fold; ++i) {
      tnt_linear_params lp = {};
      lp.a = D->ada.out_w[i]; lp.lda = d; lp.w = D->ada.proj_w[i]; lp.ldw = d; lp.c = const_cast<void*>(D->ada.fold_w[i]); lp.ldc = d;
      lp.M = c.cutoffs[i + 1] - c.cutoffs[i]; lp.N = d; lp.K = d; lp.groups = 1; lp.layout = TNT_NN;
      lp.dtype_in = c.dtype; lp.dtype_out = c.dtype; lp.epilogue = TNT_EPI_NONE;
      const int rc = tnt_linear_fwd(&lp, stream);
      if (rc != TNT_OK) return rc;
    }
  }
  TNT_CUDA(cudaGetLastError());
  // the raw (fp32 master) pointers are kept: the training backward reads weight_v / weight_g / the band tables
  D->finalized = true;
  return TNT_OK;
}
