// Decoder engine state: prepared weights (weight-norm folded, fused per-source projections, engine
// dtype) and the workspace layouts.  See engine_load.cu / engine_fwd.cu.
#pragma once
#include <map>
#include <set>
#include <string>
#include <vector>

#include "adasoftmax.cuh"

namespace tnt {

struct RawTensor {
  const float* data;
  std::vector<int64_t> shape;
  int64_t numel() const {
    int64_t n = 1;
    for (auto s : shape) n *= s;
    return n;
  }
};

struct LayerW {
  int K;
  void* w1; float* b1;            // [2d, d] GLU pair-interleaved, bias likewise
  void* wtap;                     // [H*K, d]
  void* w2; float* b2;            // [d, d]
  float* ln_conv_g; float* ln_conv_b;
  void* wq; float* bq;            // [n_src*d, d]: the query projections of all sources stacked
  void* wo; float* bo;            // [n_src][d, d], [n_src*d]
  float* bias_k; float* bias_v;   // [n_src*d]
  float* ln_src_g; float* ln_src_b;   // [n_src, d]
  void* wc; float* bc;            // [d, n_src*d]
  void* wf1; float* bf1;          // [ffn, d]
  void* wf2; float* bf2;          // [d, ffn]
  float* ln_f_g; float* ln_f_b;
};

}  // namespace tnt

struct tnt_decoder {
  tnt_decoder_config cfg;
  std::map<std::string, tnt::RawTensor> raw;
  bool finalized = false;
  char* wmem = nullptr;
  size_t wbytes = 0;
  // prepared weights
  std::vector<tnt::LayerW> layers;
  void* wkv[TNT_MAX_SOURCES] = {};     // [L*2d, source_dim]: per layer K rows then V rows
  float* bkv[TNT_MAX_SOURCES] = {};    // [L*2d]
  float* kvrow[TNT_MAX_SOURCES] = {};  // [L*2d] the learned bias_k | bias_v row of every layer (extra key S)
  void* table = nullptr;               // [vocab, d] pre-projected, pre-scaled embedding
  float* pos_table = nullptr;          // [max_positions, d]
  int* err_flags = nullptr;            // device int: TNT_DEVERR_* bits raised by the embedding kernels
  tnt::AdaSoftmax ada;
  void* scratch_dd = nullptr;          // [d, d] fp32 (finalize only)
  void* emb_scratch = nullptr;             // largest band table in the engine dtype (finalize only)
  tnt::FoldJob* fold_dev = nullptr;             // device copy of the finalize job list (one batched launch refreshes the weights)
  std::vector<tnt::FoldJob> fold_host;          //   and the list it was built from
  void* emb_proj[TNT_MAX_CLUSTERS] = {};   // band projections [d, dim_i] in the engine dtype (training backward)
  int emb_dim[TNT_MAX_CLUSTERS] = {};
  std::vector<cudaEvent_t> stage_events;   // training: gradients of stage i are final (tnt_decoder_stream_wait_stage)
  // training backward: every parameter gradient (weight-gradient GEMM, weight-norm backward, bias column sum) runs on one
  // of a few side streams beside the latency-bound data-gradient chain (engine_train.cu)
  static constexpr int kSideStreams = 4;
  int n_side = 0;                                  // side streams in use (0: everything on the step's stream)
  cudaStream_t side_streams[kSideStreams] = {};
  cudaStream_t stage_stream = nullptr;             // carries no kernels: joins chain + side streams for the stage events
  cudaEvent_t side_fork = nullptr, side_join[kSideStreams] = {};
  std::vector<cudaEvent_t> side_layer_done;        // [layer * kSideStreams + k]: side stream k has issued layer l's work
  // inference: the batch is cut into a few sub-batches whose latency-bound kernel chains run side by side on these streams
  // (engine_fwd.cu: SubBatches)
  static constexpr int kMaxSub = 4;
  cudaStream_t sub_streams[kMaxSub - 1] = {};
  cudaEvent_t sub_fork = nullptr, sub_join[kMaxSub - 1] = {};
  // teacher-forced forward: the memory K/V projections run layer by layer on this stream, on a share of the SMs, beside the
  // latency-bound layer chain (engine_fwd.cu: project_kv_beside); kv_ready[l] = layer l's K/V of every source are written
  cudaStream_t kv_stream = nullptr;
  cudaEvent_t kv_fork = nullptr;
  std::vector<cudaEvent_t> kv_ready;
  std::map<std::string, float*> grads; // training: fp32 gradient buffer per state_dict key (tnt_decoder_set_grad)
  // data-parallel training: embedding-row gradients go to a packed list instead of the dense band tables
  void* sparse_rows = nullptr; long long* sparse_meta = nullptr; int* sparse_count = nullptr; int sparse_capacity = 0;
  int article_idx = -1;   // source whose head-averaged attention the forward returns ("entity" if present, else "article")
  tnt_dropout_config drop_cfg = {};    // training dropout (tnt_decoder_set_dropout); all zero = off
  uint64_t drop_step = 0;              // training steps issued since the configuration was set
  tnt_dropout_config pending_drc = {}; bool pending_dr_on = false; void* pending_ws = nullptr;   // train_forward -> train_backward
  // pipelined host scoring (tnt_decoder_score_host_submit / _wait): copy stream + per-slot events
  cudaStream_t copy_stream = nullptr;
  cudaStream_t copy_extra[3] = {};         // further DMA queues of the live-row copies (TNT_H2D_STREAMS, default 2 in all)
  cudaEvent_t ev_cx_fork = nullptr, ev_cx_join[3] = {};
  cudaEvent_t ev_h2d[2] = {}, ev_free[2] = {}, ev_done[2] = {};
  bool slot_used[2] = {false, false};
  std::set<const void*> zeroed_staging;    // staging memory regions whose skipped (padded) rows have been zeroed once
  size_t last_h2d_bytes = 0;               // host -> device bytes of the last score_host call

  int d() const { return cfg.d; }
  int nsrc() const { return cfg.n_sources; }
  size_t esz() const { return tnt::dtype_size(cfg.dtype); }
  int ld_logits(int K) const { return (cfg.heads * K + 7) & ~7; }
  int max_ld_logits() const {
    int m = 0;
    for (int l = 0; l < cfg.n_layers; ++l) m = std::max(m, ld_logits(cfg.kernels[l]));
    return m;
  }
};

namespace tnt {

// forward workspace (teacher forcing and single-step share it; N = T*B rows)
struct FwdWs {
  void *X0, *X1, *h, *logits, *cbuf, *ybuf, *Q, *AO, *CAT, *F;
  void *ybuf2, *CATln;     // inference: aliases of ybuf / CAT; training keeps the pre-LayerNorm tensors
  void* Xd;                // training with input_dropout: the dropped-out layer input linear1 reads (else NULL)
  void* kv[TNT_MAX_SOURCES];
  float* lse[TNT_MAX_SOURCES];   // training: attention log-sum-exp [B, H, T] per source (NULL: not saved)
  uint32_t* tile_live[TNT_MAX_SOURCES];   // [B_kv] live 128-key tiles of every context entry (project_kv fills it)
  int* order[TNT_MAX_SOURCES];   // decode engines: context entries sorted by live keys, longest first (or NULL)
  int* sched;                    // 2 zeroed ints: dynamic work counters of the decode-step attention (or NULL)
  float* step_scores = nullptr;  // generation through the copy head: [B, H, S_w + 2] score dump of the decode-step attention
  float* step_lse = nullptr;     //   and its [B, H] log-sum-exp
  int kv_B = 0;                  // context entries the K/V buffers hold when the rows are a SUB-batch of them (0: B / kv_group)
  const cudaEvent_t* kv_ready = nullptr;   // [n_layers] or NULL: run_layer makes its stream wait for kv_ready[l] before layer l's attention
  size_t total;
};
// Sub-batches.  At the benchmark's batch sizes every kernel of a decoder layer is latency-bound (a 2048-row GEMM is one wave
// of tiles, most of its 12 us is launch, pipeline fill and epilogue), and the decode step alternates between one HBM-bound
// kernel (attention over the K/V cache) and ~60 such launches.  The captions are independent, so the batch is cut into
// n row ranges; each runs the same kernel chain on its own stream (sub-batch 0 on the caller's), and one range's
// latency hides behind another's work.  The K/V projections are still ONE GEMM per source over the whole batch.
struct SubBatches {
  int n = 1;
  int b0[tnt_decoder::kMaxSub + 1] = {0};          // sub-batch i = context entries [b0[i], b0[i+1])
  cudaStream_t st[tnt_decoder::kMaxSub] = {};
};
// plan: env TNT_SUB_BATCHES (default `dflt`), at least `min_rows` captions per sub-batch; creates the streams on first use
int sub_plan(tnt_decoder* d, int B, int dflt, int min_rows, cudaStream_t st, SubBatches* sb);
int sub_fork(tnt_decoder* d, const SubBatches& sb);      // streams 1.. see what has been issued on stream 0
int sub_join(tnt_decoder* d, const SubBatches& sb);      // stream 0 waits for streams 1..
// workspace / contexts of sub-batch i: rows [b0*T, b1*T) of every activation buffer (a private [T, nb, .] block), the K/V
// cache entries b0.. of the SHARED cache (kv_B = B), the slices of the masks / live-tile words / orders
FwdWs sub_ws(const tnt_decoder* d, const FwdWs& w, const SubBatches& sb, int i, int B, int T, const int* S);
tnt_contexts sub_ctx(const tnt_contexts& c, const SubBatches& sb, int i);
// B_kv >= 0: number of context entries the K/V cache holds when it differs from the row batch B (beam search)
FwdWs carve_fwd(const tnt_decoder* d, int B, int T, const int* S, void* ws, bool with_kv, int B_kv = -1);

int embed(tnt_decoder* d, const int64_t* ids, int B, int T, void* out, const int* start_pos, cudaStream_t st);
struct SubBatches;
int project_kv(tnt_decoder* d, const tnt_contexts* ctx, const FwdWs& w, cudaStream_t st, bool decode_order = false,
               const SubBatches* sb = nullptr);
// caller-owned chronological conv history of one layer (the reference's `DynamicConv1dTBC.<id>.input_buffer`,
// dynamic.py:95-99): `in` = the last `len` (<= K-1) conv inputs [len, B, C], oldest first; `out` (optional, must not
// alias `in`) receives the last min(K-1, len + T) rows of [in ; this call's conv inputs]
struct ConvHist { const void* in; int len; void* out; };
// one decoder layer on N = T*B rows; `ring`/`pos` != NULL selects the single-step conv (T must be 1); `ch` the
// step-wise protocol with caller-owned history (tnt_decoder_forward_incremental)
int run_layer(tnt_decoder* d, int l, const tnt_contexts* ctx, const FwdWs& w, const void* Xin, void* Xout, int B, int T,
              void* ring, const int* pos, float* attn_out, bool transpose_out, cudaStream_t st, int kv_group = 1,
              const tnt_dropout_config* dr = nullptr, const ConvHist* ch = nullptr);
// dr != NULL (training): the dropout sites of the layer are active with seed dr->seed (include/tnt_b200.h, TNT_DROP_SITE)
int dropout_apply(const void* in, long ldi, const void* res, long ldr, void* out, long ldo, int rows, int cols, int dtype,
                  const tnt_dropout& dr, cudaStream_t st);
inline tnt_dropout drop_site(const tnt_dropout_config* dr, float p, uint32_t site, long ld = 0, long gs = 0) {
  tnt_dropout t = {};
  if (dr && p > 0.f) {
    t.p = p; t.site = site; t.seed = dr->seed; t.ld = ld; t.group_stride = gs;
  }
  return t;
}
// kv_group > 1 (beam search): B counts beam rows; rows b*kv_group.. share context entry b (ctx->B = B / kv_group)

}  // namespace tnt
