// Copy head of DynamicConvDecoderPointer: mixes the vocabulary distribution with the last layer's
// article attention scattered onto the article's token ids.
//   ttl/tell/models/decoder_pointer.py:138-141 (p_gen = project_p_gens(cat(inp_embed, X))),
//   :151-216 output_layer: probs = sigmoid(p_gen) * softmax_vocab + scatter_add(article_ids, (1 - sigmoid(p_gen)) * attn[:, :, :-2]),
//                          lprobs = log(clamp(probs, 1e-10, 1)),
//   ttl/tell/models/transformer_context_pointer.py:128 (F.nll_loss(lprobs, target, ignore_index = pad, reduction = 'sum')).
// HBM-bound: one CTA owns one (caption, step) row of the [rows, V] fp32 log-prob matrix; the row (V = 50265 -> 196 KB)
// lives in SHARED memory between the three phases (p * exp(lp) -> shared-memory atomic scatter of the copy mass -> log),
// so the matrix is read once and written once (2 * V * 4 B per row, the algorithmic minimum of the in-place mix).
#include "common.cuh"

namespace tnt {
namespace {

constexpr int kThreads = 512;

template <typename T>
__device__ __forceinline__ float dot_row(const T* a, const float* w, int d, int tid, int nthreads) {
  float s = 0.f;
  for (int i = tid * 8; i < d; i += nthreads * 8) {
    float v[8];
    load8(a + i, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) s = fmaf(v[j], w[i + j], s);
  }
  return s;
}

template <typename T>
__global__ void __launch_bounds__(kThreads) pointer_mix_kernel(tnt_pointer_params p, int use_smem) {
  extern __shared__ float sm[];
  __shared__ float red[kThreads / 32];
  __shared__ int red_i[kThreads / 32];
  __shared__ float s_pgen;
  const int row = blockIdx.x, b = row / p.T, t = row - b * p.T;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // ---- p_gen = sigmoid(w[:d] . inp_embed[t, b] + w[d:] . X[b, t] + bias)
  {
    //      (entity-pointer head, inp_embed == nullptr: x is the hidden layer and p_gen = sigmoid(w . x + bias))
    const T* x = reinterpret_cast<const T*>(p.x) + static_cast<long>(row) * p.ldx;
    float s;
    if (p.inp_embed) {
      const T* e = reinterpret_cast<const T*>(p.inp_embed) + (static_cast<long>(t) * p.B + b) * p.lde;
      s = dot_row(e, p.w_pgen, p.d, tid, kThreads) + dot_row(x, p.w_pgen + p.d, p.d, tid, kThreads);
    } else {
      s = dot_row(x, p.w_pgen, p.d, tid, kThreads);
    }
    s = warp_sum(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    if (warp == 0) {
      float v = lane < kThreads / 32 ? red[lane] : 0.f;
      v = warp_sum(v);
      if (lane == 0) {
        s_pgen = 1.f / (1.f + __expf(-(v + p.b_pgen)));
        if (p.p_gen) p.p_gen[row] = s_pgen;
      }
    }
    __syncthreads();
  }
  const float pg = s_pgen;
  float* lp = p.logprob + static_cast<long>(row) * p.ldp;
  float* buf = use_smem ? sm : lp;
  // ---- generation part: p_gen * exp(log-prob)
  for (int v = tid; v < p.V; v += kThreads) buf[v] = pg * __expf(lp[v]);
  __syncthreads();
  // ---- copy part: (1 - p_gen) * attention mass of every article position, added at that position's token id
  const float* at = p.attn + static_cast<long>(row) * p.attn_ld;
  const int64_t* ids = p.article_ids + static_cast<long>(b) * p.S;
  for (int s = tid; s < p.S; s += kThreads) {
    const int64_t id = ids[s];
    if (id >= 0 && id < p.V) atomicAdd(buf + id, (1.f - pg) * at[s]);
  }
  __syncthreads();
  float best = -INFINITY;
  int best_v = 0x7fffffff;
  for (int v = tid; v < p.V; v += kThreads) {
    const float l = __logf(fminf(fmaxf(buf[v], 1e-10f), 1.f));
    lp[v] = l;
    if (l > best) { best = l; best_v = v; }       // ascending v per thread: the first maximum is kept
  }
  if (p.best_tok) {                                // arg-max of the mixed row, lowest index on ties
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int ov = __shfl_xor_sync(0xffffffffu, best_v, o);
      if (ob > best || (ob == best && ov < best_v)) { best = ob; best_v = ov; }
    }
    __syncthreads();
    if (lane == 0) { red[warp] = best; red_i[warp] = best_v; }
    __syncthreads();
    if (warp == 0) {
      best = lane < kThreads / 32 ? red[lane] : -INFINITY;
      best_v = lane < kThreads / 32 ? red_i[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int ov = __shfl_xor_sync(0xffffffffu, best_v, o);
        if (ob > best || (ob == best && ov < best_v)) { best = ob; best_v = ov; }
      }
      if (lane == 0) {
        p.best_tok[row] = best_v;
        if (p.best_logp) p.best_logp[row] = best;
      }
    }
  }
}

// out2 += [sum of -lp[row, target[row]] over rows with target != pad, number of such rows]
__global__ void __launch_bounds__(256) nll_rows_kernel(const float* lp, long ldp, const int64_t* target, int rows, int pad,
                                                      float* out2) {
  float s = 0.f, n = 0.f;
  for (int r = blockIdx.x * 256 + threadIdx.x; r < rows; r += gridDim.x * 256) {
    const int64_t tg = target[r];
    if (tg != pad) {
      s -= lp[static_cast<long>(r) * ldp + tg];
      n += 1.f;
    }
  }
  s = warp_sum(s);
  n = warp_sum(n);
  if ((threadIdx.x & 31) == 0 && n > 0.f) {
    atomicAdd(out2, s);
    atomicAdd(out2 + 1, n);
  }
}

}  // namespace
}  // namespace tnt

extern "C" int tnt_pointer_mix(const tnt_pointer_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p && p->x && p->w_pgen && p->attn && p->article_ids && p->logprob, "pointer_mix: null pointer");
  TNT_CHECK_ARG(p->B > 0 && p->T > 0 && p->S >= 0 && p->V > 0 && p->d > 0 && p->d % 8 == 0, "pointer_mix: bad shape");
  TNT_CHECK_ARG(p->attn_ld >= p->S && p->ldp >= p->V, "pointer_mix: leading dimensions");
  TNT_CHECK_ARG(p->dtype == TNT_F32 || p->dtype == TNT_BF16, "pointer_mix: dtype");
  TNT_CHECK_ALIGN(aligned16(p->x) && aligned16(p->inp_embed) && p->ldx % 8 == 0 && (!p->inp_embed || p->lde % 8 == 0), "pointer_mix: x / inp_embed alignment");
  const size_t row_bytes = static_cast<size_t>(p->V) * sizeof(float);
  const int use_smem = row_bytes <= 220 * 1024;
  const size_t smem = use_smem ? row_bytes : 0;
  cudaStream_t st = as_stream(stream);
  if (p->dtype == TNT_BF16) {
    if (smem > 48 * 1024) TNT_CUDA(cudaFuncSetAttribute(pointer_mix_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    pointer_mix_kernel<bf16><<<p->B * p->T, kThreads, smem, st>>>(*p, use_smem);
  } else {
    if (smem > 48 * 1024) TNT_CUDA(cudaFuncSetAttribute(pointer_mix_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    pointer_mix_kernel<float><<<p->B * p->T, kThreads, smem, st>>>(*p, use_smem);
  }
  TNT_LAUNCHED();
  return TNT_OK;
}

extern "C" int tnt_nll_rows(const float* logprob, int64_t ldp, const int64_t* target, int rows, int pad, float* out2,
                            tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(logprob && target && out2 && rows >= 0, "nll_rows: bad arguments");
  cudaStream_t st = as_stream(stream);
  TNT_CUDA(cudaMemsetAsync(out2, 0, 2 * sizeof(float), st));
  if (rows == 0) return TNT_OK;
  nll_rows_kernel<<<std::min((rows + 255) / 256, 4 * num_sms()), 256, 0, st>>>(logprob, ldp, target, rows, pad, out2);
  TNT_LAUNCHED();
  return TNT_OK;
}
