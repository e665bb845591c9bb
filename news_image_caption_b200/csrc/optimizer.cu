// Fused optimizer step over the flat fp32 parameter / gradient arenas: BertAdam as the
// reference configures it (ttl/expt/nytimes/9_transformer_objects/config.yaml:126-136 -> AllenNLP `bert_adam`
// = pytorch_pretrained_bert.BertAdam, un-vendored third party; published algorithm restated here):
//   per PARAMETER TENSOR gradient clipping to max_grad_norm, m = b1 m + (1-b1) g, v = b2 v + (1-b2) g^2,
//   update = m / (sqrt(v) + eps) + weight_decay * p  (no bias correction),  p -= lr_scheduled * update.
// Replaces the roles of apex multi_tensor_l2norm / multi_tensor_scale / multi_tensor_adam
// (ttl/libs/apex/csrc/multi_tensor_{l2norm_kernel,scale_kernel,adam}.cu) on this path.  HBM-bound:
// 16 B read + 12 B written per parameter.
#include <algorithm>

#include "common.cuh"

namespace tnt {
namespace {

constexpr int kChunk = 64;      // arena segments are padded to 64 elements; one segment id per chunk

// sum of squares per segment (segment = one parameter tensor).  Every warp owns a CONTIGUOUS run of chunks and
// flushes one atomicAdd per segment change, so a 150 M-element arena costs a few thousand atomics.
__global__ void __launch_bounds__(256) seg_sqnorm_kernel(const float* g, long c_begin, long c_end, const int* seg_of_chunk,
                                                         float* seg_sq) {
  const int lane = threadIdx.x & 31;
  const long warp = (blockIdx.x * 256L + threadIdx.x) >> 5;
  const long nw = (gridDim.x * 256L) >> 5;
  const long per = (((c_end - c_begin) + nw - 1) / nw + 7) & ~7L;       // runs of 8 chunks: four 16-byte loads per lane
  const long c0 = c_begin + warp * per, c1 = min(c_end, c0 + per);
  float acc = 0.f;
  int cur = -1;
  auto flush = [&](int next) {
    acc = warp_sum(acc);
    if (lane == 0 && cur >= 0 && acc != 0.f) atomicAdd(seg_sq + cur, acc);
    acc = 0.f;
    cur = next;
  };
  long c = c0;
  for (; c + 8 <= c1; c += 8) {
    const int sid = seg_of_chunk[c + (lane & 7)];
    const int first = __shfl_sync(0xffffffffu, sid, 0);
    if (__all_sync(0xffffffffu, sid == first)) {
      if (first != cur) flush(first);
      const float4* p = reinterpret_cast<const float4*>(g + c * kChunk);
      const float4 v0 = p[lane], v1 = p[lane + 32], v2 = p[lane + 64], v3 = p[lane + 96];
      acc = fmaf(v0.x, v0.x, fmaf(v0.y, v0.y, fmaf(v0.z, v0.z, fmaf(v0.w, v0.w, acc))));
      acc = fmaf(v1.x, v1.x, fmaf(v1.y, v1.y, fmaf(v1.z, v1.z, fmaf(v1.w, v1.w, acc))));
      acc = fmaf(v2.x, v2.x, fmaf(v2.y, v2.y, fmaf(v2.z, v2.z, fmaf(v2.w, v2.w, acc))));
      acc = fmaf(v3.x, v3.x, fmaf(v3.y, v3.y, fmaf(v3.z, v3.z, fmaf(v3.w, v3.w, acc))));
    } else {                                     // a tensor boundary inside the run: chunk by chunk
      for (int j = 0; j < 8; ++j) {
        const int sg = __shfl_sync(0xffffffffu, sid, j);
        if (sg != cur) flush(sg);
        const float2 v = *reinterpret_cast<const float2*>(g + (c + j) * kChunk + lane * 2);
        acc = fmaf(v.x, v.x, fmaf(v.y, v.y, acc));
      }
    }
  }
  for (; c < c1; ++c) {
    const int sg = seg_of_chunk[c];
    if (sg != cur) flush(sg);
    const float2 v = *reinterpret_cast<const float2*>(g + c * kChunk + lane * 2);
    acc = fmaf(v.x, v.x, fmaf(v.y, v.y, acc));
  }
  flush(-1);
}

__global__ void __launch_bounds__(256) bert_adam_kernel(float* p, const float* g, float* m, float* v, long n_chunks,
                                                        const int* seg_of_chunk, const float* seg_sq,
                                                        const float* gscale_dev, float gscale_host, float max_norm,
                                                        float lr, float b1, float b2, float eps, float wd) {
  const float gs = gscale_host * (gscale_dev ? *gscale_dev : 1.f);
  const long total4 = n_chunks * (kChunk / 4);
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total4; i += gridDim.x * 256L) {
    const long c = i / (kChunk / 4);
    float coef = gs;
    if (max_norm > 0.f && seg_sq) {
      const float norm = sqrtf(seg_sq[seg_of_chunk[c]]) * gs;
      const float clip = max_norm / (norm + 1e-6f);          // torch.nn.utils.clip_grad_norm_
      if (clip < 1.f) coef *= clip;
    }
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 pp = reinterpret_cast<float4*>(p)[i], mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    float* pa = &pp.x; float* ma = &mm.x; float* va = &vv.x;
    const float* ga = &gg.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gr = ga[j] * coef;
      ma[j] = b1 * ma[j] + (1.f - b1) * gr;
      va[j] = b2 * va[j] + (1.f - b2) * gr * gr;
      const float upd = ma[j] / (sqrtf(va[j]) + eps) + wd * pa[j];
      pa[j] -= lr * upd;
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
}
}  // namespace
}  // namespace tnt

extern "C" int tnt_bert_adam_step(const tnt_adam_params* a, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(a && a->params && a->grads && a->m && a->v, "bert_adam: null pointer");
  TNT_CHECK_ARG(a->n % kChunk == 0, "bert_adam: arena length must be a multiple of 64");
  TNT_CHECK_ARG(a->max_grad_norm <= 0.f || (a->seg_of_chunk && a->seg_sq && a->n_segments > 0),
                "bert_adam: per-tensor clipping needs the segment table");
  TNT_CHECK_ALIGN(aligned16(a->params) && aligned16(a->grads) && aligned16(a->m) && aligned16(a->v), "bert_adam: alignment");
  if (a->n == 0) return TNT_OK;
  cudaStream_t st = as_stream(stream);
  const long n_chunks = a->n / kChunk;
  int grid = static_cast<int>(std::min<long>((a->n / 4 + 255) / 256, 16L * num_sms()));
  if (a->max_ctas > 0) grid = std::min(grid, a->max_ctas);
  if (a->max_grad_norm > 0.f && !a->norms_ready) {
    TNT_CUDA(cudaMemsetAsync(a->seg_sq, 0, sizeof(float) * a->n_segments, st));
    seg_sqnorm_kernel<<<grid, 256, 0, st>>>(a->grads, 0, n_chunks, a->seg_of_chunk, a->seg_sq);
    TNT_LAUNCHED();
  }
  bert_adam_kernel<<<grid, 256, 0, st>>>(a->params, a->grads, a->m, a->v, n_chunks, a->seg_of_chunk, a->seg_sq,
                                         a->grad_scale_dev, a->grad_scale != 0.f ? a->grad_scale : 1.f, a->max_grad_norm,
                                         a->lr, a->beta1, a->beta2, a->eps, a->weight_decay);
  TNT_LAUNCHED();
  return TNT_OK;
}

extern "C" int tnt_grad_sqnorm(const float* grads, const int* seg_of_chunk, float* seg_sq, int64_t first, int64_t n,
                               int max_ctas, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(grads && seg_of_chunk && seg_sq && first >= 0 && n >= 0, "grad_sqnorm: bad arguments");
  TNT_CHECK_ARG(first % kChunk == 0 && n % kChunk == 0, "grad_sqnorm: the range must be whole 64-element chunks");
  TNT_CHECK_ALIGN(aligned16(grads), "grad_sqnorm: alignment");
  if (n == 0) return TNT_OK;
  int grid = static_cast<int>(std::min<long>((n / 4 + 255) / 256, 16L * num_sms()));
  if (max_ctas > 0) grid = std::min(grid, max_ctas);
  seg_sqnorm_kernel<<<grid, 256, 0, as_stream(stream)>>>(grads, first / kChunk, (first + n) / kChunk, seg_of_chunk, seg_sq);
  TNT_LAUNCHED();
  return TNT_OK;
}

// ------------------------------------------------------------------------------ torch.optim.Adam + accumulate
namespace tnt {
namespace {
// torch.optim.Adam (the final* scripts: Adam(lr, betas=(0.9, 0.98), eps=1e-9) under the Noam schedule,
// final_roberta2/custom_v4.py:421-430, 491-498): g' = g + wd p; m, v moments; bias-corrected step
//   p -= lr / (1 - b1^t) * m / (sqrt(v) / sqrt(1 - b2^t) + eps)
__global__ void __launch_bounds__(256) adam_kernel(float* p, const float* g, float* m, float* v, long n4, const float* gscale_dev,
                                                   float gscale, float lr, float b1, float b2, float eps, float wd, float c1, float c2) {
  const float coef = gscale * (gscale_dev ? *gscale_dev : 1.f);
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n4; i += gridDim.x * 256L) {
    const float4 gg = reinterpret_cast<const float4*>(g)[i];
    float4 pp = reinterpret_cast<float4*>(p)[i], mm = reinterpret_cast<float4*>(m)[i], vv = reinterpret_cast<float4*>(v)[i];
    float* pa = &pp.x; float* ma = &mm.x; float* va = &vv.x;
    const float* ga = &gg.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float gr = ga[j] * coef + wd * pa[j];
      ma[j] = b1 * ma[j] + (1.f - b1) * gr;
      va[j] = b2 * va[j] + (1.f - b2) * gr * gr;
      pa[j] -= (lr / c1) * ma[j] / (sqrtf(va[j]) / c2 + eps);
    }
    reinterpret_cast<float4*>(p)[i] = pp;
    reinterpret_cast<float4*>(m)[i] = mm;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
}
// dst += alpha * (*alpha_dev) * src: autograd's ".grad += upstream * local gradient" for the fused engine backward
__global__ void __launch_bounds__(256) axpy_kernel(float* dst, const float* src, long n4, float alpha, const float* alpha_dev) {
  const float a = alpha * (alpha_dev ? *alpha_dev : 1.f);
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n4; i += gridDim.x * 256L) {
    const float4 s = reinterpret_cast<const float4*>(src)[i];
    float4 d = reinterpret_cast<float4*>(dst)[i];
    d.x = fmaf(a, s.x, d.x); d.y = fmaf(a, s.y, d.y); d.z = fmaf(a, s.z, d.z); d.w = fmaf(a, s.w, d.w);
    reinterpret_cast<float4*>(dst)[i] = d;
  }
}
}  // namespace
}  // namespace tnt

extern "C" int tnt_adam_step(float* params, const float* grads, float* m, float* v, int64_t n, const float* grad_scale_dev,
                             float grad_scale, float lr, float beta1, float beta2, float eps, float weight_decay, int step,
                             tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(params && grads && m && v && n >= 0 && n % 4 == 0 && step >= 1, "adam: bad arguments (n %% 4 == 0, step >= 1)");
  TNT_CHECK_ALIGN(aligned16(params) && aligned16(grads) && aligned16(m) && aligned16(v), "adam: alignment");
  if (n == 0) return TNT_OK;
  const float c1 = 1.f - powf(beta1, static_cast<float>(step));
  const float c2 = sqrtf(1.f - powf(beta2, static_cast<float>(step)));
  const int grid = static_cast<int>(std::min<long>((n / 4 + 255) / 256, 16L * num_sms()));
  adam_kernel<<<grid, 256, 0, as_stream(stream)>>>(params, grads, m, v, n / 4, grad_scale_dev, grad_scale != 0.f ? grad_scale : 1.f, lr,
                                                   beta1, beta2, eps, weight_decay, c1, c2);
  TNT_LAUNCHED();
  return TNT_OK;
}

extern "C" int tnt_grad_accumulate(float* dst, const float* src, int64_t n, float alpha, const float* alpha_dev, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(dst && src && n >= 0 && n % 4 == 0, "grad_accumulate: bad arguments (n %% 4 == 0)");
  TNT_CHECK_ALIGN(aligned16(dst) && aligned16(src), "grad_accumulate: alignment");
  if (n == 0) return TNT_OK;
  const int grid = static_cast<int>(std::min<long>((n / 4 + 255) / 256, 16L * num_sms()));
  axpy_kernel<<<grid, 256, 0, as_stream(stream)>>>(dst, src, n / 4, alpha, alpha_dev);
  TNT_LAUNCHED();
  return TNT_OK;
}
