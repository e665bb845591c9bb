// Context assembly in front of the decoder:
//   ttl/tell/models/transformer_faces_objects.py:355-366  RoBERTa hidden-layer mix
//       X_article = sum_l softmax(bert_weight)[l] * hiddens[l]          (the reference first stacks the 25 [B,S,E]
//       tensors into [B,S,25,E] -- a full extra copy -- then multiplies and reduces), followed by .transpose(0, 1)
//   ttl/tell/models/transformer_faces_objects.py:371-377  NaN padding -> mask
//       mask = isnan(x).any(-1); x[mask] = 0                            (faces / objects), followed by .transpose(0, 1)
// Both are pure HBM streams; each input byte is read once and the result is written directly in the decoder's
// [S, B, E] layout and dtype.  Algorithmic bytes per article position: n_layers * E * e_in read + E * e_out written.
#include "common.cuh"

namespace tnt {
namespace {

constexpr int kThreads = 256;

// 8 consecutive elements kept RAW in registers (4 regs for bf16, 8 for fp32) so that many layers can be in flight per
// thread before any is consumed; streaming loads (read once: no L1 allocation)
template <typename TIn> struct Raw8;
template <> struct Raw8<bf16> {
  uint4 u;
  __device__ __forceinline__ void load(const bf16* p) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(u.x), "=r"(u.y), "=r"(u.z), "=r"(u.w)
                 : "l"(p));
  }
  __device__ __forceinline__ void fma_into(float w, float (&acc)[8]) const {
    const uint32_t q[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      acc[2 * i] = fmaf(__uint_as_float(q[i] << 16), w, acc[2 * i]);
      acc[2 * i + 1] = fmaf(__uint_as_float(q[i] & 0xffff0000u), w, acc[2 * i + 1]);
    }
  }
};
template <> struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w)
                 : "l"(p));
    asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p + 4));
  }
  __device__ __forceinline__ void fma_into(float w, float (&acc)[8]) const {
    acc[0] = fmaf(a.x, w, acc[0]); acc[1] = fmaf(a.y, w, acc[1]); acc[2] = fmaf(a.z, w, acc[2]); acc[3] = fmaf(a.w, w, acc[3]);
    acc[4] = fmaf(b.x, w, acc[4]); acc[5] = fmaf(b.y, w, acc[5]); acc[6] = fmaf(b.z, w, acc[6]); acc[7] = fmaf(b.w, w, acc[7]);
  }
};

// one thread = 8 consecutive features of one (b, s) position, all layers: n_layers independent 16/32-byte loads in
// flight per thread, fp32 accumulation in the reference's order (layer 0 first), transposed store
template <typename TIn, typename TOut, int UNROLL>
__global__ void __launch_bounds__(kThreads) context_mix_kernel(tnt_context_mix_params p) {
  __shared__ float wsm[TNT_MAX_MIX_LAYERS];
  if (threadIdx.x < 32) {                                 // softmax over the layer logits (F.softmax(bert_weight, 0))
    float mx = -INFINITY;
    for (int l = threadIdx.x; l < p.n_layers; l += 32) mx = fmaxf(mx, p.weight_logits[l]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int l = threadIdx.x; l < p.n_layers; l += 32) sum += expf(p.weight_logits[l] - mx);
    sum = warp_sum(sum);
    for (int l = threadIdx.x; l < p.n_layers; l += 32) wsm[l] = expf(p.weight_logits[l] - mx) / sum;
  }
  __syncthreads();
  const int nv = p.E >> 3;
  const long total = static_cast<long>(p.B) * p.S * nv;
  for (long i = blockIdx.x * static_cast<long>(kThreads) + threadIdx.x; i < total; i += static_cast<long>(gridDim.x) * kThreads) {
    const int vc = static_cast<int>(i % nv);
    const long pos = i / nv;                              // b * S + s
    const int s = static_cast<int>(pos % p.S), b = static_cast<int>(pos / p.S);
    const long off = pos * p.E + vc * 8;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = 0.f;
    for (int l0 = 0; l0 < p.n_layers; l0 += UNROLL) {
      Raw8<TIn> v[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u)
        if (l0 + u < p.n_layers) v[u].load(reinterpret_cast<const TIn*>(p.layers[l0 + u]) + off);
#pragma unroll
      for (int u = 0; u < UNROLL; ++u)
        if (l0 + u < p.n_layers) v[u].fma_into(wsm[l0 + u], acc);
    }
    TOut* o = reinterpret_cast<TOut*>(p.out) + (static_cast<long>(s) * p.B + b) * p.E + vc * 8;
    store8(o, acc);
  }
}

// one warp = one (b, i) row: NaN anywhere -> mask = 1 and the row is zeroed; output transposed to [n, B, d]
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(kThreads) nan_mask_kernel(tnt_nan_mask_params p) {
  const int lane = threadIdx.x & 31;
  const long row = blockIdx.x * static_cast<long>(kThreads / 32) + (threadIdx.x >> 5);
  if (row >= static_cast<long>(p.B) * p.n) return;
  const int i = static_cast<int>(row % p.n), b = static_cast<int>(row / p.n);
  const TIn* x = reinterpret_cast<const TIn*>(p.x) + row * p.d;
  bool bad = false;
  for (int c = lane; c < p.d; c += 32) bad |= isnan(to_f(x[c]));
  bad = __any_sync(0xffffffffu, bad);
  TOut* o = reinterpret_cast<TOut*>(p.out) + (static_cast<long>(i) * p.B + b) * p.d;
  for (int c = lane; c < p.d; c += 32) o[c] = from_f<TOut>(bad ? 0.f : to_f(x[c]));
  if (lane == 0) p.mask[row] = bad ? 1 : 0;
}

}  // namespace
}  // namespace tnt

using namespace tnt;

extern "C" int tnt_context_mix(const tnt_context_mix_params* p, tnt_stream stream) {
  TNT_CHECK_ARG(p && p->out && p->weight_logits, "context_mix: null pointer");
  TNT_CHECK_ARG(p->n_layers >= 1 && p->n_layers <= TNT_MAX_MIX_LAYERS, "context_mix: n_layers %d out of [1, %d]", p->n_layers,
                TNT_MAX_MIX_LAYERS);
  TNT_CHECK_ARG(p->B >= 0 && p->S >= 0 && p->E > 0 && p->E % 8 == 0, "context_mix: bad shape (E must be a multiple of 8)");
  TNT_CHECK_ARG((p->dtype_in == TNT_F32 || p->dtype_in == TNT_BF16) && (p->dtype_out == TNT_F32 || p->dtype_out == TNT_BF16),
                "context_mix: dtype");
  for (int l = 0; l < p->n_layers; ++l) {
    TNT_CHECK_ARG(p->layers[l], "context_mix: layer %d is null", l);
    TNT_CHECK_ALIGN((reinterpret_cast<uintptr_t>(p->layers[l]) & 31u) == 0, "context_mix: layer %d must be 32-byte aligned", l);
  }
  TNT_CHECK_ALIGN((reinterpret_cast<uintptr_t>(p->out) & 31u) == 0, "context_mix: out must be 32-byte aligned");
  const long total = static_cast<long>(p->B) * p->S * (p->E / 8);
  if (total == 0) return TNT_OK;
  const int grid = static_cast<int>(std::min<long>((total + kThreads - 1) / kThreads, 8L * num_sms()));
  cudaStream_t st = as_stream(stream);
  if (p->dtype_in == TNT_BF16 && p->dtype_out == TNT_BF16) context_mix_kernel<bf16, bf16, 13><<<grid, kThreads, 0, st>>>(*p);
  else if (p->dtype_in == TNT_BF16) context_mix_kernel<bf16, float, 13><<<grid, kThreads, 0, st>>>(*p);
  else if (p->dtype_out == TNT_BF16) context_mix_kernel<float, bf16, 7><<<grid, kThreads, 0, st>>>(*p);
  else context_mix_kernel<float, float, 7><<<grid, kThreads, 0, st>>>(*p);
  TNT_LAUNCHED();
  return TNT_OK;
}

extern "C" int tnt_nan_mask(const tnt_nan_mask_params* p, tnt_stream stream) {
  TNT_CHECK_ARG(p && p->x && p->out && p->mask, "nan_mask: null pointer");
  TNT_CHECK_ARG(p->B >= 0 && p->n >= 0 && p->d > 0, "nan_mask: bad shape");
  TNT_CHECK_ARG((p->dtype_in == TNT_F32 || p->dtype_in == TNT_BF16) && (p->dtype_out == TNT_F32 || p->dtype_out == TNT_BF16),
                "nan_mask: dtype");
  const long rows = static_cast<long>(p->B) * p->n;
  if (rows == 0) return TNT_OK;
  const int grid = static_cast<int>((rows + kThreads / 32 - 1) / (kThreads / 32));
  cudaStream_t st = as_stream(stream);
  if (p->dtype_in == TNT_BF16 && p->dtype_out == TNT_BF16) nan_mask_kernel<bf16, bf16><<<grid, kThreads, 0, st>>>(*p);
  else if (p->dtype_in == TNT_BF16) nan_mask_kernel<bf16, float><<<grid, kThreads, 0, st>>>(*p);
  else if (p->dtype_out == TNT_BF16) nan_mask_kernel<float, bf16><<<grid, kThreads, 0, st>>>(*p);
  else nan_mask_kernel<float, float><<<grid, kThreads, 0, st>>>(*p);
  TNT_LAUNCHED();
  return TNT_OK;
}
