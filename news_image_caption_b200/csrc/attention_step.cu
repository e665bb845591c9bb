// Decoding-step cross attention (kernel group 5): ONE query per (caption, head) against the cached K/V of a
// source -- the T == 1 case of MultiHeadAttention.forward during `_generate`
//   ttl/tell/modules/attention/multi_head.py:376-486 (scores, key_padding_mask -> -inf, fp32 softmax, PV)
//   ttl/tell/models/transformer_flattened.py:238-325 (the step loop that calls it once per layer and source).
// The work is a pure HBM stream: every K and V row of the caption is read exactly once (2 * (S+2) * d * 2 B per
// caption and layer) for ~4 FLOP per byte, so the kernel is built around coalesced 16-byte loads with many rows
// in flight, not around tensor cores:
//   CTA   = (caption b, group of 8 heads): the 8 heads' K (or V) values of one key are 1 KB contiguous in the
//           cache [S, B, L*2d]; padded keys (key_padding_mask) are never fetched (exp(-inf) = 0 exactly).
//   warp  = a strided subset of the keys; lane = (head = lane / 4, two 8-dim pieces of the 64 dims): scores by 2 shuffles,
//           running (max, sum, acc[16]) per lane in fp32 (flash-decoding), kUnroll keys loaded before use.
//   merge = the warps' partial (max, sum, acc) through shared memory; bf16 output, optional LSE.
#include <algorithm>

#include "common.cuh"

namespace tnt {
namespace {

constexpr int kHeadsPerCta = 8;
// 8 warps x kUnroll keys in flight per item: the launch lasts as long as its longest item (an article of ~514 live keys is a
// serial chain of n_live / (kWarps * kUnroll) load round trips).  Measured on the B = 256 generation: 4 warps 24.75 ms
// (4.83 TB/s in the attention launches), 8 warps 24.19 ms (5.22 TB/s), 16 warps 26.67 ms (4.14 TB/s: too few CTAs per SM).
#ifndef TNT_STEP_WARPS
#define TNT_STEP_WARPS 8
#endif
constexpr int kWarps = TNT_STEP_WARPS;
constexpr int kUnroll = 4;
constexpr int kMaxKeys = 2048;

__device__ __forceinline__ void unpack16(const uint4& a, const uint4& b, float (&v)[16]) {
  const uint32_t w[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

// up to TNT_MAX_SOURCES attention problems (the sources of one decoder layer) in ONE launch: blockIdx.z picks the source
// Work items = (source, caption, group of 8 heads), longest source first (the article's CTAs start before the image's short
// ones: 28.4 -> 27.6 ms per generation).  Optionally (`sched`, two zeroed ints the engine owns) the CTAs PULL items from an
// atomic counter and the last CTA to leave re-zeroes it; captions keep very different numbers of live keys (128..512).
struct StepBatch {
  tnt_xattn_params p[TNT_MAX_SOURCES];
  const int* order[TNT_MAX_SOURCES];      // optional per source: caption slots sorted by live keys, longest first
  int n;
  int* sched;
  float* scores;      // SCORES only: [B, H, S] log2-domain scores of source `scores_src` (live keys), for the head-averaged weights
  int scores_src;
};

template <bool SCORES>
__global__ void __launch_bounds__(kWarps * 32) xattn_step_kernel(const __grid_constant__ StepBatch batch) {
  __shared__ float sm_m[kWarps][32], sm_l[kWarps][32];
  __shared__ float sm_acc[kWarps][32][17];
  __shared__ uint8_t sm_mask[kMaxKeys];
  __shared__ uint16_t sm_live[kMaxKeys];
  __shared__ int sm_nlive, sm_item;
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_hg = batch.p[0].H / kHeadsPerCta, per_src = batch.p[0].B * n_hg, total = batch.n * per_src;
  for (int item = blockIdx.x;;) {
  if (batch.sched) {
    if (threadIdx.x == 0) sm_item = atomicAdd(&batch.sched[0], 1);
    __syncthreads();
    item = sm_item;
  }
  if (item >= total) break;
  const int src = item / per_src;
  const tnt_xattn_params& p = batch.p[src];
  const int rem = item % per_src, slot = rem / n_hg, hg = rem - slot * n_hg;
  const int b = batch.order[src] ? batch.order[src][slot] : slot;      // long captions are started first (1.7 waves of CTAs)
  // lane q of a head's 4 lanes owns dims [8q, 8q+8) and [32+8q, 32+8q+8): each of its two 16-byte loads then fills whole
  // 32-byte sectors together with its neighbours (dims [16q, 16q+16) made every sector cross L2->SM twice: ncu 770 MB per
  // launch for 385 MB of DRAM reads)
  const int h = hg * kHeadsPerCta + (lane >> 2), d0 = (lane & 3) * 8;
  const bf16* qp = reinterpret_cast<const bf16*>(p.q) + static_cast<long>(b) * p.q_stride_b + h * 64 + d0;
  const int bk = p.kv_group > 1 ? b / p.kv_group : b;     // beams of one caption share its K / V / mask
  const bf16* kp = reinterpret_cast<const bf16*>(p.k) + static_cast<long>(bk) * p.k_stride_b + h * 64 + d0;
  const bf16* vp = reinterpret_cast<const bf16*>(p.v) + static_cast<long>(bk) * p.v_stride_b + h * 64 + d0;
  const int S = p.S;
  const int SM = p.mask_mode == TNT_MASK_KEY_PADDING ? (p.mask_S > 0 ? min(p.mask_S, S) : S) : 0;   // keys the mask covers
  const uint8_t* mrow = p.mask + static_cast<long>(bk) * p.m_stride_b;

  float q[16];
  {
    const uint4 a = *reinterpret_cast<const uint4*>(qp), c = *reinterpret_cast<const uint4*>(qp + 32);
    unpack16(a, c, q);
#pragma unroll
    for (int i = 0; i < 16; ++i) q[i] *= p.scale * 1.4426950408889634f;     // scores in the log2 domain
  }
  float m = -INFINITY, l = 0.f, acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) acc[i] = 0.f;

  // live (un-padded) keys of this caption, compacted once per CTA: the mask bytes are fetched by all threads at
  // once (one memory latency), warp 0 turns them into an index list with ballots
  for (int i = threadIdx.x; i < S; i += kWarps * 32) sm_mask[i] = i < SM ? mrow[i] : uint8_t(0);
  __syncthreads();
  if (warp == 0) {
    int cnt = 0;
    for (int base = 0; base < S; base += 32) {
      const int kx = base + lane;
      const bool lv = kx < S && sm_mask[kx] == 0;
      const unsigned bal = __ballot_sync(0xffffffffu, lv);
      if (lv) sm_live[cnt + __popc(bal & ((1u << lane) - 1u))] = static_cast<uint16_t>(kx);
      cnt += __popc(bal);
    }
    if (lane == 0) sm_nlive = cnt;
  }
  __syncthreads();
  const int n_live = sm_nlive;

  // warp w takes live keys w, w + kWarps, ...; the K and V rows of kUnroll keys are requested before any is consumed
  for (int j0 = warp; j0 < n_live; j0 += kWarps * kUnroll) {
    uint4 kr[kUnroll][2], vr[kUnroll][2];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const int j = j0 + u * kWarps;
      if (j < n_live) {
        const long ky = sm_live[j];
        const bf16* kk = kp + ky * p.k_stride_s;
        const bf16* vv = vp + ky * p.v_stride_s;
        kr[u][0] = ldg_stream(kk);
        kr[u][1] = ldg_stream(kk + 32);
        vr[u][0] = ldg_stream(vv);
        vr[u][1] = ldg_stream(vv + 32);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (j0 + u * kWarps < n_live) {                    // warp-uniform
        float kf[16];
        unpack16(kr[u][0], kr[u][1], kf);
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < 16; ++i) s = fmaf(q[i], kf[i], s);
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        if (SCORES && src == batch.scores_src && (lane & 3) == 0)
          batch.scores[(static_cast<long>(b) * p.H + h) * S + sm_live[j0 + u * kWarps]] = s;
        const float mn = fmaxf(m, s);
        const float corr = exp2f(m - mn), pe = exp2f(s - mn);
        l = l * corr + pe;
        m = mn;
        float vf[16];
        unpack16(vr[u][0], vr[u][1], vf);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fmaf(pe, vf[i], acc[i] * corr);
      }
    }
  }

  sm_m[warp][lane] = m;
  sm_l[warp][lane] = l;
#pragma unroll
  for (int i = 0; i < 16; ++i) sm_acc[warp][lane][i] = acc[i];
  __syncthreads();
  if (warp == 0) {
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) M = fmaxf(M, sm_m[w][lane]);
    float Ls = 0.f, o[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i] = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) {
      const float mw = sm_m[w][lane];
      const float f = mw == -INFINITY ? 0.f : exp2f(mw - M);
      Ls += sm_l[w][lane] * f;
#pragma unroll
      for (int i = 0; i < 16; ++i) o[i] = fmaf(sm_acc[w][lane][i], f, o[i]);
    }
    const float inv = Ls > 0.f ? 1.f / Ls : 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) o[i] *= inv;
    bf16* op = reinterpret_cast<bf16*>(p.out) + static_cast<long>(b) * p.o_stride_b + h * 64 + d0;
    float t0[8], t1[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      t0[i] = o[i];
      t1[i] = o[8 + i];
    }
    store8(op, t0);
    store8(op + 32, t1);
    if (p.lse && (lane & 3) == 0)                        // natural-log LSE of the scaled, masked scores
      p.lse[static_cast<long>(b) * p.H + h] = Ls > 0.f ? (M + log2f(Ls)) * 0.6931471805599453f : INFINITY;
  }
  if (!batch.sched) break;
  __syncthreads();                                       // shared memory is reused by the next item
  }
  if (batch.sched && threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&batch.sched[1], 1) == static_cast<int>(gridDim.x) - 1) {
      batch.sched[0] = 0;
      batch.sched[1] = 0;
    }
  }
}

}  // namespace

int xattn_fwd_step_multi(const tnt_xattn_params* ps, int n, cudaStream_t st, int* sched = nullptr, const int* const* orders = nullptr,
                         float* scores = nullptr, int scores_src = -1, float* weights_out = nullptr);

bool xattn_step_supported(const tnt_xattn_params* p) {
  if (p->T != 1 || p->dtype != TNT_BF16 || p->D != 64 || (p->H % kHeadsPerCta) != 0 || p->n_extra != 0 || p->weights || p->drop.p > 0.f)
    return false;
  if (p->mask_mode != TNT_MASK_NONE && p->mask_mode != TNT_MASK_KEY_PADDING) return false;
  const long strides = p->q_stride_b | p->k_stride_s | p->k_stride_b | p->v_stride_s | p->v_stride_b | p->o_stride_b;
  return p->S >= 1 && p->S <= kMaxKeys && (strides % 8) == 0 && aligned16(p->q) && aligned16(p->k) && aligned16(p->v) && aligned16(p->out);
}

int xattn_fwd_step(const tnt_xattn_params* p, cudaStream_t st) { return xattn_fwd_step_multi(p, 1, st); }

namespace {
// order[rank] = b, rows ranked by their live-key count (tile_live[b] >> 16, see tile_bits_row), longest first; one CTA
__global__ void __launch_bounds__(1024) live_order_kernel(const uint32_t* tile_live, int B, int* order) {
  for (int b = threadIdx.x; b < B; b += 1024) {
    const uint32_t mine = tile_live[b] >> 16;
    int rank = 0;
    for (int o = 0; o < B; ++o) {
      const uint32_t other = tile_live[o] >> 16;
      rank += (other > mine) || (other == mine && o < b);
    }
    order[rank] = b;
  }
}

// head-averaged attention weights of a decode step (multi_head.py:478-484) from the step kernel's score dump:
// weights[b, s] = (1/H) sum_h exp2(score[b, h, s] - lse2[b, h]); 0 for padded keys
__global__ void __launch_bounds__(256) step_weights_kernel(const float* scores, const float* lse, const uint8_t* mask, long m_stride_b,
                                                           int B, int H, int S, int SM, float* weights) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= B * S) return;
  const int b = i / S, sx = i - b * S;
  float w = 0.f;
  if (!(sx < SM && mask[static_cast<long>(b) * m_stride_b + sx])) {
    for (int h = 0; h < H; ++h) {
      const float l = lse[static_cast<long>(b) * H + h];
      w += exp2f(scores[(static_cast<long>(b) * H + h) * S + sx] - l * 1.4426950408889634f);
    }
    w /= static_cast<float>(H);
  }
  weights[i] = w;
}
}  // namespace

int xattn_live_order(const uint32_t* tile_live, int B, int* order, cudaStream_t st) {
  live_order_kernel<<<1, 1024, 0, st>>>(tile_live, B, order);
  TNT_LAUNCHED();
  return TNT_OK;
}

// n problems with the same B and H (the sources of one decoder layer), one launch
// `scores` (scratch [B, H, S] fp32) + `weights_out` [B, S]: also return the head-averaged weights of problem `scores_src`
int xattn_fwd_step_multi(const tnt_xattn_params* ps, int n, cudaStream_t st, int* sched, const int* const* orders, float* scores,
                         int scores_src, float* weights_out) {
  TNT_CHECK_ARG(n >= 1 && n <= TNT_MAX_SOURCES, "xattn_step: %d problems", n);
  StepBatch batch;
  int order[TNT_MAX_SOURCES];
  for (int i = 0; i < n; ++i) {
    TNT_CHECK_ARG(ps[i].B == ps[0].B && ps[i].H == ps[0].H, "xattn_step: problems of one launch must share B and H");
    order[i] = i;
  }
  std::sort(order, order + n, [&](int a, int b) { return ps[a].S > ps[b].S; });      // longest source first
  for (int i = 0; i < n; ++i) {
    batch.p[i] = ps[order[i]];
    batch.order[i] = (orders && ps[order[i]].kv_group <= 1) ? orders[order[i]] : nullptr;
  }
  batch.n = n;
  batch.scores = nullptr;
  batch.scores_src = -1;
  if (scores && weights_out && scores_src >= 0) {
    TNT_CHECK_ARG(scores_src < n && ps[scores_src].kv_group <= 1 && ps[scores_src].lse, "xattn_step: weights need kv_group 1 and an LSE buffer");
    batch.scores = scores;
    for (int i = 0; i < n; ++i)
      if (order[i] == scores_src) batch.scores_src = i;
  }
  // measured (B = 256, generation bench): pulling items is no faster than one CTA per item (27.7 vs 27.5 ms) -- all 1024 CTAs
  // are co-resident anyway -- so the static mapping stays the default; TNT_STEP_ATTN_DYNAMIC=1 switches the queue on
  static const bool dyn = [] { const char* e = getenv("TNT_STEP_ATTN_DYNAMIC"); return e && e[0] == '1'; }();
  batch.sched = dyn ? sched : nullptr;
  variant_hit(V_XATTN_STEP);
  const int total = n * ps[0].B * (ps[0].H / kHeadsPerCta);
  static const int per_sm = [] { const char* e = getenv("TNT_STEP_ATTN_CTAS_PER_SM"); return e ? std::max(1, atoi(e)) : 4; }();
  dim3 grid(batch.sched ? std::min(total, per_sm * num_sms()) : total);
  if (batch.scores) {
    TNT_CUDA(launch_pdl(xattn_step_kernel<true>, grid, dim3(kWarps * 32), 0, st, 1, batch));
    TNT_LAUNCHED();
    const tnt_xattn_params& q = ps[scores_src];
    const int SM = q.mask_mode == TNT_MASK_KEY_PADDING ? (q.mask_S > 0 ? std::min(q.mask_S, q.S) : q.S) : 0;
    step_weights_kernel<<<(q.B * q.S + 255) / 256, 256, 0, st>>>(scores, q.lse, q.mask, q.m_stride_b, q.B, q.H, q.S, SM, weights_out);
    TNT_LAUNCHED();
    return TNT_OK;
  }
  TNT_CUDA(launch_pdl(xattn_step_kernel<false>, grid, dim3(kWarps * 32), 0, st, 1, batch));
  TNT_LAUNCHED();
  return TNT_OK;
}

}  // namespace tnt
