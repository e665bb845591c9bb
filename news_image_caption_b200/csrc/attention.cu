// Multi-head (cross / causal self) attention over projected q, k, v -- flash-style online softmax,
// fp32 statistics.  SIMT version: exact fp32 math for the parity mode and the general fallback for
// head sizes / layouts the tcgen05 kernel (attention_tc.cu) does not take.
// Reference: ttl/tell/modules/attention/multi_head.py:376-486 (scores, masks, fp32 softmax, PV, head-
// averaged weights) and :355-374,416-427 (bias_k/bias_v row + zero row);
// final_roberta2/transformer_v2.py:213-222 (masked_fill(mask == 0, -1e9)).
#include "common.cuh"

namespace tnt {
namespace {

constexpr int kTQ = 32;       // queries per CTA
constexpr int kTK = 64;       // keys per tile
constexpr int kQW = 8;        // queries per warp
constexpr int kThreads = 128;
constexpr int kMaxDD = 4;     // head dim <= 128

struct Smem {
  float* qs; float* ks; float* vs; float* ps;
};

__device__ __forceinline__ int kv_batch(const tnt_xattn_params& p, int b) { return p.kv_group > 1 ? b / p.kv_group : b; }

template <typename T>
__device__ __forceinline__ void load_kv_tile(const tnt_xattn_params& p, const T* kp, const T* vp, int b, int h,
                                             int key0, float* ks, float* vs) {
  const int D = p.D, nv = D >> 3, S = p.S, S2 = p.S + p.n_extra;
  for (int i = threadIdx.x; i < kTK * nv; i += kThreads) {
    const int kk = i / nv, vc = i - kk * nv, key = key0 + kk;
    float a[8], c[8];
    if (key < S) {
      load8(kp + static_cast<long>(key) * p.k_stride_s + static_cast<long>(b) * p.k_stride_b + h * D + vc * 8, a);
      load8(vp + static_cast<long>(key) * p.v_stride_s + static_cast<long>(b) * p.v_stride_b + h * D + vc * 8, c);
    } else if (key == S && key < S2) {                 // learned bias row (multi_head.py:355-358)
      load8(p.bias_k + h * D + vc * 8, a);
      load8(p.bias_v + h * D + vc * 8, c);
    } else {                                           // zero row (multi_head.py:416-421) or tile padding
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] = c[j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) ks[kk * (D + 1) + vc * 8 + j] = a[j];
    *reinterpret_cast<float4*>(vs + kk * D + vc * 8) = make_float4(c[0], c[1], c[2], c[3]);
    *reinterpret_cast<float4*>(vs + kk * D + vc * 8 + 4) = make_float4(c[4], c[5], c[6], c[7]);
  }
}

// scores of this lane's two keys (lane, lane+32 of the tile) against the warp's kQW queries, masks applied
__device__ __forceinline__ void tile_scores(const tnt_xattn_params& p, const float* qs, const float* ks, int b,
                                            int t0, int qw, int nq, int key0, int lane, float (&s)[kQW][2]) {
  const int D = p.D, S2 = p.S + p.n_extra;
  const int S = p.mask_S > 0 ? min(p.mask_S, p.S) : p.S;     // keys the mask covers
#pragma unroll
  for (int qi = 0; qi < kQW; ++qi) s[qi][0] = s[qi][1] = 0.f;
  const float* k0 = ks + lane * (D + 1);
  const float* k1 = ks + (lane + 32) * (D + 1);
  for (int d = 0; d < D; ++d) {
    const float a = k0[d], c = k1[d];
#pragma unroll
    for (int qi = 0; qi < kQW; ++qi) {
      const float qv = qs[(qw + qi) * D + d];
      s[qi][0] = fmaf(qv, a, s[qi][0]);
      s[qi][1] = fmaf(qv, c, s[qi][1]);
    }
  }
#pragma unroll
  for (int e = 0; e < 2; ++e) {
    const int key = key0 + lane + 32 * e;
    if (key >= S2) {
#pragma unroll
      for (int qi = 0; qi < kQW; ++qi) s[qi][e] = -INFINITY;
    } else if (key < S && p.mask_mode == TNT_MASK_KEY_PADDING) {
      if (p.mask[static_cast<long>(kv_batch(p, b)) * p.m_stride_b + key]) {
#pragma unroll
        for (int qi = 0; qi < kQW; ++qi) s[qi][e] = -INFINITY;
      }
    } else if (key < S && p.mask_mode == TNT_MASK_V2) {
#pragma unroll
      for (int qi = 0; qi < kQW; ++qi) {
        if (qi < nq &&
            p.mask[static_cast<long>(b) * p.m_stride_b + static_cast<long>(t0 + qw + qi) * p.m_stride_t + key] == 0)
          s[qi][e] = -1e9f;
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads) xattn_fwd_kernel(tnt_xattn_params p, DropSpec dr) {
  extern __shared__ float sm[];
  const int D = p.D, ndd = (D + 31) / 32;
  float* qs = sm;
  float* ks = qs + kTQ * D;
  float* vs = ks + kTK * (D + 1);
  float* ps = vs + kTK * D;
  const int b = blockIdx.x, h = blockIdx.y, t0 = blockIdx.z * kTQ;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int TQ = min(kTQ, p.T - t0);
  const int qw = warp * kQW;
  const int nq = max(0, min(kQW, TQ - qw));
  const T* qp = reinterpret_cast<const T*>(p.q);
  const T* kp = reinterpret_cast<const T*>(p.k);
  const T* vp = reinterpret_cast<const T*>(p.v);
  const int S2 = p.S + p.n_extra;

  for (int i = threadIdx.x; i < kTQ * (D >> 3); i += kThreads) {
    const int qi = i / (D >> 3), vc = i - qi * (D >> 3);
    float a[8];
    if (qi < TQ) {
      load8(qp + static_cast<long>(t0 + qi) * p.q_stride_t + static_cast<long>(b) * p.q_stride_b + h * D + vc * 8, a);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] *= p.scale;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) qs[qi * D + vc * 8 + j] = a[j];
  }

  float m[kQW], l[kQW], acc[kQW][kMaxDD];
#pragma unroll
  for (int qi = 0; qi < kQW; ++qi) {
    m[qi] = -INFINITY;
    l[qi] = 0.f;
#pragma unroll
    for (int dd = 0; dd < kMaxDD; ++dd) acc[qi][dd] = 0.f;
  }
  float* pw = ps + warp * kQW * kTK;

  for (int key0 = 0; key0 < S2; key0 += kTK) {
    __syncthreads();
    load_kv_tile<T>(p, kp, vp, kv_batch(p, b), h, key0, ks, vs);
    __syncthreads();
    if (nq == 0) continue;
    float s[kQW][2];
    tile_scores(p, qs, ks, b, t0, qw, nq, key0, lane, s);
#pragma unroll
    for (int qi = 0; qi < kQW; ++qi) {
      const float tm = warp_max(fmaxf(s[qi][0], s[qi][1]));
      const float mn = fmaxf(m[qi], tm);
      const float ms = (mn == -INFINITY) ? 0.f : mn;
      const float corr = __expf(m[qi] - ms);
      const float p0 = __expf(s[qi][0] - ms), p1 = __expf(s[qi][1] - ms);
      l[qi] = l[qi] * corr + warp_sum(p0 + p1);
      m[qi] = mn;
#pragma unroll
      for (int dd = 0; dd < kMaxDD; ++dd) acc[qi][dd] *= corr;
      if (dr.thr) {      // dropout on the probabilities: the normaliser l stays that of the un-dropped soft-max
        const unsigned long long i0 =
            ((static_cast<unsigned long long>(b) * p.H + h) * p.T + t0 + qw + qi) * S2 + key0 + lane;
        pw[qi * kTK + lane] = drop_val(dr, i0, p0);
        pw[qi * kTK + lane + 32] = drop_val(dr, i0 + 32, p1);
      } else {
        pw[qi * kTK + lane] = p0;
        pw[qi * kTK + lane + 32] = p1;
      }
    }
    __syncwarp();
    for (int kk = 0; kk < kTK; ++kk) {
      float vv[kMaxDD];
#pragma unroll
      for (int dd = 0; dd < kMaxDD; ++dd) vv[dd] = (dd < ndd && lane + 32 * dd < D) ? vs[kk * D + lane + 32 * dd] : 0.f;
#pragma unroll
      for (int qi = 0; qi < kQW; ++qi) {
        const float pv = pw[qi * kTK + kk];
#pragma unroll
        for (int dd = 0; dd < kMaxDD; ++dd) acc[qi][dd] = fmaf(pv, vv[dd], acc[qi][dd]);
      }
    }
    __syncwarp();
  }

  T* op = reinterpret_cast<T*>(p.out);
#pragma unroll
  for (int qi = 0; qi < kQW; ++qi) {
    if (qi < nq) {
      const float inv = l[qi] > 0.f ? 1.f / l[qi] : 0.f;
      if (p.lse && lane == 0)
        p.lse[(static_cast<long>(b) * p.H + h) * p.T + t0 + qw + qi] = l[qi] > 0.f ? m[qi] + __logf(l[qi]) : INFINITY;
      T* o = op + static_cast<long>(t0 + qw + qi) * p.o_stride_t + static_cast<long>(b) * p.o_stride_b + h * D;
#pragma unroll
      for (int dd = 0; dd < kMaxDD; ++dd)
        if (dd < ndd && lane + 32 * dd < D) o[lane + 32 * dd] = from_f<T>(acc[qi][dd] * inv);
    }
  }

  // optional head-averaged probabilities (multi_head.py:478-482): second sweep over the keys
  if (p.weights) {
    const float invH = 1.f / p.H;
    for (int key0 = 0; key0 < S2; key0 += kTK) {
      __syncthreads();
      load_kv_tile<T>(p, kp, vp, kv_batch(p, b), h, key0, ks, vs);
      __syncthreads();
      if (nq == 0) continue;
      float s[kQW][2];
      tile_scores(p, qs, ks, b, t0, qw, nq, key0, lane, s);
#pragma unroll
      for (int qi = 0; qi < kQW; ++qi) {
        if (qi >= nq) continue;
        const float ms = (m[qi] == -INFINITY) ? 0.f : m[qi];
        const float inv = l[qi] > 0.f ? invH / l[qi] : 0.f;
        float* w = p.weights + (static_cast<long>(b) * p.T + t0 + qw + qi) * S2;
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int key = key0 + lane + 32 * e;
          if (key < S2) {
            float pv = __expf(s[qi][e] - ms) * inv;
            if (dr.thr)      // training: the reference averages the DROPPED probabilities (multi_head.py:463-482)
              pv = drop_val(dr, ((static_cast<unsigned long long>(b) * p.H + h) * p.T + t0 + qw + qi) * S2 + key, pv);
            atomicAdd(w + key, pv);
          }
        }
      }
    }
  }
}

}  // namespace

int xattn_fwd_simt(const tnt_xattn_params* p, cudaStream_t st) {
  const int D = p->D;
  const size_t smem = (static_cast<size_t>(kTQ) * D + kTK * (D + 1) + kTK * D + (kThreads / 32) * kQW * kTK) * 4;
  dim3 grid(p->B, p->H, (p->T + kTQ - 1) / kTQ);
  if (p->weights) TNT_CUDA(cudaMemsetAsync(p->weights, 0, sizeof(float) * p->B * p->T * (p->S + p->n_extra), st));
  if (p->dtype == TNT_BF16) {
    if (smem > 48 * 1024)
      TNT_CUDA(cudaFuncSetAttribute(xattn_fwd_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    xattn_fwd_kernel<bf16><<<grid, kThreads, smem, st>>>(*p, make_drop(p->drop));
  } else {
    if (smem > 48 * 1024)
      TNT_CUDA(cudaFuncSetAttribute(xattn_fwd_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    xattn_fwd_kernel<float><<<grid, kThreads, smem, st>>>(*p, make_drop(p->drop));
  }
  TNT_LAUNCHED();
  return TNT_OK;
}

int xattn_check(const tnt_xattn_params* p) {
  TNT_CHECK_ARG(p && p->q && p->k && p->v && p->out, "xattn: null pointer");
  TNT_CHECK_ARG(p->B > 0 && p->T >= 0 && p->S >= 0 && p->H > 0, "xattn: bad shape");
  TNT_CHECK_ARG(p->D % 8 == 0 && p->D <= 128, "xattn: head dim %d must be a multiple of 8, <= 128", p->D);
  TNT_CHECK_ARG(p->drop.p >= 0.f && p->drop.p < 1.f, "xattn: dropout p = %f", p->drop.p);
  TNT_CHECK_ARG(p->n_extra == 0 || p->n_extra == 2, "xattn: n_extra must be 0 or 2");
  TNT_CHECK_ARG(p->n_extra == 0 || (p->bias_k && p->bias_v), "xattn: bias_k / bias_v missing");
  TNT_CHECK_ARG(p->mask_mode == TNT_MASK_NONE || p->mask, "xattn: mask missing");
  TNT_CHECK_ARG(p->S + p->n_extra > 0, "xattn: no keys");
  const long m8 = p->q_stride_t | p->q_stride_b | p->k_stride_s | p->k_stride_b | p->v_stride_s | p->v_stride_b;
  TNT_CHECK_ALIGN(m8 % 8 == 0 && aligned16(p->q) && aligned16(p->k) && aligned16(p->v),
                  "xattn: q/k/v strides must be multiples of 8 elements and 16-byte aligned");
  return TNT_OK;
}

bool xattn_tc_supported(const tnt_xattn_params* p);
int xattn_fwd_tc(const tnt_xattn_params* p, cudaStream_t st);
bool xattn_step_supported(const tnt_xattn_params* p);
int xattn_fwd_step(const tnt_xattn_params* p, cudaStream_t st);
static bool g_force_simt_attn = false;
bool xattn_force_simt() { return g_force_simt_attn; }
bool xattn_short_supported(const tnt_xattn_params* p);                                  // attention_short.cu
int xattn_short_multi(const tnt_xattn_params* ps, int n, cudaStream_t st);

int xattn_fwd_dispatch(const tnt_xattn_params* p, cudaStream_t st) {
  if (!g_force_simt_attn && xattn_step_supported(p)) return xattn_fwd_step(p, st);      // T == 1: HBM-streaming decode kernel
  if (!g_force_simt_attn && xattn_short_supported(p)) return xattn_short_multi(p, 1, st);   // <= 128 keys: warp-level mma kernel
  if (!g_force_simt_attn && xattn_tc_supported(p)) {
    variant_hit(V_XATTN_TC);
    if (p->tile_live) variant_hit(V_XATTN_TC_TILESKIP);
    if (p->weights) variant_hit(V_XATTN_TC_WEIGHTS);
    return xattn_fwd_tc(p, st);
  }
  variant_hit(V_XATTN_SIMT);
  return xattn_fwd_simt(p, st);
}

}  // namespace tnt

extern "C" void tnt_debug_force_simt_attention(int on) { tnt::g_force_simt_attn = on != 0; }

extern "C" int tnt_xattn_fwd(const tnt_xattn_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_TRY(xattn_check(p));
  if (p->T == 0) return TNT_OK;
  return xattn_fwd_dispatch(p, as_stream(stream));
}
