// Multi-head attention backward (kernel group 4), SIMT fp32 math, any head dim <= 128.
// Differentiates ttl/tell/modules/attention/multi_head.py:376-477 (scaled QK^T, masks, fp32 softmax, PV) and
// final_roberta2/transformer_v2.py:213-222.  Per (batch row b, head h) one CTA:
//   pass A  row statistics: lse[t] = logsumexp_k(s[t,k]) and delta[t] = sum_d dO[t,d] O[t,d]
//   pass B  per 32-key chunk, per 32-query tile:  p = exp(s - lse);  dp = dO V^T;  ds = p (dp - delta);
//           dV += p^T dO;  dK += scale ds^T Q;  dQ += scale ds K
// dK / dV of a key chunk live in registers across the query tiles; dQ is accumulated in a CTA-private fp32
// scratch slice (no atomics -> deterministic) and converted at the end.
#include "backward.cuh"

namespace tnt {
namespace {

constexpr int kT = 32;          // tile edge (queries and keys)
constexpr int kThreads = 256;

struct BwdArgs {
  tnt_xattn_params f;
  const void* dout; void* dq; void* dk; void* dv;
  float* scratch;               // [B, H, T, D] fp32
  DropSpec drop;                // the forward's dropout on the probabilities
  const float* dweights;        // optional [B, T, S] fp32: gradient w.r.t. the head-AVERAGED (dropped) probabilities the
                                // forward returned as `weights` (the copy heads differentiate through them)
};

template <typename T>
__device__ __forceinline__ void load_tile(const T* base, long stride_row, int row0, int nrows_valid, int D, float* dst,
                                          float mul) {
  const int nv = D >> 3, ld = D + 1;
  for (int i = threadIdx.x; i < kT * nv; i += kThreads) {
    const int r = i / nv, vc = i - r * nv;
    float v[8];
    if (r < nrows_valid) {
      load8(base + static_cast<long>(row0 + r) * stride_row + vc * 8, v);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) dst[r * ld + vc * 8 + j] = v[j] * mul;
  }
}

// masked score replacement for (query t, key): returns true if the key contributes nothing (-inf)
__device__ __forceinline__ float apply_mask(const tnt_xattn_params& p, int b, int t, int key, float s) {
  if (key >= p.S) return -INFINITY;
  const int MS = p.mask_S > 0 ? min(p.mask_S, p.S) : p.S;
  if (key < MS) {
    if (p.mask_mode == TNT_MASK_KEY_PADDING) {
      if (p.mask[static_cast<long>(b) * p.m_stride_b + key]) return -INFINITY;
    } else if (p.mask_mode == TNT_MASK_V2) {
      if (p.mask[static_cast<long>(b) * p.m_stride_b + static_cast<long>(t) * p.m_stride_t + key] == 0) return -1e9f;
    }
  }
  return s;
}

template <typename T, int MAXI>
__global__ void __launch_bounds__(kThreads) xattn_bwd_kernel(BwdArgs a) {
  extern __shared__ float sm[];
  const tnt_xattn_params& p = a.f;
  const int D = p.D, ld = D + 1, Tn = p.T, S = p.S;
  float* Qs = sm;
  float* dOs = Qs + kT * ld;
  float* Ks = dOs + kT * ld;
  float* Vs = Ks + kT * ld;
  float* Ps = Vs + kT * ld;          // [32][33]
  float* dSs = Ps + kT * 33;
  float* lse = dSs + kT * 33;        // [T]
  float* delta = lse + Tn;           // [T]
  const int b = blockIdx.x, h = blockIdx.y;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const T* qp = reinterpret_cast<const T*>(p.q) + static_cast<long>(b) * p.q_stride_b + h * D;
  const T* kp = reinterpret_cast<const T*>(p.k) + static_cast<long>(b) * p.k_stride_b + h * D;
  const T* vp = reinterpret_cast<const T*>(p.v) + static_cast<long>(b) * p.v_stride_b + h * D;
  const T* op = reinterpret_cast<const T*>(p.out) + static_cast<long>(b) * p.o_stride_b + h * D;
  const T* dop = reinterpret_cast<const T*>(a.dout) + static_cast<long>(b) * p.o_stride_b + h * D;
  float* scr = a.scratch + (static_cast<long>(b) * p.H + h) * Tn * D;
  const int n_qt = (Tn + kT - 1) / kT, n_kc = (S + kT - 1) / kT;
  const int qi = tid >> 3, kg = tid & 7;         // score mapping: query qi, keys kg*4 .. kg*4+3

  // ---------------------------------------------------------------- pass A: lse and delta per query
  for (int qt = 0; qt < n_qt; ++qt) {
    const int t0 = qt * kT, nq = min(kT, Tn - t0);
    __syncthreads();
    load_tile<T>(qp, p.q_stride_t, t0, nq, D, Qs, p.scale);
    load_tile<T>(op, p.o_stride_t, t0, nq, D, Ks, 1.f);       // O staged in Ks, dO in dOs
    load_tile<T>(dop, p.o_stride_t, t0, nq, D, dOs, 1.f);
    __syncthreads();
    {
      // delta: warp w handles queries w*4 .. w*4+3
      for (int q = warp * 4; q < warp * 4 + 4; ++q) {
        float s = 0.f;
        for (int d = lane; d < D; d += 32) s = fmaf(dOs[q * ld + d], Ks[q * ld + d], s);
        s = warp_sum(s);
        if (lane == 0 && q < nq) delta[t0 + q] = s;
      }
    }
    float m = -INFINITY, l = 0.f;                // running stats of query qi over this thread's keys
    for (int kc = 0; kc < n_kc; ++kc) {
      __syncthreads();
      load_tile<T>(kp, p.k_stride_s, kc * kT, min(kT, S - kc * kT), D, Ks, 1.f);
      __syncthreads();
      float s[4] = {0.f, 0.f, 0.f, 0.f};
      for (int d = 0; d < D; ++d) {
        const float qv = Qs[qi * ld + d];
#pragma unroll
        for (int e = 0; e < 4; ++e) s[e] = fmaf(qv, Ks[(kg * 4 + e) * ld + d], s[e]);
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float z = qi < nq ? apply_mask(p, b, t0 + qi, kc * kT + kg * 4 + e, s[e]) : -INFINITY;
        if (z > m) {
          l = l * __expf(m - z) + 1.f;
          m = z;
        } else if (z != -INFINITY) {
          l += __expf(z - m);
        }
      }
    }
    // merge the 8 key groups of query qi (consecutive lanes)
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o), l2 = __shfl_xor_sync(0xffffffffu, l, o);
      const float mn = fmaxf(m, m2);
      const float ms = mn == -INFINITY ? 0.f : mn;
      l = l * __expf(m - ms) + l2 * __expf(m2 - ms);
      m = mn;
    }
    if (kg == 0 && qi < nq) lse[t0 + qi] = l > 0.f ? m + __logf(l) : INFINITY;   // +inf: fully masked row -> p = 0
    if (a.dweights) {
      // the returned weights w[t,k] = (1/H) sum_h pd_h[t,k] add dW[t,k] / H to every head's dPd; the softmax backward's
      // row term then gains sum_k pd[t,k] dW[t,k] / H -- one more sweep over the keys now that lse is known
      const float Lq = l > 0.f ? m + __logf(l) : INFINITY;
      float extra = 0.f;
      for (int kc = 0; kc < n_kc; ++kc) {
        __syncthreads();
        load_tile<T>(kp, p.k_stride_s, kc * kT, min(kT, S - kc * kT), D, Ks, 1.f);
        __syncthreads();
        float s[4] = {0.f, 0.f, 0.f, 0.f};
        for (int d = 0; d < D; ++d) {
          const float qv = Qs[qi * ld + d];
#pragma unroll
          for (int e = 0; e < 4; ++e) s[e] = fmaf(qv, Ks[(kg * 4 + e) * ld + d], s[e]);
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int key = kc * kT + kg * 4 + e;
          const float z = qi < nq ? apply_mask(p, b, t0 + qi, key, s[e]) : -INFINITY;
          if (qi < nq && z != -INFINITY && Lq != INFINITY) {
            float pr = __expf(z - Lq);
            if (a.drop.thr) pr = drop_val(a.drop, ((static_cast<unsigned long long>(b) * p.H + h) * Tn + t0 + qi) * S + key, pr);
            extra = fmaf(pr, a.dweights[(static_cast<long>(b) * Tn + t0 + qi) * S + key], extra);
          }
        }
      }
#pragma unroll
      for (int o = 1; o < 8; o <<= 1) extra += __shfl_xor_sync(0xffffffffu, extra, o);
      __syncthreads();
      if (kg == 0 && qi < nq) delta[t0 + qi] += extra / static_cast<float>(p.H);
    }
  }

  // ---------------------------------------------------------------- pass B
  const int kr = tid >> 3, dg = tid & 7;          // gradient mapping: row kr (key or query), dims dg + 8 i
  for (int kc = 0; kc < n_kc; ++kc) {
    const int key0 = kc * kT, nk = min(kT, S - key0);
    float dk[MAXI], dv[MAXI];
#pragma unroll
    for (int i = 0; i < MAXI; ++i) dk[i] = dv[i] = 0.f;
    __syncthreads();
    load_tile<T>(kp, p.k_stride_s, key0, nk, D, Ks, 1.f);
    load_tile<T>(vp, p.v_stride_s, key0, nk, D, Vs, 1.f);
    for (int qt = 0; qt < n_qt; ++qt) {
      const int t0 = qt * kT, nq = min(kT, Tn - t0);
      __syncthreads();
      load_tile<T>(qp, p.q_stride_t, t0, nq, D, Qs, p.scale);
      load_tile<T>(dop, p.o_stride_t, t0, nq, D, dOs, 1.f);
      __syncthreads();
      {
        float s[4] = {0.f, 0.f, 0.f, 0.f}, dp[4] = {0.f, 0.f, 0.f, 0.f};
        for (int d = 0; d < D; ++d) {
          const float qv = Qs[qi * ld + d], gv = dOs[qi * ld + d];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            s[e] = fmaf(qv, Ks[(kg * 4 + e) * ld + d], s[e]);
            dp[e] = fmaf(gv, Vs[(kg * 4 + e) * ld + d], dp[e]);
          }
        }
        const bool qok = qi < nq;
        const float L = qok ? lse[t0 + qi] : INFINITY, dl = qok ? delta[t0 + qi] : 0.f;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int kk = kg * 4 + e;
          const float z = qok ? apply_mask(p, b, t0 + qi, key0 + kk, s[e]) : -INFINITY;
          const float pr = (qok && z != -INFINITY && L != INFINITY) ? __expf(z - L) : 0.f;
          // dropout mask m (0 or 1/(1-p)) of the forward: out = (P . m) V, so dV takes P.m and dS = P (m dP - delta)
          float mk = 1.f;
          if (a.drop.thr)
            mk = drop_val(a.drop, ((static_cast<unsigned long long>(b) * p.H + h) * Tn + t0 + qi) * S + key0 + kk, 1.f);
          Ps[qi * 33 + kk] = pr * mk;
          float dpe = dp[e];
          if (a.dweights && qok && key0 + kk < S) dpe += a.dweights[(static_cast<long>(b) * Tn + t0 + qi) * S + key0 + kk] / static_cast<float>(p.H);
          dSs[qi * 33 + kk] = pr * (mk * dpe - dl);
        }
      }
      __syncthreads();
      // dV[k] += sum_q P[q][k] dO[q];  dK[k] += sum_q dS[q][k] Q[q] (Q already carries the scale)
      for (int q = 0; q < kT; ++q) {
        const float pr = Ps[q * 33 + kr], ds = dSs[q * 33 + kr];
#pragma unroll
        for (int i = 0; i < MAXI; ++i) {
          const int d = dg + 8 * i;
          if (d < D) {
            dv[i] = fmaf(pr, dOs[q * ld + d], dv[i]);
            dk[i] = fmaf(ds, Qs[q * ld + d], dk[i]);
          }
        }
      }
      // dQ[q] += scale * sum_k dS[q][k] K[k]
      float dq[MAXI];
#pragma unroll
      for (int i = 0; i < MAXI; ++i) dq[i] = 0.f;
      for (int k = 0; k < kT; ++k) {
        const float ds = dSs[kr * 33 + k];
#pragma unroll
        for (int i = 0; i < MAXI; ++i) {
          const int d = dg + 8 * i;
          if (d < D) dq[i] = fmaf(ds, Ks[k * ld + d], dq[i]);
        }
      }
      if (kr < nq) {
        float* o = scr + static_cast<long>(t0 + kr) * D;
#pragma unroll
        for (int i = 0; i < MAXI; ++i) {
          const int d = dg + 8 * i;
          if (d < D) o[d] = (kc == 0 ? 0.f : o[d]) + dq[i] * p.scale;
        }
      }
    }
    if (kr < nk) {
      T* dko = reinterpret_cast<T*>(a.dk) + static_cast<long>(key0 + kr) * p.k_stride_s + static_cast<long>(b) * p.k_stride_b + h * D;
      T* dvo = reinterpret_cast<T*>(a.dv) + static_cast<long>(key0 + kr) * p.v_stride_s + static_cast<long>(b) * p.v_stride_b + h * D;
#pragma unroll
      for (int i = 0; i < MAXI; ++i) {
        const int d = dg + 8 * i;
        if (d < D) {
          dko[d] = from_f<T>(dk[i]);
          dvo[d] = from_f<T>(dv[i]);
        }
      }
    }
  }
  __syncthreads();
  // ---------------------------------------------------------------- dQ: scratch -> output dtype
  T* dqo = reinterpret_cast<T*>(a.dq) + static_cast<long>(b) * p.q_stride_b + h * D;
  for (int i = tid; i < Tn * (D >> 3); i += kThreads) {
    const int t = i / (D >> 3), vc = i - t * (D >> 3);
    float v[8];
    load8(scr + static_cast<long>(t) * D + vc * 8, v);
    if (n_kc == 0) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
    }
    store8(dqo + static_cast<long>(t) * p.q_stride_t + vc * 8, v);
  }
}

template <typename T, int MAXI>
int launch(const BwdArgs& a, size_t smem, cudaStream_t st) {
  auto kern = xattn_bwd_kernel<T, MAXI>;
  if (smem > 48 * 1024) TNT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
  dim3 grid(a.f.B, a.f.H);
  kern<<<grid, kThreads, smem, st>>>(a);
  TNT_LAUNCHED();
  return TNT_OK;
}
}  // namespace

int xattn_check(const tnt_xattn_params* p);
bool xattn_bwd_tc_supported(const XattnBwd* p);
int xattn_bwd_tc(const XattnBwd* p, cudaStream_t st);
static bool g_force_simt_attn_bwd = false;

int xattn_bwd(const XattnBwd* p, cudaStream_t st) {
  TNT_CHECK_ARG(p && p->dout && p->dq && p->dk && p->dv, "xattn_bwd: null pointer");
  TNT_TRY(xattn_check(&p->f));
  TNT_CHECK_ARG(p->f.n_extra == 0, "xattn_bwd: materialise the bias_k / zero rows in k, v (n_extra must be 0)");
  TNT_CHECK_ARG(p->scratch != nullptr, "xattn_bwd: fp32 scratch [B,H,T,D] missing");
  if (p->f.T == 0 || p->f.B == 0) return TNT_OK;
  if (!g_force_simt_attn_bwd && !p->dweights && xattn_bwd_tc_supported(p)) {
    variant_hit(V_XATTN_BWD_TC);
    return xattn_bwd_tc(p, st);
  }
  variant_hit(V_XATTN_BWD_SIMT);
  BwdArgs a;
  a.f = p->f;
  a.scratch = p->scratch;
  a.f.weights = nullptr;
  a.drop = make_drop(p->f.drop);
  a.dweights = p->dweights;
  a.dout = p->dout; a.dq = p->dq; a.dk = p->dk; a.dv = p->dv;
  const int D = p->f.D;
  const size_t smem = (4 * static_cast<size_t>(kT) * (D + 1) + 2 * kT * 33 + 2 * static_cast<size_t>(p->f.T)) * 4;
  if (p->f.dtype == TNT_BF16) return D <= 64 ? launch<bf16, 8>(a, smem, st) : launch<bf16, 16>(a, smem, st);
  return D <= 64 ? launch<float, 8>(a, smem, st) : launch<float, 16>(a, smem, st);
}

}  // namespace tnt

extern "C" void tnt_debug_force_simt_attention_bwd(int on) { tnt::g_force_simt_attn_bwd = on != 0; }

extern "C" size_t tnt_xattn_bwd_scratch_bytes(int B, int T, int H, int D) {
  return sizeof(float) * static_cast<size_t>(B) * T * H * D;
}

extern "C" int tnt_xattn_bwd(const tnt_xattn_bwd_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p, "xattn_bwd: null params");
  TNT_CHECK_ARG(p->scratch_bytes >= tnt_xattn_bwd_scratch_bytes(p->fwd.B, p->fwd.T, p->fwd.H, p->fwd.D),
                "xattn_bwd: scratch too small");
  XattnBwd b;
  b.f = p->fwd;
  b.dout = p->dout; b.dq = p->dq; b.dk = p->dk; b.dv = p->dv;
  b.scratch = static_cast<float*>(p->scratch);
  b.dweights = p->dweights;
  return xattn_bwd(&b, as_stream(stream));
}
