// Cross attention forward over SHORT memory sources (<= 128 keys: the image regions 49 + 2, faces 4 + 2, objects 64 + 2
// of ttl/tell/models/decoder_faces_objects.py:283-393; math of ttl/tell/modules/attention/multi_head.py:376-477).
// The tcgen05 kernel (attention_tc.cu) works on 128-key score tiles through a TMA / TMEM pipeline that is built for the
// article (514 keys, 5 tiles): for 6 .. 66 keys its M = 128 key rows are mostly padding and every work item still pays
// the whole pipeline's latency chain -- 24 us per launch at B = 128 whatever the source, three launches per layer.
// Here ONE launch takes all short sources of a layer, and a WARP owns a (source, batch row, head, 16 queries) item end
// to end with warp-level mma.sync.m16n8k16 (bf16 in, fp32 accumulate) and no block-level synchronisation:
//   scores   S[q, key] = Q[q, :] . K[key, :]      A = Q fragments, B = K through ldmatrix from the warp's shared-memory copy
//   soft-max in the accumulator fragments (quad shuffles), exact: all <= 128 scores of a row are in registers
//   output   O[q, :]   = P[q, :] . V              A = P re-packed from the score fragments, B = V through ldmatrix.trans
//                                                 from the warp's own swizzled shared-memory copy of V
// Same numerics contract as the tcgen05 kernel: probabilities rounded to bf16 for the PV product, normaliser and LSE in
// fp32 from the un-rounded, un-dropped probabilities; dropout index ((b*H + h)*T + t)*S + key.
#include <algorithm>

#include "common.cuh"

namespace tnt {
namespace {

constexpr int kWarps = 2;       // small CTAs: the shared-memory copies (2 x S_pad x 128 B per warp) bound the occupancy, 2-warp CTAs pack the SM better
constexpr int kMaxKeys = 128;
constexpr int kMaxN8 = kMaxKeys / 8;

struct ShortBatch {
  tnt_xattn_params p[TNT_MAX_SOURCES];
  DropSpec dr[TNT_MAX_SOURCES];
  int first[TNT_MAX_SOURCES + 1];      // warp items of problem i: [first[i], first[i + 1])
  int n;
  int s_pad_max;                       // shared-memory rows per warp
};

__device__ __forceinline__ void mma_bf16_16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack2_bf16(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t ldg32(const bf16* p) { return __ldg(reinterpret_cast<const unsigned int*>(p)); }

__device__ __forceinline__ void ldmatrix_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

template <bool DROP>
__global__ void __launch_bounds__(kWarps * 32) xattn_short_kernel(const __grid_constant__ ShortBatch batch) {
  extern __shared__ __align__(16) uint8_t smem_kv[];
  pdl_wait();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int item = blockIdx.x * kWarps + warp;
  if (item >= batch.first[batch.n]) return;          // (no block-level barrier below: a warp may leave alone)
  int src = 0;
  while (src + 1 < batch.n && item >= batch.first[src + 1]) ++src;
  const tnt_xattn_params& p = batch.p[src];
  const int n_qt = (p.T + 15) >> 4;                  // a warp item = 16 queries
  int r = item - batch.first[src];
  const int qt = r % n_qt;
  r /= n_qt;
  const int h = r % p.H, b = r / p.H;
  const int S = p.S, n16 = (S + 15) >> 4, n8 = n16 * 2;
  const int SM = p.mask_mode == TNT_MASK_KEY_PADDING ? (p.mask_S > 0 ? min(p.mask_S, S) : S) : 0;   // keys the mask covers
  const uint8_t* mrow = p.mask + static_cast<long>(b) * p.m_stride_b;
  const bf16* qp = reinterpret_cast<const bf16*>(p.q) + static_cast<long>(b) * p.q_stride_b + h * 64;
  const bf16* kp = reinterpret_cast<const bf16*>(p.k) + static_cast<long>(b) * p.k_stride_b + h * 64;
  const bf16* vp = reinterpret_cast<const bf16*>(p.v) + static_cast<long>(b) * p.v_stride_b + h * 64;
  const int q0 = qt * 16, tq0 = q0 + g, tq1 = q0 + g + 8;   // the two query rows of this lane's fragments

  // ---- every global load of the item is issued up front (one memory latency, not one per fragment):
  // K and V of this (b, h) -> the warp's shared-memory blocks [key][64 dims] bf16, the 16-byte chunk c of row `key` stored
  // at chunk c ^ (key & 7) (ldmatrix reads 8 rows at one chunk index: the XOR spreads them over all banks), rows >= S zero;
  // the key-padding bytes as ballot words; the Q fragments (4 k-steps of 16 dims, rows beyond T read as zeros)
  uint8_t* ks_ = smem_kv + static_cast<size_t>(warp) * 2 * batch.s_pad_max * 128;
  uint8_t* vs_ = ks_ + static_cast<size_t>(batch.s_pad_max) * 128;
  // (cp.async: no register staging, so all rows are in flight at once -- a load / store loop through registers waited one
  // memory latency per 32 chunks, 16 round trips for 64 keys: 29 us per launch instead of a few)
  const uint32_t ks_addr = static_cast<uint32_t>(__cvta_generic_to_shared(ks_));
  const uint32_t vs_addr = static_cast<uint32_t>(__cvta_generic_to_shared(vs_));
  for (int i = lane; i < n16 * 16 * 8; i += 32) {
    const int key = i >> 3, c = i & 7;
    const int kc = min(key, S - 1);                    // rows >= S: zero-filled (source size 0), the address stays valid
    const uint32_t nbytes = key < S ? 16u : 0u;
    const uint32_t off = key * 128 + ((c ^ (key & 7)) << 4);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(ks_addr + off),
                 "l"(kp + static_cast<long>(kc) * p.k_stride_s + c * 8), "r"(nbytes) : "memory");
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(vs_addr + off),
                 "l"(vp + static_cast<long>(kc) * p.v_stride_s + c * 8), "r"(nbytes) : "memory");
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  uint32_t masked[kMaxKeys / 32];                    // bit l of word w: key 32 w + l is padding (or beyond S)
#pragma unroll
  for (int w = 0; w < kMaxKeys / 32; ++w) {
    const int key = 32 * w + lane;
    const bool m = key >= S || (key < SM && mrow[key] != 0);
    masked[w] = __ballot_sync(0xffffffffu, m);
  }
  uint32_t qa[4][4];
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) {
    const bf16* q0p = qp + static_cast<long>(tq0) * p.q_stride_t + ks * 16 + 2 * t4;
    const bf16* q1p = qp + static_cast<long>(tq1) * p.q_stride_t + ks * 16 + 2 * t4;
    qa[ks][0] = tq0 < p.T ? ldg32(q0p) : 0u;
    qa[ks][1] = tq1 < p.T ? ldg32(q1p) : 0u;
    qa[ks][2] = tq0 < p.T ? ldg32(q0p + 8) : 0u;
    qa[ks][3] = tq1 < p.T ? ldg32(q1p + 8) : 0u;
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  __syncwarp();
  const float sl2 = p.scale * 1.4426950408889634f;

  // ---- scores: one 16 x 8 accumulator fragment per 8 keys.  B fragments (K[key = 8 j + g][16 ks + 2 t4 ..]) by plain
  // ldmatrix: matrix m of an x4 = keys 8 j .. 8 j + 7 at chunk 4 half + m (half 0: k-steps 0, 1; half 1: k-steps 2, 3)
  float sc[kMaxN8][4];
#pragma unroll
  for (int j = 0; j < kMaxN8; ++j) {
    sc[j][0] = sc[j][1] = sc[j][2] = sc[j][3] = 0.f;
    if (j < n8) {
      const int key = 8 * j + (lane & 7);
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int chunk = 4 * half + (lane >> 3);
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4(ks_addr + key * 128 + ((chunk ^ (key & 7)) << 4), b0, b1, b2, b3);
        mma_bf16_16816(sc[j], qa[2 * half], b0, b1);
        mma_bf16_16816(sc[j], qa[2 * half + 1], b2, b3);
      }
    }
  }
  // ---- exact soft-max over the row (log2 domain): max, exp2, sum; quad shuffles combine the 4 lanes of a row.
  // Score column (j, e) of this lane = key 8 j + 2 t4 + e
  float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
  for (int j = 0; j < kMaxN8; ++j) {
    if (j < n8) {
      const uint32_t mw = masked[j >> 2] >> ((j & 3) * 8 + 2 * t4);
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const bool ok = !((mw >> e) & 1u);
        sc[j][e] = ok ? sc[j][e] * sl2 : -INFINITY;
        sc[j][2 + e] = ok ? sc[j][2 + e] * sl2 : -INFINITY;
        mx0 = fmaxf(mx0, sc[j][e]);
        mx1 = fmaxf(mx1, sc[j][2 + e]);
      }
    }
  }
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
  mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
  mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
  const float sub0 = mx0 == -INFINITY ? 0.f : mx0, sub1 = mx1 == -INFINITY ? 0.f : mx1;   // fully masked row: all p = 0
  float l0 = 0.f, l1 = 0.f;
  uint32_t pa[kMaxN8][2];                            // probabilities as bf16 pairs: [j][0] row tq0, [j][1] row tq1
#pragma unroll
  for (int j = 0; j < kMaxN8; ++j) {
    pa[j][0] = pa[j][1] = 0u;
    if (j < n8) {
      float p00 = ex2_approx(sc[j][0] - sub0), p01 = ex2_approx(sc[j][1] - sub0);
      float p10 = ex2_approx(sc[j][2] - sub1), p11 = ex2_approx(sc[j][3] - sub1);
      l0 += p00 + p01;
      l1 += p10 + p11;
      if (DROP) {
        const DropSpec& dr = batch.dr[src];
        const int key = 8 * j + 2 * t4;
        const unsigned long long i0 = ((static_cast<unsigned long long>(b) * p.H + h) * p.T + tq0) * S + key;
        const unsigned long long i1 = ((static_cast<unsigned long long>(b) * p.H + h) * p.T + tq1) * S + key;
        p00 = drop_val(dr, i0, p00);
        p01 = drop_val(dr, i0 + 1, p01);
        p10 = drop_val(dr, i1, p10);
        p11 = drop_val(dr, i1 + 1, p11);
      }
      pa[j][0] = pack2_bf16(p00, p01);
      pa[j][1] = pack2_bf16(p10, p11);
    }
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1);
  l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  // ---- O = P . V: k-steps of 16 keys; per step the 8 d-tiles' B fragments come from four ldmatrix.x4.trans
  float o[8][4];
#pragma unroll
  for (int nd = 0; nd < 8; ++nd) o[nd][0] = o[nd][1] = o[nd][2] = o[nd][3] = 0.f;
  const int lrow = (lane & 7) + ((lane >> 3) & 1) * 8, lchunk = lane >> 4;   // this lane's row address: key offset, chunk offset
#pragma unroll
  for (int kk = 0; kk < kMaxN8 / 2; ++kk) {
    if (kk < n16) {
      const uint32_t a[4] = {pa[2 * kk][0], pa[2 * kk][1], pa[2 * kk + 1][0], pa[2 * kk + 1][1]};
      const int key = kk * 16 + lrow;
#pragma unroll
      for (int c2 = 0; c2 < 4; ++c2) {                // d-tiles 2 c2 and 2 c2 + 1
        const int chunk = 2 * c2 + lchunk;
        uint32_t b0, b1, b2, b3;
        ldmatrix_x4_trans(vs_addr + key * 128 + ((chunk ^ (key & 7)) << 4), b0, b1, b2, b3);
        mma_bf16_16816(o[2 * c2], a, b0, b1);
        mma_bf16_16816(o[2 * c2 + 1], a, b2, b3);
      }
    }
  }
  // ---- epilogue: O / l -> bf16; optional natural-log LSE of the scaled, masked scores
  const float inv0 = l0 > 0.f ? 1.f / l0 : 0.f, inv1 = l1 > 0.f ? 1.f / l1 : 0.f;
  bf16* op = reinterpret_cast<bf16*>(p.out) + static_cast<long>(b) * p.o_stride_b + h * 64 + 2 * t4;
#pragma unroll
  for (int nd = 0; nd < 8; ++nd) {
    if (tq0 < p.T)
      *reinterpret_cast<uint32_t*>(op + static_cast<long>(tq0) * p.o_stride_t + nd * 8) = pack2_bf16(o[nd][0] * inv0, o[nd][1] * inv0);
    if (tq1 < p.T)
      *reinterpret_cast<uint32_t*>(op + static_cast<long>(tq1) * p.o_stride_t + nd * 8) = pack2_bf16(o[nd][2] * inv1, o[nd][3] * inv1);
  }
  if (p.lse && t4 == 0) {
    float* lp = p.lse + (static_cast<long>(b) * p.H + h) * p.T;
    if (tq0 < p.T) lp[tq0] = l0 > 0.f ? (mx0 + log2f(l0)) * 0.6931471805599453f : INFINITY;
    if (tq1 < p.T) lp[tq1] = l1 > 0.f ? (mx1 + log2f(l1)) * 0.6931471805599453f : INFINITY;
  }
}

}  // namespace

bool xattn_short_supported(const tnt_xattn_params* p) {
  static const bool on = [] { const char* e = getenv("TNT_XATTN_SHORT"); return !(e && e[0] == '0'); }();
  if (!on || p->dtype != TNT_BF16 || p->D != 64 || p->n_extra != 0 || p->weights || p->T < 2 || p->kv_group > 1) return false;
  if (p->S < 1 || p->S > kMaxKeys) return false;
  if (p->mask_mode != TNT_MASK_NONE && p->mask_mode != TNT_MASK_KEY_PADDING) return false;
  const long strides = p->q_stride_t | p->q_stride_b | p->k_stride_s | p->k_stride_b | p->v_stride_s | p->v_stride_b | p->o_stride_t |
                       p->o_stride_b;
  return (strides % 8) == 0 && aligned16(p->q) && aligned16(p->k) && aligned16(p->v) && aligned16(p->out);
}

// n short problems (the short sources of one decoder layer), one launch; longest source first
int xattn_short_multi(const tnt_xattn_params* ps, int n, cudaStream_t st) {
  TNT_CHECK_ARG(n >= 1 && n <= TNT_MAX_SOURCES, "xattn_short: %d problems", n);
  ShortBatch batch = {};
  int order[TNT_MAX_SOURCES];
  for (int i = 0; i < n; ++i) order[i] = i;
  std::sort(order, order + n, [&](int a, int b) { return ps[a].S > ps[b].S; });
  bool drop = false;
  int items = 0, s_pad = 16;
  for (int i = 0; i < n; ++i) {
    const tnt_xattn_params& p = ps[order[i]];
    batch.p[i] = p;
    batch.dr[i] = make_drop(p.drop);
    drop = drop || batch.dr[i].thr != 0;
    batch.first[i] = items;
    items += p.B * p.H * ((p.T + 15) / 16);
    s_pad = std::max(s_pad, (p.S + 15) / 16 * 16);
  }
  for (int i = n; i <= TNT_MAX_SOURCES; ++i) batch.first[i] = items;
  batch.n = n;
  batch.s_pad_max = s_pad;
  if (items == 0) return TNT_OK;
  const int smem = kWarps * 2 * s_pad * 128;
  static PerDeviceOnce attr_once;
  if (attr_once.first()) {
    TNT_CUDA(cudaFuncSetAttribute(xattn_short_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWarps * 2 * kMaxKeys * 128));
    TNT_CUDA(cudaFuncSetAttribute(xattn_short_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, kWarps * 2 * kMaxKeys * 128));
  }
  const int grid = (items + kWarps - 1) / kWarps;
  if (drop) TNT_CUDA(launch_pdl(xattn_short_kernel<true>, dim3(grid), dim3(kWarps * 32), smem, st, 1, batch));
  else TNT_CUDA(launch_pdl(xattn_short_kernel<false>, dim3(grid), dim3(kWarps * 32), smem, st, 1, batch));
  TNT_LAUNCHED();
  variant_hit(V_XATTN_SHORT);
  return TNT_OK;
}

}  // namespace tnt
