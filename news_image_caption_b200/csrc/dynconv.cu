// Causal DynamicConv1d (kernel group 1), its backward (group 4) and the single-step ring-buffer
// form (group 5).  Reference: ttl/tell/modules/convolutions/dynamic.py:85-120,285-336,338-348.
//
// One CTA owns (batch row b, a group of heads, a tile of timesteps).  The conv-input window
// [t0-(K-1), t0+TT) of that channel group is staged ONCE in shared memory (each input row is used by
// K outputs), the K tap logits of every (t, head) are soft-maxed by a warp with shuffles into shared
// memory, and each thread then produces 8 consecutive channels per timestep with 128-bit HBM
// loads/stores.  HBM traffic = x once (+ a K-1 row halo per extra time tile) + logits once + y once.
#include "common.cuh"
#include "ptx.cuh"

namespace tnt {

namespace {
constexpr int kThreads = 256;
constexpr int kTT = 32;           // timesteps per CTA tile

// softmax over K logits of one (t, head) by one warp; result to smem taps[K]
template <typename T>
__device__ __forceinline__ void warp_softmax_taps(const T* lg, int K, float* out, int lane) {
  float m = -INFINITY;
  for (int j = lane; j < K; j += 32) m = fmaxf(m, to_f(lg[j]));
  m = warp_max(m);
  float s = 0.f;
  for (int j = lane; j < K; j += 32) {
    const float e = __expf(to_f(lg[j]) - m);
    out[j] = e;
    s += e;
  }
  s = warp_sum(s);
  const float inv = 1.f / s;
  for (int j = lane; j < K; j += 32) out[j] *= inv;
}

// Tap soft-max of a whole CTA tile in two passes with ONE global-memory latency: every thread copies raw logits
// (rows x hg*K values, contiguous per row) into taps[(r*HG + hl)*K + j], then one thread per (row, head) normalises
// its K values in place.  Rows at or beyond `T_end` are zero filled.  Caller syncs before and after phase 2.
template <typename T>
__device__ __forceinline__ void stage_tap_logits(const T* lg, long ldl, int B, int b, int t0, int T_end, int rows, int hg,
                                                 int HG, int K, int h0, float* taps) {
  const int per_row = hg * K;
  for (int i = threadIdx.x; i < rows * per_row; i += kThreads) {
    const int r = i / per_row, rem = i - r * per_row;
    taps[r * HG * K + rem] = (t0 + r < T_end) ? to_f(lg[(static_cast<long>(t0 + r) * B + b) * ldl + h0 * K + rem]) : 0.f;
  }
}
// Tap dropout (weight_dropout, dynamic.py:305; element index ((t*B + b)*H + h)*K + j): the forward stores the dropped
// taps (w / (1-p) or 0); the backward keeps w and marks dropped taps with the SIGN bit (w >= 0), because the soft-max
// backward needs every un-dropped w while the value path and d(taps) need the mask.
__device__ __forceinline__ void normalise_taps(int t0, int T_end, int rows, int hg, int HG, int K, float* taps,
                                               const DropSpec& dr = DropSpec(), int B = 0, int b = 0, int H = 0, int h0 = 0,
                                               bool sign_encode = false) {
  for (int i = threadIdx.x; i < rows * hg; i += kThreads) {
    const int r = i / hg, hl = i - r * hg;
    if (t0 + r >= T_end) continue;
    float* o = taps + (r * HG + hl) * K;
    float m = -INFINITY;
    for (int j = 0; j < K; ++j) m = fmaxf(m, o[j]);
    float s = 0.f;
    for (int j = 0; j < K; ++j) {
      const float e = __expf(o[j] - m);
      o[j] = e;
      s += e;
    }
    const float inv = 1.f / s;
    if (!dr.thr) {
      for (int j = 0; j < K; ++j) o[j] *= inv;
    } else {
      const unsigned long long i0 = ((static_cast<unsigned long long>(t0 + r) * B + b) * H + h0 + hl) * K;
      for (int j = 0; j < K; ++j) {
        const float w = o[j] * inv;
        const bool keep = drop_bits(dr.k0, dr.k1, i0 + j) >= dr.thr;
        o[j] = sign_encode ? (keep ? w : -w) : (keep ? w * dr.scale : 0.f);
      }
    }
  }
}
__device__ __forceinline__ float tap_kept(float enc, float scale) {      // sign-encoded tap -> dropped-out value
  return (__float_as_uint(enc) >> 31) ? 0.f : enc * scale;
}

// ------------------------------------------------------------------------------ forward
// smem: taps [kTT][HG][K] fp32, then xs [(kTT+K-1)][CC] in the INPUT dtype (bf16 halves the shared-memory traffic).
// Each thread produces 8 channels x kNT consecutive timesteps: an input row is read from shared memory once per
// kNT outputs instead of once per output (the kernel was shared-memory-bandwidth bound at K = 31).
constexpr int kNT = 4;
template <typename T>
__device__ __forceinline__ void copy8(const T* src, T* dst) {            // 8 elements, 16-byte aligned (bf16) / 32 (fp32)
  if (sizeof(T) == 2) {
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
  } else {
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(src);
    *(reinterpret_cast<uint4*>(dst) + 1) = *(reinterpret_cast<const uint4*>(src) + 1);
  }
}
template <typename T>
__device__ __forceinline__ void zero8(T* dst) {
  *reinterpret_cast<uint4*>(dst) = make_uint4(0, 0, 0, 0);
  if (sizeof(T) == 4) *(reinterpret_cast<uint4*>(dst) + 1) = make_uint4(0, 0, 0, 0);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) dynconv_fwd_kernel(tnt_dynconv_params p, int CC, int HG, DropSpec dr) {
  pdl_wait();
  extern __shared__ float sm[];
  const int K = p.K, R = p.C / p.H, B = p.B;
  const int b = blockIdx.x, c0 = blockIdx.y * CC, t0 = blockIdx.z * kTT;
  const int TT = min(kTT, p.T - t0);
  const int rows = TT + K - 1;
  float* taps = sm;
  T* xs = reinterpret_cast<T*>(sm + static_cast<size_t>(kTT) * HG * K);
  const T* x = reinterpret_cast<const T*>(p.x);
  const T* hist = reinterpret_cast<const T*>(p.hist);
  const T* lg = reinterpret_cast<const T*>(p.logits);
  T* y = reinterpret_cast<T*>(p.y);
  const int cc = min(CC, p.C - c0);
  const int nv = cc >> 3;

  // stage the input window (zeros / history before t = 0), raw
  for (int i = threadIdx.x; i < rows * nv; i += kThreads) {
    const int r = i / nv, vc = i - r * nv;
    const int ts = t0 - (K - 1) + r;
    T* d = xs + r * CC + vc * 8;
    if (ts >= 0) copy8(x + (static_cast<long>(ts) * B + b) * p.ldx + c0 + vc * 8, d);
    else if (hist && p.hist_len + ts >= 0) copy8(hist + (static_cast<long>(p.hist_len + ts) * B + b) * p.C + c0 + vc * 8, d);
    else zero8(d);
  }
  // taps: raw logits -> smem together with the input window (one memory latency), then normalised in place
  const int hg = cc / R, h0 = c0 / R;
  stage_tap_logits(lg, p.ldl, B, b, t0, p.T, TT, hg, HG, K, h0, taps);
  __syncthreads();
  normalise_taps(t0, p.T, TT, hg, HG, K, taps, dr, B, b, p.H, h0, false);
  __syncthreads();
  const int ngrp = (TT + kNT - 1) / kNT;
  for (int i = threadIdx.x; i < ngrp * nv; i += kThreads) {
    const int tg = i / nv, vc = i - tg * nv;
    const int tb = tg * kNT;                               // outputs tb .. tb+kNT-1 (local timesteps)
    const float* w0 = taps + (tb * HG + (vc * 8) / R) * K;
    const T* xr = xs + tb * CC + vc * 8;                   // output t, tap j reads local row t + j
    float acc[kNT][8];
#pragma unroll
    for (int tt = 0; tt < kNT; ++tt)
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[tt][e] = 0.f;
    const int nrr = min(K + kNT - 1, rows - tb);
    for (int rr = 0; rr < nrr; ++rr) {
      float xv[8];
      load8(xr + rr * CC, xv);
#pragma unroll
      for (int tt = 0; tt < kNT; ++tt) {
        const int j = rr - tt;
        if (j >= 0 && j < K && tb + tt < TT) {
          const float wj = w0[tt * HG * K + j];
#pragma unroll
          for (int e = 0; e < 8; ++e) acc[tt][e] = fmaf(wj, xv[e], acc[tt][e]);
        }
      }
    }
#pragma unroll
    for (int tt = 0; tt < kNT; ++tt)
      if (tb + tt < TT) store8(y + (static_cast<long>(t0 + tb + tt) * B + b) * p.ldy + c0 + vc * 8, acc[tt]);
  }
}

// ------------------------------------------------------------------------------ forward on tensor cores (bf16)
// The SIMT forward is bound by its FMA / shared-memory instruction stream (K = 31: 31 FMAs per output, 15 % of the HBM
// roofline), not by HBM.  This form does what the reference's own fast path does (dynamic.py:285-336, "band matrices"):
// per (b, head) Y[32 x 64] = W_band[32 x KW] . X_window[KW x 64] -- on warp-level tensor-core MMAs (m16n8k16, bf16 in,
// fp32 accumulate): the tile is far too small for a tcgen05 (128-row) instruction, and the kernel's job is to get out
// of the way of the HBM stream.  CTA = (b, 4 heads = 256 channels, 32 timesteps), 8 warps, warp = (head, 32 channels).
//   smem: window rows [KW][256] bf16 (row stride padded by 16 B against ldmatrix bank conflicts), taps fp32 (soft-max,
//   dropout), band matrices [4][32][KW] bf16.  Zero k-tiles of the band are skipped (3 of 4 remain at K = 31).
namespace tcv {
constexpr int TT = 32, HG = 4, R = 64, CC = HG * R;
constexpr int XS_STRIDE = CC * 2 + 16;                    // bytes per window row

__device__ __forceinline__ uint32_t ptx_smem(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }
__device__ __forceinline__ void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async4(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// NJ = taps per thread (two threads per (t, head) row); KT = the tap count when known at compile time (the reference
// layers' 3 / 7 / 15 / 31: every `j < K` test but the last folds away), 0 = p.K
template <int NJ, int KT>
__device__ __forceinline__ void band_softmax(const tnt_dynconv_params& p, const DropSpec& dr, const bf16* lraw, uint8_t* As,
                                             int AS, int LS, int TTl, int t0, int b, int h0) {
  const int K = KT ? KT : p.K, B = p.B;
  // a warp covers 16 consecutive timesteps of ONE head (head-major: the 4 heads' band rows sit 32 * AS bytes apart,
  // a multiple of 128, so a timestep-major warp stored 4-way bank-conflicted); the thread's <= 17 taps stay in
  // registers between the max, the sum and the store (the three-pass form re-read and re-exponentiated them: the
  // kernel was bound by instruction issue, 64 % active, with half of its instructions in this block)
  const int i = threadIdx.x >> 1, half = threadIdx.x & 1;          // TT * HG = 128 (t, head) rows, two threads each
  const int hl = i / TT, t_raw = i - hl * TT;
  const bool on = t_raw < TTl;                                       // rows beyond the tile: compute on row 0, write nothing
  const int t = on ? t_raw : 0;
  const bf16* l = lraw + t * LS + hl * K + half;
  float e[NJ];
  float m = -INFINITY;
#pragma unroll
  for (int jj = 0; jj < NJ; ++jj) {
    e[jj] = half + 2 * jj < K ? __bfloat162float(l[2 * jj]) : -INFINITY;
    m = fmaxf(m, e[jj]);
  }
  m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
  float sum = 0.f;
#pragma unroll
  for (int jj = 0; jj < NJ; ++jj) {
    e[jj] = half + 2 * jj < K ? __expf(e[jj] - m) : 0.f;
    sum += e[jj];
  }
  sum += __shfl_xor_sync(0xffffffffu, sum, 1);
  const float inv = 1.f / sum;
  bf16* arow = reinterpret_cast<bf16*>(As + (hl * TT + t) * AS) + t + half;
  if (!on) return;
  if (dr.thr) {
    const unsigned long long i0 = ((static_cast<unsigned long long>(t0 + t) * B + b) * p.H + h0 + hl) * K + half;
#pragma unroll
    for (int jj = 0; jj < NJ; ++jj)
      if (half + 2 * jj < K) arow[2 * jj] = __float2bfloat16_rn(drop_val(dr, i0 + 2 * jj, e[jj] * inv));
  } else {
#pragma unroll
    for (int jj = 0; jj < NJ; ++jj)
      if (half + 2 * jj < K) arow[2 * jj] = __float2bfloat16_rn(e[jj] * inv);
  }
}

// soft-max backward of one (t, head) row in place: l[j] <- w_j (g_j - sum_i w_i g_i), w from the stored (max, 1/sum)
template <int NJ>
__device__ __forceinline__ void tap_softmax_bwd(bf16* l, const float* g, float2 st, int K) {
  float w[NJ];
  float dot = 0.f;
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    if (j < K) {
      w[j] = __expf(__bfloat162float(l[j]) - st.x) * st.y;
      dot = fmaf(w[j], g[j], dot);
    }
  }
#pragma unroll
  for (int j = 0; j < NJ; ++j)
    if (j < K) l[j] = __float2bfloat16_rn(w[j] * (g[j] - dot));
}

// One (gradient row r, head): soft-max statistics (max, 1/sum) of its K taps and its column of the dx band matrix
// (tap j -> band row s = r - (K-1) + j when 0 <= s < 32); the taps are exponentiated once and kept in registers.
template <int NJ>
__device__ __forceinline__ float2 band_column(const bf16* l, int K, int KM, int r, unsigned long long i0, const DropSpec& dr,
                                              uint8_t* acol, int AS) {
  float e[NJ];
  float m = -INFINITY;
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    e[j] = j < K ? __bfloat162float(l[j]) : -INFINITY;
    m = fmaxf(m, e[j]);
  }
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    e[j] = j < K ? __expf(e[j] - m) : 0.f;
    sum += e[j];
  }
  const float inv = 1.f / sum;
  const int j_lo = max(0, KM - r), j_hi = min(K - 1, TT - 1 + KM - r);
#pragma unroll
  for (int j = 0; j < NJ; ++j) {
    if (j >= j_lo && j <= j_hi) {
      float w = e[j] * inv;
      if (dr.thr) w = drop_val(dr, i0 + j, w);
      *reinterpret_cast<bf16*>(acol + j * AS) = __float2bfloat16_rn(w);
    }
  }
  return make_float2(m, inv);
}

// All global reads of the CTA (window rows, raw tap logits) are issued as cp.async up front: the earlier version's
// staging loops (load -> store per iteration) serialised one memory latency per iteration and per thread, which --
// not the arithmetic -- was what kept the kernel at 15 % of the HBM roofline.
__global__ void __launch_bounds__(kThreads) dynconv_fwd_tc_kernel(tnt_dynconv_params p, int KW, DropSpec dr) {
  pdl_wait();
  extern __shared__ __align__(16) uint8_t smb[];
  const int K = p.K, B = p.B;
  const int b = blockIdx.x, c0 = blockIdx.y * CC, t0 = blockIdx.z * TT;
  const int TTl = min(TT, p.T - t0);
  const int rows = TTl + K - 1;
  const int AS = KW * 2 + 16;                            // bytes per band-matrix row
  const int HK = HG * K;                                 // tap logits of the CTA's heads per timestep (even)
  uint8_t* xs = smb;                                     // [KW][XS_STRIDE]
  uint8_t* As = xs + KW * XS_STRIDE;                     // [HG][TT][AS]
  bf16* lraw = reinterpret_cast<bf16*>(As + HG * TT * AS);     // [TT][LS] raw tap logits
  const int LS = (((HK * 2 + 15) & ~15) + 16) >> 1;     // row stride (elements): the group's taps + 16-byte alignment slack
  const bf16* x = reinterpret_cast<const bf16*>(p.x);
  const bf16* hist = reinterpret_cast<const bf16*>(p.hist);
  const bf16* lg = reinterpret_cast<const bf16*>(p.logits);
  bf16* y = reinterpret_cast<bf16*>(p.y);
  constexpr int nv = CC / 8;
  const int h0 = c0 / R;

  const uint32_t xs_s = ptx_smem(xs), lraw_s = ptx_smem(lraw);
  // window rows: ONE bulk copy (512 B, cp.async.bulk -> mbarrier) per row, issued by thread r -- the 16-byte cp.async
  // form spent 8 loop iterations of 64-bit address arithmetic per thread (a fifth of the kernel's instructions).
  // Rows [r_h, rows) exist (history before t = 0, then x); the others are zero.
  uint64_t* bar = reinterpret_cast<uint64_t*>(lraw + TT * LS);
  // tap logits: the group's H*K-slice of a timestep row starts 8-byte aligned at best, so the bulk copy takes the
  // enclosing 16-byte-aligned span (it stays inside the row when rows are 16-byte multiples) and the soft-max reads at
  // the offset `loff`; otherwise 4-byte cp.async
  const int g0 = h0 * K * 2, a0 = g0 & ~15, lbytes = ((g0 + HK * 2 + 15) & ~15) - a0;
  const bool lbulk = p.ldl % 8 == 0 && (reinterpret_cast<uintptr_t>(lg) & 15) == 0 && (p.H * K) % 8 == 0;
  const int loff = lbulk ? (g0 - a0) >> 1 : 0;
  const int r_x = max(0, K - 1 - t0);                                    // first row that comes from x
  const int r_h = hist ? max(0, K - 1 - t0 - p.hist_len) : r_x;          // first row that exists at all
  if (threadIdx.x == 0) {
    ptx::mbar_init(bar, 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) ptx::mbar_arrive_expect_tx(bar, static_cast<uint32_t>(rows - r_h) * CC * 2 + (lbulk ? TTl * lbytes : 0));
  if (lbulk && threadIdx.x >= 64 && threadIdx.x < 64 + TTl) {
    const int r = threadIdx.x - 64;
    const uint8_t* src = reinterpret_cast<const uint8_t*>(lg + (static_cast<long>(t0 + r) * B + b) * p.ldl) + a0;
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(lraw_s + r * LS * 2), "l"(src), "r"(lbytes), "r"(ptx_smem(bar)) : "memory");
  }
  if (threadIdx.x >= r_h && threadIdx.x < rows) {
    const int r = threadIdx.x, ts = t0 - (K - 1) + r;
    const bf16* src = r >= r_x ? x + (static_cast<long>(ts) * B + b) * p.ldx + c0
                               : hist + (static_cast<long>(p.hist_len + ts) * B + b) * p.C + c0;
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(xs_s + r * XS_STRIDE), "l"(src), "r"(CC * 2), "r"(ptx_smem(bar)) : "memory");
  }
  for (int i = threadIdx.x; i < (r_h + KW - rows) * nv; i += kThreads) {     // the rows that do not exist
    const int zr = i / nv, vc = i - zr * nv;
    const int r = zr < r_h ? zr : rows + (zr - r_h);
    *reinterpret_cast<uint4*>(xs + r * XS_STRIDE + vc * 16) = make_uint4(0, 0, 0, 0);
  }
  if (!lbulk) {
    const int wpr = HK >> 1;                             // 4-byte words per timestep
    const int w0 = threadIdx.x >> 5, l0 = threadIdx.x & 31;          // warp = timestep row, lane = word: no division by K
    if (wpr > 16) {
      for (int r = w0; r < TTl; r += kThreads / 32) {
        const bf16* src = lg + (static_cast<long>(t0 + r) * B + b) * p.ldl + h0 * K;
        for (int wd = l0; wd < wpr; wd += 32) cp_async4(lraw_s + (r * LS + 2 * wd) * 2, src + 2 * wd);
      }
    } else {                                             // few taps: a row would leave most of a warp idle
      for (int i = threadIdx.x; i < TTl * wpr; i += kThreads) {
        const int r = i / wpr, wd = i - r * wpr;
        cp_async4(lraw_s + (r * LS + 2 * wd) * 2, lg + (static_cast<long>(t0 + r) * B + b) * p.ldl + h0 * K + 2 * wd);
      }
    }
  }
  for (int i = threadIdx.x; i < (HG * TT * AS) / 16; i += kThreads) reinterpret_cast<uint4*>(As)[i] = make_uint4(0, 0, 0, 0);
  cp_async_wait_all();
  ptx::mbar_wait(bar, 0);                                // window rows and tap logits have landed (async proxy -> visible)
  __syncthreads();
  // soft-max of the K taps of every (t, head) into its row of the band matrix: A[hl][t][t + j] = w_j (window row c holds
  // time t0 - (K-1) + c, so output t reads rows t .. t+K-1), unrolled for the layer's tap count
  switch (K) {
    case 3: band_softmax<2, 3>(p, dr, lraw + loff, As, AS, LS, TTl, t0, b, h0); break;
    case 7: band_softmax<4, 7>(p, dr, lraw + loff, As, AS, LS, TTl, t0, b, h0); break;
    case 15: band_softmax<8, 15>(p, dr, lraw + loff, As, AS, LS, TTl, t0, b, h0); break;
    case 31: band_softmax<16, 31>(p, dr, lraw + loff, As, AS, LS, TTl, t0, b, h0); break;
    default:
      if (K <= 8) band_softmax<4, 0>(p, dr, lraw + loff, As, AS, LS, TTl, t0, b, h0);
      else if (K <= 16) band_softmax<8, 0>(p, dr, lraw + loff, As, AS, LS, TTl, t0, b, h0);
      else band_softmax<17, 0>(p, dr, lraw + loff, As, AS, LS, TTl, t0, b, h0);
  }
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hl = warp >> 1, chb = (warp & 1) * 32;       // channels [hl*64 + chb, +32) of the CTA's 256
  float acc[2][4][4];
#pragma unroll
  for (int mi = 0; mi < 2; ++mi)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[mi][nt][e] = 0.f;
  const uint32_t a_base = ptx_smem(As + hl * TT * AS) + ((lane & 7) + 8 * ((lane >> 3) & 1)) * AS + 16 * (lane >> 4);
  const uint32_t b_base = ptx_smem(xs) + ((lane & 7) + 8 * ((lane >> 3) & 1)) * XS_STRIDE + (hl * R + chb + 8 * (lane >> 4)) * 2;
  const int nkt = KW >> 4;
#pragma unroll
  for (int mi = 0; mi < 2; ++mi) {
    const int kt_hi = min(nkt, ((16 * mi + 15 + K - 1) >> 4) + 1);
    for (int kt = mi; kt < kt_hi; ++kt) {
      uint32_t a[4];
      ldsm_x4(a, a_base + mi * 16 * AS + kt * 32);
#pragma unroll
      for (int jp = 0; jp < 2; ++jp) {
        uint32_t bq[4];
        ldsm_x4_trans(bq, b_base + kt * 16 * XS_STRIDE + jp * 32);
        mma_bf16(acc[mi][2 * jp], a, bq[0], bq[1]);
        mma_bf16(acc[mi][2 * jp + 1], a, bq[2], bq[3]);
      }
    }
  }
  __syncthreads();                                       // every warp is done with the window: reuse it for the output tile
#pragma unroll
  for (int mi = 0; mi < 2; ++mi)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int col = hl * R + chb + nt * 8 + 2 * (lane & 3);
      const int r0 = mi * 16 + (lane >> 2);
      *reinterpret_cast<uint32_t*>(xs + r0 * XS_STRIDE + col * 2) = pack_bf16x2(acc[mi][nt][0], acc[mi][nt][1]);
      *reinterpret_cast<uint32_t*>(xs + (r0 + 8) * XS_STRIDE + col * 2) = pack_bf16x2(acc[mi][nt][2], acc[mi][nt][3]);
    }
  __syncthreads();
  for (int i = threadIdx.x; i < TTl * nv; i += kThreads) {
    const int t = i / nv, vc = i - t * nv;
    *reinterpret_cast<uint4*>(y + (static_cast<long>(t0 + t) * B + b) * p.ldy + c0 + vc * 8) =
        *reinterpret_cast<const uint4*>(xs + t * XS_STRIDE + vc * 16);
  }
}

// ------------------------------------------------------------------------------ backward on tensor cores (bf16)
// Same tiling as the forward.  Windows: XW rows = times t0-(K-1) .. t0+31 (conv inputs), DW rows = times t0 .. t0+31+(K-1)
// (output gradients; the tile's inputs feed K-1 later outputs), raw tap logits of the DW rows.
//   dx [32 x 64]  = A_dx [32 x KW] . DW [KW x 64],  A_dx[s][r] = w_dropped[r][j = s + (K-1) - r]        (value path)
//   P  [32 x KW]  = dY [32 x 64] . XW^T,             dw[t][j] = P[t][t + j] (x dropout mask)              (tap path)
//   dlogits[t][j] = w_j (dw_j - sum_i w_i dw_i)      soft-max backward with the un-dropped w (recomputed from the raw
//                                                     logits and the per-row (max, 1/sum) kept from the first pass)
__global__ void __launch_bounds__(kThreads) dynconv_bwd_tc_kernel(tnt_dynconv_bwd_params p, int KW, DropSpec dr) {
  extern __shared__ __align__(16) uint8_t smb[];
  const int K = p.K, B = p.B, KM = K - 1;
  const int b = blockIdx.x, c0 = blockIdx.y * CC, t0 = blockIdx.z * TT;
  const int TTl = min(TT, p.T - t0);
  const int rows_x = TTl + KM;                           // XW rows that hold data
  const int rows_y = min(TT + KM, p.T - t0);             // DW / logits rows that exist
  const int AS = KW * 2 + 16;
  const int HK = HG * K;
  uint8_t* xw = smb;                                     // [KW][XS_STRIDE]
  uint8_t* dw = xw + KW * XS_STRIDE;                     // [KW][XS_STRIDE]
  uint8_t* As = dw + KW * XS_STRIDE;                     // [HG][TT][AS]  band matrix, later dws fp32 [TT][HG][K]
  bf16* lraw = reinterpret_cast<bf16*>(As + HG * TT * AS);     // [KW][HK] raw logits, later dlogits of rows < 32
  float2* stat = reinterpret_cast<float2*>(reinterpret_cast<uint8_t*>(lraw) + ((KW * HK * 2 + 15) & ~15));   // [KW][HG] (max, 1/sum)
  const bf16* x = reinterpret_cast<const bf16*>(p.x);
  const bf16* dy = reinterpret_cast<const bf16*>(p.dy);
  const bf16* lg = reinterpret_cast<const bf16*>(p.logits);
  constexpr int nv = CC / 8;
  const int h0 = c0 / R;

  const uint32_t xw_s = ptx_smem(xw), dw_s = ptx_smem(dw), lraw_s = ptx_smem(lraw);
  // both windows: one 512-byte bulk copy per existing row (threads 0..KW-1: XW rows, threads 128..128+KW-1: DW rows),
  // completion counted on one mbarrier; rows that do not exist are zero-filled (see the forward kernel)
  uint64_t* bar = reinterpret_cast<uint64_t*>(stat + KW * HG);
  const int r_x = max(0, KM - t0);                                       // first XW row that exists
  if (threadIdx.x == 0) {
    ptx::mbar_init(bar, 1);
    ptx::fence_barrier_init();
  }
  __syncthreads();
  if (threadIdx.x == 0) ptx::mbar_arrive_expect_tx(bar, static_cast<uint32_t>(rows_x - r_x + rows_y) * CC * 2);
  if (threadIdx.x < 128) {
    const int r = threadIdx.x;
    if (r >= r_x && r < rows_x)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(xw_s + r * XS_STRIDE), "l"(x + (static_cast<long>(t0 - KM + r) * B + b) * p.ldx + c0), "r"(CC * 2),
                     "r"(ptx_smem(bar)) : "memory");
  } else {
    const int r = threadIdx.x - 128;
    if (r < rows_y)
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dw_s + r * XS_STRIDE), "l"(dy + (static_cast<long>(t0 + r) * B + b) * p.lddy + c0), "r"(CC * 2),
                     "r"(ptx_smem(bar)) : "memory");
  }
  for (int i = threadIdx.x; i < (r_x + KW - rows_x) * nv; i += kThreads) {
    const int zr = i / nv, vc = i - zr * nv;
    const int r = zr < r_x ? zr : rows_x + (zr - r_x);
    *reinterpret_cast<uint4*>(xw + r * XS_STRIDE + vc * 16) = make_uint4(0, 0, 0, 0);
  }
  for (int i = threadIdx.x; i < (KW - rows_y) * nv; i += kThreads) {
    const int zr = i / nv, vc = i - zr * nv;
    *reinterpret_cast<uint4*>(dw + (rows_y + zr) * XS_STRIDE + vc * 16) = make_uint4(0, 0, 0, 0);
  }
  {
    const int wpr = HK >> 1;
    for (int i = threadIdx.x; i < rows_y * wpr; i += kThreads) {
      const int r = i / wpr, wd = i - r * wpr;
      cp_async4(lraw_s + (r * HK + 2 * wd) * 2, lg + (static_cast<long>(t0 + r) * B + b) * p.ldl + h0 * K + 2 * wd);
    }
  }
  for (int i = threadIdx.x; i < (HG * TT * AS) / 16; i += kThreads) reinterpret_cast<uint4*>(As)[i] = make_uint4(0, 0, 0, 0);
  cp_async_wait_all();
  __syncthreads();
  // pass 1: soft-max statistics of every (DW row, head) and its column of the dx band matrix
  for (int i = threadIdx.x; i < rows_y * HG; i += kThreads) {
    const int r = i / HG, hl = i - r * HG;
    const bf16* l = lraw + r * HK + hl * K;
    const unsigned long long i0 = ((static_cast<unsigned long long>(t0 + r) * B + b) * p.H + h0 + hl) * K;
    uint8_t* acol = As + (hl * TT + r - KM) * AS + r * 2;              // tap j lands in band row s = r - (K-1) + j, column r
    float2 st;                                                           // taps stay in registers between the passes
    if (K <= 4) st = band_column<4>(l, K, KM, r, i0, dr, acol, AS);
    else if (K <= 8) st = band_column<8>(l, K, KM, r, i0, dr, acol, AS);
    else if (K <= 16) st = band_column<16>(l, K, KM, r, i0, dr, acol, AS);
    else st = band_column<33>(l, K, KM, r, i0, dr, acol, AS);
    stat[r * HG + hl] = st;
  }
  ptx::mbar_wait(bar, 0);                                // both windows have landed
  __syncthreads();

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int hl = warp >> 1, half = warp & 1, chb = half * 32;
  const int lrow = (lane & 7) + 8 * ((lane >> 3) & 1);
  // ---- dx = A_dx . DW (the forward's MMA structure with the gradient window)
  float accx[2][4][4];
#pragma unroll
  for (int mi = 0; mi < 2; ++mi)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) accx[mi][nt][e] = 0.f;
  {
    const uint32_t a_base = ptx_smem(As + hl * TT * AS) + lrow * AS + 16 * (lane >> 4);
    const uint32_t b_base = dw_s + lrow * XS_STRIDE + (hl * R + chb + 8 * (lane >> 4)) * 2;
    const int nkt = KW >> 4;
#pragma unroll
    for (int mi = 0; mi < 2; ++mi) {
      const int kt_hi = min(nkt, ((16 * mi + 15 + KM) >> 4) + 1);
      for (int kt = mi; kt < kt_hi; ++kt) {
        uint32_t a[4];
        ldsm_x4(a, a_base + mi * 16 * AS + kt * 32);
#pragma unroll
        for (int jp = 0; jp < 2; ++jp) {
          uint32_t bq[4];
          ldsm_x4_trans(bq, b_base + kt * 16 * XS_STRIDE + jp * 32);
          mma_bf16(accx[mi][2 * jp], a, bq[0], bq[1]);
          mma_bf16(accx[mi][2 * jp + 1], a, bq[2], bq[3]);
        }
      }
    }
  }
  // ---- P = dY . XW^T over the head's 64 channels; this warp's half of the KW columns
  const int NT = KW >> 3, nt0 = half * (NT >> 1), ntn = NT >> 1;     // n-tiles [nt0, nt0 + ntn): 4 (KW = 64) or 3 (KW = 48)
  float accp[2][4][4];
#pragma unroll
  for (int mi = 0; mi < 2; ++mi)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) accp[mi][nt][e] = 0.f;
  {
    const uint32_t a_base = dw_s + lrow * XS_STRIDE + (hl * R + 8 * (lane >> 4)) * 2;
    const uint32_t b_base = xw_s + ((lane & 7) + 8 * (lane >> 4)) * XS_STRIDE + (hl * R + 8 * ((lane >> 3) & 1)) * 2;
#pragma unroll
    for (int kt = 0; kt < R / 16; ++kt) {
      uint32_t a0[4], a1[4];
      ldsm_x4(a0, a_base + kt * 32);
      ldsm_x4(a1, a_base + 16 * XS_STRIDE + kt * 32);
#pragma unroll
      for (int jp = 0; jp < 2; ++jp) {
        if (2 * jp < ntn) {
          uint32_t bq[4];
          ldsm_x4(bq, b_base + (nt0 + 2 * jp) * 8 * XS_STRIDE + kt * 32);
          mma_bf16(accp[0][2 * jp], a0, bq[0], bq[1]);
          mma_bf16(accp[1][2 * jp], a1, bq[0], bq[1]);
          if (2 * jp + 1 < ntn) {
            mma_bf16(accp[0][2 * jp + 1], a0, bq[2], bq[3]);
            mma_bf16(accp[1][2 * jp + 1], a1, bq[2], bq[3]);
          }
        }
      }
    }
  }
  __syncthreads();                                       // band matrix and windows are no longer read
  // dw[t][hl][j] = P[t][t + j] * dropout multiplier -> fp32 in the band matrix's place
  float* dws = reinterpret_cast<float*>(As);
#pragma unroll
  for (int mi = 0; mi < 2; ++mi)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int row = mi * 16 + (lane >> 2) + 8 * (e >> 1);
        const int col = (nt0 + nt) * 8 + 2 * (lane & 3) + (e & 1);
        const int j = col - row;
        if (nt < ntn && row < TTl && j >= 0 && j < K) {
          float v = accp[mi][nt][e];
          if (dr.thr) v = drop_val(dr, ((static_cast<unsigned long long>(t0 + row) * B + b) * p.H + h0 + hl) * K + j, v);
          dws[(row * HG + hl) * K + j] = v;
        }
      }
  // dx tile -> smem (the x window's place) for coalesced stores
#pragma unroll
  for (int mi = 0; mi < 2; ++mi)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt) {
      const int col = hl * R + chb + nt * 8 + 2 * (lane & 3);
      const int r0 = mi * 16 + (lane >> 2);
      *reinterpret_cast<uint32_t*>(xw + r0 * XS_STRIDE + col * 2) = pack_bf16x2(accx[mi][nt][0], accx[mi][nt][1]);
      *reinterpret_cast<uint32_t*>(xw + (r0 + 8) * XS_STRIDE + col * 2) = pack_bf16x2(accx[mi][nt][2], accx[mi][nt][3]);
    }
  __syncthreads();
  // soft-max backward of the tile's own rows; dlogits overwrite the raw logits in place
  for (int i = threadIdx.x; i < TTl * HG; i += kThreads) {
    const int t = i / HG, hq = i - t * HG;
    bf16* l = lraw + t * HK + hq * K;
    const float2 st = stat[t * HG + hq];
    const float* g = dws + (t * HG + hq) * K;
    if (K <= 4) tap_softmax_bwd<4>(l, g, st, K);          // taps exponentiated once, kept in registers
    else if (K <= 8) tap_softmax_bwd<8>(l, g, st, K);
    else if (K <= 16) tap_softmax_bwd<16>(l, g, st, K);
    else tap_softmax_bwd<33>(l, g, st, K);
  }
  bf16* dxo = reinterpret_cast<bf16*>(p.dx);
  for (int i = threadIdx.x; i < TTl * nv; i += kThreads) {
    const int t = i / nv, vc = i - t * nv;
    *reinterpret_cast<uint4*>(dxo + (static_cast<long>(t0 + t) * B + b) * p.lddx + c0 + vc * 8) =
        *reinterpret_cast<const uint4*>(xw + t * XS_STRIDE + vc * 16);
  }
  __syncthreads();
  {
    bf16* dlo = reinterpret_cast<bf16*>(p.dlogits);
    const int wpr = HK >> 1;
    for (int i = threadIdx.x; i < TTl * wpr; i += kThreads) {
      const int t = i / wpr, wd = i - t * wpr;
      *reinterpret_cast<uint32_t*>(dlo + (static_cast<long>(t0 + t) * B + b) * p.lddl + h0 * K + 2 * wd) =
          *reinterpret_cast<const uint32_t*>(lraw + t * HK + 2 * wd);
    }
  }
}
}  // namespace tcv

// ------------------------------------------------------------------------------ single step
// ring [K-1][B][C]; slot of step s = s mod (K-1).  One thread = 8 channels of one batch row.
template <typename T>
__global__ void __launch_bounds__(kThreads) dynconv_step_kernel(tnt_dynconv_step_params p) {
  pdl_wait();
  extern __shared__ float sm[];               // taps [H][K] of this batch row
  const int K = p.K, R = p.C / p.H, b = blockIdx.x;
  const int pos = p.pos ? *p.pos : 0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const T* lg = reinterpret_cast<const T*>(p.logits) + static_cast<long>(b) * p.ldl;
  for (int h = warp; h < p.H; h += kThreads / 32) warp_softmax_taps(lg + h * K, K, sm + h * K, lane);
  __syncthreads();
  const T* x = reinterpret_cast<const T*>(p.x) + static_cast<long>(b) * p.ldx;
  T* y = reinterpret_cast<T*>(p.y) + static_cast<long>(b) * p.ldy;
  T* ring = reinterpret_cast<T*>(p.ring);
  const int nv = p.C >> 3, KM = K - 1;
  for (int vc = threadIdx.x; vc < nv; vc += kThreads) {
    const float* w = sm + ((vc * 8) / R) * K;
    float xv[8], acc[8];
    load8(x + vc * 8, xv);
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = w[K - 1] * xv[i];
    // tap j < K-1 multiplies the input of step pos-(K-1)+j
    for (int j = 0; j < KM; ++j) {
      const int s = pos - KM + j;
      if (s < 0) continue;
      float hv[8];
      load8(ring + (static_cast<long>(s % KM) * p.B + b) * p.C + vc * 8, hv);
      const float wj = w[j];
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] = fmaf(wj, hv[i], acc[i]);
    }
    store8(y + vc * 8, acc);
    if (KM > 0) store8(ring + (static_cast<long>(pos % KM) * p.B + b) * p.C + vc * 8, xv);
  }
}

// bf16 decode step with every ring read in flight at once: thread = (8-channel vector, one of 4 tap groups); its <= 8
// ring rows are requested as raw 16-byte loads BEFORE the tap soft-max runs (the version above walks the K-1 rows in a
// dependent loop on half of its threads: 19 % of the HBM roofline).  K <= 33, C <= 2048.
constexpr int kStepGroups = 4, kStepTaps = 8;
__global__ void __launch_bounds__(1024) dynconv_step_fast_kernel(tnt_dynconv_step_params p) {
  pdl_wait();
  extern __shared__ float sm[];                          // taps [H][K] fp32, then partial sums [4][C]
  const int K = p.K, KM = K - 1, R = p.C / p.H, b = blockIdx.x;
  const int nv = p.C >> 3;
  const int vc = threadIdx.x % nv, g = threadIdx.x / nv;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  float* part = sm + p.H * K;
  const int pos = p.pos ? *p.pos : 0;
  const bf16* ring = reinterpret_cast<const bf16*>(p.ring);
  uint4 rv[kStepTaps];
#pragma unroll
  for (int u = 0; u < kStepTaps; ++u) {
    const int j = g + kStepGroups * u, s = pos - KM + j;
    rv[u] = make_uint4(0, 0, 0, 0);
    if (j < KM && s >= 0) rv[u] = *reinterpret_cast<const uint4*>(ring + (static_cast<long>(s % KM) * p.B + b) * p.C + vc * 8);
  }
  uint4 xraw = make_uint4(0, 0, 0, 0);
  if (g == 0) xraw = *reinterpret_cast<const uint4*>(reinterpret_cast<const bf16*>(p.x) + static_cast<long>(b) * p.ldx + vc * 8);
  const bf16* lg = reinterpret_cast<const bf16*>(p.logits) + static_cast<long>(b) * p.ldl;
  for (int h = warp; h < p.H; h += nwarps) warp_softmax_taps(lg + h * K, K, sm + h * K, lane);
  __syncthreads();
  const float* w = sm + ((vc * 8) / R) * K;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int u = 0; u < kStepTaps; ++u) {
    const int j = g + kStepGroups * u;
    if (j < KM) {
      const float wj = w[j];
      const uint32_t q[4] = {rv[u].x, rv[u].y, rv[u].z, rv[u].w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        acc[2 * i] = fmaf(wj, __uint_as_float(q[i] << 16), acc[2 * i]);
        acc[2 * i + 1] = fmaf(wj, __uint_as_float(q[i] & 0xffff0000u), acc[2 * i + 1]);
      }
    }
  }
  if (g == 0) {
    const float wl = w[K - 1];
    const uint32_t q[4] = {xraw.x, xraw.y, xraw.z, xraw.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      acc[2 * i] = fmaf(wl, __uint_as_float(q[i] << 16), acc[2 * i]);
      acc[2 * i + 1] = fmaf(wl, __uint_as_float(q[i] & 0xffff0000u), acc[2 * i + 1]);
    }
    if (KM > 0) *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(p.ring) + (static_cast<long>(pos % KM) * p.B + b) * p.C + vc * 8) = xraw;
  }
  float* mine = part + g * p.C + vc * 8;
  *reinterpret_cast<float4*>(mine) = make_float4(acc[0], acc[1], acc[2], acc[3]);
  *reinterpret_cast<float4*>(mine + 4) = make_float4(acc[4], acc[5], acc[6], acc[7]);
  __syncthreads();
  if (g == 0) {
    float o[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) o[i] = acc[i];
#pragma unroll
    for (int gg = 1; gg < kStepGroups; ++gg) {
      const float4 a0 = *reinterpret_cast<const float4*>(part + gg * p.C + vc * 8);
      const float4 a1 = *reinterpret_cast<const float4*>(part + gg * p.C + vc * 8 + 4);
      o[0] += a0.x; o[1] += a0.y; o[2] += a0.z; o[3] += a0.w;
      o[4] += a1.x; o[5] += a1.y; o[6] += a1.z; o[7] += a1.w;
    }
    store8(reinterpret_cast<bf16*>(p.y) + static_cast<long>(b) * p.ldy + vc * 8, o);
  }
}

// ------------------------------------------------------------------------------ backward
// smem: xs [(TT+K-1)][CCP] (inputs t0-(K-1) .. t0+TT-1), dys [(TT+K-1)][CCP] (grads t0 .. t0+TT+K-2),
//       taps [(TT+K-1)][HG][K] for t0 .. t0+TT+K-2, dws [TT][HG][K]; CCP = CC + 1 (bank spread).
template <typename T>
__global__ void __launch_bounds__(kThreads) dynconv_bwd_kernel(tnt_dynconv_bwd_params p, int CC, int HG, DropSpec dr) {
  extern __shared__ float sm[];
  const int K = p.K, R = p.C / p.H, B = p.B, KM = K - 1;
  const int b = blockIdx.x, c0 = blockIdx.y * CC, t0 = blockIdx.z * kTT;
  const int TT = min(kTT, p.T - t0);
  const int CCP = CC + 1;
  const int rows = TT + KM;
  float* xs = sm;
  float* dys = xs + static_cast<size_t>(kTT + KM) * CCP;
  float* taps = dys + static_cast<size_t>(kTT + KM) * CCP;
  float* dws = taps + static_cast<size_t>(kTT + KM) * HG * K;
  const T* x = reinterpret_cast<const T*>(p.x);
  const T* dy = reinterpret_cast<const T*>(p.dy);
  const T* lg = reinterpret_cast<const T*>(p.logits);
  const int cc = min(CC, p.C - c0), nv = cc >> 3, hg = cc / R, h0 = c0 / R;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  for (int i = threadIdx.x; i < rows * nv; i += kThreads) {
    const int r = i / nv, vc = i - r * nv;
    float v[8];
    const int ts = t0 - KM + r;
    if (ts >= 0) {
      load8(x + (static_cast<long>(ts) * B + b) * p.ldx + c0 + vc * 8, v);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) xs[r * CCP + vc * 8 + j] = v[j];
    const int tf = t0 + r;
    if (tf < p.T) {
      load8(dy + (static_cast<long>(tf) * B + b) * p.lddy + c0 + vc * 8, v);
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) dys[r * CCP + vc * 8 + j] = v[j];
  }
  stage_tap_logits(lg, p.ldl, B, b, t0, p.T, rows, hg, HG, K, h0, taps);
  __syncthreads();
  normalise_taps(t0, p.T, rows, hg, HG, K, taps, dr, B, b, p.H, h0, true);      // dropped taps carry the sign bit
  __syncthreads();
  const float dscale = dr.thr ? dr.scale : 1.f;

  // dx[ts] = sum_j w[ts+KM-j][j] * dy[ts+KM-j]   (value path)
  T* dx = reinterpret_cast<T*>(p.dx);
  for (int i = threadIdx.x; i < TT * nv; i += kThreads) {
    const int t = i / nv, vc = i - t * nv;
    const int hl = (vc * 8) / R;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int j = 0; j < K; ++j) {
      const int r = t + KM - j;                     // local row of the consuming timestep
      const float wj = tap_kept(taps[(r * HG + hl) * K + j], dscale);
      const float* d = dys + r * CCP + vc * 8;
#pragma unroll
      for (int e = 0; e < 8; ++e) acc[e] = fmaf(wj, d[e], acc[e]);
    }
    store8(dx + (static_cast<long>(t0 + t) * B + b) * p.lddx + c0 + vc * 8, acc);
  }
  // dw[t][h][j] = sum_r dy[t][hR+r] * x[t-KM+j][hR+r] : one warp per (t, head), one lane per tap
  for (int i = warp; i < TT * hg; i += kThreads / 32) {
    const int t = i / hg, hl = i - t * hg;
    for (int jb = 0; jb < K; jb += 32) {
      const int j = jb + lane;
      float s = 0.f;
      if (j < K) {
        const float* xr = xs + (t + j) * CCP + hl * R;     // local row of input t-KM+j is t+j
        const float* d = dys + t * CCP + hl * R;
        for (int r = 0; r < R; ++r) s = fmaf(d[r], xr[r], s);
        dws[(t * HG + hl) * K + j] = (__float_as_uint(taps[(t * HG + hl) * K + j]) >> 31) ? 0.f : s * dscale;
      }
    }
    __syncwarp();
    // softmax backward: dl_j = w_j * (dw_j - sum_i w_i dw_i)
    const float* w = taps + (t * HG + hl) * K;
    const float* dwp = dws + (t * HG + hl) * K;
    float dot = 0.f;
    for (int j = lane; j < K; j += 32) dot = fmaf(fabsf(w[j]), dwp[j], dot);
    dot = warp_sum(dot);
    T* dl = reinterpret_cast<T*>(p.dlogits) + (static_cast<long>(t0 + t) * B + b) * p.lddl + (h0 + hl) * K;
    for (int j = lane; j < K; j += 32) dl[j] = from_f<T>(fabsf(w[j]) * (dwp[j] - dot));
  }
}

int pick_group(int C, int H, int target, int* CC, int* HG) {
  const int R = C / H;
  int hg = target / R;
  if (hg < 1) hg = 1;
  if (hg > H) hg = H;
  *HG = hg;
  *CC = hg * R;
  return (H + hg - 1) / hg;
}

template <typename K>
int set_smem(K kern, size_t bytes) {
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes));
    if (e != cudaSuccess) {
      set_error("dynconv: %zu bytes of shared memory unavailable (%s)", bytes, cudaGetErrorString(e));
      return TNT_ERR_CUDA;
    }
  }
  return TNT_OK;
}
}  // namespace

static int check_common(const void* x, long ldx, int T, int B, int C, int H, int K, int dtype) {
  TNT_CHECK_ARG(x && T >= 0 && B > 0 && C > 0 && H > 0 && K > 0, "dynconv: bad arguments");
  TNT_CHECK_ARG(C % H == 0 && (C / H) % 8 == 0, "dynconv: channels per head (%d/%d) must be a multiple of 8", C, H);
  TNT_CHECK_ARG(dtype == TNT_F32 || dtype == TNT_BF16, "dynconv: dtype");
  TNT_CHECK_ALIGN(aligned16(x) && ldx % 8 == 0, "dynconv: x must be 16-byte aligned, ldx %% 8 == 0");
  return TNT_OK;
}

}  // namespace tnt

extern "C" int tnt_dynconv_fwd(const tnt_dynconv_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p, "dynconv: null params");
  TNT_TRY(check_common(p->x, p->ldx, p->T, p->B, p->C, p->H, p->K, p->dtype));
  TNT_CHECK_ARG(p->logits && p->y, "dynconv: null logits / y");
  TNT_CHECK_ALIGN(aligned16(p->y) && p->ldy % 8 == 0, "dynconv: y alignment");
  TNT_CHECK_ARG(!p->hist || p->hist_len <= p->K - 1, "dynconv: hist_len > K-1");
  if (p->T == 0) return TNT_OK;
  int CC, HG;
  const int ngroups = pick_group(p->C, p->H, 256, &CC, &HG);
  const size_t smem = static_cast<size_t>(kTT + p->K - 1) * CC * dtype_size(p->dtype) + static_cast<size_t>(kTT) * HG * p->K * 4;
  dim3 grid(p->B, ngroups, (p->T + kTT - 1) / kTT);
  TNT_CHECK_ARG(p->drop.p >= 0.f && p->drop.p < 1.f, "dynconv: dropout p = %f", p->drop.p);
  const DropSpec dr = make_drop(p->drop);
  static const bool tc_on = [] { const char* e = getenv("TNT_DYNCONV_TC"); return !(e && e[0] == '0'); }();
  if (tc_on && p->dtype == TNT_BF16 && p->C / p->H == tcv::R && p->C % tcv::CC == 0 && p->K <= 33 && p->ldy % 8 == 0 && p->ldl % 2 == 0 &&
      (reinterpret_cast<uintptr_t>(p->logits) & 3) == 0) {
    const int KW = (tcv::TT + p->K - 1 + 15) & ~15;
    const size_t sm = static_cast<size_t>(KW) * tcv::XS_STRIDE + static_cast<size_t>(tcv::HG) * tcv::TT * (KW * 2 + 16) +
                      static_cast<size_t>(tcv::TT) * (((tcv::HG * p->K * 2 + 15) & ~15) + 16) + 16;   // tap rows + the mbarrier
    TNT_TRY(set_smem(tcv::dynconv_fwd_tc_kernel, sm));
    dim3 g2(p->B, p->C / tcv::CC, (p->T + tcv::TT - 1) / tcv::TT);
    TNT_CUDA(launch_pdl(tcv::dynconv_fwd_tc_kernel, g2, dim3(kThreads), sm, as_stream(stream), 1, *p, KW, dr));
    TNT_LAUNCHED();
    variant_hit(V_DYNCONV_FWD_TC);
    return TNT_OK;
  }
  if (p->dtype == TNT_BF16) {
    TNT_TRY(set_smem(dynconv_fwd_kernel<bf16>, smem));
    TNT_CUDA(launch_pdl(dynconv_fwd_kernel<bf16>, grid, dim3(kThreads), smem, as_stream(stream), 1, *p, CC, HG, dr));
  } else {
    TNT_TRY(set_smem(dynconv_fwd_kernel<float>, smem));
    TNT_CUDA(launch_pdl(dynconv_fwd_kernel<float>, grid, dim3(kThreads), smem, as_stream(stream), 1, *p, CC, HG, dr));
  }
  TNT_LAUNCHED();
  variant_hit(V_DYNCONV_FWD_SIMT);
  return TNT_OK;
}

extern "C" int tnt_dynconv_step(const tnt_dynconv_step_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p, "dynconv_step: null params");
  TNT_TRY(check_common(p->x, p->ldx, 1, p->B, p->C, p->H, p->K, p->dtype));
  TNT_CHECK_ARG(p->logits && p->y && (p->ring || p->K == 1), "dynconv_step: null logits / y / ring");
  const size_t smem = static_cast<size_t>(p->H) * p->K * 4;
  static const bool fast_on = [] { const char* e = getenv("TNT_DYNCONV_STEP_FAST"); return !(e && e[0] == '0'); }();
  const int nv = p->C / 8;
  if (fast_on && p->dtype == TNT_BF16 && p->K <= kStepGroups * kStepTaps + 1 && nv * kStepGroups <= 1024 && (nv * kStepGroups) % 32 == 0 &&
      (p->H * p->K) % 4 == 0 && p->ldy % 8 == 0 && p->ring) {
    const size_t sm2 = smem + static_cast<size_t>(kStepGroups) * p->C * 4;
    TNT_TRY(set_smem(dynconv_step_fast_kernel, sm2));
    TNT_CUDA(launch_pdl(dynconv_step_fast_kernel, dim3(p->B), dim3(nv * kStepGroups), sm2, as_stream(stream), 1, *p));
    TNT_LAUNCHED();
    variant_hit(V_DYNCONV_STEP_FAST);
    return TNT_OK;
  }
  if (p->dtype == TNT_BF16)
    TNT_CUDA(launch_pdl(dynconv_step_kernel<bf16>, dim3(p->B), dim3(kThreads), smem, as_stream(stream), 1, *p));
  else
    TNT_CUDA(launch_pdl(dynconv_step_kernel<float>, dim3(p->B), dim3(kThreads), smem, as_stream(stream), 1, *p));
  TNT_LAUNCHED();
  variant_hit(V_DYNCONV_STEP_SIMT);
  return TNT_OK;
}

extern "C" int tnt_dynconv_bwd(const tnt_dynconv_bwd_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p, "dynconv_bwd: null params");
  TNT_TRY(check_common(p->x, p->ldx, p->T, p->B, p->C, p->H, p->K, p->dtype));
  TNT_CHECK_ARG(p->logits && p->dy && p->dx && p->dlogits, "dynconv_bwd: null pointer");
  TNT_CHECK_ALIGN(aligned16(p->dy) && aligned16(p->dx) && p->lddy % 8 == 0 && p->lddx % 8 == 0,
                  "dynconv_bwd: dy / dx alignment");
  if (p->T == 0) return TNT_OK;
  int CC, HG;
  const int ngroups = pick_group(p->C, p->H, 128, &CC, &HG);
  const int rows = kTT + p->K - 1;
  const size_t smem = (2 * static_cast<size_t>(rows) * (CC + 1) + static_cast<size_t>(rows) * HG * p->K +
                       static_cast<size_t>(kTT) * HG * p->K) * 4;
  dim3 grid(p->B, ngroups, (p->T + kTT - 1) / kTT);
  TNT_CHECK_ARG(p->drop.p >= 0.f && p->drop.p < 1.f, "dynconv_bwd: dropout p = %f", p->drop.p);
  const DropSpec dr = make_drop(p->drop);
  static const bool tc_on = [] { const char* e = getenv("TNT_DYNCONV_TC"); return !(e && e[0] == '0'); }();
  if (tc_on && p->dtype == TNT_BF16 && p->C / p->H == tcv::R && p->C % tcv::CC == 0 && p->K <= 33 && p->ldl % 2 == 0 &&
      p->lddl % 2 == 0 && ((reinterpret_cast<uintptr_t>(p->logits) | reinterpret_cast<uintptr_t>(p->dlogits)) & 3) == 0) {
    const int KW = (tcv::TT + p->K - 1 + 15) & ~15;
    const size_t lraw_b = (static_cast<size_t>(KW) * tcv::HG * p->K * 2 + 15) & ~size_t(15);
    const size_t band_b = std::max<size_t>(static_cast<size_t>(tcv::HG) * tcv::TT * (KW * 2 + 16),
                                           static_cast<size_t>(tcv::TT) * tcv::HG * p->K * 4);
    const size_t sm = 2 * static_cast<size_t>(KW) * tcv::XS_STRIDE + band_b + lraw_b + static_cast<size_t>(KW) * tcv::HG * 8 + 16;   // + the windows' mbarrier
    TNT_TRY(set_smem(tcv::dynconv_bwd_tc_kernel, sm));
    dim3 g2(p->B, p->C / tcv::CC, (p->T + tcv::TT - 1) / tcv::TT);
    tcv::dynconv_bwd_tc_kernel<<<g2, kThreads, sm, as_stream(stream)>>>(*p, KW, dr);
    TNT_LAUNCHED();
    variant_hit(V_DYNCONV_BWD_TC);
    return TNT_OK;
  }
  if (p->dtype == TNT_BF16) {
    TNT_TRY(set_smem(dynconv_bwd_kernel<bf16>, smem));
    dynconv_bwd_kernel<bf16><<<grid, kThreads, smem, as_stream(stream)>>>(*p, CC, HG, dr);
  } else {
    TNT_TRY(set_smem(dynconv_bwd_kernel<float>, smem));
    dynconv_bwd_kernel<float><<<grid, kThreads, smem, as_stream(stream)>>>(*p, CC, HG, dr);
  }
  TNT_LAUNCHED();
  variant_hit(V_DYNCONV_BWD_SIMT);
  return TNT_OK;
}
