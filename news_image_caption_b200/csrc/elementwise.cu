// HBM-bound glue kernels: LayerNorm (+backward), embedding gather, weight-norm fold, casts.
// One warp per row, 128-bit loads, fp32 statistics.
#include "common.cuh"

namespace tnt {
namespace {

constexpr int kWarps = 8;

// ------------------------------------------------------------------------------ LayerNorm fwd
// nn.LayerNorm (decoder_flattened.py:245-268) and final* LayerNorm (transformer_v2.py:82-94).
template <typename T, int MAXV>
__global__ void __launch_bounds__(kWarps * 32) layernorm_fwd_kernel(tnt_layernorm_params p) {
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (row >= p.rows) return;
  const T* x = reinterpret_cast<const T*>(p.x) + static_cast<long>(row) * p.ldx;
  const int nv = p.C >> 3;
  float v[MAXV > 0 ? MAXV : 1][8];
  float sum = 0.f;
  if (MAXV > 0) {
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      const int vc = lane + i * 32;
      if (vc < nv) {
        load8(x + vc * 8, v[i]);
#pragma unroll
        for (int j = 0; j < 8; ++j) sum += v[i][j];
      }
    }
  } else {
    for (int vc = lane; vc < nv; vc += 32) {
      float t[8];
      load8(x + vc * 8, t);
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += t[j];
    }
  }
  const float mean = warp_sum(sum) / p.C;
  float sq = 0.f;
  if (MAXV > 0) {
#pragma unroll
    for (int i = 0; i < MAXV; ++i) {
      if (lane + i * 32 < nv) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = v[i][j] - mean;
          sq = fmaf(d, d, sq);
        }
      }
    }
  } else {
    for (int vc = lane; vc < nv; vc += 32) {
      float t[8];
      load8(x + vc * 8, t);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float d = t[j] - mean;
        sq = fmaf(d, d, sq);
      }
    }
  }
  sq = warp_sum(sq);
  const float rstd = p.variant == 0 ? rsqrtf(sq / p.C + p.eps) : 1.f / (sqrtf(sq / (p.C - 1)) + p.eps);
  long orow = row;
  if (p.transpose_T > 0) {
    const int t = row / p.transpose_B, b = row - t * p.transpose_B;
    orow = static_cast<long>(b) * p.transpose_T + t;
  }
  T* y = reinterpret_cast<T*>(p.y) + orow * p.ldy;
  const float* gm = p.gamma + static_cast<long>(p.groups > 1 ? row % p.groups : 0) * p.C;
  const float* bt = p.beta + static_cast<long>(p.groups > 1 ? row % p.groups : 0) * p.C;
  auto emit = [&](int vc, float (&t)[8]) {
    float gg[8], bb[8], o[8];
    load8(gm + vc * 8, gg);
    load8(bt + vc * 8, bb);
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = (t[j] - mean) * rstd * gg[j] + bb[j];
    store8(y + vc * 8, o);
  };
  if (MAXV > 0) {
#pragma unroll
    for (int i = 0; i < MAXV; ++i)
      if (lane + i * 32 < nv) emit(lane + i * 32, v[i]);
  } else {
    for (int vc = lane; vc < nv; vc += 32) {
      float t[8];
      load8(x + vc * 8, t);
      emit(vc, t);
    }
  }
}

// ------------------------------------------------------------------------------ LayerNorm bwd
// dgamma/dbeta: per-thread partials over the block's rows -> one atomicAdd per (block, channel).
template <typename T>
__global__ void __launch_bounds__(kWarps * 32) layernorm_bwd_kernel(tnt_layernorm_bwd_params p) {
  extern __shared__ float sm[];                        // [2][C] block partials
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nv = p.C >> 3;
  for (int i = threadIdx.x; i < 2 * p.C; i += blockDim.x) sm[i] = 0.f;
  __syncthreads();
  for (int row = blockIdx.x * kWarps + warp; row < p.rows; row += gridDim.x * kWarps) {
    const T* x = reinterpret_cast<const T*>(p.x) + static_cast<long>(row) * p.ldx;
    const T* dy = reinterpret_cast<const T*>(p.dy) + static_cast<long>(row) * p.lddy;
    T* dx = reinterpret_cast<T*>(p.dx) + static_cast<long>(row) * p.lddx;
    float sum = 0.f;
    for (int vc = lane; vc < nv; vc += 32) {
      float t[8];
      load8(x + vc * 8, t);
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += t[j];
    }
    const float mean = warp_sum(sum) / p.C;
    float sq = 0.f;
    for (int vc = lane; vc < nv; vc += 32) {
      float t[8];
      load8(x + vc * 8, t);
#pragma unroll
      for (int j = 0; j < 8; ++j) sq = fmaf(t[j] - mean, t[j] - mean, sq);
    }
    // variant 0 (nn.LayerNorm): xhat = (x - mean) / sqrt(var_biased + eps);  variant 1 (final* LayerNorm,
    // transformer_v2.py:82-94): xhat = (x - mean) / (std_unbiased + eps).  Both: dx = rstd (dxhat - mean(dxhat)) - xhat c2
    const float ssq = warp_sum(sq);
    const float sdev = sqrtf(ssq / (p.C - 1));
    const float rstd = p.variant == 1 ? 1.f / (sdev + p.eps) : rsqrtf(ssq / p.C + p.eps);
    float s1 = 0.f, s2 = 0.f;                          // sum(dxhat), sum(dxhat * xhat)
    for (int vc = lane; vc < nv; vc += 32) {
      float t[8], d[8], g[8];
      load8(x + vc * 8, t);
      load8(dy + vc * 8, d);
      load8(p.gamma + vc * 8, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (t[j] - mean) * rstd, dxh = d[j] * g[j];
        s1 += dxh;
        s2 = fmaf(dxh, xh, s2);
        atomicAdd(&sm[vc * 8 + j], d[j] * xh);
        atomicAdd(&sm[p.C + vc * 8 + j], d[j]);
      }
    }
    s1 = warp_sum(s1) / p.C;
    s2 = warp_sum(s2);
    const float c2 = p.variant == 1 ? (sdev > 0.f ? s2 / ((p.C - 1) * sdev) : 0.f) : rstd * s2 / p.C;
    for (int vc = lane; vc < nv; vc += 32) {
      float t[8], d[8], g[8], o[8];
      load8(x + vc * 8, t);
      load8(dy + vc * 8, d);
      load8(p.gamma + vc * 8, g);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (t[j] - mean) * rstd;
        o[j] = rstd * (d[j] * g[j] - s1) - xh * c2;
      }
      store8(dx + vc * 8, o);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < p.C; i += blockDim.x) {
    atomicAdd(p.dgamma + i, sm[i]);
    atomicAdd(p.dbeta + i, sm[p.C + i]);
  }
}

// C <= 1024, column-owning form: thread j of a 128-thread CTA owns the 8 columns [8j, 8j+8) of R rows at a time, so the
// gamma / beta gradients are 16 registers per thread (the warp-per-row kernel below keeps 64 + the row itself: 1 CTA of 8
// warps per SM = 8 rows = 32 KB in flight, 30 % of the HBM peak at 32768 rows).  Here x / dy stay PACKED (bf16) between the
// two reduction phases: ~140 registers -> 3 CTAs per SM x R = 8 rows = 24 rows (96 KB) in flight per SM.  Row statistics
// are block reductions over the 4 warps: phase A {sum x, sum dy g}, phase B {sum (x - mean)^2, sum dy g (x - mean)}.
template <typename T> struct LnPack;
template <> struct LnPack<bf16> { static constexpr int R = 4; using V = uint4; };
template <> struct LnPack<float> { static constexpr int R = 4; struct V { float4 a, b; }; };
__device__ __forceinline__ void ln_unpack(const uint4& u, float (&v)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ void ln_unpack(const LnPack<float>::V& u, float (&v)[8]) {
  v[0] = u.a.x; v[1] = u.a.y; v[2] = u.a.z; v[3] = u.a.w; v[4] = u.b.x; v[5] = u.b.y; v[6] = u.b.z; v[7] = u.b.w;
}
__device__ __forceinline__ void ln_load(const bf16* p, uint4& u) { u = *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void ln_load(const float* p, LnPack<float>::V& u) {
  u.a = *reinterpret_cast<const float4*>(p);
  u.b = *reinterpret_cast<const float4*>(p + 4);
}
__device__ __forceinline__ void ln_zero(uint4& u) { u = make_uint4(0, 0, 0, 0); }
__device__ __forceinline__ void ln_zero(LnPack<float>::V& u) { u.a = u.b = make_float4(0.f, 0.f, 0.f, 0.f); }

// Per-CTA column partials of dgamma / dbeta go to a static scratch, a second tiny kernel sums them: with atomics every one of
// the 2 C addresses took one atomic from EACH CTA (592-way same-address serialisation at ~80 ns: ~50 us of a 104 us launch at
// 32768 rows, and the whole 15 -> 47 us at 2048 rows once the CTAs got smaller).  One scratch set per launch parity: launches
// on ONE stream are serialised anyway; concurrent LayerNorm backward launches on different streams are not supported.
constexpr int kLnMaxCtas = 1024;
__device__ float g_ln_scratch[2][kLnMaxCtas * 2 * 1024];

__global__ void __launch_bounds__(256) layernorm_bwd_finish_kernel(const float* scr, int n_ctas, int C, float* dgamma, float* dbeta) {
  const int i = blockIdx.x * 256 + threadIdx.x;              // column of [dgamma | dbeta]
  if (i >= 2 * C) return;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  // blockIdx.y takes a slice of the CTAs' partials (one atomic per slice and column: <= 32-way contention)
  const int per = (n_ctas + gridDim.y - 1) / gridDim.y;
  int b = blockIdx.y * per;
  n_ctas = min(n_ctas, b + per);
  for (; b + 3 < n_ctas; b += 4) {
    a0 += scr[static_cast<long>(b) * 2 * C + i];
    a1 += scr[static_cast<long>(b + 1) * 2 * C + i];
    a2 += scr[static_cast<long>(b + 2) * 2 * C + i];
    a3 += scr[static_cast<long>(b + 3) * 2 * C + i];
  }
  for (; b < n_ctas; ++b) a0 += scr[static_cast<long>(b) * 2 * C + i];
  const float s = (a0 + a1) + (a2 + a3);
  atomicAdd(i < C ? dgamma + i : dbeta + (i - C), s);         // the contract: accumulate into zeroed buffers
}

template <typename T>
__global__ void __launch_bounds__(128, 3) layernorm_bwd_cols_kernel(tnt_layernorm_bwd_params p, float* scr) {
  constexpr int R = LnPack<T>::R;
  using V = typename LnPack<T>::V;
  __shared__ float redA[4][2 * R], redB[4][2 * R];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool act = tid * 8 < p.C;
  float g[8], dg[8], db[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) g[j] = dg[j] = db[j] = 0.f;
  if (act) load8(p.gamma + tid * 8, g);
  if (p.zero_param_grads && blockIdx.x == 0)     // the finish kernel (next launch in the stream) accumulates into them
    for (int i = tid; i < p.C; i += 128) p.dgamma[i] = p.dbeta[i] = 0.f;
  const float invC = 1.f / p.C;
  auto fetch = [&](int row0, V (&xv)[R], V (&dv)[R]) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      if (act && row0 + r < p.rows) {
        ln_load(reinterpret_cast<const T*>(p.x) + static_cast<long>(row0 + r) * p.ldx + tid * 8, xv[r]);
        ln_load(reinterpret_cast<const T*>(p.dy) + static_cast<long>(row0 + r) * p.lddy + tid * 8, dv[r]);
      } else {
        ln_zero(xv[r]);
        ln_zero(dv[r]);
      }
    }
  };
  V xr[R], dr[R], xn[R], dn[R];
  fetch(blockIdx.x * R, xr, dr);
  for (int row0 = blockIdx.x * R; row0 < p.rows; row0 += gridDim.x * R) {
    // the NEXT batch's rows are requested before this batch's reductions: loads stay in flight through the two barriers
    fetch(row0 + gridDim.x * R, xn, dn);
    // ---- phase A: sum x, sum dy*g per row
    float sa[2 * R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float x[8], d[8];
      ln_unpack(xr[r], x);
      ln_unpack(dr[r], d);
      float a = 0.f, c = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        a += x[j];
        c = fmaf(d[j], g[j], c);
      }
      sa[r] = a;
      sa[R + r] = c;
    }
#pragma unroll
    for (int i = 0; i < 2 * R; ++i) sa[i] = warp_sum(sa[i]);
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 2 * R; ++i) redA[warp][i] = sa[i];
    }
    __syncthreads();
    float mean[R], s1[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      mean[r] = (redA[0][r] + redA[1][r] + redA[2][r] + redA[3][r]) * invC;
      s1[r] = (redA[0][R + r] + redA[1][R + r] + redA[2][R + r] + redA[3][R + r]) * invC;
    }
    // ---- phase B: sum (x - mean)^2, sum dy*g*(x - mean)
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float x[8], d[8];
      ln_unpack(xr[r], x);
      ln_unpack(dr[r], d);
      float b = 0.f, e = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xc = act ? x[j] - mean[r] : 0.f;
        b = fmaf(xc, xc, b);
        e = fmaf(d[j] * g[j], xc, e);
      }
      sa[r] = b;
      sa[R + r] = e;
    }
#pragma unroll
    for (int i = 0; i < 2 * R; ++i) sa[i] = warp_sum(sa[i]);
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 2 * R; ++i) redB[warp][i] = sa[i];
    }
    __syncthreads();
    // ---- dx, and the column partials of dgamma / dbeta
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float ssq = redB[0][r] + redB[1][r] + redB[2][r] + redB[3][r];
      const float e = redB[0][R + r] + redB[1][R + r] + redB[2][R + r] + redB[3][R + r];
      const float sdev = sqrtf(ssq / (p.C - 1));
      const float rstd = p.variant == 1 ? 1.f / (sdev + p.eps) : rsqrtf(ssq * invC + p.eps);
      const float s2 = rstd * e;                                   // sum dxhat * xhat
      const float c2 = p.variant == 1 ? (sdev > 0.f ? s2 / ((p.C - 1) * sdev) : 0.f) : rstd * s2 * invC;
      if (act && row0 + r < p.rows) {
        float x[8], d[8], o[8];
        ln_unpack(xr[r], x);
        ln_unpack(dr[r], d);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float xh = (x[j] - mean[r]) * rstd;
          o[j] = rstd * (d[j] * g[j] - s1[r]) - xh * c2;
          dg[j] = fmaf(d[j], xh, dg[j]);
          db[j] += d[j];
        }
        store8(reinterpret_cast<T*>(p.dx) + static_cast<long>(row0 + r) * p.lddx + tid * 8, o);
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      xr[r] = xn[r];
      dr[r] = dn[r];
    }
  }
  if (act) {
    float* mine = scr + static_cast<long>(blockIdx.x) * 2 * p.C + tid * 8;
    store8(mine, dg);
    store8(mine + p.C, db);
  }
}

// C <= 1024: the row (x, dy) lives in registers -- one HBM read of each, no re-reads -- and every lane keeps the
// dgamma / dbeta partials of ITS 32 channels in registers across all rows of its warp; the block combines the warps'
// partials through shared memory once at the end (the generic kernel above issues 2*C shared atomics per row).
template <typename T>
__global__ void __launch_bounds__(kWarps * 32) layernorm_bwd_small_kernel(tnt_layernorm_bwd_params p) {
  extern __shared__ float sm[];                        // [kWarps][2][C]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nv = p.C >> 3;
  float g[4][8], dg[4][8], db[4][8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int vc = lane + 32 * i;
#pragma unroll
    for (int j = 0; j < 8; ++j) dg[i][j] = db[i][j] = 0.f, g[i][j] = 0.f;
    if (vc < nv) load8(p.gamma + vc * 8, g[i]);
  }
  const float invC = 1.f / p.C;
  for (int row = blockIdx.x * kWarps + warp; row < p.rows; row += gridDim.x * kWarps) {
    const T* x = reinterpret_cast<const T*>(p.x) + static_cast<long>(row) * p.ldx;
    const T* dy = reinterpret_cast<const T*>(p.dy) + static_cast<long>(row) * p.lddy;
    T* dx = reinterpret_cast<T*>(p.dx) + static_cast<long>(row) * p.lddx;
    float t[4][8], d[4][8];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int vc = lane + 32 * i;
      if (vc < nv) {
        load8(x + vc * 8, t[i]);
        load8(dy + vc * 8, d[i]);
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) t[i][j] = d[i][j] = 0.f;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += t[i][j];
    }
    const float mean = warp_sum(sum) * invC;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (lane + 32 * i < nv) {
#pragma unroll
        for (int j = 0; j < 8; ++j) sq = fmaf(t[i][j] - mean, t[i][j] - mean, sq);
      }
    const float ssq = warp_sum(sq);
    const float sdev = sqrtf(ssq / (p.C - 1));
    const float rstd = p.variant == 1 ? 1.f / (sdev + p.eps) : rsqrtf(ssq * invC + p.eps);
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float xh = (t[i][j] - mean) * rstd, dxh = d[i][j] * g[i][j];
        t[i][j] = xh;                                  // keep x-hat
        s1 += dxh;
        s2 = fmaf(dxh, xh, s2);
        dg[i][j] = fmaf(d[i][j], xh, dg[i][j]);
        db[i][j] += d[i][j];
      }
    s1 = warp_sum(s1) * invC;
    s2 = warp_sum(s2);
    const float c2 = p.variant == 1 ? (sdev > 0.f ? s2 / ((p.C - 1) * sdev) : 0.f) : rstd * s2 * invC;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int vc = lane + 32 * i;
      if (vc < nv) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * (d[i][j] * g[i][j] - s1) - t[i][j] * c2;
        store8(dx + vc * 8, o);
      }
    }
  }
  float* mine = sm + static_cast<size_t>(warp) * 2 * p.C;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int vc = lane + 32 * i;
    if (vc < nv) {
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        mine[vc * 8 + j] = dg[i][j];
        mine[p.C + vc * 8 + j] = db[i][j];
      }
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * p.C; i += blockDim.x) {
    float a = 0.f;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) a += sm[static_cast<size_t>(w) * 2 * p.C + i];
    atomicAdd(i < p.C ? p.dgamma + i : p.dbeta + (i - p.C), a);
  }
}

// ------------------------------------------------------------------------------ embedding
// adaptive.py:61-76 (pre-projected bands) + positional.py:167-211 + sum_text_field_embedder.py:117
template <typename T>
__global__ void __launch_bounds__(kWarps * 32) embed_kernel(tnt_embed_params p) {
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * kWarps + (threadIdx.x >> 5);      // row = t * B + b
  if (row >= p.T * p.B) return;
  const int t = row / p.B, b = row - t * p.B;
  long id = p.ids[static_cast<long>(b) * p.T + t];
  const int start = p.start_pos ? *p.start_pos : 0;
  int pos = (id == p.pad_idx) ? p.pad_idx : t + p.pad_idx + 1 + start;
  T* o = reinterpret_cast<T*>(p.out) + static_cast<long>(row) * p.ldo;
  if (pos >= p.n_pos) {                                         // the host entry points reject T / gen_len beyond the table
    if (p.error_flags && lane == 0) atomicOr(p.error_flags, TNT_DEVERR_POSITION);
    pos = p.n_pos - 1;
  }
  if (p.vocab > 0 && (id < 0 || id >= p.vocab)) {               // nn.Embedding raises here; never read outside the table
    if (p.error_flags && lane == 0) atomicOr(p.error_flags, TNT_DEVERR_TOKEN_ID);
    const float nanv = __int_as_float(0x7fc00000);
    const float bad[8] = {nanv, nanv, nanv, nanv, nanv, nanv, nanv, nanv};
    for (int vc = lane; vc < (p.d >> 3); vc += 32) store8(o + vc * 8, bad);
    return;
  }
  const T* e = reinterpret_cast<const T*>(p.table) + id * p.d;
  const float* pe = p.pos_table + static_cast<long>(pos) * p.d;
  for (int vc = lane; vc < (p.d >> 3); vc += 32) {
    float a[8], c[8];
    load8(e + vc * 8, a);
    load8(pe + vc * 8, c);
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] += c[j];
    store8(o + vc * 8, a);
  }
}

__device__ __forceinline__ void store4(float* p, float a, float b, float c, float d) {
  *reinterpret_cast<float4*>(p) = make_float4(a, b, c, d);
}
__device__ __forceinline__ void store4(bf16* p, float a, float b, float c, float d) {
  *reinterpret_cast<uint2*>(p) = make_uint2(pack_bf16x2(a, b), pack_bf16x2(c, d));
}
__device__ __forceinline__ bool dev_aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// ------------------------------------------------------------------------------ weight-norm fold
// w = g * v / ||v||_row  (linear.py:30-33); optional GLU pair interleave of the output rows.
// One warp per row.  The row is walked in chunks of 1024 columns whose 8 float4 loads per lane are all issued before the
// first is consumed (a rolled loop serialises one memory latency per iteration: 8.7 us per 4 M-parameter matrix); the first
// chunk stays in registers between the norm pass and the scale pass, later chunks are re-read from L2.
__device__ __forceinline__ void load_chunk8(const float* row, int c0, int cols, int lane, float4 (&a)[8]) {
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int c = c0 + (u * 32 + lane) * 4;
    a[u] = c < cols ? __ldg(reinterpret_cast<const float4*>(row + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}
__device__ __forceinline__ float sumsq8(const float4 (&a)[8], float sq) {
#pragma unroll
  for (int u = 0; u < 8; ++u) sq = fmaf(a[u].x, a[u].x, fmaf(a[u].y, a[u].y, fmaf(a[u].z, a[u].z, fmaf(a[u].w, a[u].w, sq))));
  return sq;
}

template <typename T>
__device__ __forceinline__ void fold_row(const float* v, const float* g, T* out, int rows, int cols, int glu, int row, int lane) {
  const float* vr = v + static_cast<long>(row) * cols;
  const bool vec = (cols & 3) == 0 && dev_aligned16(v) && dev_aligned16(out);
  int orow = row;
  if (glu) orow = row < rows / 2 ? 2 * row : 2 * (row - rows / 2) + 1;
  T* o = out + static_cast<long>(orow) * cols;
  if (vec) {
    float4 a[8];
    load_chunk8(vr, 0, cols, lane, a);
    float scale = 1.f;
    if (g) {
      float sq = sumsq8(a, 0.f);
      for (int c0 = 1024; c0 < cols; c0 += 1024) {
        float4 t[8];
        load_chunk8(vr, c0, cols, lane, t);
        sq = sumsq8(t, sq);
      }
      scale = g[row] / sqrtf(warp_sum(sq));
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int c = (u * 32 + lane) * 4;
      if (c < cols) store4(o + c, a[u].x * scale, a[u].y * scale, a[u].z * scale, a[u].w * scale);
    }
    for (int c0 = 1024; c0 < cols; c0 += 1024) {
      load_chunk8(vr, c0, cols, lane, a);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int c = c0 + (u * 32 + lane) * 4;
        if (c < cols) store4(o + c, a[u].x * scale, a[u].y * scale, a[u].z * scale, a[u].w * scale);
      }
    }
    return;
  }
  float scale = 1.f;
  if (g) {
    float sq = 0.f;
    for (int c = lane; c < cols; c += 32) sq = fmaf(vr[c], vr[c], sq);
    scale = g[row] / sqrtf(warp_sum(sq));
  }
  for (int c = lane; c < cols; c += 32) o[c] = from_f<T>(vr[c] * scale);
}

template <typename T>
__global__ void __launch_bounds__(kWarps * 32) wn_fold_kernel(const float* v, const float* g, T* out, int rows,
                                                              int cols, int glu) {
  const int row = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (row >= rows) return;
  fold_row<T>(v, g, out, rows, cols, glu, row, threadIdx.x & 31);
}

// every job of the weight refresh in one launch: warp-row -> job by binary search over row0
template <typename T>
__global__ void __launch_bounds__(kWarps * 32) wn_fold_batch_kernel(const FoldJob* jobs, int n_jobs, int total_rows) {
  __shared__ int s_row0[1024];                 // n_jobs <= 1024 (engine_load.cu kMaxFoldJobs)
  for (int i = threadIdx.x; i < n_jobs; i += kWarps * 32) s_row0[i] = jobs[i].row0;
  __syncthreads();
  const int wr = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (wr >= total_rows) return;
  int lo = 0, hi = n_jobs - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (s_row0[mid] <= wr) lo = mid;
    else hi = mid - 1;
  }
  const FoldJob j = jobs[lo];
  const int row = wr - j.row0, lane = threadIdx.x & 31;
  if (j.out_f32) fold_row<float>(j.v, j.g, static_cast<float*>(j.out), j.rows, j.cols, j.glu, row, lane);
  else fold_row<T>(j.v, j.g, static_cast<T*>(j.out), j.rows, j.cols, j.glu, row, lane);
}

__global__ void glu_bias_kernel(const float* b, float* out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i < n / 2 ? 2 * i : 2 * (i - n / 2) + 1] = b[i];
}

template <typename S, typename D>
__global__ void cast_kernel(const S* s, D* d, long n) {
  for (long i = blockIdx.x * static_cast<long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long>(gridDim.x) * blockDim.x)
    d[i] = from_f<D>(to_f(s[i]));
}
// 8 elements per thread, 128-bit accesses on the bf16 side, 2 x 128-bit on the fp32 side (both pointers 32-byte aligned);
// the gradient exchange casts 0.6 GB per step through this kernel
template <typename S, typename D>
__global__ void __launch_bounds__(256) cast8_kernel(const S* s, D* d, long n8) {
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n8; i += gridDim.x * 256L) {
    float v[8];
    load8(s + i * 8, v);
    store8(d + i * 8, v);
  }
}

}  // namespace
}  // namespace tnt

extern "C" int tnt_layernorm_fwd(const tnt_layernorm_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p && p->x && p->y && p->gamma && p->beta, "layernorm: null pointer");
  TNT_CHECK_ARG(p->rows >= 0 && p->C > 1 && p->C % 8 == 0, "layernorm: C (%d) must be a multiple of 8", p->C);
  TNT_CHECK_ALIGN(aligned16(p->x) && aligned16(p->y) && p->ldx % 8 == 0 && p->ldy % 8 == 0, "layernorm: alignment");
  if (p->rows == 0) return TNT_OK;
  const int grid = (p->rows + kWarps - 1) / kWarps;
  cudaStream_t st = as_stream(stream);
  const bool small = p->C <= 1024;
  if (p->dtype == TNT_BF16) {
    if (small) TNT_CUDA(launch_pdl(layernorm_fwd_kernel<bf16, 4>, dim3(grid), dim3(kWarps * 32), 0, st, 1, *p));
    else TNT_CUDA(launch_pdl(layernorm_fwd_kernel<bf16, 0>, dim3(grid), dim3(kWarps * 32), 0, st, 1, *p));
  } else {
    if (small) TNT_CUDA(launch_pdl(layernorm_fwd_kernel<float, 4>, dim3(grid), dim3(kWarps * 32), 0, st, 1, *p));
    else TNT_CUDA(launch_pdl(layernorm_fwd_kernel<float, 0>, dim3(grid), dim3(kWarps * 32), 0, st, 1, *p));
  }
  TNT_LAUNCHED();
  return TNT_OK;
}

extern "C" int tnt_layernorm_bwd(const tnt_layernorm_bwd_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p && p->x && p->dy && p->dx && p->gamma && p->dgamma && p->dbeta, "layernorm_bwd: null pointer");
  TNT_CHECK_ARG(p->C % 8 == 0 && p->C <= 6000, "layernorm_bwd: C");
  const bool cols_shape = p->C <= 1024 && aligned16(p->x) && aligned16(p->dy) && aligned16(p->dx) && p->ldx % 8 == 0 &&
                          p->lddy % 8 == 0 && p->lddx % 8 == 0 && (reinterpret_cast<uintptr_t>(p->gamma) & 15) == 0;
  static const bool cols_on0 = [] { const char* e = getenv("TNT_LN_BWD_COLS"); return !(e && e[0] == '0'); }();
  if (p->zero_param_grads && (p->rows == 0 || !(cols_on0 && cols_shape))) {      // only the column kernel zeroes in-kernel
    TNT_CUDA(cudaMemsetAsync(p->dgamma, 0, sizeof(float) * p->C, as_stream(stream)));
    TNT_CUDA(cudaMemsetAsync(p->dbeta, 0, sizeof(float) * p->C, as_stream(stream)));
  }
  if (p->rows == 0) return TNT_OK;
  const int grid = std::min((p->rows + kWarps - 1) / kWarps, 2 * num_sms());
  const size_t smem = 2 * static_cast<size_t>(p->C) * 4;
  static const bool cols_on = [] { const char* e = getenv("TNT_LN_BWD_COLS"); return !(e && e[0] == '0'); }();
  if (cols_on && cols_shape) {
    const int R = 4;
    static const int per_sm_c = [] { const char* e = getenv("TNT_LN_BWD_COLS_CTAS_PER_SM"); return e ? std::max(1, atoi(e)) : 3; }();
    const int grid_c = std::min(std::min((p->rows + R - 1) / R, per_sm_c * num_sms()), kLnMaxCtas);
    static float* scr_base[kMaxDevices] = {};
    static unsigned parity = 0;
    const int dev = cur_device();
    if (!scr_base[dev]) {
      void* sp = nullptr;
      TNT_CUDA(cudaGetSymbolAddress(&sp, g_ln_scratch));
      scr_base[dev] = static_cast<float*>(sp);
    }
    float* scr = scr_base[dev] + static_cast<size_t>(parity++ & 1) * kLnMaxCtas * 2 * 1024;
    if (p->dtype == TNT_BF16) layernorm_bwd_cols_kernel<bf16><<<grid_c, 128, 0, as_stream(stream)>>>(*p, scr);
    else layernorm_bwd_cols_kernel<float><<<grid_c, 128, 0, as_stream(stream)>>>(*p, scr);
    TNT_LAUNCHED();
    static const bool skip_finish = [] { const char* e = getenv("TNT_LN_BWD_SKIP_FINISH"); return e && e[0] == '1'; }();   // timing diagnosis only
    if (!skip_finish)
      layernorm_bwd_finish_kernel<<<dim3((2 * p->C + 255) / 256, std::min(32, (grid_c + 15) / 16)), 256, 0, as_stream(stream)>>>(
          scr, grid_c, p->C, p->dgamma, p->dbeta);
  } else if (p->C <= 1024) {
    const size_t smem_s = static_cast<size_t>(kWarps) * 2 * p->C * 4;           // <= 64 KB
    // every CTA ends with 2*C atomics onto dgamma / dbeta: few CTAs with several rows per warp beat one row per warp
    static const int per_sm = [] { const char* e = getenv("TNT_LN_BWD_CTAS_PER_SM"); return e ? std::max(1, atoi(e)) : 4; }();
    const int grid_s = std::min((p->rows + kWarps - 1) / kWarps, per_sm * num_sms());
    if (p->dtype == TNT_BF16) {
      TNT_CUDA(cudaFuncSetAttribute(layernorm_bwd_small_kernel<bf16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
      layernorm_bwd_small_kernel<bf16><<<grid_s, kWarps * 32, smem_s, as_stream(stream)>>>(*p);
    } else {
      TNT_CUDA(cudaFuncSetAttribute(layernorm_bwd_small_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
      layernorm_bwd_small_kernel<float><<<grid_s, kWarps * 32, smem_s, as_stream(stream)>>>(*p);
    }
  } else if (p->dtype == TNT_BF16)
    layernorm_bwd_kernel<bf16><<<grid, kWarps * 32, smem, as_stream(stream)>>>(*p);
  else
    layernorm_bwd_kernel<float><<<grid, kWarps * 32, smem, as_stream(stream)>>>(*p);
  TNT_LAUNCHED();
  return TNT_OK;
}

extern "C" int tnt_embed_fwd(const tnt_embed_params* p, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(p && p->ids && p->table && p->pos_table && p->out, "embed: null pointer");
  TNT_CHECK_ARG(p->d % 8 == 0 && p->ldo % 8 == 0, "embed: d must be a multiple of 8");
  const int rows = p->B * p->T;
  if (rows == 0) return TNT_OK;
  const int grid = (rows + kWarps - 1) / kWarps;
  if (p->dtype == TNT_BF16) TNT_CUDA(launch_pdl(embed_kernel<bf16>, dim3(grid), dim3(kWarps * 32), 0, as_stream(stream), 1, *p));
  else TNT_CUDA(launch_pdl(embed_kernel<float>, dim3(grid), dim3(kWarps * 32), 0, as_stream(stream), 1, *p));
  TNT_LAUNCHED();
  return TNT_OK;
}

extern "C" int tnt_wn_fold(const float* v, const float* g, void* w_out, int rows, int cols, int glu_interleave,
                           int dtype_out, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(v && w_out && rows > 0 && cols > 0, "wn_fold: bad arguments");
  TNT_CHECK_ARG(!glu_interleave || rows % 2 == 0, "wn_fold: GLU interleave needs an even row count");
  const int grid = (rows + kWarps - 1) / kWarps;
  if (dtype_out == TNT_BF16)
    wn_fold_kernel<bf16><<<grid, kWarps * 32, 0, as_stream(stream)>>>(v, g, reinterpret_cast<bf16*>(w_out), rows, cols,
                                                                      glu_interleave);
  else
    wn_fold_kernel<float><<<grid, kWarps * 32, 0, as_stream(stream)>>>(v, g, reinterpret_cast<float*>(w_out), rows,
                                                                       cols, glu_interleave);
  TNT_LAUNCHED();
  return TNT_OK;
}

namespace tnt {
int wn_fold_batch(const FoldJob* jobs, int n_jobs, int total_rows, int dtype, cudaStream_t st) {
  if (n_jobs == 0 || total_rows == 0) return TNT_OK;
  const int grid = (total_rows + kWarps - 1) / kWarps;
  if (dtype == TNT_BF16) wn_fold_batch_kernel<bf16><<<grid, kWarps * 32, 0, st>>>(jobs, n_jobs, total_rows);
  else wn_fold_batch_kernel<float><<<grid, kWarps * 32, 0, st>>>(jobs, n_jobs, total_rows);
  TNT_LAUNCHED();
  return TNT_OK;
}
}  // namespace tnt

extern "C" int tnt_glu_interleave_bias(const float* b, float* b_out, int n, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(b && b_out && n > 0 && n % 2 == 0, "glu_interleave_bias: bad arguments");
  glu_bias_kernel<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(b, b_out, n);
  TNT_LAUNCHED();
  return TNT_OK;
}

namespace tnt {
namespace {
// Background form of cast8_kernel: a FEW CTAs, each thread keeping 4 vectors in flight.  The gradient exchange casts on a
// side stream while the backward pass owns the GPU: a full-width grid (148 x 16 CTAs) stalled the main stream for the
// length of every cast (50 / 46 / 20 us gaps at each stage, tools/trace_train.py); this one trickles along on <= max_ctas SMs.
template <typename S, typename D>
__global__ void __launch_bounds__(256) cast8_bg_kernel(const S* s, D* d, long n8) {
  const long stride = gridDim.x * 256L;
  long i = blockIdx.x * 256L + threadIdx.x;
  for (; i + 3 * stride < n8; i += 4 * stride) {
    float v[4][8];
#pragma unroll
    for (int u = 0; u < 4; ++u) load8(s + (i + u * stride) * 8, v[u]);
#pragma unroll
    for (int u = 0; u < 4; ++u) store8(d + (i + u * stride) * 8, v[u]);
  }
  for (; i < n8; i += stride) {
    float v[8];
    load8(s + i * 8, v);
    store8(d + i * 8, v);
  }
}
}  // namespace
}  // namespace tnt

extern "C" int tnt_cast_background(const void* src, void* dst, int64_t n, int dtype_src, int dtype_dst, int max_ctas, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(src && dst && n >= 0 && max_ctas > 0, "cast_background: bad arguments");
  if (n == 0) return TNT_OK;
  if (dtype_src == dtype_dst || n % 8 != 0 || ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 31) != 0)
    return tnt_cast(src, dst, n, dtype_src, dtype_dst, stream);
  const int grid = static_cast<int>(std::min<int64_t>((n / 8 + 255) / 256, max_ctas));
  cudaStream_t st = as_stream(stream);
  if (dtype_src == TNT_F32) cast8_bg_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(src), reinterpret_cast<bf16*>(dst), n / 8);
  else cast8_bg_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(src), reinterpret_cast<float*>(dst), n / 8);
  TNT_LAUNCHED();
  return TNT_OK;
}

extern "C" int tnt_cast(const void* src, void* dst, int64_t n, int dtype_src, int dtype_dst, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(src && dst && n >= 0, "cast: bad arguments");
  if (n == 0) return TNT_OK;
  const int grid = static_cast<int>(std::min<int64_t>((n + 255) / 256, 148 * 16));
  cudaStream_t st = as_stream(stream);
  if (dtype_src != dtype_dst && n % 8 == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 31) == 0) {
    const int g8 = static_cast<int>(std::min<int64_t>((n / 8 + 255) / 256, 148 * 16));
    if (dtype_src == TNT_F32) cast8_kernel<<<g8, 256, 0, st>>>(reinterpret_cast<const float*>(src), reinterpret_cast<bf16*>(dst), n / 8);
    else cast8_kernel<<<g8, 256, 0, st>>>(reinterpret_cast<const bf16*>(src), reinterpret_cast<float*>(dst), n / 8);
    TNT_LAUNCHED();
    return TNT_OK;
  }
  if (dtype_src == TNT_F32 && dtype_dst == TNT_BF16)
    cast_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(src), reinterpret_cast<bf16*>(dst), n);
  else if (dtype_src == TNT_BF16 && dtype_dst == TNT_F32)
    cast_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(src), reinterpret_cast<float*>(dst), n);
  else if (dtype_src == TNT_F32 && dtype_dst == TNT_F32)
    cast_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float*>(src), reinterpret_cast<float*>(dst), n);
  else
    cast_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const bf16*>(src), reinterpret_cast<bf16*>(dst), n);
  TNT_LAUNCHED();
  return TNT_OK;
}

// ------------------------------------------------------------------------------ dropout (counter based)
namespace tnt {
namespace {
// out[r, c] = dropout(in[r, c]) (+ res[r, c]); 8 columns per thread (cols % 8 == 0), index = r * ld_idx + c
template <typename T>
__global__ void __launch_bounds__(256) dropout_kernel(const T* in, long ldi, const T* res, long ldr, T* out, long ldo, int rows,
                                                     int cols, DropSpec d) {
  const int nv = cols >> 3;
  const long n = static_cast<long>(rows) * nv;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n; i += gridDim.x * 256L) {
    const long r = i / nv;
    const int c = static_cast<int>(i - r * nv) * 8;
    float v[8];
    load8(in + r * ldi + c, v);
    const unsigned long long i0 = static_cast<unsigned long long>(r) * d.ld + c;
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = drop_val(d, i0 + j, v[j]);
    if (res) {
      float t[8];
      load8(res + r * ldr + c, t);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] += t[j];
    }
    store8(out + r * ldo + c, v);
  }
}
}  // namespace

int dropout_apply(const void* in, long ldi, const void* res, long ldr, void* out, long ldo, int rows, int cols, int dtype,
                  const tnt_dropout& dr, cudaStream_t st) {
  TNT_CHECK_ARG(in && out && rows >= 0 && cols > 0 && cols % 8 == 0, "dropout: bad arguments (cols must be a multiple of 8)");
  TNT_CHECK_ARG(dr.p >= 0.f && dr.p < 1.f, "dropout: p = %f", dr.p);
  const size_t E = dtype_size(dtype);
  TNT_CHECK_ALIGN(aligned16(in) && aligned16(out) && (!res || aligned16(res)) && (ldi * E) % 16 == 0 && (ldo * E) % 16 == 0 &&
                      (!res || (ldr * E) % 16 == 0),
                  "dropout: rows must be 16-byte aligned");
  if (rows == 0) return TNT_OK;
  DropSpec d = make_drop(dr);
  if (d.ld == 0) d.ld = cols;
  if (!d.thr) { d.thr = 0; d.scale = 1.f; }             // p == 0: plain copy (+ residual); drop_bits >= 0 always keeps
  const long n = static_cast<long>(rows) * (cols / 8);
  const int grid = static_cast<int>(std::min<long>((n + 255) / 256, 8L * num_sms()));
  if (dtype == TNT_BF16)
    dropout_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<const bf16*>(in), ldi, static_cast<const bf16*>(res), ldr,
                                               static_cast<bf16*>(out), ldo, rows, cols, d);
  else
    dropout_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(in), ldi, static_cast<const float*>(res), ldr,
                                                static_cast<float*>(out), ldo, rows, cols, d);
  TNT_LAUNCHED();
  return TNT_OK;
}
}  // namespace tnt

extern "C" uint32_t tnt_dropout_bits(uint64_t seed, uint32_t site, uint64_t index) {
  uint32_t k0, k1;
  tnt::drop_keys(seed, site, &k0, &k1);
  return tnt::drop_bits(k0, k1, index);
}

extern "C" int tnt_dropout_apply(const void* in, int64_t ldi, const void* residual, int64_t ldr, void* out, int64_t ldo,
                                 int rows, int cols, int dtype, const tnt_dropout* d, tnt_stream stream) {
  TNT_CHECK_ARG(d, "dropout: null descriptor");
  return tnt::dropout_apply(in, ldi, residual, ldr, out, ldo, rows, cols, dtype, *d, tnt::as_stream(stream));
}
