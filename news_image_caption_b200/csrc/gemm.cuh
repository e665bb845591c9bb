// Device-side argument block + epilogues shared by the tcgen05 GEMM (bf16) and the SIMT GEMM (fp32).
#pragma once
#include "common.cuh"

namespace tnt {

enum : int { EPI_NONE = 0, EPI_RELU = 1, EPI_GLU = 2, EPI_RESIDUAL = 3, EPI_LSE = 4, EPI_RELU_BWD = 5, EPI_SMGRAD = 6 };

struct GemmArgs {
  const void* a; const void* w; const float* bias; const void* res; void* c;
  long lda, ldw, ldr, ldc;
  int M, N, K;
  const int* m_dyn;
  int groups;
  long sa, sw, sbias, sres, sc;
  int vec_ok;               // C (and residual) rows allow 16-byte vector access
  int nacc;                 // tcgen05 single-CTA tiles: interleaved TMEM accumulators (1, 2 or 4; set by the launcher)
  // EPI_LSE only
  const int* target;        // [M] cluster-local target column or < 0
  float4* partials;         // [M, n_tiles] (max, sumexp, argmax bits, target logit or 0)
  int n_tiles;
  int lse_halves;           // tcgen05 kernel: partial records per column tile (2: one per half of the tile's columns)
  int n_lse;                // columns [0, n_lse) take part in max / sum / argmax ...
  float* side; long side_ld;  // ... columns [n_lse, N) are written raw to side[row, col - n_lse]
  // operand layouts: a_mn = A stored [K, M] (TN), b_mn = W stored [K, N] (NN / TN)
  int a_mn, b_mn;
  float alpha;
  // EPI_SMGRAD: out[m, n] = gscale * (exp(z - row_lse[m]) - (n == target[m])); rows in [M_dyn, M) get zeros
  const float* row_lse;
  float gscale;
  const float* row_w;       // optional per-row factor on top of gscale (weighted softmax gradient of the copy heads)
  // dynamic tile scheduler (single-CTA tcgen05 kernel): tile_ctr[0] = next tile, tile_ctr[1] = CTAs done (both return to 0
  // when the launch ends); NULL = static striding
  int* tile_ctr;
  // L2 prefetch hint (weights of the next GEMM of a chain), issued by the prologue
  const char* pf; unsigned long long pf_bytes;
  // dropout on the value the RESIDUAL epilogue adds the residual to / on the RELU epilogue's output (thr == 0: off)
  DropSpec drop;
};

struct LseExtra {
  const int* target; float4* partials; int n_lse; float* side; long side_ld;
  const float* row_lse = nullptr; float gscale = 1.f;     // EPI_SMGRAD
  const float* row_w = nullptr;
};

// (fast reciprocal: the IEEE division compiled to ~20 instructions per gate -- with one epilogue warp per scheduler the GLU
// GEMM of a layer took as long as the twice larger fc1 GEMM: 21.5 vs 12.2 us for the same shape without the gate)
__device__ __forceinline__ float sigmoidf_(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }

// One thread owns NC consecutive accumulator columns [n0, n0+NC) of row m (n0 % NC == 0).
// `bias_n0` points at the bias of column n0 (or is null); with `bias_padded` it is a zero-padded staging copy
// (shared memory in the tcgen05 kernel) that may be read NC wide without a bounds check.
template <int EPI, typename TOut, int NC>
__device__ __forceinline__ void epilogue_store(const GemmArgs& g, int grp, int m, int n0, float (&acc)[NC],
                                               const float* bias_n0, bool bias_padded) {
  static_assert(NC % 8 == 0 || NC == 4, "NC");
  if (g.alpha != 1.f) {
#pragma unroll
    for (int i = 0; i < NC; ++i) acc[i] *= g.alpha;
  }
  if (bias_n0) {
    if (bias_padded) {
#pragma unroll
      for (int i = 0; i < NC; i += 4) {
        const float4 b4 = *reinterpret_cast<const float4*>(bias_n0 + i);
        acc[i] += b4.x; acc[i + 1] += b4.y; acc[i + 2] += b4.z; acc[i + 3] += b4.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (n0 + i < g.N) acc[i] += __ldg(bias_n0 + i);
    }
  }
  TOut* crow = reinterpret_cast<TOut*>(g.c) + grp * g.sc + static_cast<long>(m) * g.ldc;
  if (EPI == EPI_GLU) {
    constexpr int NO = NC / 2;
    float o[NO];
#pragma unroll
    for (int i = 0; i < NO; ++i) o[i] = acc[2 * i] * sigmoidf_(acc[2 * i + 1]);
    const int c0 = n0 >> 1, NOUT = g.N >> 1;
    if (NO % 8 == 0 && g.vec_ok && c0 + NO <= NOUT) {
#pragma unroll
      for (int i = 0; i < NO; i += 8) {
        float t[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) t[j] = o[i + j];
        store8(crow + c0 + i, t);
      }
    } else {
#pragma unroll
      for (int i = 0; i < NO; ++i)
        if (c0 + i < NOUT) crow[c0 + i] = from_f<TOut>(o[i]);
    }
    return;
  }
  if (EPI == EPI_SMGRAD) {
    // softmax gradient of a cluster (softmax.py:169-191 + cross_entropy backward): p - onehot(target)
    const float lse = g.row_lse[m];
    const int tgt = g.target[m];
    const float gs = g.row_w ? g.gscale * g.row_w[m] : g.gscale;
#pragma unroll
    for (int i = 0; i < NC; ++i)
      acc[i] = gs * (__expf(acc[i] - lse) - ((n0 + i) == tgt ? 1.f : 0.f));   // lse = +inf -> ignored row
  }
  if (EPI == EPI_RELU_BWD) {
    const TOut* rrow = reinterpret_cast<const TOut*>(g.res) + grp * g.sres + static_cast<long>(m) * g.ldr;
    // 16-byte loads of the saved activation (scalar loads touch 32 sectors per instruction: the epilogue was 3x the k-loop)
    if (NC % 8 == 0 && g.vec_ok && n0 + NC <= g.N) {
#pragma unroll
      for (int i = 0; i < NC; i += 8) {
        float t[8];
        load8(rrow + n0 + i, t);
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (!(t[j] > 0.f)) acc[i + j] = 0.f;
      }
    } else {
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (n0 + i < g.N && !(to_f(rrow[n0 + i]) > 0.f)) acc[i] = 0.f;
    }
  }
  if (EPI == EPI_RESIDUAL) {
    if (g.drop.thr) {
      const unsigned long long i0 = static_cast<unsigned long long>(grp) * g.drop.gs + static_cast<unsigned long long>(m) * g.drop.ld + n0;
#pragma unroll
      for (int i = 0; i < NC; ++i) acc[i] = drop_val(g.drop, i0 + i, acc[i]);
    }
    const TOut* rrow = reinterpret_cast<const TOut*>(g.res) + grp * g.sres + static_cast<long>(m) * g.ldr;
    if (NC % 8 == 0 && g.vec_ok && n0 + NC <= g.N) {
#pragma unroll
      for (int i = 0; i < NC; i += 8) {
        float t[8];
        load8(rrow + n0 + i, t);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i + j] += t[j];
      }
    } else {
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (n0 + i < g.N) acc[i] += to_f(rrow[n0 + i]);
    }
  }
  if (EPI == EPI_RELU) {
#pragma unroll
    for (int i = 0; i < NC; ++i) acc[i] = fmaxf(acc[i], 0.f);
    if (g.drop.thr) {
      const unsigned long long i0 = static_cast<unsigned long long>(grp) * g.drop.gs + static_cast<unsigned long long>(m) * g.drop.ld + n0;
#pragma unroll
      for (int i = 0; i < NC; ++i) acc[i] = drop_val(g.drop, i0 + i, acc[i]);
    }
  }
  if (NC % 8 == 0 && g.vec_ok && n0 + NC <= g.N) {
#pragma unroll
    for (int i = 0; i < NC; i += 8) {
      float t[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) t[j] = acc[i + j];
      store8(crow + n0 + i, t);
    }
  } else {
#pragma unroll
    for (int i = 0; i < NC; ++i)
      if (n0 + i < g.N) crow[n0 + i] = from_f<TOut>(acc[i]);
  }
}

// In-register 32 x 32 transpose across a warp: before, lane l holds v[j] = element (row l, column j); after, lane l holds
// v[j] = element (row j, column l).  Five butterfly stages of 16 exchanges (static register indices only).
__device__ __forceinline__ void warp_transpose32(float (&v)[32], int lane) {
#pragma unroll
  for (int k = 16; k >= 1; k >>= 1) {
    const bool hi = (lane & k) != 0;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      if ((i & k) == 0) {
        const float send = hi ? v[i] : v[i + k];
        const float recv = __shfl_xor_sync(0xffffffffu, send, k);
        if (hi) v[i] = recv;
        else v[i + k] = recv;
      }
    }
  }
}

// fp32 output whose leading dimension allows no 16-byte stores (the [rows, 50265] log-probability matrix of
// AdaptiveSoftmax.get_log_prob): one thread per row writing 32 scalars put 32 partial sectors on the bus per store
// instruction (196 TFLOP/s for the M = 2048, N = 30265 tail GEMM: store bound).  The warp's 32 x 32 block is transposed in
// registers instead and written row by row: every store instruction covers 128 CONTIGUOUS bytes of one output row.
// Warp-collective: every lane calls it (rows >= M and columns >= N are skipped inside).  row0 = row of lane 0.
__device__ __forceinline__ void store_block_transposed(const GemmArgs& g, int grp, int row0, int M, int n0, float (&acc)[32], int lane) {
  warp_transpose32(acc, lane);
  float* c0 = reinterpret_cast<float*>(g.c) + grp * g.sc + static_cast<long>(row0) * g.ldc + n0 + lane;
  if (n0 + lane < g.N) {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (row0 + j < M) c0[static_cast<long>(j) * g.ldc] = acc[j];
  }
}

// zeros into columns [n0, n0+NC) of output row m (pad rows of a dynamic-M problem)
template <typename TOut, int NC>
__device__ __forceinline__ void zero_store(const GemmArgs& g, int grp, int m, int n0) {
  TOut* crow = reinterpret_cast<TOut*>(g.c) + grp * g.sc + static_cast<long>(m) * g.ldc;
#pragma unroll
  for (int i = 0; i < NC; ++i)
    if (n0 + i < g.N) crow[n0 + i] = from_f<TOut>(0.f);
}

// running (max, sumexp, argmax, target-logit) over a run of columns of one row
// packed f32x2 arithmetic (sm_100: FFMA2 / FADD2) and the bare MUFU.EX2
__device__ __forceinline__ uint64_t pack_f32x2(float a, float b) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ void unpack_f32x2(uint64_t v, float& a, float& b) { asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); }
__device__ __forceinline__ uint64_t fma_f32x2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t add_f32x2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ float ptx_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

struct LseAcc {
  float m, s, tz;
  int arg;
  __device__ __forceinline__ void init() { m = -INFINITY; s = 0.f; tz = 0.f; arg = -1; }
  __device__ __forceinline__ void add(float z, int col, int tgt) {
    if (z > m) {
      s = s * __expf(m - z) + 1.f;      // m = -inf -> exp(-inf) = 0
      m = z;
      arg = col;
    } else {
      s += __expf(z - m);
    }
    if (col == tgt) tz = z;
  }
  // NC consecutive columns starting at col0, all valid.  2.5 instructions per logit: the chunk maximum is a chain of 3-input
  // FMNMX, the exponent arguments and the running sum are packed f32x2 (FFMA2 / FADD2, sm_100), one MUFU.EX2 per logit.
  // The arg-max index is only looked for when the chunk raised the row maximum (a handful of chunks per row).  The first
  // form of this function (select-based max / argmax, scalar FFMA + exp2f + FADD: 13 instructions per logit) made the
  // adaptive-softmax GEMMs issue-bound: 25 M warp instructions, 51 us for 2048 x 30265 x 64 (ncu, profiles/).
  template <int NC>
  __device__ __forceinline__ void add_chunk(const float (&z)[NC], int col0, int tgt) {
    static_assert(NC % 2 == 0, "NC");
    float cm = z[0];
#pragma unroll
    for (int i = 1; i < NC; ++i) cm = fmaxf(cm, z[i]);
    if (cm > m) {
      s *= ptx_ex2((m - cm) * 1.4426950408889634f);     // m = -inf -> 0
      m = cm;
      int ca = 0;
#pragma unroll
      for (int i = NC - 1; i >= 0; --i) ca = z[i] == cm ? i : ca;   // lowest column wins a tie
      arg = col0 + ca;
    }
    const float mb = m * -1.4426950408889634f;
    const uint64_t l2 = pack_f32x2(1.4426950408889634f, 1.4426950408889634f), nb = pack_f32x2(mb, mb);
    uint64_t acc = pack_f32x2(0.f, 0.f);
#pragma unroll
    for (int i = 0; i < NC; i += 2) {
      float a, b;
      unpack_f32x2(fma_f32x2(pack_f32x2(z[i], z[i + 1]), l2, nb), a, b);
      acc = add_f32x2(acc, pack_f32x2(ptx_ex2(a), ptx_ex2(b)));
    }
    float a, b;
    unpack_f32x2(acc, a, b);
    s += a + b;
    const int ti = tgt - col0;
    if (ti >= 0 && ti < NC) {
#pragma unroll
      for (int i = 0; i < NC; ++i)
        if (i == ti) tz = z[i];
    }
  }
  __device__ __forceinline__ void merge(float m2, float s2, int arg2, float tz2) {
    if (m2 > m || (m2 == m && arg2 >= 0 && (arg < 0 || arg2 < arg))) {
      s = s * ((m == -INFINITY) ? 0.f : __expf(m - m2)) + s2;
      m = m2;
      arg = arg2;
    } else if (m2 != -INFINITY) {
      s += s2 * __expf(m2 - m);
    }
    tz += tz2;
  }
};

// host dispatch (gemm.cu)
int launch_gemm(const tnt_linear_params* p, int epi_override, const LseExtra* lse, int* n_tiles_out,
                cudaStream_t st);
int gemm_tile_n(int dtype, int M, int N);     // N-tile width the dispatcher will use (for partial sizing)

}  // namespace tnt
