// Backward-pass helper kernels (HBM-bound glue of kernel group 4): GLU gradient, bias (column) sums,
// weight-norm gradient, row permutations / gathers / scatters, embedding-table scatter, bias_k/bias_v rows.
// Reference ops they differentiate: decoder_flattened.py:281-345 (layer), linear.py:7-33 (weight_norm),
// adaptive.py:61-76 (embedding bands), softmax.py:144-191 (adaptive softmax).
#include "backward.cuh"

namespace tnt {
namespace {

// ------------------------------------------------------------------------------ GLU backward
template <typename T>
__global__ void __launch_bounds__(256) glu_bwd_kernel(const T* raw, long ldraw, const T* dh, long lddh, T* draw,
                                                      long lddraw, int N, int C) {
  const int nv = C >> 3;                       // 8 outputs <-> 16 interleaved pre-activations
  const long total = static_cast<long>(N) * nv;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
    const int r = static_cast<int>(i / nv), vc = static_cast<int>(i - static_cast<long>(r) * nv);
    float g[8], p0[8], p1[8], o0[8], o1[8];
    load8(dh + r * lddh + vc * 8, g);
    load8(raw + r * ldraw + vc * 16, p0);
    load8(raw + r * ldraw + vc * 16 + 8, p1);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float a = j < 4 ? p0[2 * j] : p1[2 * (j - 4)];
      const float b = j < 4 ? p0[2 * j + 1] : p1[2 * (j - 4) + 1];
      const float s = sigmoidf_(b);
      const float da = g[j] * s, db = g[j] * a * s * (1.f - s);
      if (j < 4) { o0[2 * j] = da; o0[2 * j + 1] = db; }
      else { o1[2 * (j - 4)] = da; o1[2 * (j - 4) + 1] = db; }
    }
    store8(draw + r * lddraw + vc * 16, o0);
    store8(draw + r * lddraw + vc * 16 + 8, o1);
  }
}

// ------------------------------------------------------------------------------ column sums
// block = 32 x 8: threadIdx.x owns 8 consecutive columns, threadIdx.y strides the rows of this block's chunk
// rows per CTA are chosen by the host so that the grid covers the GPU about twice (a [2048, 1024] bias gradient ran on 32
// CTAs with 32 dependent row loads per thread: 19 us; with 32-row chunks 256 CTAs: a few us)
template <typename T>
__global__ void __launch_bounds__(256) colsum_kernel(const T* x, long ldx, int M, int N, const int* m_dyn, float* out,
                                                     int deinterleave, int vec, int kCsRows) {
  __shared__ float red[8][32][8];
  const int Me = m_dyn ? min(M, *m_dyn) : M;
  const int c0 = (blockIdx.x * 32 + threadIdx.x) * 8;
  const int r0 = blockIdx.y * kCsRows;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (c0 < N) {
    const int rend = min(Me, r0 + kCsRows);
#pragma unroll 4
    for (int r = r0 + threadIdx.y; r < rend; r += 8) {
      if (vec && c0 + 8 <= N) {
        float v[8];
        load8(x + static_cast<long>(r) * ldx + c0, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += v[j];
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (c0 + j < N) acc[j] += to_f(x[static_cast<long>(r) * ldx + c0 + j]);
      }
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.y][threadIdx.x][j] = acc[j];
  __syncthreads();
  if (threadIdx.y == 0 && c0 < N) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float s = 0.f;
#pragma unroll
      for (int y = 0; y < 8; ++y) s += red[y][threadIdx.x][j];
      const int c = c0 + j;
      if (c < N) {
        const int o = deinterleave ? ((c & 1) ? (N >> 1) + (c >> 1) : (c >> 1)) : c;
        atomicAdd(out + o, s);
      }
    }
  }
}

// ------------------------------------------------------------------------------ weight-norm backward
// one warp per parameter row; rows are walked in 1024-column chunks with all 8 float4 loads per lane (of v and of dW) in
// flight at once -- the rolled scalar loop took 19 us per matrix, one memory latency per iteration
__device__ __forceinline__ void ld8(const float* row, int c0, int cols, int lane, float4 (&a)[8]) {
#pragma unroll
  for (int u = 0; u < 8; ++u) {
    const int c = c0 + (u * 32 + lane) * 4;
    a[u] = c < cols ? __ldg(reinterpret_cast<const float4*>(row + c)) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
}

__global__ void __launch_bounds__(256) wn_bwd_kernel(const float* v, const float* g, const float* dW, long lddw,
                                                     float* dv, float* dg, int rows, int cols, int glu) {
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int rw = glu ? (r < (rows >> 1) ? 2 * r : 2 * (r - (rows >> 1)) + 1) : r;
  const float* vr = v + static_cast<long>(r) * cols;
  const float* dr = dW + static_cast<long>(rw) * lddw;
  float* o = dv + static_cast<long>(r) * cols;
  const bool vec = (cols & 3) == 0 && (lddw & 3) == 0 && ((reinterpret_cast<uintptr_t>(v) | reinterpret_cast<uintptr_t>(dW) |
                                                          reinterpret_cast<uintptr_t>(dv)) & 15u) == 0;
  if (vec) {
    float4 a[8], d[8];
    if (!g) {
      for (int c0 = 0; c0 < cols; c0 += 1024) {
        ld8(dr, c0, cols, lane, d);
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int c = c0 + (u * 32 + lane) * 4;
          if (c < cols) *reinterpret_cast<float4*>(o + c) = d[u];
        }
      }
      return;
    }
    float nn = 0.f, dot = 0.f;
    for (int c0 = cols > 1024 ? ((cols - 1) / 1024) * 1024 : 0; c0 >= 0; c0 -= 1024) {     // chunk 0 last: it stays in registers
      ld8(vr, c0, cols, lane, a);
      ld8(dr, c0, cols, lane, d);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        nn = fmaf(a[u].x, a[u].x, fmaf(a[u].y, a[u].y, fmaf(a[u].z, a[u].z, fmaf(a[u].w, a[u].w, nn))));
        dot = fmaf(a[u].x, d[u].x, fmaf(a[u].y, d[u].y, fmaf(a[u].z, d[u].z, fmaf(a[u].w, d[u].w, dot))));
      }
    }
    nn = warp_sum(nn);
    dot = warp_sum(dot);
    const float n = sqrtf(nn), gg = g[r];
    const float k1 = gg / n, k2 = dot / nn;
    for (int c0 = 0; c0 < cols; c0 += 1024) {
      if (c0 > 0) {
        ld8(vr, c0, cols, lane, a);
        ld8(dr, c0, cols, lane, d);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int c = c0 + (u * 32 + lane) * 4;
        if (c < cols)
          *reinterpret_cast<float4*>(o + c) = make_float4(k1 * (d[u].x - k2 * a[u].x), k1 * (d[u].y - k2 * a[u].y),
                                                          k1 * (d[u].z - k2 * a[u].z), k1 * (d[u].w - k2 * a[u].w));
      }
    }
    if (lane == 0) dg[r] = dot / n;
    return;
  }
  if (!g) {
    for (int c = lane; c < cols; c += 32) o[c] = dr[c];
    return;
  }
  float nn = 0.f, dot = 0.f;
  for (int c = lane; c < cols; c += 32) {
    const float a = vr[c];
    nn = fmaf(a, a, nn);
    dot = fmaf(a, dr[c], dot);
  }
  nn = warp_sum(nn);
  dot = warp_sum(dot);
  const float n = sqrtf(nn), gg = g[r];
  const float k1 = gg / n, k2 = dot / nn;
  for (int c = lane; c < cols; c += 32) o[c] = k1 * (dr[c] - k2 * vr[c]);
  if (lane == 0) dg[r] = dot / n;
}

// ------------------------------------------------------------------------------ row movers (warp per row)
__device__ __forceinline__ int remap_row(int r, int mB, int mT) {   // b*T + t -> t*B + b
  return mB > 0 ? (r % mT) * mB + r / mT : r;
}

template <typename T>
__global__ void __launch_bounds__(256) permute_rows_kernel(const T* src, T* dst, int B, int Tn, int C, int to_tb) {
  const int lane = threadIdx.x & 31;
  const int rows = B * Tn;
  for (int r = blockIdx.x * 8 + (threadIdx.x >> 5); r < rows; r += gridDim.x * 8) {
    // r indexes the b*T + t order
    const int tb = (r % Tn) * B + r / Tn;
    const T* s = src + static_cast<long>(to_tb ? r : tb) * C;
    T* d = dst + static_cast<long>(to_tb ? tb : r) * C;
    for (int vc = lane; vc < (C >> 3); vc += 32) {
      float v[8];
      load8(s + vc * 8, v);
      store8(d + vc * 8, v);
    }
  }
}

// mode 0: out[j] = src[map(rows[j])];  mode 1: dst[map(rows[j])] += src[j]
template <typename T, int MODE>
__global__ void __launch_bounds__(256) move_rows_kernel(const T* src, long lds, const int* rows, const int* count,
                                                        T* dst, long ldd, int C, int mB, int mT) {
  const int lane = threadIdx.x & 31;
  const int n = *count;
  for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
    const int rr = remap_row(rows[j], mB, mT);
    const T* s = src + static_cast<long>(MODE == 0 ? rr : j) * lds;
    T* d = dst + static_cast<long>(MODE == 0 ? j : rr) * ldd;
    for (int vc = lane; vc < (C >> 3); vc += 32) {
      float v[8];
      load8(s + vc * 8, v);
      if (MODE == 1) {
        float o[8];
        load8(d + vc * 8, o);
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] += o[i];
      }
      store8(d + vc * 8, v);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) gather_table_kernel(const float* table, long ldt, const int* loc, const int* count,
                                                           T* out, long ldo, int C) {
  const int lane = threadIdx.x & 31;
  const int n = *count;
  for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
    const float* s = table + static_cast<long>(loc[j]) * ldt;
    T* d = out + static_cast<long>(j) * ldo;
    for (int vc = lane; vc < (C >> 3); vc += 32) {
      float v[8];
      load8(s + vc * 8, v);
      store8(d + vc * 8, v);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) scatter_table_kernel(const T* src, long lds, const int* loc, const int* count,
                                                            float* grad, long ldg, int C, int skip_row) {
  const int lane = threadIdx.x & 31;
  const int n = *count;
  for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
    const int l = loc[j];
    if (l == skip_row) continue;                          // nn.Embedding(padding_idx): that row gets no gradient
    const T* s = src + static_cast<long>(j) * lds;
    float* d = grad + static_cast<long>(l) * ldg;
    for (int vc = lane; vc < (C >> 3); vc += 32) {
      float v[8];
      load8(s + vc * 8, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) atomicAdd(d + vc * 8 + i, v[i]);
    }
  }
}

// Sparse form of the embedding-table gradient (data-parallel training): instead of scattering the rows into the dense band
// table, band `band`'s rows are appended to a packed list -- row j lands at offset sum_{k < band} count[k] + j with
// meta = band << 32 | band-local row -- that the ranks all-gather and every rank scatters (embed_grad_scatter).
template <typename T>
__global__ void __launch_bounds__(256) pack_band_rows_kernel(const T* src, long lds, const int* loc, const int* counts, int band,
                                                             T* rows_out, long ldo, long long* meta, int* total, int C, int n_bands) {
  const int lane = threadIdx.x & 31;
  int off = 0;
  for (int k = 0; k < band; ++k) off += counts[k];
  const int n = counts[band];
  if (blockIdx.x == 0 && threadIdx.x == 0 && band == n_bands - 1) *total = off + n;
  for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
    const T* s = src + static_cast<long>(j) * lds;
    T* o = rows_out + static_cast<long>(off + j) * ldo;
    for (int vc = lane; vc < (C >> 3); vc += 32) *reinterpret_cast<uint4*>(o + vc * 8) = *reinterpret_cast<const uint4*>(s + vc * 8);
    if (lane == 0) meta[off + j] = (static_cast<long long>(band) << 32) | static_cast<unsigned>(loc[j]);
  }
}
template <>
__global__ void __launch_bounds__(256) pack_band_rows_kernel<float>(const float* src, long lds, const int* loc, const int* counts,
                                                                    int band, float* rows_out, long ldo, long long* meta,
                                                                    int* total, int C, int n_bands) {
  const int lane = threadIdx.x & 31;
  int off = 0;
  for (int k = 0; k < band; ++k) off += counts[k];
  const int n = counts[band];
  if (blockIdx.x == 0 && threadIdx.x == 0 && band == n_bands - 1) *total = off + n;
  for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
    const float* s = src + static_cast<long>(j) * lds;
    float* o = rows_out + static_cast<long>(off + j) * ldo;
    for (int vc = lane; vc < (C >> 2); vc += 32) *reinterpret_cast<float4*>(o + vc * 4) = *reinterpret_cast<const float4*>(s + vc * 4);
    if (lane == 0) meta[off + j] = (static_cast<long long>(band) << 32) | static_cast<unsigned>(loc[j]);
  }
}

struct BandTables { float* grad[TNT_MAX_CLUSTERS]; int dim[TNT_MAX_CLUSTERS]; };
template <typename T>
__global__ void __launch_bounds__(256) embed_grad_scatter_kernel(const T* rows, long ldr, const long long* meta, const int* count,
                                                                 int max_rows, BandTables bt, int n_bands, int skip_row) {
  const int lane = threadIdx.x & 31;
  const int n = min(*count, max_rows);
  for (int j = blockIdx.x * 8 + (threadIdx.x >> 5); j < n; j += gridDim.x * 8) {
    const long long m = meta[j];
    const int band = static_cast<int>(m >> 32), l = static_cast<int>(m & 0xffffffffll);
    if (band < 0 || band >= n_bands || l == skip_row) continue;
    const T* s = rows + static_cast<long>(j) * ldr;
    float* d = bt.grad[band] + static_cast<long>(l) * bt.dim[band];
    for (int vc = lane; vc < (bt.dim[band] >> 3); vc += 32) {
      float v[8];
      load8(s + vc * 8, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) atomicAdd(d + vc * 8 + i, v[i]);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) kv_extra_grad_kernel(const T* dkv, long ld, int S, int B, int ncols, float* out) {
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (c >= ncols) return;
  float s = 0.f;
  for (int b = 0; b < B; ++b) s += to_f(dkv[(static_cast<long>(S) * B + b) * ld + c]);
  out[c] = s;
}

// dst[r, :] += sum_{s in [s0, ns)} src[r, s*C : (s+1)*C]
template <typename T>
__global__ void __launch_bounds__(256) add_col_blocks_kernel(T* dst, long ldd, const T* src, long lds, int N, int C,
                                                             int s0, int ns) {
  const int nv = C >> 3;
  const long total = static_cast<long>(N) * nv;
  for (long i = blockIdx.x * 256L + threadIdx.x; i < total; i += gridDim.x * 256L) {
    const int r = static_cast<int>(i / nv), vc = static_cast<int>(i - static_cast<long>(r) * nv);
    float a[8];
    load8(dst + r * ldd + vc * 8, a);
    for (int s = s0; s < ns; ++s) {
      float v[8];
      load8(src + r * lds + static_cast<long>(s) * C + vc * 8, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) a[j] += v[j];
    }
    store8(dst + r * ldd + vc * 8, a);
  }
}

__global__ void __launch_bounds__(256) smgrad_rows_head_kernel(const float4* hs, const float* side, const int* head_t,
                                                               int N, int nt, int cut0, int pad, float* row_lse,
                                                               int* row_tgt) {
  const int r = blockIdx.x * 256 + threadIdx.x;
  if (r >= N) return;
  const int ht = head_t[r];
  if (ht == pad) {                                        // ignore_index inside the head cluster
    row_lse[r] = INFINITY;
    row_tgt[r] = -1;
    return;
  }
  const float4 st = hs[r];
  float M = st.x;
  for (int i = 0; i < nt; ++i) M = fmaxf(M, side[static_cast<long>(r) * nt + i]);
  float s = st.y * __expf(st.x - M);
  for (int i = 0; i < nt; ++i) s += __expf(side[static_cast<long>(r) * nt + i] - M);
  row_lse[r] = M + __logf(s);
  row_tgt[r] = ht;
}

__global__ void __launch_bounds__(256) smgrad_rows_tail_kernel(const float4* ts, const int* tail_loc, const int* count,
                                                               int N, int pad, float* row_lse, int* row_tgt) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j >= N) return;
  if (j >= *count || tail_loc[j] == pad) {                // beyond the tail's rows, or ignore_index (local id 1)
    row_lse[j] = INFINITY;
    row_tgt[j] = -1;
    return;
  }
  const float4 st = ts[j];
  row_lse[j] = st.x + __logf(st.y);
  row_tgt[j] = tail_loc[j];
}

inline int grid_for(long work, int per_block) {
  return static_cast<int>(std::min<long>((work + per_block - 1) / per_block, 8L * num_sms()));
}
}  // namespace

int glu_bwd(const void* raw, long ldraw, const void* dh, long lddh, void* draw, long lddraw, int N, int C, int dtype,
            cudaStream_t st) {
  TNT_CHECK_ARG(raw && dh && draw && C % 8 == 0, "glu_bwd: bad arguments");
  TNT_CHECK_ALIGN(aligned16(raw) && aligned16(dh) && aligned16(draw) && ldraw % 8 == 0 && lddh % 8 == 0 && lddraw % 8 == 0,
                  "glu_bwd: alignment");
  if (N == 0) return TNT_OK;
  const int grid = grid_for(static_cast<long>(N) * (C >> 3), 256);
  if (dtype == TNT_BF16)
    glu_bwd_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<const bf16*>(raw), ldraw, static_cast<const bf16*>(dh), lddh,
                                               static_cast<bf16*>(draw), lddraw, N, C);
  else
    glu_bwd_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(raw), ldraw, static_cast<const float*>(dh),
                                                lddh, static_cast<float*>(draw), lddraw, N, C);
  TNT_LAUNCHED();
  return TNT_OK;
}

int colsum(const void* x, long ldx, int M, int N, const int* m_dyn, float* out, int dtype, bool deinterleave,
           bool accumulate, cudaStream_t st) {
  TNT_CHECK_ARG(x && out && N > 0, "colsum: bad arguments");
  if (!accumulate) TNT_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * N, st));
  if (M == 0) return TNT_OK;
  const int vec = aligned16(x) && ldx % 8 == 0;
  const int gx = (N + 255) / 256;
  int rows_per = 256;                                    // multiple of 8; shrink until ~2 CTAs per SM (>= 32 rows each:
                                                         // 4 independent loads per thread, <= M/32 atomics per column)
  while (rows_per > 32 && static_cast<long>(gx) * ((M + rows_per - 1) / rows_per) < 2L * num_sms()) rows_per >>= 1;
  dim3 grid(gx, (M + rows_per - 1) / rows_per), block(32, 8);
  if (dtype == TNT_BF16)
    colsum_kernel<bf16><<<grid, block, 0, st>>>(static_cast<const bf16*>(x), ldx, M, N, m_dyn, out, deinterleave, vec, rows_per);
  else
    colsum_kernel<float><<<grid, block, 0, st>>>(static_cast<const float*>(x), ldx, M, N, m_dyn, out, deinterleave, vec, rows_per);
  TNT_LAUNCHED();
  return TNT_OK;
}

int wn_bwd(const float* v, const float* g, const float* dW, long lddw, float* dv, float* dg, int rows, int cols,
           bool glu_interleave, cudaStream_t st) {
  TNT_CHECK_ARG(v && dW && dv && (!g || dg), "wn_bwd: bad arguments");
  wn_bwd_kernel<<<(rows + 7) / 8, 256, 0, st>>>(v, g, dW, lddw, dv, dg, rows, cols, glu_interleave ? 1 : 0);
  TNT_LAUNCHED();
  return TNT_OK;
}


namespace {
// dst[i] = sum_g parts[g * n + i]: the reduction over K-split partial weight gradients (n % 4 == 0)
__global__ void __launch_bounds__(256) sum_partials_kernel(float* dst, const float* parts, int G, long n4) {
  for (long i = blockIdx.x * 256L + threadIdx.x; i < n4; i += gridDim.x * 256L) {
    float4 a = __ldg(reinterpret_cast<const float4*>(parts) + i);
    for (int g = 1; g < G; ++g) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(parts) + g * n4 + i);
      a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    }
    reinterpret_cast<float4*>(dst)[i] = a;
  }
}
}  // namespace

int sum_partials(float* dst, const float* parts, int G, long n, cudaStream_t st) {
  TNT_CHECK_ARG(dst && parts && G >= 1 && n % 4 == 0, "sum_partials: bad arguments");
  const long n4 = n / 4;
  sum_partials_kernel<<<static_cast<int>(std::min<long>((n4 + 255) / 256, 8L * num_sms())), 256, 0, st>>>(dst, parts, G, n4);
  TNT_LAUNCHED();
  return TNT_OK;
}

int permute_rows(const void* src, void* dst, int B, int T, int C, int dtype, bool to_tb, cudaStream_t st) {
  TNT_CHECK_ARG(src && dst && C % 8 == 0, "permute_rows: bad arguments");
  if (B * T == 0) return TNT_OK;
  const int grid = grid_for(static_cast<long>(B) * T, 8);
  if (dtype == TNT_BF16)
    permute_rows_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<const bf16*>(src), static_cast<bf16*>(dst), B, T, C, to_tb);
  else
    permute_rows_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(src), static_cast<float*>(dst), B, T, C, to_tb);
  TNT_LAUNCHED();
  return TNT_OK;
}

int scatter_add_rows(const void* src, long lds, const int* rows, const int* count, int max_rows, void* dst, long ldd,
                     int C, int dtype, int map_B, int map_T, cudaStream_t st) {
  if (max_rows == 0) return TNT_OK;
  const int grid = grid_for(max_rows, 8);
  if (dtype == TNT_BF16)
    move_rows_kernel<bf16, 1><<<grid, 256, 0, st>>>(static_cast<const bf16*>(src), lds, rows, count, static_cast<bf16*>(dst), ldd, C, map_B, map_T);
  else
    move_rows_kernel<float, 1><<<grid, 256, 0, st>>>(static_cast<const float*>(src), lds, rows, count, static_cast<float*>(dst), ldd, C, map_B, map_T);
  TNT_LAUNCHED();
  return TNT_OK;
}

int gather_rows(const void* src, long lds, const int* rows, const int* count, int max_rows, void* out, long ldo, int C,
                int dtype, int map_B, int map_T, cudaStream_t st) {
  if (max_rows == 0) return TNT_OK;
  const int grid = grid_for(max_rows, 8);
  if (dtype == TNT_BF16)
    move_rows_kernel<bf16, 0><<<grid, 256, 0, st>>>(static_cast<const bf16*>(src), lds, rows, count, static_cast<bf16*>(out), ldo, C, map_B, map_T);
  else
    move_rows_kernel<float, 0><<<grid, 256, 0, st>>>(static_cast<const float*>(src), lds, rows, count, static_cast<float*>(out), ldo, C, map_B, map_T);
  TNT_LAUNCHED();
  return TNT_OK;
}

int gather_table_rows(const float* table, long ldt, const int* loc, const int* count, int max_rows, void* out, long ldo,
                      int C, int dtype, cudaStream_t st) {
  if (max_rows == 0) return TNT_OK;
  const int grid = grid_for(max_rows, 8);
  if (dtype == TNT_BF16)
    gather_table_kernel<bf16><<<grid, 256, 0, st>>>(table, ldt, loc, count, static_cast<bf16*>(out), ldo, C);
  else
    gather_table_kernel<float><<<grid, 256, 0, st>>>(table, ldt, loc, count, static_cast<float*>(out), ldo, C);
  TNT_LAUNCHED();
  return TNT_OK;
}

int scatter_table_grad(const void* src, long lds, const int* loc, const int* count, int max_rows, float* grad, long ldg,
                       int C, int dtype, int skip_row, cudaStream_t st) {
  if (max_rows == 0) return TNT_OK;
  const int grid = grid_for(max_rows, 8);
  if (dtype == TNT_BF16)
    scatter_table_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<const bf16*>(src), lds, loc, count, grad, ldg, C, skip_row);
  else
    scatter_table_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(src), lds, loc, count, grad, ldg, C, skip_row);
  TNT_LAUNCHED();
  return TNT_OK;
}

int pack_band_rows(const void* src, long lds, const int* loc, const int* counts, int band, int n_bands, int max_rows, void* rows_out,
                   long ldo, long long* meta, int* total, int C, int dtype, cudaStream_t st) {
  if (max_rows == 0) return TNT_OK;
  const int grid = grid_for(max_rows, 8);
  if (dtype == TNT_BF16)
    pack_band_rows_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<const bf16*>(src), lds, loc, counts, band, static_cast<bf16*>(rows_out),
                                                      ldo, meta, total, C, n_bands);
  else
    pack_band_rows_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(src), lds, loc, counts, band,
                                                       static_cast<float*>(rows_out), ldo, meta, total, C, n_bands);
  TNT_LAUNCHED();
  return TNT_OK;
}

namespace {
__global__ void __launch_bounds__(256) gather_floats_kernel(const float* src, const int* rows, const int* count, int N, float* out) {
  const int j = blockIdx.x * 256 + threadIdx.x;
  if (j < N) out[j] = j < *count ? src[rows[j]] : 0.f;
}
}  // namespace
int gather_floats(const float* src, const int* rows, const int* count, int N, float* out, cudaStream_t st) {
  if (N == 0) return TNT_OK;
  gather_floats_kernel<<<(N + 255) / 256, 256, 0, st>>>(src, rows, count, N, out);
  TNT_LAUNCHED();
  return TNT_OK;
}

int kv_extra_grad(const void* dkv, long ld, int S, int B, int ncols, float* out, int dtype, cudaStream_t st) {
  const int grid = (ncols + 255) / 256;
  if (dtype == TNT_BF16)
    kv_extra_grad_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<const bf16*>(dkv), ld, S, B, ncols, out);
  else
    kv_extra_grad_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(dkv), ld, S, B, ncols, out);
  TNT_LAUNCHED();
  return TNT_OK;
}

int smgrad_rows_head(const float4* head_stats, const float* side, const int* head_t, int N, int n_tails, int cut0,
                     int pad, float* row_lse, int* row_tgt, cudaStream_t st) {
  if (N == 0) return TNT_OK;
  smgrad_rows_head_kernel<<<(N + 255) / 256, 256, 0, st>>>(head_stats, side, head_t, N, n_tails, cut0, pad, row_lse, row_tgt);
  TNT_LAUNCHED();
  return TNT_OK;
}

int smgrad_rows_tail(const float4* tail_stats, const int* tail_loc, const int* count, int N, int pad, float* row_lse,
                     int* row_tgt, cudaStream_t st) {
  if (N == 0) return TNT_OK;
  smgrad_rows_tail_kernel<<<(N + 255) / 256, 256, 0, st>>>(tail_stats, tail_loc, count, N, pad, row_lse, row_tgt);
  TNT_LAUNCHED();
  return TNT_OK;
}

int add_col_blocks(void* dst, long ldd, const void* src, long lds, int N, int C, int s0, int ns, int dtype,
                   cudaStream_t st) {
  if (N == 0 || s0 >= ns) return TNT_OK;
  const int grid = grid_for(static_cast<long>(N) * (C >> 3), 256);
  if (dtype == TNT_BF16)
    add_col_blocks_kernel<bf16><<<grid, 256, 0, st>>>(static_cast<bf16*>(dst), ldd, static_cast<const bf16*>(src), lds, N, C, s0, ns);
  else
    add_col_blocks_kernel<float><<<grid, 256, 0, st>>>(static_cast<float*>(dst), ldd, static_cast<const float*>(src), lds, N, C, s0, ns);
  TNT_LAUNCHED();
  return TNT_OK;
}

}  // namespace tnt

// Applies one rank's packed embedding-gradient rows (tnt_decoder_set_sparse_embed_grad) to the dense fp32 band-table
// gradients: grad[band][loc] += row for every listed row; the band-local padding row is skipped (adaptive.py:42).
extern "C" int tnt_embed_grad_scatter(const void* rows, int64_t ld_rows, const int64_t* meta, const int* count, int max_rows,
                                      float* const* band_grads, const int* band_dims, int n_bands, int skip_row, int dtype,
                                      tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(rows && meta && count && band_grads && band_dims && n_bands >= 1 && n_bands <= TNT_MAX_CLUSTERS,
                "embed_grad_scatter: bad arguments");
  if (max_rows <= 0) return TNT_OK;
  BandTables bt = {};
  for (int i = 0; i < n_bands; ++i) {
    TNT_CHECK_ARG(band_grads[i] && band_dims[i] % 8 == 0, "embed_grad_scatter: band %d", i);
    bt.grad[i] = band_grads[i];
    bt.dim[i] = band_dims[i];
  }
  const int grid = grid_for(max_rows, 8);
  if (dtype == TNT_BF16)
    embed_grad_scatter_kernel<bf16><<<grid, 256, 0, as_stream(stream)>>>(static_cast<const bf16*>(rows), ld_rows,
                                                                         reinterpret_cast<const long long*>(meta), count, max_rows, bt,
                                                                         n_bands, skip_row);
  else
    embed_grad_scatter_kernel<float><<<grid, 256, 0, as_stream(stream)>>>(static_cast<const float*>(rows), ld_rows,
                                                                          reinterpret_cast<const long long*>(meta), count, max_rows, bt,
                                                                          n_bands, skip_row);
  TNT_LAUNCHED();
  return TNT_OK;
}

// nn.Embedding backward for the final* decoder (Embeddings: lut(x) * sqrt(d_model), transformer_v2.py Embeddings):
// grad[ids[r], :] += scale * dy[r, :] (fp32 atomics); ids outside [0, vocab) are skipped.
namespace tnt {
namespace {
template <typename T>
__global__ void __launch_bounds__(256) embedding_bwd_kernel(const int64_t* ids, int n, const T* dy, long lddy, float scale, float* grad,
                                                            long ldg, int C, int vocab) {
  const int lane = threadIdx.x & 31;
  for (int r = blockIdx.x * 8 + (threadIdx.x >> 5); r < n; r += gridDim.x * 8) {
    const long id = ids[r];
    if (id < 0 || id >= vocab) continue;
    const T* s = dy + static_cast<long>(r) * lddy;
    float* d = grad + id * ldg;
    for (int vc = lane; vc < (C >> 3); vc += 32) {
      float v[8];
      load8(s + vc * 8, v);
#pragma unroll
      for (int i = 0; i < 8; ++i) atomicAdd(d + vc * 8 + i, v[i] * scale);
    }
  }
}
}  // namespace
}  // namespace tnt

extern "C" int tnt_embedding_bwd(const int64_t* ids, int n, const void* dy, int64_t lddy, float scale, float* grad, int64_t ldg,
                                 int C, int vocab, int dtype, tnt_stream stream) {
  using namespace tnt;
  TNT_CHECK_ARG(ids && dy && grad && n >= 0 && C % 8 == 0 && vocab > 0, "embedding_bwd: bad arguments");
  TNT_CHECK_ALIGN(aligned16(dy) && lddy % 8 == 0, "embedding_bwd: dy alignment");
  if (n == 0) return TNT_OK;
  const int grid = grid_for(n, 8);
  if (dtype == TNT_BF16)
    embedding_bwd_kernel<bf16><<<grid, 256, 0, as_stream(stream)>>>(ids, n, static_cast<const bf16*>(dy), lddy, scale, grad, ldg, C, vocab);
  else
    embedding_bwd_kernel<float><<<grid, 256, 0, as_stream(stream)>>>(ids, n, static_cast<const float*>(dy), lddy, scale, grad, ldg, C, vocab);
  TNT_LAUNCHED();
  return TNT_OK;
}
