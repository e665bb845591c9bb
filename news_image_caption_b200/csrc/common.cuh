// Shared host/device helpers for the tnt_b200 kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/tnt_b200.h"

namespace tnt {

// ------------------------------------------------------------------ host side: errors, launches
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
// which kernel VARIANT a dispatcher picked (tnt_variant_count): lets a test pin that the code path a benchmark times is
// the one the parity tests exercised (CTA-pair GEMM, split-K plan, tile-skipping attention, ...)
enum Variant {
  V_GEMM_TC_1CTA, V_GEMM_TC_PAIR, V_GEMM_SPLITK, V_GEMM_SIMT,
  V_XATTN_TC, V_XATTN_TC_TILESKIP, V_XATTN_TC_WEIGHTS, V_XATTN_SIMT, V_XATTN_STEP, V_XATTN_SHORT,
  V_XATTN_BWD_TC, V_XATTN_BWD_SIMT,
  V_DYNCONV_FWD_TC, V_DYNCONV_FWD_SIMT, V_DYNCONV_BWD_TC, V_DYNCONV_BWD_SIMT, V_DYNCONV_STEP_FAST, V_DYNCONV_STEP_SIMT,
  V_LN_GEMM_FUSED, V_KV_BESIDE, V_COUNT
};
void variant_hit(Variant v);
bool prof_enabled();
int prof_begin(const char* name, cudaStream_t st);   // -1 when profiling is off
void prof_end(int id, cudaStream_t st);

#define TNT_CHECK_ARG(cond, ...)             \
  do {                                       \
    if (!(cond)) {                           \
      ::tnt::set_error(__VA_ARGS__);         \
      return TNT_ERR_ARG;                    \
    }                                        \
  } while (0)

#define TNT_CHECK_ALIGN(cond, ...)           \
  do {                                       \
    if (!(cond)) {                           \
      ::tnt::set_error(__VA_ARGS__);         \
      return TNT_ERR_ALIGN;                  \
    }                                        \
  } while (0)

#define TNT_CUDA(call)                                                                        \
  do {                                                                                        \
    cudaError_t e__ = (call);                                                                 \
    if (e__ != cudaSuccess) {                                                                 \
      ::tnt::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
      return TNT_ERR_CUDA;                                                                    \
    }                                                                                         \
  } while (0)

// after a kernel launch
#define TNT_LAUNCHED()                 \
  do {                                 \
    ::tnt::count_launch();             \
    TNT_CUDA(cudaPeekAtLastError());   \
  } while (0)

#define TNT_TRY(call)        \
  do {                       \
    int s__ = (call);        \
    if (s__ != TNT_OK) return s__; \
  } while (0)

// Programmatic dependent launch (PDL): a kernel launched through launch_pdl() may start while its predecessor in
// the stream is still draining -- its prologue (barrier init, TMEM allocation, descriptor prefetch) overlaps the
// predecessor's tail -- and MUST execute pdl_wait() before its first global-memory access (and before any write
// a predecessor could still read).  Kernels that never call pdl_wait() must be launched the plain way.
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, int cluster_x,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[2];
  int n = 0;
  if (pdl_enabled()) {
    at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster_x > 1) {
    at[n].id = cudaLaunchAttributeClusterDimension;
    at[n].val.clusterDim.x = cluster_x;
    at[n].val.clusterDim.y = 1;
    at[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = at;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline size_t dtype_size(int dt) { return dt == TNT_BF16 ? 2 : 4; }
// per-DEVICE caches: one process may drive several GPUs (cudaFuncSetAttribute and the SM count are per device)
constexpr int kMaxDevices = 64;
inline int cur_device() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev & (kMaxDevices - 1);
}
struct PerDeviceOnce {          // `static PerDeviceOnce once; if (once.first()) { ...set attributes... }`
  bool done[kMaxDevices] = {};
  bool first() {
    const int d = cur_device();
    if (done[d]) return false;
    done[d] = true;
    return true;
  }
};
// SM budget of the launches the calling thread issues (0 = all): sub-batch chains that run side by side on several streams
// size their grids for a share of the GPU, so that their CTAs can be resident at the same time (engine.cuh, SubBatches)
inline int& sm_budget() {
  static thread_local int b = 0;
  return b;
}
inline int num_sms() {
  static int n[kMaxDevices] = {};
  const int dev = cur_device();
  if (!n[dev]) {
    cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
    if (n[dev] <= 0) n[dev] = 148;
  }
  const int b = sm_budget();
  return b > 0 && b < n[dev] ? b : n[dev];
}
struct SmBudget {                // RAII: SmBudget guard(n) -- launches issued while it lives see n SMs
  int keep;
  explicit SmBudget(int n) : keep(sm_budget()) { sm_budget() = n; }
  ~SmBudget() { sm_budget() = keep; }
};
inline cudaStream_t as_stream(tnt_stream s) { return reinterpret_cast<cudaStream_t>(s); }

// ------------------------------------------------------------------ counter-based dropout (tnt_dropout)
// keep(index) = drop_bits(k0, k1, index) >= thr; (k0, k1) fold the seed and the site.  thr == 0: site off.
struct DropSpec {
  uint32_t thr = 0;
  float scale = 1.f;
  uint32_t k0 = 0, k1 = 0;
  long ld = 0, gs = 0;
};
__host__ __device__ __forceinline__ uint32_t drop_bits(uint32_t k0, uint32_t k1, unsigned long long idx) {
  uint32_t h = k0 ^ static_cast<uint32_t>(idx);
  h *= 0x9E3779B1u;
  h ^= h >> 16;
  h += k1 + static_cast<uint32_t>(idx >> 32) * 0x85EBCA6Bu;
  h ^= h >> 15;
  h *= 0x2C1B3C6Du;
  h ^= h >> 12;
  h *= 0x297A2D39u;
  h ^= h >> 15;
  return h;
}
inline void drop_keys(uint64_t seed, uint32_t site, uint32_t* k0, uint32_t* k1) {
  *k0 = static_cast<uint32_t>(seed) * 0x9E3779B1u + site * 0x7FEB352Du + 0x165667B1u;
  *k1 = static_cast<uint32_t>(seed >> 32) ^ (site * 0x846CA68Bu + 0x27D4EB2Fu);
}
inline DropSpec make_drop(const tnt_dropout& d) {
  DropSpec s;
  if (!(d.p > 0.f)) return s;
  const double t = static_cast<double>(d.p) * 4294967296.0;
  s.thr = t >= 4294967295.0 ? 0xFFFFFFFFu : static_cast<uint32_t>(t);
  if (s.thr == 0) s.thr = 1;
  s.scale = d.p < 1.f ? 1.f / (1.f - d.p) : 0.f;
  drop_keys(d.seed, d.site, &s.k0, &s.k1);
  s.ld = d.ld;
  s.gs = d.group_stride;
  return s;
}

// ------------------------------------------------------------------ device side
using bf16 = __nv_bfloat16;

// value after dropout: scaled when kept, 0 when dropped
__device__ __forceinline__ float drop_val(const DropSpec& s, unsigned long long idx, float v) {
  return drop_bits(s.k0, s.k1, idx) >= s.thr ? v * s.scale : 0.f;
}

// PDL: block until the preceding kernel of the stream has completed and its writes are visible (no-op for a plain
// launch), then let the NEXT kernel's CTAs be scheduled as soon as SM resources free up.
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

template <typename T> struct Vec8;   // 8 elements = 128 bit (bf16) / 2 x 128 bit (fp32)

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16_rn(v); }

// load / store 8 consecutive elements (16-byte aligned for bf16, 32-byte for fp32) as floats
__device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p);
  const float4 b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void load8(const bf16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ void store8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void store8(bf16* p, const float (&v)[8]) {
  uint4 u;
  u.x = pack_bf16x2(v[0], v[1]);
  u.y = pack_bf16x2(v[2], v[3]);
  u.z = pack_bf16x2(v[4], v[5]);
  u.w = pack_bf16x2(v[6], v[7]);
  *reinterpret_cast<uint4*>(p) = u;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// block-wide sum; `red` = shared float[32]; every thread gets the result
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  r = warp_sum(r);
  return r;
}

// out[b] = bits | live << 16: bit i (i < 16) = key tile i (128 keys) of batch row b holds at least one key that is not
// padding; live = number of such keys (saturated at 65535).  One warp per row.
__device__ __forceinline__ void tile_bits_row(const uint8_t* mask, long m_stride_b, int b, int S, int mask_S, uint32_t* out, int lane) {
  uint32_t bits = 0, live = 0;
  const uint8_t* mrow = mask + static_cast<long>(b) * m_stride_b;
  for (int k0 = 0; k0 < S; k0 += 256) {            // 8 independent byte loads per lane in flight
    bool lv[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const int key = k0 + u * 32 + lane;
      lv[u] = key < S && (key >= mask_S || mrow[key] == 0);
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const uint32_t bal = __ballot_sync(0xffffffffu, lv[u]);
      if (bal) bits |= 1u << ((k0 + u * 32) / 128);
      live += __popc(bal);
    }
  }
  if (lane == 0) out[b] = (bits & 0xFFFFu) | (min(live, 0xFFFFu) << 16);
}

// one row-wise copy of the weight refresh (tnt_decoder_finalize): out[r] = (g ? g[r] / ||v[r]|| : 1) * v[r], r < rows;
// optional GLU pair interleave of the output rows; fp32 output for vectors, the engine dtype for matrices
struct FoldJob {
  const float* v; const float* g; void* out;
  int rows, cols, glu, out_f32;
  int row0;                       // first warp-row of this job in the batched launch
  bool operator==(const FoldJob& o) const {
    return v == o.v && g == o.g && out == o.out && rows == o.rows && cols == o.cols && glu == o.glu && out_f32 == o.out_f32;
  }
};
// all jobs in one launch (warp per row); `jobs` is a device array ordered by row0
int wn_fold_batch(const FoldJob* jobs, int n_jobs, int total_rows, int dtype, cudaStream_t st);

}  // namespace tnt
