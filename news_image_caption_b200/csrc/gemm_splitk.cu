// Split-K form of the tcgen05 GEMM for the skinny, long-K problems of the decoding step (M = batch <= a few hundred rows,
// K >= 2048: context_fc, fc2).  Measured (tools/diag_skinny_gemm.py): with A read from shared memory every
// tcgen05.mma of a 128-row tile costs ~160 cycles however narrow N is, so a CTA's k-loop takes 0.33 us per 64-wide
// k-block regardless of tile width or ring depth -- the only way to shorten it is to run parts of K concurrently.
//   cluster of S CTAs = one 128 x BN output tile; CTA r reduces k-blocks [r*per, (r+1)*per) into its own TMEM;
//   CTAs 1.. push their fp32 partial tile into the LEADER's shared memory (st.shared::cluster), one cluster barrier,
//   the leader adds them to its accumulator and runs the normal epilogue (bias / ReLU / GLU / residual).
// One tile per CTA (grid = tiles * S <= SM count), NT layout, no groups.
#include <cuda.h>

#include <algorithm>

#include "gemm.cuh"
#include "ptx.cuh"

namespace tnt {
int make_tmap_bf16(CUtensorMap* tm, const void* base, long cols, long rows, long ld, long groups, long group_stride,
                   int box_rows);

namespace sk {
constexpr int BM = 128, BK = 64, UMMA_K = 16, kThreads = 192, kMaxStages = 8;
constexpr int A_STAGE_BYTES = BM * BK * 2;
constexpr int kSmemBudget = 220 * 1024;

__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void tmem_alloc_n(uint32_t* smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32(smem_dst)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_n(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

// smem: [partials (S-1) x (BN/32) x 128 x 32 fp32][ring stages x (A 16 KB + W BN x 128 B)][barriers][bias]
template <int EPI, typename TOut>
__global__ void __launch_bounds__(kThreads, 1)
gemm_splitk_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, GemmArgs g, int BN,
                   int stages, int S) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int part_bytes = (S - 1) * BM * BN * 4;                       // multiple of 1024
  float* part = reinterpret_cast<float*>(smem);
  uint8_t* ring = smem + part_bytes;
  const int stage_bytes = A_STAGE_BYTES + BN * BK * 2;
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + stages * stage_bytes);
  uint64_t* empty = full + kMaxStages;
  uint64_t* tfull = empty + kMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);
  float* sb = reinterpret_cast<float*>(ring + stages * stage_bytes + 512);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = static_cast<int>(ptx::cluster_ctarank());
  const int tile = blockIdx.x / S;
  const uint32_t tcols = BN < 32 ? 32u : static_cast<uint32_t>(BN);

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmW);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    ptx::mbar_init(tfull, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_n(tmem_slot, tcols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  pdl_wait();
  const int M = g.M;
  const int m_tiles = (M + BM - 1) / BM;
  const int mt = tile % m_tiles, nt = tile / m_tiles;
  const int num_kb = (g.K + BK - 1) / BK;
  const int per = (num_kb + S - 1) / S;
  const int kb0 = min(num_kb, rank * per), kb1 = min(num_kb, kb0 + per);
  const int nk = kb1 - kb0;

  // producer and issuer: the whole warp walks the loop, one elected lane issues (descriptors and coordinates stay in
  // uniform registers; see gemm.cu)
  if (warp == 0) {
    const bool leader = ptx::elect_one();
    int s = 0;
    uint32_t ph = 0;
    for (int kb = kb0; kb < kb1; ++kb) {
      ptx::mbar_wait(&empty[s], ph ^ 1);
      if (leader) {
        uint8_t* sa = ring + s * stage_bytes;
        ptx::mbar_arrive_expect_tx(&full[s], stage_bytes);
        ptx::tma_load_3d(sa, &tmA, &full[s], kb * BK, mt * BM, 0);
        ptx::tma_load_3d(sa + A_STAGE_BYTES, &tmW, &full[s], kb * BK, nt * BN, 0);
      }
      __syncwarp();
      if (++s == stages) { s = 0; ph ^= 1; }
    }
  } else if (warp == 1) {
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::umma_idesc_bf16(BM, BN, 0, 0);
    const uint32_t ring0 = ptx::smem_u32(ring);
    const uint64_t adesc0 = ptx::umma_smem_desc(ring0, 16, 1024), bdesc0 = ptx::umma_smem_desc(ring0 + A_STAGE_BYTES, 16, 1024);
    const uint32_t stage_step = static_cast<uint32_t>(stage_bytes) >> 4;
    int s = 0;
    uint32_t ph = 0;
    for (int i = 0; i < nk; ++i) {
      ptx::mbar_wait(&full[s], ph);
      ptx::tc_fence_after();
      if (leader) {
        const uint64_t ad = adesc0 + static_cast<uint64_t>(s * stage_step), bd = bdesc0 + static_cast<uint64_t>(s * stage_step);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k)
          ptx::umma_bf16_ss(tmem_base, ad + k * ((UMMA_K * 2) >> 4), bd + k * ((UMMA_K * 2) >> 4), idesc, (i | k) != 0);
        ptx::umma_commit(&empty[s]);
      }
      __syncwarp();
      if (++s == stages) { s = 0; ph ^= 1; }
    }
    if (leader) {
      if (nk > 0) ptx::umma_commit(tfull);
      else ptx::mbar_arrive(tfull);
    }
    __syncwarp();
  } else {
    const int q = warp & 3, rl = q * 32 + lane;                       // row of the tile this thread owns
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    if (rank == 0 && g.bias) {
      const float* bg = g.bias + nt * BN;
      for (int i = threadIdx.x - 64; i < BN; i += 128) sb[i] = (nt * BN + i < g.N) ? __ldg(bg + i) : 0.f;
      asm volatile("bar.sync 1, 128;" ::: "memory");
    }
    ptx::mbar_wait(tfull, 0);
    ptx::tc_fence_after();
    if (rank != 0) {
      // partial tile -> the leader's shared memory; float4 groups rotated by the row so that the 32 rows of a warp
      // do not all hit the same banks
      const uint32_t base = ptx::mapa(ptx::smem_u32(part), 0);
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(taddr + c * 32, v);
        ptx::tmem_ld_wait();
        const uint32_t dst = base + ((((rank - 1) * (BN / 32) + c) * BM + rl) * 32) * 4;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int jj = (j + rl) & 7;
          st_cluster_v4(dst + jj * 16, nk > 0 ? __uint_as_float(v[4 * j]) : 0.f, nk > 0 ? __uint_as_float(v[4 * j + 1]) : 0.f,
                        nk > 0 ? __uint_as_float(v[4 * j + 2]) : 0.f, nk > 0 ? __uint_as_float(v[4 * j + 3]) : 0.f);
        }
      }
    }
    ptx::tc_fence_before();
  }
  ptx::cluster_sync();                                                // partials of every CTA have landed in the leader
  if (rank == 0 && warp >= 2) {
    const int q = warp & 3, rl = q * 32 + lane;
    const int row = mt * BM + rl;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16);
    ptx::tc_fence_after();
    for (int c = 0; c < BN / 32; ++c) {
      const int n0 = nt * BN + c * 32;
      if (n0 >= g.N) break;
      uint32_t v[32];
      ptx::tmem_ld_32x32(taddr + c * 32, v);
      ptx::tmem_ld_wait();
      float acc[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) acc[i] = nk > 0 ? __uint_as_float(v[i]) : 0.f;
      for (int r = 0; r < S - 1; ++r) {
        const float* src = part + ((r * (BN / 32) + c) * BM + rl) * 32;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float4 t = *reinterpret_cast<const float4*>(src + ((j + rl) & 7) * 4);
          acc[4 * j] += t.x; acc[4 * j + 1] += t.y; acc[4 * j + 2] += t.z; acc[4 * j + 3] += t.w;
        }
      }
      if (row < M) epilogue_store<EPI, TOut, 32>(g, 0, row, n0, acc, g.bias ? sb + c * 32 : nullptr, true);
    }
    ptx::tc_fence_before();
  }
  __syncthreads();
  if (warp == 2) tmem_dealloc_n(tmem_base, tcols);
}
}  // namespace sk

static int g_split_k = -1;   // test / tuning hook: -1 = automatic, 0 = never, S >= 2 = force S where the shape allows it

// Split factor and tile width for M x N x K, or false when the problem is better served by the persistent kernel:
// K long enough that the serial k-loop dominates, and few enough tiles that S CTAs per tile still fit one wave.
bool splitk_plan(int M, int N, int K, long persistent_tiles, int* bn_out, int* s_out) {
  if (g_split_k == 0) return false;
  const int sms = num_sms();
  // K = 1024 .. 2047: only S = 2, and only when the persistent kernel would leave at least half of the SMs without a tile
  // (measured, tools/diag_splitk_k1024.py: 8.1 -> 6.6 us at M=256 N=1024; a loss once the persistent grid is well filled)
  if (g_split_k < 0 && K < 2048 && persistent_tiles * 2 > sms) return false;
  const int m_tiles = (M + sk::BM - 1) / sk::BM;
  const int num_kb = (K + sk::BK - 1) / sk::BK;
  const int s_max = g_split_k >= 2 ? g_split_k : (K >= 4096 ? 4 : (K >= 1024 ? 2 : 1));
  for (int S = std::min(s_max, 8); S >= 2; --S) {
    if (num_kb < 2 * S) continue;
    const int bns[3] = {32, 64, 128};
    for (int bn : bns) {
      if ((S - 1) * sk::BM * bn * 4 > 96 * 1024) break;              // partial tiles must leave room for the ring
      if (N < bn && bn > 32) break;
      const long ctas = static_cast<long>(m_tiles) * ((N + bn - 1) / bn) * S;
      if (ctas <= sms) {
        *bn_out = bn;
        *s_out = S;
        return true;
      }
    }
  }
  return false;
}

template <int EPI, typename TOut>
static int launch_splitk_t(const GemmArgs& g, int BN, int S, cudaStream_t st) {
  CUtensorMap tmA, tmW;
  TNT_TRY(make_tmap_bf16(&tmA, g.a, g.K, g.M, g.lda, 1, 0, sk::BM));
  TNT_TRY(make_tmap_bf16(&tmW, g.w, g.K, g.N, g.ldw, 1, 0, BN));
  const int part_bytes = (S - 1) * sk::BM * BN * 4;
  const int stage_bytes = sk::A_STAGE_BYTES + BN * sk::BK * 2;
  const int per = ((g.K + sk::BK - 1) / sk::BK + S - 1) / S;
  const int stages = std::max(2, std::min({sk::kMaxStages, (sk::kSmemBudget - 4096 - part_bytes) / stage_bytes, per}));
  const int smem = part_bytes + stages * stage_bytes + 1024 + 512 + 1024;
  const int tiles = ((g.M + sk::BM - 1) / sk::BM) * ((g.N + BN - 1) / BN);
  auto kern = sk::gemm_splitk_kernel<EPI, TOut>;
  static PerDeviceOnce attr_once;   // per instantiation and device
  if (attr_once.first()) TNT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, sk::kSmemBudget + 4096));
  TNT_CUDA(launch_pdl(kern, dim3(tiles * S), dim3(sk::kThreads), smem, st, S, tmA, tmW, g, BN, stages, S));
  TNT_LAUNCHED();
  return TNT_OK;
}

template <int EPI>
static int launch_splitk_e(const GemmArgs& g, int dt_out, int BN, int S, cudaStream_t st) {
  if (dt_out == TNT_BF16) return launch_splitk_t<EPI, bf16>(g, BN, S, st);
  return launch_splitk_t<EPI, float>(g, BN, S, st);
}

// NT layout, one group, static M; epilogues none / ReLU / GLU / residual
int launch_gemm_splitk(const GemmArgs& g, int epi, int dt_out, int BN, int S, cudaStream_t st) {
  switch (epi) {
    case EPI_NONE: return launch_splitk_e<EPI_NONE>(g, dt_out, BN, S, st);
    case EPI_RELU: return launch_splitk_e<EPI_RELU>(g, dt_out, BN, S, st);
    case EPI_GLU: return launch_splitk_e<EPI_GLU>(g, dt_out, BN, S, st);
    case EPI_RESIDUAL: return launch_splitk_e<EPI_RESIDUAL>(g, dt_out, BN, S, st);
  }
  set_error("split-K GEMM: unsupported epilogue %d", epi);
  return TNT_ERR_ARG;
}

}  // namespace tnt

extern "C" void tnt_debug_gemm_split_k(int s) { tnt::g_split_k = s; }
