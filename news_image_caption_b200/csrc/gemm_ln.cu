// Linear + residual + LayerNorm in ONE launch (bf16, tcgen05): Y = LayerNorm(A . W^T + bias + residual) * gamma + beta.
//   decoder layer sites (ttl/tell/models/decoder_flattened.py:283-291 linear2 -> +residual -> conv_layer_norm,
//   :300-330 out_proj of every source -> +residual -> context_attn_lns[source], :336-345 fc2 -> +residual -> final_layer_norm;
//   the same three sites in decoder_faces_objects.py:283-393)
// A LayerNorm needs the whole row (d = 1024 columns), a tcgen05 accumulator of 128 rows holds at most 512 fp32 columns, and a
// CTA that owned whole rows would leave only M / 128 CTAs with work.  So a CLUSTER of C = N / BN CTAs owns a 128-row block:
// CTA r reduces the full K for columns [r*BN, (r+1)*BN) into its own TMEM, adds bias + residual, and the row statistics cross
// the cluster through distributed shared memory:
//   pass A  TMEM -> x = acc + bias + residual, kept as fp32 in the (now idle) operand ring; per-row sum of the CTA's BN columns
//   pass B  per-row sum of squares around the CTA's own mean (Chan's parallel variance: no E[x^2] - mean^2 cancellation)
//   every CTA writes its (mean_r, M2_r) of each row into EVERY peer's stats[r][row] (st.shared::cluster), one cluster barrier
//   pass C  mean = sum mean_r / C, M2 = sum M2_r + BN * sum (mean_r - mean)^2, y = (x - mean) * rsqrt(M2 / N + eps) * gamma + beta
// One tile per CTA, grid = groups * ceil(M / 128) * C; warp 0 = TMA producer, warp 1 = MMA issuer, warps 2-9 = epilogue: two
// threads per row (one per half of the tile's columns), each a partial of its own in the exchange.  The residual segment of
// a thread (<= 128 bf16) is fetched into registers while the k-loop runs.  The statistics are taken on the fp32 sums (the
// two-launch form rounds the pre-LayerNorm tensor to bf16).  Cluster size: GPCs hold 16-20 SMs, so only 15 clusters of 8 CTAs
// with one CTA per SM are co-resident on a B200 (cudaOccupancyMaxActiveClusters) -- the plan asks the runtime and takes a
// wider tile / smaller cluster when the narrow one would need a second wave.
#include <cuda.h>

#include <algorithm>

#include "gemm.cuh"
#include "ptx.cuh"

namespace tnt {
int make_tmap_bf16(CUtensorMap* tm, const void* base, long cols, long rows, long ld, long groups, long group_stride,
                   int box_rows);

namespace gln {
constexpr int BM = 128, BK = 64, UMMA_K = 16, kEpiWarps = 8, kThreads = 64 + kEpiWarps * 32, kMaxStages = 8, kMaxCluster = 8;
constexpr int A_STAGE_BYTES = BM * BK * 2;
constexpr int kSmemBudget = 216 * 1024;

struct LnArgs {
  const float* gamma; const float* beta;   // [groups, N]
  long sgb;                                // group stride of gamma / beta
  float eps;
  int tT, tB;                              // transposed store: output row (t*B + b) -> (b*T + t); 0 = off
};

__device__ __forceinline__ void st_cluster_v2(uint32_t addr, float a, float b) {
  asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(a), "f"(b) : "memory");
}
__device__ __forceinline__ void tmem_alloc_n(uint32_t* smem_dst, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(ptx::smem_u32(smem_dst)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_n(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}

// smem: [ring: stages x (A 16 KB + W BN x 128 B), reused as the fp32 stash x[BN/32][128][32]][barriers 512 B]
//       [bias | gamma | beta: 3 x BN fp32][stats: 2C x 128 x (mean, M2)]
__global__ void __launch_bounds__(kThreads, 1)
gemm_ln_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, GemmArgs g, LnArgs ln, int BN,
               int stages, int C, int ring_bytes) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* ring = smem;
  const int stage_bytes = A_STAGE_BYTES + BN * BK * 2;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + ring_bytes);
  uint64_t* empty = full + kMaxStages;
  uint64_t* tfull = empty + kMaxStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tfull + 1);
  float* sb = reinterpret_cast<float*>(smem + ring_bytes + 512);      // bias [BN]
  float* sg = sb + BN;                                                // gamma [BN]
  float* sbt = sg + BN;                                               // beta [BN]
  float2* stats = reinterpret_cast<float2*>(sbt + BN);                // [2C][BM]: partial (rank, half) of every row
  float* stash = reinterpret_cast<float*>(ring);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = static_cast<int>(ptx::cluster_ctarank());
  const int cid = blockIdx.x / C;
  const uint32_t tcols = BN < 32 ? 32u : static_cast<uint32_t>(BN);

  // ---- prologue: no global data touched (overlaps the previous kernel's tail under PDL)
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmW);
  }
  if (warp == 3 && lane == 0 && g.pf_bytes) {                         // constant data: the next GEMM's weights into L2
    const unsigned long long chunk = ((g.pf_bytes + gridDim.x - 1) / gridDim.x + 15) & ~15ull;
    const unsigned long long off = blockIdx.x * chunk;
    if (off < g.pf_bytes) ptx::prefetch_l2_bulk(g.pf + off, static_cast<uint32_t>(min(chunk, g.pf_bytes - off) & ~15ull));
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
  const int grp = cid / m_tiles, mt = cid - grp * m_tiles;
  const int num_kb = (g.K + BK - 1) / BK;

  if (warp == 0) {
    // ---------------------------------------------------------------- TMA producer
    const bool leader = ptx::elect_one();
    int s = 0;
    uint32_t ph = 0;
    for (int kb = 0; kb < num_kb; ++kb) {
      ptx::mbar_wait(&empty[s], ph ^ 1);
      if (leader) {
        uint8_t* sa = ring + s * stage_bytes;
        ptx::mbar_arrive_expect_tx(&full[s], stage_bytes);
        ptx::tma_load_3d(sa, &tmA, &full[s], kb * BK, mt * BM, grp);
        ptx::tma_load_3d(sa + A_STAGE_BYTES, &tmW, &full[s], kb * BK, rank * BN, grp);
      }
      __syncwarp();
      if (++s == stages) { s = 0; ph ^= 1; }
    }
  } else if (warp == 1) {
    // ---------------------------------------------------------------- MMA issuer
    const bool leader = ptx::elect_one();
    const uint32_t idesc = ptx::umma_idesc_bf16(BM, BN, 0, 0);
    const uint32_t ring0 = ptx::smem_u32(ring);
    const uint64_t adesc0 = ptx::umma_smem_desc(ring0, 16, 1024), bdesc0 = ptx::umma_smem_desc(ring0 + A_STAGE_BYTES, 16, 1024);
    const uint32_t stage_step = static_cast<uint32_t>(stage_bytes) >> 4;
    int s = 0;
    uint32_t ph = 0;
    for (int i = 0; i < num_kb; ++i) {
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
    if (leader) ptx::umma_commit(tfull);
    __syncwarp();
  } else {
    // ---------------------------------------------------------------- epilogue, passes A and B
    const int q = warp & 3, half = (warp - 2) >> 2, rl = q * 32 + lane;   // row of the tile, half of its columns
    const int row = mt * BM + rl;
    const int HC = BN >> 1, nch = BN >> 6;                            // columns / 32-column chunks per thread
    const int col0 = half * HC;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + col0;
    // the thread's residual segment -> registers while the MMAs run (rows are 2 KB apart: fetched after the accumulator is
    // final, these loads were most of the epilogue)
    uint4 rr[16];
    {
      const bf16* rrow = g.res ? reinterpret_cast<const bf16*>(g.res) + grp * g.sres + static_cast<long>(row) * g.ldr + rank * BN + col0 : nullptr;
      const bool have = rrow && row < M;
#pragma unroll
      for (int i = 0; i < 16; ++i) rr[i] = (have && i * 8 < HC) ? __ldg(reinterpret_cast<const uint4*>(rrow) + i) : make_uint4(0, 0, 0, 0);
    }
    {
      const int e = threadIdx.x - 64;                                 // 0..255
      const long gb = static_cast<long>(grp) * ln.sgb + rank * BN;
      for (int i = e; i < BN; i += kEpiWarps * 32) {
        sb[i] = g.bias ? __ldg(g.bias + grp * g.sbias + rank * BN + i) : 0.f;
        sg[i] = __ldg(ln.gamma + gb + i);
        sbt[i] = __ldg(ln.beta + gb + i);
      }
      asm volatile("bar.sync 1, 256;" ::: "memory");
    }
    ptx::mbar_wait(tfull, 0);                                         // every MMA has completed: TMEM final, ring idle
    ptx::tc_fence_after();
    float s1 = 0.f;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (c < nch) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(taddr + c * 32, v);
        ptx::tmem_ld_wait();
        float* dst = stash + ((half * nch + c) * BM + rl) * 32;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint4 r4 = rr[c * 4 + j];
          const uint32_t rw[4] = {r4.x, r4.y, r4.z, r4.w};
          float x[8];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            x[2 * i] = __uint_as_float(v[8 * j + 2 * i]) + sb[col0 + c * 32 + 8 * j + 2 * i] + __uint_as_float(rw[i] << 16);
            x[2 * i + 1] = __uint_as_float(v[8 * j + 2 * i + 1]) + sb[col0 + c * 32 + 8 * j + 2 * i + 1] + __uint_as_float(rw[i] & 0xffff0000u);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) s1 += x[i];
          // float4 groups rotated by the row: the 32 rows of a warp do not all hit the same banks
          *reinterpret_cast<float4*>(dst + ((2 * j + rl) & 7) * 4) = make_float4(x[0], x[1], x[2], x[3]);
          *reinterpret_cast<float4*>(dst + ((2 * j + 1 + rl) & 7) * 4) = make_float4(x[4], x[5], x[6], x[7]);
        }
      }
    }
    ptx::tc_fence_before();
    const float mean_c = s1 / static_cast<float>(HC);
    float m2 = 0.f;
    for (int c = 0; c < nch; ++c) {
      const float* src = stash + ((half * nch + c) * BM + rl) * 32;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float4 t = *reinterpret_cast<const float4*>(src + ((j + rl) & 7) * 4);   // rotated like the writes (the order does not matter for the sum)
        const float d0 = t.x - mean_c, d1 = t.y - mean_c, d2 = t.z - mean_c, d3 = t.w - mean_c;
        m2 = fmaf(d0, d0, m2); m2 = fmaf(d1, d1, m2); m2 = fmaf(d2, d2, m2); m2 = fmaf(d3, d3, m2);
      }
    }
    const uint32_t mine = ptx::smem_u32(stats + (rank * 2 + half) * BM + rl);
    for (int r = 0; r < C; ++r) st_cluster_v2(ptx::mapa(mine, static_cast<uint32_t>(r)), mean_c, m2);
  }
  ptx::cluster_sync();                                                // every CTA's (mean, M2) of every row has landed everywhere
  if (warp >= 2) {
    // ---------------------------------------------------------------- pass C: normalise and store
    const int q = warp & 3, half = (warp - 2) >> 2, rl = q * 32 + lane;
    const int row = mt * BM + rl;
    const int HC = BN >> 1, nch = BN >> 6, col0 = half * HC, P = 2 * C;
    float mean = 0.f;
    for (int r = 0; r < P; ++r) mean += stats[r * BM + rl].x;
    mean /= static_cast<float>(P);
    float m2 = 0.f;
    for (int r = 0; r < P; ++r) {
      const float2 st = stats[r * BM + rl];
      const float d = st.x - mean;
      m2 += st.y + static_cast<float>(HC) * d * d;
    }
    const float rstd = rsqrtf(m2 / static_cast<float>(BN * C) + ln.eps);
    if (row < M) {
      long orow = row;
      if (ln.tT > 0) {
        const int t = row / ln.tB, b = row - t * ln.tB;
        orow = static_cast<long>(b) * ln.tT + t;
      }
      bf16* crow = reinterpret_cast<bf16*>(g.c) + grp * g.sc + orow * g.ldc + rank * BN + col0;
      for (int c = 0; c < nch; ++c) {
        const float* src = stash + ((half * nch + c) * BM + rl) * 32;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float4 a = *reinterpret_cast<const float4*>(src + ((2 * j + rl) & 7) * 4);
          const float4 b = *reinterpret_cast<const float4*>(src + ((2 * j + 1 + rl) & 7) * 4);
          const float x[8] = {a.x, a.y, a.z, a.w, b.x, b.y, b.z, b.w};
          float o[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int col = col0 + c * 32 + 8 * j + i;
            o[i] = (x[i] - mean) * rstd * sg[col] + sbt[col];
          }
          store8(crow + c * 32 + 8 * j, o);
        }
      }
    }
  }
  __syncthreads();
  if (warp == 2) tmem_dealloc_n(tmem_base, tcols);
}
}  // namespace gln

static int g_ln_fused = -1;      // -1: TNT_LN_FUSED from the environment (default on), 0 / 1: test hook

bool linear_ln_enabled() {
  if (g_ln_fused >= 0) return g_ln_fused != 0;
  static const bool env_on = [] { const char* e = getenv("TNT_LN_FUSED"); return !(e && e[0] == '0'); }();
  return env_on;
}

struct LnPlan { int BN, C, stages, ring_bytes, smem; };

static void ln_smem(int BN, int C, int K, LnPlan* pl) {
  const int stage_bytes = gln::A_STAGE_BYTES + BN * gln::BK * 2;
  const int stash_bytes = gln::BM * BN * 4;
  const int tail = 512 + 3 * BN * 4 + 2 * C * gln::BM * 8;
  const int num_kb = (K + gln::BK - 1) / gln::BK;
  pl->BN = BN;
  pl->C = C;
  pl->stages = std::max(2, std::min({gln::kMaxStages, (gln::kSmemBudget - 1024 - tail) / stage_bytes, num_kb}));
  pl->ring_bytes = (std::max(pl->stages * stage_bytes, stash_bytes) + 1023) & ~1023;
  pl->smem = pl->ring_bytes + tail + 1024;
}

// clusters of C CTAs with `smem` bytes each that the device can hold at once (GPC granularity: fewer than SMs / C)
static int ln_max_clusters(int C, int smem) {
  static int cache[kMaxDevices][gln::kMaxCluster + 1][4] = {};
  const int dev = cur_device(), slot = std::min(3, smem / (64 * 1024));
  int& c = cache[dev][C][slot];
  if (c) return c;
  static PerDeviceOnce attr_once;
  if (attr_once.first()) cudaFuncSetAttribute(gln::gemm_ln_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, gln::kSmemBudget + 4096);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(C * 64);
  cfg.blockDim = dim3(gln::kThreads);
  cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = C;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, gln::gemm_ln_kernel, &cfg) != cudaSuccess || n <= 0) {
    cudaGetLastError();
    n = std::max(1, (num_sms() / C) * 7 / 8);
  }
  c = n;
  return c;
}

static int g_ln_bn = 0;          // tuning hook: force the tile width (64 / 128 / 256) where it is legal

// tile width / cluster size for an N-wide row, or false: N = C * BN with C <= 8, BN in {64, 128, 256}.  The narrowest tile
// (widest cluster) whose clusters are all co-resident; else the shape with the least waves x tile cost.
static bool ln_plan(int M, int N, int K, int groups, LnPlan* out) {
  const long blocks = static_cast<long>(groups) * ((M + gln::BM - 1) / gln::BM);
  const int bns[3] = {64, 128, 256};
  const float cost[3] = {1.f, 1.f, 1.5f};
  float best = 0.f;
  bool found = false;
  for (int i = 0; i < 3; ++i) {
    const int bn = bns[i];
    if (N % bn) continue;
    const int c = N / bn;
    if (c < 1 || c > gln::kMaxCluster) continue;
    if (g_ln_bn && bn != g_ln_bn) continue;
    LnPlan pl;
    ln_smem(bn, c, K, &pl);
    if (pl.smem > gln::kSmemBudget + 4096) continue;
    const int cap = std::min(ln_max_clusters(c, pl.smem), std::max(1, num_sms() / c));   // (a caller's SM budget counts too)
    const long waves = (blocks + cap - 1) / cap;
    const float t = waves * cost[i];
    if (!found || t < best) {
      best = t;
      *out = pl;
      found = true;
    }
  }
  return found;
}

bool linear_ln_supported(const tnt_linear_params* p, const float* gamma, const float* beta) {
  if (!p || !gamma || !beta || !p->a || !p->w || !p->c) return false;
  if (p->dtype_in != TNT_BF16 || p->dtype_out != TNT_BF16 || p->layout != TNT_NT || p->m_dyn) return false;
  if (p->epilogue != EPI_RESIDUAL && p->epilogue != EPI_NONE) return false;
  if (p->epilogue == EPI_RESIDUAL && !p->residual) return false;
  if (p->alpha != 0.f && p->alpha != 1.f) return false;
  if (p->drop.p > 0.f) return false;
  if (p->M <= 0 || p->K <= 0 || p->K % 8) return false;
  const int groups = p->groups > 0 ? p->groups : 1;
  LnPlan pl;
  if (!ln_plan(p->M, p->N, p->K, groups, &pl)) return false;
  if (!aligned16(p->a) || !aligned16(p->w) || !aligned16(p->c) || p->lda % 8 || p->ldw % 8 || p->ldc % 8 || p->stride_a % 8 ||
      p->stride_w % 8 || p->stride_c % 8)
    return false;
  if (p->epilogue == EPI_RESIDUAL && (!aligned16(p->residual) || p->ldr % 8 || p->stride_res % 8)) return false;
  return true;
}

int launch_linear_ln(const tnt_linear_params* p, const float* gamma, const float* beta, long stride_gb, float eps, int tT, int tB,
                     cudaStream_t st) {
  TNT_CHECK_ARG(linear_ln_supported(p, gamma, beta), "linear_ln: unsupported problem (bf16 NT, N = C * BN with C <= 8, no dropout)");
  TNT_CHECK_ARG((tT == 0 && tB == 0) || (tT > 0 && tB > 0 && static_cast<long>(tT) * tB == p->M && (p->groups <= 1)),
                "linear_ln: transposed store needs T * B == M and one group");
  GemmArgs g = {};
  g.a = p->a; g.w = p->w; g.bias = p->bias; g.res = p->epilogue == EPI_RESIDUAL ? p->residual : nullptr; g.c = p->c;
  g.lda = p->lda; g.ldw = p->ldw; g.ldr = p->ldr; g.ldc = p->ldc;
  g.M = p->M; g.N = p->N; g.K = p->K;
  g.groups = p->groups > 0 ? p->groups : 1;
  g.sa = p->stride_a; g.sw = p->stride_w; g.sbias = p->stride_bias; g.sres = p->stride_res; g.sc = p->stride_c;
  g.alpha = 1.f;
  static const bool pf_on = [] { const char* e = getenv("TNT_L2_PREFETCH"); return !(e && e[0] == '0'); }();
  g.pf = (pf_on && p->prefetch && aligned16(p->prefetch)) ? static_cast<const char*>(p->prefetch) : nullptr;
  g.pf_bytes = g.pf ? (p->prefetch_bytes & ~15ull) : 0;
  LnPlan pl;
  ln_plan(g.M, g.N, g.K, g.groups, &pl);
  const int BN = pl.BN, C = pl.C, stages = pl.stages, ring_bytes = pl.ring_bytes, smem = pl.smem;
  gln::LnArgs ln = {gamma, beta, g.groups > 1 ? stride_gb : 0, eps, tT, tB};
  CUtensorMap tmA, tmW;
  TNT_TRY(make_tmap_bf16(&tmA, g.a, g.K, g.M, g.lda, g.groups, g.sa, gln::BM));
  TNT_TRY(make_tmap_bf16(&tmW, g.w, g.K, g.N, g.ldw, g.groups, g.sw, BN));
  const int grid = g.groups * ((g.M + gln::BM - 1) / gln::BM) * C;
  TNT_CUDA(launch_pdl(gln::gemm_ln_kernel, dim3(grid), dim3(gln::kThreads), smem, st, C, tmA, tmW, g, ln, BN, stages, C, ring_bytes));
  TNT_LAUNCHED();
  variant_hit(V_LN_GEMM_FUSED);
  return TNT_OK;
}

}  // namespace tnt

extern "C" int tnt_linear_ln_supported(const tnt_linear_params* p, const float* gamma, const float* beta) {
  return tnt::linear_ln_supported(p, gamma, beta) ? 1 : 0;
}
extern "C" int tnt_linear_ln_fwd(const tnt_linear_params* p, const float* gamma, const float* beta, int64_t stride_gb, float eps,
                                 int transpose_T, int transpose_B, tnt_stream stream) {
  return tnt::launch_linear_ln(p, gamma, beta, static_cast<long>(stride_gb), eps, transpose_T, transpose_B, tnt::as_stream(stream));
}
extern "C" void tnt_debug_ln_fused(int on) { tnt::g_ln_fused = on < 0 ? -1 : (on ? 1 : 0); }
extern "C" void tnt_debug_ln_tile_n(int bn) { tnt::g_ln_bn = (bn == 64 || bn == 128 || bn == 256) ? bn : 0; }
