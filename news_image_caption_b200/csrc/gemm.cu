// GEMM kernels behind tnt_linear_fwd / tnt_lse_fwd.
//   bf16: persistent warp-specialised tcgen05 kernel -- TMA (128B swizzle) -> smem ring -> tcgen05.mma
//         (single issuing thread, fp32 accumulators double-buffered in TMEM) -> tcgen05.ld epilogue.
//   fp32: plain SIMT fp32 kernel (parity mode; also the in-library cross-check for the tcgen05 path).
// Replaces GehringLinear / TiedLinear / F.linear: ttl/tell/modules/linear.py:7-50.
#include <cuda.h>
#include <stdlib.h>

#include <algorithm>

#include "gemm.cuh"
#include "ptx.cuh"

namespace tnt {

// ============================================================================ tcgen05 kernel
namespace tc {
constexpr int BM = 128, BK = 64, UMMA_K = 16;
// warp 0: TMA producer, warp 1: MMA issuer, warps 2..: epilogue (4 warps = one per TMEM lane quadrant).  The log-sum-exp
// epilogue (EPI_LSE: no output matrix, ~11 instructions and one MUFU per logit) is the whole cost of the adaptive-softmax
// GEMMs -- K = 64 .. 1024 against up to 30265 columns -- and with one warp per scheduler it ran at the latency of its
// dependent chains; it gets 8 warps, two per quadrant, each owning half of the tile's columns.
template <int EPI> struct EpiWarps { static constexpr int n = EPI == EPI_LSE ? 8 : 4; static constexpr int threads = 64 + 32 * n; };
constexpr int A_STAGE_BYTES = BM * BK * 2;
constexpr uint32_t kTmemCols = 512;       // 2 accumulator stages x 256 fp32 columns
constexpr int kAccStride = 256;
// interleaved accumulators of a single-CTA tile: nacc * BN <= kAccStride columns per accumulator stage
inline int acc_count(int cg, int bn) { return cg != 1 ? 1 : bn <= 64 ? 4 : bn <= 128 ? 2 : 1; }
constexpr int kMaxStages = 8;
constexpr int kSched = 4;                 // depth of the tile-index ring of the dynamic scheduler
constexpr int kSmemBudget = 220 * 1024;

// tile index -> (m tile, n tile, group).  Tiles are walked in bands of kBandRows output rows: inside a band the m tile
// runs fastest, then the n tile, so the CTAs running at any moment share a few A row-blocks and W column-blocks
// (both stay L2-resident while the band sweeps all of N) and A is fetched from HBM once instead of once per n tile.
constexpr int kBandRows = 2048;
__device__ __forceinline__ void decode_tile(int tile, int m_tiles, int n_tiles, int band_m, int* mt, int* nt, int* grp) {
  const int per_grp = m_tiles * n_tiles;
  *grp = tile / per_grp;
  const int t = tile - *grp * per_grp;
  const int band = t / (band_m * n_tiles);
  const int m_lo = band * band_m;
  const int bm = min(band_m, m_tiles - m_lo);
  const int r = t - band * band_m * n_tiles;
  *nt = r / bm;
  *mt = m_lo + r - *nt * bm;
}

// CG = 1: one CTA per 128 x BN tile.  CG = 2: a CTA PAIR (cluster of 2 on one TPC) per 256 x BN tile --
// each CTA stages its own 128 A rows and BN/2 of the W rows, the leader issues tcgen05.mma.cta_group::2
// (M = 256), so every CTA reads half as much B from shared memory / L2 per FLOP as in the CG = 1 form.
template <int EPI, typename TOut, int CG>
__global__ void __launch_bounds__(EpiWarps<EPI>::threads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, GemmArgs g,
               int BN, int stages, int n_tiles) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  constexpr int BMT = BM * CG;                         // output rows per tile
  const int BNL = BN / CG;                             // W rows (output columns) this CTA stages
  const int stage_bytes = A_STAGE_BYTES + BNL * BK * 2;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + stages * stage_bytes);
  uint64_t* empty = full + kMaxStages;
  uint64_t* tfull = empty + kMaxStages;
  uint64_t* tempty = tfull + 2;
  uint64_t* sfull = tempty + 2;                        // dynamic tile scheduler: tile-index ring, producer -> MMA + epilogue warps
  uint64_t* sempty = sfull + kSched;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sempty + kSched);
  volatile int* stile = reinterpret_cast<volatile int*>(tmem_slot + 2);
  // Tiles are handed out by an atomic counter instead of static striding: a CTA that is not resident yet (its SM is held by
  // another stream's kernel -- the NCCL all-reduce of the training step's gradient exchange overlaps the backward pass)
  // simply takes no tile, and the resident CTAs absorb the work.  With static striding every overlapped GEMM waited for the
  // SMs the collective held: +0.47 ms of kernel time per step at 2 GPUs, +1.1 ms at 8 (tools/trace_train.py).
  const bool dyn_on = CG == 1 && g.tile_ctr != nullptr;
  float* sbias = reinterpret_cast<float*>(smem + stages * stage_bytes + 512);   // [2 accumulator stages][256], zero padded

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = CG == 2 ? static_cast<int>(ptx::cluster_ctarank()) : 0;
  const int worker = blockIdx.x / CG, n_workers = gridDim.x / CG;

  // ---- prologue: touches no global data, so under PDL it overlaps the tail of the preceding kernel
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmW);
  }
  if (warp == 3 && lane == 0 && g.pf_bytes) {
    // constant data (the next GEMM's weights): safe before the PDL wait; every CTA requests its slice into L2
    const unsigned long long chunk = ((g.pf_bytes + gridDim.x - 1) / gridDim.x + 15) & ~15ull;
    const unsigned long long off = blockIdx.x * chunk;
    if (off < g.pf_bytes) ptx::prefetch_l2_bulk(g.pf + off, static_cast<uint32_t>(min(chunk, g.pf_bytes - off) & ~15ull));
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      ptx::mbar_init(&full[s], 1);
      ptx::mbar_init(&empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      ptx::mbar_init(&tfull[s], 1);
      ptx::mbar_init(&tempty[s], EpiWarps<EPI>::n * CG);   // the leader waits for the epilogue warps of BOTH CTAs
    }
    for (int s = 0; s < kSched; ++s) {
      ptx::mbar_init(&sfull[s], 1);
      ptx::mbar_init(&sempty[s], 1 + EpiWarps<EPI>::n);    // the MMA warp + the epilogue warps have read the slot
    }
    ptx::fence_barrier_init();
  }
  // first tile of this CTA: the counter belongs to this launch alone, so the fetch may run ahead of the PDL wait and its
  // latency hides behind the barrier initialisation / TMEM allocation
  int first_tile = 0;
  if (dyn_on && warp == 0) {
    if (lane == 0) first_tile = atomicAdd(g.tile_ctr, 1);
    first_tile = __shfl_sync(0xffffffffu, first_tile, 0);
  }
  if (warp == 2) {
    if (CG == 2) ptx::tmem_alloc_pair<kTmemCols>(tmem_slot);
    else ptx::tmem_alloc<kTmemCols>(tmem_slot);
  }
  ptx::tc_fence_before();
  if (CG == 2) ptx::cluster_sync();
  else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  pdl_wait();                                          // the preceding kernel's outputs are visible from here on
  // the device-side count bounds the token dimension: output rows (NT / NN) or the reduction (TN)
  const int dyn = g.m_dyn ? *reinterpret_cast<const volatile int*>(g.m_dyn) : 0x7fffffff;
  const int M = g.a_mn ? g.M : min(g.M, dyn);
  const int Kred = g.a_mn ? min(g.K, dyn) : g.K;
  const int m_tiles = (M + BMT - 1) / BMT;
  const int total_tiles = g.groups * m_tiles * n_tiles;
  const int num_kb = (Kred + BK - 1) / BK;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (warp 0 of every CTA)
    // the whole warp walks the loop, one elected lane issues (uniform registers for the descriptors / coordinates; a
    // single divergent lane paid an ELECT + 5 R2UR.BROADCAST per bulk copy)
    {
      const bool leader = ptx::elect_one();
      int s = 0;
      uint32_t ph = 0;
      int ss = 0;
      uint32_t sph = 0;
      for (int tile = dyn_on ? first_tile : worker;;) {
        int next_tile = 0;
        if (dyn_on) {
          const int pub = tile < total_tiles ? tile : -1;           // -1: no more tiles
          ptx::mbar_wait(&sempty[ss], sph ^ 1);
          if (lane == 0) {
            stile[ss] = pub;
            ptx::mbar_arrive(&sfull[ss]);
          }
          __syncwarp();
          if (++ss == kSched) { ss = 0; sph ^= 1; }
          if (pub < 0) break;
          if (lane == 0) next_tile = atomicAdd(g.tile_ctr, 1);       // prefetched behind this tile's loads
        } else if (tile >= total_tiles) {
          break;
        }
        int mt, nt, grp;
        decode_tile(tile, m_tiles, n_tiles, kBandRows / BMT, &mt, &nt, &grp);
        const int ga = g.sa ? grp : 0, gw = g.sw ? grp : 0;
        const int m0 = mt * BMT + rank * BM, n0 = nt * BN + rank * BNL;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&empty[s], ph ^ 1);
          if (leader) {
          uint8_t* sa = smem + s * stage_bytes;
          if (CG == 1) {
            ptx::mbar_arrive_expect_tx(&full[s], stage_bytes);
            if (!g.a_mn) {
              ptx::tma_load_3d(sa, &tmA, &full[s], kb * BK, m0, ga);
            } else {                                   // A stored [K, M]: two 64-column atoms of 64 k-rows
              ptx::tma_load_3d(sa, &tmA, &full[s], m0, kb * BK, ga);
              ptx::tma_load_3d(sa + 8192, &tmA, &full[s], m0 + 64, kb * BK, ga);
            }
            if (!g.b_mn) {
              ptx::tma_load_3d(sa + A_STAGE_BYTES, &tmW, &full[s], kb * BK, n0, gw);
            } else {                                   // W stored [K, N]: BN/64 atoms
              for (int at = 0; at < BNL / 64; ++at)
                ptx::tma_load_3d(sa + A_STAGE_BYTES + at * 8192, &tmW, &full[s], n0 + at * 64, kb * BK, gw);
            }
          } else {
            // all bytes of the pair are counted on the LEADER's full barrier
            if (rank == 0) ptx::mbar_arrive_expect_tx(&full[s], stage_bytes * CG);
            const uint32_t fb = ptx::mapa(ptx::smem_u32(&full[s]), 0);
            if (!g.a_mn) {
              ptx::tma_load_3d_pair(sa, &tmA, fb, kb * BK, m0, ga);
            } else {
              ptx::tma_load_3d_pair(sa, &tmA, fb, m0, kb * BK, ga);
              ptx::tma_load_3d_pair(sa + 8192, &tmA, fb, m0 + 64, kb * BK, ga);
            }
            if (!g.b_mn) {
              ptx::tma_load_3d_pair(sa + A_STAGE_BYTES, &tmW, fb, kb * BK, n0, gw);
            } else {
              for (int at = 0; at < BNL / 64; ++at)
                ptx::tma_load_3d_pair(sa + A_STAGE_BYTES + at * 8192, &tmW, fb, n0 + at * 64, kb * BK, gw);
            }
          }
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
        tile = dyn_on ? __shfl_sync(0xffffffffu, next_tile, 0) : tile + n_workers;
      }
      if (dyn_on && lane == 0) {                       // the last CTA to finish returns the counters to zero for the slot's next user
        __threadfence();
        if (atomicAdd(g.tile_ctr + 1, 1) == static_cast<int>(gridDim.x) - 1) {
          g.tile_ctr[0] = 0;
          g.tile_ctr[1] = 0;
          __threadfence();
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (warp 1 of the leader CTA)
    // The WHOLE warp walks the loop and one elected lane issues: with a single divergent lane every descriptor lived in
    // vector registers and each tcgen05.mma cost an ELECT + 5 R2UR.BROADCAST + ~25 predicated descriptor instructions
    // -- ~150 SASS instructions = 660 cycles per k-block, which (not the tensor pipe, not TMA) bounded every k-loop
    // (ncu source view: the producer spins on `empty`, the issuer never waits on `full`).  Warp-uniform code keeps the
    // descriptors in uniform registers; they are built once and advanced by adding to their 14-bit address field.
    if (rank == 0) {
      const bool leader = ptx::elect_one();
      const uint32_t idesc = ptx::umma_idesc_bf16(BMT, BN, g.a_mn, g.b_mn);
      // K-major: 32 bytes per K step inside the 128-byte row; MN-major: 16 k-rows x 128 B per step,
      // 64-element atoms 8192 B apart (LBO), 8-row groups 1024 B apart (SBO)
      const uint32_t smem0 = ptx::smem_u32(smem);
      const uint64_t adesc0 = g.a_mn ? ptx::umma_smem_desc(smem0, 8192, 1024) : ptx::umma_smem_desc(smem0, 16, 1024);
      const uint64_t bdesc0 = g.b_mn ? ptx::umma_smem_desc(smem0 + A_STAGE_BYTES, 8192, 1024)
                                     : ptx::umma_smem_desc(smem0 + A_STAGE_BYTES, 16, 1024);
      const uint32_t ka = (g.a_mn ? 2048u : UMMA_K * 2u) >> 4, kbs = (g.b_mn ? 2048u : UMMA_K * 2u) >> 4;   // per K step
      const uint32_t stage_step = static_cast<uint32_t>(stage_bytes) >> 4;
      const int nacc = CG == 1 ? g.nacc : 1;
      int s = 0;
      uint32_t ph = 0;
      int cs = 0;
      uint32_t cph = 0;
      for (int it = 0;; ++it) {
        if (dyn_on) {
          ptx::mbar_wait(&sfull[cs], cph);
          const int t = stile[cs];
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&sempty[cs]);
          if (++cs == kSched) { cs = 0; cph ^= 1; }
          if (t < 0) break;
        } else if (worker + it * n_workers >= total_tiles) {
          break;
        }
        const int as = it & 1;
        const uint32_t aph = (it >> 1) & 1;
        ptx::mbar_wait(&tempty[as], aph ^ 1);          // epilogue has drained this accumulator stage
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * kAccStride;
        for (int kb = 0; kb < num_kb; ++kb) {
          ptx::mbar_wait(&full[s], ph);                // TMA bytes have landed (in both CTAs)
          ptx::tc_fence_after();
          if (leader) {
            const uint64_t ad = adesc0 + static_cast<uint64_t>(s * stage_step);
            const uint64_t bd = bdesc0 + static_cast<uint64_t>(s * stage_step);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              if (CG == 2) ptx::umma_bf16_ss_pair(d_tmem, ad + k * ka, bd + k * kbs, idesc, (kb | k) != 0);
              // narrow single-CTA tiles rotate over `nacc` accumulators BN columns apart; the epilogue adds them
              else ptx::umma_bf16_ss(d_tmem + (k & (nacc - 1)) * BN, ad + k * ka, bd + k * kbs, idesc, kb != 0 || k >= nacc);
            }
            if (CG == 2) ptx::umma_commit_pair(&empty[s]);   // frees the smem slot of both CTAs when these MMAs retire
            else ptx::umma_commit(&empty[s]);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
        if (leader) {
          if (num_kb > 0) {                            // accumulator complete -> epilogue (of both CTAs)
            if (CG == 2) ptx::umma_commit_pair(&tfull[as]);
            else ptx::umma_commit(&tfull[as]);
          } else {                                     // empty reduction: the epilogue writes zeros
            ptx::mbar_arrive(&tfull[as]);
            if (CG == 2) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&tfull[as]), 1));
          }
        }
        __syncwarp();
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------ epilogue warps (TMEM -> regs -> HBM)
    const int q = warp & 3;                            // TMEM lane quadrant this warp may read
    const int half = (warp - 2) >> 2;                  // EPI_LSE: which half of the tile's columns (0 for the 4-warp epilogues)
    const int halves = EPI == EPI_LSE ? g.lse_halves : 1;
    const int nacc_e = CG == 1 ? g.nacc : 1;
    int cs = 0;
    uint32_t cph = 0;
    for (int it = 0;; ++it) {
      int tile;
      if (dyn_on) {
        ptx::mbar_wait(&sfull[cs], cph);
        tile = stile[cs];
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&sempty[cs]);
        if (++cs == kSched) { cs = 0; cph ^= 1; }
        if (tile < 0) break;
      } else {
        tile = worker + it * n_workers;
        if (tile >= total_tiles) break;
      }
      int mt, nt, grp;
      decode_tile(tile, m_tiles, n_tiles, kBandRows / BMT, &mt, &nt, &grp);
      const int as = it & 1;
      const uint32_t aph = (it >> 1) & 1;
      // this tile's bias slice -> shared memory (one coalesced read per tile instead of 32 scalar loads per thread and
      // chunk); the named barrier also orders it against the reads of tile it-2, which used the same slot
      float* sb = sbias + as * 256;
      if (g.bias) {
        const float* bg = g.bias + grp * g.sbias + nt * BN;
        for (int i = threadIdx.x - 64; i < BN; i += 32 * EpiWarps<EPI>::n) sb[i] = (nt * BN + i < g.N) ? __ldg(bg + i) : 0.f;
        asm volatile("bar.sync 1, %0;" ::"n"(32 * EpiWarps<EPI>::n) : "memory");
      }
      ptx::mbar_wait(&tfull[as], aph);
      ptx::tc_fence_after();
      const int row = mt * BMT + rank * BM + q * 32 + lane;
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * kAccStride;
      LseAcc la;
      int tgt = -1;
      if (EPI == EPI_LSE) {
        la.init();
        if (row < M && g.target) tgt = g.target[row];
      }
      // one 32-column chunk of this row: accumulator values in v -> bias / epilogue -> global
      auto process = [&](uint32_t (&v)[32], int c) {
        const int n0 = nt * BN + c * 32;
        if (EPI != EPI_LSE && EPI != EPI_GLU && g.m_dyn && !g.a_mn && row >= M && row < g.M) {
          // pad rows of a dynamic-M problem get zeros: a later TN GEMM reduces over whole 64-row blocks
          zero_store<TOut, 32>(g, grp, row, n0);
        }
        if (EPI == EPI_NONE && sizeof(TOut) == 4 && !g.vec_ok && !g.m_dyn) {
          // unaligned fp32 rows: bias, then a register transpose so that the stores are coalesced (warp-collective)
          float acc[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[i] = (num_kb > 0 ? __uint_as_float(v[i]) : 0.f) * g.alpha;
          if (g.bias) {
#pragma unroll
            for (int i = 0; i < 32; ++i) acc[i] += sb[c * 32 + i];
          }
          store_block_transposed(g, grp, row - lane, M, n0, acc, lane);
          return;
        }
        if (row < M) {
          float acc[32];
#pragma unroll
          for (int i = 0; i < 32; ++i) acc[i] = num_kb > 0 ? __uint_as_float(v[i]) : 0.f;
          if (EPI == EPI_LSE) {
            if (g.bias) {
#pragma unroll
              for (int i = 0; i < 32; ++i) acc[i] += sb[c * 32 + i];
            }
            if (n0 + 32 <= g.n_lse) {                  // common case: whole chunk inside the cluster
              la.add_chunk<32>(acc, n0, tgt);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                const int col = n0 + i;
                if (col < g.N) {
                  if (col < g.n_lse) la.add(acc[i], col, tgt);
                  else g.side[static_cast<long>(row) * g.side_ld + col - g.n_lse] = acc[i];
                }
              }
            }
          } else {
            epilogue_store<EPI, TOut, 32>(g, grp, row, n0, acc, g.bias ? sb + c * 32 : nullptr, true);
          }
        }
      };
      const int nch = min(BN / 32, (g.N - nt * BN + 31) / 32);      // chunks with live columns (warp-uniform)
      // (measured: software-pipelining these TMEM loads -- chunk c+1 in flight while chunk c is converted and stored -- is
      // SLOWER: forward 1.629 vs 1.600 ms, generation 25.55 vs 24.93 ms, training 8.29 vs 8.21 ms on one box; the epilogue is
      // not on the critical path, the extra 70 registers and the longer code are)
      // EPI_LSE with two warps per quadrant: warp-half h owns chunks [h * BN/64, (h + 1) * BN/64) and its own partial record
      const int c_lo = halves == 2 ? half * (BN / 64) : (half ? nch : 0);
      const int c_hi = halves == 2 ? min(nch, c_lo + BN / 64) : nch;
      {
        for (int c = c_lo; c < c_hi; ++c) {
          uint32_t v[32];
          ptx::tmem_ld_32x32(taddr + c * 32, v);
          ptx::tmem_ld_wait();
          for (int j = 1; j < nacc_e; ++j) {               // the other accumulators of a narrow tile (see the MMA issuer)
            uint32_t u[32];
            ptx::tmem_ld_32x32(taddr + j * BN + c * 32, u);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) + __uint_as_float(u[i]));
          }
          process(v, c);
        }
      }
      if (EPI == EPI_LSE && row < M && half < halves)
        g.partials[static_cast<long>(row) * g.n_tiles + nt * halves + half] = make_float4(la.m, la.s, __int_as_float(la.arg), la.tz);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (CG == 2) ptx::mbar_arrive_cluster_relaxed(ptx::mapa(ptx::smem_u32(&tempty[as]), 0));
        else ptx::mbar_arrive(&tempty[as]);
      }
    }
  }
  ptx::tc_fence_before();
  if (CG == 2) ptx::cluster_sync();
  else __syncthreads();
  if (warp == 2) {
    if (CG == 2) ptx::tmem_dealloc_pair<kTmemCols>(tmem_base);
    else ptx::tmem_dealloc<kTmemCols>(tmem_base);
  }
}
}  // namespace tc

// ============================================================================ SIMT kernel
namespace simt {
constexpr int TM = 64, TN = 64, TK = 16;

template <int EPI, typename TIn, typename TOut>
__global__ void __launch_bounds__(256) gemm_simt_kernel(GemmArgs g) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Ws[TK][TN + 4];
  const int grp = blockIdx.z;
  const int dyn = g.m_dyn ? __ldg(g.m_dyn) : 0x7fffffff;
  const int M = g.a_mn ? g.M : min(g.M, dyn);
  const int Kred = g.a_mn ? min(g.K, dyn) : g.K;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  if (m0 >= M && !(EPI != EPI_LSE && EPI != EPI_GLU && g.m_dyn && !g.a_mn && m0 < g.M && m0 < ((M + 127) & ~127))) return;
  const TIn* A = reinterpret_cast<const TIn*>(g.a) + grp * g.sa;
  const TIn* W = reinterpret_cast<const TIn*>(g.w) + grp * g.sw;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k0 = 0; k0 < Kred; k0 += TK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int idx = threadIdx.x + i * 256;
      {
        const int r = g.a_mn ? (idx & 63) : (idx >> 4), kk = g.a_mn ? (idx >> 6) : (idx & 15);
        const bool in = (k0 + kk) < Kred && (m0 + r) < M;
        const long off = g.a_mn ? static_cast<long>(k0 + kk) * g.lda + m0 + r : static_cast<long>(m0 + r) * g.lda + k0 + kk;
        As[kk][r] = in ? to_f(A[off]) : 0.f;
      }
      {
        const int r = g.b_mn ? (idx & 63) : (idx >> 4), kk = g.b_mn ? (idx >> 6) : (idx & 15);
        const bool in = (k0 + kk) < Kred && (n0 + r) < g.N;
        const long off = g.b_mn ? static_cast<long>(k0 + kk) * g.ldw + n0 + r : static_cast<long>(n0 + r) * g.ldw + k0 + kk;
        Ws[kk][r] = in ? to_f(W[off]) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < TK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Ws[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = m0 + ty * 4 + i;
    if (EPI == EPI_LSE) {
      LseAcc la;
      la.init();
      const int tgt = (row < M && g.target) ? g.target[row] : -1;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int col = n0 + tx * 4 + j;
        if (col < g.N) {
          const float z = acc[i][j] + (g.bias ? __ldg(g.bias + col) : 0.f);
          if (col < g.n_lse) la.add(z, col, tgt);
          else if (row < M) g.side[static_cast<long>(row) * g.side_ld + col - g.n_lse] = z;
        }
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) {
        const float m2 = __shfl_xor_sync(0xffffffffu, la.m, o), s2 = __shfl_xor_sync(0xffffffffu, la.s, o);
        const float t2 = __shfl_xor_sync(0xffffffffu, la.tz, o);
        const int a2 = __shfl_xor_sync(0xffffffffu, la.arg, o);
        la.merge(m2, s2, a2, t2);
      }
      if (tx == 0 && row < M)
        g.partials[static_cast<long>(row) * g.n_tiles + blockIdx.x] =
            make_float4(la.m, la.s, __int_as_float(la.arg), la.tz);
    } else if (row < M) {
      epilogue_store<EPI, TOut, 4>(g, grp, row, n0 + tx * 4, acc[i], g.bias ? g.bias + grp * g.sbias + n0 + tx * 4 : nullptr, false);
    } else if (EPI != EPI_LSE && EPI != EPI_GLU && g.m_dyn && !g.a_mn && row < g.M) {
      zero_store<TOut, 4>(g, grp, row, n0 + tx * 4);
    }
  }
}
}  // namespace simt

// ============================================================================ host dispatch
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 3-D bf16 map {cols (K, contiguous), rows, groups}; box {64, box_rows, 1}; 128-byte swizzle.
int make_tmap_bf16(CUtensorMap* tm, const void* base, long cols, long rows, long ld, long groups,
                   long group_stride, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point unavailable");
    return TNT_ERR_CUDA;
  }
  if (groups < 1 || group_stride == 0) groups = 1;
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows),
                        static_cast<cuuint64_t>(groups)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ld) * 2,
                           static_cast<cuuint64_t>(groups > 1 ? group_stride : ld * std::max<long>(rows, 1)) * 2};
  cuuint32_t box[3] = {64, static_cast<cuuint32_t>(box_rows), 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d): base %p cols %ld rows %ld ld %ld groups %ld gstride %ld", (int)r,
              base, cols, rows, ld, groups, group_stride);
    return TNT_ERR_CUDA;
  }
  return TNT_OK;
}

// generic 3-D bf16 map, dims {d0 (contiguous), d1, d2}, element strides, 128-byte swizzle (box0 must be 64)
int make_tmap_bf16_3d(CUtensorMap* tm, const void* base, long d0, long d1, long d2, long stride1, long stride2,
                      int box0, int box1, int box2) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point unavailable");
    return TNT_ERR_CUDA;
  }
  cuuint64_t dims[3] = {static_cast<cuuint64_t>(d0), static_cast<cuuint64_t>(d1), static_cast<cuuint64_t>(d2)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(stride1) * 2, static_cast<cuuint64_t>(stride2) * 2};
  cuuint32_t box[3] = {static_cast<cuuint32_t>(box0), static_cast<cuuint32_t>(box1), static_cast<cuuint32_t>(box2)};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled(3d) failed (%d): dims %ld %ld %ld strides %ld %ld", (int)r, d0, d1, d2, stride1,
              stride2);
    return TNT_ERR_CUDA;
  }
  return TNT_OK;
}

static int g_cta_group = 0;   // test hook: 0 = automatic, 1 / 2 = force the single-CTA / CTA-pair kernel
static int g_force_stages = 0;   // tuning hook: cap the smem ring depth
static int g_single_acc = 1;  // tuning hook: 0 = narrow tiles rotate their MMAs over 2-4 TMEM accumulators (measured: no gain --
                              // the k-loop was bound by the issuing thread's instruction stream, not by accumulator dependencies)
static int g_force_bn = 0;    // tuning hook: force the tile width (32..256; CTA pair: 64..256)

// tile shape the tcgen05 dispatcher uses: CTA-pair 256 x 256 tiles when the problem fills the 74 SM pairs for
// several waves, else single-CTA 128 x BN with BN shrunk until the grid covers the SMs
static void pick_tile(int M, int N, int groups, bool b_mn, int* bn_out, int* cg_out) {
  const int sms = num_sms();
  if (g_force_bn) {
    *cg_out = g_cta_group == 2 ? 2 : 1;
    *bn_out = g_force_bn;
    if (b_mn && *bn_out / *cg_out < 64) *bn_out = 64 * *cg_out;
    return;
  }
  if (g_cta_group != 1 && M >= 256 && N >= 256) {
    const long pair_tiles = static_cast<long>(groups) * ((M + 255) / 256) * ((N + 255) / 256);
    if (g_cta_group == 2 || pair_tiles >= 4L * (sms / 2)) {
      *bn_out = 256;
      *cg_out = 2;
      return;
    }
  }
  *cg_out = 1;
  const int m_tiles = (M + tc::BM - 1) / tc::BM;
  const int cands[4] = {256, 128, 64, 32};
  const int last = b_mn ? 2 : 3;                       // an MN-major W tile is built from 64-column atoms
  for (int i = 0; i <= last; ++i) {
    const int bn = cands[i];
    if (i < last && N <= cands[i + 1]) continue;       // tile would be mostly padding
    const long tiles = static_cast<long>(groups) * m_tiles * ((N + bn - 1) / bn);
    // measured (tools/tune_gemm.py, profiles/r1_gemm_tile_sweep.txt): the widest tile that still gives ~3/4 of the SMs a
    // tile wins -- narrower tiles re-read the 128-row A block once per n tile and become L2->SM bound
    if (tiles * 4 >= 3L * sms || i == last) {
      *bn_out = bn;
      return;
    }
  }
  *bn_out = cands[last];
}

int gemm_tile_n(int dtype, int M, int N) {
  if (dtype != TNT_BF16) return simt::TN;
  int bn, cg;
  pick_tile(M, N, 1, false, &bn, &cg);
  return bn;
}

// per-launch counter pairs of the dynamic tile scheduler: a static device array, handed out round-robin; a slot comes back
// after kCtrSlots launches, long after its previous user (same stream, or an earlier step) has reset it
constexpr int kCtrSlots = 4096;
__device__ int g_tile_ctrs[2 * kCtrSlots];
static int* next_tile_ctr() {
  static int* base[kMaxDevices] = {};
  static unsigned n = 0;
  const int dev = cur_device();
  if (!base[dev]) {
    void* p = nullptr;
    if (cudaGetSymbolAddress(&p, g_tile_ctrs) != cudaSuccess) return nullptr;
    base[dev] = static_cast<int*>(p);
  }
  return base[dev] + 2 * (n++ % kCtrSlots);
}
static int g_dynamic_tiles = -1;          // -1: TNT_GEMM_DYNAMIC from the environment (default on)

template <int EPI, typename TOut, int CG>
static int launch_tc_cg(const GemmArgs& g, int BN, cudaStream_t st) {
  CUtensorMap tmA, tmW;
  const int BNL = BN / CG;
  if (g.a_mn) TNT_TRY(make_tmap_bf16(&tmA, g.a, g.M, g.K, g.lda, g.groups, g.sa, 64));
  else TNT_TRY(make_tmap_bf16(&tmA, g.a, g.K, g.M, g.lda, g.groups, g.sa, tc::BM));
  if (g.b_mn) TNT_TRY(make_tmap_bf16(&tmW, g.w, g.N, g.K, g.ldw, g.groups, g.sw, 64));
  else TNT_TRY(make_tmap_bf16(&tmW, g.w, g.K, g.N, g.ldw, g.groups, g.sw, BNL));
  const int stage_bytes = tc::A_STAGE_BYTES + BNL * tc::BK * 2;
  int stages = std::min(tc::kMaxStages, (tc::kSmemBudget - 4096) / stage_bytes);
  if (g_force_stages > 0) stages = std::min(stages, g_force_stages);
  const int smem = stages * stage_bytes + 1024 + 512 + 2048;     // ring + alignment slack + barriers + bias staging
  const int n_tiles = (g.N + BN - 1) / BN;
  const long tiles = static_cast<long>(g.groups) * ((g.M + tc::BM * CG - 1) / (tc::BM * CG)) * n_tiles;
  const int grid = CG * static_cast<int>(std::min<long>(tiles, num_sms() / CG));
  auto kern = tc::gemm_tc_kernel<EPI, TOut, CG>;
  static PerDeviceOnce attr_once;   // per instantiation and device
  if (attr_once.first()) TNT_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tc::kSmemBudget + 4096));
  GemmArgs ga = g;
  ga.nacc = g_single_acc ? 1 : tc::acc_count(CG, BN);
  if (g_dynamic_tiles < 0) {
    const char* e = getenv("TNT_GEMM_DYNAMIC");
    g_dynamic_tiles = (e && e[0] == '0') ? 0 : 1;
  }
  // more than one tile per CTA somewhere, or fewer tiles than SMs: either way the counter decides who works
  ga.tile_ctr = (CG == 1 && g_dynamic_tiles) ? next_tile_ctr() : nullptr;
  TNT_CUDA(launch_pdl(kern, dim3(grid), dim3(tc::EpiWarps<EPI>::threads), smem, st, CG, tmA, tmW, ga, BN, stages, n_tiles));
  TNT_LAUNCHED();
  variant_hit(CG == 2 ? V_GEMM_TC_PAIR : V_GEMM_TC_1CTA);
  return TNT_OK;
}

template <int EPI, typename TOut>
static int launch_tc(const GemmArgs& g, int BN, int CG, cudaStream_t st) {
  if (CG == 2) return launch_tc_cg<EPI, TOut, 2>(g, BN, st);
  return launch_tc_cg<EPI, TOut, 1>(g, BN, st);
}

template <int EPI, typename TIn, typename TOut>
static int launch_simt(const GemmArgs& g, cudaStream_t st) {
  dim3 grid((g.N + simt::TN - 1) / simt::TN, (g.M + simt::TM - 1) / simt::TM, g.groups);
  simt::gemm_simt_kernel<EPI, TIn, TOut><<<grid, 256, 0, st>>>(g);
  TNT_LAUNCHED();
  variant_hit(V_GEMM_SIMT);
  return TNT_OK;
}

template <int EPI>
static int dispatch_dtype(const GemmArgs& g, int dt_in, int dt_out, int BN, int CG, bool force_simt, cudaStream_t st) {
  if (dt_in == TNT_BF16 && !force_simt) {
    if (dt_out == TNT_BF16) return launch_tc<EPI, bf16>(g, BN, CG, st);
    return launch_tc<EPI, float>(g, BN, CG, st);
  }
  if (dt_in == TNT_BF16) {
    if (dt_out == TNT_BF16) return launch_simt<EPI, bf16, bf16>(g, st);
    return launch_simt<EPI, bf16, float>(g, st);
  }
  if (dt_out == TNT_F32) return launch_simt<EPI, float, float>(g, st);
  return launch_simt<EPI, float, bf16>(g, st);
}

bool splitk_plan(int M, int N, int K, long persistent_tiles, int* bn_out, int* s_out);                                   // gemm_splitk.cu
int launch_gemm_splitk(const GemmArgs& g, int epi, int dt_out, int BN, int S, cudaStream_t st);

static bool g_force_simt = false;   // test hook: run bf16 problems on the SIMT kernel (cross-check)
void gemm_force_simt(bool on) { g_force_simt = on; }

int launch_gemm(const tnt_linear_params* p, int epi_override, const LseExtra* lse, int* n_tiles_out,
                cudaStream_t st) {
  TNT_CHECK_ARG(p && p->a && p->w, "linear: null operand");
  TNT_CHECK_ARG(p->M >= 0 && p->N > 0 && p->K > 0, "linear: bad shape M=%d N=%d K=%d", p->M, p->N, p->K);
  const int epi = epi_override >= 0 ? epi_override : p->epilogue;
  TNT_CHECK_ARG(epi == EPI_LSE || p->c, "linear: null output");
  TNT_CHECK_ARG(p->dtype_in == TNT_F32 || p->dtype_in == TNT_BF16, "linear: dtype_in");
  TNT_CHECK_ARG(epi != EPI_GLU || (p->N % 2 == 0), "linear: GLU needs an even N");
  TNT_CHECK_ARG((epi != EPI_RESIDUAL && epi != EPI_RELU_BWD) || p->residual, "linear: residual epilogue without residual");
  if (p->M == 0) return TNT_OK;
  GemmArgs g;
  g.tile_ctr = nullptr;
  g.lse_halves = 1;
  g.a = p->a; g.w = p->w; g.bias = p->bias; g.res = p->residual; g.c = p->c;
  g.lda = p->lda; g.ldw = p->ldw; g.ldr = p->ldr; g.ldc = p->ldc;
  g.M = p->M; g.N = p->N; g.K = p->K;
  g.m_dyn = p->m_dyn;
  g.groups = p->groups > 0 ? p->groups : 1;
  g.sa = p->stride_a; g.sw = p->stride_w; g.sbias = p->stride_bias; g.sres = p->stride_res; g.sc = p->stride_c;
  g.target = lse ? lse->target : nullptr;
  g.partials = lse ? lse->partials : nullptr;
  g.n_lse = lse ? lse->n_lse : 0;
  g.side = lse ? lse->side : nullptr;
  g.side_ld = lse ? lse->side_ld : 0;
  g.n_tiles = 0;
  TNT_CHECK_ARG(p->layout >= TNT_NT && p->layout <= TNT_TN, "linear: layout");
  g.a_mn = p->layout == TNT_TN;
  g.b_mn = p->layout != TNT_NT;
  g.alpha = p->alpha != 0.f ? p->alpha : 1.f;
  g.row_lse = lse ? lse->row_lse : nullptr;
  g.gscale = lse ? lse->gscale : 1.f;
  g.row_w = lse ? lse->row_w : nullptr;
  static const bool pf_on = [] { const char* e = getenv("TNT_L2_PREFETCH"); return !(e && e[0] == '0'); }();
  g.pf = (pf_on && p->prefetch && aligned16(p->prefetch)) ? static_cast<const char*>(p->prefetch) : nullptr;
  g.pf_bytes = g.pf ? (p->prefetch_bytes & ~15ull) : 0;
  g.drop = make_drop(p->drop);
  if (g.drop.thr) {
    TNT_CHECK_ARG(epi == EPI_RESIDUAL || epi == EPI_RELU, "linear: dropout needs the residual or ReLU epilogue");
    TNT_CHECK_ARG(p->drop.p < 1.f, "linear: dropout p = %f", p->drop.p);
    if (g.drop.ld == 0) g.drop.ld = p->N;
  }
  TNT_CHECK_ARG(epi != EPI_SMGRAD || (lse && lse->row_lse && lse->target), "smgrad: missing row statistics");
  TNT_CHECK_ARG(epi != EPI_LSE || (lse && lse->partials && (lse->n_lse >= p->N || lse->side)), "lse: missing buffers");
  const size_t eo = dtype_size(p->dtype_out);
  const long vec = 16 / static_cast<long>(eo);
  g.vec_ok = epi == EPI_LSE ? 0
                            : (aligned16(p->c) && p->ldc % vec == 0 && g.sc % vec == 0 &&
                               ((epi != EPI_RESIDUAL && epi != EPI_RELU_BWD) ||
                                (aligned16(p->residual) && p->ldr % vec == 0 && g.sres % vec == 0)));
  const bool use_tc = p->dtype_in == TNT_BF16 && !g_force_simt;
  int BN = simt::TN, CG = 1;
  if (use_tc) {
    TNT_CHECK_ALIGN(aligned16(p->a) && aligned16(p->w) && p->lda % 8 == 0 && p->ldw % 8 == 0 && g.sa % 8 == 0 &&
                        g.sw % 8 == 0,
                    "linear(bf16): A / W must be 16-byte aligned with leading dimensions multiple of 8 (lda=%ld ldw=%ld)",
                    (long)p->lda, (long)p->ldw);
    pick_tile(g.M, g.N, g.groups, g.b_mn != 0, &BN, &CG);
  }
  if (use_tc && !g_force_bn && !g.a_mn && !g.b_mn && g.groups == 1 && !g.m_dyn && !n_tiles_out &&
      (epi == EPI_NONE || epi == EPI_RELU || epi == EPI_GLU || epi == EPI_RESIDUAL)) {
    int bn_s, S;
    const long pt = static_cast<long>((g.M + tc::BM * CG - 1) / (tc::BM * CG)) * ((g.N + BN - 1) / BN);
    if (splitk_plan(g.M, g.N, g.K, pt, &bn_s, &S)) {
      variant_hit(V_GEMM_SPLITK);
      return launch_gemm_splitk(g, epi, p->dtype_out, bn_s, S, st);
    }
  }
  g.n_tiles = (g.N + BN - 1) / BN;
  // tcgen05 log-sum-exp epilogue: two partial records per tile (one per warp-half) once the tile has >= 4 column chunks;
  // 2 * ceil(N / BN) <= ceil(N / 32) records per row, the size lse_partials_bytes() reserves
  g.lse_halves = (epi == EPI_LSE && use_tc && BN >= 128) ? 2 : 1;
  g.n_tiles *= g.lse_halves;
  if (n_tiles_out) *n_tiles_out = g.n_tiles;
  switch (epi) {
    case EPI_NONE: return dispatch_dtype<EPI_NONE>(g, p->dtype_in, p->dtype_out, BN, CG, g_force_simt, st);
    case EPI_RELU: return dispatch_dtype<EPI_RELU>(g, p->dtype_in, p->dtype_out, BN, CG, g_force_simt, st);
    case EPI_GLU: return dispatch_dtype<EPI_GLU>(g, p->dtype_in, p->dtype_out, BN, CG, g_force_simt, st);
    case EPI_RESIDUAL: return dispatch_dtype<EPI_RESIDUAL>(g, p->dtype_in, p->dtype_out, BN, CG, g_force_simt, st);
    case EPI_LSE: return dispatch_dtype<EPI_LSE>(g, p->dtype_in, TNT_F32, BN, CG, g_force_simt, st);
    case EPI_RELU_BWD: return dispatch_dtype<EPI_RELU_BWD>(g, p->dtype_in, p->dtype_out, BN, CG, g_force_simt, st);
    case EPI_SMGRAD: return dispatch_dtype<EPI_SMGRAD>(g, p->dtype_in, p->dtype_out, BN, CG, g_force_simt, st);
  }
  set_error("linear: unknown epilogue %d", epi);
  return TNT_ERR_ARG;
}

}  // namespace tnt

extern "C" int tnt_linear_fwd(const tnt_linear_params* p, tnt_stream stream) {
  return tnt::launch_gemm(p, -1, nullptr, nullptr, tnt::as_stream(stream));
}
extern "C" void tnt_debug_force_simt_gemm(int on) { tnt::gemm_force_simt(on != 0); }
extern "C" void tnt_debug_gemm_cta_group(int cg) { tnt::g_cta_group = (cg == 1 || cg == 2) ? cg : 0; }
extern "C" void tnt_debug_gemm_dynamic(int on) { tnt::g_dynamic_tiles = on < 0 ? -1 : (on ? 1 : 0); }
extern "C" void tnt_debug_gemm_stages(int n) { tnt::g_force_stages = n > 0 ? n : 0; }
extern "C" void tnt_debug_gemm_single_acc(int on) { tnt::g_single_acc = on ? 1 : 0; }
extern "C" void tnt_debug_gemm_tile_n(int bn) { tnt::g_force_bn = (bn == 32 || bn == 64 || bn == 128 || bn == 256) ? bn : 0; }
