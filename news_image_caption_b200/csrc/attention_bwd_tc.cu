// tcgen05 / TMEM / TMA multi-head attention BACKWARD (kernel group 4), bf16, head dim 64, T <= 64.
// Differentiates ttl/tell/modules/attention/multi_head.py:376-477.  Inputs: q, k, v, dO, the forward's
// log-sum-exp (lse) and delta[t] = sum_d dO[t,d] O[t,d] (attn_delta kernel below).
//
// One work item = (batch row b, PAIR of heads); persistent CTAs, 10 warps (same roles as the forward kernel):
//   warp 0      TMA producer: Q and dO of the pair (all query tiles), then 128-key K|V stages
//   warp 1      one thread issues tcgen05.mma, per (key tile i, query tile j):
//                 S^T [keys, q]  = K_h . Q_h^T            dP^T[keys, q] = V_h . dO_h^T      (M=128, N=32, K=64)
//                 dV_h[keys, d] += P_h^T-tile . dO_h       dK_h[keys, d] += dS_h-tile . Q_h  (M=128, N=64, K=32)
//                 dQ^T[(h,d), (h',q)] += K^T[(h,d), keys] . dS[keys, (h',q)]                 (M=128, N=64, K=128)
//               S(n+1) is issued before the gradient MMAs of iteration n, so the tensor pipe works while the
//               softmax warps turn S(n+1), dP(n+1) into P, dS.
//   warps 2-9   lane = key: P = exp2(scale*log2e*S - lse2), dS = scale * P * (dP - delta) -> smem as bf16
//               (one [keys x (h,q)] tile serves as the K-major A operand of dV/dK and the MN-major B operand of dQ),
//               then the epilogues: dV / dK rows (contiguous 128 B per key and head) and dQ^T (coalesced channels).
// TMEM: S 64 + dP 64 + dV 128 + dK 128 + dQ 2 x 64 = 512 columns.
#include <cuda.h>

#include "backward.cuh"
#include "ptx.cuh"

namespace tnt {

int make_tmap_bf16_3d(CUtensorMap* tm, const void* base, long d0, long d1, long d2, long stride1, long stride2,
                      int box0, int box1, int box2);

namespace abt {
constexpr int kThreads = 320;
constexpr int TQ = 32, TKEYS = 128, D = 64, MAXQT = 2;
constexpr int QT_BYTES = 2 * TQ * 128;            // one query tile of two heads (Q or dO)
constexpr int KV_HALF = 2 * TKEYS * 128;          // K (or V) of two heads, 128 keys
constexpr int KV_STAGE = 2 * KV_HALF;             // K | V
constexpr int KVS = 2;
constexpr int P_BYTES = TKEYS * 128;              // [128 keys x 64 (h,q)] bf16
constexpr uint32_t COL_S = 0, COL_DP = 64, COL_DV = 128, COL_DK = 256, COL_DQ = 384;
constexpr int kListBytes = 2048 * 4 + 128;
constexpr int SMEM_BYTES = 2 * MAXQT * QT_BYTES + KVS * KV_STAGE + 2 * P_BYTES + 2048 + 1024 + 1024 + kListBytes;

struct Args {
  const uint8_t* mask; long m_stride_b;
  const float* lse; const float* delta;           // [B, H, T]
  void* dq; long q_stride_t, q_stride_b;
  void* dk; void* dv; long k_stride_s, k_stride_b;
  int B, T, S, H, mask_S;
  float scale, scale_log2e;
  int n_items, n_qt, n_tiles;
  DropSpec drop;                                  // the forward's dropout on the probabilities
  const uint32_t* tile_live;                      // optional [B]: bit i set = key tile i of batch row b holds an un-masked key
};

// sorted work list in shared memory (see attention_tc.cu): entry = batch row | (key tiles to visit) << 16; skipped tiles
// get zero dK / dV rows without being loaded
constexpr int kMaxList = 2048;
__device__ __forceinline__ void build_work_list(const uint32_t* tile_live, int B, int n_tiles, uint32_t* wl, int* hist, int lane) {
  const uint32_t all = (1u << n_tiles) - 1u;
  if (lane < 18) hist[lane] = 0;
  __syncwarp();
  for (int b = lane; b < B; b += 32) {
    uint32_t m = tile_live[b] & all;
    if (!m) m = 1u;
    atomicAdd(&hist[__popc(m)], 1);
  }
  __syncwarp();
  if (lane == 0) {
    int o = 0;
    for (int c = 17; c >= 0; --c) {
      const int n = hist[c];
      hist[c] = o;
      o += n;
    }
  }
  __syncwarp();
  for (int b = lane; b < B; b += 32) {
    uint32_t m = tile_live[b] & all;
    if (!m) m = 1u;
    wl[atomicAdd(&hist[__popc(m)], 1)] = static_cast<uint32_t>(b) | (m << 16);
  }
}
__device__ __forceinline__ uint32_t slot_entry(const uint32_t* wl, bool listed, int bi, int n_tiles) {
  return listed ? wl[bi] : (static_cast<uint32_t>(bi) | (((1u << n_tiles) - 1u) << 16));
}

struct Bars {
  uint64_t q_full, q_empty, s_ready, s_empty, p_ready, p_empty, acc_ready, acc_empty, dq_ready, dq_empty;
  uint64_t kv_full[KVS], kv_empty[KVS];
  uint32_t tmem;
};

template <bool DROP>
__global__ void __launch_bounds__(kThreads, 1)
xattn_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmDO,
                    const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV, Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                    // [n_qt][2 heads][32 q][128 B]
  uint8_t* sDO = sQ + MAXQT * QT_BYTES;
  uint8_t* sKV = sDO + MAXQT * QT_BYTES;                 // [KVS][K: 2 heads x 128 keys | V: same]
  uint8_t* sP = sKV + KVS * KV_STAGE;                    // [128 keys][(h,q) 64] bf16, 128B swizzle
  uint8_t* sDS = sP + P_BYTES;
  float* sStat = reinterpret_cast<float*>(sDS + P_BYTES);   // [2: lse2 | delta][2 heads][MAXQT*32]
  Bars* bar = reinterpret_cast<Bars*>(reinterpret_cast<uint8_t*>(sStat) + 2048);
  uint32_t* wl = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(bar) + 1024);
  int* wl_hist = reinterpret_cast<int*>(wl + kMaxList);
  const bool listed = a.tile_live != nullptr && a.B <= kMaxList;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = a.n_tiles, n_qt = a.n_qt;
  const int HP = a.H >> 1;

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmQ);
    ptx::prefetch_tmap(&tmDO);
    ptx::prefetch_tmap(&tmK);
    ptx::prefetch_tmap(&tmV);
    ptx::mbar_init(&bar->q_full, 1);
    ptx::mbar_init(&bar->q_empty, 1);
    ptx::mbar_init(&bar->s_ready, 1);
    ptx::mbar_init(&bar->s_empty, 8);
    ptx::mbar_init(&bar->p_ready, 8);
    ptx::mbar_init(&bar->p_empty, 1);
    ptx::mbar_init(&bar->acc_ready, 1);
    ptx::mbar_init(&bar->acc_empty, 8);
    ptx::mbar_init(&bar->dq_ready, 1);
    ptx::mbar_init(&bar->dq_empty, 8);
    for (int i = 0; i < KVS; ++i) { ptx::mbar_init(&bar->kv_full[i], 1); ptx::mbar_init(&bar->kv_empty[i], 1); }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc<512>(&bar->tmem);
  if (warp == 3 && listed) build_work_list(a.tile_live, a.B, a.n_tiles, wl, wl_hist, lane);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = bar->tmem;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    {   // the whole warp walks the loops, ONE elected lane issues: operands stay in uniform registers (a single divergent
        // lane paid an ELECT + 5 R2UR.BROADCAST + descriptor arithmetic in vector registers per tcgen05.mma / bulk copy)
      const bool leader = ptx::elect_one();
      int ks = 0;
      uint32_t kph = 0, it = 0;
      for (int item = blockIdx.x; item < a.n_items; item += gridDim.x, ++it) {
        const int hp = item % HP;
        const uint32_t ent = slot_entry(wl, listed, item / HP, n_tiles);
        const int b = ent & 0xFFFF;
        const uint32_t live = ent >> 16;
        const int c0 = hp * 2 * D;
        ptx::mbar_wait(&bar->q_empty, (it & 1) ^ 1);
        if (leader) ptx::mbar_arrive_expect_tx(&bar->q_full, 2 * n_qt * QT_BYTES);
        for (int j = 0; j < n_qt; ++j) {
          for (int h = 0; h < 2; ++h) {
            if (leader) ptx::tma_load_3d(sQ + j * QT_BYTES + h * TQ * 128, &tmQ, &bar->q_full, c0 + h * D, b, j * TQ);
            if (leader) ptx::tma_load_3d(sDO + j * QT_BYTES + h * TQ * 128, &tmDO, &bar->q_full, c0 + h * D, b, j * TQ);
          }
        }
        for (uint32_t m = live; m; m &= m - 1) {
          const int i = __ffs(m) - 1;
          ptx::mbar_wait(&bar->kv_empty[ks], kph ^ 1);
          if (leader) ptx::mbar_arrive_expect_tx(&bar->kv_full[ks], KV_STAGE);
          uint8_t* st = sKV + ks * KV_STAGE;
          for (int h = 0; h < 2; ++h) {
            if (leader) ptx::tma_load_3d(st + h * TKEYS * 128, &tmK, &bar->kv_full[ks], c0 + h * D, b, i * TKEYS);
            if (leader) ptx::tma_load_3d(st + KV_HALF + h * TKEYS * 128, &tmV, &bar->kv_full[ks], c0 + h * D, b, i * TKEYS);
          }
          if (++ks == KVS) { ks = 0; kph ^= 1; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    {   // the whole warp walks the loops, ONE elected lane issues: operands stay in uniform registers (a single divergent
        // lane paid an ELECT + 5 R2UR.BROADCAST + descriptor arithmetic in vector registers per tcgen05.mma / bulk copy)
      const bool leader = ptx::elect_one();
      const uint32_t idesc_s = ptx::umma_idesc_bf16(128, TQ, 0, 0);       // K-major A (keys x d), K-major B (q x d)
      const uint32_t idesc_g = ptx::umma_idesc_bf16(128, D, 0, 1);        // K-major A (keys x q), MN-major B (q x d)
      const uint32_t idesc_q = ptx::umma_idesc_bf16(128, 2 * TQ, 1, 1);   // MN-major A (keys x (h,d)), MN-major B
      int ks = 0;
      uint32_t kph = 0, it = 0, n_glob = 0, tile_glob = 0;
      auto issue_scores = [&](int kstage, int j) {
        const uint32_t k_addr = ptx::smem_u32(sKV + kstage * KV_STAGE), v_addr = k_addr + KV_HALF;
        const uint32_t q_addr = ptx::smem_u32(sQ + j * QT_BYTES), do_addr = ptx::smem_u32(sDO + j * QT_BYTES);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
#pragma unroll
          for (int k = 0; k < D / 16; ++k) {
            if (leader) ptx::umma_bf16_ss(tmem + COL_S + h * 32, ptx::umma_smem_desc(k_addr + h * TKEYS * 128 + k * 32, 16, 1024),
                              ptx::umma_smem_desc(q_addr + h * TQ * 128 + k * 32, 16, 1024), idesc_s, k != 0);
            if (leader) ptx::umma_bf16_ss(tmem + COL_DP + h * 32, ptx::umma_smem_desc(v_addr + h * TKEYS * 128 + k * 32, 16, 1024),
                              ptx::umma_smem_desc(do_addr + h * TQ * 128 + k * 32, 16, 1024), idesc_s, k != 0);
          }
        }
        if (leader) ptx::umma_commit(&bar->s_ready);
      };
      for (int item = blockIdx.x; item < a.n_items; item += gridDim.x, ++it) {
        const int n_iter = __popc(slot_entry(wl, listed, item / HP, n_tiles) >> 16) * n_qt;    // i below counts LIVE key tiles
        ptx::mbar_wait(&bar->q_full, it & 1);
        ptx::mbar_wait(&bar->dq_empty, (it & 1) ^ 1);          // previous item's dQ has been read
        // prologue: scores of iteration 0
        ptx::mbar_wait(&bar->kv_full[ks], kph);
        ptx::mbar_wait(&bar->s_empty, (n_glob & 1) ^ 1);
        ptx::tc_fence_after();
        issue_scores(ks, 0);
        for (int n = 0; n < n_iter; ++n, ++n_glob) {
          const int i = n / n_qt, j = n - i * n_qt;
          const int cur_ks = ks;
          // scores of iteration n+1 first (they only need the softmax warps to have drained S(n), dP(n))
          if (n + 1 < n_iter) {
            const int i2 = (n + 1) / n_qt, j2 = (n + 1) - i2 * n_qt;
            int nks = ks;
            if (i2 != i) {
              nks = ks + 1 == KVS ? 0 : ks + 1;
              ptx::mbar_wait(&bar->kv_full[nks], nks == 0 ? (kph ^ 1) : kph);
            }
            ptx::mbar_wait(&bar->s_empty, ((n_glob + 1) & 1) ^ 1);
            ptx::tc_fence_after();
            issue_scores(nks, j2);
          }
          // gradient MMAs of iteration n
          ptx::mbar_wait(&bar->p_ready, n_glob & 1);
          if (j == 0) ptx::mbar_wait(&bar->acc_empty, (tile_glob & 1) ^ 1);   // dV / dK of the previous key tile read
          ptx::tc_fence_after();
          const uint32_t k_addr = ptx::smem_u32(sKV + cur_ks * KV_STAGE);
          const uint32_t q_addr = ptx::smem_u32(sQ + j * QT_BYTES), do_addr = ptx::smem_u32(sDO + j * QT_BYTES);
          const uint32_t p_addr = ptx::smem_u32(sP), ds_addr = ptx::smem_u32(sDS);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
#pragma unroll
            for (int k = 0; k < TQ / 16; ++k) {
              // A: K-steps 2h+k of the [keys x 64 (h,q)] tile; B: 16 query rows x 128 B of Q_h / dO_h (MN-major)
              if (leader) ptx::umma_bf16_ss(tmem + COL_DV + h * D, ptx::umma_smem_desc(p_addr + (2 * h + k) * 32, 16, 1024),
                                ptx::umma_smem_desc(do_addr + h * TQ * 128 + k * 2048, TQ * 128, 1024), idesc_g,
                                (j | k) != 0);
              if (leader) ptx::umma_bf16_ss(tmem + COL_DK + h * D, ptx::umma_smem_desc(ds_addr + (2 * h + k) * 32, 16, 1024),
                                ptx::umma_smem_desc(q_addr + h * TQ * 128 + k * 2048, TQ * 128, 1024), idesc_g,
                                (j | k) != 0);
            }
          }
#pragma unroll
          for (int k = 0; k < TKEYS / 16; ++k) {
            if (leader) ptx::umma_bf16_ss(tmem + COL_DQ + j * 64, ptx::umma_smem_desc(k_addr + k * 2048, TKEYS * 128, 1024),
                              ptx::umma_smem_desc(ds_addr + k * 2048, 16, 1024), idesc_q, (i | k) != 0);
          }
          if (leader) ptx::umma_commit(&bar->p_empty);
          if (j == n_qt - 1) {
            if (leader) ptx::umma_commit(&bar->acc_ready);
            if (leader) ptx::umma_commit(&bar->kv_empty[cur_ks]);
            ++tile_glob;
            if (++ks == KVS) { ks = 0; kph ^= 1; }
          }
        }
        if (leader) ptx::umma_commit(&bar->dq_ready);
        if (leader) ptx::umma_commit(&bar->q_empty);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax-gradient + epilogue warps
    const int quad = warp & 3;                 // TMEM lane quadrant
    const int hd = (warp - 2) >> 2;            // head of the pair whose 32 score columns this warp owns
    const uint32_t lane_base = static_cast<uint32_t>(quad * 32) << 16;
    const int x = (warp - 2) * 32 + lane;      // 0..255
    uint32_t it = 0, n_glob = 0, tile_glob = 0;
    for (int item = blockIdx.x; item < a.n_items; item += gridDim.x, ++it) {
      const int hp = item % HP;
      const uint32_t ent = slot_entry(wl, listed, item / HP, n_tiles);
      const int b = ent & 0xFFFF;
      const uint32_t live = ent >> 16;
      // row statistics of the pair: lse2 = lse * log2(e) (+inf beyond T -> P = 0), delta
      asm volatile("bar.sync 3, 256;" ::: "memory");
      {
        const int which = x >> 7, h = (x >> 6) & 1, t = x & 63;
        float v = which == 0 ? INFINITY : 0.f;
        if (t < a.T) {
          const long idx = (static_cast<long>(b) * a.H + hp * 2 + h) * a.T + t;
          v = which == 0 ? a.lse[idx] * 1.4426950408889634f : a.delta[idx];
        }
        sStat[which * 128 + h * 64 + t] = v;
      }
      asm volatile("bar.sync 3, 256;" ::: "memory");
      const int n_iter = __popc(live) * n_qt;
      uint32_t valid = 0;             // bit of key TILE i (not of its position in the live list)
      for (uint32_t m = live; m; m &= m - 1) {
        const int i = __ffs(m) - 1;
        const int key = i * TKEYS + quad * 32 + lane;
        bool ok = key < a.S;
        if (ok && a.mask && key < a.mask_S) ok = a.mask[static_cast<long>(b) * a.m_stride_b + key] == 0;
        valid |= static_cast<uint32_t>(ok) << i;
      }
      uint32_t mleft = live;
      for (int n = 0; n < n_iter; ++n, ++n_glob) {
        const int j = n % n_qt;
        const int i = __ffs(mleft) - 1;                      // key tile of this iteration
        if (j == n_qt - 1) mleft &= mleft - 1;
        const float* lse2 = sStat + hd * 64 + j * 32;
        const float* dl = sStat + 128 + hd * 64 + j * 32;
        ptx::mbar_wait(&bar->s_ready, n_glob & 1);
        ptx::tc_fence_after();
        uint32_t sv[32], dv[32];
        ptx::tmem_ld_32x32(tmem + lane_base + COL_S + hd * 32, sv);
        ptx::tmem_ld_32x32(tmem + lane_base + COL_DP + hd * 32, dv);
        ptx::tmem_ld_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bar->s_empty);
        const bool ok = (valid >> i) & 1;
        uint32_t pk[16], dk[16];
#pragma unroll
        for (int c = 0; c < 32; c += 2) {
          float p0 = ok ? exp2f(fmaf(__uint_as_float(sv[c]), a.scale_log2e, -lse2[c])) : 0.f;
          float p1 = ok ? exp2f(fmaf(__uint_as_float(sv[c + 1]), a.scale_log2e, -lse2[c + 1])) : 0.f;
          // with dropout m (0 or 1/(1-p)): dV uses P.m, dS = scale * P * (m * dP - delta)
          float m0 = 1.f, m1 = 1.f;
          if (DROP) {
            const unsigned long long i0 =
                ((static_cast<unsigned long long>(b) * a.H + hp * 2 + hd) * a.T + j * TQ + c) * a.S + i * TKEYS + quad * 32 + lane;
            m0 = drop_val(a.drop, i0, 1.f);
            m1 = drop_val(a.drop, i0 + a.S, 1.f);
          }
          const float d0 = p0 * (m0 * __uint_as_float(dv[c]) - dl[c]) * a.scale;
          const float d1 = p1 * (m1 * __uint_as_float(dv[c + 1]) - dl[c + 1]) * a.scale;
          p0 *= m0;
          p1 *= m1;
          pk[c >> 1] = pack_bf16x2(p0, p1);
          dk[c >> 1] = pack_bf16x2(d0, d1);
        }
        ptx::mbar_wait(&bar->p_empty, (n_glob & 1) ^ 1);
        {
          const int row = quad * 32 + lane;
          uint8_t* prow = sP + row * 128;
          uint8_t* drow = sDS + row * 128;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) {
            const int chunk = (hd * 4 + jj) ^ (row & 7);           // 128-byte swizzle
            *reinterpret_cast<uint4*>(prow + chunk * 16) = make_uint4(pk[4 * jj], pk[4 * jj + 1], pk[4 * jj + 2], pk[4 * jj + 3]);
            *reinterpret_cast<uint4*>(drow + chunk * 16) = make_uint4(dk[4 * jj], dk[4 * jj + 1], dk[4 * jj + 2], dk[4 * jj + 3]);
          }
        }
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bar->p_ready);
        if (j == n_qt - 1) {
          // ---- dV / dK rows of key tile i (head hd): 64 fp32 columns each -> 128 B of bf16 per key
          ptx::mbar_wait(&bar->acc_ready, tile_glob & 1);
          ptx::tc_fence_after();
          const int key = i * TKEYS + quad * 32 + lane;
          bf16* dvo = reinterpret_cast<bf16*>(a.dv) + static_cast<long>(key) * a.k_stride_s + static_cast<long>(b) * a.k_stride_b + (hp * 2 + hd) * D;
          bf16* dko = reinterpret_cast<bf16*>(a.dk) + static_cast<long>(key) * a.k_stride_s + static_cast<long>(b) * a.k_stride_b + (hp * 2 + hd) * D;
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            uint32_t g[32];
            ptx::tmem_ld_32x32(tmem + lane_base + COL_DV + hd * D + half * 32, g);
            ptx::tmem_ld_wait();
            if (key < a.S) {
#pragma unroll
              for (int c = 0; c < 32; c += 8)
                *reinterpret_cast<uint4*>(dvo + half * 32 + c) =
                    make_uint4(pack_bf16x2(__uint_as_float(g[c]), __uint_as_float(g[c + 1])), pack_bf16x2(__uint_as_float(g[c + 2]), __uint_as_float(g[c + 3])),
                               pack_bf16x2(__uint_as_float(g[c + 4]), __uint_as_float(g[c + 5])), pack_bf16x2(__uint_as_float(g[c + 6]), __uint_as_float(g[c + 7])));
            }
            ptx::tmem_ld_32x32(tmem + lane_base + COL_DK + hd * D + half * 32, g);
            ptx::tmem_ld_wait();
            if (key < a.S) {
#pragma unroll
              for (int c = 0; c < 32; c += 8)
                *reinterpret_cast<uint4*>(dko + half * 32 + c) =
                    make_uint4(pack_bf16x2(__uint_as_float(g[c]), __uint_as_float(g[c + 1])), pack_bf16x2(__uint_as_float(g[c + 2]), __uint_as_float(g[c + 3])),
                               pack_bf16x2(__uint_as_float(g[c + 4]), __uint_as_float(g[c + 5])), pack_bf16x2(__uint_as_float(g[c + 6]), __uint_as_float(g[c + 7])));
            }
          }
          ptx::tc_fence_before();
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&bar->acc_empty);
          ++tile_glob;
        }
      }
      // ---- key tiles that were skipped (all padding): their dK / dV rows are exactly zero
      for (uint32_t m = ((1u << n_tiles) - 1u) & ~live; m; m &= m - 1) {
        const int key = (__ffs(m) - 1) * TKEYS + quad * 32 + lane;
        if (key < a.S) {
          const long o = static_cast<long>(key) * a.k_stride_s + static_cast<long>(b) * a.k_stride_b + (hp * 2 + hd) * D;
#pragma unroll
          for (int c = 0; c < D; c += 8) {
            *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(a.dv) + o + c) = make_uint4(0, 0, 0, 0);
            *reinterpret_cast<uint4*>(reinterpret_cast<bf16*>(a.dk) + o + c) = make_uint4(0, 0, 0, 0);
          }
        }
      }
      // ---- dQ^T: rows (h, d) of this quadrant belong to head (quad / 2); this warp takes 16 queries per tile
      ptx::mbar_wait(&bar->dq_ready, it & 1);
      ptx::tc_fence_after();
      const int oh = quad >> 1, q0 = hd * 16;
      const int dch = (quad & 1) * 32 + lane;
      bf16* dqo = reinterpret_cast<bf16*>(a.dq) + static_cast<long>(b) * a.q_stride_b + (hp * 2 + oh) * D + dch;
      for (int j = 0; j < n_qt; ++j) {
        uint32_t o[16];
        ptx::tmem_ld_32x16(tmem + lane_base + COL_DQ + j * 64 + oh * 32 + q0, o);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int jj = 0; jj < 16; ++jj) {
          const int t = j * TQ + q0 + jj;
          if (t < a.T) dqo[static_cast<long>(t) * a.q_stride_t] = __float2bfloat16_rn(__uint_as_float(o[jj]));
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bar->dq_empty);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<512>(tmem);
}

// delta[b, h, t] = sum_d dO[t, b, h*D + d] * O[t, b, h*D + d]: 8 lanes per (t, b, h), one 16-byte load of each tensor per lane
// (D = 64; the 4-bytes-per-lane, warp-per-row form took 13 us for 8 MB)
__global__ void __launch_bounds__(256) attn_delta_kernel(const bf16* o, const bf16* dout, long stride_t, long stride_b, int B,
                                                         int T, int H, int Dh, float* delta) {
  pdl_wait();
  const int sub = threadIdx.x & 7;
  const long w = (blockIdx.x * 256L + threadIdx.x) >> 3;
  const long total = static_cast<long>(B) * T * H;
  float s = 0.f;
  int h = 0, t = 0, b = 0;
  if (w < total) {
    h = static_cast<int>(w % H);
    t = static_cast<int>((w / H) % T);
    b = static_cast<int>(w / (static_cast<long>(H) * T));
    const long base = static_cast<long>(t) * stride_t + static_cast<long>(b) * stride_b + h * Dh;
    for (int d = sub * 8; d < Dh; d += 64) {
      float x[8], y[8];
      load8(o + base + d, x);
      load8(dout + base + d, y);
#pragma unroll
      for (int j = 0; j < 8; ++j) s = fmaf(x[j], y[j], s);
    }
  }
  s += __shfl_xor_sync(0xffffffffu, s, 1);
  s += __shfl_xor_sync(0xffffffffu, s, 2);
  s += __shfl_xor_sync(0xffffffffu, s, 4);
  if (sub == 0 && w < total) delta[(static_cast<long>(b) * H + h) * T + t] = s;
}
}  // namespace abt

bool xattn_bwd_tc_supported(const XattnBwd* p) {
  const tnt_xattn_params& f = p->f;
  if (f.dtype != TNT_BF16 || f.D != 64 || (f.H & 1) || f.n_extra != 0 || !f.lse) return false;
  if (f.mask_mode != TNT_MASK_NONE && f.mask_mode != TNT_MASK_KEY_PADDING) return false;
  if (f.T < 1 || f.T > abt::MAXQT * abt::TQ || f.S < 1 || f.S > 32 * abt::TKEYS) return false;
  const long strides = f.q_stride_t | f.q_stride_b | f.k_stride_s | f.k_stride_b | f.v_stride_s | f.v_stride_b |
                       f.o_stride_t | f.o_stride_b;
  if (f.v_stride_s != f.k_stride_s || f.v_stride_b != f.k_stride_b) return false;   // dk / dv share one stride pair
  return (strides % 8) == 0 && aligned16(f.q) && aligned16(f.k) && aligned16(f.v) && aligned16(f.out) && aligned16(p->dout) &&
         aligned16(p->dq) && aligned16(p->dk) && aligned16(p->dv);
}

// `scratch` (fp32 [B, H, T, D]) provides the delta buffer [B, H, T]
int xattn_bwd_tc(const XattnBwd* p, cudaStream_t st) {
  const tnt_xattn_params& f = p->f;
  const int E = f.H * f.D;
  {
    const long rows = static_cast<long>(f.B) * f.T * f.H;
    TNT_CUDA(launch_pdl(abt::attn_delta_kernel, dim3(static_cast<unsigned>((rows * 8 + 255) / 256)), dim3(256), 0, st, 1,
                        static_cast<const bf16*>(f.out), static_cast<const bf16*>(p->dout), static_cast<long>(f.o_stride_t),
                        static_cast<long>(f.o_stride_b), f.B, f.T, f.H, f.D, p->scratch));
    TNT_LAUNCHED();
  }
  CUtensorMap tmQ, tmDO, tmK, tmV;
  TNT_TRY(make_tmap_bf16_3d(&tmQ, f.q, E, f.B, f.T, f.q_stride_b, f.q_stride_t, 64, 1, abt::TQ));
  TNT_TRY(make_tmap_bf16_3d(&tmDO, p->dout, E, f.B, f.T, f.o_stride_b, f.o_stride_t, 64, 1, abt::TQ));
  TNT_TRY(make_tmap_bf16_3d(&tmK, f.k, E, f.B, f.S, f.k_stride_b, f.k_stride_s, 64, 1, abt::TKEYS));
  TNT_TRY(make_tmap_bf16_3d(&tmV, f.v, E, f.B, f.S, f.v_stride_b, f.v_stride_s, 64, 1, abt::TKEYS));
  abt::Args a;
  a.mask = f.mask_mode == TNT_MASK_KEY_PADDING ? f.mask : nullptr;
  a.m_stride_b = f.m_stride_b;
  a.lse = f.lse; a.delta = p->scratch;
  a.drop = make_drop(f.drop);
  a.tile_live = f.tile_live;
  a.dq = p->dq; a.q_stride_t = f.q_stride_t; a.q_stride_b = f.q_stride_b;
  a.dk = p->dk; a.dv = p->dv; a.k_stride_s = f.k_stride_s; a.k_stride_b = f.k_stride_b;
  a.B = f.B; a.T = f.T; a.S = f.S; a.H = f.H;
  a.mask_S = f.mask_S > 0 ? f.mask_S : f.S;
  a.scale = f.scale;
  a.scale_log2e = f.scale * 1.4426950408889634f;
  a.n_qt = (f.T + abt::TQ - 1) / abt::TQ;
  a.n_items = f.B * (f.H / 2);
  a.n_tiles = (f.S + abt::TKEYS - 1) / abt::TKEYS;
  static PerDeviceOnce attr_once;
  if (attr_once.first()) {
    TNT_CUDA(cudaFuncSetAttribute(abt::xattn_bwd_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, abt::SMEM_BYTES));
    TNT_CUDA(cudaFuncSetAttribute(abt::xattn_bwd_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, abt::SMEM_BYTES));
  }
  const int grid = std::min(a.n_items, num_sms());
  if (a.drop.thr) abt::xattn_bwd_tc_kernel<true><<<grid, abt::kThreads, abt::SMEM_BYTES, st>>>(tmQ, tmDO, tmK, tmV, a);
  else abt::xattn_bwd_tc_kernel<false><<<grid, abt::kThreads, abt::SMEM_BYTES, st>>>(tmQ, tmDO, tmK, tmV, a);
  TNT_LAUNCHED();
  return TNT_OK;
}

}  // namespace tnt
