// tcgen05 / TMEM / TMA multi-head attention forward (kernel group 2), bf16 inputs, head dim 64.
// Reference math: ttl/tell/modules/attention/multi_head.py:376-477 (scaled QK^T, key-padding mask, fp32
// softmax, PV).  The extra bias_k / zero rows are ordinary k/v rows here (the engine materialises them
// once per batch, engine_fwd.cu), the mask covers the first `mask_S` keys only.
//
// One work item = (batch row b, PAIR of heads, tile of <= 32 queries); persistent CTAs, 18 warps:
//   warp 0      TMA producer: Q pair, then 128-key K tiles and V tiles (128B swizzle) through mbarrier rings
//   warp 1      one thread issues tcgen05.mma:
//                 S^T[keys, q] = K_h[keys, 64] . Q_h[q, 64]^T   (M=128 keys, N=32 queries, per head)
//                 O^T[(h,d), (h',q)] += V[(h,d), keys] . P[keys, (h',q)]   (M=128 = 2 heads x 64 dims, N=64)
//               -> every score tile of the item stays in TMEM (<= 7 tiles x 64 columns), O^T in 64 more
//   warps 2-17  softmax (the serial chain of the kernel, so the 32 queries of a head are split over TWO warps of 16
//               columns each: 4 lane quadrants x 2 heads x 2 column halves): lane = key, exact two-pass softmax in fp32 straight out of TMEM (column max with
//               redux.sync + one cross-warp exchange), P written to smem as the bf16 MN-major B operand,
//               epilogue: O^T * 1/l -> global (lanes = consecutive channels, coalesced)
// Zero padding waste in M for the score MMAs (keys fill M), 2x waste (cross-head blocks) in the PV MMAs.
#include <cuda.h>

#include "common.cuh"
#include "ptx.cuh"

namespace tnt {

int make_tmap_bf16_3d(CUtensorMap* tm, const void* base, long d0, long d1, long d2, long stride1, long stride2,
                      int box0, int box1, int box2);

namespace atc {
constexpr int kSoftWarps = 16;                    // 4 TMEM lane quadrants x 2 heads x 2 halves of the 32 query columns
constexpr int kThreads = (2 + kSoftWarps) * 32;
constexpr int CW = 16;                            // score columns (queries) per softmax warp
constexpr int TQ = 32, TKEYS = 128, D = 64;
// K and V tiles share ONE ring, filled in consumption order (the item's K tiles, then its V tiles): a K tile is dead as soon
// as its score product has been issued, so the item's V tiles -- and, behind the PV products, the NEXT item's K tiles --
// stream in while the softmax warps work; with separate 3 + 2 stage rings every round of loads sat in their serial chain
constexpr int KVS = 5, PS = 2;
constexpr int Q_BYTES = 2 * TQ * 128;             // two heads
constexpr int KV_STAGE = 2 * TKEYS * 128;         // two heads x 128 keys x 128 B
constexpr int P_STAGE = TKEYS * 128;              // 128 keys x 64 (h,q) bf16
constexpr int MAX_TILES = 7;
constexpr uint32_t O_COL = 448;
constexpr int kListBytes = 2048 * 4 + 128 + 1024;  // sorted work list + histogram (+ the alignment slack of the base)
constexpr int SMEM_BYTES = Q_BYTES + KVS * KV_STAGE + PS * P_STAGE + 4096 + 2048 + kListBytes;

struct Args {
  float* lse; float scale;
  const uint8_t* mask; long m_stride_b;
  void* out; long o_stride_t, o_stride_b;
  int B, T, S, H, mask_S;
  float scale_log2e;
  int n_items, n_qt, n_tiles;
  uint32_t mul_qt, mul_hp;                        // floor(2^32 / d) + 1 for d = n_qt, head slots: item -> (slot, hp, qt) without IDIV
  DropSpec drop;                                  // dropout on the probabilities (the normaliser stays un-dropped)
  const uint32_t* tile_live;                      // optional [B]: bit i set = key tile i of batch row b holds an un-masked key
  float* weights;                                 // optional [B, T, S] fp32, ZEROED by the launcher: head-averaged probabilities
  float inv_heads;                                //   (multi_head.py:478-482), accumulated with red.global.add by a third pass
  long long* trace;                               // debug (tnt_debug_xattn_trace): clock64 stamps of CTAs 0..3, [cta][role 0..2][item 0..7][8]
  long m_stride_t;                                // D128 / V2 mask: byte stride over the query index (0 = one mask row per batch entry)
};

// Work list in shared memory: entry = batch row | (key tiles to visit) << 16.  Tiles whose keys are all padding contribute
// exactly 0 and are skipped.  Every CTA sorts the B entries by live-tile count (longest first, counting sort by one warp
// in the kernel prologue) so that the static item striding hands every CTA a similar amount of work.
constexpr int kMaxList = 2048;
__device__ __forceinline__ void build_work_list(const uint32_t* tile_live, int B, int n_tiles, uint32_t* wl, int* hist, int lane) {
  const uint32_t all = (1u << n_tiles) - 1u;
  if (lane < 18) hist[lane] = 0;
  __syncwarp();
  for (int b = lane; b < B; b += 32) {
    uint32_t m = tile_live[b] & all;
    if (!m) m = 1u;                               // a fully masked row still walks one tile (P = 0 -> out = 0)
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
// entry of batch slot bi: from the sorted list, or (bi, every tile) when there is none
__device__ __forceinline__ uint32_t slot_entry(const uint32_t* wl, bool listed, int bi, int n_tiles) {
  return listed ? wl[bi] : (static_cast<uint32_t>(bi) | (((1u << n_tiles) - 1u) << 16));
}

// x / d for x * d < 2^32 (checked by the launcher) with mul = floor(2^32 / d) + 1; d == 1 has no 32-bit multiplier
__device__ __forceinline__ int fast_div(int x, int d, uint32_t mul) { return d == 1 ? x : static_cast<int>(__umulhi(static_cast<uint32_t>(x), mul)); }
struct ItemPos { int qt, hp, slot; };
__device__ __forceinline__ ItemPos item_pos(int item, int n_qt, uint32_t mul_qt, int HP, uint32_t mul_hp) {
  ItemPos p;
  const int r = fast_div(item, n_qt, mul_qt);
  p.qt = item - r * n_qt;
  p.slot = fast_div(r, HP, mul_hp);
  p.hp = r - p.slot * HP;
  return p;
}

#define XTRACE(role, slot)                                                                          \
  do {                                                                                              \
    if (a.trace && blockIdx.x < 4 && it < 8 && lane == 0)                                           \
      a.trace[((blockIdx.x * 3 + (role)) * 8 + it) * 8 + (slot)] = clock64();                       \
  } while (0)

struct Bars {
  uint64_t q_full, q_empty, s_empty, o_ready, o_empty;
  uint64_t s_tile[MAX_TILES];       // score tile i of the current item has been written (one phase per item that HAS a tile i)
  uint64_t kv_full[KVS], kv_empty[KVS], p_ready[PS], p_empty[PS];
  uint32_t tmem;
};

__device__ __forceinline__ float redux_max(float v) {
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}
// 2^x for x <= 0 (a probability relative to the column maximum): one MUFU instruction, results below 2^-126 flush to 0
__device__ __forceinline__ float ex2_fast(float x) {
  float r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}
__device__ __forceinline__ void named_bar(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }

// DROP: dropout on the probabilities, compiled in only for the training step (zero cost for inference).
// D128: ONE head of 128 dims per item instead of a pair of 64-dim heads -- the `final*` model (8 heads of 128,
// final_roberta2/transformer_v2.py:224-266).  In memory a 128-dim head is the same two 64-column boxes as a head pair, so the
// TMA traffic, the V^T operand of the PV product (two 64-channel atoms) and the O^T epilogue carry over; the score product
// accumulates both 64-column halves into ONE 32-query accumulator (8 k-steps), only the first four softmax warps have
// scores to work on (the other four keep their half of P at zero and join the barriers), and the mask follows
// attention() of transformer_v2.py:213-222: mask[b, t, s] == 0 -> the score is REPLACED by -1e9 (a fully masked row
// becomes uniform over the S keys, it is not dropped), one validity bit per (key, query).
template <bool DROP, bool D128>
__global__ void __launch_bounds__(kThreads, 1)
xattn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                const __grid_constant__ CUtensorMap tmV, Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sKV = sQ + Q_BYTES;
  uint8_t* sP = sKV + KVS * KV_STAGE;
  float* red0 = reinterpret_cast<float*>(sP + PS * P_STAGE);         // [item parity][max, sum][2 heads][4 quadrants][32]
  Bars* bar = reinterpret_cast<Bars*>(reinterpret_cast<uint8_t*>(red0) + 4096);
  uint32_t* wl = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(bar) + 2048);
  int* wl_hist = reinterpret_cast<int*>(wl + kMaxList);
  const bool listed = a.tile_live != nullptr && a.B <= kMaxList;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    ptx::prefetch_tmap(&tmQ);
    ptx::prefetch_tmap(&tmK);
    ptx::prefetch_tmap(&tmV);
    ptx::mbar_init(&bar->q_full, 1);
    ptx::mbar_init(&bar->q_empty, 1);
    for (int i = 0; i < MAX_TILES; ++i) ptx::mbar_init(&bar->s_tile[i], 1);
    ptx::mbar_init(&bar->s_empty, kSoftWarps);
    ptx::mbar_init(&bar->o_ready, 1);
    ptx::mbar_init(&bar->o_empty, kSoftWarps);
    for (int i = 0; i < KVS; ++i) { ptx::mbar_init(&bar->kv_full[i], 1); ptx::mbar_init(&bar->kv_empty[i], 1); }
    for (int i = 0; i < PS; ++i) { ptx::mbar_init(&bar->p_ready[i], kSoftWarps); ptx::mbar_init(&bar->p_empty[i], 1); }
    ptx::fence_barrier_init();
  }
  if (warp == 2) ptx::tmem_alloc<512>(&bar->tmem);
  // tile_live was written several launches earlier; every kernel of the stream waits for its predecessor before it
  // completes, so the data is final although this prologue runs ahead of pdl_wait()
  if (warp == 3 && listed) build_work_list(a.tile_live, a.B, a.n_tiles, wl, wl_hist, lane);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = bar->tmem;
  const int HP = D128 ? a.H : a.H >> 1;      // head slots per batch entry: single 128-dim heads or pairs of 64-dim heads
  if (D128) {                        // the second head's half of every P stage stays zero for the whole kernel
    for (int i = threadIdx.x; i < PS * TKEYS * 4; i += kThreads) {
      const int st = i / (TKEYS * 4), r = (i / 4) % TKEYS, j = i & 3;
      *reinterpret_cast<uint4*>(sP + st * P_STAGE + r * 128 + ((4 + j) ^ (r & 7)) * 16) = make_uint4(0, 0, 0, 0);
    }
    ptx::fence_proxy_async();
    __syncthreads();
  }
  pdl_wait();                       // prologue above overlaps the preceding kernel's tail (PDL)

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    {   // the whole warp walks the loops, ONE elected lane issues: operands stay in uniform registers (a single divergent
        // lane paid an ELECT + 5 R2UR.BROADCAST + descriptor arithmetic in vector registers per tcgen05.mma / bulk copy)
      const bool leader = ptx::elect_one();
      int ks = 0;
      uint32_t kph = 0, it = 0;
      for (int item = blockIdx.x; item < a.n_items; item += gridDim.x, ++it) {
        const ItemPos ip = item_pos(item, a.n_qt, a.mul_qt, HP, a.mul_hp);
        const int qt = ip.qt, hp = ip.hp;
        const uint32_t ent = slot_entry(wl, listed, ip.slot, a.n_tiles);
        const int b = ent & 0xFFFF;
        const uint32_t live = ent >> 16;
        const int c0 = hp * 2 * D;
        ptx::mbar_wait(&bar->q_empty, (it & 1) ^ 1);
        XTRACE(2, 0);
        if (leader) ptx::mbar_arrive_expect_tx(&bar->q_full, Q_BYTES);
        if (leader) ptx::tma_load_3d(sQ, &tmQ, &bar->q_full, c0, b, qt * TQ);
        if (leader) ptx::tma_load_3d(sQ + TQ * 128, &tmQ, &bar->q_full, c0 + D, b, qt * TQ);
        for (uint32_t m = live; m; m &= m - 1) {
          const int i = __ffs(m) - 1;
          ptx::mbar_wait(&bar->kv_empty[ks], kph ^ 1);
          if (leader) ptx::mbar_arrive_expect_tx(&bar->kv_full[ks], KV_STAGE);
          if (leader) ptx::tma_load_3d(sKV + ks * KV_STAGE, &tmK, &bar->kv_full[ks], c0, b, i * TKEYS);
          if (leader) ptx::tma_load_3d(sKV + ks * KV_STAGE + TKEYS * 128, &tmK, &bar->kv_full[ks], c0 + D, b, i * TKEYS);
          if (++ks == KVS) { ks = 0; kph ^= 1; }
        }
        for (uint32_t m = live; m; m &= m - 1) {
          const int i = __ffs(m) - 1;
          ptx::mbar_wait(&bar->kv_empty[ks], kph ^ 1);
          if (leader) ptx::mbar_arrive_expect_tx(&bar->kv_full[ks], KV_STAGE);
          if (leader) ptx::tma_load_3d(sKV + ks * KV_STAGE, &tmV, &bar->kv_full[ks], c0, b, i * TKEYS);
          if (leader) ptx::tma_load_3d(sKV + ks * KV_STAGE + TKEYS * 128, &tmV, &bar->kv_full[ks], c0 + D, b, i * TKEYS);
          if (++ks == KVS) { ks = 0; kph ^= 1; }
          XTRACE(2, 1 + min(i, 6));
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    {   // the whole warp walks the loops, ONE elected lane issues: operands stay in uniform registers (a single divergent
        // lane paid an ELECT + 5 R2UR.BROADCAST + descriptor arithmetic in vector registers per tcgen05.mma / bulk copy)
      const bool leader = ptx::elect_one();
      const uint32_t idesc_qk = ptx::umma_idesc_bf16(128, TQ, 0, 0);
      const uint32_t idesc_pv = ptx::umma_idesc_bf16(128, 2 * TQ, 1, 1);
      int ks = 0, ps = 0;
      uint32_t kph = 0, pph = 0, it = 0;
      for (int item = blockIdx.x; item < a.n_items; item += gridDim.x, ++it) {
        const uint32_t par = it & 1;
        const int nl = __popc(slot_entry(wl, listed, item_pos(item, a.n_qt, a.mul_qt, HP, a.mul_hp).slot, a.n_tiles) >> 16);   // live key tiles: score tile j at TMEM column j*64
        XTRACE(1, 0);
        ptx::mbar_wait(&bar->q_full, par);
        XTRACE(1, 1);
        ptx::mbar_wait(&bar->s_empty, par ^ 1);            // previous item's scores have been consumed
        ptx::tc_fence_after();
        XTRACE(1, 2);
        const uint32_t q_addr = ptx::smem_u32(sQ);
        for (int i = 0; i < nl; ++i) {
          ptx::mbar_wait(&bar->kv_full[ks], kph);
          ptx::tc_fence_after();
          const uint32_t k_addr = ptx::smem_u32(sKV + ks * KV_STAGE);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
#pragma unroll
            for (int k = 0; k < D / 16; ++k) {
              const uint64_t ad = ptx::umma_smem_desc(k_addr + h * TKEYS * 128 + k * 32, 16, 1024);
              const uint64_t bd = ptx::umma_smem_desc(q_addr + h * TQ * 128 + k * 32, 16, 1024);
              // pair of heads: one accumulator per head; one 128-dim head: both 64-column halves into the same accumulator
              if (leader) ptx::umma_bf16_ss(tmem + i * 64 + (D128 ? 0 : h * 32), ad, bd, idesc_qk, D128 ? (h | k) != 0 : k != 0);
            }
          }
          if (leader) ptx::umma_commit(&bar->kv_empty[ks]);
          if (leader) ptx::umma_commit(&bar->s_tile[i]);     // pass 1 starts on tile 0 while the later K tiles are still in flight
          if (++ks == KVS) { ks = 0; kph ^= 1; }
        }
        if (leader) ptx::umma_commit(&bar->q_empty);
        XTRACE(1, 3);
        ptx::mbar_wait(&bar->o_empty, par ^ 1);            // previous item's output has been read
        ptx::tc_fence_after();
        XTRACE(1, 4);
        for (int i = 0; i < nl; ++i) {
          ptx::mbar_wait(&bar->kv_full[ks], kph);
          if (i == 0) XTRACE(1, 5);
          ptx::mbar_wait(&bar->p_ready[ps], pph);
          ptx::tc_fence_after();
          if (i == 0) XTRACE(1, 6);
          const uint32_t v_addr = ptx::smem_u32(sKV + ks * KV_STAGE);
          const uint32_t p_addr = ptx::smem_u32(sP + ps * P_STAGE);
#pragma unroll
          for (int k = 0; k < TKEYS / 16; ++k) {
            // A = V^T: MN-major, two 64-channel atoms (the two heads) TKEYS*128 B apart, 8-key groups 1024 B apart
            const uint64_t ad = ptx::umma_smem_desc(v_addr + k * 2048, TKEYS * 128, 1024);
            const uint64_t bd = ptx::umma_smem_desc(p_addr + k * 2048, 16, 1024);
            if (leader) ptx::umma_bf16_ss(tmem + O_COL, ad, bd, idesc_pv, (i | k) != 0);
          }
          if (leader) ptx::umma_commit(&bar->kv_empty[ks]);
          if (leader) ptx::umma_commit(&bar->p_empty[ps]);
          if (++ks == KVS) { ks = 0; kph ^= 1; }
          if (++ps == PS) { ps = 0; pph ^= 1; }
        }
        if (leader) ptx::umma_commit(&bar->o_ready);
        XTRACE(1, 7);
      }
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ softmax + epilogue warps
    const int quad = warp & 3;                 // TMEM lane quadrant
    const int grp = (warp - 2) >> 2;           // 0..3
    const int hd = grp >> 1;                   // which head's score columns this warp works on
    const int half = grp & 1;                  //   and which 16 of its 32 queries
    const int cb = half * CW;                  // first query column of this warp inside the head's 32
    const int l16 = lane & (CW - 1);
    const uint32_t lane_base = static_cast<uint32_t>(quad * 32) << 16;
    // red[] is double buffered by item parity: a warp that runs ahead into the next item writes the OTHER copy, and cannot
    // reach the item after that before every warp has passed this item's barriers (no end-of-item barrier needed)
    int ps = 0;
    uint32_t pph = 0, it = 0;
    uint32_t sph = 0;                          // bit i: parity of the next phase of score-tile barrier i
    // key-padding bytes of this lane's key in every live tile of an item (head pairs): fetched one item AHEAD, behind the
    // previous item's pass 2, so the global-load latency is not part of the softmax warps' serial chain (15 % of their
    // stall samples when the bytes were loaded at the top of the item)
    uint32_t mb[MAX_TILES];
    auto fetch_mask = [&](int item_n) {
      if (D128) return;
#pragma unroll
      for (int j = 0; j < MAX_TILES; ++j) mb[j] = 1;                 // 0 = this lane's key of live tile j takes part
      if (item_n >= a.n_items) return;
      const uint32_t ent_n = slot_entry(wl, listed, item_pos(item_n, a.n_qt, a.mul_qt, HP, a.mul_hp).slot, a.n_tiles);
      const uint8_t* mrow = a.mask + static_cast<long>(ent_n & 0xFFFF) * a.m_stride_b;
      const int lim = a.mask ? min(a.S, a.mask_S) : 0;                // keys below it have a mask byte
      uint32_t m = ent_n >> 16;
#pragma unroll
      for (int j = 0; j < MAX_TILES; ++j) {
        if (m) {
          const int key = (__ffs(m) - 1) * TKEYS + quad * 32 + lane;
          if (key < lim) mb[j] = mrow[key];
          else if (key < a.S) mb[j] = 0;
          m &= m - 1;
        }
      }
    };
    fetch_mask(blockIdx.x);
    for (int item = blockIdx.x; item < a.n_items; item += gridDim.x, ++it) {
      const uint32_t par = it & 1;
      float* red = red0 + par * 512;
      float* red_max = red + (hd * 4 + quad) * 32;
      float* red_sum = red + 256 + (hd * 4 + quad) * 32;
      if (warp == 2) XTRACE(0, 0);
      const ItemPos ip = item_pos(item, a.n_qt, a.mul_qt, HP, a.mul_hp);
      const int qt = ip.qt, hp = ip.hp;
      const uint32_t ent = slot_entry(wl, listed, ip.slot, a.n_tiles);
      const int b = ent & 0xFFFF;
      const uint32_t live = ent >> 16;
      // validity of this lane's key in every tile (key-padding mask: multi_head.py:441-449)
      const int nl = __popc(live);
      const bool score_warp = !D128 || hd == 0;      // D128: the second group of four warps has no score columns
      uint32_t valid = 0;             // bit j: this lane's key of the j-th LIVE tile takes part (D128: lies inside [0, S))
      uint32_t vis[D128 ? MAX_TILES : 1];            // D128: bit c of vis[j] = query c of this tile SEES the key (mask != 0)
      if (!D128) {
#pragma unroll
        for (int j = 0; j < MAX_TILES; ++j) valid |= static_cast<uint32_t>(mb[j] == 0) << j;
      } else {
        int j = 0;
        for (uint32_t m = live; m; m &= m - 1, ++j) {
          const int key = (__ffs(m) - 1) * TKEYS + quad * 32 + lane;
          bool ok = key < a.S;
          if (D128) {
            uint32_t w = 0xFFFFFFFFu;
            if (ok && a.mask && score_warp) {
              const uint8_t* mp = a.mask + static_cast<long>(b) * a.m_stride_b + key;
              if (a.m_stride_t == 0) {
                w = mp[0] ? 0xFFFFFFFFu : 0u;
              } else {
                w = 0;
                for (int c = 0; c < TQ; ++c) {
                  const int t = min(qt * TQ + c, a.T - 1);
                  w |= static_cast<uint32_t>(mp[static_cast<long>(t) * a.m_stride_t] != 0) << c;
                }
              }
            }
            vis[j] = w;
          }
          valid |= static_cast<uint32_t>(ok) << j;
        }
      }
      const float neg_raw = -1e9f / a.scale;         // raw-score value whose scaled form is the reference's -1e9 fill
      if (warp == 2) XTRACE(0, 1);
      // ---- pass 1: column maxima over the keys
      float mx[CW];
#pragma unroll
      for (int c = 0; c < CW; ++c) mx[c] = -INFINITY;
      for (int i = 0; i < nl && score_warp; ++i) {
        ptx::mbar_wait(&bar->s_tile[i], (sph >> i) & 1);
        ptx::tc_fence_after();
        if (warp == 2 && i == 0) XTRACE(0, 2);
        uint32_t v[CW];
        ptx::tmem_ld_32x16(tmem + lane_base + i * 64 + hd * 32 + cb, v);
        ptx::tmem_ld_wait();
        if ((valid >> i) & 1) {
#pragma unroll
          for (int c = 0; c < CW; ++c)
            mx[c] = fmaxf(mx[c], (!D128 || ((vis[D128 ? i : 0] >> (cb + c)) & 1)) ? __uint_as_float(v[c]) : neg_raw);
        }
      }
      sph ^= (1u << nl) - 1u;
      float mine = -INFINITY;
#pragma unroll
      for (int c = 0; c < CW; ++c) {
        const float w = redux_max(mx[c]);
        if (lane == c) mine = w;
      }
      if (lane < CW) red_max[cb + lane] = mine;
      if (warp == 2) XTRACE(0, 3);
      named_bar(1 + hd, 256);            // the 8 warps of this head
      float col_max;                     // max of the unscaled scores of column (query) cb + (lane & 15)
      {
        const float* rm = red + hd * 128 + cb + l16;
        const float m4 = fmaxf(fmaxf(rm[0], rm[32]), fmaxf(rm[64], rm[96]));
        const float ms = (m4 == -INFINITY) ? 0.f : m4;
        col_max = ms;
#pragma unroll
        for (int c = 0; c < CW; ++c) mx[c] = __shfl_sync(0xffffffffu, ms, c) * a.scale_log2e;
      }
      // ---- pass 2: probabilities -> smem (bf16, MN-major B operand), running column sums
      float ls[CW];
#pragma unroll
      for (int c = 0; c < CW; ++c) ls[c] = 0.f;
      uint32_t mleft = live;
      for (int i = 0; i < nl; ++i, mleft &= mleft - 1) {
        const int tile = __ffs(mleft) - 1;                    // key tile the i-th score tile belongs to
        if (!score_warp) {                                    // D128: nothing to write (this half of P is zero); keep the ring in step
          ptx::mbar_wait(&bar->p_empty[ps], pph ^ 1);
          __syncwarp();
          if (lane == 0) ptx::mbar_arrive(&bar->p_ready[ps]);
          if (++ps == PS) { ps = 0; pph ^= 1; }
          continue;
        }
        uint32_t v[CW];
        ptx::tmem_ld_32x16(tmem + lane_base + i * 64 + hd * 32 + cb, v);
        ptx::tmem_ld_wait();
        const bool ok = (valid >> i) & 1;
        if (D128) {                                            // masked (query, key) pairs carry the -1e9 fill, not "absent"
#pragma unroll
          for (int c = 0; c < CW; ++c)
            if (!((vis[D128 ? i : 0] >> (cb + c)) & 1)) v[c] = __float_as_uint(neg_raw);
        }
        uint32_t pk[CW / 2];
#pragma unroll
        for (int c = 0; c < CW / 2; ++c) pk[c] = 0u;
        if (ok)                                    // (a key outside the row contributes nothing: only a ragged tile diverges)
#pragma unroll
        for (int c = 0; c < CW; c += 2) {
          float p0 = ex2_fast(fmaf(__uint_as_float(v[c]), a.scale_log2e, -mx[c]));
          float p1 = ex2_fast(fmaf(__uint_as_float(v[c + 1]), a.scale_log2e, -mx[c + 1]));
          ls[c] += p0;
          ls[c + 1] += p1;
          if (DROP) {
            const unsigned long long i0 =
                ((static_cast<unsigned long long>(b) * a.H + hp * 2 + hd) * a.T + qt * TQ + cb + c) * a.S + tile * TKEYS + quad * 32 + lane;
            p0 = drop_val(a.drop, i0, p0);
            p1 = drop_val(a.drop, i0 + a.S, p1);
          }
          pk[c >> 1] = pack_bf16x2(p0, p1);
        }
        ptx::mbar_wait(&bar->p_empty[ps], pph ^ 1);
        const int row = quad * 32 + lane;
        uint8_t* prow = sP + ps * P_STAGE + row * 128;
#pragma unroll
        for (int j = 0; j < CW / 8; ++j) {
          const int chunk = (hd * 4 + half * (CW / 8) + j) ^ (row & 7);   // 128-byte swizzle
          *reinterpret_cast<uint4*>(prow + chunk * 16) = make_uint4(pk[4 * j], pk[4 * j + 1], pk[4 * j + 2], pk[4 * j + 3]);
        }
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bar->p_ready[ps]);
        if (++ps == PS) { ps = 0; pph ^= 1; }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (warp == 2) XTRACE(0, 4);
      if (!a.weights && lane == 0) ptx::mbar_arrive(&bar->s_empty);        // score columns may be overwritten
      fetch_mask(item + gridDim.x);      // the next item's mask bytes: in flight through the column sums and the epilogue
      // column sums across the warp, then across the 4 quadrant warps of this head
      // butterfly reduce-scatter: at distance ofs a lane keeps the half of its columns that its own bit selects and hands the
      // other half to its partner -- 31 shuffles for the 32 sums (a full warp_sum per column: 160); lane L ends with column L
#pragma unroll
      for (int ofs = CW / 2; ofs >= 1; ofs >>= 1) {
        const bool up = (lane & ofs) != 0;
#pragma unroll
        for (int i = 0; i < ofs; ++i) {
          const float send = up ? ls[i] : ls[i + ofs];
          const float keep = up ? ls[i + ofs] : ls[i];
          ls[i] = keep + __shfl_xor_sync(0xffffffffu, send, ofs);
        }
      }
      ls[0] += __shfl_xor_sync(0xffffffffu, ls[0], CW);     // lanes L and L ^ 16 hold the two halves of column L & 15
      if (lane < CW) red_sum[cb + lane] = ls[0];
      named_bar(3, kSoftWarps * 32);     // the epilogue reads the sums of BOTH head groups
      if (warp == 2) XTRACE(0, 5);
      if (a.weights) {
        // ---- pass 3 (last decoder layer only): head-averaged attention weights.  The scores are still resident in TMEM
        // and the normalisers are final, so every probability is rebuilt exactly as pass 2 made it, scaled by
        // 1 / (l * H) and added to weights[b, t, key]: the 8 head pairs of a caption land on the same row through
        // red.global.add (lanes = consecutive keys: one 128-byte reduction per warp and query).  Masked keys add nothing.
        const float* rsl = red + 256 + hd * 128 + cb + l16;
        const float lsum = rsl[0] + rsl[32] + rsl[64] + rsl[96];
        const float my_inv = lsum > 0.f ? a.inv_heads / lsum : 0.f;          // column (query) cb + (lane & 15)
        uint32_t ml = live;
        for (int i = 0; i < nl; ++i, ml &= ml - 1) {
          const int tile = __ffs(ml) - 1;
          uint32_t v[CW];
          ptx::tmem_ld_32x16(tmem + lane_base + i * 64 + hd * 32 + cb, v);
          ptx::tmem_ld_wait();
          const bool ok = (valid >> i) & 1;
          const int key = tile * TKEYS + quad * 32 + lane;
          float* wrow = a.weights + (static_cast<long>(b) * a.T + qt * TQ + cb) * a.S + key;
#pragma unroll
          for (int c = 0; c < CW; ++c) {
            const float inv = __shfl_sync(0xffffffffu, my_inv, c);
            float pw = ex2_fast(fmaf(__uint_as_float(v[c]), a.scale_log2e, -mx[c]));
            if (DROP) {
              const unsigned long long i0 =
                  ((static_cast<unsigned long long>(b) * a.H + hp * 2 + hd) * a.T + qt * TQ + cb + c) * a.S + tile * TKEYS + quad * 32 + lane;
              pw = drop_val(a.drop, i0, pw);
            }
            if (ok && qt * TQ + cb + c < a.T) atomicAdd(wrow + static_cast<long>(c) * a.S, pw * inv);
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bar->s_empty);
      }
      if (a.lse && quad == 0) {          // log-sum-exp of the scaled scores, saved for the backward pass
        const float* rsl = red + 256 + hd * 128 + cb + l16;
        const float l = rsl[0] + rsl[32] + rsl[64] + rsl[96];
        const int t = qt * TQ + cb + lane;
        if (lane < CW && t < a.T && score_warp)
          a.lse[(static_cast<long>(b) * a.H + (D128 ? hp : hp * 2 + hd)) * a.T + t] = l > 0.f ? col_max * a.scale + __logf(l) : INFINITY;
      }
      // ---- epilogue: O^T rows of this quadrant belong to head (quad / 2); each of the quadrant's 4 warps takes 8 queries
      const int oh = D128 ? 0 : quad >> 1;
      const int q0 = grp * 8;
      // 1 / l of query q0 + (lane & 7), handed out below with one shuffle per query (a division and four shared-memory
      // loads per ELEMENT made the epilogue a quarter of the kernel's instructions)
      float inv_mine;
      {
        const float* rs = red + 256 + oh * 128;
        const int q = q0 + (lane & 7);
        const float l = rs[q] + rs[32 + q] + rs[64 + q] + rs[96 + q];
        inv_mine = l > 0.f ? 1.f / l : 0.f;
      }
      ptx::mbar_wait(&bar->o_ready, par);
      ptx::tc_fence_after();
      if (warp == 2) XTRACE(0, 6);
      uint32_t o[8];
      ptx::tmem_ld_32x8(tmem + lane_base + O_COL + oh * 32 + q0, o);       // D128: O^T columns 0..31 = the 32 queries
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&bar->o_empty);
      const int dch = D128 ? quad * 32 + lane : (quad & 1) * 32 + lane;      // channel inside the head
      bf16* out = reinterpret_cast<bf16*>(a.out) + static_cast<long>(b) * a.o_stride_b + (D128 ? hp * 2 * D : (hp * 2 + oh) * D) + dch +
                  static_cast<long>(qt * TQ + q0) * a.o_stride_t;
      const int n_q = a.T - (qt * TQ + q0);                                  // queries of this block that exist
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float inv = __shfl_sync(0xffffffffu, inv_mine, j);
        if (j < n_q) *out = __float2bfloat16_rn(__uint_as_float(o[j]) * inv);
        out += a.o_stride_t;
      }
      if (warp == 2) XTRACE(0, 7);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc<512>(tmem);
}
}  // namespace atc

namespace atc { static long long* g_trace_buf = nullptr; }
// debug: device buffer of 4 * 3 * 8 * 8 int64 that the next attention launches stamp with clock64 (NULL switches it off)
extern "C" void tnt_debug_xattn_trace(long long* dev_buf) { atc::g_trace_buf = dev_buf; }

namespace atc {
__global__ void __launch_bounds__(256) tile_mask_kernel(const uint8_t* mask, long m_stride_b, int B, int S, int mask_S,
                                                        uint32_t* out) {
  const int b = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (b < B) tile_bits_row(mask, m_stride_b, b, S, mask_S, out, threadIdx.x & 31);
}
}  // namespace atc

int xattn_tile_mask(const uint8_t* mask, long m_stride_b, int B, int S, int mask_S, uint32_t* out, cudaStream_t st) {
  TNT_CHECK_ARG(mask && out && B > 0 && S >= 0 && S <= 16 * atc::TKEYS, "xattn_tile_mask: bad arguments (S <= 2048)");
  atc::tile_mask_kernel<<<(B + 7) / 8, 256, 0, st>>>(mask, m_stride_b, B, S, mask_S > 0 ? std::min(mask_S, S) : S, out);
  TNT_LAUNCHED();
  return TNT_OK;
}

bool xattn_tc_supported(const tnt_xattn_params* p) {
  if (p->dtype != TNT_BF16 || p->n_extra != 0 || p->kv_group > 1) return false;
  if (p->D == 128) {
    // the final* model: one 128-dim head per item, transformer_v2's `mask == 0 -> -1e9` rule; no weights / LSE / dropout
    static const bool d128_on = [] { const char* e = getenv("TNT_XATTN_TC_D128"); return !(e && e[0] == '0'); }();
    if (!d128_on || p->weights || p->lse || p->drop.p > 0.f || p->tile_live) return false;
    if (p->mask_mode != TNT_MASK_NONE && p->mask_mode != TNT_MASK_V2) return false;
  } else {
    if (p->D != 64 || (p->H & 1)) return false;
    if (p->mask_mode != TNT_MASK_NONE && p->mask_mode != TNT_MASK_KEY_PADDING) return false;
  }
  if (p->S < 1 || p->S > atc::MAX_TILES * atc::TKEYS) return false;
  const long strides = p->q_stride_t | p->q_stride_b | p->k_stride_s | p->k_stride_b | p->v_stride_s | p->v_stride_b;
  return (strides % 8) == 0 && aligned16(p->q) && aligned16(p->k) && aligned16(p->v);
}

int xattn_fwd_tc(const tnt_xattn_params* p, cudaStream_t st) {
  CUtensorMap tmQ, tmK, tmV;
  const int E = p->H * p->D;
  TNT_TRY(make_tmap_bf16_3d(&tmQ, p->q, E, p->B, p->T, p->q_stride_b, p->q_stride_t, 64, 1, atc::TQ));
  TNT_TRY(make_tmap_bf16_3d(&tmK, p->k, E, p->B, p->S, p->k_stride_b, p->k_stride_s, 64, 1, atc::TKEYS));
  TNT_TRY(make_tmap_bf16_3d(&tmV, p->v, E, p->B, p->S, p->v_stride_b, p->v_stride_s, 64, 1, atc::TKEYS));
  atc::Args a;
  const bool d128 = p->D == 128;
  a.mask = (p->mask_mode == TNT_MASK_KEY_PADDING || p->mask_mode == TNT_MASK_V2) ? p->mask : nullptr;
  a.m_stride_b = p->m_stride_b;
  a.m_stride_t = p->m_stride_t;
  a.out = p->out; a.o_stride_t = p->o_stride_t; a.o_stride_b = p->o_stride_b;
  a.B = p->B; a.T = p->T; a.S = p->S; a.H = p->H;
  a.mask_S = p->mask_S > 0 ? p->mask_S : p->S;
  a.scale_log2e = p->scale * 1.4426950408889634f;
  a.lse = p->lse; a.scale = p->scale;
  a.drop = make_drop(p->drop);
  a.tile_live = p->tile_live;
  a.weights = p->weights;
  a.inv_heads = 1.f / static_cast<float>(p->H);
  // the kernel accumulates the weights with reductions (and never visits all-padding tiles): start from zero
  if (p->weights) TNT_CUDA(cudaMemsetAsync(p->weights, 0, sizeof(float) * static_cast<size_t>(p->B) * p->T * p->S, st));
  a.n_qt = (p->T + atc::TQ - 1) / atc::TQ;
  a.n_items = p->B * (d128 ? p->H : p->H / 2) * a.n_qt;
  {
    const int hp = d128 ? p->H : p->H / 2;
    // fast_div is exact while x * d < 2^32 (x < n_items, d <= max(n_qt, hp)); xattn_tc_supported keeps B <= 65535
    TNT_CHECK_ARG(static_cast<uint64_t>(a.n_items) * std::max(a.n_qt, hp) < (1ull << 32), "xattn_tc: too many work items");
    a.mul_qt = a.n_qt > 1 ? static_cast<uint32_t>((1ull << 32) / a.n_qt) + 1u : 0u;
    a.mul_hp = hp > 1 ? static_cast<uint32_t>((1ull << 32) / hp) + 1u : 0u;
  }
  a.n_tiles = (p->S + atc::TKEYS - 1) / atc::TKEYS;
  a.trace = atc::g_trace_buf;
  static PerDeviceOnce attr_once;
  if (attr_once.first()) {
    TNT_CUDA(cudaFuncSetAttribute(atc::xattn_tc_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::SMEM_BYTES));
    TNT_CUDA(cudaFuncSetAttribute(atc::xattn_tc_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::SMEM_BYTES));
    TNT_CUDA(cudaFuncSetAttribute(atc::xattn_tc_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, atc::SMEM_BYTES));
  }
  const int grid = std::min(a.n_items, num_sms());
  if (d128)
    TNT_CUDA(launch_pdl(atc::xattn_tc_kernel<false, true>, dim3(grid), dim3(atc::kThreads), atc::SMEM_BYTES, st, 1, tmQ, tmK, tmV, a));
  else if (a.drop.thr)
    TNT_CUDA(launch_pdl(atc::xattn_tc_kernel<true, false>, dim3(grid), dim3(atc::kThreads), atc::SMEM_BYTES, st, 1, tmQ, tmK, tmV, a));
  else
    TNT_CUDA(launch_pdl(atc::xattn_tc_kernel<false, false>, dim3(grid), dim3(atc::kThreads), atc::SMEM_BYTES, st, 1, tmQ, tmK, tmV, a));
  TNT_LAUNCHED();
  return TNT_OK;
}

}  // namespace tnt

extern "C" int tnt_xattn_tile_mask(const uint8_t* mask, int64_t m_stride_b, int B, int S, int mask_S, uint32_t* out,
                                   tnt_stream stream) {
  return tnt::xattn_tile_mask(mask, m_stride_b, B, S, mask_S, out, tnt::as_stream(stream));
}
