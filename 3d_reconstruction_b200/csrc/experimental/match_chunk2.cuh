// The chunk sweep on CTA pairs (tcgen05 cta_group::2), option "kernel" = 6.
//
// Why: an SM's TMEM read port moves 256 B/clk -- exactly the rate at which K = 128 int8 MMAs produce
// int32 results -- so the chunk sweep needs the tensor pipe AND the TMEM read port busy all the time.
// chunk_sweep256_kernel (match_chunk.cuh) has two 128 x 256 accumulators: every hand-over between
// the MMA warp and the epilogue (commit -> mbarrier -> wake-up -> first tcgen05.ld) is exposed, and
// the tensor pipe idles for that long once per accumulator (87 % active). Four accumulators hide the
// hand-over, but four need N = 128 per MMA, and a single CTA then reads 128 B/clk of operands from
// shared memory (its limit). A CTA pair splits the B operand: each CTA stages half of the columns,
// 96 B/clk per SM, and every SM has four 128 x 128 accumulators (all 512 TMEM columns).
//
// A cluster of two CTAs owns a work item = (pair, block of 512 sorted query rows). Row u*256 + r*128 + i
// of the block is row i of M-half u in CTA r. Per tile of 256 sorted train columns CTA r stages columns
// [128 r, 128 r + 128) (16 KiB) and the leader issues four groups of tcgen05.mma.cta_group::2
// (M = 256 across the pair, N = 128, K = 32): group (u, j) multiplies M-half u of both CTAs by columns
// {64 j .. 64 j + 63} of both CTAs' tile halves and fills accumulator slot 2 u + j in both CTAs:
// accumulator columns [0, 64) are tile columns 64 j + [0, 64), columns [64, 128) are tile columns
// 128 + 64 j + [0, 64) -- the 64-column chunks 4 t + j and 4 t + 2 + j of the train image.
//   warp 0       TMA producer (both CTAs: own rows, own half of every column tile; bytes are
//                accounted on the leader's barriers)
//   warp 1       MMA issuer (leader CTA only; commits are multicast to both CTAs' barriers)
//   warp 2       TMEM allocator (cta_group::2)
//   warps 4-19   epilogue: warp (lane quadrant q, slot s = 2 u + j) drains slot s of every tile for
//                rows 32 q .. 32 q + 31 of M-half u: two chunks per tile, branch-free (match_chunk.cuh)
#pragma once
#include "../match_chunk.cuh"

namespace b200m {
namespace pair {

constexpr int kRowsP = 512;          // rows per work item of the pair
constexpr int kAStagesP = 2;
constexpr int kBStagesP = 5;
constexpr int kSlots = 4;
constexpr int kSlotCols = 128;

struct ChunkPairSmem {
  uint8_t a[kAStagesP][2][kMBytes];     // [stage][M-half u]: this CTA's 128 rows
  uint8_t b[kBStagesP][kMBytes];        // this CTA's 128 columns of a 256-column tile
  int4 rowmerge[2][2 * kTM];            // column group j = 1 hands its row states to j = 0
  uint64_t a_full[kAStagesP], a_empty[kAStagesP];
  uint64_t b_full[kBStagesP], b_empty[kBStagesP];
  uint64_t t_full[kSlots], t_empty[kSlots];
  uint32_t tmem_base;
};

// 128 accumulator columns (two 64-column chunks) of one row, in halves of 16 columns through two
// register buffers; the slot goes back to the leader's MMA warp as soon as the last load has landed.
template <bool kPeakOnly>
__device__ __forceinline__ void drain128(uint32_t taddr, int cm_lo, int cm_hi, int g_lo, int g_hi, ChunkState& st,
                                         uint32_t t_empty_leader, uint32_t lane) {
  int va[16], vb[16];
  tmem_ld_32x16(taddr, va);
  tmem_ld_fence16(va);
  if (kPeakOnly) {          // MMA / TMA / hand-over ceiling: one load per slot
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive_cluster(t_empty_leader);
    chunk_update(st, cm_lo - 2 * max16(va), g_lo);
    return;
  }
  tmem_ld_32x16(taddr + 16, vb);
  int m = max16(va);
  tmem_ld_fence16(vb);
  tmem_ld_32x16(taddr + 32, va);
  m = max17(vb, m);
  tmem_ld_fence16(va);
  tmem_ld_32x16(taddr + 48, vb);
  m = max17(va, m);
  tmem_ld_fence16(vb);
  tmem_ld_32x16(taddr + 64, va);
  m = max17(vb, m);
  chunk_update(st, cm_lo - 2 * m, g_lo);
  tmem_ld_fence16(va);
  tmem_ld_32x16(taddr + 80, vb);
  m = max16(va);
  tmem_ld_fence16(vb);
  tmem_ld_32x16(taddr + 96, va);
  m = max17(vb, m);
  tmem_ld_fence16(va);
  tmem_ld_32x16(taddr + 112, vb);
  m = max17(va, m);
  tmem_ld_fence16(vb);
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive_cluster(t_empty_leader);
  m = max17(vb, m);
  chunk_update(st, cm_hi - 2 * m, g_hi);
}

template <bool kPeakOnly>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
chunk_sweep_pair_kernel(const __grid_constant__ CUtensorMap tmap, const ChunkParams p) {
  extern __shared__ uint8_t smem_raw[];
  ChunkPairSmem& sm = *reinterpret_cast<ChunkPairSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int blocks_per_pair = static_cast<int>((p.k_stride + kRowsP - 1) / kRowsP);
  const int n_items = p.n_pairs * blocks_per_pair;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap);
    for (int i = 0; i < kAStagesP; ++i) { mbar_init(&sm.a_full[i], 1); mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kBStagesP; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < kSlots; ++i) { mbar_init(&sm.t_full[i], 1); mbar_init(&sm.t_empty[i], 2 * 4); }   // 4 quadrant warps x 2 CTAs
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_pair<512>(&sm.tmem_base);
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // the peer's barriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;

  // work item = (pair, block of 512 sorted query rows); nm = M-halves in use (rows are u-major)
  auto decode = [&](int item, int& pair, int& blk, int& row0, int& col_img, int& n_rows, int& ntiles, int& nm) {
    pair = item / blocks_per_pair;
    blk = item - pair * blocks_per_pair;
    const int qi = __ldg(p.pairs + 2 * pair);
    const int ti = __ldg(p.pairs + 2 * pair + 1);
    const int kq = __ldg(p.k_per_image + qi);
    const int kt = __ldg(p.k_per_image + ti);
    n_rows = min(kq - blk * kRowsP, kRowsP);
    col_img = ti;
    row0 = static_cast<int>(qi * p.k_stride + blk * kRowsP);
    ntiles = (kt + kTT - 1) / kTT;
    nm = n_rows > 2 * kTM ? 2 : 1;
    return n_rows > 0 && ntiles > 0;
  };

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer (both CTAs)
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0;
      for (int item = cluster_id; item < n_items; item += n_clusters) {
        int pair, blk, row0, col_img, n_rows, ntiles, nm;
        if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
        const uint32_t as = a_it % kAStagesP, aph = (a_it / kAStagesP) & 1;
        mbar_wait(&sm.a_empty[as], aph ^ 1);
        if (leader) mbar_arrive_expect_tx(&sm.a_full[as], 2 * nm * kMBytes);      // both CTAs' row tiles
        const uint32_t afull = mapa_u32(smem_u32(&sm.a_full[as]), 0);
        for (int u = 0; u < nm; ++u)
          tma_load_2d_pair(sm.a[as][u], &tmap, 0, row0 + u * 2 * kTM + static_cast<int>(rank) * kTM, afull);
        ++a_it;
        const int64_t crow0 = static_cast<int64_t>(col_img) * p.k_stride;
        for (int t = 0; t < ntiles; ++t) {
          const int32_t crow = static_cast<int32_t>(crow0 + t * kTT);
          const uint32_t bs = b_it % kBStagesP, bph = (b_it / kBStagesP) & 1;
          mbar_wait(&sm.b_empty[bs], bph ^ 1);
          if (leader) mbar_arrive_expect_tx(&sm.b_full[bs], 2 * kMBytes);           // both halves of the column tile
          tma_load_2d_pair(sm.b[bs], &tmap, 0, crow + static_cast<int>(rank) * kTM,
                           mapa_u32(smem_u32(&sm.b_full[bs]), 0));
          ++b_it;
        }
      }
      // the leader's last commits are multicast into this CTA's barriers: wait for them before leaving
      for (uint32_t k = 0; k < kAStagesP && k < a_it; ++k) {
        const uint32_t idx = a_it - 1 - k;
        mbar_wait(&sm.a_empty[idx % kAStagesP], (idx / kAStagesP) & 1);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer (leader CTA)
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = make_idesc_u8(2 * kTM, kSlotCols);
      uint32_t a_it = 0, b_it = 0;
      uint32_t t_cnt[kSlots] = {0, 0, 0, 0};
      for (int item = cluster_id; item < n_items; item += n_clusters) {
        int pair, blk, row0, col_img, n_rows, ntiles, nm;
        if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
        const uint32_t as = a_it % kAStagesP, aph = (a_it / kAStagesP) & 1;
        mbar_wait(&sm.a_full[as], aph);
        for (int t = 0; t < ntiles; ++t) {
          const uint32_t bs = b_it % kBStagesP, bph = (b_it / kBStagesP) & 1;
          mbar_wait(&sm.b_full[bs], bph);
#pragma unroll
          for (int s = 0; s < kSlots; ++s) {
            const int u = s >> 1, j = s & 1;
            if (u < nm) {
              mbar_wait(&sm.t_empty[s], (t_cnt[s] & 1) ^ 1);
              tc_fence_after();
              const uint64_t adesc = make_desc_k128(smem_u32(sm.a[as][u]));
              const uint64_t bdesc = make_desc_k128(smem_u32(sm.b[bs]) + j * (64 * kD));   // rows 64 j .. of this CTA's half
#pragma unroll
              for (int k = 0; k < kD / 32; ++k)
                mma_i8_ss_pair(tmem_base + s * kSlotCols, adesc + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
              tc_commit_pair(&sm.t_full[s]);
              ++t_cnt[s];
            }
          }
          tc_commit_pair(&sm.b_empty[bs]);
          ++b_it;
        }
        tc_commit_pair(&sm.a_empty[as]);
        ++a_it;
      }
    }
  } else if (warp >= kFirstEpiWarp) {
    // ------------------------------------------------------------- epilogue (both CTAs)
    const uint32_t quad = warp & 3;                      // TMEM lane quadrant this warp may read
    const int s = (warp - kFirstEpiWarp) >> 2;           // accumulator slot = (M-half u, column group j)
    const int u = s >> 1, j = s & 1;
    const int rl = quad * 32 + lane;                     // row inside the M-half of this CTA
    const int brow = u * 2 * kTM + static_cast<int>(rank) * kTM + rl;   // row inside the item's block
    const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + s * kSlotCols;
    const uint32_t te = mapa_u32(smem_u32(&sm.t_empty[s]), 0);
    uint32_t t_it = 0, item_it = 0;
    for (int item = cluster_id; item < n_items; item += n_clusters) {
      int pair, blk, row0, col_img, n_rows, ntiles, nm;
      if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
      const uint32_t par = item_it & 1;
      ++item_it;
      ChunkState st;
      st.L1 = kBig; st.g1 = 0; st.L2 = kBig;
      if (u < nm) {
        // minima of the 64-column chunks 4 t + j and 4 t + 2 + j (sorted norms: the first of their 32-column halves)
        const int32_t* cm = p.cmin + (static_cast<int64_t>(col_img) * p.k_stride) / kChunk + 2 * j;
        int c_lo = __ldg(cm), c_hi = __ldg(cm + 4);
        for (int t = 0; t < ntiles; ++t) {
          const int lo = c_lo, hi = c_hi;
          if (t + 1 < ntiles) { c_lo = __ldg(cm + 8 * (t + 1)); c_hi = __ldg(cm + 8 * (t + 1) + 4); }
          mbar_wait(&sm.t_full[s], t_it & 1);
          tc_fence_after();
          ++t_it;
          drain128<kPeakOnly>(taddr, lo, hi, 4 * t + j, 4 * t + 2 + j, st, te, lane);
        }
      }
      // the two column groups of a row meet here: j = 1 -> smem -> j = 0 -> global
      if (j == 1) sm.rowmerge[par][u * kTM + rl] = make_int4(st.L1, st.g1, st.L2, 0);
      named_bar_sync(kEpiBarrier, kEpiWarps * 32);
      if (j == 0 && brow < n_rows) {
        const int4 q = sm.rowmerge[par][u * kTM + rl];
        // the chunks of the two groups interleave: order the pair (L, g) explicitly
        const bool other = lex_less(q.x, q.y, st.L1, st.g1);
        st.L2 = min(min(q.z, st.L2), other ? st.L1 : q.x);
        st.g1 = other ? q.y : st.g1;
        st.L1 = other ? q.x : st.L1;
        p.rowres[static_cast<int64_t>(pair) * p.k_stride + blk * kRowsP + brow] = make_int4(st.L1, st.g1, st.L2, 0);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // no CTA of the pair leaves (or frees TMEM) while the other may still signal it
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc_pair<512>(tmem_base);
  }
}

}  // namespace pair
}  // namespace b200m
