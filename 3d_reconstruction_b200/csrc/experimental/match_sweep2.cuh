// The sweep kernel, CTA-pair version (tcgen05 cta_group::2): same work decomposition, same
// epilogue (process16 / RowState of match_sweep.cuh), different machine mapping.
//
// A cluster of two CTAs (two SMs of one TPC) owns a work item = (pair, block of 256 rows). CTA r of
// the pair keeps rows [128 r, 128 r + 128) of the block in shared memory and loads only rows
// [128 r, 128 r + 128) of every 256-column tile; one tcgen05.mma.cta_group::2 (M = 256, N = 256,
// K = 32, issued by the leader CTA) multiplies the pair's two row halves by both column halves and
// leaves a 128 x 256 int32 accumulator in each CTA's TMEM. Compared with the single-CTA kernel:
//   * N = 256 per MMA at half the shared-memory and L2 traffic per SM (each SM stages 16 KiB per tile),
//   * two full 128 x 256 accumulators per CTA: the MMA of tile t+1 overlaps the whole epilogue of tile t.
//   warp 0       TMA producer (own row tile per item; own half of each column tile; the tile's norms)
//   warp 1       MMA issuer (leader CTA only)
//   warp 2       TMEM allocator (both CTAs, cta_group::2)
//   warps 4-19   epilogue: warp (quadrant q, column quarter c) owns rows 32q..32q+31 and columns
//                64c..64c+63 of every tile; the four quarters of a row share their running bound
//                through shared memory and meet once per item.
#pragma once
#include "../match_sweep.cuh"

namespace b200m {
namespace pair {

constexpr int kBHalfBytes = kBBytes / 2;   // 16 KiB: this CTA's half of a column tile
constexpr int kAStages2 = 2;
constexpr int kBStages2 = 6;
constexpr int kNStages2 = 4;
constexpr int kQuarters = 4;
constexpr int kThreads2 = (kFirstEpiWarp + kEpiWarps) * 32;   // 640

struct PairSmem {
  // tiles first: SWIZZLE_128B needs 1024 B alignment
  uint8_t a[kAStages2][kMBytes];
  uint8_t b[kBStages2][kBHalfBytes];
  NormSlot norms[kNStages2];
  int4 rowmerge[2][kQuarters - 1][kTM];   // quarters 1..3 hand their row state to quarter 0
  int2 share[2][kQuarters][kTM];          // [item parity][quarter][row]: published (s1, s2)
  uint64_t a_full[kAStages2], a_empty[kAStages2];
  uint64_t b_full[kBStages2], b_empty[kBStages2];
  uint64_t n_full[kNStages2], n_empty[kNStages2];
  uint64_t t_full[2], t_empty[2];
  uint32_t tmem_base;
};

// merge (s, i) pair lists of two sorted top-2 states, lexicographic on (score, column)
__device__ __forceinline__ void merge_top2(int& s1, int& i1, int& s2, int& i2, int os1, int oi1, int os2, int oi2) {
  int r1s, r1i, r2s, r2i;
  if (lex_less(os1, oi1, s1, i1)) {
    r1s = os1; r1i = oi1;
    if (lex_less(os2, oi2, s1, i1)) { r2s = os2; r2i = oi2; } else { r2s = s1; r2i = i1; }
  } else {
    r1s = s1; r1i = i1;
    if (lex_less(os1, oi1, s2, i2)) { r2s = os1; r2i = oi1; } else { r2s = s2; r2i = i2; }
  }
  s1 = r1s; i1 = r1i; s2 = r2s; i2 = r2i;
}


template <int kKind>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, 1)
sweep_pair_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_a2,
                  const __grid_constant__ CUtensorMap tmap_b, const SweepParams p) {
  extern __shared__ uint8_t smem_raw[];
  PairSmem& sm = *reinterpret_cast<PairSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  const int n_items = (kKind == kVerify && p.vitems) ? __ldg(p.vitem_count)
                                                     : p.n_pairs * p.blocks_per_pair * (kKind == kVerify ? 2 : 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_a2);
    tma_prefetch_desc(&tmap_b);
    for (int i = 0; i < kAStages2; ++i) { mbar_init(&sm.a_full[i], 1); mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kBStages2; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < kNStages2; ++i) { mbar_init(&sm.n_full[i], 1); mbar_init(&sm.n_empty[i], kEpiWarps); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.t_full[i], 1); mbar_init(&sm.t_empty[i], 2 * kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc_pair<512>(&sm.tmem_base);
  if (warp >= kFirstEpiWarp) {
    const int t = threadIdx.x - kFirstEpiWarp * 32;          // 0..511 -> one (quarter, row) slot each
    sm.share[0][t >> 7][t & 127] = make_int2(kBig, kBig);
    sm.share[1][t >> 7][t & 127] = make_int2(kBig, kBig);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();          // the peer's barriers are initialised before anything arrives on them
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer (both CTAs)
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0, n_it = 0;
      for (int item = cluster_id; item < n_items; item += n_clusters) {
        const ItemInfo it = decode_item<kKind>(p, item);
        if (!it.valid) continue;
        const uint32_t as = a_it % kAStages2, aph = (a_it / kAStages2) & 1;
        mbar_wait(&sm.a_empty[as], aph ^ 1);
        if (leader) mbar_arrive_expect_tx(&sm.a_full[as], 2 * kMBytes);      // both CTAs' row tiles
        const CUtensorMap* ta = (kKind == kVerify && it.check == 1) ? &tmap_a2 : &tmap_a;
        tma_load_2d_pair(sm.a[as], ta, 0, it.row_img_row0 + static_cast<int>(rank) * kTM,
                         mapa_u32(smem_u32(&sm.a_full[as]), 0));
        ++a_it;
        const int64_t crow0 = static_cast<int64_t>(it.col_img) * p.k_stride;
        for (int t = 0; t < it.ntiles; ++t) {
          const int32_t crow = static_cast<int32_t>(crow0 + t * kTT);
          if (kKind != kPeak) {
            const uint32_t ns = n_it % kNStages2, nph = (n_it / kNStages2) & 1;
            mbar_wait(&sm.n_empty[ns], nph ^ 1);
            mbar_arrive_expect_tx(&sm.n_full[ns], sizeof(NormSlot));
            bulk_load(sm.norms[ns].nb, p.sqnorm + crow, kTT * 4, &sm.n_full[ns]);
            bulk_load(sm.norms[ns].nbmin, p.nbmin32 + crow / 32, 8 * 4, &sm.n_full[ns]);
            ++n_it;
          }
          const uint32_t bs = b_it % kBStages2, bph = (b_it / kBStages2) & 1;
          mbar_wait(&sm.b_empty[bs], bph ^ 1);
          if (leader) mbar_arrive_expect_tx(&sm.b_full[bs], kBBytes);        // both halves of the column tile
          tma_load_2d_pair(sm.b[bs], &tmap_b, 0, crow + static_cast<int>(rank) * kTM,
                           mapa_u32(smem_u32(&sm.b_full[bs]), 0));
          ++b_it;
        }
      }
      // the leader's last commits multicast into this CTA's barriers: wait for them before leaving
      for (uint32_t k = 0; k < kAStages2 && k < a_it; ++k) {
        const uint32_t idx = a_it - 1 - k;
        mbar_wait(&sm.a_empty[idx % kAStages2], (idx / kAStages2) & 1);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer (leader CTA)
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = make_idesc_u8(2 * kTM, kTT);
      uint32_t a_it = 0, b_it = 0, t_it = 0;
      for (int item = cluster_id; item < n_items; item += n_clusters) {
        const ItemInfo it = decode_item<kKind>(p, item);
        if (!it.valid) continue;
        const uint32_t as = a_it % kAStages2, aph = (a_it / kAStages2) & 1;
        mbar_wait(&sm.a_full[as], aph);
        const uint64_t adesc = make_desc_k128(smem_u32(sm.a[as]));
        for (int t = 0; t < it.ntiles; ++t) {
          const uint32_t bs = b_it % kBStages2, bph = (b_it / kBStages2) & 1;
          const uint32_t ts = t_it & 1, tph = (t_it >> 1) & 1;
          mbar_wait(&sm.b_full[bs], bph);
          mbar_wait(&sm.t_empty[ts], tph ^ 1);
          tc_fence_after();
          const uint64_t bdesc = make_desc_k128(smem_u32(sm.b[bs]));
          const uint32_t tacc = tmem_base + ts * kTT;
#pragma unroll
          for (int k = 0; k < kD / 32; ++k) mma_i8_ss_pair(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
          tc_commit_pair(&sm.b_empty[bs]);   // both CTAs' stage is free once these MMAs retire
          tc_commit_pair(&sm.t_full[ts]);    // both CTAs' accumulator halves are ready
          ++b_it;
          ++t_it;
        }
        tc_commit_pair(&sm.a_empty[as]);
        ++a_it;
      }
    }
  } else if (warp >= kFirstEpiWarp) {
    // ------------------------------------------------------------- epilogue (both CTAs)
    const uint32_t quad = warp & 3;                   // TMEM lane quadrant this warp may read
    const int cq = (warp - kFirstEpiWarp) >> 2;       // column quarter
    const int rl = quad * 32 + lane;                  // row inside this CTA's tile
    const int brow = static_cast<int>(rank) * kTM + rl;   // row inside the item's block
    const uint32_t te0 = mapa_u32(smem_u32(&sm.t_empty[0]), 0), te1 = mapa_u32(smem_u32(&sm.t_empty[1]), 0);
    uint32_t t_it = 0, n_it = 0, item_it = 0;
    for (int item = cluster_id; item < n_items; item += n_clusters) {
      const ItemInfo it = decode_item<kKind>(p, item);
      if (!it.valid) continue;
      const bool row_ok = brow < it.n_rows;
      const int64_t orow = static_cast<int64_t>(it.pair) * p.k_stride + it.blk * kRows + brow;

      RowState st;
      st.s1 = kBig; st.i1 = -1; st.s2 = kBig; st.i2 = -1; st.lim = kBig;
      st.thr = -kBig; st.excl = -1; st.tie = -kBig; st.viol = 0;   // padding rows can never be violated
      const uint32_t par = item_it & 1;
      ++item_it;
      st.pub = &sm.share[par][cq][rl];
      if (kKind == kVerify && row_ok) {
        if (it.check == 0) {
          st.thr = __ldg(p.cand_thr_ratio + orow);
          st.excl = __ldg(p.cand_excl + orow);
        } else {
          st.thr = __ldg(p.cand_thr_mut + orow);
          st.tie = __ldg(p.cand_q + orow);
        }
      }

      for (int t = 0; t < it.ntiles; ++t) {
        const uint32_t ts = t_it & 1, tph = (t_it >> 1) & 1;
        ++t_it;
        if (kKind == kPeak) {
          mbar_wait(&sm.t_full[ts], tph);
          tc_fence_after();
          int v[16];
          tmem_ld_32x16(tmem_base + ((quad * 32u) << 16) + ts * kTT + cq * 64, v);
          tmem_ld_fence16(v);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(ts ? te1 : te0);
          st.s1 = min(st.s1, v[0]);
          continue;
        }
        const uint32_t ns = n_it % kNStages2, nph = (n_it / kNStages2) & 1;
        ++n_it;
        mbar_wait(&sm.n_full[ns], nph);
        const NormSlot& nsl = sm.norms[ns];
        if (kKind == kTop2 || kKind == kTop1) {
          // Bounds from the other three column quarters of this row. kTop1: the row's best is
          // <= every s1'; kTop2: its second best is <= min(max(s1, s1'), s2'). Anything at or above
          // the bound (strictly above on ties: the quarters interleave in column order) cannot
          // matter. Racy reads are fine: published values only ever decrease.
          int bound = kBig;
#pragma unroll
          for (int o = 1; o < kQuarters; ++o) {
            const long long raw = *reinterpret_cast<const volatile long long*>(&sm.share[par][(cq + o) & 3][rl]);
            const int ox = static_cast<int>(raw), oy = static_cast<int>(raw >> 32);
            bound = min(bound, kKind == kTop1 ? ox : min(max(st.s1, ox), oy));
          }
          st.lim = min(st.lim, bound + 1);
        }
        mbar_wait(&sm.t_full[ts], tph);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + ts * kTT + cq * 64;
        const int col0 = t * kTT + cq * 64;
        const uint32_t nb_s = smem_u32(nsl.nb) + (cq * 64) * 4;
        const uint32_t nbm_s = smem_u32(nsl.nbmin) + (cq * 2) * 4;
        // software-pipelined drain in 16-column halves (two register buffers)
        int va[16], vb[16];
        tmem_ld_32x16(taddr, va);
        tmem_ld_fence16(va);
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          const int nbm = lds_s32(nbm_s + 4 * c);
          tmem_ld_32x16(taddr + c * 32 + 16, vb);
          process16<kKind>(va, nb_s + (c * 32) * 4, nbm, col0 + c * 32, st);
          tmem_ld_fence16(vb);
          if (c < 1) {
            tmem_ld_32x16(taddr + (c + 1) * 32, va);
          } else {
            // this warp's part of the accumulator is in registers: hand it back to the leader's MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(ts ? te1 : te0);
          }
          process16<kKind>(vb, nb_s + (c * 32 + 16) * 4, nbm, col0 + c * 32 + 16, st);
          if (c < 1) tmem_ld_fence16(va);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.n_empty[ns]);
      }

      // the four column quarters of a row meet here: quarters 1..3 -> smem -> quarter 0 -> global
      if (cq > 0)
        sm.rowmerge[par][cq - 1][rl] = kKind == kVerify ? make_int4(st.viol, 0, 0, 0) : make_int4(st.s1, st.i1, st.s2, st.i2);
      sm.share[par ^ 1][cq][rl] = make_int2(kBig, kBig);    // next item's slot; nobody reads it before the barrier
      named_bar_sync(kEpiBarrier, kEpiWarps * 32);
      if (cq == 0 && row_ok) {
        if (kKind == kVerify) {
          int viol = st.viol;
#pragma unroll
          for (int o = 0; o < kQuarters - 1; ++o) viol |= sm.rowmerge[par][o][rl].x;
          (it.check == 0 ? p.viol_ratio : p.viol_mut)[orow] = viol;
        } else {
#pragma unroll
          for (int o = 0; o < kQuarters - 1; ++o) {
            const int4 q = sm.rowmerge[par][o][rl];
            if (kKind == kTop1) {
              const bool other = lex_less(q.x, q.y, st.s1, st.i1);
              st.s2 = min(min(q.z, st.s2), other ? st.s1 : q.x);   // the losing best bounds the runner-up
              st.s1 = other ? q.x : st.s1;
              st.i1 = other ? q.y : st.i1;
            } else {
              merge_top2(st.s1, st.i1, st.s2, st.i2, q.x, q.y, q.z, q.w);
            }
          }
          const int na = __ldg(p.sqnorm + it.row_img_row0 + brow);
          p.rowres[orow] = make_int4(st.s1 + na, st.i1, st.s2 + na, kKind == kTop1 ? -1 : st.i2);
        }
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
