// The chunk sweep: the dominant kernel of the default pipeline ("pipeline" = 2).
//
// Same machine mapping as sweep_kernel (match_sweep.cuh: TMA producer warp, MMA issuer warp, 16
// epilogue warps around four 128 x 128 int32 TMEM accumulators), but the epilogue is branch-free.
// It runs on the NORM-SORTED copy of the bank (b200m_bank_sort_u8): inside a chunk of 32
// consecutive sorted columns the squared norms differ by next to nothing, so
//     LB(chunk) = min|col|^2 - 2 * max_{col in chunk} row.col
// is a lower bound of the chunk's best score |col|^2 - 2 row.col that is off by at most the
// chunk's norm spread. Per row the epilogue keeps only the two smallest chunk bounds and the
// chunk that holds the smallest:
//     16 x VIMNMX3 per 32 accumulator values (max tree) + 5 integer ops per chunk,
// no divergence, no shared-memory traffic. The refine kernel (match_refine.cuh) then evaluates the
// one winning chunk of a row exactly, and only rows whose answer is not settled by the bounds fall
// back to the exact threshold-reject sweep (sweep_kernel<kTop1G>) -- results stay bit-exact.
//
// Stands in for lightglue/lightglue.py:223 (dense contraction) + :237 (row arg-best) in the
// exact-integer L2 form named by BASELINE.json (cv2.BFMatcher NORM_L2 knnMatch).
#pragma once
#include <type_traits>

#include "match_sweep.cuh"

namespace b200m {

constexpr int kChunk = 32;                  // columns per chunk
constexpr int kChunksPerTile = kTT / kChunk;   // 8

struct ChunkSmem {
  // tiles first: SWIZZLE_128B needs 1024 B alignment
  uint8_t a[kAStages][2][kMBytes];
  uint8_t b[kBStages][kBBytes];
  int4 rowmerge[2][kRows];          // column half 1 hands its row state to column half 0
  uint64_t a_full[kAStages], a_empty[kAStages];
  uint64_t b_full[kBStages], b_empty[kBStages];
  uint64_t t_full[4], t_empty[4];   // accumulator (m, h) -> index m * 2 + h
  uint32_t tmem_base;
};

struct ChunkParams {
  const int32_t* cmin;          // [n_img * k_stride / 32] smallest squared norm of every sorted chunk
  const int32_t* k_per_image;   // [n_img]
  const int32_t* pairs;         // [n_pairs, 2] (query image, train image) of this batch
  int64_t k_stride;
  int32_t n_pairs;
  int32_t blocks_per_pair;      // k_stride / 256
  int32_t n_img;                // pair-list entries outside [0, n_img) are skipped (load_pair)
  // out, per (pair, SORTED query row): (L1, g1, L2, 0) -- the smallest chunk bound, its chunk,
  // the second smallest chunk bound (kBig when the train image has a single chunk)
  int4* rowres;
  int32_t peak;                 // != 0: MMA/TMA ceiling measurement, the epilogue reads one chunk per tile
  int32_t chunk_cols;           // 32, or 64 (chunk_sweep256_kernel only): g1 then counts 64-column chunks
};

// max of 16 values (7 three-input maxima + 1), and of 16 values and a carried one (8 three-input maxima):
// 16 ALU instructions per 32 accumulator values
__device__ __forceinline__ int max16(const int (&v)[16]) {
  const int a = __vimax3_s32(v[0], v[1], v[2]), b = __vimax3_s32(v[3], v[4], v[5]);
  const int c = __vimax3_s32(v[6], v[7], v[8]), d = __vimax3_s32(v[9], v[10], v[11]);
  const int e = __vimax3_s32(v[12], v[13], v[14]);
  return max(__vimax3_s32(a, b, c), __vimax3_s32(d, e, v[15]));
}
__device__ __forceinline__ int max17(const int (&v)[16], int carry) {
  const int a = __vimax3_s32(v[0], v[1], v[2]), b = __vimax3_s32(v[3], v[4], v[5]);
  const int c = __vimax3_s32(v[6], v[7], v[8]), d = __vimax3_s32(v[9], v[10], v[11]);
  const int e = __vimax3_s32(v[12], v[13], v[14]), f = __vimax3_s32(v[15], carry, a);
  return __vimax3_s32(__vimax3_s32(b, c, d), e, f);
}

struct ChunkState {
  int L1, g1, L2;
};

// chunk g (visited in increasing g) with bound L: strict '<' keeps the lowest chunk on ties
__device__ __forceinline__ void chunk_update(ChunkState& s, int L, int g) {
  const bool lt = L < s.L1;
  s.L2 = min(s.L2, max(s.L1, L));
  s.g1 = lt ? g : s.g1;
  s.L1 = min(s.L1, L);
}

// knnMatch form of the chunk state: the THREE smallest chunk bounds of a row and the chunks holding the two
// smallest -- the exact two nearest neighbours then lie in chunks g1 / g2 whenever the second of them scores
// below L3 (match_knn.cuh evaluates the two chunks exactly, the rest falls back to the exact top-2 sweep).
struct ChunkState3 {
  int L1, g1, L2, g2, L3;
};
__device__ __forceinline__ void chunk_update(ChunkState3& s, int L, int g) {
  const bool c1 = L < s.L1, c2 = L < s.L2;
  s.L3 = min(s.L3, max(s.L2, L));
  s.g2 = c1 ? s.g1 : (c2 ? g : s.g2);
  s.L2 = min(s.L2, max(s.L1, L));
  s.g1 = c1 ? g : s.g1;
  s.L1 = min(s.L1, L);
}
__device__ __forceinline__ void chunk_init(ChunkState& s) { s.L1 = kBig; s.g1 = 0; s.L2 = kBig; }
__device__ __forceinline__ void chunk_init(ChunkState3& s) { s.L1 = kBig; s.g1 = 0; s.L2 = kBig; s.g2 = 0; s.L3 = kBig; }
// row states travel as int4: (L1, g1, L2, 0) / (L1, g1 | g2 << 16, L2, L3)
__device__ __forceinline__ int4 chunk_pack(const ChunkState& s) { return make_int4(s.L1, s.g1, s.L2, 0); }
__device__ __forceinline__ int4 chunk_pack(const ChunkState3& s) { return make_int4(s.L1, s.g1 | (s.g2 << 16), s.L2, s.L3); }
// merge the state another column quarter handed over (chunks of the quarters interleave tile by tile: order on (L, g))
__device__ __forceinline__ void chunk_merge(ChunkState& st, const int4 q);
__device__ __forceinline__ void chunk_merge(ChunkState3& st, const int4 q) {
  chunk_update(st, q.x, q.y & 0xffff);
  chunk_update(st, q.z, q.y >> 16);
  chunk_update(st, q.w, 0);          // third bound of the other quarter: can only land in L3 (q.w >= q.z >= q.x)
}

__global__ void __launch_bounds__(kThreads, 1)
chunk_sweep_kernel(const __grid_constant__ CUtensorMap tmap, const ChunkParams p) {
  extern __shared__ uint8_t smem_raw[];
  ChunkSmem& sm = *reinterpret_cast<ChunkSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int n_items = p.n_pairs * p.blocks_per_pair;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap);
    for (int i = 0; i < kAStages; ++i) { mbar_init(&sm.a_full[i], 1); mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kBStages; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < 4; ++i) { mbar_init(&sm.t_full[i], 1); mbar_init(&sm.t_empty[i], 4); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&sm.tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;

  // work item = (pair, block of 256 sorted query rows); rows and columns both come from the sorted bank
  auto decode = [&](int item, int& pair, int& blk, int& row0, int& col_img, int& n_rows, int& ntiles, int& nm) {
    pair = item / p.blocks_per_pair;
    blk = item - pair * p.blocks_per_pair;
    const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
    const int qi = pi.qi, ti = pi.ti, kq = pi.kq, kt = pi.kt;
    n_rows = min(kq - blk * kRows, kRows);
    col_img = ti;
    row0 = static_cast<int>(qi * p.k_stride + blk * kRows);
    ntiles = (kt + kTT - 1) / kTT;
    nm = n_rows > kTM ? 2 : 1;
    return n_rows > 0 && ntiles > 0;
  };

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int pair, blk, row0, col_img, n_rows, ntiles, nm;
        if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        mbar_wait(&sm.a_empty[as], aph ^ 1);
        mbar_arrive_expect_tx(&sm.a_full[as], nm * kMBytes);
        for (int m = 0; m < nm; ++m) tma_load_2d(sm.a[as][m], &tmap, 0, row0 + m * kTM, &sm.a_full[as]);
        ++a_it;
        const int64_t crow0 = static_cast<int64_t>(col_img) * p.k_stride;
        for (int t = 0; t < ntiles; ++t) {
          const int32_t crow = static_cast<int32_t>(crow0 + t * kTT);
          const uint32_t bs = b_it % kBStages, bph = (b_it / kBStages) & 1;
          mbar_wait(&sm.b_empty[bs], bph ^ 1);
          mbar_arrive_expect_tx(&sm.b_full[bs], kBBytes);
          tma_load_2d(sm.b[bs], &tmap, 0, crow, &sm.b_full[bs]);
          tma_load_2d(sm.b[bs] + kBBytes / 2, &tmap, 0, crow + 128, &sm.b_full[bs]);
          ++b_it;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_u8(kTM, kTN);
      uint32_t a_it = 0, b_it = 0, t_it[4] = {0, 0, 0, 0};
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int pair, blk, row0, col_img, n_rows, ntiles, nm;
        if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        mbar_wait(&sm.a_full[as], aph);
        for (int t = 0; t < ntiles; ++t) {
          const uint32_t bs = b_it % kBStages, bph = (b_it / kBStages) & 1;
          mbar_wait(&sm.b_full[bs], bph);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint64_t bdesc = make_desc_k128(smem_u32(sm.b[bs] + h * (kBBytes / 2)));
#pragma unroll
            for (int m = 0; m < 2; ++m) {
              if (m >= nm) continue;
              const int acc = m * 2 + h;
              mbar_wait(&sm.t_empty[acc], (t_it[acc] & 1) ^ 1);
              tc_fence_after();
              const uint64_t adesc = make_desc_k128(smem_u32(sm.a[as][m]));
              const uint32_t tacc = tmem_base + m * kTT + h * kTN;
#pragma unroll
              for (int k = 0; k < kD / 32; ++k) mma_i8_ss(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
              tc_commit(&sm.t_full[acc]);
              ++t_it[acc];
            }
          }
          tc_commit(&sm.b_empty[bs]);
          ++b_it;
        }
        tc_commit(&sm.a_empty[as]);
        ++a_it;
      }
    }
  } else if (warp >= kFirstEpiWarp) {
    // ------------------------------------------------------------- epilogue
    const uint32_t ew = warp - kFirstEpiWarp;       // 0..15
    const uint32_t quad = warp & 3;                 // TMEM lane quadrant this warp may read
    const int m = ew >> 3;                          // M-tile
    const int h = (ew >> 2) & 1;                    // column half
    const int acc = m * 2 + h;
    const int rl = m * kTM + quad * 32 + lane;      // row inside the item's block
    const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + m * kTT + h * kTN;
    uint32_t t_it = 0, item_it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      int pair, blk, row0, col_img, n_rows, ntiles, nm;
      if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
      const bool active = m < nm;
      const uint32_t par = item_it & 1;
      ++item_it;
      ChunkState st;
      st.L1 = kBig; st.g1 = 0; st.L2 = kBig;
      if (active) {
        // this thread's four chunk minima of tile t: chunks t * 8 + h * 4 .. + 3 of the column image
        const int4* cm = reinterpret_cast<const int4*>(p.cmin + (static_cast<int64_t>(col_img) * p.k_stride) / kChunk) + h;
        int4 cm_next = __ldg(cm);
        for (int t = 0; t < ntiles; ++t) {
          const int4 c4 = cm_next;
          if (t + 1 < ntiles) cm_next = __ldg(cm + 2 * (t + 1));
          const int g0 = t * kChunksPerTile + h * 4;
          mbar_wait(&sm.t_full[acc], t_it & 1);
          tc_fence_after();
          ++t_it;
          // software-pipelined drain in halves of 16 columns (two register buffers)
          int va[16], vb[16];
          tmem_ld_32x16(taddr, va);
          tmem_ld_fence16(va);
          tmem_ld_32x16(taddr + 16, vb);
          if (p.peak) {
            tmem_ld_fence16(vb);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.t_empty[acc]);
            chunk_update(st, c4.x - 2 * max17(vb, max16(va)), g0);
            continue;
          }
          const int cmv[4] = {c4.x, c4.y, c4.z, c4.w};
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int m0 = max16(va);
            tmem_ld_fence16(vb);
            if (c < 3) {
              tmem_ld_32x16(taddr + (c + 1) * 32, va);
            } else {
              // accumulator fully drained into registers: hand it back to the MMA warp
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&sm.t_empty[acc]);
            }
            const int m1 = max17(vb, m0);
            if (c < 3) {
              tmem_ld_fence16(va);
              tmem_ld_32x16(taddr + (c + 1) * 32 + 16, vb);
            }
            chunk_update(st, cmv[c] - 2 * m1, g0 + c);
          }
        }
      }
      // the two column halves of a row meet here: half 1 -> smem -> half 0 -> global
      if (h == 1 && active) sm.rowmerge[par][rl] = make_int4(st.L1, st.g1, st.L2, 0);
      named_bar_sync(kEpiBarrier, kEpiWarps * 32);
      if (h == 0 && active && rl < n_rows) {
        const int4 o = sm.rowmerge[par][rl];
        // chunks of the two halves interleave tile by tile: order the pair (L, g) explicitly
        const bool other = lex_less(o.x, o.y, st.L1, st.g1);
        const int L1 = other ? o.x : st.L1;
        const int g1 = other ? o.y : st.g1;
        const int L2 = min(min(o.z, st.L2), other ? st.L1 : o.x);
        p.rowres[static_cast<int64_t>(pair) * p.k_stride + blk * kRows + rl] = make_int4(L1, g1, L2, 0);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}


// ---------------------------------------------------------------------------------------------
// N = 256 variant: one tcgen05.mma covers a whole 256-column tile (M128 N256 K32), which needs 96 B
// of shared-memory operand reads per clock instead of the 128 B of two N = 128 halves. Two
// accumulators (one per M-tile of the row block) fill TMEM; the MMA of one M-tile overlaps the
// epilogue of the other. All 16 epilogue warps drain every accumulator: warp (quadrant q, column
// quarter c) owns rows 32q..32q+31 of BOTH M-tiles and chunks 2c, 2c+1 of every tile.
struct Chunk256Smem {
  uint8_t a[kAStages][2][kMBytes];
  uint8_t b[kBStages][kBBytes];
  int4 rowmerge[2][3][kRows];       // column quarters 1..3 hand their row states to quarter 0
  uint64_t a_full[kAStages], a_empty[kAStages];
  uint64_t b_full[kBStages], b_empty[kBStages];
  uint64_t t_full[2], t_empty[2];   // accumulator m
  uint32_t tmem_base;
};

// Registers are allocated per SM sub-partition: 5 warps each at 640 threads -> 96 per thread, enough
// for the 64-register drain buffer below. (setmaxnreg cannot help: the CTA's pool is what it was
// launched with, 5 x 96 per sub-partition.)
constexpr int kFirstEpiWarp256 = kFirstEpiWarp;
constexpr int kThreads256 = kThreads;

// 64 columns (two chunks) of one row, drained in halves of 16 columns through two register buffers so
// that the tcgen05.ld of the next half is in flight while the current one goes through the max tree.
// (Measured alternatives, all slower: pulling all 64 columns first and releasing the accumulator before
// the max trees, -17 %; two loads of 32 columns through two 32-register buffers, -5..9 %; issuing the first
// load of the NEXT accumulator before the last max tree of this one, -6 %.)
__device__ __forceinline__ void chunk_merge(ChunkState& st, const int4 q) {
  const bool other = lex_less(q.x, q.y, st.L1, st.g1);
  st.L2 = min(min(q.z, st.L2), other ? st.L1 : q.x);
  st.g1 = other ? q.y : st.g1;
  st.L1 = other ? q.x : st.L1;
}

template <bool kPeakOnly, int kChunkCols, class State>
__device__ __forceinline__ void drain64(uint32_t taddr, int cm0, int cm1, int g, State& st, uint64_t* t_empty,
                                        uint32_t lane) {
  int va[16], vb[16];
  tmem_ld_32x16(taddr, va);
  tmem_ld_fence16(va);
  if (kPeakOnly) {
    tc_fence_before();
    __syncwarp();
    if (lane == 0) mbar_arrive(t_empty);
    chunk_update(st, cm0 - 2 * max16(va), g);
    return;
  }
  tmem_ld_32x16(taddr + 16, vb);
  const int m0 = max16(va);
  tmem_ld_fence16(vb);
  tmem_ld_32x16(taddr + 32, va);
  const int m1 = max17(vb, m0);
  tmem_ld_fence16(va);
  tmem_ld_32x16(taddr + 48, vb);
  if (kChunkCols == 32) chunk_update(st, cm0 - 2 * m1, g);
  const int m2 = kChunkCols == 32 ? max16(va) : max17(va, m1);
  tmem_ld_fence16(vb);
  // this warp's part of the accumulator is in registers: hand it back to the MMA warp
  tc_fence_before();
  __syncwarp();
  if (lane == 0) mbar_arrive(t_empty);
  const int m3 = max17(vb, m2);
  if (kChunkCols == 32) chunk_update(st, cm1 - 2 * m3, g + 1);
  else chunk_update(st, cm0 - 2 * m3, g);       // sorted norms: the 64-column chunk's minimum is its first
}

template <bool kPeakOnly, int kChunkCols, bool kTop3 = false>
__global__ void __launch_bounds__(kThreads256, 1)
chunk_sweep256_kernel(const __grid_constant__ CUtensorMap tmap, const ChunkParams p) {
  using State = typename std::conditional<kTop3, ChunkState3, ChunkState>::type;
  extern __shared__ uint8_t smem_raw[];
  Chunk256Smem& sm = *reinterpret_cast<Chunk256Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int n_items = p.n_pairs * p.blocks_per_pair;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap);
    for (int i = 0; i < kAStages; ++i) { mbar_init(&sm.a_full[i], 1); mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kBStages; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.t_full[i], 1); mbar_init(&sm.t_empty[i], kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&sm.tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;

  auto decode = [&](int item, int& pair, int& blk, int& row0, int& col_img, int& n_rows, int& ntiles, int& nm) {
    pair = item / p.blocks_per_pair;
    blk = item - pair * p.blocks_per_pair;
    const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
    const int qi = pi.qi, ti = pi.ti, kq = pi.kq, kt = pi.kt;
    n_rows = min(kq - blk * kRows, kRows);
    col_img = ti;
    row0 = static_cast<int>(qi * p.k_stride + blk * kRows);
    ntiles = (kt + kTT - 1) / kTT;
    nm = n_rows > kTM ? 2 : 1;
    return n_rows > 0 && ntiles > 0;
  };

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int pair, blk, row0, col_img, n_rows, ntiles, nm;
        if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        mbar_wait_ahead(&sm.a_empty[as], aph ^ 1);
        mbar_arrive_expect_tx(&sm.a_full[as], nm * kMBytes);
        for (int m = 0; m < nm; ++m) tma_load_2d(sm.a[as][m], &tmap, 0, row0 + m * kTM, &sm.a_full[as]);
        ++a_it;
        const int64_t crow0 = static_cast<int64_t>(col_img) * p.k_stride;
        for (int t = 0; t < ntiles; ++t) {
          const int32_t crow = static_cast<int32_t>(crow0 + t * kTT);
          const uint32_t bs = b_it % kBStages, bph = (b_it / kBStages) & 1;
          mbar_wait_ahead(&sm.b_empty[bs], bph ^ 1);
          mbar_arrive_expect_tx(&sm.b_full[bs], kBBytes);
          tma_load_2d(sm.b[bs], &tmap, 0, crow, &sm.b_full[bs]);
          tma_load_2d(sm.b[bs] + kBBytes / 2, &tmap, 0, crow + 128, &sm.b_full[bs]);
          ++b_it;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_u8(kTM, kTT);
      uint32_t a_it = 0, b_it = 0, t_it0 = 0, t_it1 = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int pair, blk, row0, col_img, n_rows, ntiles, nm;
        if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        mbar_wait(&sm.a_full[as], aph);
        const uint64_t adesc0 = make_desc_k128(smem_u32(sm.a[as][0]));
        const uint64_t adesc1 = make_desc_k128(smem_u32(sm.a[as][1]));
        for (int t = 0; t < ntiles; ++t) {
          const uint32_t bs = b_it % kBStages, bph = (b_it / kBStages) & 1;
          mbar_wait(&sm.b_full[bs], bph);
          const uint64_t bdesc = make_desc_k128(smem_u32(sm.b[bs]));
          mbar_wait(&sm.t_empty[0], (t_it0 & 1) ^ 1);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < kD / 32; ++k) mma_i8_ss(tmem_base, adesc0 + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
          tc_commit(&sm.t_full[0]);
          ++t_it0;
          if (nm > 1) {
            mbar_wait(&sm.t_empty[1], (t_it1 & 1) ^ 1);
            tc_fence_after();
#pragma unroll
            for (int k = 0; k < kD / 32; ++k)
              mma_i8_ss(tmem_base + kTT, adesc1 + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
            tc_commit(&sm.t_full[1]);
            ++t_it1;
          }
          tc_commit(&sm.b_empty[bs]);
          ++b_it;
        }
        tc_commit(&sm.a_empty[as]);
        ++a_it;
      }
    }
  } else if (warp >= kFirstEpiWarp256) {
    // ------------------------------------------------------------- epilogue
    const uint32_t quad = warp & 3;                   // TMEM lane quadrant this warp may read
    const int cq = (warp - kFirstEpiWarp256) >> 2;       // column quarter: chunks 2 cq, 2 cq + 1 of every tile
    const int rl = quad * 32 + lane;                  // row inside an M-tile
    const uint32_t taddr0 = tmem_base + ((quad * 32u) << 16) + cq * 64;
    uint32_t t_it0 = 0, t_it1 = 0, item_it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      int pair, blk, row0, col_img, n_rows, ntiles, nm;
      if (!decode(item, pair, blk, row0, col_img, n_rows, ntiles, nm)) continue;
      const uint32_t par = item_it & 1;
      ++item_it;
      State s0, s1;
      chunk_init(s0);
      chunk_init(s1);
      const int2* cm = reinterpret_cast<const int2*>(p.cmin + (static_cast<int64_t>(col_img) * p.k_stride) / kChunk) + cq;
      int2 cm_next = __ldg(cm);
      for (int t = 0; t < ntiles; ++t) {
        const int2 c2 = cm_next;
        if (t + 1 < ntiles) cm_next = __ldg(cm + 4 * (t + 1));
        const int g = kChunkCols == 32 ? t * kChunksPerTile + cq * 2 : t * (kTT / 64) + cq;
        mbar_wait(&sm.t_full[0], t_it0 & 1);
        tc_fence_after();
        ++t_it0;
        drain64<kPeakOnly, kChunkCols, State>(taddr0, c2.x, c2.y, g, s0, &sm.t_empty[0], lane);
        if (nm > 1) {
          mbar_wait(&sm.t_full[1], t_it1 & 1);
          tc_fence_after();
          ++t_it1;
          drain64<kPeakOnly, kChunkCols, State>(taddr0 + kTT, c2.x, c2.y, g, s1, &sm.t_empty[1], lane);
        }
      }
      // the four column quarters of a row meet here: quarters 1..3 -> smem -> quarter 0 -> global
      if (cq > 0) {
        sm.rowmerge[par][cq - 1][rl] = chunk_pack(s0);
        sm.rowmerge[par][cq - 1][kTM + rl] = chunk_pack(s1);
      }
      named_bar_sync(kEpiBarrier, kEpiWarps * 32);
      if (cq == 0) {
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          const int row = m * kTM + rl;
          if (row >= n_rows) continue;
          State st = m == 0 ? s0 : s1;
#pragma unroll
          for (int o = 0; o < 3; ++o) chunk_merge(st, sm.rowmerge[par][o][row]);
          p.rowres[static_cast<int64_t>(pair) * p.k_stride + blk * kRows + row] = chunk_pack(st);
        }
      }
    }
  }


  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc<512>(tmem_base);
  }
}

}  // namespace b200m
