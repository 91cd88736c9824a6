// The sweep kernel: exhaustive squared-L2 distances between the query and train descriptors of
// every scheduled image pair on tcgen05 int8 tensor cores, with the 2-NN (per query row) and the
// nearest-query (per train column) reductions fused into the TMEM epilogue, so the Kq x Kt
// distance matrix never leaves the SM.
//
// Stands in for the dense contraction + arg-best passes of the reference matcher head
//   lightglue/lightglue.py:223  sim = einsum('bmd,bnd->bmn')      (dense contraction)
//   lightglue/lightglue.py:237  scores.max(2), scores.max(1)      (row / column arg-best)
// in the exact-integer L2 form BASELINE.json names (cv2.BFMatcher NORM_L2 knnMatch k=2 + crossCheck).
//
// One CTA = one SM, persistent over work items (pair, query tile of 128 rows). For an item the CTA
// keeps the 128 x 128 B query tile resident in shared memory and sweeps the train image in tiles of
// 256 rows:
//   warp 0      TMA producer   (query tile per item, train tiles through a 5-deep ring)
//   warp 1      MMA issuer     (4 x tcgen05.mma.kind::i8 M128 N256 K32 per train tile -> TMEM)
//   warp 2      TMEM allocator
//   warps 4-11  epilogue       (tcgen05.ld -> registers; row top-2 stays in registers for the
//                               whole sweep, column minima go out as packed 64-bit atomicMin)
// Two 128 x 256 int32 accumulators (all 512 TMEM columns) double-buffer MMA against the epilogue.
#pragma once
#include "ptx_sm100.cuh"

namespace b200m {

constexpr int kD = 128;             // descriptor bytes
constexpr int kTQ = 128;            // query rows per tile  (UMMA M)
constexpr int kTT = 256;            // train rows per tile  (UMMA N)
constexpr int kABytes = kTQ * kD;   // 16 KiB
constexpr int kBBytes = kTT * kD;   // 32 KiB
constexpr int kAStages = 2;
constexpr int kBStages = 5;
constexpr int kAccStages = 2;       // TMEM accumulator buffers of 256 columns
constexpr int kEpiWarps = 8;
constexpr int kFirstEpiWarp = 4;
constexpr int kThreads = (kFirstEpiWarp + kEpiWarps) * 32;  // 384
constexpr int kPadNorm = 1 << 23;
constexpr int kIntMax = 0x7fffffff;

struct SweepSmem {
  // tiles first: SWIZZLE_128B needs 1024 B alignment
  uint8_t a[kAStages][kABytes];
  uint8_t b[kBStages][kBBytes];
  uint64_t a_full[kAStages], a_empty[kAStages];
  uint64_t b_full[kBStages], b_empty[kBStages];
  uint64_t t_full[kAccStages], t_empty[kAccStages];
  uint32_t tmem_base;
  uint32_t pad_[3];
  int32_t nbw[kEpiWarps][2][128];      // per epilogue warp: train norms of its 128 columns, double-buffered
  int32_t nbmin[kEpiWarps][2][4];      // per 32-column chunk minimum of the above
  int4 rowmerge[2][kTQ];               // half-1 warps hand their row top-2 to half-0 warps
};

struct SweepParams {
  const int32_t* sqnorm;        // [n_img, k_stride]
  const int32_t* k_per_image;   // [n_img]
  const int32_t* pairs;         // [n_pairs, 2] for this batch
  int64_t k_stride;
  int32_t n_pairs;
  int32_t qtiles;               // k_stride / 128
  int4* rowres;                 // [n_pairs, k_stride] (d1^2, i1, d2^2, i2)
  unsigned long long* colbest;  // [n_pairs, k_stride] (d^2 << 32) | query row
};

struct ItemInfo {
  int pair, qt, qi, ti, kq, kt, ntiles;
  bool valid;
};

__device__ __forceinline__ ItemInfo decode_item(const SweepParams& p, int item) {
  ItemInfo it;
  it.pair = item / p.qtiles;
  it.qt = item - it.pair * p.qtiles;
  it.qi = __ldg(p.pairs + 2 * it.pair);
  it.ti = __ldg(p.pairs + 2 * it.pair + 1);
  it.kq = __ldg(p.k_per_image + it.qi);
  it.kt = __ldg(p.k_per_image + it.ti);
  it.ntiles = (it.kt + kTT - 1) / kTT;
  it.valid = (it.qt * kTQ < it.kq) && (it.ntiles > 0);
  return it;
}

// Exact insertion of (s, n) into a row's running top-2, n strictly increasing within a thread:
// strict '<' keeps the lowest index on ties, and a duplicate of the best becomes the second best
// (the stable-argsort rule cv2.BFMatcher follows).
__device__ __forceinline__ void top2_insert(int s, int n, int& s1, int& i1, int& s2, int& i2) {
  const bool lt1 = s < s1;
  const bool lt2 = s < s2;
  s2 = lt1 ? s1 : (lt2 ? s : s2);
  i2 = lt1 ? i1 : (lt2 ? n : i2);
  s1 = lt1 ? s : s1;
  i1 = lt1 ? n : i1;
}

__device__ __forceinline__ bool lex_less(int sa, int ia, int sb, int ib) {
  return sa < sb || (sa == sb && ia < ib);
}

// Transposing min-reduction: on entry lane r holds v[j] = key(row r, column j), j < 32;
// on exit lane L holds in v[0] the minimum over the warp's 32 rows for column L.
__device__ __forceinline__ void warp_colmin_32(int (&v)[32], uint32_t lane) {
#pragma unroll
  for (int off = 16; off >= 1; off >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int j = 0; j < off; ++j) {
      const int send = up ? v[j] : v[j + off];
      const int keep = up ? v[j + off] : v[j];
      const int recv = __shfl_xor_sync(0xffffffffu, send, off);
      v[j] = min(keep, recv);
    }
  }
}

// Publish the warp's column minima for 32 columns. kc = (na - 2 a.b) * 128 + local query row.
__device__ __forceinline__ void publish_colmin(int kc, int nb_col, int n_col, int kt, int qrow0,
                                               unsigned long long* colbest_pair) {
  const int score = kc >> 7;
  if (n_col < kt && score < kPadNorm) {
    const unsigned long long key =
        (static_cast<unsigned long long>(static_cast<unsigned>(score + nb_col)) << 32) |
        static_cast<unsigned>(qrow0 + (kc & 127));
    unsigned long long* dst = colbest_pair + n_col;
    const unsigned long long cur = *reinterpret_cast<volatile unsigned long long*>(dst);
    if (key < cur) atomicMin(dst, key);
  }
}

template <int kMode>
__global__ void __launch_bounds__(kThreads, 1)
sweep_kernel(const __grid_constant__ CUtensorMap tmap, const SweepParams p) {
  extern __shared__ uint8_t smem_raw[];
  SweepSmem& sm = *reinterpret_cast<SweepSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int n_items = p.n_pairs * p.qtiles;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap);
    for (int i = 0; i < kAStages; ++i) { mbar_init(&sm.a_full[i], 1); mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kBStages; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < kAccStages; ++i) { mbar_init(&sm.t_full[i], 1); mbar_init(&sm.t_empty[i], kEpiWarps); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&sm.tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const ItemInfo it = decode_item(p, item);
        if (!it.valid) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        mbar_wait(&sm.a_empty[as], aph ^ 1);
        mbar_arrive_expect_tx(&sm.a_full[as], kABytes);
        tma_load_2d(sm.a[as], &tmap, 0, static_cast<int32_t>(it.qi * p.k_stride + it.qt * kTQ), &sm.a_full[as]);
        ++a_it;
        const int32_t trow0 = static_cast<int32_t>(it.ti * p.k_stride);
        for (int t = 0; t < it.ntiles; ++t) {
          const uint32_t bs = b_it % kBStages, bph = (b_it / kBStages) & 1;
          mbar_wait(&sm.b_empty[bs], bph ^ 1);
          mbar_arrive_expect_tx(&sm.b_full[bs], kBBytes);
          tma_load_2d(sm.b[bs], &tmap, 0, trow0 + t * kTT, &sm.b_full[bs]);
          tma_load_2d(sm.b[bs] + kBBytes / 2, &tmap, 0, trow0 + t * kTT + 128, &sm.b_full[bs]);
          ++b_it;
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_u8(kTQ, kTT);
      uint32_t a_it = 0, b_it = 0, t_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const ItemInfo it = decode_item(p, item);
        if (!it.valid) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        mbar_wait(&sm.a_full[as], aph);
        const uint64_t adesc = make_desc_k128(smem_u32(sm.a[as]));
        for (int t = 0; t < it.ntiles; ++t) {
          const uint32_t bs = b_it % kBStages, bph = (b_it / kBStages) & 1;
          const uint32_t ts = t_it % kAccStages, tph = (t_it / kAccStages) & 1;
          mbar_wait(&sm.b_full[bs], bph);
          mbar_wait(&sm.t_empty[ts], tph ^ 1);
          tc_fence_after();
          const uint64_t bdesc = make_desc_k128(smem_u32(sm.b[bs]));
          const uint32_t tacc = tmem_base + ts * kTT;
#pragma unroll
          for (int k = 0; k < kD / 32; ++k) {
            // +32 B along K inside the 128 B swizzle atom = +2 in the (addr >> 4) field
            mma_i8_ss(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
          }
          tc_commit(&sm.b_empty[bs]);   // smem stage free once these MMAs retire
          tc_commit(&sm.t_full[ts]);    // accumulator ready for the epilogue
          ++b_it;
          ++t_it;
        }
        tc_commit(&sm.a_empty[as]);
        ++a_it;
      }
    }
  } else if (warp >= kFirstEpiWarp) {
    // ------------------------------------------------------------- epilogue
    const uint32_t ew = warp - kFirstEpiWarp;       // 0..7
    const uint32_t quad = warp & 3;                 // TMEM lane quadrant this warp may read
    const uint32_t half = ew >> 2;                  // which 128 of the 256 tile columns
    const int rl = quad * 32 + lane;                // query row inside the tile
    uint32_t t_it = 0, item_it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const ItemInfo it = decode_item(p, item);
      if (!it.valid) continue;
      const int qrow0 = it.qt * kTQ;
      const int32_t* nb_img = p.sqnorm + it.ti * p.k_stride;
      unsigned long long* colbest_pair = p.colbest + static_cast<int64_t>(it.pair) * p.k_stride;
      const int na = __ldg(p.sqnorm + it.qi * p.k_stride + qrow0 + rl);   // kPadNorm on padding rows
      const int rowc = na * 128 + rl;
      int s1 = kIntMax, i1 = kIntMax, s2 = kIntMax, i2 = kIntMax;        // scores are nb - 2 a.b

      // stage the norms of tile 0
      {
        const int4 nbv = __ldg(reinterpret_cast<const int4*>(nb_img + half * 128) + lane);
        *reinterpret_cast<int4*>(&sm.nbw[ew][0][lane * 4]) = nbv;
        int m = min(min(nbv.x, nbv.y), min(nbv.z, nbv.w));
        m = min(m, __shfl_xor_sync(0xffffffffu, m, 1));
        m = min(m, __shfl_xor_sync(0xffffffffu, m, 2));
        m = min(m, __shfl_xor_sync(0xffffffffu, m, 4));
        if ((lane & 7) == 0) sm.nbmin[ew][0][lane >> 3] = m;
        __syncwarp();
      }

      for (int t = 0; t < it.ntiles; ++t) {
        const uint32_t ts = t_it % kAccStages, tph = (t_it / kAccStages) & 1;
        const int nbuf = t & 1;
        // prefetch next tile's norms into registers while this tile is processed
        int4 nbv_next = make_int4(kPadNorm, kPadNorm, kPadNorm, kPadNorm);
        if (t + 1 < it.ntiles)
          nbv_next = __ldg(reinterpret_cast<const int4*>(nb_img + (t + 1) * kTT + half * 128) + lane);

        mbar_wait(&sm.t_full[ts], tph);
        tc_fence_after();
        const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + ts * kTT + half * 128;
        const int col0 = t * kTT + half * 128;
        const int32_t* nbw = sm.nbw[ew][nbuf];

#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          int v[32];
          tmem_ld_32x32(taddr + c * 32, v);
          tmem_ld_wait();
          if (c == 3) {
            // accumulator fully drained into registers: hand the TMEM buffer back to the MMA warp
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.t_empty[ts]);
          }
          const int nbase = col0 + c * 32;
          if (kMode == 0) {
            // exact reference epilogue: every element goes through the top-2 insertion
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const int s = nbw[c * 32 + j] - 2 * v[j];
              top2_insert(s, nbase + j, s1, i1, s2, i2);
            }
          } else {
            // threshold-reject epilogue: s < s2  <=>  2*acc > nb - s2; a chunk can only change the
            // row's top-2 if 2*max(acc) + s2 > min(nb over the chunk). One 3-input max per two
            // elements, the exact insertion runs only for chunks that pass.
            int mx = v[0];
#pragma unroll
            for (int j = 1; j < 31; j += 2) mx = __vimax3_s32(mx, v[j], v[j + 1]);
            mx = max(mx, v[31]);
            const int nbm = sm.nbmin[ew][nbuf][c];
            // s2 may be kIntMax: 2*mx + s2 would overflow, compare in 64 bit only on that path
            const bool hit = (s2 == kIntMax) || (2 * mx + s2 > nbm);
            if (hit) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const int s = nbw[c * 32 + j] - 2 * v[j];
                if (s < s2) top2_insert(s, nbase + j, s1, i1, s2, i2);
              }
            }
          }
          // column minima over this warp's 32 rows
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = rowc - (v[j] << 8);
          warp_colmin_32(v, lane);
          publish_colmin(v[0], nbw[c * 32 + lane], nbase + lane, it.kt, qrow0, colbest_pair);
        }

        // stage next tile's norms (other buffer; this warp is the only reader/writer)
        if (t + 1 < it.ntiles) {
          *reinterpret_cast<int4*>(&sm.nbw[ew][nbuf ^ 1][lane * 4]) = nbv_next;
          int m = min(min(nbv_next.x, nbv_next.y), min(nbv_next.z, nbv_next.w));
          m = min(m, __shfl_xor_sync(0xffffffffu, m, 1));
          m = min(m, __shfl_xor_sync(0xffffffffu, m, 2));
          m = min(m, __shfl_xor_sync(0xffffffffu, m, 4));
          if ((lane & 7) == 0) sm.nbmin[ew][nbuf ^ 1][lane >> 3] = m;
        }
        __syncwarp();
        ++t_it;
      }

      // merge the two column halves of every row and store the row result
      const uint32_t par = item_it & 1;
      if (half == 1) sm.rowmerge[par][rl] = make_int4(s1, i1, s2, i2);
      named_bar_sync(1, kEpiWarps * 32);
      if (half == 0) {
        const int4 o = sm.rowmerge[par][rl];
        // half 0 holds lower column indices inside each tile but not globally: merge lexicographically
        int r1s, r1i, r2s, r2i;
        if (lex_less(o.x, o.y, s1, i1)) {
          r1s = o.x; r1i = o.y;
          if (lex_less(o.z, o.w, s1, i1)) { r2s = o.z; r2i = o.w; } else { r2s = s1; r2i = i1; }
        } else {
          r1s = s1; r1i = i1;
          if (lex_less(o.x, o.y, s2, i2)) { r2s = o.x; r2i = o.y; } else { r2s = s2; r2i = i2; }
        }
        if (qrow0 + rl < it.kq) {
          const int d1 = (r1s == kIntMax) ? kIntMax : r1s + na;
          const int d2 = (r2s == kIntMax) ? kIntMax : r2s + na;
          p.rowres[static_cast<int64_t>(it.pair) * p.k_stride + qrow0 + rl] = make_int4(d1, r1i, d2, r2i);
        }
      }
      ++item_it;
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
