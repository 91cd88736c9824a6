// The sweep kernel: exhaustive squared-L2 distances between two descriptor sets on tcgen05 int8
// tensor cores with the nearest-neighbour reductions fused into the TMEM epilogue, so the distance
// matrix never leaves the SM.
//
// Stands in for the dense contraction + arg-best passes of the reference matcher head
//   lightglue/lightglue.py:223      sim = einsum('bmd,bnd->bmn')     (dense contraction)
//   lightglue/lightglue.py:237      scores.max(2), scores.max(1)     (row / column arg-best)
//   lightglue/lightglue.py:239-242  mutual check
// in the exact-integer L2 form BASELINE.json names (cv2.BFMatcher NORM_L2 knnMatch k=2 + crossCheck).
//
// One CTA = one SM, persistent over work items. An item keeps a block of 256 "row" descriptors
// (two UMMA M-tiles of 128) resident in shared memory and sweeps a "column" image in tiles of 256
// columns, each tile issued as two N=128 halves:
//   warp 0       TMA producer  (row block per item; column tiles + their norms through rings)
//   warp 1       MMA issuer    (per column tile: 4 accumulators x 4 tcgen05.mma.kind::i8 M128 N128 K32)
//   warp 2       TMEM allocator
//   warps 4-19   epilogue: 4 warps (one per TMEM lane quadrant) for each accumulator (m, h),
//                m = M-tile, h = column half. TMEM columns of (m, h): m*256 + h*128 .. +127.
// With four 128 x 128 int32 accumulators (all 512 TMEM columns) the MMA of one column half overlaps
// the epilogue of the other half of the same rows, and every column tile fetched from L2 serves
// 256 rows. Each epilogue thread owns one row (tcgen05.ld 32x32b: lane == row) and half of the
// columns for the whole sweep; the two halves of a row meet once per item through shared memory.
//
// Several passes share the kernel (kKind):
//   kTop1    rows = query descriptors of a pair, columns = train descriptors: exact nearest column
//            per row (lowest index on ties) + an upper bound on the second-nearest distance.
//            A threshold-reject epilogue does one 3-input integer max per two accumulator values and
//            looks at exact scores only for 4-column groups that can beat the row's running best.
//   kTop2    same sweep with the exact second-nearest as well (knnMatch k=2 output).
//   kVerify  lazy checks on the few candidate rows that survive the first sweep, both of the form
//            "does any column beat this row's threshold":
//              ratio  rows = candidate query descriptors (gathered), columns = the train image:
//                     violated when another train row lies inside the Lowe-ratio radius;
//              mutual rows = the candidates' train descriptors (gathered), columns = the query
//                     image: violated when some query is nearer (or equally near, lower index).
//   kTop1G   kTop1 over GATHERED query rows (the exact fallback of the chunk pipeline): rows come from a
//            gathered buffer, results go to the rows' original slots in (d1^2, i1, d2^2 bound, state) form.
//   kTop2G   kTop2 over gathered query rows (the exact fallback of the chunk-based knnMatch, match_knn.cuh):
//            results go to the rows' original slots in (d1^2, i1, d2^2, i2) form.
//   kTop2Exact / kPeak  debugging (every element through the insertion) and MMA-only ceiling.
#pragma once
#include "match_common.cuh"
#include "ptx_sm100.cuh"

namespace b200m {

constexpr int kD = 128;             // descriptor bytes
constexpr int kTM = 128;            // rows per UMMA M-tile
constexpr int kRows = 256;          // rows per work item (two M-tiles)
constexpr int kTT = 256;            // columns per tile
constexpr int kTN = 128;            // columns per MMA / accumulator (UMMA N)
constexpr int kMBytes = kTM * kD;   // 16 KiB per M-tile
constexpr int kBBytes = kTT * kD;   // 32 KiB per column tile
constexpr int kAStages = 2;
constexpr int kBStages = 4;
constexpr int kNStages = 4;         // norms ring
constexpr int kEpiWarps = 16;
constexpr int kFirstEpiWarp = 4;
constexpr int kThreads = (kFirstEpiWarp + kEpiWarps) * 32;  // 640
constexpr int kPadNorm = 1 << 23;   // > 128*255^2: norm of a padding row
constexpr int kBig = 1 << 28;       // "no candidate yet" score; 2*acc + kBig cannot overflow
constexpr uint32_t kEpiBarrier = 1; // named barrier of the 16 epilogue warps

enum SweepKind { kTop2 = 0, kVerify = 1, kPeak = 2, kTop2Exact = 3, kTop1 = 4, kTop1G = 5, kTop2G = 6 };

__host__ __device__ constexpr bool is_top1(int kind) { return kind == kTop1 || kind == kTop1G; }
__host__ __device__ constexpr bool is_top2(int kind) { return kind == kTop2 || kind == kTop2G; }
__host__ __device__ constexpr bool is_gathered(int kind) { return kind == kTop1G || kind == kTop2G; }

struct NormSlot {
  int32_t nb[kTT];      // squared norms of the tile's 256 columns
  int32_t nbmin[8];     // minimum per 32-column chunk
};

struct SweepSmem {
  // tiles first: SWIZZLE_128B needs 1024 B alignment
  uint8_t a[kAStages][2][kMBytes];
  uint8_t b[kBStages][kBBytes];
  NormSlot norms[kNStages];
  int4 rowmerge[2][kRows];          // column half 1 hands its row state to column half 0
  int2 share[2][2][kRows];          // [item parity][h][row]: running (s1, s2) published to the other half
  uint64_t a_full[kAStages], a_empty[kAStages];
  uint64_t b_full[kBStages], b_empty[kBStages];
  uint64_t n_full[kNStages], n_empty[kNStages];
  uint64_t t_full[4], t_empty[4];   // accumulator (m, h) -> index m * 2 + h
  uint32_t tmem_base;
};

struct SweepParams {
  const int32_t* sqnorm;        // [n_img * k_stride] squared norms (kPadNorm on padding rows)
  const int32_t* nbmin32;       // [n_img * k_stride / 32] minima of the above per 32 rows
  const int32_t* k_per_image;   // [n_img]
  const int32_t* pairs;         // [n_pairs, 2] (query image, train image) of this batch
  int64_t k_stride;
  int32_t n_pairs;
  int32_t blocks_per_pair;      // k_stride / 256
  int32_t n_img;                // pair-list entries outside [0, n_img) are skipped (load_pair)
  // kTop1 / kTop2 output per (pair, query row): (d1^2, i1, d2^2, i2); for kTop1 d2^2 is an upper
  // bound of the second-nearest squared distance and i2 is unused
  int4* rowres;
  // kVerify: candidates of the first sweep, slot s of pair p lives at p * k_stride + s
  const int32_t* cand_count;      // [n_pairs] candidates of the mutual check
  const int32_t* cand_count_r;    // [n_pairs] candidates of the ratio check (same list in pipeline 1)
  const int32_t* cand_q;          // query row of the candidate
  const int32_t* cand_t;          // its nearest train row
  const int32_t* cand_excl;       // ratio check: the train row that is allowed inside the radius
  const int32_t* cand_thr_ratio;  // ratio check: a train column with score < this kills the row
  const int32_t* cand_thr_mut;    // mutual check: d1^2 - |t1|^2, the score a query must beat
  int32_t* viol_ratio;            // out: 1 when the ratio test fails
  int32_t* viol_mut;              // out: 1 when a better (or equal, lower-index) query exists
  int32_t do_ratio, do_mutual;    // which of the two checks to run
  const int32_t* vitems;          // optional compact work list [n][2] = (pair, blk * 2 + check); NULL = enumerate all
  const int32_t* vitem_count;     // [1] entries of vitems
  // kTop1G: rows gathered by the refine kernel, slot s of pair p at p * k_stride + s
  const int32_t* fb_count;        // [n_pairs]
  const int32_t* fb_total;        // [1] 0 = no row of the batch fell back: the launch has nothing to do
  const int32_t* fb_q;            // original query row of the gathered row
  int4* rowres_o;                 // out, original query rows: (d1^2, i1, d2^2 bound, kRowFallback)
};

struct ItemInfo {
  int pair, blk;
  int row_img_row0;   // first A row in tensor-map A coordinates
  int col_img;        // image whose rows are the columns
  int n_rows;         // valid rows in this block (1..256)
  int n_cols;         // valid columns in the sweep
  int ntiles, nm;
  int check;          // kVerify: 0 = ratio, 1 = mutual
  bool valid;
};

template <int kKind>
__device__ __forceinline__ ItemInfo decode_item(const SweepParams& p, int item) {
  ItemInfo it;
  it.check = 0;
  if (kKind == kVerify && p.vitems) {
    // compact work list written by select2_kernel: (pair, blk * 2 + check) of the non-empty blocks only
    const int2 w = __ldg(reinterpret_cast<const int2*>(p.vitems) + item);
    it.pair = w.x;
    it.blk = w.y >> 1;
    it.check = w.y & 1;
  } else {
    if (kKind == kVerify) {
      it.check = item & 1;
      item >>= 1;
    }
    if (is_gathered(kKind)) {
      // block-major: the few non-empty items of a fallback sweep (block 0 of most pairs, seldom more) are
      // consecutive, so the persistent CTAs share them evenly (pair-major put them 64 apart: 37 of 148 CTAs)
      it.blk = item / p.n_pairs;
      it.pair = item - it.blk * p.n_pairs;
    } else {
      it.pair = item / p.blocks_per_pair;
      it.blk = item - it.pair * p.blocks_per_pair;
    }
  }
  const PairInfo pi = load_pair(p.pairs, p.k_per_image, it.pair, p.n_img, p.k_stride);
  const int qi = pi.qi, ti = pi.ti, kq = pi.kq, kt = pi.kt;
  if (kKind == kVerify) {
    const bool on = it.check == 0 ? p.do_ratio != 0 : p.do_mutual != 0;
    const int s = on ? __ldg((it.check == 0 ? p.cand_count_r : p.cand_count) + it.pair) : 0;
    it.n_rows = s - it.blk * kRows;
    it.n_cols = it.check == 0 ? kt : kq;
    it.col_img = it.check == 0 ? ti : qi;
    it.row_img_row0 = static_cast<int>(it.pair * p.k_stride + it.blk * kRows);   // gathered buffers
  } else if (is_gathered(kKind)) {
    it.n_rows = __ldg(p.fb_count + it.pair) - it.blk * kRows;
    it.n_cols = kt;
    it.col_img = ti;
    it.row_img_row0 = static_cast<int>(it.pair * p.k_stride + it.blk * kRows);   // gathered buffer
  } else {
    it.n_rows = kq - it.blk * kRows;
    it.n_cols = kt;
    it.col_img = ti;
    it.row_img_row0 = static_cast<int>(qi * p.k_stride + it.blk * kRows);        // bank
  }
  if (it.n_rows > kRows) it.n_rows = kRows;
  it.ntiles = (it.n_cols + kTT - 1) / kTT;
  it.nm = it.n_rows > kTM ? 2 : 1;
  it.valid = it.n_rows > 0 && it.ntiles > 0;
  return it;
}

// Exact insertion of (s, n), s < s2, into a row's running top-2. Within a thread n is strictly
// increasing, so strict '<' keeps the lowest index on ties and a duplicate of the best becomes the
// second best (the stable-argsort rule cv2.BFMatcher follows).
__device__ __forceinline__ void top2_insert(int s, int n, int& s1, int& i1, int& s2, int& i2) {
  const bool lt1 = s < s1;
  s2 = lt1 ? s1 : s;
  i2 = lt1 ? i1 : n;
  s1 = lt1 ? s : s1;
  i1 = lt1 ? n : i1;
}

__device__ __forceinline__ bool lex_less(int sa, int ia, int sb, int ib) {
  return sa < sb || (sa == sb && ia < ib);
}

__device__ __forceinline__ int max4(const int* v) { return max(__vimax3_s32(v[0], v[1], v[2]), v[3]); }

// Per-row epilogue state. Scores are  |col|^2 - 2 row.col  (the row's own norm is added at the end).
struct RowState {
  int s1, i1, s2, i2;    // running best / second best (score, column); kTop1: s2 is only an upper bound
  int lim;               // kTop1 / kTop2: only scores < lim can matter (own state and the other column half)
  int2* pub;             // where this thread publishes (s1, s2) for the other half
  int thr, excl, tie;    // kVerify: violated by a column != excl with score < thr, or == thr and column < tie
  int viol;
};

__device__ __forceinline__ int4 lds_v4(uint32_t saddr) {
  int4 r;
  asm volatile("ld.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(saddr));
  return r;
}
__device__ __forceinline__ int lds_s32(uint32_t saddr) {
  int r;
  asm volatile("ld.shared.s32 %0, [%1];" : "=r"(r) : "r"(saddr));
  return r;
}

// Sixteen columns of one row. v = raw int32 dot products, nb_s = shared-memory address of the
// columns' squared norms, nbm = a lower bound of those norms (minimum over the enclosing 32-column
// chunk), nbase = index of the first column.
template <int kKind>
__device__ __forceinline__ void process16(const int (&v)[16], uint32_t nb_s, int nbm, int nbase, RowState& st) {
  if (kKind == kTop2Exact) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int4 nb4 = lds_v4(nb_s + 16 * k);
      const int sv[4] = {nb4.x - 2 * v[4 * k], nb4.y - 2 * v[4 * k + 1], nb4.z - 2 * v[4 * k + 2], nb4.w - 2 * v[4 * k + 3]};
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (sv[j] < st.s2) top2_insert(sv[j], nbase + 4 * k + j, st.s1, st.i1, st.s2, st.i2);
    }
    return;
  }
  // Threshold-reject main path, shared by all kinds:  score < lim  <=>  2*acc > nb - lim  <=>
  // acc > floor((nb - lim) / 2).  A group of columns can only matter if
  // max(acc) > floor((min(nb) - lim) / 2): one 3-input max per two accumulator values, one compare
  // per 4 columns of a half-chunk that passes, exact scores only for the groups that pass.
  int g[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) g[k] = max4(v + 4 * k);
  const int mx = max(__vimax3_s32(g[0], g[1], g[2]), g[3]);
  if (kKind == kVerify) {
    // violated  <=>  exists column != excl with s < thr, or with s == thr and column index < tie
    //           =>   2*acc >= nb - thr  =>  acc > floor((min(nb) - thr - 1) / 2)
    const int T = (nbm - st.thr - 1) >> 1;
    if (mx > T) {
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (g[k] > T) {
          const int4 nb4 = lds_v4(nb_s + 16 * k);
          const int sv[4] = {nb4.x - 2 * v[4 * k], nb4.y - 2 * v[4 * k + 1], nb4.z - 2 * v[4 * k + 2],
                             nb4.w - 2 * v[4 * k + 3]};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int col = nbase + 4 * k + j;
            if ((sv[j] < st.thr && col != st.excl) || (sv[j] == st.thr && col < st.tie)) st.viol = 1;
          }
        }
      }
    }
    return;
  }
  int T = (nbm - st.lim) >> 1;
  if (mx > T) {
    const int lim_p = st.lim;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (g[k] > T) {
        const int4 nb4 = lds_v4(nb_s + 16 * k);
        if (is_top1(kKind)) {
          // packed (score << 2 | j) minimum: IMADs on the FMA pipe, two integer minima on the ALU pipe
          const int ka = (nb4.x * 4 + 0) - 8 * v[4 * k], kb = (nb4.y * 4 + 1) - 8 * v[4 * k + 1];
          const int kc = (nb4.z * 4 + 2) - 8 * v[4 * k + 2], kd = (nb4.w * 4 + 3) - 8 * v[4 * k + 3];
          const int mk = min(__vimin3_s32(ka, kb, kc), kd);     // ties -> lowest j
          const int sc = mk >> 2;
          if (sc < st.lim) {
            st.s2 = min(st.s2, st.s1);                          // the displaced best bounds the runner-up
            st.s1 = sc;
            st.i1 = nbase + 4 * k + (mk & 3);
            st.lim = sc;
            T = (nbm - sc) >> 1;
          }
        } else {   // kTop2; lim <= s2 always, so s < lim implies the insertion precondition s < s2
          const int sv[4] = {nb4.x - 2 * v[4 * k], nb4.y - 2 * v[4 * k + 1], nb4.z - 2 * v[4 * k + 2],
                             nb4.w - 2 * v[4 * k + 3]};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            if (sv[j] < st.lim) {
              top2_insert(sv[j], nbase + 4 * k + j, st.s1, st.i1, st.s2, st.i2);
              st.lim = min(st.lim, st.s2);
            }
          }
          T = (nbm - st.lim) >> 1;
        }
      }
    }
    if (st.lim != lim_p) *st.pub = make_int2(st.s1, st.s2);
  }
}

template <int kKind>
__global__ void __launch_bounds__(kThreads, 1)
sweep_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_a2,
             const __grid_constant__ CUtensorMap tmap_b, const SweepParams p) {
  extern __shared__ uint8_t smem_raw[];
  SweepSmem& sm = *reinterpret_cast<SweepSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  if (is_gathered(kKind) && __ldg(p.fb_total) == 0) return;   // the usual case: the bounds settled every row

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int n_items = (kKind == kVerify && p.vitems) ? __ldg(p.vitem_count)
                                                     : p.n_pairs * p.blocks_per_pair * (kKind == kVerify ? 2 : 1);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap_a);
    tma_prefetch_desc(&tmap_a2);
    tma_prefetch_desc(&tmap_b);
    for (int i = 0; i < kAStages; ++i) { mbar_init(&sm.a_full[i], 1); mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kBStages; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < kNStages; ++i) { mbar_init(&sm.n_full[i], 1); mbar_init(&sm.n_empty[i], kEpiWarps); }
    for (int i = 0; i < 4; ++i) { mbar_init(&sm.t_full[i], 1); mbar_init(&sm.t_empty[i], 4); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&sm.tmem_base);
  if (warp >= kFirstEpiWarp) {
    const int t = threadIdx.x - kFirstEpiWarp * 32;          // 0..511 -> one (h, row) slot each
    sm.share[0][t >> 8][t & 255] = make_int2(kBig, kBig);
    sm.share[1][t >> 8][t & 255] = make_int2(kBig, kBig);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0, n_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        const ItemInfo it = decode_item<kKind>(p, item);
        if (!it.valid) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        mbar_wait(&sm.a_empty[as], aph ^ 1);
        mbar_arrive_expect_tx(&sm.a_full[as], it.nm * kMBytes);
        const CUtensorMap* ta = (kKind == kVerify && it.check == 1) ? &tmap_a2 : &tmap_a;
        for (int m = 0; m < it.nm; ++m) tma_load_2d(sm.a[as][m], ta, 0, it.row_img_row0 + m * kTM, &sm.a_full[as]);
        ++a_it;
        const int64_t crow0 = static_cast<int64_t>(it.col_img) * p.k_stride;
        for (int t = 0; t < it.ntiles; ++t) {
          const int32_t crow = static_cast<int32_t>(crow0 + t * kTT);
          if (kKind != kPeak) {
            const uint32_t ns = n_it % kNStages, nph = (n_it / kNStages) & 1;
            mbar_wait(&sm.n_empty[ns], nph ^ 1);
            mbar_arrive_expect_tx(&sm.n_full[ns], sizeof(NormSlot));
            bulk_load(sm.norms[ns].nb, p.sqnorm + crow, kTT * 4, &sm.n_full[ns]);
            bulk_load(sm.norms[ns].nbmin, p.nbmin32 + crow / 32, 8 * 4, &sm.n_full[ns]);
            ++n_it;
          }
          const uint32_t bs = b_it % kBStages, bph = (b_it / kBStages) & 1;
          mbar_wait(&sm.b_empty[bs], bph ^ 1);
          mbar_arrive_expect_tx(&sm.b_full[bs], kBBytes);
          tma_load_2d(sm.b[bs], &tmap_b, 0, crow, &sm.b_full[bs]);
          tma_load_2d(sm.b[bs] + kBBytes / 2, &tmap_b, 0, crow + 128, &sm.b_full[bs]);
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
        const ItemInfo it = decode_item<kKind>(p, item);
        if (!it.valid) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        mbar_wait(&sm.a_full[as], aph);
        for (int t = 0; t < it.ntiles; ++t) {
          const uint32_t bs = b_it % kBStages, bph = (b_it / kBStages) & 1;
          mbar_wait(&sm.b_full[bs], bph);
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const uint64_t bdesc = make_desc_k128(smem_u32(sm.b[bs] + h * (kBBytes / 2)));
            for (int m = 0; m < it.nm; ++m) {
              const int acc = m * 2 + h;
              mbar_wait(&sm.t_empty[acc], (t_it[acc] & 1) ^ 1);
              tc_fence_after();
              const uint64_t adesc = make_desc_k128(smem_u32(sm.a[as][m]));
              const uint32_t tacc = tmem_base + m * kTT + h * kTN;
#pragma unroll
              for (int k = 0; k < kD / 32; ++k) {
                // +32 B along K inside the 128 B swizzle atom = +2 in the (addr >> 4) field
                mma_i8_ss(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
              }
              tc_commit(&sm.t_full[acc]);   // accumulator (m, h) ready for its epilogue warps
              ++t_it[acc];
            }
          }
          tc_commit(&sm.b_empty[bs]);   // smem stage free once these MMAs retire
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
    uint32_t t_it = 0, n_it = 0, item_it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      const ItemInfo it = decode_item<kKind>(p, item);
      if (!it.valid) continue;
      const bool active = m < it.nm;
      const int64_t orow = static_cast<int64_t>(it.pair) * p.k_stride + it.blk * kRows + rl;

      RowState st;
      st.s1 = kBig; st.i1 = -1; st.s2 = kBig; st.i2 = -1; st.lim = kBig;
      const uint32_t par = item_it & 1;
      ++item_it;
      st.pub = &sm.share[par][h][rl];
      const int2* partner = &sm.share[par][h ^ 1][rl];
      st.thr = -kBig; st.excl = -1; st.tie = -kBig; st.viol = 0;   // padding rows of a kVerify block can never be violated
      if (kKind == kVerify && active && rl < it.n_rows) {
        if (it.check == 0) {
          st.thr = __ldg(p.cand_thr_ratio + orow);
          st.excl = __ldg(p.cand_excl + orow);
        } else {
          st.thr = __ldg(p.cand_thr_mut + orow);
          st.tie = __ldg(p.cand_q + orow);
        }
      }

      for (int t = 0; t < it.ntiles; ++t) {
        const uint32_t ns = n_it % kNStages, nph = (n_it / kNStages) & 1;
        ++n_it;
        if (kKind == kPeak) {
          if (active) {
            mbar_wait(&sm.t_full[acc], t_it & 1);
            tc_fence_after();
            int v[32];
            tmem_ld_32x32(tmem_base + ((quad * 32u) << 16) + m * kTT + h * kTN, v);
            tmem_ld_fence(v);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.t_empty[acc]);
            st.s1 = min(st.s1, v[0]);
            ++t_it;
          }
          continue;
        }
        mbar_wait(&sm.n_full[ns], nph);
        if (active) {
          const NormSlot& nsl = sm.norms[ns];
          if (is_top2(kKind) || is_top1(kKind)) {
            // Bound from the other column half of this row. kTop2: the union's second best is
            // <= min(max(s1, s1'), s2'); kTop1: the union's best is <= s1'. Anything at or above the
            // bound (strictly above on ties, the halves interleave in column order) cannot matter.
            // Racy reads are fine: published values only ever decrease.
            const long long raw = *reinterpret_cast<const volatile long long*>(partner);
            const int ox = static_cast<int>(raw), oy = static_cast<int>(raw >> 32);
            st.lim = min(st.lim, (is_top1(kKind) ? ox : min(max(st.s1, ox), oy)) + 1);
          }
          mbar_wait(&sm.t_full[acc], t_it & 1);
          tc_fence_after();
          const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + m * kTT + h * kTN;
          const int col0 = t * kTT + h * kTN;
          // Software-pipelined drain in half-chunks of 16 columns: the tcgen05.ld of the next half
          // is in flight while the current one is processed (two 16-register buffers).
          const uint32_t nb_s = smem_u32(nsl.nb) + (h * kTN) * 4;
          const uint32_t nbm_s = smem_u32(nsl.nbmin) + (h * 4) * 4;
          int va[16], vb[16];
          tmem_ld_32x16(taddr, va);
          tmem_ld_fence16(va);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            const int nbm = lds_s32(nbm_s + 4 * c);
            tmem_ld_32x16(taddr + c * 32 + 16, vb);
            process16<kKind>(va, nb_s + (c * 32) * 4, nbm, col0 + c * 32, st);
            tmem_ld_fence16(vb);
            if (c < 3) {
              tmem_ld_32x16(taddr + (c + 1) * 32, va);
            } else {
              // accumulator fully drained into registers: hand it back to the MMA warp
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&sm.t_empty[acc]);
            }
            process16<kKind>(vb, nb_s + (c * 32 + 16) * 4, nbm, col0 + c * 32 + 16, st);
            if (c < 3) tmem_ld_fence16(va);
          }
          ++t_it;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&sm.n_empty[ns]);
      }

      // the two column halves of a row meet here: half 1 -> smem -> half 0 -> global
      if (h == 1 && active)
        sm.rowmerge[par][rl] = kKind == kVerify ? make_int4(st.viol, 0, 0, 0) : make_int4(st.s1, st.i1, st.s2, st.i2);
      sm.share[par ^ 1][h][rl] = make_int2(kBig, kBig);     // next item's slot; nobody reads it before the barrier
      named_bar_sync(kEpiBarrier, kEpiWarps * 32);
      if (h == 0 && active && rl < it.n_rows) {
        const int4 o = sm.rowmerge[par][rl];
        if (kKind == kVerify) {
          (it.check == 0 ? p.viol_ratio : p.viol_mut)[orow] = st.viol | o.x;
        } else {
          // merge the two halves; columns of the halves interleave tile by tile, so ties are
          // resolved on the column index explicitly
          int r1s, r1i, r2s, r2i;
          if (is_top1(kKind)) {
            const bool other = lex_less(o.x, o.y, st.s1, st.i1);
            r1s = other ? o.x : st.s1;
            r1i = other ? o.y : st.i1;
            r2s = min(min(o.z, st.s2), other ? st.s1 : o.x);   // the losing best bounds the runner-up
            r2i = -1;
          } else if (lex_less(o.x, o.y, st.s1, st.i1)) {
            r1s = o.x; r1i = o.y;
            if (lex_less(o.z, o.w, st.s1, st.i1)) { r2s = o.z; r2i = o.w; } else { r2s = st.s1; r2i = st.i1; }
          } else {
            r1s = st.s1; r1i = st.i1;
            if (lex_less(o.x, o.y, st.s2, st.i2)) { r2s = o.x; r2i = o.y; } else { r2s = st.s2; r2i = st.i2; }
          }
          if (is_gathered(kKind)) {
            const int qi = __ldg(p.pairs + 2 * it.pair);
            const int qrow = __ldg(p.fb_q + orow);
            const int na = __ldg(p.sqnorm + qi * p.k_stride + qrow);
            p.rowres_o[static_cast<int64_t>(it.pair) * p.k_stride + qrow] =
                make_int4(r1s + na, r1i, r2s + na, kKind == kTop2G ? r2i : 3);
          } else {
            const int na = __ldg(p.sqnorm + it.row_img_row0 + rl);
            p.rowres[orow] = make_int4(r1s + na, r1i, r2s + na, r2i);
          }
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
