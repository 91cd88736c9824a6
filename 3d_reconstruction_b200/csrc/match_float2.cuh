// Float path, ONE tensor sweep per pair: row arg-max AND column maximum of the similarity (or of the log
// assignment score) from the same accumulators, so that the mutual check of lightglue/lightglue.py:239-242
// no longer costs a second, transposed sweep (match_float.cuh sweeps every pair in both directions).
//
//   lightglue/lightglue.py:223      sim = einsum('bmd,bnd->bmn')              one dense contraction per pair
//   lightglue/lightglue.py:237      scores.max(2)  -> per-thread running (best, column) as before
//                                   scores.max(1)  -> per column: max over the two rows a thread owns (FMNMX),
//                                                     redux.sync.max.f32 over the warp's 32 lanes (CREDUX.MAX.F32,
//                                                     one warp instruction per column), order-preserving integer
//                                                     key of the result and one 32-bit atomicMax per (warp, column)
//                                                     into colkey[pair][column]
//   lightglue/lightglue.py:239-250  mutual check + threshold in ip_finalize2_kernel: row r keeps its partner j iff
//                                   its value IS the column maximum (same accumulator value on both sides, so the
//                                   comparison is exact) and r is the lowest such row (claim array, atomicMin).
//
// Work item = (pair, block of 256 query rows): two UMMA M-tiles stay in shared memory and share every streamed
// column tile, which halves the TMA write traffic into shared memory against the 128-row items of match_float.cuh
// (the bound of that kernel: 128 B/clk of MMA operand reads + 64 B/clk of column-tile writes asked of a 128 B/clk
// port; here 128 + 32). TMEM: two pairs of 128 x 128 fp32 accumulators (tile parity x M-tile), the MMAs of one
// column tile overlap the epilogue of the previous one.
//   warp 0       TMA producer: row block per item (per K-block barriers, so the reload overlaps the last tile's
//                MMAs), column image as a stream of (128 columns x 64 elements) K-blocks through a ring
//   warp 1       MMA issuer: per column tile and K-block 2 x 4 tcgen05.mma.kind::f16 (M128 N128 K16)
//   warps 4-19   epilogue: warp (lane quadrant q, tile parity, column half) drains 64 columns of BOTH M-tiles of
//                every second column tile for rows 32q..32q+31 (+128)
// Exact ties between two DIFFERENT rows of one column (bit-equal fp32 scores) are resolved among the rows whose
// own arg-max is that column; duplicates of a row behave as in the reference (lowest row wins).
#pragma once
#include "match_float.cuh"

namespace b200m {
namespace ip {

constexpr int kRows2 = 256;                       // rows per work item (two M-tiles)
constexpr int kBStages2 = 5;
// cudaMemset(0x80). As an ordered key this is the score -3.39e38 (float bits 0xff7f7f7f), NOT below every finite key
// (key(-FLT_MAX) = 0x80800000 is smaller): a column whose largest score lay under -3.39e38 would keep the pre-set value
// and simply find no row equal to it, i.e. stay unmatched. Scores of that size do not come out of these sweeps (inner
// products of bf16 rows with finite norms, log-assignments of them), so it serves as "no row seen yet".
constexpr int kColKeyNone = static_cast<int>(0x80808080u);
constexpr int kClaimNone = 0x7f7f7f7f;                        // cudaMemset(0x7f): above every row index

template <int kDim>
struct Ip2Smem {
  static constexpr int kNKB = kDim / kKB;
  static constexpr int kAStages = kDim == 128 ? 2 : 1;        // D = 256: the 128 KB row block is single buffered
  uint8_t a[kAStages][kNKB][2][kBlockBytes];                  // [stage][K-block][M-tile]
  uint8_t b[kBStages2][kBlockBytes];
  float2 rowmerge[2][3][kRows2];                              // (value, column as int bits) of epilogue sets 1..3
  int colx[kEpiWarps][2][32];                                 // per warp: column maxima of a 32-column group (warp-uniform -> one per lane)
  uint64_t a_full[kAStages][kNKB], a_empty[kAStages][kNKB];
  uint64_t b_full[kBStages2], b_empty[kBStages2];
  uint64_t t_full[2], t_empty[2];                             // accumulator pair of a tile parity
  uint32_t tmem_base;
  int lse_cnt[16];                                            // kIpLse: arrivals of the four lane quadrants per (round parity, tile parity, column half, group)
};

struct Ip2Params {
  const int32_t* k_per_image;   // [n_img]
  const int32_t* pairs;         // [n_pairs, 2] (query image, train image)
  int64_t k_stride;
  int32_t n_pairs;
  int32_t blocks_per_image;     // k_stride / 256
  int32_t n_img;
  float2* rowbest;              // out [n_pairs][k_stride]: (best score, best column) of every query row
  int32_t* colkey;              // out [n_pairs][k_stride]: ordered key of every train column's maximum (pre-set to kColKeyNone)
  const float* nb;              // head only: [2][n_pairs][k_stride] logsigmoid(z) - log-sum-exp (written by the kIpLse sweep)
  // kIpLse (single-sweep log-sum-exp, see the epilogue) only:
  const float* z;               // [n_img, k_stride] matchability logits
  const float* maxnorm;         // [n_img] largest squared row norm of every image (ip_maxnorm_kernel)
  float* nbw;                   // out: nb, direction 0 (rows); direction 1 is written by ip_lse_finalize_kernel
  float* colpart;               // out [n_pairs][blocks_per_image][k_stride]: column sums of exp over the 256 rows of a block
  int32_t* lse_flag;            // out: set when a sum left the fp32 range (the caller re-runs the batch on the two-direction sweep)
};

// Smallest sum the single-sweep log-sum-exp accepts, PER TERM of the sum: ex2.approx.ftz flushes every exponential below
// 2^-126 to zero, so a sum of n terms can silently lose up to n * 2^-126; asking for sum >= n * 2^-102 keeps that loss
// below 2^-24 of the sum (fp32 rounding) whatever the number of columns / rows. Smaller sums raise lse_flag and the
// batch is redone with per-row maxima. (n = 16,384: 2^-88, i.e. a largest term ~61 nats below the reference point.)
constexpr float kLseTinyPerTerm = 0x1p-102f;

// Column sums over the 32 lanes of a warp for 32 columns at once: c[k] (k = 0..31) holds this lane's value of
// column k; afterwards c[0] of lane l is the sum over all lanes of column l. Butterfly: at every step a lane keeps
// one half of its columns and hands the other half to its partner -- 31 shuffles for 32 columns instead of 5 each.
__device__ __forceinline__ float warp_colsum32(float (&c)[32], uint32_t lane) {
#pragma unroll
  for (int w = 16; w >= 1; w >>= 1) {
    const bool up = (lane & w) != 0;
#pragma unroll
    for (int k = 0; k < w; ++k) {
      const float keep = up ? c[k + w] : c[k];
      const float send = up ? c[k] : c[k + w];
      c[k] = keep + __shfl_xor_sync(0xffffffffu, send, w);
    }
  }
  return c[0];
}


// redux.sync.max.f32 (sm_100a: CREDUX.MAX.F32): the maximum over the warp's 32 lanes in one instruction, no integer key
__device__ __forceinline__ float warp_max_f32(float v) {
  float r;
  asm volatile("redux.sync.max.f32 %0, %1, 0xffffffff;" : "=f"(r) : "f"(v));
  return r;
}

// order-preserving map fp32 -> int32 (compares like the floats, -0 just below +0; callers add +0.0f first, so that
// the two zeros -- equal as scores -- share a key)
__device__ __forceinline__ int ordered_key(float f) {
  const int b = __float_as_int(f);
  return b ^ ((b >> 31) & 0x7fffffff);
}

__device__ __forceinline__ float fmax8(const int* v) {
  const float a = fmaxf(fmaxf(__int_as_float(v[0]), __int_as_float(v[1])), __int_as_float(v[2]));
  const float b = fmaxf(fmaxf(__int_as_float(v[3]), __int_as_float(v[4])), __int_as_float(v[5]));
  return fmaxf(fmaxf(a, b), fmaxf(__int_as_float(v[6]), __int_as_float(v[7])));
}

// eight columns of one row starting at column c0: a new maximum records the group and a copy of its values
__device__ __forceinline__ void ip_process8(const int* v, int c0, float& best, int& bidx, int (&keep)[8]) {
  const float m = fmax8(v);
  if (m > best) {           // strict: an equal value further right never replaces the lowest column
    best = m;
    bidx = c0;
#pragma unroll
    for (int j = 0; j < 8; ++j) keep[j] = v[j];
  }
}
__device__ __forceinline__ int ip_locate8(const int (&keep)[8], float best, int bidx) {
  if (bidx < 0) return bidx;
  int j = 7;
#pragma unroll
  for (int jj = 6; jj >= 0; --jj)
    if (__int_as_float(keep[jj]) == best) j = jj;
  return bidx + j;
}

template <int kDim, int kMode>
__global__ void __launch_bounds__(kThreads, 1)
ip_sweep2_kernel(const __grid_constant__ CUtensorMap tmap, const Ip2Params p) {
  static_assert(kMode == kIpArgmax || kMode == kIpBiasArgmax || kMode == kIpLse, "arg-max, head arg-max or log-sum-exp");
  using Smem = Ip2Smem<kDim>;
  constexpr int kNKB = Smem::kNKB;
  constexpr int kAStages = Smem::kAStages;
  extern __shared__ uint8_t smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int n_items = p.n_pairs * p.blocks_per_image;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap);
    for (int s = 0; s < kAStages; ++s)
      for (int kb = 0; kb < kNKB; ++kb) { mbar_init(&sm.a_full[s][kb], 1); mbar_init(&sm.a_empty[s][kb], 1); }
    for (int i = 0; i < kBStages2; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&sm.t_full[i], 1); mbar_init(&sm.t_empty[i], 8); }
    for (int i = 0; i < 16; ++i) sm.lse_cnt[i] = 0;
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&sm.tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;

  // item -> (pair, row block); rows = query image, columns = train image
  auto decode = [&](int item, int& pair, int& blk, int& row_img, int& col_img, int& n_rows, int& n_cols) {
    pair = item / p.blocks_per_image;
    blk = item - pair * p.blocks_per_image;
    const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
    row_img = pi.qi;
    col_img = pi.ti;
    n_rows = min(pi.kq - blk * kRows2, kRows2);
    n_cols = pi.kt;
    return n_rows > 0 && n_cols > 0;
  };

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0;
      auto load_b_tile = [&](int col_img, int t) {
        const int32_t crow = static_cast<int32_t>(col_img * p.k_stride + t * kColsF);
        for (int kb = 0; kb < kNKB; ++kb) {
          const uint32_t bs = b_it % kBStages2, bph = (b_it / kBStages2) & 1;
          mbar_wait(&sm.b_empty[bs], bph ^ 1);
          mbar_arrive_expect_tx(&sm.b_full[bs], kBlockBytes);
          tma_load_2d(sm.b[bs], &tmap, kb * kKB, crow, &sm.b_full[bs]);
          ++b_it;
        }
      };
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int pair, blk, row_img, col_img, n_rows, n_cols;
        if (!decode(item, pair, blk, row_img, col_img, n_rows, n_cols)) continue;
        const int ntiles = (n_cols + kColsF - 1) / kColsF;
        // the first column tile goes first: it does not depend on the row block of the previous item being free
        load_b_tile(col_img, 0);
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        const int32_t arow = static_cast<int32_t>(row_img * p.k_stride + blk * kRows2);
        for (int kb = 0; kb < kNKB; ++kb) {
          mbar_wait(&sm.a_empty[as][kb], aph ^ 1);
          mbar_arrive_expect_tx(&sm.a_full[as][kb], 2 * kBlockBytes);
          tma_load_2d(sm.a[as][kb][0], &tmap, kb * kKB, arow, &sm.a_full[as][kb]);
          tma_load_2d(sm.a[as][kb][1], &tmap, kb * kKB, arow + kRowsF, &sm.a_full[as][kb]);
        }
        ++a_it;
        for (int t = 1; t < ntiles; ++t) load_b_tile(col_img, t);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kRowsF, kColsF);
      uint32_t a_it = 0, b_it = 0, t_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int pair, blk, row_img, col_img, n_rows, n_cols;
        if (!decode(item, pair, blk, row_img, col_img, n_rows, n_cols)) continue;
        const uint32_t as = a_it % kAStages, aph = (a_it / kAStages) & 1;
        const int ntiles = (n_cols + kColsF - 1) / kColsF;
        for (int t = 0; t < ntiles; ++t) {
          const uint32_t tp = t_it & 1, tph = (t_it >> 1) & 1;
          mbar_wait(&sm.t_empty[tp], tph ^ 1);
          tc_fence_after();
          const uint32_t tacc = tmem_base + tp * 2 * kColsF;
#pragma unroll
          for (int kb = 0; kb < kNKB; ++kb) {
            if (t == 0) mbar_wait(&sm.a_full[as][kb], aph);
            const uint32_t bs = b_it % kBStages2, bph = (b_it / kBStages2) & 1;
            mbar_wait(&sm.b_full[bs], bph);
            tc_fence_after();
            const uint64_t bdesc = make_desc_k128(smem_u32(sm.b[bs]));
#pragma unroll
            for (int m = 0; m < 2; ++m) {
              const uint64_t adesc = make_desc_k128(smem_u32(sm.a[as][kb][m]));
#pragma unroll
              for (int k = 0; k < kKB / 16; ++k)
                mma_bf16_ss(tacc + m * kColsF, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) > 0 ? 1u : 0u);
            }
            tc_commit(&sm.b_empty[bs]);
            if (t == ntiles - 1) tc_commit(&sm.a_empty[as][kb]);   // this K-block of the row block is free for the next item
            ++b_it;
          }
          tc_commit(&sm.t_full[tp]);
          ++t_it;
        }
        ++a_it;
      }
    }
  } else if (warp >= kFirstEpiWarp) {
    // ------------------------------------------------------------- epilogue
    const uint32_t quad = warp & 3;                   // TMEM lane quadrant this warp may read
    const uint32_t e = (warp - kFirstEpiWarp) >> 2;   // 0..3: (tile parity, column half)
    const uint32_t tp = e & 1, h = e >> 1;
    const int r0 = quad * 32 + lane, r1 = kRowsF + r0;   // this thread's two rows inside the item's block
    const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + tp * 2 * kColsF + h * 64;
    uint32_t t_seen = 0, my_it = 0, item_it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      int pair, blk, row_img, col_img, n_rows, n_cols;
      if (!decode(item, pair, blk, row_img, col_img, n_rows, n_cols)) continue;
      const uint32_t par = item_it & 1;
      ++item_it;
      const int ntiles = (n_cols + kColsF - 1) / kColsF;
      const bool ok0 = r0 < n_rows, ok1 = r1 < n_rows;
      const bool full_rows = n_rows == kRows2;          // warp-uniform
      const int64_t obase = static_cast<int64_t>(pair) * p.k_stride;
      if constexpr (kMode == kIpLse) {
        // ---- log-sum-exp of every row AND every column of sim (lightglue.py:196-208) from ONE sweep.
        // Reference point of the exponentials: ref = |longest query row| * |longest train row| >= every sim[i, j]
        // (Cauchy-Schwarz), the same for the whole pair, so that exp(sim - ref) <= 1 serves the row sums and the
        // column sums alike -- one ex2 per score. A row / column whose largest term sits more than ~ 61-69 nats
        // (kLseTinyPerTerm times the number of terms) below ref would lose part of its sum to the fp32 range: such
        // sums raise lse_flag and the caller falls back.
        const float ref = sqrtf(__ldg(p.maxnorm + row_img) * __ldg(p.maxnorm + col_img));
        const float ref2 = ref * kLog2e;
        float rs0 = 0.f, rs1 = 0.f;
        float csum[32];
        uint32_t rnd = my_it;                              // tiles this warp has drained so far (same for the eight warps of a pair)
        float* cpart = p.colpart + (static_cast<int64_t>(pair) * p.blocks_per_image + blk) * p.k_stride;
        // The four lane quadrants of a column group meet in shared memory: every warp parks its 32 column sums, the
        // last of the four to arrive adds them in a fixed order (deterministic) and writes ONE partial per (row block,
        // column). Slots are double buffered by the parity of the warp's tile count: a warp can be at most one of its
        // tiles ahead of the slowest warp of its accumulator pair (t_empty needs all eight arrivals). In this mode the
        // row sums use the first half of `rowmerge` as plain floats; its second half and `colx` hold the slots.
        float* rsum = reinterpret_cast<float*>(sm.rowmerge);                     // [par][3][256]
        volatile float* cq0 = reinterpret_cast<volatile float*>(sm.colx);      // [tp][h][grp][quad][32], even rounds
        volatile float* cq1 = rsum + 2 * 3 * kRows2;                           // the same, odd rounds
        for (int t = static_cast<int>((tp + 2 - (t_seen & 1)) & 1); t < ntiles; t += 2) {
          mbar_wait(&sm.t_full[tp], my_it & 1);
          tc_fence_after();
          rnd = my_it;
          ++my_it;
#pragma unroll
          for (int s = 0; s < 4; ++s) {
            const int c0 = t * kColsF + h * 64 + s * 16;
            int va[16], vb[16];
            tmem_ld_32x16(taddr + s * 16, va);
            tmem_ld_32x16(taddr + kColsF + s * 16, vb);
            tmem_ld_fence16(va);
            tmem_ld_fence16(vb);
            if (s == 3) {
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive(&sm.t_empty[tp]);
            }
            if (c0 + 16 > n_cols) {                      // ragged end: exp -> 0
#pragma unroll
              for (int j = 0; j < 16; ++j)
                if (c0 + j >= n_cols) { va[j] = __float_as_int(kLseFloor); vb[j] = __float_as_int(kLseFloor); }
            }
            if (!full_rows) {                            // rows past the image add nothing to a column
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                if (!ok0) va[j] = __float_as_int(kLseFloor);
                if (!ok1) vb[j] = __float_as_int(kLseFloor);
              }
            }
            float a0 = 0.f, a1 = 0.f, b0 = 0.f, b1 = 0.f;
#pragma unroll
            for (int j = 0; j < 16; j += 2) {
              const float ea0 = ex2_ftz(fmaf(__int_as_float(va[j]), kLog2e, -ref2));
              const float ea1 = ex2_ftz(fmaf(__int_as_float(va[j + 1]), kLog2e, -ref2));
              const float eb0 = ex2_ftz(fmaf(__int_as_float(vb[j]), kLog2e, -ref2));
              const float eb1 = ex2_ftz(fmaf(__int_as_float(vb[j + 1]), kLog2e, -ref2));
              a0 += ea0; a1 += ea1; b0 += eb0; b1 += eb1;
              csum[(s & 1) * 16 + j] = ea0 + eb0;
              csum[(s & 1) * 16 + j + 1] = ea1 + eb1;
            }
            rs0 += a0 + a1;
            rs1 += b0 + b1;
            if (s & 1) {
              const float mine = warp_colsum32(csum, lane);
              const int col = t * kColsF + h * 64 + (s >> 1) * 32 + static_cast<int>(lane);
              const uint32_t rp = rnd & 1, g4 = ((tp * 2 + h) * 2 + (s >> 1));
              volatile float* slot = (rp ? cq1 : cq0) + g4 * 128;
              slot[quad * 32 + lane] = mine;
              __syncwarp();
              int old = 0;
              if (lane == 0) {
                __threadfence_block();
                old = atomicAdd(&sm.lse_cnt[rp * 8 + g4], 1);
              }
              old = __shfl_sync(0xffffffffu, old, 0);
              if (old == 3) {                            // the other three quadrants are in
                __threadfence_block();
                const float tot = ((slot[lane] + slot[32 + lane]) + slot[64 + lane]) + slot[96 + lane];
                if (col < n_cols) cpart[col] = tot;      // every (pair, block, column) is written exactly once
                if (lane == 0) atomicExch(&sm.lse_cnt[rp * 8 + g4], 0);
              }
            }
          }
        }
        t_seen += ntiles;
        if (e > 0) {
          rsum[(par * 3 + (e - 1)) * kRows2 + r0] = rs0;
          rsum[(par * 3 + (e - 1)) * kRows2 + r1] = rs1;
        }
        named_bar_sync(kEpiBarrier, kEpiWarps * 32);
        if (e == 0) {
#pragma unroll
          for (int m = 0; m < 2; ++m) {
            const int r = m ? r1 : r0;
            if (r >= n_rows) continue;
            float ss = m ? rs1 : rs0;
#pragma unroll
            for (int oo = 0; oo < 3; ++oo) ss += rsum[(par * 3 + oo) * kRows2 + r];
            if (!(ss >= kLseTinyPerTerm * static_cast<float>(n_cols)) || ss > 3.0e38f) atomicOr(p.lse_flag, 1);
            const int64_t row = static_cast<int64_t>(blk) * kRows2 + r;
            p.nbw[obase + row] = log_sigmoid(__ldg(p.z + static_cast<int64_t>(row_img) * p.k_stride + row)) - (ref + logf(ss));
          }
        }
        continue;
      }
      float best0 = kNegInf, best1 = kNegInf;
      int bidx0 = -1, bidx1 = -1;
      int keep0[8], keep1[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) { keep0[j] = 0; keep1[j] = 0; }
      float own0 = 0.f, own1 = 0.f;                      // head: the rows' own term nb0[i]
      const float* bias = nullptr;                       // head: nb1[j], the per-column term
      if constexpr (kMode == kIpBiasArgmax) {
        bias = p.nb + (static_cast<int64_t>(p.n_pairs) + pair) * p.k_stride;
        if (ok0) own0 = __ldg(p.nb + obase + blk * kRows2 + r0);
        if (ok1) own1 = __ldg(p.nb + obase + blk * kRows2 + r1);
      }
      int32_t* colkey = p.colkey + obase;
      int (*cx)[32] = sm.colx[warp - kFirstEpiWarp];

      for (int t = static_cast<int>((tp + 2 - (t_seen & 1)) & 1); t < ntiles; t += 2) {
        mbar_wait(&sm.t_full[tp], my_it & 1);
        tc_fence_after();
        ++my_it;
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          const int c0 = t * kColsF + h * 64 + s * 16;
          int va[16], vb[16];                            // the two rows' sixteen columns
          tmem_ld_32x16(taddr + s * 16, va);
          tmem_ld_32x16(taddr + kColsF + s * 16, vb);
          tmem_ld_fence16(va);
          tmem_ld_fence16(vb);
          if (s == 3) {                                  // both accumulators are in registers: hand the pair back
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.t_empty[tp]);
          }
          if constexpr (kMode == kIpBiasArgmax) {        // S = 2 sim + nb1[j] + nb0[i], the full log assignment score
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(bias + c0 + j));
              const float bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                va[j + u] = __float_as_int(fmaf(2.f, __int_as_float(va[j + u]), bb[u]) + own0);
                vb[j + u] = __float_as_int(fmaf(2.f, __int_as_float(vb[j + u]), bb[u]) + own1);
              }
            }
          }
          if (c0 + 16 > n_cols) {                        // ragged end (and padding columns: zero vectors must not compete)
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (c0 + j >= n_cols) { va[j] = __float_as_int(kNegInf); vb[j] = __float_as_int(kNegInf); }
          }
          if (!full_rows) {                              // rows past the image must not win a column
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              if (!ok0) va[j] = __float_as_int(kNegInf);
              if (!ok1) vb[j] = __float_as_int(kNegInf);
            }
          }
          // rows: running (best, group) per row
          ip_process8(va, c0, best0, bidx0, keep0);
          ip_process8(va + 8, c0 + 8, best0, bidx0, keep0);
          ip_process8(vb, c0, best1, bidx1, keep1);
          ip_process8(vb + 8, c0 + 8, best1, bidx1, keep1);
          // columns: max over this thread's two rows, then over the warp's 32 lanes (REDUX: the result is warp-uniform).
          // Lane 0 parks the 32 maxima of a column group in shared memory (a store on the idle LSU pipe instead of a
          // compare + select per column on the busy ALU pipe), then every lane picks up one column for the atomic.
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const float m = warp_max_f32(fmaxf(__int_as_float(va[j]), __int_as_float(vb[j])));
            if (lane == 0) cx[s >> 1][(s & 1) * 16 + j] = __float_as_int(m);
          }
          if (s & 1) {
            __syncwarp();
            const int col = t * kColsF + h * 64 + (s >> 1) * 32 + static_cast<int>(lane);
            if (col < n_cols) atomicMax(colkey + col, ordered_key(__int_as_float(cx[s >> 1][lane]) + 0.0f));
          }
        }
      }
      t_seen += ntiles;
      bidx0 = ip_locate8(keep0, best0, bidx0);
      bidx1 = ip_locate8(keep1, best1, bidx1);
      // the four partial results of a row meet here
      if (e > 0) {
        sm.rowmerge[par][e - 1][r0] = make_float2(best0, __int_as_float(bidx0));
        sm.rowmerge[par][e - 1][r1] = make_float2(best1, __int_as_float(bidx1));
      }
      named_bar_sync(kEpiBarrier, kEpiWarps * 32);
      if (e == 0) {
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          const int r = m ? r1 : r0;
          if (r >= n_rows) continue;
          float best = m ? best1 : best0;
          int bidx = m ? bidx1 : bidx0;
#pragma unroll
          for (int oo = 0; oo < 3; ++oo) {
            const float2 q = sm.rowmerge[par][oo][r];
            const int qi = __float_as_int(q.y);
            if (qi >= 0 && (q.x > best || (q.x == best && qi < bidx) || bidx < 0)) { best = q.x; bidx = qi; }
          }
          p.rowbest[obase + blk * kRows2 + r] = make_float2(best, __int_as_float(bidx));
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

// Mutual check + threshold + ascending-query packing for the single-sweep kernel, one CTA per pair.
// claim[pair][j] (pre-set to kClaimNone): the lowest row whose arg-max is column j AND whose score is j's maximum.
template <int kBlock>
__global__ void __launch_bounds__(kBlock)
ip_finalize2_kernel(const float2* __restrict__ rowbest, const int32_t* __restrict__ colkey, int32_t* __restrict__ claim,
                    const int32_t* __restrict__ pairs, const int32_t* __restrict__ k_per_image, int n_img, int64_t k_stride,
                    float threshold, int use_threshold, int mutual, int exp_scores, int32_t* __restrict__ out_q,
                    int32_t* __restrict__ out_t, float* __restrict__ out_s, int32_t* __restrict__ out_count, int64_t cap) {
  __shared__ int warp_sums[kBlock / 32];
  __shared__ int running;
  const int pair = blockIdx.x;
  const PairInfo pi = load_pair(pairs, k_per_image, pair, n_img, k_stride);
  const int kq = pi.kq, kt = pi.kt;
  if (threadIdx.x == 0) running = 0;
  const int64_t base = static_cast<int64_t>(pair) * k_stride;
  const float2* fwd = rowbest + base;
  if (mutual && kt > 0) {
    for (int r = threadIdx.x; r < kq; r += kBlock) {
      const float2 v = fwd[r];
      const int j = __float_as_int(v.y);
      if (j >= 0 && j < kt && ordered_key(v.x + 0.0f) == colkey[base + j]) atomicMin(claim + base + j, r);
    }
    __threadfence();
  }
  __syncthreads();
  const int64_t obase = static_cast<int64_t>(pair) * cap;
  for (int r0 = 0; r0 < kq && kt > 0; r0 += kBlock) {
    const int r = r0 + threadIdx.x;
    bool keep = false;
    float s = 0.f;
    int j = -1;
    if (r < kq) {
      const float2 v = fwd[r];
      s = v.x;
      j = __float_as_int(v.y);
      keep = j >= 0 && j < kt;
      if (exp_scores) s = expf(s);      // log assignment -> matching score (lightglue.py:243)
      if (keep && use_threshold) keep = s > threshold;
      if (keep && mutual) keep = claim[base + j] == r;
    }
    const int slot = block_compact_slot<kBlock>(keep, warp_sums, &running);
    if (keep && slot < cap) {
      out_q[obase + slot] = r;
      out_t[obase + slot] = j;
      out_s[obase + slot] = s;
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) out_count[pair] = running;
}

// Largest squared row norm per image (bf16 rows, fp32 sum): the reference point of the single-sweep log-sum-exp.
// maxnorm must be zeroed first; non-negative floats order like their bit patterns.
template <int kDim>
__global__ void __launch_bounds__(256)
ip_maxnorm_kernel(const uint16_t* __restrict__ desc, const int32_t* __restrict__ k_per_image, int64_t k_stride,
                  int32_t n_img, float* __restrict__ maxnorm) {
  const int64_t row = static_cast<int64_t>(blockIdx.x) * 8 + (threadIdx.x >> 5);       // one warp per row
  const int img = static_cast<int>(row / k_stride);
  if (img >= n_img) return;
  const int r = static_cast<int>(row - static_cast<int64_t>(img) * k_stride);
  if (r >= min(max(__ldg(k_per_image + img), 0), static_cast<int>(k_stride))) return;
  const uint32_t lane = threadIdx.x & 31;
  const uint2* src = reinterpret_cast<const uint2*>(desc + row * kDim);                   // four bf16 per lane and step
  float acc = 0.f;
#pragma unroll
  for (int i = 0; i < kDim / 128; ++i) {
    const uint2 v = __ldg(src + i * 32 + lane);
    const float x0 = __uint_as_float(v.x << 16), x1 = __uint_as_float(v.x & 0xffff0000u);
    const float x2 = __uint_as_float(v.y << 16), x3 = __uint_as_float(v.y & 0xffff0000u);
    acc += x0 * x0 + x1 * x1 + x2 * x2 + x3 * x3;
  }
#pragma unroll
  for (int d = 16; d >= 1; d >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, d);
  if (lane == 0) atomicMax(reinterpret_cast<int*>(maxnorm + img), __float_as_int(acc));
}

// Columns of the single-sweep log-sum-exp: the partial sums of the row blocks in a fixed order
// -> nb[1][pair][column] = logsigmoid(z_column) - log-sum-exp over the rows.
__global__ void __launch_bounds__(256)
ip_lse_finalize_kernel(const float* __restrict__ colpart, const float* __restrict__ maxnorm, const float* __restrict__ z,
                       const int32_t* __restrict__ pairs, const int32_t* __restrict__ k_per_image, int n_img, int64_t k_stride,
                       int n_pairs, float* __restrict__ nb, int32_t* __restrict__ lse_flag) {
  const int pair = blockIdx.y;
  const PairInfo pi = load_pair(pairs, k_per_image, pair, n_img, k_stride);
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= pi.kt || pi.kq <= 0) return;
  const int parts_all = static_cast<int>(k_stride / kRows2);
  const int parts = (pi.kq + kRows2 - 1) / kRows2;
  const float* src = colpart + static_cast<int64_t>(pair) * parts_all * k_stride + col;
  float ss = 0.f;
  for (int q = 0; q < parts; ++q) ss += src[static_cast<int64_t>(q) * k_stride];
  if (!(ss >= kLseTinyPerTerm * static_cast<float>(pi.kq)) || ss > 3.0e38f) atomicOr(lse_flag, 1);
  const float ref = sqrtf(__ldg(maxnorm + pi.qi) * __ldg(maxnorm + pi.ti));
  nb[(static_cast<int64_t>(n_pairs) + pair) * k_stride + col] =
      log_sigmoid(__ldg(z + static_cast<int64_t>(pi.ti) * k_stride + col)) - (ref + logf(ss));
}

}  // namespace ip
}  // namespace b200m
