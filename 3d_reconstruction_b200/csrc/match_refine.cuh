// Refinement of the chunk sweep (pipeline 2): turns the per-row chunk bounds (L1, g1, L2) of
// chunk_sweep_kernel into exact answers, row by row, in ORIGINAL query order.
//
// For a query row with squared norm na the sweep guarantees, in score units (d^2 - na):
//   * every train column of chunk g1 scores >= L1, every column outside g1 scores >= L2,
//   * chunk g1 holds a column scoring <= L1 + spread, another chunk one scoring <= L2 + spread
//     (spread = largest norm spread of a chunk of the train image; tiny after the norm sort).
// Step 1 (all rows, no descriptor traffic): the Lowe test needs a second neighbour outside the
//   radius D*(d1); d1^2 >= L1 + na and two distinct columns lie within L2 + spread + na, so
//   L2 + spread + na < D*(L1 + na) settles "rejected" for the bulk of the rows.
// Step 2 (the remaining rows, one warp per row): the 32 columns of chunk g1 are evaluated exactly
//   (dp4a): best (score, original index), runner-up inside the chunk. If best < L2 the row's nearest
//   neighbour is known exactly, and the Lowe test is settled by the chunk's runner-up and L2 in all
//   but a sliver of cases; the sliver goes to the ratio verify sweep, rows with best >= L2 to the
//   exact fallback sweep (sweep_kernel<kTop1G>).
// Output per (pair, original query row): int4 (d1^2, i1 (original train row), d2^2 bound, state).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "match_finalize.cuh"

namespace b200m {

enum RowStateCode {
  kRowReject = 0,        // cannot be a match
  kRowPass = 1,          // nearest neighbour exact, Lowe test settled: passed
  kRowNeedRatio = 2,     // nearest neighbour exact, Lowe test needs the verify sweep
  kRowFallback = 3       // written by the fallback sweep: (d1^2, i1, upper bound of d2^2)
};

struct RefineParams {
  const uint8_t* bank;          // original order (query rows are read from here)
  const uint8_t* sbank;         // norm-sorted copy (train chunks are read from here)
  const int32_t* sqnorm;        // original order
  const int32_t* perm;          // sorted position -> original row
  const int32_t* inv;           // original row -> sorted position
  const int32_t* snorm;         // sorted norms
  const int32_t* spread;        // [n_img] largest chunk spread
  const int32_t* k_per_image;
  const int32_t* pairs;
  int64_t k_stride;
  int32_t n_img;                // pair-list entries outside [0, n_img) are skipped (load_pair)
  const int4* rowres_s;         // chunk sweep output (sorted query rows)
  int4* rowres_o;               // out (original query rows)
  int32_t* fb_count;            // [n_pairs] rows sent to the fallback sweep (zeroed by the caller)
  int32_t* fb_total;            // [1] != 0 when any row of the batch fell back (zeroed by the caller)
  int32_t* fb_q;                // [n_pairs, k_stride] their original query rows
  uint8_t* fb_desc;             // [n_pairs, k_stride, 128] their descriptors (rows of the fallback sweep)
  double ratio;
  double inv_ratio_sq;          // 1 / ratio^2 (0 when the ratio test is off)
  int32_t chunk_cols;           // 32 or 64: columns per chunk of the sweep that produced rowres_s
  unsigned long long* stats;    // counters of the call (match_common.cuh), may be NULL
  // split form (refine_classify_kernel + refine_eval_kernel) only:
  int32_t* need_count;          // [n_pairs] rows whose winning chunk has to be evaluated (zeroed by the caller)
  int32_t* need_list;           // [n_pairs, k_stride] their original query rows (any order)
};

constexpr int kRefineThreads = 256;
constexpr int kBigScore = 1 << 28;
constexpr int kPadScore = 1 << 23;   // bound of an all-padding chunk (B200M_PAD_NORM - 0): no real column behind it

// Step 1 for one row: does its winning chunk have to be evaluated? v = (L1, g1, L2), na = the row's squared norm.
__device__ __forceinline__ bool refine_needs_eval(const RefineParams& p, const int4 v, int na, int kt, int ti, double inv_ratio) {
  if (!(p.ratio > 0.0)) return true;
  if (kt < 2) return false;                             // no second neighbour: dropped by the ratio test
  if (v.z < kPadScore) {                                // a second chunk with real columns exists
    const int spread = p.spread[ti];
    const int d1lb = max(v.x + na, 0);
    const long long ub2 = static_cast<long long>(v.z) + spread + na;
    // D*(d1) = d1^2 / ratio^2 up to float rounding (relative 3e-7): a cheap test with a wide margin
    // settles almost every row, the exact radius is only computed for the rest
    if (static_cast<double>(ub2) < static_cast<double>(d1lb) * p.inv_ratio_sq * 0.999 - 2.0) return false;
    if (ub2 < ratio_radius_sq(dist_from_sq(d1lb), p.ratio, inv_ratio)) return false;   // a second column lies inside the radius for sure
  }
  return true;
}

__global__ void __launch_bounds__(kRefineThreads)
refine_kernel(const RefineParams p) {
  const int pair = blockIdx.y;
  const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
  const int qi = pi.qi, ti = pi.ti, kq = pi.kq, kt = pi.kt;
  const int lane = threadIdx.x & 31;
  const int r_warp0 = (blockIdx.x * kRefineThreads + threadIdx.x) & ~31;
  if (r_warp0 >= kq) return;
  const int r = r_warp0 + lane;
  const int64_t base = static_cast<int64_t>(pair) * p.k_stride;
  const int64_t qbase = static_cast<int64_t>(qi) * p.k_stride;
  const int64_t tbase = static_cast<int64_t>(ti) * p.k_stride;
  const bool use_ratio = p.ratio > 0.0;
  const double inv_ratio = use_ratio ? 1.0 / p.ratio : 0.0;

  int4 v = make_int4(kBigScore, 0, kBigScore, 0);
  int na = 0;
  bool need = false;
  if (r < kq && kt > 0) {
    na = p.sqnorm[qbase + r];
    v = p.rowres_s[base + p.inv[qbase + r]];
    need = refine_needs_eval(p, v, na, kt, ti, inv_ratio);
  }
  int4 out = make_int4(0, -1, 0, kRowReject);
  bool fallback = false;

  unsigned todo = __ballot_sync(0xffffffffu, need);
  if (lane == 0) {
    stat_add(p.stats, 0, static_cast<unsigned>(min(kq - r_warp0, 32)));
    stat_add(p.stats, 1, __popc(todo));
  }
  while (todo) {
    const int src = __ffs(todo) - 1;
    todo &= todo - 1;
    const int rr = r_warp0 + src;
    const int g1 = __shfl_sync(0xffffffffu, v.y, src);
    const uint4* rowp = reinterpret_cast<const uint4*>(p.bank + (qbase + rr) * 128);
    int bs = 0x7fffffff, bo = 0x7fffffff, s2 = 0x7fffffff;   // best (score, original index), runner-up score
    for (int sub = 0; sub < p.chunk_cols; sub += 32) {
      const int c = g1 * p.chunk_cols + sub + lane;          // sorted train column of this lane
      int s = 0x7fffffff, orig = 0x7fffffff;
      if (c < kt) {
        const uint4* colp = reinterpret_cast<const uint4*>(p.sbank + (tbase + c) * 128);
        uint32_t dot = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const uint4 a = __ldg(rowp + i), b = __ldg(colp + i);
          dot = __dp4a(a.x, b.x, dot);
          dot = __dp4a(a.y, b.y, dot);
          dot = __dp4a(a.z, b.z, dot);
          dot = __dp4a(a.w, b.w, dot);
        }
        s = p.snorm[tbase + c] - 2 * static_cast<int>(dot);
        orig = p.perm[tbase + c];
      }
      // best of these 32 columns by (score, original index), then the runner-up among the other lanes: three
      // redux.sync.min on offset keys (scores lie in (-2^25, 2^23)) instead of three 5-step shuffle butterflies
      const unsigned ks = c < kt ? static_cast<unsigned>(s + (1 << 25)) : 0xffffffffu;
      const unsigned km = __reduce_min_sync(0xffffffffu, ks);
      const unsigned ko = __reduce_min_sync(0xffffffffu, ks == km ? static_cast<unsigned>(orig) : 0xffffffffu);
      const unsigned k2 = __reduce_min_sync(0xffffffffu, (ks == km && static_cast<unsigned>(orig) == ko) ? 0xffffffffu : ks);
      const int ws = km == 0xffffffffu ? 0x7fffffff : static_cast<int>(km) - (1 << 25);
      const int wo = km == 0xffffffffu ? 0x7fffffff : static_cast<int>(ko);
      const int w2 = k2 == 0xffffffffu ? 0x7fffffff : static_cast<int>(k2) - (1 << 25);
      // merge into the chunk's running (best, runner-up)
      if (ws < bs || (ws == bs && wo < bo)) {
        s2 = min(min(s2, w2), bs);
        bs = ws;
        bo = wo;
      } else {
        s2 = min(min(s2, w2), ws);
      }
    }
    if (lane == src) {
      // v = (L1, g1, L2): columns outside chunk g1 score >= L2
      if (!(bs < v.z)) {
        fallback = true;                                 // another chunk may hold the nearest (or an equal) column
      } else {
        const int d1 = bs + na;
        int state = kRowPass;
        if (use_ratio) {
          const int thr = ratio_radius_sq(dist_from_sq(d1), p.ratio, inv_ratio) - na;   // a second column scoring < thr spoils the test
          if (s2 < thr) state = kRowReject;
          else if (v.z >= thr) state = kRowPass;
          else if (v.z < kPadScore && static_cast<long long>(v.z) + p.spread[ti] < thr) state = kRowReject;
          else state = kRowNeedRatio;
        }
        out = make_int4(d1, bo, 0, state);
      }
    }
  }
  const unsigned fbm = __ballot_sync(0xffffffffu, fallback);
  if (lane == 0) stat_add(p.stats, 2, __popc(fbm));
  if (r < kq) {
    if (fallback) {
      const int slot = atomicAdd(p.fb_count + pair, 1);
      if (slot == 0) atomicOr(p.fb_total, 1);
      p.fb_q[base + slot] = r;
      const uint4* src = reinterpret_cast<const uint4*>(p.bank + (qbase + r) * 128);
      uint4* dst = reinterpret_cast<uint4*>(p.fb_desc + (base + slot) * 128);
#pragma unroll
      for (int i = 0; i < 8; ++i) dst[i] = src[i];
      // rowres_o of this row is written by the fallback sweep
    } else {
      p.rowres_o[base + r] = out;
    }
  }
}

// ---- The same refinement as two kernels (option "refine" = 1). refine_kernel walks the rows that need their chunk evaluated one after
// the other inside each warp (a serial chain of memory round trips per row: fine when 2 % of the rows need it, 0.6 ms
// per 16 pairs when a third do, as on strongly overlapping real images). Split: a streaming pass settles step 1 for
// every row and lists the rest; a persistent pass gives every listed row its own warp, so rows are evaluated side by
// side. Same decisions, same outputs (the order of the fallback list differs, which nothing depends on).
__global__ void __launch_bounds__(kRefineThreads)
refine_classify_kernel(const RefineParams p) {
  const int pair = blockIdx.y;
  const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
  const int kq = pi.kq, kt = pi.kt;
  const int lane = threadIdx.x & 31;
  const int r_warp0 = (blockIdx.x * kRefineThreads + threadIdx.x) & ~31;
  if (r_warp0 >= kq) return;
  const int r = r_warp0 + lane;
  const int64_t base = static_cast<int64_t>(pair) * p.k_stride;
  const int64_t qbase = static_cast<int64_t>(pi.qi) * p.k_stride;
  const double inv_ratio = p.ratio > 0.0 ? 1.0 / p.ratio : 0.0;
  bool need = false;
  if (r < kq && kt > 0) {
    const int na = p.sqnorm[qbase + r];
    const int4 v = p.rowres_s[base + p.inv[qbase + r]];
    need = refine_needs_eval(p, v, na, kt, pi.ti, inv_ratio);
  }
  const unsigned todo = __ballot_sync(0xffffffffu, need);
  int slot0 = 0;
  if (lane == 0) {
    stat_add(p.stats, 0, static_cast<unsigned>(min(kq - r_warp0, 32)));
    stat_add(p.stats, 1, __popc(todo));
    if (todo) slot0 = atomicAdd(p.need_count + pair, __popc(todo));
  }
  slot0 = __shfl_sync(0xffffffffu, slot0, 0);
  if (r < kq) {
    if (need) p.need_list[base + slot0 + __popc(todo & ((1u << lane) - 1u))] = r;
    else p.rowres_o[base + r] = make_int4(0, -1, 0, kRowReject);
  }
}

// Persistent grid; warp w of the grid takes the listed rows w, w + W, w + 2 W, ... of every pair of the batch.
__global__ void __launch_bounds__(kRefineThreads)
refine_eval_kernel(const RefineParams p, int n_pairs) {
  const int lane = threadIdx.x & 31;
  const int warps = gridDim.x * (kRefineThreads / 32);
  const int gw = blockIdx.x * (kRefineThreads / 32) + (threadIdx.x >> 5);
  const bool use_ratio = p.ratio > 0.0;
  const double inv_ratio = use_ratio ? 1.0 / p.ratio : 0.0;
  unsigned n_fb = 0;
  int carry = gw;                                        // position of this warp's next row in the concatenated lists
  for (int pair = 0; pair < n_pairs; ++pair) {
    const int n_need = __ldg(p.need_count + pair);
    if (carry >= n_need) { carry -= n_need; continue; }
    const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
    const int kt = pi.kt;
    const int64_t base = static_cast<int64_t>(pair) * p.k_stride;
    const int64_t qbase = static_cast<int64_t>(pi.qi) * p.k_stride;
    const int64_t tbase = static_cast<int64_t>(pi.ti) * p.k_stride;
    int idx = carry;
    for (; idx < n_need; idx += warps) {
      const int r = __ldg(p.need_list + base + idx);
      const int na = p.sqnorm[qbase + r];
      const int4 v = p.rowres_s[base + p.inv[qbase + r]];
      const uint4* rowp = reinterpret_cast<const uint4*>(p.bank + (qbase + r) * 128);
      int bs = 0x7fffffff, bo = 0x7fffffff, s2 = 0x7fffffff;   // best (score, original index), runner-up score
      for (int sub = 0; sub < p.chunk_cols; sub += 32) {
        const int c = v.y * p.chunk_cols + sub + lane;         // sorted train column of this lane
        int s = 0x7fffffff, orig = 0x7fffffff;
        if (c < kt) {
          const uint4* colp = reinterpret_cast<const uint4*>(p.sbank + (tbase + c) * 128);
          uint4 a[8], b[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) b[i] = __ldg(colp + i);
#pragma unroll
          for (int i = 0; i < 8; ++i) a[i] = __ldg(rowp + i);
          uint32_t d0 = 0, d1 = 0;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            d0 = __dp4a(a[i].x, b[i].x, d0);
            d1 = __dp4a(a[i].y, b[i].y, d1);
            d0 = __dp4a(a[i].z, b[i].z, d0);
            d1 = __dp4a(a[i].w, b[i].w, d1);
          }
          s = p.snorm[tbase + c] - 2 * static_cast<int>(d0 + d1);
          orig = p.perm[tbase + c];
        }
        const unsigned ks = c < kt ? static_cast<unsigned>(s + (1 << 25)) : 0xffffffffu;
        const unsigned km = __reduce_min_sync(0xffffffffu, ks);
        const unsigned ko = __reduce_min_sync(0xffffffffu, ks == km ? static_cast<unsigned>(orig) : 0xffffffffu);
        const unsigned k2 = __reduce_min_sync(0xffffffffu, (ks == km && static_cast<unsigned>(orig) == ko) ? 0xffffffffu : ks);
        const int ws = km == 0xffffffffu ? 0x7fffffff : static_cast<int>(km) - (1 << 25);
        const int wo = km == 0xffffffffu ? 0x7fffffff : static_cast<int>(ko);
        const int w2 = k2 == 0xffffffffu ? 0x7fffffff : static_cast<int>(k2) - (1 << 25);
        if (ws < bs || (ws == bs && wo < bo)) {
          s2 = min(min(s2, w2), bs);
          bs = ws;
          bo = wo;
        } else {
          s2 = min(min(s2, w2), ws);
        }
      }
      if (!(bs < v.z)) {                                       // another chunk may hold the nearest (or an equal) column
        int slot = 0;
        if (lane == 0) {
          slot = atomicAdd(p.fb_count + pair, 1);
          if (slot == 0) atomicOr(p.fb_total, 1);
          p.fb_q[base + slot] = r;
          ++n_fb;
        }
        slot = __shfl_sync(0xffffffffu, slot, 0);
        if (lane < 8) reinterpret_cast<uint4*>(p.fb_desc + (base + slot) * 128)[lane] = __ldg(rowp + lane);
      } else if (lane == 0) {
        const int d1 = bs + na;
        int state = kRowPass;
        if (use_ratio) {
          const int thr = ratio_radius_sq(dist_from_sq(d1), p.ratio, inv_ratio) - na;   // a second column scoring < thr spoils the test
          if (s2 < thr) state = kRowReject;
          else if (v.z >= thr) state = kRowPass;
          else if (v.z < kPadScore && static_cast<long long>(v.z) + p.spread[pi.ti] < thr) state = kRowReject;
          else state = kRowNeedRatio;
        }
        p.rowres_o[base + r] = make_int4(d1, bo, 0, state);
      }
    }
    carry = idx - n_need;
  }
  if (lane == 0) stat_add(p.stats, 2, n_fb);
}

// Candidate selection of pipeline 2, ascending original query order (one CTA per pair).
//   list B (all surviving rows): cand_q / cand_t / cand_d / cand_thr_mut (+ gathered train rows for the
//           mutual verify sweep) and cand_rslot = the row's slot in list A or -1;
//   list A (rows whose Lowe test is still open): thr_ratio / excl (+ gathered query rows).
// Every warp owns a contiguous run of rows, so ballots inside the warp plus ONE block-wide scan of the
// warp totals give every kept row its ordered slot (one barrier per pair instead of six per 1024 rows). The non-empty
// 256-row blocks of both lists are appended to the verify sweep's compact work list.
struct RowClass {
  bool keep, open;
  float d1f;
};

__device__ __forceinline__ RowClass classify_row(const int4 v, int kt, double ratio) {
  RowClass c;
  c.keep = false; c.open = false; c.d1f = 0.f;
  const bool has1 = v.y >= 0 && v.y < kt;
  if (has1 && v.w != kRowReject) {
    c.d1f = dist_from_sq(v.x);
    if (v.w == kRowFallback) {
      // (d1^2, i1, upper bound of d2^2): rows passing against the bound are a superset of the survivors
      if (ratio > 0.0) {
        c.keep = kt >= 2 && static_cast<double>(c.d1f) < ratio * static_cast<double>(dist_from_sq(v.z));
        c.open = c.keep;
      } else {
        c.keep = true;
      }
    } else {
      c.keep = true;
      c.open = v.w == kRowNeedRatio;
    }
  }
  return c;
}

// exclusive block scan of (x, y); returns this thread's offsets, totals through tot_x / tot_y
template <int kBlock>
__device__ __forceinline__ void block_scan2(int x, int y, int& off_x, int& off_y, int& tot_x, int& tot_y, int2* warp_tot) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int ix = x, iy = y;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int ox = __shfl_up_sync(0xffffffffu, ix, d), oy = __shfl_up_sync(0xffffffffu, iy, d);
    if (lane >= d) { ix += ox; iy += oy; }
  }
  if (lane == 31) warp_tot[warp] = make_int2(ix, iy);
  __syncthreads();
  int bx = 0, by = 0, sx = 0, sy = 0;
#pragma unroll
  for (int w = 0; w < kBlock / 32; ++w) {
    const int2 t = warp_tot[w];
    if (w < warp) { bx += t.x; by += t.y; }
    sx += t.x; sy += t.y;
  }
  off_x = bx + ix - x;
  off_y = by + iy - y;
  tot_x = sx;
  tot_y = sy;
}

template <int kBlock>
__global__ void __launch_bounds__(kBlock)
select2_kernel(const int4* __restrict__ rowres_o, const int32_t* __restrict__ pairs,
               const int32_t* __restrict__ k_per_image, int n_img, int64_t k_stride, double ratio,
               const uint8_t* __restrict__ bank, const int32_t* __restrict__ sqnorm, int do_ratio, int gather_t,
               int32_t* __restrict__ cand_q, int32_t* __restrict__ cand_t, float* __restrict__ cand_d,
               int32_t* __restrict__ cand_thr_mut, int32_t* __restrict__ cand_rslot, int32_t* __restrict__ cand_count,
               int32_t* __restrict__ a_thr_ratio, int32_t* __restrict__ a_excl, int32_t* __restrict__ a_count,
               uint8_t* __restrict__ gath_q, uint8_t* __restrict__ gath_t, int32_t* __restrict__ vitems,
               int32_t* __restrict__ vitem_count, unsigned long long* __restrict__ stats) {
  __shared__ int2 warp_tot[kBlock / 32];
  const int pair = blockIdx.x;
  const PairInfo pi = load_pair(pairs, k_per_image, pair, n_img, k_stride);
  const int qi = pi.qi, ti = pi.ti, kq = pi.kq, kt = pi.kt;
  const int64_t base = static_cast<int64_t>(pair) * k_stride;
  const int64_t qbase = static_cast<int64_t>(qi) * k_stride;
  const int64_t tbase = static_cast<int64_t>(ti) * k_stride;
  // warp w owns the contiguous rows [w * seg, (w + 1) * seg), seg = 32 * iters; iteration i of a warp
  // covers 32 consecutive rows (coalesced), ordered slots come from ballots + one block scan of warp totals
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int iters = kt > 0 ? (kq + kBlock - 1) / kBlock : 0;
  const int row0 = warp * iters * 32;
  int n_keep = 0, n_open = 0;                                   // warp totals (same value in every lane)
#pragma unroll 4
  for (int i = 0; i < iters; ++i) {
    const int r = row0 + i * 32 + lane;
    RowClass c;
    c.keep = false; c.open = false;
    if (r < kq) c = classify_row(rowres_o[base + r], kt, ratio);
    n_keep += __popc(__ballot_sync(0xffffffffu, c.keep));
    n_open += __popc(__ballot_sync(0xffffffffu, c.open));
  }
  int slot, aslot, tot_b, tot_a;
  block_scan2<kBlock>(lane == 0 ? n_keep : 0, lane == 0 ? n_open : 0, slot, aslot, tot_b, tot_a, warp_tot);
  slot = __shfl_sync(0xffffffffu, slot, 0);                     // offset of this warp's first kept row
  aslot = __shfl_sync(0xffffffffu, aslot, 0);
  if (n_keep) {
    const unsigned below = (1u << lane) - 1u;
    for (int i = 0; i < iters; ++i) {
      const int r = row0 + i * 32 + lane;
      int4 v = make_int4(0, -1, 0, kRowReject);
      RowClass c;
      c.keep = false; c.open = false; c.d1f = 0.f;
      if (r < kq) {
        v = rowres_o[base + r];
        c = classify_row(v, kt, ratio);
      }
      const unsigned mk = __ballot_sync(0xffffffffu, c.keep), mo = __ballot_sync(0xffffffffu, c.open);
      if (c.keep) {
        const int64_t o = base + slot + __popc(mk & below);
        const int my_a = aslot + __popc(mo & below);
        cand_q[o] = r;
        cand_t[o] = v.y;
        cand_d[o] = c.d1f;
        cand_thr_mut[o] = v.x - sqnorm[tbase + v.y];
        cand_rslot[o] = c.open ? my_a : -1;
        if (gather_t) {
          const uint4* src = reinterpret_cast<const uint4*>(bank + (tbase + v.y) * 128);
          uint4* dst = reinterpret_cast<uint4*>(gath_t + o * 128);
#pragma unroll
          for (int k = 0; k < 8; ++k) dst[k] = src[k];
        }
        if (c.open) {
          const int64_t oa = base + my_a;
          a_thr_ratio[oa] = ratio_radius_sq(c.d1f, ratio) - sqnorm[qbase + r];
          a_excl[oa] = v.y;
          const uint4* src = reinterpret_cast<const uint4*>(bank + (qbase + r) * 128);
          uint4* dst = reinterpret_cast<uint4*>(gath_q + oa * 128);
#pragma unroll
          for (int k = 0; k < 8; ++k) dst[k] = src[k];
        }
      }
      slot += __popc(mk);
      aslot += __popc(mo);
    }
  }
  if (threadIdx.x == 0) {
    cand_count[pair] = tot_b;
    a_count[pair] = tot_a;
    stat_add(stats, 3, tot_b);
    stat_add(stats, 4, tot_a);
    // verify work list: 256-row blocks of list A (check 0, ratio) and list B (check 1, mutual)
    const int na = do_ratio ? (tot_a + 255) / 256 : 0;
    const int nb = gather_t ? (tot_b + 255) / 256 : 0;
    if (na + nb > 0) {
      int w = atomicAdd(vitem_count, na + nb);
      for (int b = 0; b < na; ++b, ++w) reinterpret_cast<int2*>(vitems)[w] = make_int2(pair, b * 2);
      for (int b = 0; b < nb; ++b, ++w) reinterpret_cast<int2*>(vitems)[w] = make_int2(pair, b * 2 + 1);
    }
  }
}

}  // namespace b200m
