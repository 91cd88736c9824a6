// Float path: mutual nearest neighbours by inner product for SuperPoint / DISK style descriptors
// (D = 128 or 256), bf16 operands on tcgen05 (kind::f16, fp32 accumulate in TMEM).
//
// Stands in for the reference's matcher head in its raw-similarity form:
//   lightglue/lightglue.py:223      sim = einsum('bmd,bnd->bmn')              (dense contraction)
//   lightglue/lightglue.py:237      scores.max(2), scores.max(1)              (row / column arg-best)
//   lightglue/lightglue.py:239-250  mutual check + score threshold            (filter_matches)
//   lightglue/lightglue.py:417-426  match list, ascending query index
// The column arg-best of a pair is the row arg-best of the transposed pair, so every pair is swept in
// both directions by the same kernel (row arg-max fused into the TMEM epilogue, the similarity matrix
// never reaches HBM) and ip_finalize_kernel applies mutual check, threshold and packing.
//
// ip_sweep_kernel<kDim>: one CTA per SM, persistent over work items (pair direction, block of 128 rows).
//   warp 0       TMA producer: the row block (kDim/64 K-blocks of 128 rows x 128 B, double buffered) and
//                the column image as a stream of (128 columns x 64 elements) K-blocks through a ring
//   warp 1       MMA issuer: per column tile kDim/16 tcgen05.mma.kind::f16 (M128 N128 K16, bf16 -> f32)
//                into one of four 128 x 128 fp32 accumulators (all 512 TMEM columns), round robin
//   warps 4-19   epilogue: warp (lane quadrant q, accumulator e) drains every 4th column tile for rows
//                32q..32q+31: running (best value, lowest column) per row; the four partial results of a
//                row meet through shared memory at the end of the item.
// Bound: the shared-memory port (128 B/clk). An N = 128 bf16 MMA reads 8 KiB of operands per 64 clk (128 B/clk)
// and the column stream is written at 64 B/clk (one 128-row block per column image): 192 B/clk asked of 128,
// a ceiling of 67 % of the tensor peak -- where the D = 256 sweep sits (1.48 of 2.25 PFLOP/s). N = 256 tiles
// (96 + 64 B/clk, 80 %) would need the row block single-buffered at D = 256 and expose a TMA round trip per
// 2,048-column item; two row blocks per column stream do not fit next to a useful column ring.
#pragma once
#include <cuda_bf16.h>

#include "match_sweep.cuh"

namespace b200m {
namespace ip {

constexpr int kKB = 64;                     // bf16 elements per K-block (one 128-byte swizzle atom)
constexpr int kRowsF = 128;                 // rows per work item (one UMMA M-tile)
constexpr int kColsF = 128;                 // columns per tile / accumulator (UMMA N)
constexpr int kBlockBytes = 128 * 128;      // 16 KiB: 128 rows x one K-block
constexpr int kAStagesF = 2;
constexpr int kBStagesF = 5;
constexpr int kAccs = 4;
constexpr float kNegInf = -3.0e38f;
constexpr float kLseFloor = -1.0e30f;       // "minus infinity" of the log-sum-exp sweep (its products with log2(e) stay finite)
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

// epilogue of ip_sweep_kernel
constexpr int kIpArgmax = 0;      // row arg-max of the raw similarity
constexpr int kIpLse = 1;         // row log-sum-exp of the similarity -> nb = logsigmoid(z_row) - lse
constexpr int kIpBiasArgmax = 2;  // row arg-max of 2*sim + nb_other[column], + nb_own[row] at the end

template <int kDim>
struct IpSmem {
  uint8_t a[kAStagesF][kDim / kKB][kBlockBytes];
  uint8_t b[kBStagesF][kBlockBytes];
  float2 rowmerge[2][kAccs - 1][kRowsF];    // (value, column as int bits)
  uint64_t a_full[kAStagesF], a_empty[kAStagesF];
  uint64_t b_full[kBStagesF], b_empty[kBStagesF];
  uint64_t t_full[kAccs], t_empty[kAccs];
  uint32_t tmem_base;
};

struct IpParams {
  const int32_t* k_per_image;   // [n_img]
  const int32_t* pairs;         // [n_pairs, 2] (query image, train image)
  int64_t k_stride;
  int32_t n_pairs;
  int32_t n_dirs;               // 2: both directions (mutual check), 1: rows = query image only
  int32_t blocks_per_image;     // k_stride / 128
  int32_t n_img;                // pair-list entries outside [0, n_img) are skipped (load_pair)
  // out: [n_dirs][n_pairs][k_stride] (best similarity, best column); direction 0: rows = query image
  float2* rowbest;
  // dual-softmax head only (kIpLse writes, kIpBiasArgmax reads):
  const float* z;               // [n_img, k_stride] matchability logits (lightglue.py:224-225)
  float* nb;                    // [2][n_pairs][k_stride] logsigmoid(z_row) - log-sum-exp of the row
  const int32_t* gate;          // optional: the launch does nothing unless *gate != 0 (fallback of the single-sweep log-sum-exp)
};

// kind::f16 instruction descriptor: bf16 A and B (K-major), fp32 accumulator
__host__ __device__ constexpr uint32_t make_idesc_bf16(uint32_t m, uint32_t n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

__device__ __forceinline__ void mma_bf16_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                            uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

__device__ __forceinline__ float fmax16(const int (&v)[16]) {
  float m[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) m[i] = fmaxf(__int_as_float(v[2 * i]), __int_as_float(v[2 * i + 1]));
#pragma unroll
  for (int i = 0; i < 4; ++i) m[i] = fmaxf(m[i], m[i + 4]);
  return fmaxf(fmaxf(m[0], m[1]), fmaxf(m[2], m[3]));
}

// sixteen columns of one row starting at column c0; n_cols = valid columns of the sweep.
// A new row maximum only records the group (c0) and keeps a copy of its sixteen values; the column inside
// the group is looked up once per work item (ip_locate). With 32 independent rows per warp nearly every
// group holds a new maximum for SOME lane, so the per-group cost of that path is what matters: sixteen
// predicated moves instead of a sixteen-step compare / select chain.
__device__ __forceinline__ void ip_process16(int (&v)[16], int c0, int n_cols, float& best, int& bidx, int (&keep)[16]) {
  if (c0 + 16 > n_cols) {   // ragged end (and padding columns: zero vectors must not compete)
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (c0 + j >= n_cols) v[j] = __float_as_int(kNegInf);
  }
  const float m = fmax16(v);
  if (m > best) {           // strict: an equal value further right never replaces the lowest column
    best = m;
    bidx = c0;
#pragma unroll
    for (int j = 0; j < 16; ++j) keep[j] = v[j];
  }
}
// column of the maximum: lowest index inside the recorded group
__device__ __forceinline__ int ip_locate(const int (&keep)[16], float best, int bidx) {
  if (bidx < 0) return bidx;
  int j = 15;
#pragma unroll
  for (int jj = 14; jj >= 0; --jj)
    if (__int_as_float(keep[jj]) == best) j = jj;
  return bidx + j;
}

// ex2.approx.ftz.f32, one MUFU instruction: exp2f() without -use_fast_math wraps it in a range test and two scalings
// for results below 2^-126; such terms are below fp32 resolution of every sum they enter here.
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// sixteen columns of one row: online log-sum-exp, (m, s) with sum_j exp(v_j) = s * exp(m)
__device__ __forceinline__ void lse_process16(int (&v)[16], int c0, int n_cols, float& m, float& s) {
  if (c0 + 16 > n_cols) {
#pragma unroll
    for (int j = 0; j < 16; ++j)
      if (c0 + j >= n_cols) v[j] = __float_as_int(kLseFloor);
  }
  const float nm = fmaxf(m, fmax16(v));
  s *= exp2f((m - nm) * kLog2e);
  m = nm;
  const float mb = nm * kLog2e;
  float acc0 = 0.f, acc1 = 0.f;
#pragma unroll
  for (int j = 0; j < 16; j += 2) {
    acc0 += ex2_ftz(fmaf(__int_as_float(v[j]), kLog2e, -mb));
    acc1 += ex2_ftz(fmaf(__int_as_float(v[j + 1]), kLog2e, -mb));
  }
  s += acc0 + acc1;
}

// sixteen columns: v <- 2 v + bias[column] (bias is the same for every row: one broadcast load per four columns)
__device__ __forceinline__ void bias16(int (&v)[16], const float* __restrict__ bias) {
#pragma unroll
  for (int j = 0; j < 16; j += 4) {
    const float4 b = __ldg(reinterpret_cast<const float4*>(bias + j));
    v[j] = __float_as_int(fmaf(2.f, __int_as_float(v[j]), b.x));
    v[j + 1] = __float_as_int(fmaf(2.f, __int_as_float(v[j + 1]), b.y));
    v[j + 2] = __float_as_int(fmaf(2.f, __int_as_float(v[j + 2]), b.z));
    v[j + 3] = __float_as_int(fmaf(2.f, __int_as_float(v[j + 3]), b.w));
  }
}

__device__ __forceinline__ float log_sigmoid(float z) {   // F.logsigmoid
  return fminf(z, 0.f) - log1pf(expf(-fabsf(z)));
}

template <int kDim, int kMode>
__global__ void __launch_bounds__(kThreads, 1)
ip_sweep_kernel(const __grid_constant__ CUtensorMap tmap, const IpParams p) {
  constexpr int kNKB = kDim / kKB;          // K-blocks per descriptor
  extern __shared__ uint8_t smem_raw[];
  IpSmem<kDim>& sm = *reinterpret_cast<IpSmem<kDim>*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  if (p.gate && __ldg(p.gate) == 0) return;

  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t lane = lane_id();
  const int n_items = p.n_dirs * p.n_pairs * p.blocks_per_image;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmap);
    for (int i = 0; i < kAStagesF; ++i) { mbar_init(&sm.a_full[i], 1); mbar_init(&sm.a_empty[i], 1); }
    for (int i = 0; i < kBStagesF; ++i) { mbar_init(&sm.b_full[i], 1); mbar_init(&sm.b_empty[i], 1); }
    for (int i = 0; i < kAccs; ++i) { mbar_init(&sm.t_full[i], 1); mbar_init(&sm.t_empty[i], 4); }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc<512>(&sm.tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;

  // item -> (pair, direction, row block); rows of direction 0 are the query image's
  auto decode = [&](int item, int& pair, int& dir, int& blk, int& row_img, int& col_img, int& n_rows, int& n_cols) {
    const int pd = item / p.blocks_per_image;
    blk = item - pd * p.blocks_per_image;
    pair = pd / p.n_dirs;
    dir = pd - pair * p.n_dirs;
    const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
    row_img = dir ? pi.ti : pi.qi;
    col_img = dir ? pi.qi : pi.ti;
    n_rows = min((dir ? pi.kt : pi.kq) - blk * kRowsF, kRowsF);
    n_cols = dir ? pi.kq : pi.kt;
    return n_rows > 0 && n_cols > 0;
  };

  if (warp == 0) {
    // ------------------------------------------------------------- TMA producer
    if (lane == 0) {
      uint32_t a_it = 0, b_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int pair, dir, blk, row_img, col_img, n_rows, n_cols;
        if (!decode(item, pair, dir, blk, row_img, col_img, n_rows, n_cols)) continue;
        const uint32_t as = a_it % kAStagesF, aph = (a_it / kAStagesF) & 1;
        mbar_wait(&sm.a_empty[as], aph ^ 1);
        mbar_arrive_expect_tx(&sm.a_full[as], kNKB * kBlockBytes);
        const int32_t arow = static_cast<int32_t>(row_img * p.k_stride + blk * kRowsF);
        for (int kb = 0; kb < kNKB; ++kb) tma_load_2d(sm.a[as][kb], &tmap, kb * kKB, arow, &sm.a_full[as]);
        ++a_it;
        const int ntiles = (n_cols + kColsF - 1) / kColsF;
        for (int t = 0; t < ntiles; ++t) {
          const int32_t crow = static_cast<int32_t>(col_img * p.k_stride + t * kColsF);
          for (int kb = 0; kb < kNKB; ++kb) {
            const uint32_t bs = b_it % kBStagesF, bph = (b_it / kBStagesF) & 1;
            mbar_wait(&sm.b_empty[bs], bph ^ 1);
            mbar_arrive_expect_tx(&sm.b_full[bs], kBlockBytes);
            tma_load_2d(sm.b[bs], &tmap, kb * kKB, crow, &sm.b_full[bs]);
            ++b_it;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------- MMA issuer
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(kRowsF, kColsF);
      uint32_t a_it = 0, b_it = 0, t_it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
        int pair, dir, blk, row_img, col_img, n_rows, n_cols;
        if (!decode(item, pair, dir, blk, row_img, col_img, n_rows, n_cols)) continue;
        const uint32_t as = a_it % kAStagesF, aph = (a_it / kAStagesF) & 1;
        mbar_wait(&sm.a_full[as], aph);
        const int ntiles = (n_cols + kColsF - 1) / kColsF;
        for (int t = 0; t < ntiles; ++t) {
          const uint32_t acc = t_it % kAccs, tph = (t_it / kAccs) & 1;
          mbar_wait(&sm.t_empty[acc], tph ^ 1);
          tc_fence_after();
          const uint32_t tacc = tmem_base + acc * kColsF;
#pragma unroll
          for (int kb = 0; kb < kNKB; ++kb) {
            const uint32_t bs = b_it % kBStagesF, bph = (b_it / kBStagesF) & 1;
            mbar_wait(&sm.b_full[bs], bph);
            tc_fence_after();
            const uint64_t adesc = make_desc_k128(smem_u32(sm.a[as][kb]));
            const uint64_t bdesc = make_desc_k128(smem_u32(sm.b[bs]));
#pragma unroll
            for (int k = 0; k < kKB / 16; ++k)
              mma_bf16_ss(tacc, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) > 0 ? 1u : 0u);
            tc_commit(&sm.b_empty[bs]);
            ++b_it;
          }
          tc_commit(&sm.t_full[acc]);
          ++t_it;
        }
        tc_commit(&sm.a_empty[as]);
        ++a_it;
      }
    }
  } else if (warp >= kFirstEpiWarp) {
    // ------------------------------------------------------------- epilogue
    const uint32_t quad = warp & 3;                   // TMEM lane quadrant this warp may read
    const uint32_t e = (warp - kFirstEpiWarp) >> 2;   // accumulator this warp drains
    const int rl = quad * 32 + lane;                  // row inside the item's block
    const uint32_t taddr = tmem_base + ((quad * 32u) << 16) + e * kColsF;
    uint32_t t_seen = 0;                              // tiles of all items so far (round robin over accumulators)
    uint32_t my_it = 0, item_it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x) {
      int pair, dir, blk, row_img, col_img, n_rows, n_cols;
      if (!decode(item, pair, dir, blk, row_img, col_img, n_rows, n_cols)) continue;
      const uint32_t par = item_it & 1;
      ++item_it;
      const int ntiles = (n_cols + kColsF - 1) / kColsF;
      float best = kMode == kIpLse ? kLseFloor : kNegInf;   // kIpLse: running maximum
      int bidx = -1;                                          // kIpLse: unused
      float lsum = 0.f;                                       // kIpLse: running sum of exp(v - best)
      int keep[16];                                           // arg-max modes: the sixteen values of the best group
#pragma unroll
      for (int j = 0; j < 16; ++j) keep[j] = 0;
      const float* bias = nullptr;                            // kIpBiasArgmax: nb of the opposite direction (= per column)
      if constexpr (kMode == kIpBiasArgmax)
        bias = p.nb + (static_cast<int64_t>(1 - dir) * p.n_pairs + pair) * p.k_stride;
      auto process = [&](int (&v)[16], int c0) {
        if constexpr (kMode == kIpLse) {
          lse_process16(v, c0, n_cols, best, lsum);
        } else {
          if constexpr (kMode == kIpBiasArgmax) bias16(v, bias + c0);
          ip_process16(v, c0, n_cols, best, bidx, keep);
        }
      };
      // this warp owns the tiles whose global sequence number is e modulo 4
      int t = static_cast<int>((e + kAccs - (t_seen % kAccs)) % kAccs);
      for (; t < ntiles; t += kAccs) {
        mbar_wait(&sm.t_full[e], my_it & 1);
        tc_fence_after();
        ++my_it;
        const int col0 = t * kColsF;
        int va[16], vb[16];
        tmem_ld_32x16(taddr, va);
        tmem_ld_fence16(va);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          tmem_ld_32x16(taddr + c * 32 + 16, vb);
          process(va, col0 + c * 32);
          tmem_ld_fence16(vb);
          if (c < 3) {
            tmem_ld_32x16(taddr + (c + 1) * 32, va);
          } else {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm.t_empty[e]);
          }
          process(vb, col0 + c * 32 + 16);
          if (c < 3) tmem_ld_fence16(va);
        }
      }
      t_seen += ntiles;
      if constexpr (kMode != kIpLse) bidx = ip_locate(keep, best, bidx);
      // the four partial results of a row meet here
      if (e > 0) sm.rowmerge[par][e - 1][rl] = make_float2(best, kMode == kIpLse ? lsum : __int_as_float(bidx));
      named_bar_sync(kEpiBarrier, kEpiWarps * 32);
      if (e == 0 && rl < n_rows) {
        const int64_t o = (static_cast<int64_t>(dir) * p.n_pairs + pair) * p.k_stride + blk * kRowsF + rl;
        if constexpr (kMode == kIpLse) {
          float mm = best;
#pragma unroll
          for (int oo = 0; oo < kAccs - 1; ++oo) mm = fmaxf(mm, sm.rowmerge[par][oo][rl].x);
          float ss = lsum * exp2f((best - mm) * kLog2e);
#pragma unroll
          for (int oo = 0; oo < kAccs - 1; ++oo) {
            const float2 q = sm.rowmerge[par][oo][rl];
            ss += q.y * exp2f((q.x - mm) * kLog2e);
          }
          const float lse = mm + logf(ss);
          p.nb[o] = log_sigmoid(__ldg(p.z + static_cast<int64_t>(row_img) * p.k_stride + blk * kRowsF + rl)) - lse;
        } else {
#pragma unroll
          for (int oo = 0; oo < kAccs - 1; ++oo) {
            const float2 q = sm.rowmerge[par][oo][rl];
            const int qi = __float_as_int(q.y);
            if (qi >= 0 && (q.x > best || (q.x == best && qi < bidx) || bidx < 0)) { best = q.x; bidx = qi; }
          }
          if constexpr (kMode == kIpBiasArgmax) best += p.nb[o];   // the row's own term: the full log assignment score
          p.rowbest[o] = make_float2(best, __int_as_float(bidx));
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

// Mutual check + threshold + ascending-query packing (lightglue.py:239-250, 417-426), one CTA per pair.
// exp_scores: rowbest holds log assignment scores (dual-softmax head); the threshold applies to exp(score).
template <int kBlock>
__global__ void __launch_bounds__(kBlock)
ip_finalize_kernel(const float2* __restrict__ rowbest, const int32_t* __restrict__ pairs,
                   const int32_t* __restrict__ k_per_image, int n_img, int64_t k_stride, int32_t n_pairs, float threshold,
                   int use_threshold, int mutual, int exp_scores, int32_t* __restrict__ out_q, int32_t* __restrict__ out_t,
                   float* __restrict__ out_s, int32_t* __restrict__ out_count, int64_t cap) {
  __shared__ int warp_sums[kBlock / 32];
  __shared__ int running;
  const int pair = blockIdx.x;
  const PairInfo pi = load_pair(pairs, k_per_image, pair, n_img, k_stride);
  const int kq = pi.kq, kt = pi.kt;
  if (threadIdx.x == 0) running = 0;
  __syncthreads();
  const float2* fwd = rowbest + static_cast<int64_t>(pair) * k_stride;
  const float2* bwd = rowbest + (static_cast<int64_t>(n_pairs) + pair) * k_stride;
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
      if (keep && mutual) keep = __float_as_int(bwd[j].y) == r;
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

}  // namespace ip
}  // namespace b200m
