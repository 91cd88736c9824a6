// Small HBM-bound kernels either side of the sweep: bank preparation (norms + padding) and the
// finalisation of row/column results into cv2-shaped outputs.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "match_common.cuh"

namespace b200m {

// One warp per descriptor row: lane l owns bytes [4l, 4l+4). Padding rows are zeroed and get the
// sentinel norm so that they can never win a row or column minimum.
// (replaces the per-pair re-packing at matching.py:107-120 with a one-off pass over the bank)
__global__ void bank_prepare_kernel(uint8_t* __restrict__ bank, int64_t n_rows, int64_t k_stride,
                                    const int32_t* __restrict__ k_per_image, int32_t* __restrict__ sqnorm,
                                    int pad_norm) {
  const int64_t warp_global = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const int lane = threadIdx.x & 31;
  for (int64_t row = warp_global; row < n_rows; row += n_warps) {
    const int64_t img = row / k_stride;
    const int64_t r = row - img * k_stride;
    uint32_t* p = reinterpret_cast<uint32_t*>(bank + row * 128) + lane;
    if (r >= k_per_image[img]) {
      *p = 0u;
      if (lane == 0) sqnorm[row] = pad_norm;
    } else {
      const uint32_t w = *p;
      // sum of squares of four u8: dp4a(w, w, 0) on unsigned operands
      uint32_t s = __dp4a(w, w, 0u);
#pragma unroll
      for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
      if (lane == 0) sqnorm[row] = static_cast<int32_t>(s);
    }
  }
}

// Minimum squared norm of every 32 consecutive rows (one warp per chunk): the sweep's
// threshold-reject test needs min |col|^2 over a 32-column chunk.
__global__ void norm_chunk_min_kernel(const int32_t* __restrict__ sqnorm, int64_t n_rows,
                                      int32_t* __restrict__ nbmin32) {
  const int64_t warp_global = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const int lane = threadIdx.x & 31;
  for (int64_t c = warp_global; c < n_rows / 32; c += n_warps) {
    int m = sqnorm[c * 32 + lane];
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) m = min(m, __shfl_xor_sync(0xffffffffu, m, off));
    if (lane == 0) nbmin32[c] = m;
  }
}

__device__ __forceinline__ float dist_from_sq(int d2) {
  // exact: d2 < 2^24 is representable, sqrt is IEEE round-to-nearest == cv2's std::sqrt(float)
  return __fsqrt_rn(static_cast<float>(d2));
}

// knnMatch(k=2) layout: out_train_idx / out_dist are [n_pairs_total, k_stride, 2]
__global__ void knn2_finalize_kernel(const int4* __restrict__ rowres, const int32_t* __restrict__ pairs,
                                     const int32_t* __restrict__ k_per_image, int n_img, int64_t k_stride,
                                     int32_t* __restrict__ out_idx, float* __restrict__ out_dist) {
  const int pair = blockIdx.y;
  const PairInfo pi = load_pair(pairs, k_per_image, pair, n_img, k_stride);
  const int kq = pi.kq, kt = pi.kt;
  const float inf = __int_as_float(0x7f800000);
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < kq; r += gridDim.x * blockDim.x) {
    const int64_t o = static_cast<int64_t>(pair) * k_stride + r;
    int4 v = make_int4(0, -1, 0, -1);
    if (kt > 0) v = rowres[o];
    const bool has1 = kt > 0 && v.y >= 0 && v.y < kt;
    const bool has2 = kt > 0 && v.w >= 0 && v.w < kt;
    reinterpret_cast<int2*>(out_idx)[o] = make_int2(has1 ? v.y : -1, has2 ? v.w : -1);
    reinterpret_cast<float2*>(out_dist)[o] =
        make_float2(has1 ? dist_from_sq(v.x) : inf, has2 ? dist_from_sq(v.z) : inf);
  }
}

// Order-preserving block compaction helper: returns this thread's slot among the kept ones and
// advances the CTA-wide running total. All threads of the CTA must call it.
template <int kBlock>
__device__ __forceinline__ int block_compact_slot(bool keep, int* warp_sums, int* running) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned ballot = __ballot_sync(0xffffffffu, keep);
  const int prefix = __popc(ballot & ((1u << lane) - 1u));
  if (lane == 0) warp_sums[warp] = __popc(ballot);
  __syncthreads();
  int woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kBlock / 32; ++w) {
    const int c = warp_sums[w];
    if (w < warp) woff += c;
    total += c;
  }
  const int start = *running;
  __syncthreads();
  if (threadIdx.x == 0) *running = start + total;
  __syncthreads();
  return start + woff + prefix;
}

// Smallest squared distance D* at which a second neighbour no longer spoils the Lowe test of a row
// whose best float32 distance is d1f:  D* = min { D : double(d1f) < ratio * double(sqrtf(D)) }.
// The row passes  <=>  every other train row has d^2 >= D*  (sqrtf is monotone), which turns the
// float-exact `m.distance < ratio * n.distance` into an integer threshold for the verify sweep.
// inv_ratio = 1 / ratio, computed once by the caller: the double division is the most expensive step here and the
// estimate it feeds is corrected by the two loops anyway.
__device__ __forceinline__ int ratio_radius_sq(float d1f, double ratio, double inv_ratio) {
  const double d1 = static_cast<double>(d1f);
  double est = d1 * inv_ratio;
  est = est * est;
  long long D = est > 33554432.0 ? 33554432ll : static_cast<long long>(est);
  auto passes = [&](long long x) { return d1 < ratio * static_cast<double>(__fsqrt_rn(static_cast<float>(x))); };
  while (D > 0 && passes(D - 1)) --D;
  while (D < 33554432ll && !passes(D)) ++D;
  return static_cast<int>(D);
}
__device__ __forceinline__ int ratio_radius_sq(float d1f, double ratio) { return ratio_radius_sq(d1f, ratio, 1.0 / ratio); }

// Candidate selection after the first sweep; candidates are compacted in ascending query order.
//   exact2 != 0: rowres holds the exact second neighbour (kTop2 sweep): the Lowe test is decided here.
//   exact2 == 0: rowres holds an upper bound (kTop1 sweep): rows that pass against the bound are a
//                superset of the true survivors; the verify sweep decides with the exact radius.
// The compare is done in double on the float32 distances: exactly Python's
// `m.distance < 0.75 * n.distance` on cv2 DMatch objects. For the verify sweeps the candidates'
// query / train descriptors are gathered so they can be streamed with plain TMA.
template <int kBlock>
__global__ void __launch_bounds__(kBlock)
select_kernel(const int4* __restrict__ rowres, const int32_t* __restrict__ pairs,
              const int32_t* __restrict__ k_per_image, int n_img, int64_t k_stride, double ratio, int exact2,
              const uint8_t* __restrict__ bank, const int32_t* __restrict__ sqnorm, int gather_q, int gather_t,
              int32_t* __restrict__ cand_q, int32_t* __restrict__ cand_t, float* __restrict__ cand_d,
              int32_t* __restrict__ cand_thr_ratio, int32_t* __restrict__ cand_thr_mut,
              int32_t* __restrict__ cand_count, uint8_t* __restrict__ gath_q, uint8_t* __restrict__ gath_t) {
  __shared__ int warp_sums[kBlock / 32];
  __shared__ int running;
  const int pair = blockIdx.x;
  const PairInfo pi = load_pair(pairs, k_per_image, pair, n_img, k_stride);
  const int qi = pi.qi, ti = pi.ti, kq = pi.kq, kt = pi.kt;
  if (threadIdx.x == 0) running = 0;
  __syncthreads();
  const int64_t base = static_cast<int64_t>(pair) * k_stride;
  const int64_t qbase = static_cast<int64_t>(qi) * k_stride;
  const int64_t tbase = static_cast<int64_t>(ti) * k_stride;
  for (int r0 = 0; r0 < kq && kt > 0; r0 += kBlock) {
    const int r = r0 + threadIdx.x;
    bool keep = false;
    int4 v = make_int4(0, -1, 0, -1);
    float d1f = 0.f;
    if (r < kq) {
      v = rowres[base + r];
      const bool has1 = v.y >= 0 && v.y < kt;
      if (has1) {
        d1f = dist_from_sq(v.x);
        if (ratio > 0.0) {
          // v.z may be a sentinel / padding value above 2^24: float rounding keeps it above every real distance
          const bool has2 = exact2 ? (v.w >= 0 && v.w < kt) : (kt >= 2);
          keep = has2 && static_cast<double>(d1f) < ratio * static_cast<double>(dist_from_sq(v.z));
        } else {
          keep = true;
        }
      }
    }
    const int slot = block_compact_slot<kBlock>(keep, warp_sums, &running);
    if (keep) {
      const int64_t o = base + slot;
      cand_q[o] = r;
      cand_t[o] = v.y;
      cand_d[o] = d1f;
      cand_thr_mut[o] = v.x - sqnorm[tbase + v.y];
      if (gather_q) {
        cand_thr_ratio[o] = ratio_radius_sq(d1f, ratio) - sqnorm[qbase + r];
        const uint4* src = reinterpret_cast<const uint4*>(bank + (qbase + r) * 128);
        uint4* dst = reinterpret_cast<uint4*>(gath_q + o * 128);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = src[i];
      }
      if (gather_t) {
        const uint4* src = reinterpret_cast<const uint4*>(bank + (tbase + v.y) * 128);
        uint4* dst = reinterpret_cast<uint4*>(gath_t + o * 128);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = src[i];
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) cand_count[pair] = running;
}

// Final order-preserving compaction of the candidates that passed the verify sweeps.
template <int kBlock>
__global__ void __launch_bounds__(kBlock)
emit_kernel(const int32_t* __restrict__ cand_q, const int32_t* __restrict__ cand_t,
            const float* __restrict__ cand_d, const int32_t* __restrict__ viol_ratio,
            const int32_t* __restrict__ viol_mut, const int32_t* __restrict__ cand_count,
            const int32_t* __restrict__ cand_rslot, int64_t k_stride, int use_ratio, int use_mutual, int32_t* __restrict__ out_q, int32_t* __restrict__ out_t,
            float* __restrict__ out_d, int32_t* __restrict__ out_count, int64_t cap,
            unsigned long long* __restrict__ stats = nullptr) {
  __shared__ int warp_sums[kBlock / 32];
  __shared__ int running;
  const int pair = blockIdx.x;
  const int n = cand_count[pair];
  if (threadIdx.x == 0) running = 0;
  __syncthreads();
  const int64_t base = static_cast<int64_t>(pair) * k_stride;
  const int64_t obase = static_cast<int64_t>(pair) * cap;
  for (int s0 = 0; s0 < n; s0 += kBlock) {
    const int s = s0 + threadIdx.x;
    bool keep = s < n;
    if (keep && use_ratio) {
      // pipeline 2 keeps the rows with an open Lowe test in a list of their own (slot in cand_rslot, -1 = settled)
      const int rs = cand_rslot ? cand_rslot[base + s] : s;
      if (rs >= 0) keep = viol_ratio[base + rs] == 0;
    }
    if (keep && use_mutual) keep = viol_mut[base + s] == 0;
    const int slot = block_compact_slot<kBlock>(keep, warp_sums, &running);
    if (keep && slot < cap) {
      out_q[obase + slot] = cand_q[base + s];
      out_t[obase + slot] = cand_t[base + s];
      out_d[obase + slot] = cand_d[base + s];
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    out_count[pair] = running;
    stat_add(stats, 5, running);
  }
}

// [n_pairs, cap] slots + counts -> flat lists at offsets[pair] (exclusive scan of min(count, cap)).
__global__ void compact_kernel(const int32_t* __restrict__ q, const int32_t* __restrict__ t,
                               const float* __restrict__ d, const int32_t* __restrict__ count,
                               const int64_t* __restrict__ offsets, int64_t cap, int32_t* __restrict__ fq,
                               int32_t* __restrict__ ft, float* __restrict__ fd) {
  const int64_t pair = blockIdx.x;
  int64_t n = count[pair];
  if (n > cap) n = cap;
  const int64_t src = pair * cap, dst = offsets[pair];
  for (int64_t i = threadIdx.x; i < n; i += blockDim.x) {
    fq[dst + i] = q[src + i];
    ft[dst + i] = t[src + i];
    fd[dst + i] = d[src + i];
  }
}

}  // namespace b200m
