// Small HBM-bound kernels either side of the sweep: bank preparation (norms + padding) and the
// finalisation of row/column results into cv2-shaped outputs.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

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

__device__ __forceinline__ float dist_from_sq(int d2) {
  // exact: d2 < 2^24 is representable, sqrt is IEEE round-to-nearest == cv2's std::sqrt(float)
  return __fsqrt_rn(static_cast<float>(d2));
}

// knnMatch(k=2) layout: out_train_idx / out_dist are [n_pairs_total, k_stride, 2]
__global__ void knn2_finalize_kernel(const int4* __restrict__ rowres, const int32_t* __restrict__ pairs,
                                     const int32_t* __restrict__ k_per_image, int64_t k_stride,
                                     int32_t* __restrict__ out_idx, float* __restrict__ out_dist) {
  const int pair = blockIdx.y;
  const int kq = k_per_image[pairs[2 * pair]];
  const int kt = k_per_image[pairs[2 * pair + 1]];
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

// Ratio test + mutual check + order-preserving compaction. One CTA per pair.
//   keep q  <=>  has second neighbour and double(d1) < ratio * double(d2)   (ratio > 0)
//               and colbest[t1].row == q                                    (mutual)
// The compare is done in double on the float32 distances: exactly Python's
// `m.distance < 0.75 * n.distance` on cv2 DMatch objects.
template <int kBlock>
__global__ void __launch_bounds__(kBlock)
match_finalize_kernel(const int4* __restrict__ rowres, const unsigned long long* __restrict__ colbest,
                      const int32_t* __restrict__ pairs, const int32_t* __restrict__ k_per_image,
                      int64_t k_stride, double ratio, int mutual, int32_t* __restrict__ out_q,
                      int32_t* __restrict__ out_t, float* __restrict__ out_d, int32_t* __restrict__ out_count,
                      int64_t cap) {
  __shared__ int warp_sums[kBlock / 32];
  __shared__ int running;
  const int pair = blockIdx.x;
  const int kq = k_per_image[pairs[2 * pair]];
  const int kt = k_per_image[pairs[2 * pair + 1]];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) running = 0;
  __syncthreads();
  const int64_t base = static_cast<int64_t>(pair) * k_stride;
  const int64_t obase = static_cast<int64_t>(pair) * cap;
  for (int r0 = 0; r0 < kq && kt > 0; r0 += kBlock) {
    const int r = r0 + threadIdx.x;
    bool keep = false;
    int t1 = -1;
    float d1f = 0.f;
    if (r < kq) {
      const int4 v = rowres[base + r];
      const bool has1 = v.y >= 0 && v.y < kt;
      const bool has2 = v.w >= 0 && v.w < kt;
      if (has1) {
        t1 = v.y;
        d1f = dist_from_sq(v.x);
        if (ratio > 0.0) {
          keep = has2 && (static_cast<double>(d1f) < ratio * static_cast<double>(dist_from_sq(v.z)));
        } else {
          keep = true;
        }
        if (keep && mutual) {
          const unsigned long long cb = colbest[base + t1];
          keep = static_cast<uint32_t>(cb & 0xffffffffull) == static_cast<uint32_t>(r);
        }
      }
    }
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
    const int start = running;
    if (keep) {
      const int64_t slot = static_cast<int64_t>(start) + woff + prefix;
      if (slot < cap) {
        out_q[obase + slot] = r;
        out_t[obase + slot] = t1;
        out_d[obase + slot] = d1f;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) running = start + total;
    __syncthreads();
  }
  if (threadIdx.x == 0) out_count[pair] = running;
}

}  // namespace b200m
