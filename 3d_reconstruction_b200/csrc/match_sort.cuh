// Norm-sorted copy of the descriptor bank (one-off, HBM-bound): per image the rows are ordered by
// (squared norm, original row) so that 32 consecutive rows -- one chunk of the chunk sweep -- have
// nearly equal norms. Bitonic sort of 64-bit keys (norm << 32 | row) per image: shared-memory
// passes for strides < 2048, global passes above. Keys are unique, so the order is total and
// deterministic. Padding rows (norm = B200M_PAD_NORM) sort behind every real row.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200m {

constexpr int kSortChunk = 2048;      // keys per shared-memory block
constexpr int kSortThreads = 1024;

// keys [n_img, n2]; n2 = power of two >= k_stride, fillers = ~0
__global__ void sort_init_kernel(const int32_t* __restrict__ sqnorm, int64_t k_stride, int64_t n2, int64_t total,
                                 unsigned long long* __restrict__ keys, int64_t key_stride) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t img = i / n2, r = i - img * n2;
    unsigned long long k = ~0ull;
    if (r < k_stride)
      k = (static_cast<unsigned long long>(static_cast<uint32_t>(sqnorm[img * k_stride + r])) << 32) | static_cast<uint32_t>(r);
    keys[img * key_stride + r] = k;
  }
}

__device__ __forceinline__ void cmp_swap(unsigned long long& a, unsigned long long& b, bool asc) {
  if ((a > b) == asc) { const unsigned long long t = a; a = b; b = t; }
}

// One CTA per block of 2048 keys. kFull: all bitonic stages up to size 2048; otherwise the tail of
// the merge of `size` (strides 1024 .. 1). Direction of element i: ascending iff (i & size) == 0.
template <bool kFull>
__global__ void __launch_bounds__(kSortThreads)
sort_smem_kernel(unsigned long long* __restrict__ keys, int64_t n2, int64_t key_stride, int64_t size_in) {
  __shared__ unsigned long long s[kSortChunk];
  const int64_t blocks_per_img = n2 / kSortChunk;
  const int64_t img = blockIdx.x / blocks_per_img, blk = blockIdx.x - img * blocks_per_img;
  unsigned long long* g = keys + img * key_stride + blk * kSortChunk;
  const int t = threadIdx.x;
  s[t] = g[t];
  s[t + kSortThreads] = g[t + kSortThreads];
  __syncthreads();
  const int64_t i0 = blk * kSortChunk;
  for (int64_t size = kFull ? 2 : size_in; size <= (kFull ? kSortChunk : size_in); size <<= 1) {
    for (int stride = static_cast<int>(size < kSortChunk ? size / 2 : kSortChunk / 2); stride >= 1; stride >>= 1) {
      const int lo = 2 * t - (t & (stride - 1));     // index of the lower element of this thread's pair
      const bool asc = ((i0 + lo) & size) == 0;
      unsigned long long a = s[lo], b = s[lo + stride];
      cmp_swap(a, b, asc);
      s[lo] = a;
      s[lo + stride] = b;
      __syncthreads();
    }
  }
  g[t] = s[t];
  g[t + kSortThreads] = s[t + kSortThreads];
}

// one compare-exchange pass at a stride >= 2048
__global__ void sort_global_kernel(unsigned long long* __restrict__ keys, int64_t n2, int64_t key_stride, int64_t n_img,
                                   int64_t size, int64_t stride) {
  const int64_t half = n2 / 2, total = n_img * half;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t img = i / half, t = i - img * half;
    const int64_t lo = 2 * t - (t & (stride - 1));
    const bool asc = (lo & size) == 0;
    unsigned long long* g = keys + img * key_stride;
    unsigned long long a = g[lo], b = g[lo + stride];
    if ((a > b) == asc) { g[lo] = b; g[lo + stride] = a; }
  }
}

// keys -> perm / inv / sorted norms
__global__ void sort_extract_kernel(const unsigned long long* __restrict__ keys, int64_t key_stride, int64_t k_stride,
                                    int64_t total, int32_t* __restrict__ perm, int32_t* __restrict__ inv,
                                    int32_t* __restrict__ snorm) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t img = i / k_stride, pos = i - img * k_stride;
    const unsigned long long k = keys[img * key_stride + pos];
    const int32_t row = static_cast<int32_t>(k & 0xffffffffu);
    perm[i] = row;
    inv[img * k_stride + row] = static_cast<int32_t>(pos);
    snorm[i] = static_cast<int32_t>(k >> 32);
  }
}

// sorted_bank[img][pos] = bank[img][perm[pos]]; one warp per row, lane l moves bytes [4l, 4l+4)
__global__ void sort_gather_kernel(const uint8_t* __restrict__ bank, const int32_t* __restrict__ perm, int64_t k_stride,
                                   int64_t n_rows, uint8_t* __restrict__ sbank) {
  const int64_t warp_global = (static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const int64_t n_warps = (static_cast<int64_t>(gridDim.x) * blockDim.x) >> 5;
  const int lane = threadIdx.x & 31;
  for (int64_t row = warp_global; row < n_rows; row += n_warps) {
    const int64_t img = row / k_stride;
    const int64_t src = img * k_stride + perm[row];
    reinterpret_cast<uint32_t*>(sbank + row * 128)[lane] = reinterpret_cast<const uint32_t*>(bank + src * 128)[lane];
  }
}

// chunk minima (= first norm of the sorted chunk) and the largest spread of a 32-row / 64-row chunk over real rows per image
__global__ void sort_chunk_stats_kernel(const int32_t* __restrict__ snorm, const int32_t* __restrict__ k_per_image,
                                        int64_t k_stride, int64_t n_chunks, int32_t* __restrict__ cmin,
                                        int32_t* __restrict__ spread, int32_t* __restrict__ spread64) {
  for (int64_t c = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; c < n_chunks;
       c += static_cast<int64_t>(gridDim.x) * blockDim.x) {
    const int64_t row0 = c * 32;
    const int64_t img = row0 / k_stride;
    const int pos0 = static_cast<int>(row0 - img * k_stride);
    const int k = min(k_per_image[img], static_cast<int>(k_stride));
    const int first = snorm[row0];
    cmin[c] = first;
    if (pos0 < k) {
      const int last_pos = min(pos0 + 31, k - 1);
      atomicMax(spread + img, snorm[img * k_stride + last_pos] - first);
      if ((c & 1) == 0) {          // the 64-row chunk starting here (chunk sweep option "chunk" = 64)
        const int last64 = min(pos0 + 63, k - 1);
        atomicMax(spread64 + img, snorm[img * k_stride + last64] - first);
      }
    }
  }
}

}  // namespace b200m
