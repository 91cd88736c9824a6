// Shared by every kernel that walks the pair list: one place where a (query image, train image) entry and
// the per-image row counts are read and sanitised, so that a bad pair list or row count coming through the
// C-ABI can never turn into an out-of-bounds TMA coordinate or table index.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200m {

struct PairInfo {
  int qi, ti;   // query / train image (0 when the entry is invalid)
  int kq, kt;   // their valid rows, clamped to [0, k_stride]; both 0 when the entry is invalid
};

// An entry with an image index outside [0, n_img) behaves like a pair of EMPTY images: no work item, no
// candidates, no matches. mark_invalid_pairs_kernel then reports it as out_count = -1 (the host layer raises).
__device__ __forceinline__ PairInfo load_pair(const int32_t* __restrict__ pairs, const int32_t* __restrict__ k_per_image,
                                              int pair, int n_img, int64_t k_stride) {
  PairInfo pi;
  pi.qi = __ldg(pairs + 2 * pair);
  pi.ti = __ldg(pairs + 2 * pair + 1);
  const bool ok = static_cast<unsigned>(pi.qi) < static_cast<unsigned>(n_img) &&
                  static_cast<unsigned>(pi.ti) < static_cast<unsigned>(n_img);
  if (!ok) {
    pi.qi = pi.ti = pi.kq = pi.kt = 0;
    return pi;
  }
  const int ks = static_cast<int>(k_stride);
  pi.kq = min(max(__ldg(k_per_image + pi.qi), 0), ks);
  pi.kt = min(max(__ldg(k_per_image + pi.ti), 0), ks);
  return pi;
}

// Counters of one b200m_match_pairs_u8 call (first block of the workspace, read with b200m_match_stats_read):
// kStatSlots copies of kStatFields 64-bit counters; a CTA adds into copy (blockIdx.x % kStatSlots) so that the
// atomics of a launch do not pile up on one address. Fields:
//   0 query rows seen by the refine kernel        1 rows whose winning chunk was evaluated exactly (refine step 2)
//   2 rows sent to the exact fallback sweep        3 candidates (rows surviving the Lowe test, mutual check pending)
//   4 rows whose Lowe test went to the verify sweep 5 matches emitted
constexpr int kStatSlots = 32;
constexpr int kStatFields = 8;
__device__ __forceinline__ void stat_add(unsigned long long* stats, int field, unsigned long long v) {
  if (stats && v) atomicAdd(stats + (blockIdx.x % kStatSlots) * kStatFields + field, v);
}

// out_count[p] = -1 for every pair-list entry that names an image outside the bank
__global__ void mark_invalid_pairs_kernel(const int32_t* __restrict__ pairs, int64_t n_pairs, int n_img,
                                          int32_t* __restrict__ out_count) {
  const int64_t p = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (p >= n_pairs) return;
  const int qi = pairs[2 * p], ti = pairs[2 * p + 1];
  if (static_cast<unsigned>(qi) >= static_cast<unsigned>(n_img) || static_cast<unsigned>(ti) >= static_cast<unsigned>(n_img))
    out_count[p] = -1;
}

}  // namespace b200m
