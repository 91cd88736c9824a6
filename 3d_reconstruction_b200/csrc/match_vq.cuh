// Visual-word assignment of the BoW pair scheduler (matching.py:26-28: `vq(descriptors, codebook)`),
// the step right before the matcher: nearest codebook row in L2 for every descriptor. HBM/L2-bound
// and tiny next to the matching itself (290 images x 2,048 descriptors x 200 words), so plain CUDA
// cores in fp64 -- the arithmetic scipy's vq uses (|x|^2 + |c|^2 - 2 x.c in double, first minimum
// wins), which keeps the word histograms, and with them the view graph, equal to the reference's.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200m {

constexpr int kVqObsPerBlock = 32;
constexpr int kVqLanesPerObs = 8;
constexpr int kVqThreads = kVqObsPerBlock * kVqLanesPerObs;   // 256

// obs float32 [n_obs, dim] (dim <= 256), codebook float64 [n_codes, dim], code_sqr float64 [n_codes]
__global__ void __launch_bounds__(kVqThreads)
vq_argmin_kernel(const float* __restrict__ obs, int64_t n_obs, int dim, const double* __restrict__ codebook,
                 const double* __restrict__ code_sqr, int n_codes, int32_t* __restrict__ out_code,
                 double* __restrict__ out_dist) {
  extern __shared__ double s_obs[];                 // [kVqObsPerBlock][dim]
  __shared__ double s_sqr[kVqObsPerBlock];
  const int64_t o0 = static_cast<int64_t>(blockIdx.x) * kVqObsPerBlock;
  for (int i = threadIdx.x; i < kVqObsPerBlock * dim; i += kVqThreads) {
    const int64_t o = o0 + i / dim;
    s_obs[i] = o < n_obs ? static_cast<double>(obs[o * dim + i % dim]) : 0.0;
  }
  __syncthreads();
  const int lo = threadIdx.x / kVqLanesPerObs, j = threadIdx.x % kVqLanesPerObs;
  const double* x = s_obs + lo * dim;
  if (j == 0) {
    double s = 0.0;
    for (int d = 0; d < dim; ++d) s += x[d] * x[d];
    s_sqr[lo] = s;
  }
  __syncthreads();
  const double xs = s_sqr[lo];
  double best = 1.0e300;
  int bidx = 0x7fffffff;
  for (int c = j; c < n_codes; c += kVqLanesPerObs) {
    const double* cb = codebook + static_cast<int64_t>(c) * dim;
    double dot = 0.0;
#pragma unroll 8
    for (int d = 0; d < dim; ++d) dot += x[d] * cb[d];
    const double dist = (-2.0 * dot + xs) + code_sqr[c];
    if (dist < best) { best = dist; bidx = c; }       // increasing c: the first minimum wins
  }
  // the 8 lanes of an observation are consecutive lanes of one warp
#pragma unroll
  for (int off = kVqLanesPerObs / 2; off >= 1; off >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, off);
    const int oi = __shfl_xor_sync(0xffffffffu, bidx, off);
    if (ob < best || (ob == best && oi < bidx)) { best = ob; bidx = oi; }
  }
  const int64_t o = o0 + lo;
  if (j == 0 && o < n_obs) {
    out_code[o] = bidx;
    if (out_dist) out_dist[o] = sqrt(best > 0.0 ? best : 0.0);
  }
}

__global__ void vq_code_sqr_kernel(const double* __restrict__ codebook, int dim, int n_codes, double* __restrict__ code_sqr) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_codes) return;
  double s = 0.0;
  for (int d = 0; d < dim; ++d) s += codebook[static_cast<int64_t>(c) * dim + d] * codebook[static_cast<int64_t>(c) * dim + d];
  code_sqr[c] = s;
}

}  // namespace b200m
