// Microbenchmark that sizes the fused column arg-max of the float sweep (development aid; not part of the product):
// cycles per warp-instruction of REDUX.MAX.S32, VOTE.BALLOT, SHFL, and of the whole per-column sequence
// (ordered key, redux, compare, ballot, owner select) with 1..4 warps per SM sub-partition.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_redux tools/ubench_redux.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int kMode>
__global__ void __launch_bounds__(512, 1) bench(int iters, int warps_per_smsp, long long* cycles, int* sink) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const bool on = (warp >> 2) < warps_per_smsp;
  int v[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = (threadIdx.x * 2654435761u + j * 40503u) ^ (j << 20);
  int acc = 0, keep_m = 0, keep_w = 0;
  __syncthreads();
  const long long t0 = clock64();
  if (on) {
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        if (kMode == 0) {
          acc += __reduce_max_sync(0xffffffffu, v[j] + acc);
        } else if (kMode == 1) {
          acc += __ballot_sync(0xffffffffu, (v[j] + acc) & 1);
        } else if (kMode == 2) {
          acc += __shfl_xor_sync(0xffffffffu, v[j] + acc, 16);
        } else if (kMode == 3) {      // independent redux (throughput, not latency)
          v[j] = __reduce_max_sync(0xffffffffu, v[j]) + it;
        } else if (kMode == 4) {      // the per-column sequence: key, redux, compare, ballot, owner select
          const int b = v[j] + it;
          const int key = b ^ ((b >> 31) & 0x7fffffff);
          const int m = __reduce_max_sync(0xffffffffu, key);
          const unsigned who = __ballot_sync(0xffffffffu, key == m);
          if (lane == j) { keep_m = m; keep_w = who; }
        } else if (kMode == 5) {      // two rows per thread: keys, max, select bit, redux, compare, ballot, owner select
          const int b0 = v[j] + it, b1 = v[(j + 5) & 15] - it;
          const int k0 = b0 ^ ((b0 >> 31) & 0x7fffffff), k1 = b1 ^ ((b1 >> 31) & 0x7fffffff);
          const int kk = max(k0, k1);
          acc |= (k1 > k0) << j;
          const int m = __reduce_max_sync(0xffffffffu, kk);
          const unsigned who = __ballot_sync(0xffffffffu, kk == m);
          if (lane == j) { keep_m = m; keep_w = who; }
        } else if (kMode == 6) {      // independent ballots
          v[j] = __ballot_sync(0xffffffffu, v[j] & 1) + it;
        }
      }
    }
  }
  const long long t1 = clock64();
  __syncthreads();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
  int s = acc + keep_m + keep_w;
#pragma unroll
  for (int j = 0; j < 16; ++j) s += v[j];
  if (s == 0x12345678) sink[0] = s;
}

template <int kMode>
void run(const char* name) {
  long long* cyc;
  int* sink;
  cudaMalloc(&cyc, 148 * sizeof(long long));
  cudaMalloc(&sink, 4);
  const int iters = 2000;
  for (int w = 1; w <= 4; ++w) {
    bench<kMode><<<148, 512>>>(iters, w, cyc, sink);
    cudaDeviceSynchronize();
    long long h[148];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0;
    for (int i = 0; i < 148; ++i) avg += h[i];
    avg /= 148;
    // per SMSP: w warps each issue iters*16 column steps
    printf("%-34s warps/SMSP %d: %.2f clk per column step per SMSP\n", name, w, avg / (double(iters) * 16 * w));
  }
  cudaFree(cyc);
  cudaFree(sink);
}

int main() {
  run<0>("redux.max dependent chain");
  run<3>("redux.max independent");
  run<1>("ballot dependent chain");
  run<6>("ballot independent");
  run<2>("shfl.xor dependent chain");
  run<4>("column step, 1 row/thread");
  run<5>("column step, 2 rows/thread");
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
