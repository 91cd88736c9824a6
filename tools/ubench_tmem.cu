// Microbenchmarks that size the sweep epilogue (development aid; not part of the product):
//   1. tcgen05.ld throughput per SM for x16/x32/x64 shapes with 1..4 warps per SM sub-partition
//   2. VIMNMX3 (3-input integer max) throughput, alone and interleaved with tcgen05.ld
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench_tmem tools/ubench_tmem.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../3d_reconstruction_b200/csrc/ptx_sm100.cuh"

using namespace b200m;

__device__ __forceinline__ void tmem_ld_32x64(uint32_t taddr, int (&v)[64]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x64.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, "
      "%32, %33, %34, %35, %36, %37, %38, %39, %40, %41, %42, %43, %44, %45, %46, %47, "
      "%48, %49, %50, %51, %52, %53, %54, %55, %56, %57, %58, %59, %60, %61, %62, %63}, [%64];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
        "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
        "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]),
        "=r"(v[32]), "=r"(v[33]), "=r"(v[34]), "=r"(v[35]), "=r"(v[36]), "=r"(v[37]), "=r"(v[38]), "=r"(v[39]),
        "=r"(v[40]), "=r"(v[41]), "=r"(v[42]), "=r"(v[43]), "=r"(v[44]), "=r"(v[45]), "=r"(v[46]), "=r"(v[47]),
        "=r"(v[48]), "=r"(v[49]), "=r"(v[50]), "=r"(v[51]), "=r"(v[52]), "=r"(v[53]), "=r"(v[54]), "=r"(v[55]),
        "=r"(v[56]), "=r"(v[57]), "=r"(v[58]), "=r"(v[59]), "=r"(v[60]), "=r"(v[61]), "=r"(v[62]), "=r"(v[63])
      : "r"(taddr)
      : "memory");
}

template <int N>
__device__ __forceinline__ int maxtree(const int (&v)[N], int m) {
  // running 3-input max: two new values per instruction, two independent chains
  int m0 = m, m1 = m;
#pragma unroll
  for (int i = 0; i < N; i += 4) {
    m0 = __vimax3_s32(m0, v[i], v[i + 1]);
    m1 = __vimax3_s32(m1, v[i + 2], v[i + 3]);
  }
  return max(m0, m1);
}

// mode 0: x16 loads only, 1: x32, 2: x64, 3: x32 + max tree, 4: max tree only (on registers), 5: x64 + max tree
template <int kMode>
__global__ void __launch_bounds__(512, 1) bench(int iters, int warps_per_smsp, long long* cycles, int* sink) {
  __shared__ uint32_t tmem_base_s;
  const uint32_t warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<512>(&tmem_base_s);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tmem_base_s;
  const uint32_t quad = warp & 3;
  const bool on = (warp >> 2) < (uint32_t)warps_per_smsp;
  int acc = threadIdx.x;
  __syncthreads();
  const long long t0 = clock64();
  if (on) {
    const uint32_t taddr = tb + ((quad * 32u) << 16) + (warp >> 2) * 128;
    if (kMode == 0) {
      int v[16];
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < 8; ++c) { tmem_ld_32x16(taddr + c * 16, v); }
        tmem_ld_fence16(v);
        acc ^= v[0];
      }
    } else if (kMode == 1) {
      int v[32];
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < 4; ++c) { tmem_ld_32x32(taddr + c * 32, v); }
        tmem_ld_fence(v);
        acc ^= v[0];
      }
    } else if (kMode == 2) {
      int v[64];
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < 2; ++c) { tmem_ld_32x64(taddr + c * 64, v); }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc ^= v[0];
      }
    } else if (kMode == 3) {
      int va[32], vb[32];
      tmem_ld_32x32(taddr, va);
      for (int i = 0; i < iters; ++i) {
        tmem_ld_fence(va);
        tmem_ld_32x32(taddr + 32, vb);
        acc = maxtree<32>(va, acc);
        tmem_ld_fence(vb);
        tmem_ld_32x32(taddr + 64, va);
        acc = maxtree<32>(vb, acc);
        tmem_ld_fence(va);
        tmem_ld_32x32(taddr + 96, vb);
        acc = maxtree<32>(va, acc);
        tmem_ld_fence(vb);
        tmem_ld_32x32(taddr, va);
        acc = maxtree<32>(vb, acc);
      }
      tmem_ld_fence(va);
      acc ^= va[0];
    } else if (kMode == 4) {
      int v[32];
#pragma unroll
      for (int j = 0; j < 32; ++j) v[j] = acc * (j + 3);
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          acc = maxtree<32>(v, acc);
          asm volatile("" : "+r"(v[0]), "+r"(v[5]), "+r"(v[17]), "+r"(acc));
        }
      }
    } else if (kMode == 5) {
      int va[64];
      for (int i = 0; i < iters; ++i) {
        tmem_ld_32x64(taddr, va);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc = maxtree<64>(va, acc);
        tmem_ld_32x64(taddr + 64, va);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        acc = maxtree<64>(va, acc);
      }
    }
  }
  const long long t1 = clock64();
  if (threadIdx.x % 32 == 0) cycles[blockIdx.x * 16 + warp] = on ? (t1 - t0) : 0;
  sink[blockIdx.x * 512 + threadIdx.x] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) { tc_fence_after(); tmem_dealloc<512>(tb); }
}

template <int kMode>
void run(const char* name, int wps) {
  const int iters = 2000, grid = 148;
  long long* cyc; int* sink;
  cudaMalloc(&cyc, grid * 16 * 8);
  cudaMalloc(&sink, grid * 512 * 4);
  bench<kMode><<<grid, 512>>>(iters, wps, cyc, sink);
  bench<kMode><<<grid, 512>>>(iters, wps, cyc, sink);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("%s: %s\n", name, cudaGetErrorString(e)); return; }
  long long h[16];
  cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0;
  for (int i = 0; i < 16; ++i) mx = h[i] > mx ? h[i] : mx;
  // every active warp handles 128 columns x 32 lanes x 4 B per iteration
  const double bytes = 4.0 * wps * 128.0 * 32 * 4 * iters;
  printf("%-28s warps/SMSP=%d cycles=%lld  values/clk/SM=%.1f  B/clk/SM=%.1f\n", name, wps, mx, bytes / 4 / mx, bytes / mx);
  cudaFree(cyc); cudaFree(sink);
}

int main() {
  for (int wps = 1; wps <= 4; wps *= 2) {
    run<0>("ld x16", wps);
    run<1>("ld x32", wps);
    run<2>("ld x64", wps);
    run<3>("ld x32 + vimax3 tree", wps);
    run<5>("ld x64 + vimax3 tree (serial)", wps);
    run<4>("vimax3 tree only", wps);
  }
  return 0;
}
