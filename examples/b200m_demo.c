/* The C-ABI of include/b200m.h used from plain C, no Python and no torch: what a maintainer's own host code (or a
 * cgo / JNI / ctypes stub) does around one image pair -- the per-pair unit of the reference at matching.py:107-128.
 *
 *   b200m_demo <in.bin> <out.bin> [ratio] [mutual]
 *   in.bin : int32 kq, int32 kt, uint8 q[kq][128], uint8 t[kt][128]
 *   out.bin: int32 count, then count x (int32 query_idx, int32 train_idx, float32 dist), then the knnMatch(k=2) table:
 *            kq x (int32 idx0, int32 idx1, float32 d0, float32 d1)
 * Build (tests/test_gpu_c_abi.py does exactly this):
 *   gcc -O2 -std=c11 examples/b200m_demo.c -Iinclude -I/usr/local/cuda/include -L3d_reconstruction_b200/lib -lb200m \
 *       -L/usr/local/cuda/lib64 -lcudart -Wl,-rpath,$PWD/3d_reconstruction_b200/lib -o b200m_demo
 */
#include <cuda_runtime_api.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "b200m.h"

#define CK(call)                                                                                     \
  do {                                                                                               \
    cudaError_t e_ = (call);                                                                         \
    if (e_ != cudaSuccess) { fprintf(stderr, "CUDA: %s (%s:%d)\n", cudaGetErrorString(e_), __FILE__, __LINE__); return 2; } \
  } while (0)
#define B200(call)                                                                                   \
  do {                                                                                               \
    int rc_ = (call);                                                                                \
    if (rc_ != B200M_OK) { fprintf(stderr, "b200m: %d %s (%s:%d)\n", rc_, b200m_last_error_string(), __FILE__, __LINE__); return 3; } \
  } while (0)

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: %s in.bin out.bin [ratio] [mutual]\n", argv[0]); return 1; }
  const double ratio = argc > 3 ? atof(argv[3]) : 0.75;
  const int mutual = argc > 4 ? atoi(argv[4]) : 1;
  FILE* f = fopen(argv[1], "rb");
  if (!f) { perror(argv[1]); return 1; }
  int32_t k[2];
  if (fread(k, sizeof(int32_t), 2, f) != 2 || k[0] < 0 || k[1] < 0) { fprintf(stderr, "bad header\n"); return 1; }
  const int64_t kmax = k[0] > k[1] ? k[0] : k[1];
  const int64_t k_stride = ((kmax > 0 ? kmax : 1) + B200M_TILE_T - 1) / B200M_TILE_T * B200M_TILE_T;
  const int32_t n_img = 2;
  const size_t bank_bytes = (size_t)n_img * k_stride * 128;
  uint8_t* h_bank = (uint8_t*)calloc(bank_bytes, 1);
  for (int i = 0; i < 2; ++i)
    if (k[i] && fread(h_bank + (size_t)i * k_stride * 128, 128, (size_t)k[i], f) != (size_t)k[i]) { fprintf(stderr, "short file\n"); return 1; }
  fclose(f);

  B200(b200m_device_check());
  cudaStream_t st;
  CK(cudaStreamCreate(&st));
  uint8_t *bank, *sbank;
  int32_t *kpi, *norms, *stab, *pairs, *oq, *ot, *ocnt, *kidx;
  float *od, *kdist;
  void* ws;
  const int64_t cap = k_stride;
  const size_t ws_bytes = b200m_workspace_bytes(k_stride, 1);
  CK(cudaMalloc((void**)&bank, bank_bytes));
  CK(cudaMalloc((void**)&sbank, bank_bytes));
  CK(cudaMalloc((void**)&kpi, sizeof(int32_t) * n_img));
  CK(cudaMalloc((void**)&norms, sizeof(int32_t) * (size_t)b200m_norm_table_elems(n_img, k_stride)));
  CK(cudaMalloc((void**)&stab, sizeof(int32_t) * (size_t)b200m_sort_table_elems(n_img, k_stride)));
  CK(cudaMalloc((void**)&pairs, sizeof(int32_t) * 2));
  CK(cudaMalloc((void**)&oq, sizeof(int32_t) * (size_t)cap));
  CK(cudaMalloc((void**)&ot, sizeof(int32_t) * (size_t)cap));
  CK(cudaMalloc((void**)&od, sizeof(float) * (size_t)cap));
  CK(cudaMalloc((void**)&ocnt, sizeof(int32_t)));
  CK(cudaMalloc((void**)&kidx, sizeof(int32_t) * 2 * (size_t)k_stride));
  CK(cudaMalloc((void**)&kdist, sizeof(float) * 2 * (size_t)k_stride));
  CK(cudaMalloc(&ws, ws_bytes));
  const int32_t h_pairs[2] = {0, 1};
  CK(cudaMemcpyAsync(bank, h_bank, bank_bytes, cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(kpi, k, sizeof(k), cudaMemcpyHostToDevice, st));
  CK(cudaMemcpyAsync(pairs, h_pairs, sizeof(h_pairs), cudaMemcpyHostToDevice, st));

  /* one-off per bank: norms + padding, norm-sorted copy */
  B200(b200m_bank_prepare_u8(bank, n_img, k_stride, kpi, norms, st));
  B200(b200m_bank_sort_u8(bank, norms, kpi, n_img, k_stride, sbank, stab, st));
  /* the rule: 2-NN + Lowe ratio + cross-check, and the plain knnMatch(k=2) table */
  B200(b200m_match_pairs_u8(bank, norms, sbank, stab, kpi, n_img, k_stride, pairs, 1, ratio, mutual, oq, ot, od, ocnt, cap, ws,
                            ws_bytes, st));
  B200(b200m_knn2_sorted_u8(bank, norms, sbank, stab, kpi, n_img, k_stride, pairs, 1, kidx, kdist, ws, ws_bytes, st));

  int32_t count = 0;
  CK(cudaMemcpyAsync(&count, ocnt, sizeof(count), cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  if (count < 0 || count > cap) { fprintf(stderr, "pair overflowed or invalid: count %d cap %lld\n", count, (long long)cap); return 4; }
  int32_t* hq = (int32_t*)malloc(sizeof(int32_t) * (size_t)(count + 1));
  int32_t* ht = (int32_t*)malloc(sizeof(int32_t) * (size_t)(count + 1));
  float* hd = (float*)malloc(sizeof(float) * (size_t)(count + 1));
  int32_t* hki = (int32_t*)malloc(sizeof(int32_t) * 2 * (size_t)(k[0] + 1));
  float* hkd = (float*)malloc(sizeof(float) * 2 * (size_t)(k[0] + 1));
  CK(cudaMemcpy(hq, oq, sizeof(int32_t) * (size_t)count, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(ht, ot, sizeof(int32_t) * (size_t)count, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hd, od, sizeof(float) * (size_t)count, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hki, kidx, sizeof(int32_t) * 2 * (size_t)k[0], cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(hkd, kdist, sizeof(float) * 2 * (size_t)k[0], cudaMemcpyDeviceToHost));

  FILE* o = fopen(argv[2], "wb");
  if (!o) { perror(argv[2]); return 1; }
  fwrite(&count, sizeof(count), 1, o);
  for (int32_t i = 0; i < count; ++i) {
    fwrite(&hq[i], 4, 1, o);
    fwrite(&ht[i], 4, 1, o);
    fwrite(&hd[i], 4, 1, o);
  }
  for (int32_t r = 0; r < k[0]; ++r) {
    fwrite(&hki[2 * r], 4, 2, o);
    fwrite(&hkd[2 * r], 4, 2, o);
  }
  fclose(o);
  printf("b200m %d on %d SMs: %d x %d descriptors -> %d matches (ratio %.2f, mutual %d)\n", b200m_version(), b200m_sm_count(),
         k[0], k[1], count, ratio, mutual);
  cudaFree(bank); cudaFree(sbank); cudaFree(kpi); cudaFree(norms); cudaFree(stab); cudaFree(pairs); cudaFree(oq); cudaFree(ot);
  cudaFree(od); cudaFree(ocnt); cudaFree(kidx); cudaFree(kdist); cudaFree(ws);
  cudaStreamDestroy(st);
  free(h_bank); free(hq); free(ht); free(hd); free(hki); free(hkd);
  return 0;
}
