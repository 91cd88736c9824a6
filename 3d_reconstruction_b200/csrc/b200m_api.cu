// C-ABI of the B200 matcher (see include/b200m.h). Host side: argument checks, TMA descriptor
// encoding, batching of pairs through a bounded workspace, kernel launches. No CPU fallback.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <mutex>
#include <set>
#include <utility>
#include <vector>

#include "../../include/b200m.h"
#include "match_finalize.cuh"
#include "match_sweep.cuh"
#include "match_chunk.cuh"
#ifdef B200M_EXPERIMENTAL   // cta_group::2 variants, measured 19-39 % slower: not in the shipped library
#include "experimental/match_sweep2.cuh"
#include "experimental/match_chunk2.cuh"
#endif
#include "match_refine.cuh"
#include "match_knn.cuh"
#include "match_sort.cuh"
#include "match_float.cuh"
#include "match_float2.cuh"
#include "match_vq.cuh"
#include "match_ransac.cuh"

namespace {

thread_local char g_err[512] = "";

// Process-global tuning knobs; every field is an atomic so that a set_option from one host thread and a
// launch from another never race. Entry points do not read these directly: see OptScope / opt() below.
struct Options {
  std::atomic<int64_t> epilogue{1};
  std::atomic<int64_t> grid{0};
  std::atomic<int64_t> batch{0};
  // chunk sweep (pipeline 2): 5 = N256 MMAs / two accumulators (default), otherwise N128 MMAs / four accumulators.
  // (-DB200M_EXPERIMENTAL builds add the slower cta_group::2 variants: 6 = CTA-pair chunk sweep, 4 = CTA-pair exact sweeps.)
  std::atomic<int64_t> kernel{5};
  // 2 = chunk sweep on the norm-sorted bank + refine + exact fallback (default, needs the sorted bank),
  // 1 = top-1 sweep + ratio/mutual verify sweeps, 0 = top-2 sweep + mutual verify
  std::atomic<int64_t> pipeline{2};
  std::atomic<int64_t> chunk{64};     // columns per chunk of the chunk sweep (64 needs kernel 5, else 32 is used)
  // float path: 0 = one sweep per pair with the column maximum fused into it (match_float2.cuh, default),
  // 1 = one sweep per direction (match_float.cuh; resolves exact ties between different rows like the reference)
  std::atomic<int64_t> float_sweep{0};
  // knnMatch chunk pipeline, exact evaluation of the two winning chunks: 0 = tensor cores over gathered rows
  // (knn_eval_tc_kernel, default), 1 = dp4a (knn_eval_kernel)
  std::atomic<int64_t> knn_eval{0};
  // refinement of the chunk bounds (pipeline 2): 0 = one fused kernel (default: cheapest when few rows need their chunk
  // evaluated, as on the headline workload), 1 = streaming classification + one warp per row to evaluate (banks where
  // many rows need it: 1.8x faster refinement on uniform random bytes, 23 us more per batch otherwise)
  std::atomic<int64_t> refine{0};
} g_opt;

// One consistent view of the knobs per entry-point call. A public entry point opens an OptScope first thing; the
// outermost scope on this host thread copies every knob ONCE, and everything below it -- the workspace size check, the
// batch size the workspace is carved for, the per-batch kernel choices, nested public calls such as
// b200m_workspace_bytes -- reads that copy through opt(). A b200m_set_option from another thread therefore takes effect
// at the next call, never in the middle of one (where a larger "batch" between the size check and the carve would make
// the kernels write past the checked workspace).
struct OptSnap {
  int64_t epilogue, grid, batch, kernel, pipeline, chunk, float_sweep, knn_eval, refine;
};
thread_local OptSnap t_opt;
thread_local int t_opt_depth = 0;

void snapshot_options() {
  t_opt.epilogue = g_opt.epilogue.load();
  t_opt.grid = g_opt.grid.load();
  t_opt.batch = g_opt.batch.load();
  t_opt.kernel = g_opt.kernel.load();
  t_opt.pipeline = g_opt.pipeline.load();
  t_opt.chunk = g_opt.chunk.load();
  t_opt.float_sweep = g_opt.float_sweep.load();
  t_opt.knn_eval = g_opt.knn_eval.load();
  t_opt.refine = g_opt.refine.load();
}

struct OptScope {
  OptScope() { if (t_opt_depth++ == 0) snapshot_options(); }
  ~OptScope() { --t_opt_depth; }
  OptScope(const OptScope&) = delete;
  OptScope& operator=(const OptScope&) = delete;
};

// the knobs as the current entry-point call sees them (outside any scope: as they are right now)
const OptSnap& opt() {
  if (t_opt_depth == 0) snapshot_options();
  return t_opt;
}

// Optional per-launch timing of the sweeps with CUDA events on the launching stream ("profile" option).
// The event pool and the lists are guarded by `mu` (two host threads may launch concurrently); the pool is
// capped: once kMaxEvents are in flight without a b200m_profile_read, further launches are counted but not timed.
struct Profile {
  std::atomic<bool> on{false};
  std::atomic<int64_t> launches{0};                        // every kernel this library launched
  std::mutex mu;
  std::vector<cudaEvent_t> pool;
  std::vector<std::pair<int, int>> first, verify;          // (start, stop) indices into pool
  std::vector<std::pair<int, int>> fallback, refine, select, emit;
  size_t used = 0;
  int64_t dropped = 0;                                     // launches not timed because the pool was full
  static constexpr size_t kMaxEvents = size_t(1) << 16;
  // a (start, stop) pair of events, or false when the pool is exhausted; call with mu held
  bool take2(int* e0, int* e1) {
    if (used + 2 > kMaxEvents) { ++dropped; return false; }
    while (pool.size() < used + 2) {
      cudaEvent_t e;
      if (cudaEventCreate(&e) != cudaSuccess) { cudaGetLastError(); ++dropped; return false; }
      pool.push_back(e);
    }
    *e0 = static_cast<int>(used++);
    *e1 = static_cast<int>(used++);
    return true;
  }
} g_prof;

// brackets one launch with events when profiling is on
struct ProfScope {
  std::vector<std::pair<int, int>>* list;
  cudaStream_t st;
  int e0 = -1, e1 = -1;
  ProfScope(std::vector<std::pair<int, int>>* l, cudaStream_t s) : list(l), st(s) {
    if (g_prof.on.load(std::memory_order_relaxed) && list) {
      std::lock_guard<std::mutex> lk(g_prof.mu);
      if (g_prof.take2(&e0, &e1)) cudaEventRecord(g_prof.pool[e0], st);
    }
  }
  ~ProfScope() {
    g_prof.launches.fetch_add(1, std::memory_order_relaxed);
    if (e0 >= 0) {
      std::lock_guard<std::mutex> lk(g_prof.mu);
      cudaEventRecord(g_prof.pool[e1], st);
      list->push_back({e0, e1});
    }
  }
};

// cudaFuncAttributeMaxDynamicSharedMemorySize is per device (and context): remember which (kernel, device)
// pairs have been opted in, so that a process driving several GPUs from one thread works on every one of them.
template <typename F>
cudaError_t ensure_smem_attr(F* func, size_t bytes) {
  static std::mutex mu;
  static std::set<std::pair<const void*, int>> done;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  const std::pair<const void*, int> key(reinterpret_cast<const void*>(func), dev);
  std::lock_guard<std::mutex> lk(mu);
  if (done.count(key)) return cudaSuccess;
  e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(bytes));
  if (e == cudaSuccess) done.insert(key);
  return e;
}

int fail(int code, const char* fmt, const char* detail = "") {
  snprintf(g_err, sizeof(g_err), fmt, detail);
  return code;
}

#define CUDA_TRY(expr)                                                            \
  do {                                                                            \
    cudaError_t e_ = (expr);                                                      \
    if (e_ != cudaSuccess) {                                                      \
      snprintf(g_err, sizeof(g_err), "%s failed: %s", #expr, cudaGetErrorString(e_)); \
      return B200M_E_CUDA;                                                        \
    }                                                                             \
  } while (0)

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* p = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(p);
  return fn;
}

// [n_rows, 128] u8 row-major, box = 128 rows x 128 B, 128-byte swizzle (matches make_desc_k128).
int make_row_tmap(CUtensorMap* tm, const void* base, uint64_t n_rows) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail(B200M_E_NO_DEVICE, "cuTensorMapEncodeTiled entry point unavailable%s");
  cuuint64_t dims[2] = {128, n_rows};
  cuuint64_t strides[1] = {128};
  cuuint32_t box[2] = {128, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_err, sizeof(g_err), "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
    return B200M_E_CUDA;
  }
  return B200M_OK;
}

// [n_rows, dim] bf16 row-major, box = 128 rows x 64 elements (one 128-byte swizzle atom along K)
int make_row_tmap_bf16(CUtensorMap* tm, const void* base, uint64_t n_rows, uint32_t dim) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail(B200M_E_NO_DEVICE, "cuTensorMapEncodeTiled entry point unavailable%s");
  cuuint64_t dims[2] = {dim, n_rows};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(dim) * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    snprintf(g_err, sizeof(g_err), "cuTensorMapEncodeTiled (bf16) failed with CUresult %d", static_cast<int>(r));
    return B200M_E_CUDA;
  }
  return B200M_OK;
}

struct DeviceInfo {
  int ok = 0;   // 0 unknown, 1 good, -1 bad
  int sms = 0;
};

DeviceInfo& device_info() {
  static thread_local DeviceInfo info;
  static thread_local int cached_dev = -1;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    info.ok = -1;
    cudaGetLastError();
    return info;
  }
  if (dev == cached_dev && info.ok != 0) return info;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
    info.ok = -1;
    cudaGetLastError();
    return info;
  }
  cached_dev = dev;
  info.ok = (prop.major == 10) ? 1 : -1;
  info.sms = prop.multiProcessorCount;
  return info;
}

constexpr size_t kSweepSmemBytes = sizeof(b200m::SweepSmem) + 1024;

template <int kKind>
int launch_sweep(const CUtensorMap& tmap_a, const CUtensorMap& tmap_a2, const CUtensorMap& tmap_b,
                 const b200m::SweepParams& sp, int grid, cudaStream_t st) {
  CUDA_TRY(ensure_smem_attr(b200m::sweep_kernel<kKind>, kSweepSmemBytes));
  {
    ProfScope ps(kKind == b200m::kVerify ? &g_prof.verify : b200m::is_gathered(kKind) ? &g_prof.fallback : &g_prof.first, st);
    b200m::sweep_kernel<kKind><<<grid, b200m::kThreads, kSweepSmemBytes, st>>>(tmap_a, tmap_a2, tmap_b, sp);
  }
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}

constexpr size_t kChunkSmemBytes = sizeof(b200m::ChunkSmem) + 1024;
constexpr size_t kChunk256SmemBytes = sizeof(b200m::Chunk256Smem) + 1024;

int launch_chunk_sweep(const CUtensorMap& tmap, const b200m::ChunkParams& cp, int grid, int64_t kernel, cudaStream_t st,
                       bool top3 = false) {
  if (top3) {      // knnMatch form: three bounds + two chunks per row (64-column chunks, N = 256 MMAs)
    CUDA_TRY((ensure_smem_attr(b200m::chunk_sweep256_kernel<false, 64, true>, kChunk256SmemBytes)));
    {
      ProfScope ps(&g_prof.first, st);
      b200m::chunk_sweep256_kernel<false, 64, true><<<grid, b200m::kThreads256, kChunk256SmemBytes, st>>>(tmap, cp);
    }
    CUDA_TRY(cudaGetLastError());
    return B200M_OK;
  }
  CUDA_TRY(ensure_smem_attr(b200m::chunk_sweep_kernel, kChunkSmemBytes));
  CUDA_TRY(ensure_smem_attr(b200m::chunk_sweep256_kernel<false, 32>, kChunk256SmemBytes));
  CUDA_TRY(ensure_smem_attr(b200m::chunk_sweep256_kernel<false, 64>, kChunk256SmemBytes));
  CUDA_TRY(ensure_smem_attr(b200m::chunk_sweep256_kernel<true, 32>, kChunk256SmemBytes));
#ifdef B200M_EXPERIMENTAL
  constexpr size_t kChunkPairSmemBytes = sizeof(b200m::pair::ChunkPairSmem) + 1024;
  CUDA_TRY(ensure_smem_attr(b200m::pair::chunk_sweep_pair_kernel<false>, kChunkPairSmemBytes));
  CUDA_TRY(ensure_smem_attr(b200m::pair::chunk_sweep_pair_kernel<true>, kChunkPairSmemBytes));
#endif
  {
    ProfScope ps(&g_prof.first, st);
#ifdef B200M_EXPERIMENTAL
    if (kernel == 6) {
      // whole CTA pairs (the kernel carries __cluster_dims__(2, 1, 1)); work items are blocks of 512 rows
      const int64_t items = static_cast<int64_t>(cp.n_pairs) * ((cp.k_stride + 511) / 512);
      const int pgrid = 2 * static_cast<int>(std::min<int64_t>(std::max<int64_t>(items, 1), std::max(grid / 2, 1)));
      if (cp.peak)
        b200m::pair::chunk_sweep_pair_kernel<true><<<pgrid, b200m::kThreads, kChunkPairSmemBytes, st>>>(tmap, cp);
      else
        b200m::pair::chunk_sweep_pair_kernel<false><<<pgrid, b200m::kThreads, kChunkPairSmemBytes, st>>>(tmap, cp);
    } else
#endif
    if (kernel == 5 && cp.peak)
      b200m::chunk_sweep256_kernel<true, 32><<<grid, b200m::kThreads256, kChunk256SmemBytes, st>>>(tmap, cp);
    else if (kernel == 5 && cp.chunk_cols == 64)
      b200m::chunk_sweep256_kernel<false, 64><<<grid, b200m::kThreads256, kChunk256SmemBytes, st>>>(tmap, cp);
    else if (kernel == 5)
      b200m::chunk_sweep256_kernel<false, 32><<<grid, b200m::kThreads256, kChunk256SmemBytes, st>>>(tmap, cp);
    else
      b200m::chunk_sweep_kernel<<<grid, b200m::kThreads, kChunkSmemBytes, st>>>(tmap, cp);
  }
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}

#ifdef B200M_EXPERIMENTAL
constexpr size_t kPairSmemBytes = sizeof(b200m::pair::PairSmem) + 1024;

template <int kKind>
int launch_sweep_pair(const CUtensorMap& tmap_a, const CUtensorMap& tmap_a2, const CUtensorMap& tmap_b,
                      const b200m::SweepParams& sp, int grid, cudaStream_t st) {
  CUDA_TRY(ensure_smem_attr(b200m::pair::sweep_pair_kernel<kKind>, kPairSmemBytes));
  grid = (grid + 1) & ~1;   // whole CTA pairs (the kernel carries __cluster_dims__(2, 1, 1))
  {
    ProfScope ps(kKind == b200m::kVerify ? &g_prof.verify : &g_prof.first, st);
    b200m::pair::sweep_pair_kernel<kKind><<<grid, b200m::pair::kThreads2, kPairSmemBytes, st>>>(tmap_a, tmap_a2, tmap_b, sp);
  }
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}
#endif

// dispatch on the "kernel" option (the CTA-pair variant exists in -DB200M_EXPERIMENTAL builds only)
template <int kKind>
int launch_any(const CUtensorMap& tmap_a, const CUtensorMap& tmap_a2, const CUtensorMap& tmap_b,
               const b200m::SweepParams& sp, int grid, cudaStream_t st) {
#ifdef B200M_EXPERIMENTAL
  if (opt().kernel == 4) return launch_sweep_pair<kKind>(tmap_a, tmap_a2, tmap_b, sp, grid, st);
#endif
  return launch_sweep<kKind>(tmap_a, tmap_a2, tmap_b, sp, grid, st);
}

constexpr int64_t kDefaultBatchRows = int64_t(1) << 22;   // pairs * k_stride per internal batch

// Largest internal batch a call over n_pairs pairs may use: what the workspace is sized and carved for.
int64_t batch_cap(int64_t k_stride, int64_t n_pairs) {
  const int64_t opt_batch = opt().batch;
  int64_t b = opt_batch > 0 ? opt_batch : (kDefaultBatchRows / (k_stride > 0 ? k_stride : 1));
  if (b < 1) b = 1;
  if (b > n_pairs) b = n_pairs;
  if (b < 1) b = 1;
  return b;
}

// Pairs per internal batch: equal batches under the cap, so 4,112 pairs run as 17 x 242 instead of 16 x 256 + a
// launch sequence carrying 16 pairs. (The workspace of a longer list is large enough for any shorter one.)
int64_t batch_pairs(int64_t k_stride, int64_t n_pairs) {
  int64_t b = batch_cap(k_stride, n_pairs);
  const int64_t n_batches = (n_pairs + b - 1) / b;
  if (n_batches > 1) b = (n_pairs + n_batches - 1) / n_batches;
  return b;
}

// Workspace carved per batch of B pairs (R = B * k_stride rows):
struct Workspace {
  unsigned long long* stats;  // kStatSlots x kStatFields counters of the call (first block of the workspace)
  int4* rowres;             // R
  int32_t* cand_q;          // R
  int32_t* cand_t;          // R
  float* cand_d;            // R
  int32_t* cand_thr_ratio;  // R
  int32_t* cand_thr_mut;    // R
  int32_t* viol_ratio;      // R
  int32_t* viol_mut;        // R
  int32_t* cand_count;      // B (padded to 64)
  uint8_t* gath_q;          // R * 128
  uint8_t* gath_t;          // R * 128
  // pipeline 2
  int4* rowres_o;           // R   refined rows, original query order
  int32_t* fb_q;            // R   rows of the fallback sweep
  int32_t* cand_rslot;      // R   slot of a candidate in the open-ratio list, -1 = settled
  int32_t* cand_excl;       // R   open-ratio list: the nearest train row
  int32_t* fb_count;        // B (padded to 64)
  int32_t* cand_count_a;    // B (padded to 64)
  int32_t* vitems;          // 2 * (2 * B * k_stride / 256): compact work list of the verify sweep
};

size_t carve(Workspace* w, void* base, int64_t B, int64_t k_stride) {
  const int64_t R = B * k_stride;
  uintptr_t p0 = (reinterpret_cast<uintptr_t>(base) + 255) & ~uintptr_t(255);
  uintptr_t p = p0;
  auto take = [&](size_t bytes) { uintptr_t r = p; p += (bytes + 255) & ~size_t(255); return r; };
  uintptr_t s0 = take(sizeof(unsigned long long) * b200m::kStatSlots * b200m::kStatFields);
  uintptr_t a = take(sizeof(int4) * R);
  uintptr_t b = take(4 * R), c = take(4 * R), d = take(4 * R), e = take(4 * R), f = take(4 * R);
  uintptr_t f2 = take(4 * R), f3 = take(4 * R);
  uintptr_t g = take(4 * (B + 64));
  uintptr_t h = take(static_cast<size_t>(128) * R);
  uintptr_t h2 = take(static_cast<size_t>(128) * R);
  uintptr_t i1 = take(sizeof(int4) * R);
  uintptr_t i2 = take(4 * R), i3 = take(4 * R), i4 = take(4 * R);
  uintptr_t i5 = take(4 * (B + 64)), i6 = take(4 * (B + 64));
  uintptr_t i7 = take(sizeof(int32_t) * 2 * (2 * R / 256 + 64));
  if (w) {
    w->stats = reinterpret_cast<unsigned long long*>(s0);
    w->vitems = reinterpret_cast<int32_t*>(i7);
    w->rowres_o = reinterpret_cast<int4*>(i1);
    w->fb_q = reinterpret_cast<int32_t*>(i2);
    w->cand_rslot = reinterpret_cast<int32_t*>(i3);
    w->cand_excl = reinterpret_cast<int32_t*>(i4);
    w->fb_count = reinterpret_cast<int32_t*>(i5);
    w->cand_count_a = reinterpret_cast<int32_t*>(i6);
    w->rowres = reinterpret_cast<int4*>(a);
    w->cand_q = reinterpret_cast<int32_t*>(b);
    w->cand_t = reinterpret_cast<int32_t*>(c);
    w->cand_d = reinterpret_cast<float*>(d);
    w->cand_thr_ratio = reinterpret_cast<int32_t*>(e);
    w->cand_thr_mut = reinterpret_cast<int32_t*>(f);
    w->viol_ratio = reinterpret_cast<int32_t*>(f2);
    w->viol_mut = reinterpret_cast<int32_t*>(f3);
    w->cand_count = reinterpret_cast<int32_t*>(g);
    w->gath_q = reinterpret_cast<uint8_t*>(h);
    w->gath_t = reinterpret_cast<uint8_t*>(h2);
  }
  return (p - p0) + 256;
}

int check_common(const void* bank, const void* sqnorm, const void* kpi, int32_t n_img, int64_t k_stride,
                 const void* pairs, int64_t n_pairs) {
  if (!bank || !sqnorm || !kpi || n_img <= 0) return fail(B200M_E_BAD_ARG, "null bank/norm_table/k_per_image or n_img <= 0%s");
  if (k_stride <= 0 || (k_stride % B200M_TILE_T) != 0)
    return fail(B200M_E_BAD_ARG, "k_stride must be a positive multiple of 256%s");
  if (n_pairs < 0 || (n_pairs > 0 && !pairs)) return fail(B200M_E_BAD_ARG, "bad pairs / n_pairs%s");
  if (static_cast<int64_t>(n_img) * k_stride >= (int64_t(1) << 31))
    return fail(B200M_E_BAD_ARG, "bank has more than 2^31 rows%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  return B200M_OK;
}

// sort table (int32): perm[R] | inv[R] | snorm[R] | cmin[R/32] | spread[n4] | spread64[n4], R = n_img * k_stride, n4 = n_img rounded up to 4
struct SortTable {
  const int32_t *perm, *inv, *snorm, *cmin, *spread, *spread64;
};

int64_t sort_table_elems(int64_t n_img, int64_t k_stride) {
  const int64_t R = n_img * k_stride;
  return 3 * R + R / 32 + 2 * ((n_img + 3) & ~int64_t(3));
}

SortTable sort_table_view(const int32_t* t, int64_t n_img, int64_t k_stride) {
  const int64_t R = n_img * k_stride;
  SortTable v;
  v.perm = t;
  v.inv = t + R;
  v.snorm = t + 2 * R;
  v.cmin = t + 3 * R;
  v.spread = t + 3 * R + R / 32;
  v.spread64 = v.spread + ((n_img + 3) & ~int64_t(3));
  return v;
}

int grid_for(int64_t work, int block) {
  int64_t blocks = (work + block - 1) / block;
  const int64_t max_blocks = static_cast<int64_t>(device_info().sms) * 16;
  if (blocks > max_blocks) blocks = max_blocks;
  return blocks < 1 ? 1 : static_cast<int>(blocks);
}

int sweep_grid(int64_t items) {
  const int64_t opt_grid = opt().grid;
  int grid = opt_grid > 0 ? static_cast<int>(opt_grid) : device_info().sms;
  if (grid > items) grid = static_cast<int>(items);
  return grid < 1 ? 1 : grid;
}

b200m::SweepParams base_params(const int32_t* norm_table, int32_t n_img, const int32_t* kpi, int64_t k_stride,
                               const int32_t* pairs_batch, int64_t nb, const Workspace& w) {
  b200m::SweepParams sp;
  sp.sqnorm = norm_table;
  sp.nbmin32 = norm_table + static_cast<int64_t>(n_img) * k_stride;
  sp.k_per_image = kpi;
  sp.pairs = pairs_batch;
  sp.k_stride = k_stride;
  sp.n_pairs = static_cast<int32_t>(nb);
  sp.blocks_per_pair = static_cast<int32_t>(k_stride / b200m::kRows);
  sp.n_img = n_img;
  sp.rowres = w.rowres;
  sp.cand_count = w.cand_count;
  sp.cand_count_r = w.cand_count;
  sp.cand_q = w.cand_q;
  sp.cand_t = w.cand_t;
  sp.cand_excl = w.cand_t;
  sp.fb_count = w.fb_count;
  sp.fb_total = nullptr;
  sp.fb_q = w.fb_q;
  sp.rowres_o = w.rowres_o;
  sp.cand_thr_ratio = w.cand_thr_ratio;
  sp.cand_thr_mut = w.cand_thr_mut;
  sp.viol_ratio = w.viol_ratio;
  sp.viol_mut = w.viol_mut;
  sp.do_ratio = 0;
  sp.do_mutual = 0;
  sp.vitems = nullptr;
  sp.vitem_count = nullptr;
  return sp;
}

// First sweep over every query row of every pair of the batch -> w.rowres.
// exact2: the second neighbour must be exact (knnMatch output / top-2 pipeline).
int sweep_rows(const CUtensorMap& tmap_bank, const b200m::SweepParams& sp, bool exact2, cudaStream_t st) {
  const int grid = sweep_grid(static_cast<int64_t>(sp.n_pairs) * sp.blocks_per_pair);
  switch (opt().epilogue) {
    case 0: return launch_any<b200m::kTop2Exact>(tmap_bank, tmap_bank, tmap_bank, sp, grid, st);
    case 2: return launch_any<b200m::kPeak>(tmap_bank, tmap_bank, tmap_bank, sp, grid, st);
    default:
      return exact2 ? launch_any<b200m::kTop2>(tmap_bank, tmap_bank, tmap_bank, sp, grid, st)
                    : launch_any<b200m::kTop1>(tmap_bank, tmap_bank, tmap_bank, sp, grid, st);
  }
}

// out_count[p] = -1 for pair-list entries that name an image outside the bank (the kernels treat them as empty)
int mark_invalid_pairs(const int32_t* pairs, int64_t n_pairs, int32_t n_img, int32_t* out_count, cudaStream_t st) {
  b200m::mark_invalid_pairs_kernel<<<static_cast<unsigned>((n_pairs + 255) / 256), 256, 0, st>>>(pairs, n_pairs, n_img, out_count);
  CUDA_TRY(cudaGetLastError());
  g_prof.launches.fetch_add(1);
  return B200M_OK;
}

template <int kDim, int kMode = b200m::ip::kIpArgmax>
int launch_ip_sweep(const CUtensorMap& tmap, const b200m::ip::IpParams& ip, int grid, cudaStream_t st) {
  constexpr size_t smem = sizeof(b200m::ip::IpSmem<kDim>) + 1024;
  CUDA_TRY((ensure_smem_attr(b200m::ip::ip_sweep_kernel<kDim, kMode>, smem)));
  {
    ProfScope ps(kMode == b200m::ip::kIpLse ? &g_prof.verify : &g_prof.first, st);
    b200m::ip::ip_sweep_kernel<kDim, kMode><<<grid, b200m::kThreads, smem, st>>>(tmap, ip);
  }
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}

template <int kDim, int kMode>
int launch_ip_sweep2(const CUtensorMap& tmap, const b200m::ip::Ip2Params& ip, int grid, cudaStream_t st) {
  constexpr size_t smem = sizeof(b200m::ip::Ip2Smem<kDim>) + 1024;
  CUDA_TRY((ensure_smem_attr(b200m::ip::ip_sweep2_kernel<kDim, kMode>, smem)));
  {
    ProfScope ps(kMode == b200m::ip::kIpLse ? &g_prof.verify : &g_prof.first, st);
    b200m::ip::ip_sweep2_kernel<kDim, kMode><<<grid, b200m::kThreads, smem, st>>>(tmap, ip);
  }
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}

// One batch of the single-sweep float path: (row arg-max, column maxima) sweep + mutual / threshold / packing.
// ws_rows: float2 [nb * k_stride] row results, followed by int32 column keys and int32 claims of the same length.
int float_single_sweep_batch(const CUtensorMap& tmap, const int32_t* k_per_image, int32_t n_img, int64_t k_stride, int32_t dim,
                             const int32_t* pairs_batch, int64_t nb, int64_t batch_cap_rows, float2* ws_rows, const float* nbuf,
                             float threshold, int use_threshold, int mutual, int32_t* out_q, int32_t* out_t, float* out_s,
                             int32_t* out_count, int64_t cap, cudaStream_t st) {
  int32_t* colkey = reinterpret_cast<int32_t*>(ws_rows + batch_cap_rows);
  int32_t* claim = colkey + batch_cap_rows;
  CUDA_TRY(cudaMemsetAsync(colkey, 0x80, sizeof(int32_t) * nb * k_stride, st));
  CUDA_TRY(cudaMemsetAsync(claim, 0x7f, sizeof(int32_t) * nb * k_stride, st));
  b200m::ip::Ip2Params ip;
  ip.k_per_image = k_per_image;
  ip.pairs = pairs_batch;
  ip.k_stride = k_stride;
  ip.n_pairs = static_cast<int32_t>(nb);
  ip.blocks_per_image = static_cast<int32_t>(k_stride / b200m::ip::kRows2);
  ip.n_img = n_img;
  ip.rowbest = ws_rows;
  ip.colkey = colkey;
  ip.nb = nbuf;
  const int grid = sweep_grid(nb * ip.blocks_per_image);
  int rc;
  if (nbuf)
    rc = dim == 256 ? launch_ip_sweep2<256, b200m::ip::kIpBiasArgmax>(tmap, ip, grid, st)
                    : launch_ip_sweep2<128, b200m::ip::kIpBiasArgmax>(tmap, ip, grid, st);
  else
    rc = dim == 256 ? launch_ip_sweep2<256, b200m::ip::kIpArgmax>(tmap, ip, grid, st)
                    : launch_ip_sweep2<128, b200m::ip::kIpArgmax>(tmap, ip, grid, st);
  if (rc) return rc;
  {
    ProfScope ps(&g_prof.emit, st);
    b200m::ip::ip_finalize2_kernel<1024><<<static_cast<unsigned>(nb), 1024, 0, st>>>(
        ws_rows, colkey, claim, pairs_batch, k_per_image, n_img, k_stride, threshold, use_threshold, mutual, nbuf ? 1 : 0,
        out_q, out_t, out_s, out_count, cap);
  }
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}

}  // namespace

extern "C" {

int b200m_version(void) { return B200M_VERSION; }
const char* b200m_last_error_string(void) { return g_err; }

int b200m_device_check(void) {
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available%s");
  return B200M_OK;
}

int b200m_sm_count(void) {
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available%s");
  return device_info().sms;
}

int b200m_set_option(const char* name, int64_t value) {
  if (!name) return fail(B200M_E_BAD_ARG, "null option name%s");
  if (!strcmp(name, "epilogue")) { g_opt.epilogue = value; return B200M_OK; }
  if (!strcmp(name, "grid")) { g_opt.grid = value; return B200M_OK; }
  if (!strcmp(name, "batch")) { g_opt.batch = value; return B200M_OK; }
  if (!strcmp(name, "pipeline")) { g_opt.pipeline = value; return B200M_OK; }
  if (!strcmp(name, "kernel")) {
#ifndef B200M_EXPERIMENTAL
    if (value == 4 || value == 6)
      return fail(B200M_E_BAD_ARG, "kernel variants 4 / 6 (cta_group::2) exist in -DB200M_EXPERIMENTAL builds only%s");
#endif
    g_opt.kernel = value;
    return B200M_OK;
  }
  if (!strcmp(name, "chunk")) {
    if (value != 32 && value != 64) return fail(B200M_E_BAD_ARG, "chunk must be 32 or 64%s");
    g_opt.chunk = value;
    return B200M_OK;
  }
  if (!strcmp(name, "profile")) { g_prof.on = value != 0; return B200M_OK; }
  if (!strcmp(name, "float_sweep")) {
    if (value != 0 && value != 1) return fail(B200M_E_BAD_ARG, "float_sweep must be 0 (single sweep) or 1 (one sweep per direction)%s");
    g_opt.float_sweep = value;
    return B200M_OK;
  }
  if (!strcmp(name, "knn_eval")) {
    if (value != 0 && value != 1) return fail(B200M_E_BAD_ARG, "knn_eval must be 0 (tensor cores) or 1 (dp4a)%s");
    g_opt.knn_eval = value;
    return B200M_OK;
  }
  if (!strcmp(name, "refine")) {
    if (value != 0 && value != 1) return fail(B200M_E_BAD_ARG, "refine must be 0 (fused kernel) or 1 (classify + warp per row)%s");
    g_opt.refine = value;
    return B200M_OK;
  }
  return fail(B200M_E_BAD_ARG, "unknown option %s", name);
}

int64_t b200m_get_option(const char* name) {
  if (!name) return -1;
  if (!strcmp(name, "epilogue")) return g_opt.epilogue;
  if (!strcmp(name, "grid")) return g_opt.grid;
  if (!strcmp(name, "batch")) return g_opt.batch;
  if (!strcmp(name, "pipeline")) return g_opt.pipeline;
  if (!strcmp(name, "kernel")) return g_opt.kernel;
  if (!strcmp(name, "chunk")) return g_opt.chunk;
  if (!strcmp(name, "profile")) return g_prof.on ? 1 : 0;
  if (!strcmp(name, "float_sweep")) return g_opt.float_sweep;
  if (!strcmp(name, "knn_eval")) return g_opt.knn_eval;
  if (!strcmp(name, "refine")) return g_opt.refine;
  return -1;
}

int64_t b200m_norm_table_elems(int32_t n_img, int64_t k_stride) {
  if (n_img <= 0 || k_stride <= 0) return 0;
  const int64_t rows = static_cast<int64_t>(n_img) * k_stride;
  return rows + rows / 32;
}

int b200m_bank_prepare_range_u8(uint8_t* bank, int32_t n_img, int64_t k_stride, const int32_t* k_per_image,
                                int32_t* norm_table, int32_t img_lo, int32_t img_count, void* stream) {
  if (!bank || !k_per_image || !norm_table || n_img <= 0) return fail(B200M_E_BAD_ARG, "null argument or n_img <= 0%s");
  if (k_stride <= 0 || (k_stride % B200M_TILE_T) != 0)
    return fail(B200M_E_BAD_ARG, "k_stride must be a positive multiple of 256%s");
  if (img_lo < 0 || img_count < 0 || static_cast<int64_t>(img_lo) + img_count > n_img)
    return fail(B200M_E_BAD_ARG, "image range outside [0, n_img)%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  if (img_count == 0) return B200M_OK;
  // every kernel below indexes relative to its first row, so a range is the same launch on offset pointers;
  // only the section offsets of the tables come from the size of the WHOLE bank
  const int64_t all_rows = static_cast<int64_t>(n_img) * k_stride;
  const int64_t row_lo = static_cast<int64_t>(img_lo) * k_stride;
  const int64_t n_rows = static_cast<int64_t>(img_count) * k_stride;
  const int block = 256;
  int64_t blocks = (n_rows * 32 + block - 1) / block;
  const int64_t max_blocks = static_cast<int64_t>(device_info().sms) * 16;
  if (blocks > max_blocks) blocks = max_blocks;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  b200m::bank_prepare_kernel<<<static_cast<int>(blocks), block, 0, st>>>(bank + row_lo * 128, n_rows, k_stride,
                                                                         k_per_image + img_lo, norm_table + row_lo,
                                                                         B200M_PAD_NORM);
  CUDA_TRY(cudaGetLastError());
  g_prof.launches.fetch_add(2);
  int64_t blocks2 = (n_rows + block - 1) / block;
  if (blocks2 > max_blocks) blocks2 = max_blocks;
  b200m::norm_chunk_min_kernel<<<static_cast<int>(blocks2), block, 0, st>>>(norm_table + row_lo, n_rows,
                                                                            norm_table + all_rows + row_lo / 32);
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}

int b200m_bank_prepare_u8(uint8_t* bank, int32_t n_img, int64_t k_stride, const int32_t* k_per_image,
                          int32_t* norm_table, void* stream) {
  return b200m_bank_prepare_range_u8(bank, n_img, k_stride, k_per_image, norm_table, 0, n_img, stream);
}

int64_t b200m_sort_table_elems(int32_t n_img, int64_t k_stride) {
  if (n_img <= 0 || k_stride <= 0) return 0;
  return sort_table_elems(n_img, k_stride);
}

int b200m_bank_sort_range_u8(const uint8_t* bank, const int32_t* norm_table, const int32_t* k_per_image, int32_t n_img,
                             int64_t k_stride, uint8_t* sorted_bank, int32_t* sort_table, int32_t img_lo,
                             int32_t img_count, void* stream) {
  if (!bank || !norm_table || !k_per_image || !sorted_bank || !sort_table || n_img <= 0)
    return fail(B200M_E_BAD_ARG, "null argument or n_img <= 0%s");
  if (k_stride <= 0 || (k_stride % B200M_TILE_T) != 0)
    return fail(B200M_E_BAD_ARG, "k_stride must be a positive multiple of 256%s");
  if (img_lo < 0 || img_count < 0 || static_cast<int64_t>(img_lo) + img_count > n_img)
    return fail(B200M_E_BAD_ARG, "image range outside [0, n_img)%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  if (img_count == 0) return B200M_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t row_lo = static_cast<int64_t>(img_lo) * k_stride;
  const int64_t R = static_cast<int64_t>(img_count) * k_stride;      // rows of the range
  int64_t n2 = b200m::kSortChunk;
  while (n2 < k_stride) n2 <<= 1;
  // the 64-bit sort keys live in the (not yet written) sorted bank: 8 * n2 <= 16 * k_stride bytes of each 128 * k_stride slab
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(sorted_bank + row_lo * 128);
  const int64_t key_stride = k_stride * 16;
  // sections of the WHOLE bank's table, every pointer moved to the first image of the range (the kernels index
  // relative to their first row / image)
  const SortTable tv = sort_table_view(sort_table, n_img, k_stride);
  int32_t* perm = const_cast<int32_t*>(tv.perm) + row_lo;
  int32_t* inv = const_cast<int32_t*>(tv.inv) + row_lo;
  int32_t* snorm = const_cast<int32_t*>(tv.snorm) + row_lo;
  int32_t* cmin = const_cast<int32_t*>(tv.cmin) + row_lo / 32;
  int32_t* spread = const_cast<int32_t*>(tv.spread) + img_lo;
  int32_t* spread64 = const_cast<int32_t*>(tv.spread64) + img_lo;
  const int32_t* norms = norm_table + row_lo;
  const int64_t total = static_cast<int64_t>(img_count) * n2;
  b200m::sort_init_kernel<<<grid_for(total, 256), 256, 0, st>>>(norms, k_stride, n2, total, keys, key_stride);
  CUDA_TRY(cudaGetLastError());
  const unsigned smem_blocks = static_cast<unsigned>(total / b200m::kSortChunk);
  b200m::sort_smem_kernel<true><<<smem_blocks, b200m::kSortThreads, 0, st>>>(keys, n2, key_stride, 0);
  CUDA_TRY(cudaGetLastError());
  g_prof.launches.fetch_add(2);
  for (int64_t size = 2 * b200m::kSortChunk; size <= n2; size <<= 1) {
    for (int64_t stride = size / 2; stride >= b200m::kSortChunk; stride >>= 1) {
      b200m::sort_global_kernel<<<grid_for(total / 2, 256), 256, 0, st>>>(keys, n2, key_stride, img_count, size, stride);
      CUDA_TRY(cudaGetLastError());
      g_prof.launches.fetch_add(1);
    }
    b200m::sort_smem_kernel<false><<<smem_blocks, b200m::kSortThreads, 0, st>>>(keys, n2, key_stride, size);
    CUDA_TRY(cudaGetLastError());
    g_prof.launches.fetch_add(1);
  }
  b200m::sort_extract_kernel<<<grid_for(R, 256), 256, 0, st>>>(keys, key_stride, k_stride, R, perm, inv, snorm);
  CUDA_TRY(cudaGetLastError());
  b200m::sort_gather_kernel<<<grid_for(R * 32, 256), 256, 0, st>>>(bank + row_lo * 128, perm, k_stride, R,
                                                                   sorted_bank + row_lo * 128);
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemsetAsync(spread, 0, sizeof(int32_t) * img_count, st));
  CUDA_TRY(cudaMemsetAsync(spread64, 0, sizeof(int32_t) * img_count, st));
  b200m::sort_chunk_stats_kernel<<<grid_for(R / 32, 256), 256, 0, st>>>(snorm, k_per_image + img_lo, k_stride, R / 32,
                                                                      cmin, spread, spread64);
  CUDA_TRY(cudaGetLastError());
  g_prof.launches.fetch_add(3);
  return B200M_OK;
}

int b200m_bank_sort_u8(const uint8_t* bank, const int32_t* norm_table, const int32_t* k_per_image, int32_t n_img,
                       int64_t k_stride, uint8_t* sorted_bank, int32_t* sort_table, void* stream) {
  if (sort_table && n_img > 0 && k_stride > 0)      // the padding entries of the two spread arrays (n_img rounded up to 4) stay defined
    cudaMemsetAsync(sort_table + 3 * static_cast<int64_t>(n_img) * k_stride + static_cast<int64_t>(n_img) * k_stride / 32, 0,
                    sizeof(int32_t) * 2 * ((n_img + 3) & ~3), static_cast<cudaStream_t>(stream));
  return b200m_bank_sort_range_u8(bank, norm_table, k_per_image, n_img, k_stride, sorted_bank, sort_table, 0, n_img, stream);
}

size_t b200m_workspace_bytes(int64_t k_stride, int64_t n_pairs) {
  OptScope opt_scope;
  if (k_stride <= 0 || n_pairs <= 0) return 256;
  return carve(nullptr, nullptr, batch_cap(k_stride, n_pairs), k_stride);
}

int b200m_knn2_u8(const uint8_t* bank, const int32_t* norm_table, const int32_t* k_per_image, int32_t n_img,
                  int64_t k_stride, const int32_t* pairs, int64_t n_pairs, int32_t* out_train_idx,
                  float* out_dist, void* workspace, size_t workspace_bytes, void* stream) {
  OptScope opt_scope;
  int rc = check_common(bank, norm_table, k_per_image, n_img, k_stride, pairs, n_pairs);
  if (rc) return rc;
  if (n_pairs == 0) return B200M_OK;
  if (!out_train_idx || !out_dist || !workspace) return fail(B200M_E_BAD_ARG, "null output / workspace%s");
  if (workspace_bytes < b200m_workspace_bytes(k_stride, n_pairs)) return fail(B200M_E_WORKSPACE, "workspace too small%s");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUtensorMap tmap;
  rc = make_row_tmap(&tmap, bank, static_cast<uint64_t>(n_img) * k_stride);
  if (rc) return rc;
  const int64_t B = batch_pairs(k_stride, n_pairs);
  Workspace w;
  carve(&w, workspace, B, k_stride);
  for (int64_t p0 = 0; p0 < n_pairs; p0 += B) {
    const int64_t nb = (n_pairs - p0 < B) ? (n_pairs - p0) : B;
    const b200m::SweepParams sp = base_params(norm_table, n_img, k_per_image, k_stride, pairs + 2 * p0, nb, w);
    rc = sweep_rows(tmap, sp, true, st);
    if (rc) return rc;
    dim3 grid(static_cast<unsigned>((k_stride + 255) / 256), static_cast<unsigned>(nb));
    b200m::knn2_finalize_kernel<<<grid, 256, 0, st>>>(w.rowres, pairs + 2 * p0, k_per_image, n_img, k_stride,
                                                      out_train_idx + 2 * p0 * k_stride, out_dist + 2 * p0 * k_stride);
    CUDA_TRY(cudaGetLastError());
    g_prof.launches.fetch_add(1);
  }
  return B200M_OK;
}

int b200m_knn2_sorted_u8(const uint8_t* bank, const int32_t* norm_table, const uint8_t* sorted_bank,
                         const int32_t* sort_table, const int32_t* k_per_image, int32_t n_img, int64_t k_stride,
                         const int32_t* pairs, int64_t n_pairs, int32_t* out_train_idx, float* out_dist, void* workspace,
                         size_t workspace_bytes, void* stream) {
  OptScope opt_scope;
  int rc = check_common(bank, norm_table, k_per_image, n_img, k_stride, pairs, n_pairs);
  if (rc) return rc;
  if (n_pairs == 0) return B200M_OK;
  if (!sorted_bank || !sort_table) return fail(B200M_E_BAD_ARG, "null sorted_bank / sort_table (use b200m_knn2_u8 without them)%s");
  if (!out_train_idx || !out_dist || !workspace) return fail(B200M_E_BAD_ARG, "null output / workspace%s");
  if (workspace_bytes < b200m_workspace_bytes(k_stride, n_pairs)) return fail(B200M_E_WORKSPACE, "workspace too small%s");
  if (k_stride / b200m::kKnnChunk > b200m::kKnnMaxChunks)      // beyond 262,144 rows per image: the exact sweep
    return b200m_knn2_u8(bank, norm_table, k_per_image, n_img, k_stride, pairs, n_pairs, out_train_idx, out_dist, workspace,
                         workspace_bytes, stream);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t B = batch_pairs(k_stride, n_pairs);
  Workspace w;
  carve(&w, workspace, B, k_stride);
  CUtensorMap tmap_bank, tmap_sbank, tmap_gq;
  rc = make_row_tmap(&tmap_bank, bank, static_cast<uint64_t>(n_img) * k_stride);
  if (rc) return rc;
  rc = make_row_tmap(&tmap_sbank, sorted_bank, static_cast<uint64_t>(n_img) * k_stride);
  if (rc) return rc;
  rc = make_row_tmap(&tmap_gq, w.gath_q, static_cast<uint64_t>(B) * k_stride);
  if (rc) return rc;
  const SortTable tv = sort_table_view(sort_table, n_img, k_stride);
  const int64_t R = B * k_stride;
  CUDA_TRY(cudaMemsetAsync(w.stats, 0, sizeof(unsigned long long) * b200m::kStatSlots * b200m::kStatFields, st));
  for (int64_t p0 = 0; p0 < n_pairs; p0 += B) {
    const int64_t nb = (n_pairs - p0 < B) ? (n_pairs - p0) : B;
    const int32_t* pb = pairs + 2 * p0;
    b200m::ChunkParams cp;
    cp.cmin = tv.cmin;
    cp.k_per_image = k_per_image;
    cp.pairs = pb;
    cp.k_stride = k_stride;
    cp.n_pairs = static_cast<int32_t>(nb);
    cp.blocks_per_pair = static_cast<int32_t>(k_stride / b200m::kRows);
    cp.n_img = n_img;
    cp.rowres = w.rowres;
    cp.peak = 0;
    cp.chunk_cols = 64;
    rc = launch_chunk_sweep(tmap_sbank, cp, sweep_grid(nb * cp.blocks_per_pair), 5, st, true);
    if (rc) return rc;
    CUDA_TRY(cudaMemsetAsync(w.fb_count, 0, sizeof(int32_t) * (B + 64), st));
    b200m::KnnParams kp;
    kp.bank = bank;
    kp.sbank = sorted_bank;
    kp.perm = tv.perm;
    kp.snorm = tv.snorm;
    kp.k_per_image = k_per_image;
    kp.pairs = pb;
    kp.k_stride = k_stride;
    kp.n_img = n_img;
    kp.n_chunks = static_cast<int32_t>(k_stride / b200m::kKnnChunk);
    kp.rowres_s = w.rowres;
    kp.boff = reinterpret_cast<int32_t*>(w.cand_d);                       // nb * (n_chunks + 1) <= R entries
    kp.tasks = reinterpret_cast<int32_t*>(w.gath_t);                       // 2 R entries
    kp.taskres = reinterpret_cast<int4*>(w.gath_t + 8 * R);                // 2 R records of 16 bytes (40 R of the buffer's 128 R bytes)
    kp.rowres_o = w.rowres_o;
    kp.fb_count = w.fb_count;
    kp.fb_total = w.fb_count + B;
    kp.fb_q = w.fb_q;
    kp.fb_desc = w.gath_q;
    kp.stats = w.stats;
    {
      ProfScope ps(&g_prof.refine, st);
      b200m::knn_bucket_kernel<<<static_cast<unsigned>(nb), 1024, 0, st>>>(kp);
      if (opt().knn_eval == 0)
        b200m::knn_eval_tc_kernel<<<dim3(static_cast<unsigned>(kp.n_chunks), static_cast<unsigned>(nb)), b200m::kKnnTcThreads, 0, st>>>(kp);
      else
        b200m::knn_eval_kernel<<<dim3(static_cast<unsigned>(kp.n_chunks), static_cast<unsigned>(nb)), b200m::kKnnEvalThreads, 0, st>>>(kp);
      b200m::knn_combine_kernel<<<dim3(static_cast<unsigned>((k_stride + 255) / 256), static_cast<unsigned>(nb)), 256, 0, st>>>(kp);
    }
    CUDA_TRY(cudaGetLastError());
    g_prof.launches.fetch_add(2);
    // exact top-2 sweep over the rows the bounds could not settle (an outside column may tie or beat the runner-up)
    b200m::SweepParams sp = base_params(norm_table, n_img, k_per_image, k_stride, pb, nb, w);
    sp.fb_total = w.fb_count + B;
    rc = launch_sweep<b200m::kTop2G>(tmap_gq, tmap_gq, tmap_bank, sp, sweep_grid(nb * sp.blocks_per_pair), st);
    if (rc) return rc;
    dim3 grid(static_cast<unsigned>((k_stride + 255) / 256), static_cast<unsigned>(nb));
    b200m::knn2_finalize_kernel<<<grid, 256, 0, st>>>(w.rowres_o, pb, k_per_image, n_img, k_stride,
                                                      out_train_idx + 2 * p0 * k_stride, out_dist + 2 * p0 * k_stride);
    CUDA_TRY(cudaGetLastError());
    g_prof.launches.fetch_add(1);
  }
  return B200M_OK;
}

int b200m_match_pairs_u8(const uint8_t* bank, const int32_t* norm_table, const uint8_t* sorted_bank,
                         const int32_t* sort_table, const int32_t* k_per_image, int32_t n_img, int64_t k_stride,
                         const int32_t* pairs, int64_t n_pairs, double ratio, int mutual, int32_t* out_query_idx,
                         int32_t* out_train_idx, float* out_dist, int32_t* out_count, int64_t cap, void* workspace,
                         size_t workspace_bytes, void* stream) {
  OptScope opt_scope;
  int rc = check_common(bank, norm_table, k_per_image, n_img, k_stride, pairs, n_pairs);
  if (rc) return rc;
  if (n_pairs == 0) return B200M_OK;
  if (!out_query_idx || !out_train_idx || !out_dist || !out_count || !workspace || cap <= 0)
    return fail(B200M_E_BAD_ARG, "null output / workspace or cap <= 0%s");
  if (workspace_bytes < b200m_workspace_bytes(k_stride, n_pairs)) return fail(B200M_E_WORKSPACE, "workspace too small%s");
  if ((sorted_bank == nullptr) != (sort_table == nullptr))
    return fail(B200M_E_BAD_ARG, "sorted_bank and sort_table must be given together%s");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t B = batch_pairs(k_stride, n_pairs);
  Workspace w;
  carve(&w, workspace, B, k_stride);
  CUtensorMap tmap_bank, tmap_sbank, tmap_gq, tmap_gt;
  rc = make_row_tmap(&tmap_bank, bank, static_cast<uint64_t>(n_img) * k_stride);
  if (rc) return rc;
  rc = make_row_tmap(&tmap_gq, w.gath_q, static_cast<uint64_t>(B) * k_stride);
  if (rc) return rc;
  rc = make_row_tmap(&tmap_gt, w.gath_t, static_cast<uint64_t>(B) * k_stride);
  if (rc) return rc;
  const int do_mutual = mutual ? 1 : 0;
  CUDA_TRY(cudaMemsetAsync(w.stats, 0, sizeof(unsigned long long) * b200m::kStatSlots * b200m::kStatFields, st));

  // ---------------------------------------------------------------- pipeline 2: chunk sweep on the sorted bank
  if (opt().pipeline == 2 && sorted_bank && opt().epilogue != 0) {
    rc = make_row_tmap(&tmap_sbank, sorted_bank, static_cast<uint64_t>(n_img) * k_stride);
    if (rc) return rc;
    const SortTable tv = sort_table_view(sort_table, n_img, k_stride);
    const int do_ratio = ratio > 0.0 ? 1 : 0;
    const int64_t opt_kernel = opt().kernel;
    for (int64_t p0 = 0; p0 < n_pairs; p0 += B) {
      const int64_t nb = (n_pairs - p0 < B) ? (n_pairs - p0) : B;
      const int32_t* pb = pairs + 2 * p0;
      b200m::ChunkParams cp;
      cp.cmin = tv.cmin;
      cp.k_per_image = k_per_image;
      cp.pairs = pb;
      cp.k_stride = k_stride;
      cp.n_pairs = static_cast<int32_t>(nb);
      cp.blocks_per_pair = static_cast<int32_t>(k_stride / b200m::kRows);
      cp.n_img = n_img;
      cp.rowres = w.rowres;
      cp.peak = opt().epilogue == 2 ? 1 : 0;
      cp.chunk_cols = opt_kernel == 6 ? 64 : (opt_kernel == 5 && opt().chunk == 64 && !cp.peak) ? 64 : 32;
      rc = launch_chunk_sweep(tmap_sbank, cp, opt_kernel == 6 ? sweep_grid(1 << 20) : sweep_grid(nb * cp.blocks_per_pair),
                              opt_kernel, st);
      if (rc) return rc;
      CUDA_TRY(cudaMemsetAsync(w.fb_count, 0, sizeof(int32_t) * (B + 64), st));   // per-pair counts, the batch total at [B], the verify work-list length at [B + 1]
      b200m::RefineParams rp;
      rp.bank = bank;
      rp.sbank = sorted_bank;
      rp.sqnorm = norm_table;
      rp.perm = tv.perm;
      rp.inv = tv.inv;
      rp.snorm = tv.snorm;
      rp.spread = cp.chunk_cols == 64 ? tv.spread64 : tv.spread;
      rp.chunk_cols = cp.chunk_cols;
      rp.k_per_image = k_per_image;
      rp.pairs = pb;
      rp.k_stride = k_stride;
      rp.n_img = n_img;
      rp.rowres_s = w.rowres;
      rp.rowres_o = w.rowres_o;
      rp.fb_count = w.fb_count;
      rp.fb_total = w.fb_count + B;
      rp.fb_q = w.fb_q;
      rp.fb_desc = w.gath_q;
      rp.ratio = ratio;
      rp.inv_ratio_sq = ratio > 0.0 ? 1.0 / (ratio * ratio) : 0.0;
      rp.stats = w.stats;
      rp.need_count = w.cand_count_a;       // free until select2_kernel writes it
      rp.need_list = w.cand_q;
      {
        dim3 grid(static_cast<unsigned>((k_stride + b200m::kRefineThreads - 1) / b200m::kRefineThreads), static_cast<unsigned>(nb));
        if (opt().refine == 1) {
          CUDA_TRY(cudaMemsetAsync(rp.need_count, 0, sizeof(int32_t) * nb, st));
          ProfScope ps(&g_prof.refine, st);
          b200m::refine_classify_kernel<<<grid, b200m::kRefineThreads, 0, st>>>(rp);
          b200m::refine_eval_kernel<<<static_cast<unsigned>(device_info().sms * 5), b200m::kRefineThreads, 0, st>>>(rp, static_cast<int>(nb));
        } else {
          ProfScope ps(&g_prof.refine, st);
          b200m::refine_kernel<<<grid, b200m::kRefineThreads, 0, st>>>(rp);
        }
      }
      CUDA_TRY(cudaGetLastError());
      // exact fallback for the rows the bounds could not settle (usually none: the kernel finds empty lists)
      b200m::SweepParams sp = base_params(norm_table, n_img, k_per_image, k_stride, pb, nb, w);
      sp.fb_total = w.fb_count + B;
      rc = launch_sweep<b200m::kTop1G>(tmap_gq, tmap_gq, tmap_bank, sp, sweep_grid(nb * sp.blocks_per_pair), st);
      if (rc) return rc;
      {
        ProfScope ps(&g_prof.select, st);
        b200m::select2_kernel<1024><<<static_cast<unsigned>(nb), 1024, 0, st>>>(
            w.rowres_o, pb, k_per_image, n_img, k_stride, ratio, bank, norm_table, do_ratio, do_mutual, w.cand_q, w.cand_t,
            w.cand_d, w.cand_thr_mut, w.cand_rslot, w.cand_count, w.cand_thr_ratio, w.cand_excl, w.cand_count_a,
            w.gath_q, w.gath_t, w.vitems, w.fb_count + B + 1, w.stats);
      }
      CUDA_TRY(cudaGetLastError());
      if (do_ratio || do_mutual) {
        sp.do_ratio = do_ratio;
        sp.do_mutual = do_mutual;
        sp.cand_count_r = w.cand_count_a;
        sp.cand_excl = w.cand_excl;
        sp.vitems = w.vitems;                 // only the non-empty blocks (written by select2_kernel)
        sp.vitem_count = w.fb_count + B + 1;
        rc = launch_sweep<b200m::kVerify>(tmap_gq, tmap_gt, tmap_bank, sp, sweep_grid(nb * sp.blocks_per_pair * 2), st);
        if (rc) return rc;
      }
      {
        ProfScope ps(&g_prof.emit, st);
        b200m::emit_kernel<1024><<<static_cast<unsigned>(nb), 1024, 0, st>>>(
            w.cand_q, w.cand_t, w.cand_d, w.viol_ratio, w.viol_mut, w.cand_count, w.cand_rslot, k_stride, do_ratio,
            do_mutual, out_query_idx + p0 * cap, out_train_idx + p0 * cap, out_dist + p0 * cap, out_count + p0, cap,
            w.stats);
      }
      CUDA_TRY(cudaGetLastError());
    }
    return mark_invalid_pairs(pairs, n_pairs, n_img, out_count, st);
  }

  // ---------------------------------------------------------------- pipelines 0 / 1 (original bank only)
  // pipeline 1: top-1 sweep, the Lowe test is settled by a verify sweep over the candidates;
  // pipeline 0 (and the debug epilogue): exact top-2 sweep, only the mutual check is verified.
  const bool exact2 = opt().pipeline == 0 || opt().epilogue == 0;
  const int do_ratio = (!exact2 && ratio > 0.0) ? 1 : 0;
  for (int64_t p0 = 0; p0 < n_pairs; p0 += B) {
    const int64_t nb = (n_pairs - p0 < B) ? (n_pairs - p0) : B;
    const int32_t* pb = pairs + 2 * p0;
    b200m::SweepParams sp = base_params(norm_table, n_img, k_per_image, k_stride, pb, nb, w);
    rc = sweep_rows(tmap_bank, sp, exact2, st);
    if (rc) return rc;
    {
      ProfScope ps(&g_prof.select, st);
      b200m::select_kernel<1024><<<static_cast<unsigned>(nb), 1024, 0, st>>>(
          w.rowres, pb, k_per_image, n_img, k_stride, ratio, exact2 ? 1 : 0, bank, norm_table, do_ratio, do_mutual, w.cand_q,
          w.cand_t, w.cand_d, w.cand_thr_ratio, w.cand_thr_mut, w.cand_count, w.gath_q, w.gath_t);
    }
    CUDA_TRY(cudaGetLastError());
    if (do_ratio || do_mutual) {
      sp.do_ratio = do_ratio;
      sp.do_mutual = do_mutual;
      const int grid = sweep_grid(nb * sp.blocks_per_pair * 2);
      rc = launch_any<b200m::kVerify>(tmap_gq, tmap_gt, tmap_bank, sp, grid, st);
      if (rc) return rc;
    }
    {
      ProfScope ps(&g_prof.emit, st);
      b200m::emit_kernel<1024><<<static_cast<unsigned>(nb), 1024, 0, st>>>(
          w.cand_q, w.cand_t, w.cand_d, w.viol_ratio, w.viol_mut, w.cand_count, nullptr, k_stride, do_ratio, do_mutual,
          out_query_idx + p0 * cap, out_train_idx + p0 * cap, out_dist + p0 * cap, out_count + p0, cap, w.stats);
    }
    CUDA_TRY(cudaGetLastError());
  }
  return mark_invalid_pairs(pairs, n_pairs, n_img, out_count, st);
}

size_t b200m_workspace_bytes_bf16(int64_t k_stride, int64_t n_pairs) {
  OptScope opt_scope;
  if (k_stride <= 0 || n_pairs <= 0) return 256;
  return static_cast<size_t>(2) * batch_cap(k_stride, n_pairs) * k_stride * sizeof(float2) + 512;
}

int b200m_match_pairs_bf16(const void* bank_bf16, const int32_t* k_per_image, int32_t n_img, int64_t k_stride,
                           int32_t dim, const int32_t* pairs, int64_t n_pairs, float score_threshold,
                           int use_threshold, int mutual, int32_t* out_query_idx, int32_t* out_train_idx,
                           float* out_score, int32_t* out_count, int64_t cap, void* workspace,
                           size_t workspace_bytes, void* stream) {
  OptScope opt_scope;
  if (!bank_bf16 || !k_per_image || n_img <= 0) return fail(B200M_E_BAD_ARG, "null bank / k_per_image or n_img <= 0%s");
  if (dim != 128 && dim != 256) return fail(B200M_E_BAD_DIM, "the bf16 path handles descriptor lengths 128 and 256%s");
  if (k_stride <= 0 || (k_stride % B200M_TILE_T) != 0)
    return fail(B200M_E_BAD_ARG, "k_stride must be a positive multiple of 256%s");
  if (n_pairs < 0 || (n_pairs > 0 && !pairs)) return fail(B200M_E_BAD_ARG, "bad pairs / n_pairs%s");
  if (static_cast<int64_t>(n_img) * k_stride >= (int64_t(1) << 31)) return fail(B200M_E_BAD_ARG, "bank has more than 2^31 rows%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  if (n_pairs == 0) return B200M_OK;
  if (!out_query_idx || !out_train_idx || !out_score || !out_count || !workspace || cap <= 0)
    return fail(B200M_E_BAD_ARG, "null output / workspace or cap <= 0%s");
  if (workspace_bytes < b200m_workspace_bytes_bf16(k_stride, n_pairs)) return fail(B200M_E_WORKSPACE, "workspace too small%s");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUtensorMap tmap;
  int rc = make_row_tmap_bf16(&tmap, bank_bf16, static_cast<uint64_t>(n_img) * k_stride, static_cast<uint32_t>(dim));
  if (rc) return rc;
  const int64_t B = batch_pairs(k_stride, n_pairs);
  float2* rowbest = reinterpret_cast<float2*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  for (int64_t p0 = 0; p0 < n_pairs; p0 += B) {
    const int64_t nb = (n_pairs - p0 < B) ? (n_pairs - p0) : B;
    if (opt().float_sweep == 0) {
      rc = float_single_sweep_batch(tmap, k_per_image, n_img, k_stride, dim, pairs + 2 * p0, nb, B * k_stride, rowbest, nullptr,
                                    score_threshold, use_threshold, mutual ? 1 : 0, out_query_idx + p0 * cap,
                                    out_train_idx + p0 * cap, out_score + p0 * cap, out_count + p0, cap, st);
      if (rc) return rc;
      continue;
    }
    b200m::ip::IpParams ip;
    ip.k_per_image = k_per_image;
    ip.pairs = pairs + 2 * p0;
    ip.k_stride = k_stride;
    ip.n_pairs = static_cast<int32_t>(nb);
    ip.n_dirs = mutual ? 2 : 1;
    ip.blocks_per_image = static_cast<int32_t>(k_stride / b200m::ip::kRowsF);
    ip.n_img = n_img;
    ip.rowbest = rowbest;
    ip.z = nullptr;
    ip.nb = nullptr;
    ip.gate = nullptr;
    const int grid = sweep_grid(static_cast<int64_t>(ip.n_dirs) * nb * ip.blocks_per_image);
    rc = dim == 256 ? launch_ip_sweep<256>(tmap, ip, grid, st) : launch_ip_sweep<128>(tmap, ip, grid, st);
    if (rc) return rc;
    {
      ProfScope ps(&g_prof.emit, st);
      b200m::ip::ip_finalize_kernel<1024><<<static_cast<unsigned>(nb), 1024, 0, st>>>(
          rowbest, pairs + 2 * p0, k_per_image, n_img, k_stride, static_cast<int32_t>(nb), score_threshold, use_threshold,
          mutual ? 1 : 0, 0, out_query_idx + p0 * cap, out_train_idx + p0 * cap, out_score + p0 * cap, out_count + p0, cap);
    }
    CUDA_TRY(cudaGetLastError());
  }
  return mark_invalid_pairs(pairs, n_pairs, n_img, out_count, st);
}

// Head workspace: row results float2 [2][B][k_stride] (pass 2 reuses the second half for column keys and claims) |
// nb float [2][B][k_stride] | largest row norm per image float [kMaxNormImgs] | flag | column partial sums of the
// single-sweep log-sum-exp float [B][k_stride / 256][k_stride].
constexpr int64_t kMaxNormImgs = 1 << 18;
size_t b200m_workspace_bytes_assign(int64_t k_stride, int64_t n_pairs) {
  OptScope opt_scope;
  if (k_stride <= 0 || n_pairs <= 0) return 256;
  const size_t B = static_cast<size_t>(batch_cap(k_stride, n_pairs));
  // the column partial sums -- the one term that grows with k_stride^2 (1 GiB at k_stride = 16,384, B = 256) -- are
  // only needed when the single-sweep log-sum-exp can run at all, i.e. with "float_sweep" = 0
  const size_t colpart = opt().float_sweep == 0 ? B * static_cast<size_t>(k_stride / 256) * k_stride * sizeof(float) : 0;
  return 2 * B * k_stride * (sizeof(float2) + sizeof(float)) + kMaxNormImgs * sizeof(float) + 1024 + colpart + 1024;
}

int b200m_assign_pairs_bf16(const void* mdesc_bf16, const float* matchability, const int32_t* k_per_image,
                            int32_t n_img, int64_t k_stride, int32_t dim, const int32_t* pairs, int64_t n_pairs,
                            float filter_threshold, int use_threshold, int32_t* out_query_idx, int32_t* out_train_idx,
                            float* out_score, int32_t* out_count, int64_t cap, void* workspace, size_t workspace_bytes,
                            void* stream) {
  OptScope opt_scope;
  if (!mdesc_bf16 || !matchability || !k_per_image || n_img <= 0)
    return fail(B200M_E_BAD_ARG, "null descriptors / matchability / k_per_image or n_img <= 0%s");
  if (dim != 128 && dim != 256) return fail(B200M_E_BAD_DIM, "the bf16 path handles descriptor lengths 128 and 256%s");
  if (k_stride <= 0 || (k_stride % B200M_TILE_T) != 0)
    return fail(B200M_E_BAD_ARG, "k_stride must be a positive multiple of 256%s");
  if (n_pairs < 0 || (n_pairs > 0 && !pairs)) return fail(B200M_E_BAD_ARG, "bad pairs / n_pairs%s");
  if (static_cast<int64_t>(n_img) * k_stride >= (int64_t(1) << 31)) return fail(B200M_E_BAD_ARG, "bank has more than 2^31 rows%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  if (n_pairs == 0) return B200M_OK;
  if (!out_query_idx || !out_train_idx || !out_score || !out_count || !workspace || cap <= 0)
    return fail(B200M_E_BAD_ARG, "null output / workspace or cap <= 0%s");
  if (workspace_bytes < b200m_workspace_bytes_assign(k_stride, n_pairs)) return fail(B200M_E_WORKSPACE, "workspace too small%s");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUtensorMap tmap;
  int rc = make_row_tmap_bf16(&tmap, mdesc_bf16, static_cast<uint64_t>(n_img) * k_stride, static_cast<uint32_t>(dim));
  if (rc) return rc;
  const int64_t B = batch_pairs(k_stride, n_pairs);
  float2* rowbest = reinterpret_cast<float2*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  float* nbuf = reinterpret_cast<float*>(rowbest + 2 * B * k_stride);
  float* maxnorm = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(nbuf + 2 * B * k_stride) + 255) & ~uintptr_t(255));
  int32_t* lse_flag = reinterpret_cast<int32_t*>(maxnorm + kMaxNormImgs);
  float* colpart = reinterpret_cast<float*>(lse_flag + 64);
  // pass 1 as ONE sweep per pair too (row and column log-sum-exp from the same exponentials) unless the two-direction
  // sweeps are asked for; needs the largest row norm of every image as the common reference point
  const bool lse_single = opt().float_sweep == 0 && n_img <= kMaxNormImgs;
  if (lse_single) {
    CUDA_TRY(cudaMemsetAsync(maxnorm, 0, sizeof(float) * n_img, st));
    const int64_t rows = static_cast<int64_t>(n_img) * k_stride;
    const unsigned blocks = static_cast<unsigned>((rows + 7) / 8);
    if (dim == 256)
      b200m::ip::ip_maxnorm_kernel<256><<<blocks, 256, 0, st>>>(static_cast<const uint16_t*>(mdesc_bf16), k_per_image, k_stride, n_img, maxnorm);
    else
      b200m::ip::ip_maxnorm_kernel<128><<<blocks, 256, 0, st>>>(static_cast<const uint16_t*>(mdesc_bf16), k_per_image, k_stride, n_img, maxnorm);
    CUDA_TRY(cudaGetLastError());
  }
  for (int64_t p0 = 0; p0 < n_pairs; p0 += B) {
    const int64_t nb = (n_pairs - p0 < B) ? (n_pairs - p0) : B;
    b200m::ip::IpParams ip;
    ip.k_per_image = k_per_image;
    ip.pairs = pairs + 2 * p0;
    ip.k_stride = k_stride;
    ip.n_pairs = static_cast<int32_t>(nb);
    ip.n_dirs = 2;
    ip.blocks_per_image = static_cast<int32_t>(k_stride / b200m::ip::kRowsF);
    ip.n_img = n_img;
    ip.rowbest = rowbest;
    ip.z = matchability;
    ip.nb = nbuf;
    ip.gate = nullptr;
    const int grid = sweep_grid(static_cast<int64_t>(2) * nb * ip.blocks_per_image);
    if (lse_single) {
      CUDA_TRY(cudaMemsetAsync(lse_flag, 0, sizeof(int32_t), st));
      b200m::ip::Ip2Params lp;
      lp.k_per_image = k_per_image;
      lp.pairs = pairs + 2 * p0;
      lp.k_stride = k_stride;
      lp.n_pairs = static_cast<int32_t>(nb);
      lp.blocks_per_image = static_cast<int32_t>(k_stride / b200m::ip::kRows2);
      lp.n_img = n_img;
      lp.rowbest = nullptr;
      lp.colkey = nullptr;
      lp.nb = nullptr;
      lp.z = matchability;
      lp.maxnorm = maxnorm;
      lp.nbw = nbuf;
      lp.colpart = colpart;
      lp.lse_flag = lse_flag;
      const int grid2 = sweep_grid(nb * lp.blocks_per_image);
      rc = dim == 256 ? launch_ip_sweep2<256, b200m::ip::kIpLse>(tmap, lp, grid2, st)
                      : launch_ip_sweep2<128, b200m::ip::kIpLse>(tmap, lp, grid2, st);
      if (rc) return rc;
      b200m::ip::ip_lse_finalize_kernel<<<dim3(static_cast<unsigned>((k_stride + 255) / 256), static_cast<unsigned>(nb)), 256, 0, st>>>(
          colpart, maxnorm, matchability, pairs + 2 * p0, k_per_image, n_img, k_stride, static_cast<int32_t>(nb), nbuf, lse_flag);
      CUDA_TRY(cudaGetLastError());
      ip.gate = lse_flag;      // a sum left the fp32 range around the common reference: redo the batch with per-row maxima
    }
    // pass 1 (two-direction form): row and column log-sum-exp -> nb; pass 2: arg-max of the log assignment
    rc = dim == 256 ? launch_ip_sweep<256, b200m::ip::kIpLse>(tmap, ip, grid, st)
                    : launch_ip_sweep<128, b200m::ip::kIpLse>(tmap, ip, grid, st);
    if (rc) return rc;
    if (opt().float_sweep == 0) {      // pass 2 as ONE sweep per pair: row arg-max and column maxima of the full score
      rc = float_single_sweep_batch(tmap, k_per_image, n_img, k_stride, dim, pairs + 2 * p0, nb, B * k_stride, rowbest, nbuf,
                                    filter_threshold, use_threshold, 1, out_query_idx + p0 * cap, out_train_idx + p0 * cap,
                                    out_score + p0 * cap, out_count + p0, cap, st);
      if (rc) return rc;
      continue;
    }
    rc = dim == 256 ? launch_ip_sweep<256, b200m::ip::kIpBiasArgmax>(tmap, ip, grid, st)
                    : launch_ip_sweep<128, b200m::ip::kIpBiasArgmax>(tmap, ip, grid, st);
    if (rc) return rc;
    {
      ProfScope ps(&g_prof.emit, st);
      b200m::ip::ip_finalize_kernel<1024><<<static_cast<unsigned>(nb), 1024, 0, st>>>(
          rowbest, pairs + 2 * p0, k_per_image, n_img, k_stride, static_cast<int32_t>(nb), filter_threshold, use_threshold, 1,
          1, out_query_idx + p0 * cap, out_train_idx + p0 * cap, out_score + p0 * cap, out_count + p0, cap);
    }
    CUDA_TRY(cudaGetLastError());
  }
  return mark_invalid_pairs(pairs, n_pairs, n_img, out_count, st);
}

size_t b200m_workspace_bytes_ransac(int64_t n_pairs, int32_t iters) {
  if (n_pairs <= 0 || iters <= 0) return 256;
  const int64_t bpp = (iters + b200m::geo::kRansacThreads - 1) / b200m::geo::kRansacThreads;
  return static_cast<size_t>(n_pairs * bpp) * (sizeof(int32_t) + 9 * sizeof(double)) + 512;
}

int b200m_essential_ransac_f32(const float* matches_xyxy, const int64_t* offsets, int64_t n_pairs, int32_t iters,
                               float threshold_norm, uint64_t seed, float max_depth, double* out_E, uint8_t* out_mask,
                               int32_t* out_inliers, double* out_R, double* out_t, uint8_t* out_pose_mask,
                               int32_t* out_pose_inliers, void* workspace, size_t workspace_bytes, void* stream) {
  if (n_pairs == 0) return B200M_OK;
  if (!matches_xyxy || !offsets || n_pairs < 0 || iters <= 0 || !(threshold_norm > 0.f))
    return fail(B200M_E_BAD_ARG, "null matches / offsets, or bad n_pairs / iters / threshold%s");
  if (!out_E || !out_mask || !out_inliers || !out_R || !out_t || !out_pose_mask || !out_pose_inliers || !workspace)
    return fail(B200M_E_BAD_ARG, "null output / workspace%s");
  if ((reinterpret_cast<uintptr_t>(matches_xyxy) & 15) != 0) return fail(B200M_E_BAD_ARG, "matches must be 16-byte aligned%s");
  if (workspace_bytes < b200m_workspace_bytes_ransac(n_pairs, iters)) return fail(B200M_E_WORKSPACE, "workspace too small%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int bpp = (iters + b200m::geo::kRansacThreads - 1) / b200m::geo::kRansacThreads;
  double* blk_E = reinterpret_cast<double*>((reinterpret_cast<uintptr_t>(workspace) + 255) & ~uintptr_t(255));
  int32_t* blk_count = reinterpret_cast<int32_t*>(blk_E + static_cast<size_t>(n_pairs) * bpp * 9);
  b200m::geo::RansacParams rp;
  rp.matches = reinterpret_cast<const float4*>(matches_xyxy);
  rp.offsets = offsets;
  rp.n_pairs = static_cast<int32_t>(n_pairs);
  rp.blocks_per_pair = bpp;
  rp.thr2 = threshold_norm * threshold_norm;
  rp.seed = seed;
  rp.blk_count = blk_count;
  rp.blk_E = blk_E;
  rp.pair0 = 0;
  for (int64_t p0 = 0; p0 < n_pairs; p0 += 65535) {      // gridDim.y limit
    const int64_t np = std::min<int64_t>(65535, n_pairs - p0);
    rp.pair0 = static_cast<int32_t>(p0);
    ProfScope ps(&g_prof.first, st);
    b200m::geo::ransac_kernel<<<dim3(bpp, static_cast<unsigned>(np)), b200m::geo::kRansacThreads, 0, st>>>(rp);
  }
  CUDA_TRY(cudaGetLastError());
  b200m::geo::PoseParams pp;
  pp.matches = rp.matches;
  pp.offsets = offsets;
  pp.n_pairs = rp.n_pairs;
  pp.blocks_per_pair = bpp;
  pp.thr2 = rp.thr2;
  pp.max_depth = max_depth > 0.f ? max_depth : 50.f;
  pp.blk_count = blk_count;
  pp.blk_E = blk_E;
  pp.out_E = out_E;
  pp.out_mask = out_mask;
  pp.out_inliers = out_inliers;
  pp.out_R = out_R;
  pp.out_t = out_t;
  pp.out_pose_mask = out_pose_mask;
  pp.out_pose_inliers = out_pose_inliers;
  {
    ProfScope ps(&g_prof.emit, st);
    b200m::geo::pose_kernel<<<static_cast<unsigned>(n_pairs), b200m::geo::kPoseThreads, 0, st>>>(pp);
  }
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}

int b200m_five_point_f64(const double* x0, const double* x1, double* out_E, int32_t* out_n, void* stream) {
  if (!x0 || !x1 || !out_E || !out_n) return fail(B200M_E_BAD_ARG, "null argument%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  b200m::geo::five_point_debug_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(x0, x1, out_E, out_n);
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}

int b200m_vq_f32(const float* obs, int64_t n_obs, int32_t dim, const double* codebook, int32_t n_codes,
                 int32_t* out_code, double* out_dist, double* code_sqr_scratch, void* stream) {
  if (n_obs == 0) return B200M_OK;
  if (!obs || !codebook || !out_code || !code_sqr_scratch || n_obs < 0 || n_codes <= 0)
    return fail(B200M_E_BAD_ARG, "null argument / bad sizes%s");
  if (dim <= 0 || dim > 256) return fail(B200M_E_BAD_DIM, "vq handles descriptor lengths up to 256%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  b200m::vq_code_sqr_kernel<<<(n_codes + 127) / 128, 128, 0, st>>>(codebook, dim, n_codes, code_sqr_scratch);
  CUDA_TRY(cudaGetLastError());
  const size_t smem = sizeof(double) * b200m::kVqObsPerBlock * dim;
  CUDA_TRY(ensure_smem_attr(b200m::vq_argmin_kernel, 65536));
  const int64_t blocks = (n_obs + b200m::kVqObsPerBlock - 1) / b200m::kVqObsPerBlock;
  b200m::vq_argmin_kernel<<<static_cast<unsigned>(blocks), b200m::kVqThreads, smem, st>>>(
      obs, n_obs, dim, codebook, code_sqr_scratch, n_codes, out_code, out_dist);
  CUDA_TRY(cudaGetLastError());
  g_prof.launches.fetch_add(2);
  return B200M_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// b200m_vq_f32_static: the vq entry point for callers that cannot hand us device scratch (plugins whose engine owns
// the memory pool, bindings without the CUDA runtime, services forbidden to cudaMalloc after start-up). The squared
// codebook norms go into an array that is part of this library's device image instead of caller memory: every GPU that
// loads the library carries these 32 KiB until the process exits. Nothing on this path allocates or frees.
//
// g_vq_static_code_sqr is a __device__ variable. Device code may use the name directly; in HOST code the name is only
// a handle for the runtime, not a device address, so it must never be passed to a kernel as is. The host gets the
// real address from cudaGetSymbolAddress, per device (each device has its own copy at its own address).
static __device__ __align__(256) double g_vq_static_code_sqr[B200M_VQ_STATIC_MAX_CODES];

namespace {

constexpr int kVqStaticMaxDevices = 64;

// Device address of g_vq_static_code_sqr on the CURRENT device; looked up on the first call per device and kept.
// cudaGetSymbolAddress allocates nothing (the array came with the module), but it is not a stream operation, which
// is why the header asks for the first call on a device to happen outside stream capture.
int vq_static_scratch(double** out) {
  static std::mutex mu;
  static double* addr[kVqStaticMaxDevices] = {};
  int dev = -1;
  CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kVqStaticMaxDevices) return fail(B200M_E_BAD_ARG, "b200m_vq_f32_static: device ordinal out of range%s");
  std::lock_guard<std::mutex> lk(mu);
  if (!addr[dev]) {
    void* p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress(&p, g_vq_static_code_sqr));
    if (!p) return fail(B200M_E_CUDA, "cudaGetSymbolAddress returned a null address for the built-in vq scratch%s");
    addr[dev] = static_cast<double*>(p);
  }
  *out = addr[dev];
  return B200M_OK;
}

}  // namespace

int b200m_vq_f32_static(const float* obs, int64_t n_obs, int32_t dim, const double* codebook, int32_t n_codes,
                        int32_t* out_code, double* out_dist, void* stream) {
  if (n_obs == 0) return B200M_OK;
  if (!obs || !codebook || !out_code || n_obs < 0 || n_codes <= 0) return fail(B200M_E_BAD_ARG, "null argument / bad sizes%s");
  if (n_codes > B200M_VQ_STATIC_MAX_CODES) {
    snprintf(g_err, sizeof(g_err),
             "b200m_vq_f32_static: n_codes = %d exceeds the built-in scratch of %d codebook rows "
             "(B200M_VQ_STATIC_MAX_CODES); call b200m_vq_f32 with caller-allocated scratch instead",
             static_cast<int>(n_codes), B200M_VQ_STATIC_MAX_CODES);
    return B200M_E_BAD_ARG;
  }
  if (dim <= 0 || dim > 256) return fail(B200M_E_BAD_DIM, "vq handles descriptor lengths up to 256%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  double* code_sqr = nullptr;
  if (int rc = vq_static_scratch(&code_sqr)) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // entries [0, n_codes) are rewritten here before the arg-min kernel (same stream) reads them, so whatever an earlier
  // call left in the shared array does not matter and nothing has to be cleared
  b200m::vq_code_sqr_kernel<<<(n_codes + 127) / 128, 128, 0, st>>>(codebook, dim, n_codes, code_sqr);
  CUDA_TRY(cudaGetLastError());
  const size_t smem = sizeof(double) * b200m::kVqObsPerBlock * dim;
  CUDA_TRY(ensure_smem_attr(b200m::vq_argmin_kernel, 65536));
  const int64_t blocks = (n_obs + b200m::kVqObsPerBlock - 1) / b200m::kVqObsPerBlock;
  b200m::vq_argmin_kernel<<<static_cast<unsigned>(blocks), b200m::kVqThreads, smem, st>>>(
      obs, n_obs, dim, codebook, code_sqr, n_codes, out_code, out_dist);
  CUDA_TRY(cudaGetLastError());
  g_prof.launches.fetch_add(2);
  return B200M_OK;
}

int b200m_compact_matches(const int32_t* query_idx, const int32_t* train_idx, const float* dist,
                          const int32_t* count, const int64_t* offsets, int64_t n_pairs, int64_t cap,
                          int32_t* flat_query_idx, int32_t* flat_train_idx, float* flat_dist, void* stream) {
  if (n_pairs == 0) return B200M_OK;
  if (!query_idx || !train_idx || !dist || !count || !offsets || !flat_query_idx || !flat_train_idx || !flat_dist ||
      n_pairs < 0 || cap <= 0)
    return fail(B200M_E_BAD_ARG, "null argument / bad sizes%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  b200m::compact_kernel<<<static_cast<unsigned>(n_pairs), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      query_idx, train_idx, dist, count, offsets, cap, flat_query_idx, flat_train_idx, flat_dist);
  CUDA_TRY(cudaGetLastError());
  g_prof.launches.fetch_add(1);
  return B200M_OK;
}

int b200m_profile_read(double* first_sweep_ms, double* verify_sweep_ms, int64_t* first_sweeps, int64_t* launches) {
  std::lock_guard<std::mutex> lk(g_prof.mu);
  double a = 0.0, b = 0.0;
  for (auto& pr : g_prof.first) {
    float ms = 0.f;
    CUDA_TRY(cudaEventSynchronize(g_prof.pool[pr.second]));
    CUDA_TRY(cudaEventElapsedTime(&ms, g_prof.pool[pr.first], g_prof.pool[pr.second]));
    a += ms;
  }
  for (auto& pr : g_prof.verify) {
    float ms = 0.f;
    CUDA_TRY(cudaEventSynchronize(g_prof.pool[pr.second]));
    CUDA_TRY(cudaEventElapsedTime(&ms, g_prof.pool[pr.first], g_prof.pool[pr.second]));
    b += ms;
  }
  if (first_sweep_ms) *first_sweep_ms = a;
  if (verify_sweep_ms) *verify_sweep_ms = b;
  if (first_sweeps) *first_sweeps = static_cast<int64_t>(g_prof.first.size());
  if (launches) *launches = g_prof.launches.load();
  g_prof.first.clear();
  g_prof.verify.clear();
  g_prof.fallback.clear();
  g_prof.refine.clear();
  g_prof.select.clear();
  g_prof.emit.clear();
  g_prof.used = 0;
  g_prof.dropped = 0;
  g_prof.launches = 0;
  return B200M_OK;
}

int b200m_match_stats_read(const void* workspace, int64_t* out_host, int n, void* stream) {
  if (!workspace || !out_host || n < 6) return fail(B200M_E_BAD_ARG, "null workspace / out_host must hold 6 int64%s");
  Workspace w;
  carve(&w, const_cast<void*>(workspace), 1, B200M_TILE_T);
  unsigned long long host[b200m::kStatSlots * b200m::kStatFields];
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  CUDA_TRY(cudaMemcpyAsync(host, w.stats, sizeof(host), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  for (int f = 0; f < 6; ++f) {
    unsigned long long s = 0;
    for (int i = 0; i < b200m::kStatSlots; ++i) s += host[i * b200m::kStatFields + f];
    out_host[f] = static_cast<int64_t>(s);
  }
  return B200M_OK;
}

int b200m_profile_detail(double* out_ms, int n) {
  if (!out_ms || n < 6) return fail(B200M_E_BAD_ARG, "out_ms must hold 6 doubles%s");
  std::lock_guard<std::mutex> lk(g_prof.mu);
  std::vector<std::pair<int, int>>* lists[6] = {&g_prof.first, &g_prof.verify, &g_prof.fallback,
                                                &g_prof.refine, &g_prof.select, &g_prof.emit};
  for (int i = 0; i < 6; ++i) {
    double a = 0.0;
    for (auto& pr : *lists[i]) {
      float ms = 0.f;
      CUDA_TRY(cudaEventSynchronize(g_prof.pool[pr.second]));
      CUDA_TRY(cudaEventElapsedTime(&ms, g_prof.pool[pr.first], g_prof.pool[pr.second]));
      a += ms;
    }
    out_ms[i] = a;
  }
  return B200M_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ diagnostics
namespace b200m {

struct DebugSmem {
  uint8_t a[kMBytes];
  uint8_t b[kBBytes];
  uint64_t full, done;
  uint32_t tmem_base;
};

__global__ void __launch_bounds__(128, 1)
debug_dot_tile_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                      int32_t* __restrict__ out) {
  extern __shared__ uint8_t smem_raw[];
  DebugSmem& sm = *reinterpret_cast<DebugSmem*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const uint32_t warp = threadIdx.x >> 5, lane = lane_id();
  if (threadIdx.x == 0) {
    mbar_init(&sm.full, 1);
    mbar_init(&sm.done, 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<256>(&sm.tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = sm.tmem_base;
  if (threadIdx.x == 0) {
    mbar_arrive_expect_tx(&sm.full, kMBytes + kBBytes);
    tma_load_2d(sm.a, &tmap_a, 0, 0, &sm.full);
    tma_load_2d(sm.b, &tmap_b, 0, 0, &sm.full);
    tma_load_2d(sm.b + kBBytes / 2, &tmap_b, 0, 128, &sm.full);
    mbar_wait(&sm.full, 0);
    tc_fence_after();
    const uint64_t adesc = make_desc_k128(smem_u32(sm.a));
    const uint64_t bdesc = make_desc_k128(smem_u32(sm.b));
    constexpr uint32_t idesc = make_idesc_u8(kTM, kTT);
#pragma unroll
    for (int k = 0; k < kD / 32; ++k) mma_i8_ss(tmem_base, adesc + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
    tc_commit(&sm.done);
  }
  __syncwarp();
  mbar_wait(&sm.done, 0);
  tc_fence_after();
  const int row = warp * 32 + lane;
  for (int c = 0; c < 8; ++c) {
    int v[32];
    tmem_ld_32x32(tmem_base + ((warp * 32u) << 16) + c * 32, v);
    tmem_ld_wait();
#pragma unroll
    for (int j = 0; j < 32; ++j) out[row * kTT + c * 32 + j] = v[j];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<256>(tmem_base);
  }
}

}  // namespace b200m

extern "C" int b200m_debug_dot_tile_u8(const uint8_t* a_rows, const uint8_t* b_rows, int32_t* out, void* stream) {
  if (!a_rows || !b_rows || !out) return fail(B200M_E_BAD_ARG, "null argument%s");
  if (device_info().ok != 1) return fail(B200M_E_NO_DEVICE, "no sm_100 CUDA device available (no CPU fallback)%s");
  CUtensorMap ta, tb;
  int rc = make_row_tmap(&ta, a_rows, 128);
  if (rc) return rc;
  rc = make_row_tmap(&tb, b_rows, 256);
  if (rc) return rc;
  const size_t smem = sizeof(b200m::DebugSmem) + 1024;
  CUDA_TRY(cudaFuncSetAttribute(b200m::debug_dot_tile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                static_cast<int>(smem)));
  b200m::debug_dot_tile_kernel<<<1, 128, smem, static_cast<cudaStream_t>(stream)>>>(ta, tb, out);
  CUDA_TRY(cudaGetLastError());
  return B200M_OK;
}
