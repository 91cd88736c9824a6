// knnMatch(k = 2) on the chunk pipeline: from the three smallest chunk bounds of every query row
// (chunk_sweep256_kernel<.., 64, kTop3>: L1, g1, L2, g2, L3 on the norm-sorted bank) to the exact two nearest
// neighbours with cv2's stable order (lowest train index on equal distances).
//
// Every train column outside chunks g1 / g2 scores >= L3. So: evaluate the 2 x 64 columns of g1 and g2 exactly,
// take the two best by (score, original train row); if the second scores < L3 both are the row's exact two
// nearest neighbours, otherwise (an outside column may tie or win) the row goes to the exact top-2 sweep over
// gathered rows (sweep_kernel<kTop2G>). Evaluating 128 columns for EVERY query row one row at a time would pull
// 16 KB per row through L2 (68 GB per batch of 256 pairs x 16,384 rows); instead the (row, chunk) tasks of a
// pair are bucketed by chunk, a CTA stages one chunk in shared memory and runs all of its tasks against it
// (descriptor traffic ~8 MB per pair):
//   knn_bucket_kernel   per pair: histogram of the tasks over chunks, exclusive scan, task lists per chunk
//   knn_eval_kernel     per (pair, chunk): dp4a dot products of every task row with the chunk's 64 columns,
//                       packed key (score << 6 | rank of the column's original index inside the chunk),
//                       two REDUX.MIN -> best and runner-up of the task
//   knn_combine_kernel  per query row: merge its two tasks, test against L3, write (d1^2, i1, d2^2, i2) in
//                       original row order or queue the row for the fallback sweep
// Stands in for cv2.BFMatcher(NORM_L2).knnMatch(query, train, k=2) (the call BASELINE.json names).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "match_refine.cuh"
#include "ptx_sm100.cuh"

namespace b200m {

constexpr int kKnnChunk = 64;            // columns per chunk (chunk sweep option chunk = 64)
constexpr int kKnnMaxChunks = 4096;      // k_stride <= 262,144 (chunk ids travel as 16 bits, the histogram lives in shared memory)
constexpr int kKnnScoreOff = 1 << 25;    // scores |col|^2 - 2 row.col lie in (-2^25, 2^23): offset keys are positive 26-bit numbers
constexpr int kKnnEvalThreads = 256;

struct KnnParams {
  const uint8_t* bank;          // original order (fallback rows are gathered from here)
  const uint8_t* sbank;         // norm-sorted copy: query rows and train chunks are read from here
  const int32_t* perm;          // sorted position -> original row
  const int32_t* snorm;         // sorted norms
  const int32_t* k_per_image;
  const int32_t* pairs;
  int64_t k_stride;
  int32_t n_img;
  int32_t n_chunks;             // k_stride / 64
  const int4* rowres_s;         // chunk sweep output per (pair, SORTED query row): (L1, g1 | g2 << 16, L2, L3)
  int32_t* boff;                // [n_pairs][n_chunks + 1] first task of every chunk
  int32_t* tasks;               // [n_pairs][2 * k_stride] sorted query row | slot << 30
  int4* taskres;                // [n_pairs][k_stride][2] (s1, i1, s2, i2) of the task: scores without the row norm, original train rows
  int4* rowres_o;               // out per (pair, ORIGINAL query row): (d1^2, i1, d2^2, i2)
  int32_t* fb_count;            // [n_pairs] rows sent to the fallback sweep (zeroed by the caller)
  int32_t* fb_total;            // [1]
  int32_t* fb_q;                // [n_pairs][k_stride] their original query rows
  uint8_t* fb_desc;             // [n_pairs][k_stride][128] their descriptors
  unsigned long long* stats;    // call counters (match_common.cuh): 0 rows, 1 settled by the two chunks, 2 fallback
};

// One CTA per pair: tasks (row, chunk g1) and (row, chunk g2) bucketed by chunk.
__global__ void __launch_bounds__(1024)
knn_bucket_kernel(const KnnParams p) {
  __shared__ int hist[kKnnMaxChunks];
  __shared__ int warp_tot[32];
  const int pair = blockIdx.x;
  const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
  const int kq = pi.kt > 0 ? pi.kq : 0;
  const int64_t base = static_cast<int64_t>(pair) * p.k_stride;
  for (int g = threadIdx.x; g < p.n_chunks; g += blockDim.x) hist[g] = 0;
  __syncthreads();
  for (int pos = threadIdx.x; pos < kq; pos += blockDim.x) {
    const int4 v = p.rowres_s[base + pos];
    atomicAdd(&hist[v.y & 0xffff], 1);
    if (v.z < kPadScore) atomicAdd(&hist[v.y >> 16], 1);     // a second chunk with real columns exists
  }
  __syncthreads();
  // exclusive scan of hist[0 .. n_chunks): every thread owns four consecutive entries
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int v4[4], sum = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int g = threadIdx.x * 4 + i;
    v4[i] = g < p.n_chunks ? hist[g] : 0;
    sum += v4[i];
  }
  int inc = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) warp_tot[warp] = inc;
  __syncthreads();
  int before = 0;
  for (int w = 0; w < warp; ++w) before += warp_tot[w];
  int off = before + inc - sum;
  int32_t* boff = p.boff + static_cast<int64_t>(pair) * (p.n_chunks + 1);
  __syncthreads();                                           // every thread has read its hist entries
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int g = threadIdx.x * 4 + i;
    if (g < p.n_chunks) {
      boff[g] = off;
      hist[g] = off;                                         // becomes the bucket's write cursor
      off += v4[i];
    }
    if (g == p.n_chunks - 1) boff[p.n_chunks] = off;
  }
  __syncthreads();
  int32_t* tasks = p.tasks + 2 * base;
  for (int pos = threadIdx.x; pos < kq; pos += blockDim.x) {
    const int4 v = p.rowres_s[base + pos];
    tasks[atomicAdd(&hist[v.y & 0xffff], 1)] = pos;
    if (v.z < kPadScore) tasks[atomicAdd(&hist[v.y >> 16], 1)] = pos | (1 << 30);
  }
}

// One CTA per (chunk, pair): the chunk's 64 columns staged in shared memory, one warp per task.
__global__ void __launch_bounds__(kKnnEvalThreads)
knn_eval_kernel(const KnnParams p) {
  __shared__ uint4 colT[8][kKnnChunk];       // [16-byte word][column]: lanes read consecutive columns, conflict-free
  __shared__ int s_norm[kKnnChunk], s_orig[kKnnChunk], s_rank[kKnnChunk], s_byrank[kKnnChunk];
  const int g = blockIdx.x, pair = blockIdx.y;
  const int32_t* boff = p.boff + static_cast<int64_t>(pair) * (p.n_chunks + 1);
  const int t0 = boff[g], n_tasks = boff[g + 1] - t0;
  if (n_tasks <= 0) return;
  const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
  const int64_t qbase = static_cast<int64_t>(pi.qi) * p.k_stride, tbase = static_cast<int64_t>(pi.ti) * p.k_stride;
  const int c0 = g * kKnnChunk;
  const int n_valid = min(pi.kt - c0, kKnnChunk);            // real columns of this chunk (>= 1 for a chunk with tasks)
  for (int i = threadIdx.x; i < 8 * kKnnChunk; i += kKnnEvalThreads) {
    const int col = i >> 3, w = i & 7;
    colT[w][col] = __ldg(reinterpret_cast<const uint4*>(p.sbank + (tbase + c0 + col) * 128) + w);
  }
  if (threadIdx.x < kKnnChunk) {
    const int c = threadIdx.x;
    s_norm[c] = p.snorm[tbase + c0 + c];
    s_orig[c] = c < n_valid ? p.perm[tbase + c0 + c] : 0x7fffffff - (kKnnChunk - c);   // padding ranks last, distinct
  }
  __syncthreads();
  if (threadIdx.x < kKnnChunk) {
    const int c = threadIdx.x, mine = s_orig[c];
    int rank = 0;
#pragma unroll 8
    for (int o = 0; o < kKnnChunk; ++o) rank += s_orig[o] < mine;
    s_rank[c] = rank;
    s_byrank[rank] = mine;
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int ca = lane, cb = lane + 32;
  const int na = s_norm[ca] + kKnnScoreOff, nb = s_norm[cb] + kKnnScoreOff;
  const unsigned ra = s_rank[ca], rb = s_rank[cb];
  const bool va = ca < n_valid, vb = cb < n_valid;
  const int32_t* tasks = p.tasks + 2 * static_cast<int64_t>(pair) * p.k_stride + t0;
  for (int ti = warp; ti < n_tasks; ti += kKnnEvalThreads / 32) {
    const int task = __ldg(tasks + ti);
    const int pos = task & 0x3fffffff, slot = task >> 30;
    const uint4* rowp = reinterpret_cast<const uint4*>(p.sbank + (qbase + pos) * 128);
    uint32_t da = 0, db = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
      const uint4 r = __ldg(rowp + w);                       // the same 16 bytes for every lane: one broadcast transaction
      const uint4 a = colT[w][ca], b = colT[w][cb];
      da = __dp4a(r.x, a.x, da); da = __dp4a(r.y, a.y, da); da = __dp4a(r.z, a.z, da); da = __dp4a(r.w, a.w, da);
      db = __dp4a(r.x, b.x, db); db = __dp4a(r.y, b.y, db); db = __dp4a(r.z, b.z, db); db = __dp4a(r.w, b.w, db);
    }
    // (score + 2^25) << 6 | rank of the original train row inside the chunk: unsigned order == (score, original row)
    const unsigned ka = va ? (static_cast<unsigned>(na - 2 * static_cast<int>(da)) << 6) | ra : 0xffffffffu;
    const unsigned kb = vb ? (static_cast<unsigned>(nb - 2 * static_cast<int>(db)) << 6) | rb : 0xffffffffu;
    const unsigned lo = min(ka, kb), hi = max(ka, kb);
    const unsigned m1 = __reduce_min_sync(0xffffffffu, lo);
    const unsigned m2 = __reduce_min_sync(0xffffffffu, lo == m1 ? hi : lo);   // keys are distinct: one lane gives up its best
    if (lane == 0) {
      int4 out;
      out.x = static_cast<int>(m1 >> 6) - kKnnScoreOff;
      out.y = s_byrank[m1 & 63];
      out.z = m2 != 0xffffffffu ? static_cast<int>(m2 >> 6) - kKnnScoreOff : kBigScore;
      out.w = m2 != 0xffffffffu ? s_byrank[m2 & 63] : -1;
      p.taskres[(static_cast<int64_t>(pair) * p.k_stride + pos) * 2 + slot] = out;
    }
  }
}

// Tensor-core version of knn_eval_kernel (the one the library launches): one CTA of 128 threads per (chunk, pair),
// several CTAs per SM. The chunk's 64 columns are staged once as a K-major SWIZZLE_128B operand (the layout TMA
// would produce: 16-byte word w of row r at r * 128 + ((w ^ (r & 7)) << 4)); the task rows are GATHERED by index
// into the same layout, 128 at a time, with plain 16-byte loads / stores (eight lanes per row: whole 128-byte lines
// both ways), made visible to the tensor core with fence.proxy.async; one elected thread issues the four
// tcgen05.mma.kind::i8 of a 128 x 64 x 128 tile into 64 TMEM columns and every thread then reduces ITS task's 64
// dot products to the two smallest keys (score << 6 | rank of the original train row) -- no cross-lane traffic.
constexpr int kKnnTcThreads = 128;

__global__ void __launch_bounds__(kKnnTcThreads)
knn_eval_tc_kernel(const KnnParams p) {
  __shared__ alignas(1024) uint8_t a_tile[128 * 128];
  __shared__ alignas(1024) uint8_t b_tile[kKnnChunk * 128];
  __shared__ alignas(16) uint32_t s_nkey[kKnnChunk];           // ((norm + 2^25) << 6) | rank; >= 0xffffffc0 for padding
  __shared__ int s_orig[kKnnChunk], s_byrank[kKnnChunk], s_task[128];
  __shared__ alignas(8) uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int g = blockIdx.x, pair = blockIdx.y;
  const int32_t* boff = p.boff + static_cast<int64_t>(pair) * (p.n_chunks + 1);
  const int t0 = __ldg(boff + g), n_tasks = __ldg(boff + g + 1) - t0;
  if (n_tasks <= 0) return;
  const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
  const int64_t qbase = static_cast<int64_t>(pi.qi) * p.k_stride, tbase = static_cast<int64_t>(pi.ti) * p.k_stride;
  const int c0 = g * kKnnChunk;
  const int n_valid = min(pi.kt - c0, kKnnChunk);              // real columns of this chunk (>= 1 for a chunk with tasks)
  const int t = threadIdx.x, warp = t >> 5;
  if (warp == 0) tmem_alloc<64>(&tmem_slot);
  if (t == 32) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  // the chunk: 64 rows x 128 B, padding columns as zeros (their dot products are 0, their keys stay >= 0xffffffc0)
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    const int idx = it * kKnnTcThreads + t, r = idx >> 3, w = idx & 7;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r < n_valid) v = __ldg(reinterpret_cast<const uint4*>(p.sbank + (tbase + c0 + r) * 128) + w);
    *reinterpret_cast<uint4*>(b_tile + r * 128 + ((w ^ (r & 7)) << 4)) = v;
  }
  if (t < kKnnChunk) s_orig[t] = t < n_valid ? __ldg(p.perm + tbase + c0 + t) : 0x7fffffff - (kKnnChunk - t);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  if (t < kKnnChunk) {
    const int mine = s_orig[t];
    int rank = 0;
#pragma unroll 8
    for (int o = 0; o < kKnnChunk; ++o) rank += s_orig[o] < mine;
    s_byrank[rank] = mine;
    s_nkey[t] = t < n_valid ? (static_cast<uint32_t>(__ldg(p.snorm + tbase + c0 + t) + kKnnScoreOff) << 6) | rank
                            : 0xffffffc0u | rank;
  }
  const int32_t* tasks = p.tasks + 2 * static_cast<int64_t>(pair) * p.k_stride + t0;
  constexpr uint32_t idesc = make_idesc_u8(128, kKnnChunk);
  const uint64_t adesc = make_desc_k128(smem_u32(a_tile)), bdesc = make_desc_k128(smem_u32(b_tile));
  uint32_t phase = 0;
  for (int base = 0; base < n_tasks; base += 128) {
    const int mytask = base + t < n_tasks ? __ldg(tasks + base + t) : -1;
    s_task[t] = mytask;
    __syncthreads();                                           // also: s_nkey / s_byrank written, previous tile drained
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int idx = it * kKnnTcThreads + t, r = idx >> 3, w = idx & 7;
      const int task = s_task[r];
      if (task >= 0) {                                         // rows past the last task keep stale bytes: never read back
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(p.sbank + (qbase + (task & 0x3fffffff)) * 128) + w);
        *reinterpret_cast<uint4*>(a_tile + r * 128 + ((w ^ (r & 7)) << 4)) = v;
      }
    }
    fence_proxy_async_smem();                                  // generic-proxy stores -> visible to the tensor core
    tc_fence_before();
    __syncthreads();
    if (t == 0) {
      tc_fence_after();
#pragma unroll
      for (int k = 0; k < 4; ++k) mma_i8_ss(tmem, adesc + 2 * k, bdesc + 2 * k, idesc, k > 0 ? 1u : 0u);
      tc_commit(&bar);
    }
    mbar_wait(&bar, phase);
    phase ^= 1;
    tc_fence_after();
    uint32_t m1 = 0xffffffffu, m2 = 0xffffffffu;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      int v[32];
      tmem_ld_32x32(tmem + (static_cast<uint32_t>(warp * 32) << 16) + h * 32, v);
      tmem_ld_fence(v);
#pragma unroll
      for (int j4 = 0; j4 < 8; ++j4) {
        const uint4 nk = *reinterpret_cast<const uint4*>(s_nkey + h * 32 + j4 * 4);
        const uint32_t n4[4] = {nk.x, nk.y, nk.z, nk.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t key = n4[j] - (static_cast<uint32_t>(v[j4 * 4 + j]) << 7);   // (norm - 2 dot + 2^25) << 6 | rank
          const uint32_t hi = max(m1, key);
          m1 = min(m1, key);
          m2 = min(m2, hi);
        }
      }
    }
    if (mytask >= 0) {
      const int pos = mytask & 0x3fffffff, slot = mytask >> 30;
      const bool second = m2 < 0xffffffc0u;
      int4 out;
      out.x = static_cast<int>(m1 >> 6) - kKnnScoreOff;
      out.y = s_byrank[m1 & 63];
      out.z = second ? static_cast<int>(m2 >> 6) - kKnnScoreOff : kBigScore;
      out.w = second ? s_byrank[m2 & 63] : -1;
      p.taskres[(static_cast<int64_t>(pair) * p.k_stride + pos) * 2 + slot] = out;
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 0) tmem_dealloc<64>(tmem);
}

// Per (pair, sorted query row): merge the row's two tasks, settle or fall back.
__global__ void __launch_bounds__(256)
knn_combine_kernel(const KnnParams p) {
  const int pair = blockIdx.y;
  const PairInfo pi = load_pair(p.pairs, p.k_per_image, pair, p.n_img, p.k_stride);
  const int pos = blockIdx.x * blockDim.x + threadIdx.x;
  if (pos >= pi.kq || pi.kt <= 0) return;
  const int64_t base = static_cast<int64_t>(pair) * p.k_stride;
  const int64_t qbase = static_cast<int64_t>(pi.qi) * p.k_stride;
  const int4 v = p.rowres_s[base + pos];
  const int4 a = p.taskres[(base + pos) * 2];
  int s1 = a.x, i1 = a.y, s2 = a.z, i2 = a.w;
  if (v.z < kPadScore) {
    const int4 b = p.taskres[(base + pos) * 2 + 1];
    // two sorted pairs -> the two smallest by (score, original train row)
    if (lex_less(b.x, b.y, s1, i1)) {
      if (b.w >= 0 && lex_less(b.z, b.w, s1, i1)) { s2 = b.z; i2 = b.w; } else { s2 = s1; i2 = i1; }
      s1 = b.x; i1 = b.y;
    } else if (i2 < 0 || lex_less(b.x, b.y, s2, i2)) {
      s2 = b.x; i2 = b.y;
    }
  }
  const int row = p.perm[qbase + pos];                       // original query row
  const int nq = p.snorm[qbase + pos];
  // every column outside the two chunks scores >= L3: with s2 < L3 nothing outside can tie or beat the runner-up
  // (a row with a single train column has no runner-up and no outside column: L3 stays at its initial value)
  const bool settled = i2 < 0 ? v.w >= kPadScore : s2 < v.w;
  {
    const unsigned act = __activemask(), ok = __ballot_sync(act, settled);
    if ((threadIdx.x & 31) == __ffs(act) - 1) {
      stat_add(p.stats, 0, __popc(act));
      stat_add(p.stats, 1, __popc(ok));
      stat_add(p.stats, 2, __popc(act & ~ok));
    }
  }
  if (settled) {
    p.rowres_o[base + row] = make_int4(s1 + nq, i1, i2 >= 0 ? s2 + nq : 0, i2);
  } else {
    const int slot = atomicAdd(p.fb_count + pair, 1);
    if (slot == 0) atomicOr(p.fb_total, 1);
    p.fb_q[base + slot] = row;
    const uint4* src = reinterpret_cast<const uint4*>(p.bank + (qbase + row) * 128);
    uint4* dst = reinterpret_cast<uint4*>(p.fb_desc + (base + slot) * 128);
#pragma unroll
    for (int i = 0; i < 8; ++i) dst[i] = src[i];
  }
}

}  // namespace b200m
