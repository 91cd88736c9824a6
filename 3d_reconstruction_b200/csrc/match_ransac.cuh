// Batched geometric verification of match lists: essential matrix by 5-point RANSAC + pose recovery.
//
// The step right after the matcher in the reference (matching.py:133-142):
//     E, mask = cv2.findEssentialMat(k0, k1, K, method=cv2.RANSAC, prob=0.999, threshold=1)
//     _, R, t, inl = cv2.recoverPose(E, k0[mask], k1[mask], K)
// once per visited edge on the host. Here every pair of a batch is verified in one launch: one thread per
// RANSAC hypothesis (minimal 5-point solve in fp64, up to ten candidate matrices, each scored against all
// matches of the pair with the Sampson distance in fp32), block-wide arg-best, then one CTA per pair marks
// the inliers, decomposes the winner and counts the cheirality inliers of its four (R, t) candidates.
// RANSAC is randomised: parity with cv2 is statistical (inlier sets, pose), see tests/test_gpu_geometry.py.
//
// The minimal solver is Nister's five-point algorithm ("An efficient solution to the five-point relative
// pose problem", PAMI 2004), restated from the paper: null space of the 5 x 9 epipolar system, the ten cubic
// constraints det(E) = 0 and 2 E E^T E - tr(E E^T) E = 0 in the monomials of (x, y, z), Gauss-Jordan
// elimination, a 3 x 3 polynomial matrix in z whose determinant is a degree-10 polynomial, its real roots,
// back-substitution.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200m {
namespace geo {

// monomial bookkeeping. linear: [x, y, z, 1]; quadratic: [x2, xy, xz, x, y2, yz, y, z2, z, 1];
// cubic (Nister's elimination order): [x3, y3, x2y, xy2, x2z, x2, y2z, y2, xyz, xy | xz2, xz, x, yz2, yz, y, z3, z2, z, 1]
__constant__ int8_t kQ2[4][4] = {{0, 1, 2, 3}, {1, 4, 5, 6}, {2, 5, 7, 8}, {3, 6, 8, 9}};
__constant__ int8_t kQ3[10][4] = {{0, 2, 4, 5},   {2, 3, 8, 9},   {4, 8, 10, 11},   {5, 9, 11, 12},   {3, 1, 6, 7},
                                  {8, 6, 13, 14}, {9, 7, 14, 15}, {10, 13, 16, 17}, {11, 14, 17, 18}, {12, 15, 18, 19}};

// q += s * (a * b), a and b linear
__device__ __forceinline__ void mul_ll(const double* a, const double* b, double* q, double s) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) q[kQ2[i][j]] += s * a[i] * b[j];
}
// c += s * (q * l), q quadratic, l linear
__device__ __forceinline__ void mul_ql(const double* q, const double* l, double* c, double s) {
#pragma unroll
  for (int i = 0; i < 10; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) c[kQ3[i][j]] += s * q[i] * l[j];
}

// r (degree da + db) = a (degree da) * b (degree db), coefficients by ascending power
__device__ __forceinline__ void poly_mul(const double* a, int da, const double* b, int db, double* r) {
  for (int i = 0; i <= da + db; ++i) r[i] = 0.0;
  for (int i = 0; i <= da; ++i)
    for (int j = 0; j <= db; ++j) r[i + j] += a[i] * b[j];
}
__device__ __forceinline__ double poly_val(const double* a, int d, double z) {
  double r = a[d];
  for (int i = d - 1; i >= 0; --i) r = r * z + a[i];
  return r;
}

// Null space of the 5 x 9 epipolar system by Gauss-Jordan elimination with full pivoting.
// x0, x1: five normalised image points each; basis: four 9-vectors (row-major 3 x 3). false = degenerate sample.
__device__ inline bool epipolar_null_space(const double (*x0)[2], const double (*x1)[2], double (*basis)[9]) {
  double M[5][9];
  for (int i = 0; i < 5; ++i) {
    const double a = x0[i][0], b = x0[i][1], c = x1[i][0], d = x1[i][1];
    M[i][0] = c * a; M[i][1] = c * b; M[i][2] = c;
    M[i][3] = d * a; M[i][4] = d * b; M[i][5] = d;
    M[i][6] = a;     M[i][7] = b;     M[i][8] = 1.0;
  }
  int pivcol[5];
  bool used[9] = {false, false, false, false, false, false, false, false, false};
  for (int r = 0; r < 5; ++r) {
    int br = r, bc = -1;
    double best = 0.0;
    for (int i = r; i < 5; ++i)
      for (int c = 0; c < 9; ++c)
        if (!used[c] && fabs(M[i][c]) > best) { best = fabs(M[i][c]); br = i; bc = c; }
    if (bc < 0 || best < 1e-12) return false;
    if (br != r)
      for (int c = 0; c < 9; ++c) { const double t = M[r][c]; M[r][c] = M[br][c]; M[br][c] = t; }
    used[bc] = true;
    pivcol[r] = bc;
    const double inv = 1.0 / M[r][bc];
    for (int c = 0; c < 9; ++c) M[r][c] *= inv;
    for (int i = 0; i < 5; ++i) {
      if (i == r) continue;
      const double f = M[i][bc];
      if (f != 0.0)
        for (int c = 0; c < 9; ++c) M[i][c] -= f * M[r][c];
    }
  }
  int k = 0;
  for (int f = 0; f < 9; ++f) {
    if (used[f]) continue;
    for (int c = 0; c < 9; ++c) basis[k][c] = 0.0;
    basis[k][f] = 1.0;
    for (int r = 0; r < 5; ++r) basis[k][pivcol[r]] = -M[r][f];
    ++k;
  }
  return k == 4;
}

// All real roots of a real polynomial of degree 10 (coefficients by ascending power, c[10] != 0):
// Durand-Kerner iteration on the ten complex roots, then a Newton polish of the nearly real ones.
__device__ inline int real_roots_deg10(const double* c, double* roots) {
  double a[11];
  const double inv = 1.0 / c[10];
  double bound = 0.0;
  for (int i = 0; i < 10; ++i) { a[i] = c[i] * inv; bound = fmax(bound, fabs(a[i])); }
  a[10] = 1.0;
  const double radius = 1.0 + bound;            // Cauchy bound
  double zr[10], zi[10];
  // start on a circle well inside the bound (roots of these polynomials are O(1) for normalised points)
  const double r0 = fmin(radius, 1.5 + 0.05 * bound);
  for (int k = 0; k < 10; ++k) {
    double s, co;
    sincos(0.4 + 6.283185307179586 * k / 10.0, &s, &co);
    zr[k] = r0 * co;
    zi[k] = r0 * s;
  }
  for (int it = 0; it < 80; ++it) {
    double moved = 0.0;
    for (int k = 0; k < 10; ++k) {
      // p(z_k) by Horner in complex arithmetic
      double pr = 1.0, pi = 0.0;
      for (int i = 9; i >= 0; --i) {
        const double tr = pr * zr[k] - pi * zi[k] + a[i];
        pi = pr * zi[k] + pi * zr[k];
        pr = tr;
      }
      double qr = 1.0, qi = 0.0;      // prod_{j != k} (z_k - z_j)
      for (int j = 0; j < 10; ++j) {
        if (j == k) continue;
        const double dr = zr[k] - zr[j], di = zi[k] - zi[j];
        const double tr = qr * dr - qi * di;
        qi = qr * di + qi * dr;
        qr = tr;
      }
      const double den = qr * qr + qi * qi;
      if (den < 1e-300) continue;
      const double wr = (pr * qr + pi * qi) / den, wi = (pi * qr - pr * qi) / den;
      zr[k] -= wr;
      zi[k] -= wi;
      moved = fmax(moved, fabs(wr) + fabs(wi));
    }
    if (moved < 1e-13) break;
  }
  int n = 0;
  for (int k = 0; k < 10; ++k) {
    if (fabs(zi[k]) > 1e-5 * (1.0 + fabs(zr[k]))) continue;
    double z = zr[k];
    for (int it = 0; it < 3; ++it) {            // Newton polish on the real polynomial
      double p = 1.0, dp = 0.0;
      for (int i = 9; i >= 0; --i) { dp = dp * z + p; p = p * z + a[i]; }
      if (fabs(dp) < 1e-300) break;
      z -= p / dp;
    }
    if (!isfinite(z)) continue;
    bool dup = false;                           // a conjugate pair that collapsed onto the axis counts once
    for (int j = 0; j < n; ++j) dup |= fabs(roots[j] - z) < 1e-9 * (1.0 + fabs(z));
    if (!dup) roots[n++] = z;
  }
  return n;
}

// Five-point minimal solver. x0[i] <-> x1[i] normalised image points with x1^T E x0 = 0.
// Writes up to ten candidate matrices (row-major, unit Frobenius norm), returns how many.
__device__ inline int five_point(const double (*x0)[2], const double (*x1)[2], double (*Es)[9]) {
  double N[4][9];                                // X, Y, Z, W: E = x X + y Y + z Z + W
  if (!epipolar_null_space(x0, x1, N)) return 0;
  double E[9][4];                                // entry e of E as a linear polynomial in (x, y, z, 1)
  for (int e = 0; e < 9; ++e)
    for (int k = 0; k < 4; ++k) E[e][k] = N[k][e];

  double A[10][20];
  for (int r = 0; r < 10; ++r)
    for (int c = 0; c < 20; ++c) A[r][c] = 0.0;
  {
    // EEt (symmetric, quadratic) and L = EEt - 1/2 trace(EEt) I
    double EEt[3][3][10];
    for (int i = 0; i < 3; ++i)
      for (int j = i; j < 3; ++j) {
        for (int m = 0; m < 10; ++m) EEt[i][j][m] = 0.0;
        for (int k = 0; k < 3; ++k) mul_ll(E[3 * i + k], E[3 * j + k], EEt[i][j], 1.0);
      }
    for (int m = 0; m < 10; ++m) {
      const double half_tr = 0.5 * (EEt[0][0][m] + EEt[1][1][m] + EEt[2][2][m]);
      EEt[0][0][m] -= half_tr;
      EEt[1][1][m] -= half_tr;
      EEt[2][2][m] -= half_tr;
      EEt[1][0][m] = EEt[0][1][m];
      EEt[2][0][m] = EEt[0][2][m];
      EEt[2][1][m] = EEt[1][2][m];
    }
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        for (int k = 0; k < 3; ++k) mul_ql(EEt[i][k], E[3 * k + j], A[3 * i + j], 1.0);
    // det(E): cofactor expansion along the first row
    double q[10];
    for (int m = 0; m < 10; ++m) q[m] = 0.0;
    mul_ll(E[4], E[8], q, 1.0); mul_ll(E[5], E[7], q, -1.0);
    mul_ql(q, E[0], A[9], 1.0);
    for (int m = 0; m < 10; ++m) q[m] = 0.0;
    mul_ll(E[3], E[8], q, 1.0); mul_ll(E[5], E[6], q, -1.0);
    mul_ql(q, E[1], A[9], -1.0);
    for (int m = 0; m < 10; ++m) q[m] = 0.0;
    mul_ll(E[3], E[7], q, 1.0); mul_ll(E[4], E[6], q, -1.0);
    mul_ql(q, E[2], A[9], 1.0);
  }
  // Gauss-Jordan on the ten leading monomials (partial pivoting over the rows)
  for (int c = 0; c < 10; ++c) {
    int br = c;
    double best = fabs(A[c][c]);
    for (int r = c + 1; r < 10; ++r)
      if (fabs(A[r][c]) > best) { best = fabs(A[r][c]); br = r; }
    if (best < 1e-13) return 0;
    if (br != c)
      for (int k = c; k < 20; ++k) { const double t = A[c][k]; A[c][k] = A[br][k]; A[br][k] = t; }
    const double inv = 1.0 / A[c][c];
    for (int k = c; k < 20; ++k) A[c][k] *= inv;
    for (int r = 0; r < 10; ++r) {
      if (r == c) continue;
      const double f = A[r][c];
      if (f != 0.0)
        for (int k = c; k < 20; ++k) A[r][k] -= f * A[c][k];
    }
  }
  // rows 4..9 lead with x2z, x2, y2z, y2, xyz, xy; their tails live on [xz2, xz, x, yz2, yz, y, z3, z2, z, 1].
  // row(2k) - z * row(2k+1) has no leading term left: it reads  x * bx(z) + y * by(z) + b1(z) = 0.
  double bx[3][4], by[3][4], b1[3][5];
  for (int k = 0; k < 3; ++k) {
    const double* u = &A[4 + 2 * k][10];
    const double* v = &A[5 + 2 * k][10];
    bx[k][3] = -v[0]; bx[k][2] = u[0] - v[1]; bx[k][1] = u[1] - v[2]; bx[k][0] = u[2];
    by[k][3] = -v[3]; by[k][2] = u[3] - v[4]; by[k][1] = u[4] - v[5]; by[k][0] = u[5];
    b1[k][4] = -v[6]; b1[k][3] = u[6] - v[7]; b1[k][2] = u[7] - v[8]; b1[k][1] = u[8] - v[9]; b1[k][0] = u[9];
  }
  // det of the 3 x 3 polynomial matrix: degree 10 in z
  double det[11];
  for (int i = 0; i <= 10; ++i) det[i] = 0.0;
  {
    double t1[8], t2[8], t3[11];
    // bx[0] * (by[1] b1[2] - b1[1] by[2])
    poly_mul(by[1], 3, b1[2], 4, t1); poly_mul(b1[1], 4, by[2], 3, t2);
    for (int i = 0; i < 8; ++i) t1[i] -= t2[i];
    poly_mul(bx[0], 3, t1, 7, t3);
    for (int i = 0; i <= 10; ++i) det[i] += t3[i];
    // - by[0] * (bx[1] b1[2] - b1[1] bx[2])
    poly_mul(bx[1], 3, b1[2], 4, t1); poly_mul(b1[1], 4, bx[2], 3, t2);
    for (int i = 0; i < 8; ++i) t1[i] -= t2[i];
    poly_mul(by[0], 3, t1, 7, t3);
    for (int i = 0; i <= 10; ++i) det[i] -= t3[i];
    // + b1[0] * (bx[1] by[2] - by[1] bx[2])
    poly_mul(bx[1], 3, by[2], 3, t1); poly_mul(by[1], 3, bx[2], 3, t2);
    for (int i = 0; i < 7; ++i) t1[i] -= t2[i];
    poly_mul(b1[0], 4, t1, 6, t3);
    for (int i = 0; i <= 10; ++i) det[i] += t3[i];
  }
  double scale = 0.0;
  for (int i = 0; i <= 10; ++i) scale = fmax(scale, fabs(det[i]));
  if (!(scale > 0.0) || fabs(det[10]) < 1e-12 * scale) return 0;
  double roots[10];
  const int nr = real_roots_deg10(det, roots);
  int n = 0;
  for (int r = 0; r < nr; ++r) {
    const double z = roots[r];
    double B[3][3];
    for (int k = 0; k < 3; ++k) { B[k][0] = poly_val(bx[k], 3, z); B[k][1] = poly_val(by[k], 3, z); B[k][2] = poly_val(b1[k], 4, z); }
    // (x, y, 1) spans the null space of B(z): the cross product of two rows; take the best-conditioned pair
    double bestn[3] = {0, 0, 0};
    for (int p = 0; p < 3; ++p) {
      const double* u = B[p];
      const double* v = B[(p + 1) % 3];
      const double cx = u[1] * v[2] - u[2] * v[1], cy = u[2] * v[0] - u[0] * v[2], cz = u[0] * v[1] - u[1] * v[0];
      if (fabs(cz) > fabs(bestn[2])) { bestn[0] = cx; bestn[1] = cy; bestn[2] = cz; }
    }
    if (fabs(bestn[2]) < 1e-300) continue;
    const double x = bestn[0] / bestn[2], y = bestn[1] / bestn[2];
    double nrm = 0.0;
    for (int e = 0; e < 9; ++e) {
      const double v = x * N[0][e] + y * N[1][e] + z * N[2][e] + N[3][e];
      Es[n][e] = v;
      nrm += v * v;
    }
    if (!(nrm > 0.0) || !isfinite(nrm)) continue;
    const double inv = rsqrt(nrm);
    for (int e = 0; e < 9; ++e) Es[n][e] *= inv;
    ++n;
  }
  return n;
}

// ------------------------------------------------------------------------------------------------ RANSAC
constexpr int kRansacThreads = 128;     // hypotheses per CTA

struct RansacParams {
  const float4* matches;     // [total] (x0, y0, x1, y1): normalised image coordinates of every match of every pair
  const int64_t* offsets;    // [n_pairs + 1] extent of each pair's matches
  int32_t n_pairs;
  int32_t pair0;             // first pair of this launch (gridDim.y covers at most 65535 pairs)
  int32_t blocks_per_pair;   // hypotheses per pair = blocks_per_pair * kRansacThreads
  float thr2;                // squared Sampson threshold in normalised units: (threshold_px / focal)^2
  uint64_t seed;
  int32_t* blk_count;        // [n_pairs, blocks_per_pair] best inlier count of each CTA
  double* blk_E;             // [n_pairs, blocks_per_pair, 9] and its matrix
};

__device__ __forceinline__ uint64_t splitmix64(uint64_t& s) {
  uint64_t z = (s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}

// squared Sampson distance of a match to E (row-major), fp32
__device__ __forceinline__ float sampson2(const float* e, const float4 m) {
  const float a = fmaf(e[0], m.x, fmaf(e[1], m.y, e[2]));
  const float b = fmaf(e[3], m.x, fmaf(e[4], m.y, e[5]));
  const float c = fmaf(e[6], m.x, fmaf(e[7], m.y, e[8]));
  const float at = fmaf(e[0], m.z, fmaf(e[3], m.w, e[6]));
  const float bt = fmaf(e[1], m.z, fmaf(e[4], m.w, e[7]));
  const float num = fmaf(m.z, a, fmaf(m.w, b, c));
  return num * num / (a * a + b * b + at * at + bt * bt);
}
// the same test without the division: sampson2 < thr2
__device__ __forceinline__ bool sampson_inlier(const float* e, const float4 m, float thr2) {
  const float a = fmaf(e[0], m.x, fmaf(e[1], m.y, e[2]));
  const float b = fmaf(e[3], m.x, fmaf(e[4], m.y, e[5]));
  const float c = fmaf(e[6], m.x, fmaf(e[7], m.y, e[8]));
  const float at = fmaf(e[0], m.z, fmaf(e[3], m.w, e[6]));
  const float bt = fmaf(e[1], m.z, fmaf(e[4], m.w, e[7]));
  const float num = fmaf(m.z, a, fmaf(m.w, b, c));
  return num * num < thr2 * (a * a + b * b + at * at + bt * bt);
}

__global__ void __launch_bounds__(kRansacThreads)
ransac_kernel(const RansacParams p) {
  const int pair = p.pair0 + blockIdx.y;
  const int64_t o0 = p.offsets[pair];
  const int S = static_cast<int>(p.offsets[pair + 1] - o0);
  const float4* m = p.matches + o0;
  const int hyp = blockIdx.x * kRansacThreads + threadIdx.x;
  int best = -1;
  double bestE[9];
  for (int e = 0; e < 9; ++e) bestE[e] = 0.0;
  // best count any thread of this CTA has reached so far: a candidate that can no longer reach it stops
  // being scored (it could not win; candidates that can still TIE go on, so the winner does not depend on timing)
  __shared__ int s_blk_best;
  if (threadIdx.x == 0) s_blk_best = 0;
  __syncthreads();
  if (S >= 5) {
    uint64_t rng = p.seed ^ (static_cast<uint64_t>(pair) << 32) ^ static_cast<uint64_t>(hyp) * 0xD1342543DE82EF95ull;
    int idx[5];
    for (int k = 0; k < 5; ++k) {
      for (int tries = 0; tries < 64; ++tries) {
        idx[k] = static_cast<int>((splitmix64(rng) >> 32) * static_cast<uint64_t>(S) >> 32);
        bool dup = false;
        for (int j = 0; j < k; ++j) dup |= idx[j] == idx[k];
        if (!dup) break;
      }
    }
    double x0[5][2], x1[5][2];
    for (int k = 0; k < 5; ++k) {
      const float4 v = __ldg(m + idx[k]);
      x0[k][0] = v.x; x0[k][1] = v.y; x1[k][0] = v.z; x1[k][1] = v.w;
    }
    double Es[10][9];
    const int n = five_point(x0, x1, Es);
    for (int c = 0; c < n; ++c) {
      float e[9];
      for (int k = 0; k < 9; ++k) e[k] = static_cast<float>(Es[c][k]);
      int cnt = 0;
      for (int i0 = 0; i0 < S; i0 += 64) {
        const int lim = min(S, i0 + 64);
        for (int i = i0; i < lim; ++i) cnt += sampson_inlier(e, __ldg(m + i), p.thr2) ? 1 : 0;
        if (cnt + (S - lim) < max(best, *reinterpret_cast<volatile int*>(&s_blk_best))) { cnt = -1; break; }
      }
      if (cnt > best) {
        best = cnt;
        for (int k = 0; k < 9; ++k) bestE[k] = Es[c][k];
        atomicMax(&s_blk_best, cnt);
      }
    }
  }
  // arg-best over the CTA: (count, lowest thread)
  __shared__ int s_cnt[kRansacThreads / 32], s_tid[kRansacThreads / 32];
  int bc = best, bt = threadIdx.x;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const int oc = __shfl_xor_sync(0xffffffffu, bc, o), ot = __shfl_xor_sync(0xffffffffu, bt, o);
    if (oc > bc || (oc == bc && ot < bt)) { bc = oc; bt = ot; }
  }
  if ((threadIdx.x & 31) == 0) { s_cnt[threadIdx.x >> 5] = bc; s_tid[threadIdx.x >> 5] = bt; }
  __syncthreads();
  bc = s_cnt[0]; bt = s_tid[0];
#pragma unroll
  for (int w = 1; w < kRansacThreads / 32; ++w)
    if (s_cnt[w] > bc) { bc = s_cnt[w]; bt = s_tid[w]; }
  if (static_cast<int>(threadIdx.x) == bt) {
    const int64_t slot = static_cast<int64_t>(pair) * p.blocks_per_pair + blockIdx.x;
    p.blk_count[slot] = best;
    for (int k = 0; k < 9; ++k) p.blk_E[slot * 9 + k] = bestE[k];
  }
}

// ------------------------------------------------------------------------------------------------ pose
// Eigen-decomposition of a symmetric 3 x 3 matrix by cyclic Jacobi rotations: A = V diag(w) V^T.
__device__ inline void jacobi3(double A[3][3], double V[3][3], double w[3]) {
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) V[i][j] = i == j ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    const double off = fabs(A[0][1]) + fabs(A[0][2]) + fabs(A[1][2]);
    if (off < 1e-300) break;
    for (int pq = 0; pq < 3; ++pq) {
      const int p = pq == 2 ? 1 : 0, q = pq == 0 ? 1 : 2;
      if (fabs(A[p][q]) < 1e-300) continue;
      const double theta = (A[q][q] - A[p][p]) / (2.0 * A[p][q]);
      const double t = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
      const double c = 1.0 / sqrt(t * t + 1.0), s = t * c;
      for (int k = 0; k < 3; ++k) {       // A <- A J
        const double akp = A[k][p], akq = A[k][q];
        A[k][p] = c * akp - s * akq;
        A[k][q] = s * akp + c * akq;
      }
      for (int k = 0; k < 3; ++k) {       // A <- J^T A
        const double apk = A[p][k], aqk = A[q][k];
        A[p][k] = c * apk - s * aqk;
        A[q][k] = s * apk + c * aqk;
      }
      for (int k = 0; k < 3; ++k) {
        const double vkp = V[k][p], vkq = V[k][q];
        V[k][p] = c * vkp - s * vkq;
        V[k][q] = s * vkp + c * vkq;
      }
    }
  }
  for (int i = 0; i < 3; ++i) w[i] = A[i][i];
}

// The four (R, t) candidates of an essential matrix (Hartley & Zisserman 9.6.2): E = U diag(s, s, 0) V^T,
// R = U W V^T or U W^T V^T, t = +-u3; cand[c] = (R row-major, t), order (R1, t), (R2, t), (R1, -t), (R2, -t).
__device__ inline void decompose_essential(const double* E, double (*cand)[12]) {
  double EtE[3][3], V[3][3], w[3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) EtE[i][j] = E[0 + i] * E[0 + j] + E[3 + i] * E[3 + j] + E[6 + i] * E[6 + j];
  jacobi3(EtE, V, w);
  // sort columns by decreasing eigenvalue
  int o[3] = {0, 1, 2};
  for (int a = 0; a < 2; ++a)
    for (int b = a + 1; b < 3; ++b)
      if (w[o[b]] > w[o[a]]) { const int t = o[a]; o[a] = o[b]; o[b] = t; }
  double v[3][3], u[3][3];      // v[k] = k-th right singular vector, u[k] = k-th left singular vector
  for (int k = 0; k < 3; ++k)
    for (int i = 0; i < 3; ++i) v[k][i] = V[i][o[k]];
  for (int k = 0; k < 2; ++k) {
    double n = 0.0;
    for (int i = 0; i < 3; ++i) {
      u[k][i] = E[3 * i] * v[k][0] + E[3 * i + 1] * v[k][1] + E[3 * i + 2] * v[k][2];
      n += u[k][i] * u[k][i];
    }
    n = n > 0.0 ? rsqrt(n) : 0.0;
    for (int i = 0; i < 3; ++i) u[k][i] *= n;
  }
  // re-orthogonalise u1 against u0 (equal singular values: E v0, E v1 are orthogonal up to rounding)
  {
    const double d = u[0][0] * u[1][0] + u[0][1] * u[1][1] + u[0][2] * u[1][2];
    double n = 0.0;
    for (int i = 0; i < 3; ++i) { u[1][i] -= d * u[0][i]; n += u[1][i] * u[1][i]; }
    n = n > 0.0 ? rsqrt(n) : 0.0;
    for (int i = 0; i < 3; ++i) u[1][i] *= n;
  }
  // third vectors as cross products: det(U) = det(V) = +1
  u[2][0] = u[0][1] * u[1][2] - u[0][2] * u[1][1];
  u[2][1] = u[0][2] * u[1][0] - u[0][0] * u[1][2];
  u[2][2] = u[0][0] * u[1][1] - u[0][1] * u[1][0];
  v[2][0] = v[0][1] * v[1][2] - v[0][2] * v[1][1];
  v[2][1] = v[0][2] * v[1][0] - v[0][0] * v[1][2];
  v[2][2] = v[0][0] * v[1][1] - v[0][1] * v[1][0];
  // U W V^T = -u0 v1^T... written out: W = [[0,-1,0],[1,0,0],[0,0,1]] -> U W = [u1, -u0, u2]
  for (int c = 0; c < 4; ++c) {
    const double sgn = (c & 1) ? -1.0 : 1.0;      // W^T: [-u1, u0, u2]
    for (int i = 0; i < 3; ++i)
      for (int j = 0; j < 3; ++j)
        cand[c][3 * i + j] = sgn * (u[1][i] * v[0][j] - u[0][i] * v[1][j]) + u[2][i] * v[2][j];
    const double ts = c >= 2 ? -1.0 : 1.0;
    for (int i = 0; i < 3; ++i) cand[c][9 + i] = ts * u[2][i];
  }
}

struct PoseParams {
  const float4* matches;
  const int64_t* offsets;
  int32_t n_pairs;
  int32_t blocks_per_pair;
  float thr2;
  float max_depth;           // cv2.recoverPose distanceThresh (50)
  const int32_t* blk_count;
  const double* blk_E;
  double* out_E;             // [n_pairs, 9]
  uint8_t* out_mask;         // [total] RANSAC inliers of the winning matrix
  int32_t* out_inliers;      // [n_pairs]
  double* out_R;             // [n_pairs, 9]
  double* out_t;             // [n_pairs, 3]
  uint8_t* out_pose_mask;    // [total] inliers in front of both cameras (recoverPose's mask)
  int32_t* out_pose_inliers; // [n_pairs]
};

constexpr int kPoseThreads = 256;

__device__ __forceinline__ int block_sum(int v, int* scratch) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) scratch[threadIdx.x >> 5] = v;
  __syncthreads();
  int t = 0;
  for (int w = 0; w < kPoseThreads / 32; ++w) t += scratch[w];
  return t;
}

// depths of a match in both cameras for the pose (R, t): lambda1 x1 = lambda0 R x0 + t in the least-squares sense
__device__ __forceinline__ bool in_front(const float* Rt, const float4 m, float max_depth) {
  const float ax = fmaf(Rt[0], m.x, fmaf(Rt[1], m.y, Rt[2]));
  const float ay = fmaf(Rt[3], m.x, fmaf(Rt[4], m.y, Rt[5]));
  const float az = fmaf(Rt[6], m.x, fmaf(Rt[7], m.y, Rt[8]));
  const float aa = ax * ax + ay * ay + az * az, bb = m.z * m.z + m.w * m.w + 1.f;
  const float ab = ax * m.z + ay * m.w + az;
  const float at = ax * Rt[9] + ay * Rt[10] + az * Rt[11], bt = m.z * Rt[9] + m.w * Rt[10] + Rt[11];
  const float det = aa * bb - ab * ab;
  if (!(det > 1e-12f * aa * bb)) return false;        // parallel rays: at infinity
  const float l0 = (ab * bt - bb * at) / det, l1 = (aa * bt - ab * at) / det;
  return l0 > 0.f && l0 < max_depth && l1 > 0.f && l1 < max_depth;
}

__global__ void __launch_bounds__(kPoseThreads)
pose_kernel(const PoseParams p) {
  __shared__ float sE[9];
  __shared__ float sRt[4][12];
  __shared__ double sCand[4][12];
  __shared__ int scratch[kPoseThreads / 32];
  __shared__ int s_best;
  const int pair = blockIdx.x;
  const int64_t o0 = p.offsets[pair];
  const int S = static_cast<int>(p.offsets[pair + 1] - o0);
  const float4* m = p.matches + o0;
  if (threadIdx.x == 0) {
    int best = -1, bb = 0;
    for (int b = 0; b < p.blocks_per_pair; ++b) {
      const int c = p.blk_count[static_cast<int64_t>(pair) * p.blocks_per_pair + b];
      if (c > best) { best = c; bb = b; }
    }
    s_best = best;
    const double* E = p.blk_E + (static_cast<int64_t>(pair) * p.blocks_per_pair + bb) * 9;
    for (int k = 0; k < 9; ++k) {
      sE[k] = static_cast<float>(E[k]);
      p.out_E[static_cast<int64_t>(pair) * 9 + k] = best >= 0 ? E[k] : 0.0;
    }
    if (best >= 0) {
      decompose_essential(E, sCand);
      for (int c = 0; c < 4; ++c)
        for (int k = 0; k < 12; ++k) sRt[c][k] = static_cast<float>(sCand[c][k]);
    }
  }
  __syncthreads();
  const bool have = s_best >= 0;
  int n_in = 0, good[4] = {0, 0, 0, 0};
  for (int i = threadIdx.x; i < S; i += kPoseThreads) {
    const float4 v = __ldg(m + i);
    const bool in = have && sampson2(sE, v) < p.thr2;
    p.out_mask[o0 + i] = in ? 1 : 0;
    n_in += in ? 1 : 0;
    if (in) {
#pragma unroll
      for (int c = 0; c < 4; ++c) good[c] += in_front(sRt[c], v, p.max_depth) ? 1 : 0;
    }
  }
  n_in = block_sum(n_in, scratch);
  int g[4];
#pragma unroll
  for (int c = 0; c < 4; ++c) g[c] = block_sum(good[c], scratch);
  int bc = 0;
#pragma unroll
  for (int c = 1; c < 4; ++c)
    if (g[c] > g[bc]) bc = c;
  for (int i = threadIdx.x; i < S; i += kPoseThreads) {
    const float4 v = __ldg(m + i);
    const bool in = have && sampson2(sE, v) < p.thr2;
    p.out_pose_mask[o0 + i] = (in && in_front(sRt[bc], v, p.max_depth)) ? 1 : 0;
  }
  if (threadIdx.x == 0) {
    p.out_inliers[pair] = have ? n_in : 0;
    p.out_pose_inliers[pair] = have ? g[bc] : 0;
    for (int k = 0; k < 9; ++k) p.out_R[static_cast<int64_t>(pair) * 9 + k] = have ? sCand[bc][k] : (k % 4 == 0 ? 1.0 : 0.0);
    for (int k = 0; k < 3; ++k) p.out_t[static_cast<int64_t>(pair) * 3 + k] = have ? sCand[bc][9 + k] : 0.0;
  }
}

// one minimal solve (tests: the solver against cv2.findEssentialMat on exactly five points)
__global__ void five_point_debug_kernel(const double* x0, const double* x1, double* out_E, int* out_n) {
  double a[5][2], b[5][2], Es[10][9];
  for (int k = 0; k < 5; ++k) { a[k][0] = x0[2 * k]; a[k][1] = x0[2 * k + 1]; b[k][0] = x1[2 * k]; b[k][1] = x1[2 * k + 1]; }
  const int n = five_point(a, b, Es);
  for (int c = 0; c < n; ++c)
    for (int k = 0; k < 9; ++k) out_E[9 * c + k] = Es[c][k];
  *out_n = n;
}

}  // namespace geo
}  // namespace b200m
