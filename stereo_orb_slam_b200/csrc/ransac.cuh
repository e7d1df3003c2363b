// Stereo RANSAC rigid alignment, hypothesis-parallel.
//
// Replaces RANSAC::Solver<ObservationPair, Matrix4f>::Solve over StereoTransModel
// (reference: ransac_solver/include/ransac_solver.h:51-166, ransac_solver/include/stereo_reprojection.h:28-134),
// the step that hands the visual odometer (src/visual_odometer.cpp:93-103) and the loop closer
// (src/pose_graph_optimizer.cpp:189) their initial relative pose and inlier mask.
//
// The reference runs its 100 hypotheses one after the other and re-triangulates every element inside every
// inlier test (100 x n calls of cv::triangulatePoints on one point).  Here:
//   k_ransac_triangulate   one thread per observation: 4x4 DLT null vector by one-sided Jacobi in double, once
//   k_ransac_hypotheses    one CTA per (hypothesis, problem): thread 0 fits the 3-point Umeyama model in double,
//                          the CTA counts the inliers of all elements in float
//   k_ransac_finish        one CTA per problem: first-best hypothesis (or the first above the consensus
//                          threshold), optional final fit over its inliers, labels and losses of every element
// Several independent problems (frame pairs, loop candidates) go through one call.  All reductions are integer or
// fixed-order: results are bit-reproducible.  Float operations follow oracle/ransac_oracle.py's stated order
// (left-to-right sums, no fused multiply-add) so labels and losses are bit-comparable with it.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace sba {

struct RansacDev {
  int32_t nprob, iters;
  const int64_t* off;            // [nprob+1] element ranges
  const float4* first;           // [n] (u_l, v_l, u_r, v_r): triangulated, transformed, projected ...
  const float4* second;          // [n] ... and compared with these pixels
  float Pl[12], Pr[12];          // 3x4 projection matrices, row-major
  const int32_t* samples;        // [nprob, iters, 3] element indices local to the problem, or null
  uint64_t seed;                 // for the built-in draw when samples is null
  float thr_px, ratio;
  int32_t final_fit, early;
  float4* src;                   // [n] points triangulated from `first`
  float4* dst;                   // [n] from `second`
  float* Thyp;                   // [nprob*iters*12] models of the hypotheses (3x4 row-major)
  int32_t* counts;               // [nprob*iters]
  float* T_out;                  // [nprob*16]
  uint8_t* labels;               // [n]
  float* losses;                 // [n]
  int32_t* best_it;              // [nprob]
};

// ---------------------------------------------------------------------------------------------------------------
// One-sided Jacobi (Hestenes): rotates the columns of G (N x N, column j = G[j]) until they are orthogonal and
// applies the same rotations to V.  On return |G[j]| are the singular values and V[j] the right singular vectors.
template <int N>
__device__ __forceinline__ void jacobi_columns(double (&G)[N][N], double (&V)[N][N]) {
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j < N; ++j) V[i][j] = (i == j) ? 1.0 : 0.0;
  for (int sweep = 0; sweep < 30; ++sweep) {
    bool rotated = false;
#pragma unroll
    for (int p = 0; p < N - 1; ++p) {
#pragma unroll
      for (int q = p + 1; q < N; ++q) {
        double a = 0.0, b = 0.0, g = 0.0;
#pragma unroll
        for (int k = 0; k < N; ++k) { a += G[p][k] * G[p][k]; b += G[q][k] * G[q][k]; g += G[p][k] * G[q][k]; }
        if (g == 0.0 || fabs(g) <= 1e-17 * sqrt(a * b)) continue;
        rotated = true;
        const double zeta = (b - a) / (2.0 * g);
        const double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
        const double c = rsqrt(1.0 + t * t), s = c * t;
#pragma unroll
        for (int k = 0; k < N; ++k) {
          const double gp = G[p][k], gq = G[q][k];
          G[p][k] = c * gp - s * gq; G[q][k] = s * gp + c * gq;
          const double vp = V[p][k], vq = V[q][k];
          V[p][k] = c * vp - s * vq; V[q][k] = s * vp + c * vq;
        }
      }
    }
    if (!rotated) break;
  }
}

// cv::triangulatePoints for one point + the reference's float division by w (stereo_reprojection.h:79-92)
__device__ __forceinline__ float4 dev_triangulate(const float* Pl, const float* Pr, float4 o) {
  double G[4][4], V[4][4];                       // G[j] = column j of the DLT matrix
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    G[k][0] = (double)o.x * (double)Pl[8 + k] - (double)Pl[k];
    G[k][1] = (double)o.y * (double)Pl[8 + k] - (double)Pl[4 + k];
    G[k][2] = (double)o.z * (double)Pr[8 + k] - (double)Pr[k];
    G[k][3] = (double)o.w * (double)Pr[8 + k] - (double)Pr[4 + k];
  }
  jacobi_columns<4>(G, V);
  int jm = 0; double nm = 0.0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    double nn = 0.0;
#pragma unroll
    for (int k = 0; k < 4; ++k) nn += G[j][k] * G[j][k];
    if (j == 0 || nn < nm) { nm = nn; jm = j; }
  }
  double X[4] = {0, 0, 0, 0};
#pragma unroll
  for (int j = 0; j < 4; ++j) if (j == jm) {
#pragma unroll
    for (int k = 0; k < 4; ++k) X[k] = V[j][k];
  }
  const float w = (float)X[3];
  return make_float4(__fdiv_rn((float)X[0], w), __fdiv_rn((float)X[1], w), __fdiv_rn((float)X[2], w), 0.f);
}

// threads [0, n) triangulate `first` into src, threads [n, total) `second` into dst (total = n or 2n)
__global__ void __launch_bounds__(128) k_ransac_triangulate(RansacDev d, int64_t n, int64_t total) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  if (i < n) d.src[i] = dev_triangulate(d.Pl, d.Pr, d.first[i]);
  else d.dst[i - n] = dev_triangulate(d.Pl, d.Pr, d.second[i - n]);
}

// Eigen::umeyama(src, dst, false) from accumulated sums: n, sum src, sum dst and sum dst src^T (about the
// means ms / md the caller subtracted).  A proper rotation with R v1 = u1, R v2 = u2 for the two leading singular
// pairs of sigma is unique, R v3 = det(V) (u1 x u2): the third singular value (zero for 3 points) is never used.
// Rank-deficient samples (repeated or collinear picks) get a deterministic completion; the reference's result for
// those depends on how Eigen completes the basis and is not defined by its source.
__device__ void dev_umeyama_from_cov(const double* sig, const double* ms, const double* md, float* T12) {
  double G[3][3], V[3][3];
#pragma unroll
  for (int j = 0; j < 3; ++j)
#pragma unroll
    for (int k = 0; k < 3; ++k) G[j][k] = sig[3 * k + j];          // column j
  jacobi_columns<3>(G, V);
  double nn[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) nn[j] = G[j][0] * G[j][0] + G[j][1] * G[j][1] + G[j][2] * G[j][2];
  int i1 = 0;
  if (nn[1] > nn[i1]) i1 = 1;
  if (nn[2] > nn[i1]) i1 = 2;
  int i2 = (i1 == 0) ? 1 : 0;
#pragma unroll
  for (int j = 0; j < 3; ++j) if (j != i1 && nn[j] > nn[i2]) i2 = j;
  const int i3 = 3 - i1 - i2;
  double u1[3], u2[3], v1[3], v2[3], v3[3];
  auto pick = [&](int j, double* g, double* v) {
#pragma unroll
    for (int c = 0; c < 3; ++c) if (c == j) {
#pragma unroll
      for (int k = 0; k < 3; ++k) { g[k] = G[c][k]; v[k] = V[c][k]; }
    }
  };
  double g3[3];
  pick(i1, u1, v1); pick(i2, u2, v2); pick(i3, g3, v3);
  double R[9];
  const double s1 = sqrt(nn[i1]), s2 = sqrt(nn[i2]);
  if (!(s1 > 1e-20 * (ms[0] * ms[0] + ms[1] * ms[1] + ms[2] * ms[2] + md[0] * md[0] + md[1] * md[1] + md[2] * md[2])) || !(s1 > 0.0)) {   // all picks coincide
#pragma unroll
    for (int k = 0; k < 9; ++k) R[k] = (k % 4 == 0) ? 1.0 : 0.0;
  } else {
#pragma unroll
    for (int k = 0; k < 3; ++k) u1[k] /= s1;
    if (s2 > 1e-12 * s1) {
#pragma unroll
      for (int k = 0; k < 3; ++k) u2[k] /= s2;
    } else {                                                        // rank 1: any unit vector orthogonal to u1 / v1
      auto ortho = [](const double* a, double* o) {
        int m = 0;
        if (fabs(a[1]) < fabs(a[m])) m = 1;
        if (fabs(a[2]) < fabs(a[m])) m = 2;
        double e[3] = {m == 0 ? 1.0 : 0.0, m == 1 ? 1.0 : 0.0, m == 2 ? 1.0 : 0.0};
        const double dt = a[0] * e[0] + a[1] * e[1] + a[2] * e[2];
        double nrm = 0.0;
        for (int k = 0; k < 3; ++k) { o[k] = e[k] - dt * a[k]; nrm += o[k] * o[k]; }
        nrm = rsqrt(nrm);
        for (int k = 0; k < 3; ++k) o[k] *= nrm;
      };
      ortho(u1, u2); ortho(v1, v2);
      v3[0] = v1[1] * v2[2] - v1[2] * v2[1]; v3[1] = v1[2] * v2[0] - v1[0] * v2[2]; v3[2] = v1[0] * v2[1] - v1[1] * v2[0];
    }
    const double c3[3] = {u1[1] * u2[2] - u1[2] * u2[1], u1[2] * u2[0] - u1[0] * u2[2], u1[0] * u2[1] - u1[1] * u2[0]};
    const double w3[3] = {v1[1] * v2[2] - v1[2] * v2[1], v1[2] * v2[0] - v1[0] * v2[2], v1[0] * v2[1] - v1[1] * v2[0]};
    const double detV = (w3[0] * v3[0] + w3[1] * v3[1] + w3[2] * v3[2]) >= 0.0 ? 1.0 : -1.0;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) R[3 * r + c] = u1[r] * v1[c] + u2[r] * v2[c] + detV * c3[r] * v3[c];
  }
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    T12[4 * r + 0] = (float)R[3 * r]; T12[4 * r + 1] = (float)R[3 * r + 1]; T12[4 * r + 2] = (float)R[3 * r + 2];
    T12[4 * r + 3] = (float)(md[r] - (R[3 * r] * ms[0] + R[3 * r + 1] * ms[1] + R[3 * r + 2] * ms[2]));
  }
}

// ((m0 x + m1 y) + m2 z) + m3 w in float, no contraction (oracle/ransac_oracle.py _mat_vec)
__device__ __forceinline__ float row4(const float* m, float x, float y, float z, float w) {
  return __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(m[0], x), __fmul_rn(m[1], y)), __fmul_rn(m[2], z)), __fmul_rn(m[3], w));
}

// StereoTransModel::IsInlier's dist (stereo_reprojection.h:112-124)
__device__ __forceinline__ float dev_reproj_dist(const float* T12, const float* Pl, const float* Pr, float4 p, float4 o) {
  const float s0 = row4(T12, p.x, p.y, p.z, 1.f), s1 = row4(T12 + 4, p.x, p.y, p.z, 1.f), s2 = row4(T12 + 8, p.x, p.y, p.z, 1.f);
  const float s3 = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(0.f, p.x), __fmul_rn(0.f, p.y)), __fmul_rn(0.f, p.z)), 1.f);
  const float l0 = row4(Pl, s0, s1, s2, s3), l1 = row4(Pl + 4, s0, s1, s2, s3), l2 = row4(Pl + 8, s0, s1, s2, s3);
  const float r0 = row4(Pr, s0, s1, s2, s3), r1 = row4(Pr + 4, s0, s1, s2, s3), r2 = row4(Pr + 8, s0, s1, s2, s3);
  const float d0 = __fsub_rn(o.x, __fdiv_rn(l0, l2)), d1 = __fsub_rn(o.y, __fdiv_rn(l1, l2));
  const float d2 = __fsub_rn(o.z, __fdiv_rn(r0, r2)), d3 = __fsub_rn(o.w, __fdiv_rn(r1, r2));
  float acc = __fmul_rn(d0, d0);
  acc = __fadd_rn(acc, __fmul_rn(d1, d1));
  acc = __fadd_rn(acc, __fmul_rn(d2, d2));
  acc = __fadd_rn(acc, __fmul_rn(d3, d3));
  return __fsqrt_rn(acc);
}

__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}

constexpr int kRansacThreads = 128;

__global__ void __launch_bounds__(kRansacThreads) k_ransac_hypotheses(RansacDev d) {
  const int it = blockIdx.x, pb = blockIdx.y;
  const int64_t e0 = d.off[pb], n = d.off[pb + 1] - e0;
  __shared__ float T12[12];
  __shared__ int wsum[kRansacThreads / 32];
  const int64_t hyp = (int64_t)pb * d.iters + it;
  if (n <= 0) {                        // no elements: count 0 and a zero model (sba.h: such a problem gets T = 0)
    if (threadIdx.x == 0) d.counts[hyp] = 0;
    if (threadIdx.x < 12) d.Thyp[hyp * 12 + threadIdx.x] = 0.f;
    return;
  }
  if (threadIdx.x == 0) {
    int64_t idx[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      idx[k] = d.samples ? (int64_t)d.samples[hyp * 3 + k] : (int64_t)(splitmix64(d.seed + (uint64_t)(hyp * 3 + k)) % (uint64_t)n);
      if (idx[k] < 0 || idx[k] >= n) idx[k] = 0;                     // host validates; never index out of range
    }
    double ms[3] = {0, 0, 0}, md[3] = {0, 0, 0}, ps[3][3], pd[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const float4 a = d.src[e0 + idx[k]], b = d.dst[e0 + idx[k]];
      ps[k][0] = a.x; ps[k][1] = a.y; ps[k][2] = a.z; pd[k][0] = b.x; pd[k][1] = b.y; pd[k][2] = b.z;
#pragma unroll
      for (int c = 0; c < 3; ++c) { ms[c] += ps[k][c]; md[c] += pd[k][c]; }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) { ms[c] /= 3.0; md[c] /= 3.0; }
    double sig[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < 3; ++k) s += (pd[k][r] - md[r]) * (ps[k][c] - ms[c]);
        sig[3 * r + c] = s / 3.0;
      }
    float T[12];
    dev_umeyama_from_cov(sig, ms, md, T);
#pragma unroll
    for (int k = 0; k < 12; ++k) { T12[k] = T[k]; d.Thyp[hyp * 12 + k] = T[k]; }
  }
  __syncthreads();
  float T[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) T[k] = T12[k];
  int cnt = 0;
  for (int64_t i = threadIdx.x; i < n; i += kRansacThreads)
    cnt += dev_reproj_dist(T, d.Pl, d.Pr, d.src[e0 + i], d.second[e0 + i]) < d.thr_px ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
#pragma unroll
    for (int w = 0; w < kRansacThreads / 32; ++w) t += wsum[w];
    d.counts[hyp] = t;
  }
}

constexpr int kRansacFinishThreads = 256;

// fixed-order block sum of K doubles per thread (tree over shared memory); result in every thread's out[]
template <int K>
__device__ __forceinline__ void ransac_block_sum(double (&v)[K], double* sh /* [K * blockDim] */) {
  const int t = threadIdx.x;
#pragma unroll
  for (int k = 0; k < K; ++k) sh[k * kRansacFinishThreads + t] = v[k];
  __syncthreads();
  for (int s = kRansacFinishThreads / 2; s > 0; s >>= 1) {
    if (t < s) {
#pragma unroll
      for (int k = 0; k < K; ++k) sh[k * kRansacFinishThreads + t] += sh[k * kRansacFinishThreads + t + s];
    }
    __syncthreads();
  }
#pragma unroll
  for (int k = 0; k < K; ++k) v[k] = sh[k * kRansacFinishThreads];
  __syncthreads();
}

__global__ void __launch_bounds__(kRansacFinishThreads) k_ransac_finish(RansacDev d) {
  const int pb = blockIdx.x, t = threadIdx.x;
  const int64_t e0 = d.off[pb], n = d.off[pb + 1] - e0;
  __shared__ float T12[12];
  __shared__ int s_best, s_cnt;
  __shared__ double sh[9 * kRansacFinishThreads];
  if (t == 0) {
    // ransac_solver.h:111-125: strictly larger count wins (the first of equals); with early termination the loop
    // stops at the first hypothesis above the consensus threshold and later ones never ran (count -1)
    const unsigned thr = (unsigned)ceilf(__fmul_rn((float)(unsigned)n, d.ratio));
    int best = -1, bc = -1;
    int stop = d.iters;
    for (int i = 0; i < d.iters; ++i) {
      const int c = d.counts[(int64_t)pb * d.iters + i];
      if (c > bc) { bc = c; best = i; }
      if (d.early && (unsigned)c > thr) { stop = i + 1; break; }
    }
    for (int i = stop; i < d.iters; ++i) d.counts[(int64_t)pb * d.iters + i] = -1;
    s_best = best; s_cnt = 0;
    if (d.best_it) d.best_it[pb] = best;
    if (best >= 0) {
      for (int k = 0; k < 12; ++k) T12[k] = d.Thyp[((int64_t)pb * d.iters + best) * 12 + k];
    } else {
      for (int k = 0; k < 12; ++k) T12[k] = 0.f;                      // max_iterations 0: Matrix4f left unset by the reference
    }
  }
  __syncthreads();
  if (d.final_fit && n > 0) {
    float T[12];
#pragma unroll
    for (int k = 0; k < 12; ++k) T[k] = T12[k];
    // inliers of the best model (recomputed: same arithmetic, same answer as in k_ransac_hypotheses)
    int c = 0;
    for (int64_t i = t; i < n; i += kRansacFinishThreads)
      c += dev_reproj_dist(T, d.Pl, d.Pr, d.src[e0 + i], d.second[e0 + i]) < d.thr_px ? 1 : 0;
    atomicAdd(&s_cnt, c);                                            // integer: order-free
    __syncthreads();
    const bool use_all = !(s_cnt > 3);                               // ransac_solver.h:143-150
    const double cnt = use_all ? (double)n : (double)s_cnt;
    double m6[6] = {0, 0, 0, 0, 0, 0};
    for (int64_t i = t; i < n; i += kRansacFinishThreads) {
      const float4 a = d.src[e0 + i], b = d.dst[e0 + i];
      if (use_all || dev_reproj_dist(T, d.Pl, d.Pr, a, d.second[e0 + i]) < d.thr_px) {
        m6[0] += a.x; m6[1] += a.y; m6[2] += a.z; m6[3] += b.x; m6[4] += b.y; m6[5] += b.z;
      }
    }
    ransac_block_sum<6>(m6, sh);
    double ms[3] = {m6[0] / cnt, m6[1] / cnt, m6[2] / cnt}, md[3] = {m6[3] / cnt, m6[4] / cnt, m6[5] / cnt};
    double sg[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int64_t i = t; i < n; i += kRansacFinishThreads) {
      const float4 a = d.src[e0 + i], b = d.dst[e0 + i];
      if (use_all || dev_reproj_dist(T, d.Pl, d.Pr, a, d.second[e0 + i]) < d.thr_px) {
        const double sa[3] = {a.x - ms[0], a.y - ms[1], a.z - ms[2]}, sb[3] = {b.x - md[0], b.y - md[1], b.z - md[2]};
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int cc = 0; cc < 3; ++cc) sg[3 * r + cc] += sb[r] * sa[cc];
      }
    }
    ransac_block_sum<9>(sg, sh);
    if (t == 0) {
#pragma unroll
      for (int k = 0; k < 9; ++k) sg[k] /= cnt;
      float Tn[12];
      dev_umeyama_from_cov(sg, ms, md, Tn);
      for (int k = 0; k < 12; ++k) T12[k] = Tn[k];
    }
    __syncthreads();
  }
  float T[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) T[k] = T12[k];
  if (t < 16) d.T_out[(int64_t)pb * 16 + t] = t < 12 ? T12[t] : (t == 15 && s_best >= 0 ? 1.f : 0.f);
  for (int64_t i = t; i < n; i += kRansacFinishThreads) {            // ransac_solver.h:153-161
    const float dist = dev_reproj_dist(T, d.Pl, d.Pr, d.src[e0 + i], d.second[e0 + i]);
    d.losses[e0 + i] = dist;
    d.labels[e0 + i] = dist < d.thr_px ? 1 : 0;
  }
}

}  // namespace sba
