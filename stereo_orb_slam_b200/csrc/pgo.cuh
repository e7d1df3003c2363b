// pgo.cuh — pose-graph optimisation kernels, fp64, sm_100a.
//
// Replaces what PoseGraphOptimizer::Optimize hands to g2o (reference: src/pose_graph_optimizer.cpp:34-96; vertices and
// odometry edges :98-170, loop-closure edges :172-217): SE3 vertices (world-from-camera poses), relative-pose edges with
// information diag(0.01, 0.01, 0.01, 1, 1, 1) (:23-26) and a Huber kernel (:162, :207), vertex 0 fixed (:123-126),
// 10 Levenberg iterations (:66).  g2o's arithmetic restated (oracle/pgo_oracle.py is the CPU twin):
//   error      e = toVectorMQT(Z^-1 X_i^-1 X_j)      (translation, vector part of the unit quaternion with w >= 0)
//   update     X <- X * fromVectorMQT(delta)
//   quadratic  H += rho' J^T Omega J,  b -= rho' J^T Omega e,  chi2 = rho(e^T Omega e)
// The reduced system is the same 6x6-block sparse matrix family as the BA's camera system and goes through the same
// exact factorisation (k_spchol).  No atomics: every edge writes its five partial blocks, every block / vertex sums
// its static list in order.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace sba {

struct PgoDev {
  int32_t n, nf, m, nbu;
  const int32_t *fv;               // [n] free index or -1
  const int32_t *ei, *ej;          // [m]
  const double* Z;                 // [m*12] R_z (9, row-major), t_z (3)
  double *X, *Xn;                  // [n*12] current / candidate poses: R (9), t (3)
  double info[6], huber_delta;
  double* epart;                   // [m*120] per edge: H_ii (36) H_jj (36) H_ij (36) b_i (6) b_j (6)
  double* echi;                    // [m] rho(e^T Omega e) at X (linearise) or Xn (chi2)
  const int32_t *blk_part_off, *blk_part;   // per upper block: entries edge * 4 + kind (0 H_ii, 1 H_jj, 2 H_ij, 3 H_ij^T)
  const int32_t *vtx_part_off, *vtx_part;   // per free vertex: entries edge * 2 + side
  const int32_t *blk_row, *blk_col;
  double *H0, *b0;                 // [nbu*36], [nf*6] assembled, undamped
  double *Su, *rhs, *x;            // what k_spchol solves: H0 + lambda I, b0 -> x
  double* red;                     // [4] chi2, scale = x.(lambda x + b), max diag(H0)
};

// Eigen::Quaterniond(Matrix3d) (Shepperd's method), normalised, w >= 0 (g2o toCompactQuaternion).  q = (x, y, z, w).
__device__ __forceinline__ void pgo_quat_from_R(const double* R, double* q) {
  const double t = R[0] + R[4] + R[8];
  if (t > 0.0) {
    double s = sqrt(t + 1.0);
    q[3] = 0.5 * s;
    s = 0.5 / s;
    q[0] = (R[7] - R[5]) * s; q[1] = (R[2] - R[6]) * s; q[2] = (R[3] - R[1]) * s;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    double s = sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    q[i] = 0.5 * s;
    s = 0.5 / s;
    q[3] = (R[3 * k + j] - R[3 * j + k]) * s;
    q[j] = (R[3 * j + i] + R[3 * i + j]) * s;
    q[k] = (R[3 * k + i] + R[3 * i + k]) * s;
  }
  const double nrm = 1.0 / sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  const double sg = (q[3] < 0.0) ? -nrm : nrm;
  q[0] *= sg; q[1] *= sg; q[2] *= sg; q[3] *= sg;
}
__device__ __forceinline__ void pgo_quat_mul(const double* a, const double* b, double* o) {
  o[0] = a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1];
  o[1] = a[3] * b[1] - a[0] * b[2] + a[1] * b[3] + a[2] * b[0];
  o[2] = a[3] * b[2] + a[0] * b[1] - a[1] * b[0] + a[2] * b[3];
  o[3] = a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2];
}
// C = A^T B (3x3, row-major)
__device__ __forceinline__ void pgo_mulT(const double* A, const double* B, double* C) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = A[i] * B[j] + A[3 + i] * B[3 + j] + A[6 + i] * B[6 + j];
}

// e = toVectorMQT(Z^-1 X_i^-1 X_j); also R_e, q_e, R_b, t_b
__device__ __forceinline__ void pgo_error(const double* Xi, const double* Xj, const double* Zk, double* e, double* Re, double* q,
                                          double* Rb, double* tb) {
  pgo_mulT(Xi, Xj, Rb);
  const double d[3] = {Xj[9] - Xi[9], Xj[10] - Xi[10], Xj[11] - Xi[11]};
#pragma unroll
  for (int i = 0; i < 3; ++i) tb[i] = Xi[i] * d[0] + Xi[3 + i] * d[1] + Xi[6 + i] * d[2];
  pgo_mulT(Zk, Rb, Re);
  const double u[3] = {tb[0] - Zk[9], tb[1] - Zk[10], tb[2] - Zk[11]};
#pragma unroll
  for (int i = 0; i < 3; ++i) e[i] = Zk[i] * u[0] + Zk[3 + i] * u[1] + Zk[6 + i] * u[2];
  pgo_quat_from_R(Re, q);
  e[3] = q[0]; e[4] = q[1]; e[5] = q[2];
}

__device__ __forceinline__ void pgo_huber(double e2, double delta, double& rho, double& rho1) {
  const double d2 = delta * delta;
  if (e2 <= d2) { rho = e2; rho1 = 1.0; }
  else { const double s = sqrt(e2); rho = 2.0 * s * delta - d2; rho1 = delta / s; }
}

// thread per edge: error, the two 6x6 Jacobians (exact derivatives at delta = 0, see oracle/pgo_oracle.py
// edge_jacobians), Huber weight, and the edge's five partial blocks
__global__ void k_pgo_linearize(PgoDev g) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= g.m) return;
  const int i = g.ei[k], j = g.ej[k];
  const double* Xi = g.X + (size_t)i * 12; const double* Xj = g.X + (size_t)j * 12; const double* Zk = g.Z + (size_t)k * 12;
  double e[6], Re[9], q[4], Rb[9], tb[3];
  pgo_error(Xi, Xj, Zk, e, Re, q, Rb, tb);
  double Ji[36], Jj[36];
#pragma unroll
  for (int a = 0; a < 36; ++a) { Ji[a] = 0.0; Jj[a] = 0.0; }
  // X_j <- X_j D:  t' = t_e + R_e dt ;  q' = q_e (x) (dq, 1)  ->  w I + [v]x
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) Jj[6 * r + c] = Re[3 * r + c];
  Jj[6 * 3 + 3] = q[3];  Jj[6 * 3 + 4] = -q[2]; Jj[6 * 3 + 5] = q[1];
  Jj[6 * 4 + 3] = q[2];  Jj[6 * 4 + 4] = q[3];  Jj[6 * 4 + 5] = -q[0];
  Jj[6 * 5 + 3] = -q[1]; Jj[6 * 5 + 4] = q[0];  Jj[6 * 5 + 5] = q[3];
  // X_i <- X_i D:  t' = R_z^T ((I - 2 [dq]x) t_b - dt - t_z) ;  q' = q_z^-1 (x) (-dq, 1) (x) q_b
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) Ji[6 * r + c] = -Zk[3 * c + r];
  {
    const double S[9] = {0.0, -2.0 * tb[2], 2.0 * tb[1], 2.0 * tb[2], 0.0, -2.0 * tb[0], -2.0 * tb[1], 2.0 * tb[0], 0.0};   // 2 [t_b]x
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int c = 0; c < 3; ++c) Ji[6 * r + 3 + c] = Zk[r] * S[c] + Zk[3 + r] * S[3 + c] + Zk[6 + r] * S[6 + c];
  }
  {
    double qz[4], qb[4], qzi[4], p[4], t1[4], t2[4];
    pgo_quat_from_R(Zk, qz); pgo_quat_from_R(Rb, qb);
    qzi[0] = -qz[0]; qzi[1] = -qz[1]; qzi[2] = -qz[2]; qzi[3] = qz[3];
    pgo_quat_mul(qzi, qb, p);
    const double sg = (p[0] * q[0] + p[1] * q[1] + p[2] * q[2] + p[3] * q[3] > 0.0) ? 1.0 : -1.0;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      double ek[4] = {0.0, 0.0, 0.0, 0.0};
      ek[c] = -1.0;
      pgo_quat_mul(qzi, ek, t1); pgo_quat_mul(t1, qb, t2);
      Ji[6 * 3 + 3 + c] = sg * t2[0]; Ji[6 * 4 + 3 + c] = sg * t2[1]; Ji[6 * 5 + 3 + c] = sg * t2[2];
    }
  }
  double e2 = 0.0;
#pragma unroll
  for (int a = 0; a < 6; ++a) e2 += g.info[a] * e[a] * e[a];
  double rho, rho1;
  pgo_huber(e2, g.huber_delta, rho, rho1);
  g.echi[k] = rho;
  double w[6];
#pragma unroll
  for (int a = 0; a < 6; ++a) w[a] = rho1 * g.info[a];
  double* out = g.epart + (size_t)k * 120;
#pragma unroll
  for (int r = 0; r < 6; ++r) {
#pragma unroll
    for (int c = 0; c < 6; ++c) {
      double hii = 0.0, hjj = 0.0, hij = 0.0;
#pragma unroll
      for (int a = 0; a < 6; ++a) {
        hii += Ji[6 * a + r] * w[a] * Ji[6 * a + c];
        hjj += Jj[6 * a + r] * w[a] * Jj[6 * a + c];
        hij += Ji[6 * a + r] * w[a] * Jj[6 * a + c];
      }
      out[6 * r + c] = hii; out[36 + 6 * r + c] = hjj; out[72 + 6 * r + c] = hij;
    }
    double bi = 0.0, bj = 0.0;
#pragma unroll
    for (int a = 0; a < 6; ++a) { bi -= Ji[6 * a + r] * w[a] * e[a]; bj -= Jj[6 * a + r] * w[a] * e[a]; }
    out[108 + r] = bi; out[114 + r] = bj;
  }
}

// CTA of 64 threads per upper block (threads 0..35: one element each, list order) and per free vertex (b)
__global__ void __launch_bounds__(64) k_pgo_assemble(PgoDev g) {
  const int b = blockIdx.x, t = threadIdx.x;
  if (b < g.nbu) {
    if (t >= 36) return;
    const int r = t / 6, c = t - 6 * r;
    double acc = 0.0;
    for (int l = g.blk_part_off[b]; l < g.blk_part_off[b + 1]; ++l) {
      const int code = g.blk_part[l], k = code >> 2, kind = code & 3;
      const double* p = g.epart + (size_t)k * 120;
      acc += (kind == 0) ? p[t] : (kind == 1) ? p[36 + t] : (kind == 2) ? p[72 + t] : p[72 + 6 * c + r];
    }
    g.H0[(size_t)b * 36 + t] = acc;
    return;
  }
  const int f = b - g.nbu;
  if (f >= g.nf || t >= 6) return;
  double acc = 0.0;
  for (int l = g.vtx_part_off[f]; l < g.vtx_part_off[f + 1]; ++l) {
    const int code = g.vtx_part[l];
    acc += g.epart[(size_t)(code >> 1) * 120 + 108 + 6 * (code & 1) + t];
  }
  g.b0[(size_t)f * 6 + t] = acc;
}

// Su = H0 + lambda I on the diagonal blocks, rhs = b0
__global__ void k_pgo_damp(PgoDev g, double lambda) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < g.nbu * 36) {
    const int b = t / 36, e = t - 36 * b;
    double v = g.H0[t];
    if (g.blk_row[b] == g.blk_col[b] && (e % 7) == 0) v += lambda;
    g.Su[t] = v;
  }
  if (t < g.nf * 6) g.rhs[t] = g.b0[t];
}

// thread per vertex: Xn = X * fromVectorMQT(x_f) (g2o VertexSE3::oplusImpl); fixed vertices are copied
__global__ void k_pgo_update(PgoDev g) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= g.n) return;
  const double* X = g.X + (size_t)v * 12;
  double* O = g.Xn + (size_t)v * 12;
  const int f = g.fv[v];
  if (f < 0) {
#pragma unroll
    for (int a = 0; a < 12; ++a) O[a] = X[a];
    return;
  }
  const double* d = g.x + (size_t)f * 6;
  const double qx = d[3], qy = d[4], qz = d[5];
  const double w2 = 1.0 - (qx * qx + qy * qy + qz * qz);
  double D[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1};
  if (!(w2 < 0.0)) {
    const double qw = sqrt(w2);
    D[0] = 1 - 2 * (qy * qy + qz * qz); D[1] = 2 * (qx * qy - qz * qw); D[2] = 2 * (qx * qz + qy * qw);
    D[3] = 2 * (qx * qy + qz * qw); D[4] = 1 - 2 * (qx * qx + qz * qz); D[5] = 2 * (qy * qz - qx * qw);
    D[6] = 2 * (qx * qz - qy * qw); D[7] = 2 * (qy * qz + qx * qw); D[8] = 1 - 2 * (qx * qx + qy * qy);
  }
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c) O[3 * r + c] = X[3 * r] * D[c] + X[3 * r + 1] * D[3 + c] + X[3 * r + 2] * D[6 + c];
    O[9 + r] = X[9 + r] + X[3 * r] * d[0] + X[3 * r + 1] * d[1] + X[3 * r + 2] * d[2];
  }
}

// thread per edge: rho(e^T Omega e) at the candidate
__global__ void k_pgo_chi2(PgoDev g) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= g.m) return;
  double e[6], Re[9], q[4], Rb[9], tb[3];
  pgo_error(g.Xn + (size_t)g.ei[k] * 12, g.Xn + (size_t)g.ej[k] * 12, g.Z + (size_t)k * 12, e, Re, q, Rb, tb);
  double e2 = 0.0;
#pragma unroll
  for (int a = 0; a < 6; ++a) e2 += g.info[a] * e[a] * e[a];
  double rho, rho1;
  pgo_huber(e2, g.huber_delta, rho, rho1);
  g.echi[k] = rho;
}

// one CTA of 256: red[0] = sum echi (fixed order: strided per-thread sums, then a fixed tree), red[1] = x.(lambda x + b0),
// red[2] = max diagonal entry of H0 over the diagonal blocks
__global__ void __launch_bounds__(256) k_pgo_reduce(PgoDev g, double lambda) {
  __shared__ double sm[3][256];
  const int t = threadIdx.x;
  double c = 0.0, s = 0.0, mx = 0.0;
  for (int k = t; k < g.m; k += 256) c += g.echi[k];
  for (int a = t; a < g.nf * 6; a += 256) s += g.x[a] * (lambda * g.x[a] + g.b0[a]);
  for (int a = t; a < g.nbu * 6; a += 256) {
    const int b = a / 6, d = a - 6 * b;
    if (g.blk_row[b] == g.blk_col[b]) mx = fmax(mx, g.H0[(size_t)b * 36 + 7 * d]);
  }
  sm[0][t] = c; sm[1][t] = s; sm[2][t] = mx;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (t < o) { sm[0][t] += sm[0][t + o]; sm[1][t] += sm[1][t + o]; sm[2][t] = fmax(sm[2][t], sm[2][t + o]); }
    __syncthreads();
  }
  if (t == 0) { g.red[0] = sm[0][0]; g.red[1] = sm[1][0]; g.red[2] = sm[2][0]; }
}

// accepted: X <- Xn
__global__ void k_pgo_accept(PgoDev g) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < g.n * 12) g.X[t] = g.Xn[t];
}

}  // namespace sba
