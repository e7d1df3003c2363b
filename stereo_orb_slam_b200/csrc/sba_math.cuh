// sba_math.cuh — per-observation algebra of the stereo BA hot path (fp64), shared by every kernel.
//
// Replaces, analytically, what the reference obtains from Ceres autodiff of
// ReprojectionError::operator() (reference: src/reprojection_error.h:12-41, :58) plus
// ceres::HuberLoss(1.0) (src/bundle_adjuster.cpp:100).
//
//   p = R(w) X + t                                   (src/reprojection_error.h:19-24)
//   for P in {P_l, P_r}:  c = P[2,:]·[p;1],  u = P[0,:]·[p;1]/c,  v = P[1,:]·[p;1]/c   (:27-33)
//   r = (u_l-ul, v_l-vl, u_r-ur, v_r-vr)             (:35-38)
//   D = dr/dp (4x3):  rows (P[0,0:3] - u P[2,0:3])/c , (P[1,0:3] - v P[2,0:3])/c
//   dp/dt = I,  dp/dX = R,  dp/dw = G = -[R X]x J_l(w),  J_l = left Jacobian of SO(3)
// so J_cam = [D G | D] (4x6), J_pt = D R (4x3), and with the Huber weight rho' every normal-
// equation block of one observation derives from  A = rho' D^T D (3x3 sym)  and  h = rho' D^T r:
//   U_tt = A, U_wt = G^T A, U_ww = G^T A G ;  g_t = h, g_w = G^T h
//   V    = R^T A R ;  g_p = R^T h ;  W = [G^T A R ; A R]  (6x3)
// Jacobians are never materialised.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define SBA_HD __host__ __device__ __forceinline__
#else
#define SBA_HD inline
#endif

namespace sba {

// per-camera precompute (K0): R (row-major 9), J_l (row-major 9), t (3)  => 21 doubles
struct CamPre {
  double R[9];
  double Jl[9];
  double t[3];
};

SBA_HD void cam_precompute(const double* pose, CamPre& c) {
  const double wx = pose[0], wy = pose[1], wz = pose[2];
  const double th2 = wx * wx + wy * wy + wz * wz;
  double A, B, C;  // sin(th)/th, (1-cos th)/th^2, (th - sin th)/th^3
  if (th2 < 1e-8) {
    A = 1.0 - th2 * (1.0 / 6.0) + th2 * th2 * (1.0 / 120.0);
    B = 0.5 - th2 * (1.0 / 24.0) + th2 * th2 * (1.0 / 720.0);
    C = (1.0 / 6.0) - th2 * (1.0 / 120.0) + th2 * th2 * (1.0 / 5040.0);
  } else {
    const double th = sqrt(th2);
    const double s = sin(th), co = cos(th);
    A = s / th;
    B = (1.0 - co) / th2;
    C = (th - s) / (th2 * th);
  }
  // K = [w]x ; K2 = w w^T - th2 I
  const double xx = wx * wx, yy = wy * wy, zz = wz * wz, xy = wx * wy, xz = wx * wz, yz = wy * wz;
  const double K2[9] = {xx - th2, xy, xz, xy, yy - th2, yz, xz, yz, zz - th2};
  const double K[9] = {0.0, -wz, wy, wz, 0.0, -wx, -wy, wx, 0.0};
#pragma unroll
  for (int i = 0; i < 9; ++i) {
    const double I = (i == 0 || i == 4 || i == 8) ? 1.0 : 0.0;
    c.R[i] = I + A * K[i] + B * K2[i];
    c.Jl[i] = I + B * K[i] + C * K2[i];
  }
  c.t[0] = pose[3]; c.t[1] = pose[4]; c.t[2] = pose[5];
}

// Huber on the squared block norm s: returns rho(s) and rho'(s)   (ceres::HuberLoss restated)
SBA_HD void huber(double delta, double s, double& rho, double& rho1) {
  const double b = delta * delta;
  if (s > b) {
    const double rt = sqrt(s);
    rho = 2.0 * delta * rt - b;
    rho1 = delta / rt;
  } else {
    rho = s;
    rho1 = 1.0;
  }
}

// residual (uncorrected) + D.  rdim 4 or 3.  Returns squared norm.
SBA_HD double project(const double* __restrict__ pl, const double* __restrict__ pr, const CamPre& c,
                      const double X[3], const float uv[4], int rdim, double q[3], double r[4],
                      double D[12]) {
  q[0] = c.R[0] * X[0] + c.R[1] * X[1] + c.R[2] * X[2];
  q[1] = c.R[3] * X[0] + c.R[4] * X[1] + c.R[5] * X[2];
  q[2] = c.R[6] * X[0] + c.R[7] * X[1] + c.R[8] * X[2];
  const double p0 = q[0] + c.t[0], p1 = q[1] + c.t[1], p2 = q[2] + c.t[2];
  {
    const double a = pl[0] * p0 + pl[1] * p1 + pl[2] * p2 + pl[3];
    const double b = pl[4] * p0 + pl[5] * p1 + pl[6] * p2 + pl[7];
    const double cc = pl[8] * p0 + pl[9] * p1 + pl[10] * p2 + pl[11];
    const double inv = 1.0 / cc;
    const double u = a * inv, v = b * inv;
    r[0] = u - (double)uv[0];
    r[1] = v - (double)uv[1];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      D[k] = (pl[k] - u * pl[8 + k]) * inv;
      D[3 + k] = (pl[4 + k] - v * pl[8 + k]) * inv;
    }
  }
  {
    const double a = pr[0] * p0 + pr[1] * p1 + pr[2] * p2 + pr[3];
    const double b = pr[4] * p0 + pr[5] * p1 + pr[6] * p2 + pr[7];
    const double cc = pr[8] * p0 + pr[9] * p1 + pr[10] * p2 + pr[11];
    const double inv = 1.0 / cc;
    const double u = a * inv, v = b * inv;
    r[2] = u - (double)uv[2];
    r[3] = (rdim == 4) ? (v - (double)uv[3]) : 0.0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      D[6 + k] = (pr[k] - u * pr[8 + k]) * inv;
      D[9 + k] = (rdim == 4) ? (pr[4 + k] - v * pr[8 + k]) * inv : 0.0;
    }
  }
  return r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3];
}

// residual only (candidate cost, K6): returns 1/2 rho(||r||^2)
SBA_HD double obs_cost(const double* __restrict__ pl, const double* __restrict__ pr, const CamPre& c,
                       const double X[3], const float uv[4], int rdim, double delta) {
  const double q0 = c.R[0] * X[0] + c.R[1] * X[1] + c.R[2] * X[2] + c.t[0];
  const double q1 = c.R[3] * X[0] + c.R[4] * X[1] + c.R[5] * X[2] + c.t[1];
  const double q2 = c.R[6] * X[0] + c.R[7] * X[1] + c.R[8] * X[2] + c.t[2];
  const double il = 1.0 / (pl[8] * q0 + pl[9] * q1 + pl[10] * q2 + pl[11]);
  const double r0 = (pl[0] * q0 + pl[1] * q1 + pl[2] * q2 + pl[3]) * il - (double)uv[0];
  const double r1 = (pl[4] * q0 + pl[5] * q1 + pl[6] * q2 + pl[7]) * il - (double)uv[1];
  const double ir = 1.0 / (pr[8] * q0 + pr[9] * q1 + pr[10] * q2 + pr[11]);
  const double r2 = (pr[0] * q0 + pr[1] * q1 + pr[2] * q2 + pr[3]) * ir - (double)uv[2];
  const double r3 = (rdim == 4) ? ((pr[4] * q0 + pr[5] * q1 + pr[6] * q2 + pr[7]) * ir - (double)uv[3]) : 0.0;
  double rho, rho1;
  huber(delta, r0 * r0 + r1 * r1 + r2 * r2 + r3 * r3, rho, rho1);
  return 0.5 * rho;
}

// A = w D^T D (sym, 6: 00 01 02 11 12 22), h = w D^T r
SBA_HD void form_A_h(const double D[12], const double r[4], double w, double A[6], double h[3]) {
  double a00 = 0, a01 = 0, a02 = 0, a11 = 0, a12 = 0, a22 = 0, h0 = 0, h1 = 0, h2 = 0;
#pragma unroll
  for (int m = 0; m < 4; ++m) {
    const double d0 = D[3 * m], d1 = D[3 * m + 1], d2 = D[3 * m + 2];
    a00 += d0 * d0; a01 += d0 * d1; a02 += d0 * d2; a11 += d1 * d1; a12 += d1 * d2; a22 += d2 * d2;
    h0 += d0 * r[m]; h1 += d1 * r[m]; h2 += d2 * r[m];
  }
  A[0] = w * a00; A[1] = w * a01; A[2] = w * a02; A[3] = w * a11; A[4] = w * a12; A[5] = w * a22;
  h[0] = w * h0; h[1] = w * h1; h[2] = w * h2;
}

// G = -[q]x Jl  : column j of G = Jl[:,j] x q
SBA_HD void form_G(const CamPre& c, const double q[3], double G[9]) {
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const double a0 = c.Jl[j], a1 = c.Jl[3 + j], a2 = c.Jl[6 + j];
    G[j] = a1 * q[2] - a2 * q[1];
    G[3 + j] = a2 * q[0] - a0 * q[2];
    G[6 + j] = a0 * q[1] - a1 * q[0];
  }
}

// full symmetric 3x3 from 6
SBA_HD void sym3_full(const double A[6], double F[9]) {
  F[0] = A[0]; F[1] = A[1]; F[2] = A[2]; F[3] = A[1]; F[4] = A[3]; F[5] = A[4]; F[6] = A[2]; F[7] = A[4]; F[8] = A[5];
}

// C = A(3x3 full) * B(3x3)
SBA_HD void mul33(const double A[9], const double B[9], double C[9]) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = A[3 * i] * B[j] + A[3 * i + 1] * B[3 + j] + A[3 * i + 2] * B[6 + j];
}
// C = A^T * B
SBA_HD void mulT33(const double A[9], const double B[9], double C[9]) {
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C[3 * i + j] = A[i] * B[j] + A[3 + i] * B[3 + j] + A[6 + i] * B[6 + j];
}

// inverse of SPD 3x3 given as 6 (00 01 02 11 12 22). Returns false if det <= 0 / non-finite.
SBA_HD bool inv_sym3(const double M[6], double Mi[6]) {
  const double a = M[0], b = M[1], c = M[2], d = M[3], e = M[4], f = M[5];
  const double c00 = d * f - e * e, c01 = c * e - b * f, c02 = b * e - c * d;
  const double det = a * c00 + b * c01 + c * c02;
  if (!(det > 0.0) || !isfinite(det)) {
    Mi[0] = Mi[1] = Mi[2] = Mi[3] = Mi[4] = Mi[5] = 0.0;
    return false;
  }
  const double id = 1.0 / det;
  Mi[0] = c00 * id; Mi[1] = c01 * id; Mi[2] = c02 * id;
  Mi[3] = (a * f - c * c) * id; Mi[4] = (b * c - a * e) * id; Mi[5] = (a * d - b * b) * id;
  return true;
}

SBA_HD double clampd(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }

// index of (i,j), i<=j, in the packed upper triangle of a 6x6 (21 entries, row-major upper)
SBA_HD int up6(int i, int j) { return i * 6 - (i * (i - 1)) / 2 + (j - i); }

// Accumulate one observation's camera-side blocks: U (21 upper) and gc (6).
SBA_HD void accum_camera(const double A[6], const double h[3], const double G[9], double U[21], double gc[6]) {
  double Af[9]; sym3_full(A, Af);
  double GtA[9]; mulT33(G, Af, GtA);          // G^T A   (3x3)  = U_wt
  // U_ww = GtA * G (sym)
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = i; j < 3; ++j)
      U[up6(i, j)] += GtA[3 * i] * G[j] + GtA[3 * i + 1] * G[3 + j] + GtA[3 * i + 2] * G[6 + j];
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) U[up6(i, 3 + j)] += GtA[3 * i + j];
  U[up6(3, 3)] += A[0]; U[up6(3, 4)] += A[1]; U[up6(3, 5)] += A[2];
  U[up6(4, 4)] += A[3]; U[up6(4, 5)] += A[4]; U[up6(5, 5)] += A[5];
#pragma unroll
  for (int i = 0; i < 3; ++i) gc[i] += G[i] * h[0] + G[3 + i] * h[1] + G[6 + i] * h[2];
  gc[3] += h[0]; gc[4] += h[1]; gc[5] += h[2];
}

// One observation's point-side blocks: V (6 sym) += R^T A R, gp += R^T h, W (6x3 row-major) = [G^T A R; A R]
SBA_HD void accum_point(const CamPre& c, const double A[6], const double h[3], const double G[9], double V[6],
                        double gp[3], double W[18]) {
  double Af[9]; sym3_full(A, Af);
  double AR[9]; mul33(Af, c.R, AR);           // A R
  // V += R^T (A R)  (sym)
  V[0] += c.R[0] * AR[0] + c.R[3] * AR[3] + c.R[6] * AR[6];
  V[1] += c.R[0] * AR[1] + c.R[3] * AR[4] + c.R[6] * AR[7];
  V[2] += c.R[0] * AR[2] + c.R[3] * AR[5] + c.R[6] * AR[8];
  V[3] += c.R[1] * AR[1] + c.R[4] * AR[4] + c.R[7] * AR[7];
  V[4] += c.R[1] * AR[2] + c.R[4] * AR[5] + c.R[7] * AR[8];
  V[5] += c.R[2] * AR[2] + c.R[5] * AR[5] + c.R[8] * AR[8];
#pragma unroll
  for (int i = 0; i < 3; ++i) gp[i] += c.R[i] * h[0] + c.R[3 + i] * h[1] + c.R[6 + i] * h[2];
  mulT33(G, AR, W);                            // rows 0..2 : G^T A R
#pragma unroll
  for (int i = 0; i < 9; ++i) W[9 + i] = AR[i];  // rows 3..5 : A R
}

// in-place inverse of an SPD 6x6 (full, row-major) by Cholesky; returns false if not SPD.
// Every loop is unrolled so that L and its inverse live in registers (rolled, they sit in local memory and the
// single thread that runs this per camera takes > 10 us).
SBA_HD bool inv_spd6(double* M) {
  // all loops run 0..5 with the triangular ranges as (compile-time after unrolling) predicates
  double L[36];
  bool ok = true;
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double d = M[7 * j];
#pragma unroll
    for (int k = 0; k < 6; ++k)
      if (k < j) d -= L[6 * j + k] * L[6 * j + k];
    if (!(d > 0.0) || !isfinite(d)) { ok = false; d = 1.0; }
    d = sqrt(d);
    L[7 * j] = d;
    const double id = 1.0 / d;
#pragma unroll
    for (int i = 0; i < 6; ++i)
      if (i > j) {
        double s = M[6 * i + j];
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (k < j) s -= L[6 * i + k] * L[6 * j + k];
        L[6 * i + j] = s * id;
      }
  }
  if (!ok) return false;
  // invert L (lower) -> Li
  double Li[36];
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    Li[7 * j] = 1.0 / L[7 * j];
#pragma unroll
    for (int i = 0; i < 6; ++i)
      if (i > j) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (k >= j && k < i) s -= L[6 * i + k] * Li[6 * k + j];
        Li[6 * i + j] = s / L[7 * i];
      }
  }
  // M^-1 = Li^T Li
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j < 6; ++j)
      if (j <= i) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < 6; ++k)
          if (k >= i) s += Li[6 * k + i] * Li[6 * k + j];
        M[6 * i + j] = s;
        M[6 * j + i] = s;
      }
  return true;
}

}  // namespace sba
