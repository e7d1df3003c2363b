// =====================================================================================
// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// CPU restatement of the stereo bundle-adjustment path that the reference hands to
// Ceres Solver (reference: src/bundle_adjuster.cpp:39-133, src/reprojection_error.h:12-62).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
// may load this library.  The product (libsba.so) never links or calls it.
//
// PARITY UNPINNED: the arithmetic of the path lives in Ceres Solver, which is not vendored
// and not version-pinned by the reference (CMakeLists.txt:44,104) and is absent from this
// image, and the reference holds no tests / golden vectors for this path.
// What is pinned by reference source and restated here line by line:
//   * the residual            src/reprojection_error.h:12-41  (4 residuals, general 3x4 P)
//   * autodiff dimensions     src/reprojection_error.h:58     (<4, 6, 3>)
//   * robust loss             src/bundle_adjuster.cpp:100     (HuberLoss(1.0) per block)
//   * point bounds            src/bundle_adjuster.cpp:104-108 (+-1e4)
//   * gauge                   src/bundle_adjuster.cpp:113     (first camera constant)
//   * solver options          src/bundle_adjuster.cpp:14-36   (LM, SPARSE_SCHUR, 50 its, tolerances)
// What is restated from Ceres's *published* algorithm (ceres-solver.org docs, >= 2.0):
//   * AngleAxisRotatePoint incl. its small-angle branch (ceres/rotation.h)
//   * Jet-based forward autodiff (ceres/jet.h) -> evaluated here with a 9-wide Jet
//   * HuberLoss rho/rho'/rho'' and the Triggs corrector (loss_function.cc, corrector.cc)
//   * trust-region LM loop: Jacobi column scaling, LM diagonal clamped to [1e-6,1e32],
//     radius update rules, step acceptance (min_relative_decrease 1e-3), parameter /
//     function / gradient tolerance tests, bounds projection + projected Armijo line
//     search (trust_region_minimizer.cc, levenberg_marquardt_strategy.cc)
//   * exact Schur elimination of all points + Cholesky of the reduced camera matrix
//     (schur_complement_solver.cc; dense LLT here instead of CHOLMOD — same solution).
// The KATs in tests/golden/ (50-digit mpmath) pin the residual/Jacobian arithmetic.
// =====================================================================================
#include <algorithm>
#include <chrono>
#include <cmath>
#include <complex>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <set>
#include <utility>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ---------------------------------------------------------------- Jet (ceres/jet.h restated)
template <int N>
struct Jet {
  double a;
  double v[N];
  Jet() : a(0.0) { for (int i = 0; i < N; ++i) v[i] = 0.0; }
  Jet(double s) : a(s) { for (int i = 0; i < N; ++i) v[i] = 0.0; }  // NOLINT
  Jet(double s, int k) : a(s) { for (int i = 0; i < N; ++i) v[i] = 0.0; v[k] = 1.0; }
};
template <int N> inline Jet<N> operator+(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a + g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] + g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a - g.a; for (int i = 0; i < N; ++i) h.v[i] = f.v[i] - g.v[i]; return h; }
template <int N> inline Jet<N> operator*(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; h.a = f.a * g.a; for (int i = 0; i < N; ++i) h.v[i] = f.a * g.v[i] + f.v[i] * g.a; return h; }
template <int N> inline Jet<N> operator/(const Jet<N>& f, const Jet<N>& g) {
  Jet<N> h; const double gi = 1.0 / g.a; const double fg = f.a * gi; h.a = fg;
  for (int i = 0; i < N; ++i) h.v[i] = (f.v[i] - fg * g.v[i]) * gi; return h; }
template <int N> inline Jet<N> operator*(double s, const Jet<N>& g) {
  Jet<N> h; h.a = s * g.a; for (int i = 0; i < N; ++i) h.v[i] = s * g.v[i]; return h; }
template <int N> inline Jet<N> operator-(double s, const Jet<N>& g) {
  Jet<N> h; h.a = s - g.a; for (int i = 0; i < N; ++i) h.v[i] = -g.v[i]; return h; }
template <int N> inline Jet<N> operator-(const Jet<N>& f, double s) {
  Jet<N> h = f; h.a -= s; return h; }
template <int N> inline Jet<N> operator+(const Jet<N>& f, double s) {
  Jet<N> h = f; h.a += s; return h; }
template <int N> inline Jet<N> operator/(double s, const Jet<N>& g) {
  Jet<N> h; const double gi = 1.0 / g.a; h.a = s * gi; const double m = -s * gi * gi;
  for (int i = 0; i < N; ++i) h.v[i] = m * g.v[i]; return h; }
template <int N> inline Jet<N> jsqrt(const Jet<N>& f) {
  Jet<N> h; h.a = std::sqrt(f.a); const double m = 1.0 / (2.0 * h.a);
  for (int i = 0; i < N; ++i) h.v[i] = m * f.v[i]; return h; }
template <int N> inline Jet<N> jsin(const Jet<N>& f) {
  Jet<N> h; h.a = std::sin(f.a); const double c = std::cos(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = c * f.v[i]; return h; }
template <int N> inline Jet<N> jcos(const Jet<N>& f) {
  Jet<N> h; h.a = std::cos(f.a); const double s = -std::sin(f.a);
  for (int i = 0; i < N; ++i) h.v[i] = s * f.v[i]; return h; }
inline double jsqrt(double x) { return std::sqrt(x); }
inline double jsin(double x) { return std::sin(x); }
inline double jcos(double x) { return std::cos(x); }
inline double scalar_of(double x) { return x; }
template <int N> inline double scalar_of(const Jet<N>& x) { return x.a; }

// ------------------------------------------------ ceres::AngleAxisRotatePoint (rotation.h restated)
template <typename T>
inline void angle_axis_rotate_point(const T aa[3], const T pt[3], T result[3]) {
  const T theta2 = aa[0] * aa[0] + aa[1] * aa[1] + aa[2] * aa[2];
  if (scalar_of(theta2) > std::numeric_limits<double>::epsilon()) {
    const T theta = jsqrt(theta2);
    const T costheta = jcos(theta);
    const T sintheta = jsin(theta);
    const T theta_inverse = 1.0 / theta;
    const T w[3] = {aa[0] * theta_inverse, aa[1] * theta_inverse, aa[2] * theta_inverse};
    const T w_cross_pt[3] = {w[1] * pt[2] - w[2] * pt[1], w[2] * pt[0] - w[0] * pt[2],
                             w[0] * pt[1] - w[1] * pt[0]};
    const T tmp = (w[0] * pt[0] + w[1] * pt[1] + w[2] * pt[2]) * (1.0 - costheta);
    result[0] = pt[0] * costheta + w_cross_pt[0] * sintheta + w[0] * tmp;
    result[1] = pt[1] * costheta + w_cross_pt[1] * sintheta + w[1] * tmp;
    result[2] = pt[2] * costheta + w_cross_pt[2] * sintheta + w[2] * tmp;
  } else {
    // first-order Taylor: R ~ I + [aa]x   (the branch autodiff differentiates near zero)
    const T w_cross_pt[3] = {aa[1] * pt[2] - aa[2] * pt[1], aa[2] * pt[0] - aa[0] * pt[2],
                             aa[0] * pt[1] - aa[1] * pt[0]};
    result[0] = pt[0] + w_cross_pt[0];
    result[1] = pt[1] + w_cross_pt[1];
    result[2] = pt[2] + w_cross_pt[2];
  }
}

// ------------------------------------------------ ReprojectionError::operator() restated
// reference: src/reprojection_error.h:12-41
template <typename T>
inline void reprojection_error(const double* p_l, const double* p_r, const double uv[4],
                               const T* camera, const T* point, T* residuals) {
  T p[3];
  angle_axis_rotate_point(camera, point, p);
  p[0] = p[0] + camera[3];
  p[1] = p[1] + camera[4];
  p[2] = p[2] + camera[5];
  const T s_l_hat = 1.0 / (p_l[8] * p[0] + p_l[9] * p[1] + p_l[10] * p[2] + p_l[11] * 1.0);
  const T u_l_hat = (p_l[0] * p[0] + p_l[1] * p[1] + p_l[2] * p[2] + p_l[3] * 1.0) * s_l_hat;
  const T v_l_hat = (p_l[4] * p[0] + p_l[5] * p[1] + p_l[6] * p[2] + p_l[7] * 1.0) * s_l_hat;
  const T s_r_hat = 1.0 / (p_r[8] * p[0] + p_r[9] * p[1] + p_r[10] * p[2] + p_r[11] * 1.0);
  const T u_r_hat = (p_r[0] * p[0] + p_r[1] * p[1] + p_r[2] * p[2] + p_r[3] * 1.0) * s_r_hat;
  const T v_r_hat = (p_r[4] * p[0] + p_r[5] * p[1] + p_r[6] * p[2] + p_r[7] * 1.0) * s_r_hat;
  residuals[0] = u_l_hat - uv[0];
  residuals[1] = v_l_hat - uv[1];
  residuals[2] = u_r_hat - uv[2];
  residuals[3] = v_r_hat - uv[3];
}

// AutoDiffCostFunction<ReprojectionError, 4, 6, 3> (src/reprojection_error.h:58):
// r[4], Jc[4x6] row-major, Jp[4x3] row-major.
inline void eval_autodiff(const double* p_l, const double* p_r, const double uv[4],
                          const double* cam, const double* pt, double* r, double* Jc, double* Jp) {
  typedef Jet<9> J9;
  J9 c[6], x[3], res[4];
  for (int i = 0; i < 6; ++i) c[i] = J9(cam[i], i);
  for (int i = 0; i < 3; ++i) x[i] = J9(pt[i], 6 + i);
  reprojection_error<J9>(p_l, p_r, uv, c, x, res);
  for (int k = 0; k < 4; ++k) {
    r[k] = res[k].a;
    if (Jc) for (int i = 0; i < 6; ++i) Jc[k * 6 + i] = res[k].v[i];
    if (Jp) for (int i = 0; i < 3; ++i) Jp[k * 3 + i] = res[k].v[6 + i];
  }
}

inline void eval_residual_only(const double* p_l, const double* p_r, const double uv[4],
                               const double* cam, const double* pt, double* r) {
  reprojection_error<double>(p_l, p_r, uv, cam, pt, r);
}

// ------------------------------------------------ analytic Jacobian (cross-check of the formulas
// the CUDA kernels use; NOT how the reference computes it — the reference autodiffs).
//   p = R(w) X + t ;  d p / d t = I ; d p / d X = R ; d p / d w = -[R X]x J_l(w)
//   J_l(w) = I + (1-cos th)/th^2 [w]x + (th - sin th)/th^3 [w]x^2
inline void rot_and_left_jacobian(const double w[3], double R[9], double Jl[9]) {
  const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2];
  double A, B, C;  // sin th / th, (1-cos th)/th^2, (th - sin th)/th^3
  if (th2 < 1e-8) {
    A = 1.0 - th2 / 6.0 + th2 * th2 / 120.0;
    B = 0.5 - th2 / 24.0 + th2 * th2 / 720.0;
    C = 1.0 / 6.0 - th2 / 120.0 + th2 * th2 / 5040.0;
  } else {
    const double th = std::sqrt(th2);
    const double s = std::sin(th), c = std::cos(th);
    A = s / th;
    B = (1.0 - c) / th2;
    C = (th - s) / (th2 * th);
  }
  const double K[9] = {0, -w[2], w[1], w[2], 0, -w[0], -w[1], w[0], 0};
  double K2[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += K[i * 3 + k] * K[k * 3 + j];
      K2[i * 3 + j] = s;
    }
  for (int i = 0; i < 9; ++i) {
    const double I = (i % 4 == 0) ? 1.0 : 0.0;
    R[i] = I + A * K[i] + B * K2[i];
    Jl[i] = I + B * K[i] + C * K2[i];
  }
}

inline void eval_analytic(const double* p_l, const double* p_r, const double uv[4],
                          const double* cam, const double* pt, double* r, double* Jc, double* Jp) {
  double R[9], Jl[9];
  rot_and_left_jacobian(cam, R, Jl);
  double q[3], p[3];
  for (int i = 0; i < 3; ++i) {
    q[i] = R[i * 3 + 0] * pt[0] + R[i * 3 + 1] * pt[1] + R[i * 3 + 2] * pt[2];
    p[i] = q[i] + cam[3 + i];
  }
  double D[12];  // 4x3 d(residual)/d p
  const double* P[2] = {p_l, p_r};
  for (int s = 0; s < 2; ++s) {
    const double* M = P[s];
    const double a = M[0] * p[0] + M[1] * p[1] + M[2] * p[2] + M[3];
    const double b = M[4] * p[0] + M[5] * p[1] + M[6] * p[2] + M[7];
    const double c = M[8] * p[0] + M[9] * p[1] + M[10] * p[2] + M[11];
    const double inv = 1.0 / c;
    const double u = a * inv, v = b * inv;
    r[2 * s + 0] = u - uv[2 * s + 0];
    r[2 * s + 1] = v - uv[2 * s + 1];
    for (int k = 0; k < 3; ++k) {
      D[(2 * s + 0) * 3 + k] = (M[k] - u * M[8 + k]) * inv;
      D[(2 * s + 1) * 3 + k] = (M[4 + k] - v * M[8 + k]) * inv;
    }
  }
  // G = -[q]x Jl  (3x3)
  const double Q[9] = {0, -q[2], q[1], q[2], 0, -q[0], -q[1], q[0], 0};
  double G[9];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double s = 0;
      for (int k = 0; k < 3; ++k) s += Q[i * 3 + k] * Jl[k * 3 + j];
      G[i * 3 + j] = -s;
    }
  for (int k = 0; k < 4; ++k)
    for (int j = 0; j < 3; ++j) {
      double sw = 0, sx = 0;
      for (int i = 0; i < 3; ++i) {
        sw += D[k * 3 + i] * G[i * 3 + j];
        sx += D[k * 3 + i] * R[i * 3 + j];
      }
      if (Jc) { Jc[k * 6 + j] = sw; Jc[k * 6 + 3 + j] = D[k * 3 + j]; }
      if (Jp) Jp[k * 3 + j] = sx;
    }
}

// ------------------------------------------------ ceres::HuberLoss + Corrector restated
// rho(s): s<=d^2 -> s ; else 2 d sqrt(s) - d^2.  Objective = 1/2 sum rho(||r||^2).
// rho'' <= 0 => corrector scales r and J by sqrt(rho') (corrector.cc, alpha = 0 branch).
inline void huber(double delta, double s, double* rho, double* sqrt_rho1) {
  const double b = delta * delta;
  if (s > b) {
    const double rt = std::sqrt(s);
    *rho = 2.0 * delta * rt - b;
    const double rho1 = std::max(std::numeric_limits<double>::min(), delta / rt);
    *sqrt_rho1 = std::sqrt(rho1);
  } else {
    *rho = s;
    *sqrt_rho1 = 1.0;
  }
}

struct Problem {
  int Nc, Np, Nobs;
  const int32_t* obs_cam;
  const int32_t* obs_pt;
  const float* obs_uv;  // [Nobs*4], stored f32 upstream (src/observation.h:11-16)
  const double* p_l;
  const double* p_r;
  const uint8_t* cam_fixed;  // [Nc]; reference: only camera 0 (src/bundle_adjuster.cpp:113)
  double huber_delta;
  int rdim;  // 4 (reference) or 3 (north_star wording)
  bool analytic;
};

struct Linearization {
  std::vector<double> r;   // Nobs*4 corrected residuals
  std::vector<double> Jc;  // Nobs*24 corrected
  std::vector<double> Jp;  // Nobs*12 corrected
  double cost;
};

inline double eval_obs(const Problem& P, const double* poses, const double* points, int k,
                       double* r, double* Jc, double* Jp) {
  const double uv[4] = {(double)P.obs_uv[4 * k + 0], (double)P.obs_uv[4 * k + 1],
                        (double)P.obs_uv[4 * k + 2], (double)P.obs_uv[4 * k + 3]};
  const double* cam = poses + 6 * P.obs_cam[k];
  const double* pt = points + 3 * P.obs_pt[k];
  if (Jc) {
    if (P.analytic) eval_analytic(P.p_l, P.p_r, uv, cam, pt, r, Jc, Jp);
    else eval_autodiff(P.p_l, P.p_r, uv, cam, pt, r, Jc, Jp);
  } else {
    eval_residual_only(P.p_l, P.p_r, uv, cam, pt, r);
  }
  if (P.rdim == 3) {
    r[3] = 0.0;
    if (Jc) { for (int i = 0; i < 6; ++i) Jc[18 + i] = 0.0; for (int i = 0; i < 3; ++i) Jp[9 + i] = 0.0; }
  }
  const double s = r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3];
  double rho, sq;
  huber(P.huber_delta, s, &rho, &sq);
  for (int i = 0; i < 4; ++i) r[i] *= sq;
  if (Jc) { for (int i = 0; i < 24; ++i) Jc[i] *= sq; for (int i = 0; i < 12; ++i) Jp[i] *= sq; }
  return 0.5 * rho;
}

double evaluate_cost(const Problem& P, const double* poses, const double* points) {
  double cost = 0.0;
#pragma omp parallel for reduction(+ : cost) schedule(static)
  for (int k = 0; k < P.Nobs; ++k) {
    double r[4];
    cost += eval_obs(P, poses, points, k, r, nullptr, nullptr);
  }
  return cost;
}

void linearize(const Problem& P, const double* poses, const double* points, Linearization* L) {
  L->r.resize((size_t)P.Nobs * 4);
  L->Jc.resize((size_t)P.Nobs * 24);
  L->Jp.resize((size_t)P.Nobs * 12);
  double cost = 0.0;
#pragma omp parallel for reduction(+ : cost) schedule(static)
  for (int k = 0; k < P.Nobs; ++k)
    cost += eval_obs(P, poses, points, k, &L->r[(size_t)k * 4], &L->Jc[(size_t)k * 24], &L->Jp[(size_t)k * 12]);
  L->cost = cost;
}

// point-order and camera-order adjacency (stable: input order inside each segment)
struct Adjacency {
  std::vector<int> pt_off, pt_obs, cam_off, cam_obs;
  void build(const Problem& P) {
    pt_off.assign(P.Np + 1, 0); cam_off.assign(P.Nc + 1, 0);
    for (int k = 0; k < P.Nobs; ++k) { pt_off[P.obs_pt[k] + 1]++; cam_off[P.obs_cam[k] + 1]++; }
    for (int i = 0; i < P.Np; ++i) pt_off[i + 1] += pt_off[i];
    for (int i = 0; i < P.Nc; ++i) cam_off[i + 1] += cam_off[i];
    pt_obs.resize(P.Nobs); cam_obs.resize(P.Nobs);
    std::vector<int> pc(pt_off.begin(), pt_off.end() - 1), cc(cam_off.begin(), cam_off.end() - 1);
    for (int k = 0; k < P.Nobs; ++k) { pt_obs[pc[P.obs_pt[k]]++] = k; cam_obs[cc[P.obs_cam[k]]++] = k; }
  }
};

// Normal-equation blocks: U (Nc x 6x6), V (Np x 3x3), W per observation (6x3), gc, gp.
struct Blocks {
  std::vector<double> U, V, W, gc, gp;
};

void build_blocks(const Problem& P, const Adjacency& A, const Linearization& L, Blocks* B) {
  B->U.assign((size_t)P.Nc * 36, 0.0); B->V.assign((size_t)P.Np * 9, 0.0);
  B->W.assign((size_t)P.Nobs * 18, 0.0); B->gc.assign((size_t)P.Nc * 6, 0.0); B->gp.assign((size_t)P.Np * 3, 0.0);
#pragma omp parallel for schedule(static)
  for (int k = 0; k < P.Nobs; ++k) {
    const double* Jc = &L.Jc[(size_t)k * 24]; const double* Jp = &L.Jp[(size_t)k * 12];
    double* W = &B->W[(size_t)k * 18];
    for (int i = 0; i < 6; ++i) for (int j = 0; j < 3; ++j) {
      double s = 0; for (int m = 0; m < 4; ++m) s += Jc[m * 6 + i] * Jp[m * 3 + j]; W[i * 3 + j] = s; }
  }
#pragma omp parallel for schedule(dynamic, 1)
  for (int c = 0; c < P.Nc; ++c) {
    double* U = &B->U[(size_t)c * 36]; double* g = &B->gc[(size_t)c * 6];
    for (int e = A.cam_off[c]; e < A.cam_off[c + 1]; ++e) {
      const int k = A.cam_obs[e];
      const double* Jc = &L.Jc[(size_t)k * 24]; const double* r = &L.r[(size_t)k * 4];
      for (int i = 0; i < 6; ++i) {
        for (int j = 0; j < 6; ++j) { double s = 0; for (int m = 0; m < 4; ++m) s += Jc[m * 6 + i] * Jc[m * 6 + j]; U[i * 6 + j] += s; }
        double s = 0; for (int m = 0; m < 4; ++m) s += Jc[m * 6 + i] * r[m]; g[i] += s;
      }
    }
  }
#pragma omp parallel for schedule(static)
  for (int p = 0; p < P.Np; ++p) {
    double* V = &B->V[(size_t)p * 9]; double* g = &B->gp[(size_t)p * 3];
    for (int e = A.pt_off[p]; e < A.pt_off[p + 1]; ++e) {
      const int k = A.pt_obs[e];
      const double* Jp = &L.Jp[(size_t)k * 12]; const double* r = &L.r[(size_t)k * 4];
      for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) { double s = 0; for (int m = 0; m < 4; ++m) s += Jp[m * 3 + i] * Jp[m * 3 + j]; V[i * 3 + j] += s; }
        double s = 0; for (int m = 0; m < 4; ++m) s += Jp[m * 3 + i] * r[m]; g[i] += s;
      }
    }
  }
}

inline bool invert_spd3(const double* M, double* Mi) {
  const double a = M[0], b = M[1], c = M[2], d = M[4], e = M[5], f = M[8];
  const double c00 = d * f - e * e, c01 = c * e - b * f, c02 = b * e - c * d;
  const double det = a * c00 + b * c01 + c * c02;
  if (!(det > 0.0) || !std::isfinite(det)) return false;
  const double id = 1.0 / det;
  Mi[0] = c00 * id; Mi[1] = c01 * id; Mi[2] = c02 * id;
  Mi[3] = Mi[1]; Mi[4] = (a * f - c * c) * id; Mi[5] = (b * c - a * e) * id;
  Mi[6] = Mi[2]; Mi[7] = Mi[5]; Mi[8] = (a * d - b * b) * id;
  return true;
}

// dense in-place Cholesky (lower), row-major n x n. Returns false if not SPD.
bool cholesky(std::vector<double>& A, int n) {
  for (int j = 0; j < n; ++j) {
    double d = A[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) d -= A[(size_t)j * n + k] * A[(size_t)j * n + k];
    if (!(d > 0.0) || !std::isfinite(d)) return false;
    d = std::sqrt(d);
    A[(size_t)j * n + j] = d;
    const double id = 1.0 / d;
#pragma omp parallel for schedule(static) if (n - j > 256)
    for (int i = j + 1; i < n; ++i) {
      double s = A[(size_t)i * n + j];
      const double* ai = &A[(size_t)i * n]; const double* aj = &A[(size_t)j * n];
      for (int k = 0; k < j; ++k) s -= ai[k] * aj[k];
      A[(size_t)i * n + j] = s * id;
    }
  }
  return true;
}
void cholesky_solve(const std::vector<double>& Lm, int n, double* b) {
  for (int i = 0; i < n; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= Lm[(size_t)i * n + k] * b[k]; b[i] = s / Lm[(size_t)i * n + i]; }
  for (int i = n - 1; i >= 0; --i) { double s = b[i]; for (int k = i + 1; k < n; ++k) s -= Lm[(size_t)k * n + i] * b[k]; b[i] = s / Lm[(size_t)i * n + i]; }
}

// Schur complement with per-parameter damping Dc (6 per camera), Dp (3 per point):
//   S = (U + Dc) - sum_p W (V+Dp)^-1 W^T   over FREE cameras ;  rhs = -(gc - sum W (V+Dp)^-1 gp)
// dense S (n = 6*Nfree). Returns false on singular point block.
bool schur(const Problem& P, const Adjacency& A, const Blocks& B, const std::vector<int>& fc, int Nf,
           const double* Dc, const double* Dp, std::vector<double>* S, std::vector<double>* rhs,
           std::vector<double>* Vinv) {
  const int n = 6 * Nf;
  S->assign((size_t)n * n, 0.0); rhs->assign(n, 0.0); Vinv->assign((size_t)P.Np * 9, 0.0);
  for (int c = 0; c < P.Nc; ++c) {
    if (fc[c] < 0) continue;
    const int o = 6 * fc[c];
    for (int i = 0; i < 6; ++i) {
      for (int j = 0; j < 6; ++j) (*S)[(size_t)(o + i) * n + o + j] = B.U[(size_t)c * 36 + i * 6 + j];
      (*S)[(size_t)(o + i) * n + o + i] += Dc[6 * c + i];
      (*rhs)[o + i] = -B.gc[6 * c + i];
    }
  }
  bool ok = true;
#pragma omp parallel for schedule(static)
  for (int p = 0; p < P.Np; ++p) {
    double M[9];
    for (int i = 0; i < 9; ++i) M[i] = B.V[(size_t)p * 9 + i];
    for (int i = 0; i < 3; ++i) M[i * 4] += Dp[3 * p + i];
    if (!invert_spd3(M, &(*Vinv)[(size_t)p * 9])) {
#pragma omp atomic write
      ok = false;
    }
  }
  if (!ok) return false;
  int nthreads = 1;
#ifdef _OPENMP
  nthreads = omp_get_max_threads();
#endif
  if ((size_t)n * n * nthreads > (size_t)1 << 27) nthreads = 1;  // keep private copies small
  std::vector<std::vector<double>> Sp(nthreads), rp(nthreads);
#pragma omp parallel num_threads(nthreads)
  {
    int t = 0;
#ifdef _OPENMP
    t = omp_get_thread_num();
#endif
    std::vector<double>& Sl = Sp[t]; std::vector<double>& rl = rp[t];
    Sl.assign((size_t)n * n, 0.0); rl.assign(n, 0.0);
#pragma omp for schedule(static)
    for (int p = 0; p < P.Np; ++p) {
      const double* Vi = &(*Vinv)[(size_t)p * 9];
      const double* gp = &B.gp[(size_t)p * 3];
      double tp[3];
      for (int i = 0; i < 3; ++i) tp[i] = Vi[i * 3] * gp[0] + Vi[i * 3 + 1] * gp[1] + Vi[i * 3 + 2] * gp[2];
      for (int ea = A.pt_off[p]; ea < A.pt_off[p + 1]; ++ea) {
        const int ka = A.pt_obs[ea]; const int fa = fc[P.obs_cam[ka]];
        if (fa < 0) continue;
        const double* Wa = &B.W[(size_t)ka * 18];
        double Y[18];  // Wa * Vinv
        for (int i = 0; i < 6; ++i) for (int j = 0; j < 3; ++j)
          Y[i * 3 + j] = Wa[i * 3] * Vi[j] + Wa[i * 3 + 1] * Vi[3 + j] + Wa[i * 3 + 2] * Vi[6 + j];
        for (int i = 0; i < 6; ++i) rl[6 * fa + i] += Wa[i * 3] * tp[0] + Wa[i * 3 + 1] * tp[1] + Wa[i * 3 + 2] * tp[2];
        for (int eb = A.pt_off[p]; eb < A.pt_off[p + 1]; ++eb) {
          const int kb = A.pt_obs[eb]; const int fb = fc[P.obs_cam[kb]];
          if (fb < 0) continue;
          const double* Wb = &B.W[(size_t)kb * 18];
          for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j)
            Sl[(size_t)(6 * fa + i) * n + 6 * fb + j] -= Y[i * 3] * Wb[j * 3] + Y[i * 3 + 1] * Wb[j * 3 + 1] + Y[i * 3 + 2] * Wb[j * 3 + 2];
        }
      }
    }
  }
  for (int t = 0; t < nthreads; ++t) {
    for (size_t i = 0; i < (size_t)n * n; ++i) (*S)[i] += Sp[t][i];
    for (int i = 0; i < n; ++i) (*rhs)[i] += rp[t][i];
  }
  return true;
}

inline double clampd(double x, double lo, double hi) { return x < lo ? lo : (x > hi ? hi : x); }

// ---- envelope (skyline) form of the reduced camera matrix, for global problems (thousands of cameras) -----------
// Ceres factors the reduced camera matrix with a sparse Cholesky (CHOLMOD, src/bundle_adjuster.cpp:26-29); the dense
// LLT above is the same solution but O(n^2) memory.  In the reference's frame order the matrix is banded
// (co-visibility reaches a few frames back) plus loop-closure rows, so a row-wise envelope — row i stores columns
// first[i]..i of the lower triangle — holds it exactly, fill-in included, and Jennings' row Cholesky factors it in
// place.  Same arithmetic as the dense LLT up to summation order.
struct Envelope {
  int n = 0;
  std::vector<int> first;        // [n] first stored column of row i
  std::vector<size_t> off;       // [n+1]
  std::vector<double> v;
  double* row(int i) { return v.data() + off[i] - first[i]; }             // row(i)[j] for first[i] <= j <= i
  const double* row(int i) const { return v.data() + off[i] - first[i]; }
  void build(const Problem& P, const Adjacency& A, const std::vector<int>& fc, int Nf) {
    n = 6 * Nf;
    std::vector<int> fb(Nf);
    for (int f = 0; f < Nf; ++f) fb[f] = f;
    for (int p = 0; p < P.Np; ++p) {
      int mn = Nf;
      for (int e = A.pt_off[p]; e < A.pt_off[p + 1]; ++e) { const int f = fc[P.obs_cam[A.pt_obs[e]]]; if (f >= 0) mn = std::min(mn, f); }
      for (int e = A.pt_off[p]; e < A.pt_off[p + 1]; ++e) { const int f = fc[P.obs_cam[A.pt_obs[e]]]; if (f >= 0) fb[f] = std::min(fb[f], mn); }
    }
    first.resize(n); off.assign(n + 1, 0);
    for (int i = 0; i < n; ++i) { first[i] = 6 * fb[i / 6]; off[i + 1] = off[i] + (size_t)(i - first[i] + 1); }
    v.assign(off[n], 0.0);
  }
};

bool envelope_cholesky(Envelope& E) {
  for (int i = 0; i < E.n; ++i) {
    double* ri = E.row(i);
    for (int j = E.first[i]; j <= i; ++j) {
      const double* rj = E.row(j);
      double s = ri[j];
      for (int k = std::max(E.first[i], E.first[j]); k < j; ++k) s -= ri[k] * rj[k];
      if (j < i) ri[j] = s / rj[j];
      else { if (!(s > 0.0) || !std::isfinite(s)) return false; ri[i] = std::sqrt(s); }
    }
  }
  return true;
}
void envelope_solve(const Envelope& E, double* b) {
  for (int i = 0; i < E.n; ++i) { const double* ri = E.row(i); double s = b[i]; for (int k = E.first[i]; k < i; ++k) s -= ri[k] * b[k]; b[i] = s / ri[i]; }
  for (int i = E.n - 1; i >= 0; --i) { const double* ri = E.row(i); const double x = b[i] / ri[i]; b[i] = x; for (int k = E.first[i]; k < i; ++k) b[k] -= ri[k] * x; }
}

// the same damped Schur complement as schur(), accumulated into the envelope (lower triangle only)
bool schur_envelope(const Problem& P, const Adjacency& A, const Blocks& B, const std::vector<int>& fc, int Nf,
                    const double* Dc, const double* Dp, Envelope* E, std::vector<double>* rhs, std::vector<double>* Vinv) {
  const int n = 6 * Nf;
  std::fill(E->v.begin(), E->v.end(), 0.0);
  rhs->assign(n, 0.0); Vinv->assign((size_t)P.Np * 9, 0.0);
  for (int c = 0; c < P.Nc; ++c) {
    if (fc[c] < 0) continue;
    const int o = 6 * fc[c];
    for (int i = 0; i < 6; ++i) {
      for (int j = 0; j <= i; ++j) E->row(o + i)[o + j] = B.U[(size_t)c * 36 + i * 6 + j];
      E->row(o + i)[o + i] += Dc[6 * c + i];
      (*rhs)[o + i] = -B.gc[6 * c + i];
    }
  }
  bool ok = true;
#pragma omp parallel for schedule(static)
  for (int p = 0; p < P.Np; ++p) {
    double M[9];
    for (int i = 0; i < 9; ++i) M[i] = B.V[(size_t)p * 9 + i];
    for (int i = 0; i < 3; ++i) M[i * 4] += Dp[3 * p + i];
    if (!invert_spd3(M, &(*Vinv)[(size_t)p * 9])) {
#pragma omp atomic write
      ok = false;
    }
  }
  if (!ok) return false;
  int nthreads = 1;
#ifdef _OPENMP
  nthreads = omp_get_max_threads();
#endif
  if (E->v.size() * (size_t)nthreads > (size_t)1 << 27) nthreads = 1;
  std::vector<Envelope> Ep(nthreads, *E);
  std::vector<std::vector<double>> rp(nthreads);
#pragma omp parallel num_threads(nthreads)
  {
    int t = 0;
#ifdef _OPENMP
    t = omp_get_thread_num();
#endif
    Envelope& El = Ep[t]; std::vector<double>& rl = rp[t];
    std::fill(El.v.begin(), El.v.end(), 0.0); rl.assign(n, 0.0);
#pragma omp for schedule(static)
    for (int p = 0; p < P.Np; ++p) {
      const double* Vi = &(*Vinv)[(size_t)p * 9];
      const double* gp = &B.gp[(size_t)p * 3];
      double tp[3];
      for (int i = 0; i < 3; ++i) tp[i] = Vi[i * 3] * gp[0] + Vi[i * 3 + 1] * gp[1] + Vi[i * 3 + 2] * gp[2];
      for (int ea = A.pt_off[p]; ea < A.pt_off[p + 1]; ++ea) {
        const int ka = A.pt_obs[ea]; const int fa = fc[P.obs_cam[ka]];
        if (fa < 0) continue;
        const double* Wa = &B.W[(size_t)ka * 18];
        double Y[18];
        for (int i = 0; i < 6; ++i) for (int j = 0; j < 3; ++j)
          Y[i * 3 + j] = Wa[i * 3] * Vi[j] + Wa[i * 3 + 1] * Vi[3 + j] + Wa[i * 3 + 2] * Vi[6 + j];
        for (int i = 0; i < 6; ++i) rl[6 * fa + i] += Wa[i * 3] * tp[0] + Wa[i * 3 + 1] * tp[1] + Wa[i * 3 + 2] * tp[2];
        for (int eb = A.pt_off[p]; eb < A.pt_off[p + 1]; ++eb) {
          const int kb = A.pt_obs[eb]; const int fb = fc[P.obs_cam[kb]];
          if (fb < 0 || fb > fa) continue;                      // lower triangle
          const double* Wb = &B.W[(size_t)kb * 18];
          for (int i = 0; i < 6; ++i) {
            double* r = El.row(6 * fa + i);
            const int jmax = (fb == fa) ? i : 5;
            for (int j = 0; j <= jmax; ++j) r[6 * fb + j] -= Y[i * 3] * Wb[j * 3] + Y[i * 3 + 1] * Wb[j * 3 + 1] + Y[i * 3 + 2] * Wb[j * 3 + 2];
          }
        }
      }
    }
  }
  for (int t = 0; t < nthreads; ++t) {
    for (size_t i = 0; i < E->v.size(); ++i) E->v[i] += Ep[t].v[i];
    for (int i = 0; i < n; ++i) (*rhs)[i] += rp[t][i];
  }
  return true;
}

// reduced systems above this size go through the envelope (ORACLE_ENVELOPE=1 / 0 forces it on / off: tests)
inline bool use_envelope(int n) {
  if (const char* e = std::getenv("ORACLE_ENVELOPE")) return std::atoi(e) != 0;
  return n > 1500;
}

// ---- polynomial interpolation for the Armijo line search (ceres/polynomial.cc + line_search.cc restated) ----
// FunctionSample: value always valid here; gradient = directional derivative along the search direction.
struct Sample { double x, f, g; bool has_g; };

// FindInterpolatingPolynomial: lowest-degree polynomial through all values / gradients, highest power first.
bool fit_polynomial(const std::vector<Sample>& s, std::vector<double>* coef) {
  int neq = 0;
  for (auto& a : s) neq += a.has_g ? 2 : 1;
  const int deg = neq - 1;
  std::vector<double> M((size_t)neq * neq, 0.0), rhs(neq, 0.0);
  int row = 0;
  for (auto& a : s) {
    for (int j = 0; j <= deg; ++j) M[(size_t)row * neq + j] = std::pow(a.x, deg - j);
    rhs[row++] = a.f;
    if (a.has_g) {
      for (int j = 0; j < deg; ++j) M[(size_t)row * neq + j] = (deg - j) * std::pow(a.x, deg - j - 1);
      rhs[row++] = a.g;
    }
  }
  // LU with full pivoting (Ceres: lhs.fullPivLu().solve(rhs))
  std::vector<int> colperm(neq);
  for (int i = 0; i < neq; ++i) colperm[i] = i;
  for (int k = 0; k < neq; ++k) {
    int pr = k, pc = k; double best = -1.0;
    for (int i = k; i < neq; ++i) for (int j2 = k; j2 < neq; ++j2) { const double v = std::fabs(M[(size_t)i * neq + j2]); if (v > best) { best = v; pr = i; pc = j2; } }
    if (!(best > 0.0)) return false;
    if (pr != k) { for (int j2 = 0; j2 < neq; ++j2) std::swap(M[(size_t)k * neq + j2], M[(size_t)pr * neq + j2]); std::swap(rhs[k], rhs[pr]); }
    if (pc != k) { for (int i = 0; i < neq; ++i) std::swap(M[(size_t)i * neq + k], M[(size_t)i * neq + pc]); std::swap(colperm[k], colperm[pc]); }
    for (int i = k + 1; i < neq; ++i) {
      const double f = M[(size_t)i * neq + k] / M[(size_t)k * neq + k];
      for (int j2 = k; j2 < neq; ++j2) M[(size_t)i * neq + j2] -= f * M[(size_t)k * neq + j2];
      rhs[i] -= f * rhs[k];
    }
  }
  std::vector<double> y(neq);
  for (int i = neq - 1; i >= 0; --i) { double acc = rhs[i]; for (int j2 = i + 1; j2 < neq; ++j2) acc -= M[(size_t)i * neq + j2] * y[j2]; y[i] = acc / M[(size_t)i * neq + i]; }
  coef->assign(neq, 0.0);
  for (int i = 0; i < neq; ++i) (*coef)[colperm[i]] = y[i];
  return true;
}
inline double poly_eval(const std::vector<double>& c, double x) { double v = 0; for (double a : c) v = v * x + a; return v; }

// FindPolynomialRoots: REAL PARTS of all roots (Ceres keeps the real part of every companion-matrix eigenvalue).
// Degrees 1-2 closed form as in polynomial.cc; higher degrees by Laguerre iteration with deflation (complex).
void polynomial_root_real_parts(std::vector<double> c, std::vector<double>* out) {
  out->clear();
  while (c.size() > 1 && c[0] == 0.0) c.erase(c.begin());
  const int deg = (int)c.size() - 1;
  if (deg <= 0) return;
  if (deg == 1) { out->push_back(-c[1] / c[0]); return; }
  if (deg == 2) {
    const double a = c[0], b = c[1], cc = c[2];
    const double D = b * b - 4 * a * cc, sD = std::sqrt(std::fabs(D));
    if (D >= 0) {
      if (b >= 0) { out->push_back((-b - sD) / (2 * a)); out->push_back((2 * cc) / (-b - sD)); }
      else { out->push_back((2 * cc) / (-b + sD)); out->push_back((-b + sD) / (2 * a)); }
    } else { out->push_back(-b / (2 * a)); out->push_back(-b / (2 * a)); }
    return;
  }
  typedef std::complex<double> cd;
  std::vector<cd> a(c.begin(), c.end());           // highest power first
  std::vector<cd> work = a;
  std::vector<cd> roots;
  for (int m = deg; m >= 1; --m) {
    cd x(0.0, 0.0);
    for (int it = 0; it < 200; ++it) {
      cd b = work[0], d(0, 0), f(0, 0);
      for (int j = 1; j <= m; ++j) { f = x * f + d; d = x * d + b; b = x * b + work[j]; }
      if (std::abs(b) == 0.0) break;
      const cd g = d / b, g2 = g * g, h = g2 - 2.0 * f / b;
      const cd sq = std::sqrt(double(m - 1) * (double(m) * h - g2));
      cd gp = g + sq, gm = g - sq;
      if (std::abs(gp) < std::abs(gm)) gp = gm;
      const cd dx = (std::abs(gp) > 0.0) ? double(m) / gp : std::polar(1.0 + std::abs(x), (double)it);
      const cd x1 = x - dx;
      if (x == x1) break;
      x = x1;
    }
    // polish on the undeflated polynomial (two Newton steps)
    for (int it = 0; it < 2; ++it) {
      cd b = a[0], d(0, 0);
      for (int j = 1; j <= deg; ++j) { d = x * d + b; b = x * b + a[j]; }
      if (std::abs(d) > 0.0) x -= b / d;
    }
    roots.push_back(x);
    std::vector<cd> q(m);                             // deflate: synthetic division by (z - x)
    q[0] = work[0];
    for (int j = 1; j < m; ++j) q[j] = work[j] + x * q[j - 1];
    work = q;
  }
  for (auto& r : roots) out->push_back(r.real());
}

// MinimizeInterpolatingPolynomial (MinimizePolynomial): midpoint, both ends, real parts of the derivative's roots.
double minimize_interpolating(const std::vector<Sample>& s, double xmin, double xmax) {
  std::vector<double> c;
  double best_x = 0.5 * (xmin + xmax);
  if (!fit_polynomial(s, &c)) return best_x;
  double best_v = poly_eval(c, best_x);
  const double vmin = poly_eval(c, xmin), vmax = poly_eval(c, xmax);
  if (vmin < best_v) { best_v = vmin; best_x = xmin; }
  if (vmax < best_v) { best_v = vmax; best_x = xmax; }
  const int deg = (int)c.size() - 1;
  if (deg < 2) return best_x;
  std::vector<double> dc(deg), roots;
  for (int j = 0; j < deg; ++j) dc[j] = (deg - j) * c[j];
  polynomial_root_real_parts(dc, &roots);
  for (double r0 : roots) {
    if (!(r0 >= xmin) || !(r0 <= xmax)) continue;
    const double v = poly_eval(c, r0);
    if (v < best_v) { best_v = v; best_x = r0; }
  }
  return best_x;
}

}  // namespace

extern "C" {

struct oracle_options {
  int32_t max_iterations;       // 50   (src/params.h:34)
  double max_time_s;            // 1.0  (src/params.h:41); <=0 => unlimited (the :40 alternative)
  double huber_delta;           // 1.0  (src/bundle_adjuster.cpp:100)
  double point_lower_bound;     // -1e4 (src/params.h:44)
  double point_upper_bound;     // +1e4 (src/params.h:47)
  double parameter_tolerance;   // 1e-8 (Ceres default)
  double function_tolerance;    // 1e-16 (src/bundle_adjuster.cpp:36)
  double gradient_tolerance;    // 1e-16 (src/bundle_adjuster.cpp:35)
  int32_t residual_dim;         // 4
  double initial_radius;        // 1e4 (Ceres default)
  int32_t line_search;          // 1: projected Armijo line search as Ceres does when bounds are set
  int32_t num_threads;          // 4 = reference (src/params.h:37); 0 => all cores
  int32_t analytic_jacobian;    // 0: Jet autodiff (reference behaviour); 1: analytic (cross-check)
  int32_t verbose;
};

struct oracle_summary {
  double initial_cost, final_cost;
  int32_t iterations;        // LM iterations performed (excluding iteration 0)
  int32_t successful_steps;
  int32_t linearizations;    // Jacobian evaluations
  int32_t termination;       // 0 CONVERGENCE, 1 NO_CONVERGENCE, 2 FAILURE
  int32_t reason;            // 1 param tol, 2 function tol, 3 gradient tol, 4 max iter, 5 max time, 6 min radius, 7 invalid steps
  double time_total_s, time_linearize_s, time_schur_s, time_solve_s;
  double final_radius, final_gradient_max_norm;
};

void oracle_default_options(oracle_options* o) {
  o->max_iterations = 50; o->max_time_s = 1.0; o->huber_delta = 1.0;
  o->point_lower_bound = -10000.0; o->point_upper_bound = 10000.0;
  o->parameter_tolerance = 1e-8; o->function_tolerance = 1e-16; o->gradient_tolerance = 1e-16;
  o->residual_dim = 4; o->initial_radius = 1e4; o->line_search = 1; o->num_threads = 4;
  o->analytic_jacobian = 0; o->verbose = 0;
}

int oracle_num_procs() {
#ifdef _OPENMP
  return omp_get_num_procs();
#else
  return 1;
#endif
}

// single residual block, uncorrected (no loss): r[4], Jc[24], Jp[12].  mode 0 = Jet autodiff, 1 = analytic
void oracle_residual(const double* cam, const double* pt, const double* uv, const double* p_l,
                     const double* p_r, int mode, double* r, double* Jc, double* Jp) {
  if (mode == 0) eval_autodiff(p_l, p_r, uv, cam, pt, r, Jc, Jp);
  else eval_analytic(p_l, p_r, uv, cam, pt, r, Jc, Jp);
}

void oracle_huber(double delta, double s, double* rho, double* sqrt_rho1) { huber(delta, s, rho, sqrt_rho1); }

// line-search step selection on explicit samples (tests): returns argmin of the interpolating polynomial
double oracle_ls_minimize(int n, const double* x, const double* f, const double* g, const int32_t* has_g, double xmin, double xmax) {
  std::vector<Sample> s;
  for (int i = 0; i < n; ++i) s.push_back(Sample{x[i], f[i], g[i], has_g[i] != 0});
  return minimize_interpolating(s, xmin, xmax);
}

// Linearise: corrected residuals and normal-equation blocks.
int oracle_linearize(int Nc, int Np, int Nobs, const double* poses, const double* points,
                     const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv,
                     const double* p_l, const double* p_r, double huber_delta, int rdim, int analytic,
                     double* r_out, double* cost_out, double* U, double* V, double* W, double* gc, double* gp) {
  Problem P{Nc, Np, Nobs, obs_cam, obs_pt, obs_uv, p_l, p_r, nullptr, huber_delta, rdim, analytic != 0};
  Adjacency A; A.build(P);
  Linearization L; linearize(P, poses, points, &L);
  Blocks B; build_blocks(P, A, L, &B);
  if (r_out) std::memcpy(r_out, L.r.data(), L.r.size() * 8);
  if (cost_out) *cost_out = L.cost;
  if (U) std::memcpy(U, B.U.data(), B.U.size() * 8);
  if (V) std::memcpy(V, B.V.data(), B.V.size() * 8);
  if (W) std::memcpy(W, B.W.data(), B.W.size() * 8);
  if (gc) std::memcpy(gc, B.gc.data(), B.gc.size() * 8);
  if (gp) std::memcpy(gp, B.gp.data(), B.gp.size() * 8);
  return 0;
}

double oracle_cost(int Nc, int Np, int Nobs, const double* poses, const double* points,
                   const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv,
                   const double* p_l, const double* p_r, double huber_delta, int rdim) {
  Problem P{Nc, Np, Nobs, obs_cam, obs_pt, obs_uv, p_l, p_r, nullptr, huber_delta, rdim, false};
  return evaluate_cost(P, poses, points);
}

// Dense damped Schur complement over free cameras for a given radius (no Jacobi scaling:
// damping = clamp(diag, 1e-6, 1e32) / radius).  S is (6Nf)^2 row-major, rhs 6Nf.
int oracle_schur(int Nc, int Np, int Nobs, const double* poses, const double* points,
                 const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv,
                 const double* p_l, const double* p_r, const uint8_t* cam_fixed, double huber_delta,
                 int rdim, double radius, double* S_out, double* rhs_out) {
  Problem P{Nc, Np, Nobs, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed, huber_delta, rdim, false};
  Adjacency A; A.build(P);
  Linearization L; linearize(P, poses, points, &L);
  Blocks B; build_blocks(P, A, L, &B);
  std::vector<int> fc(Nc); int Nf = 0;
  for (int c = 0; c < Nc; ++c) fc[c] = cam_fixed[c] ? -1 : Nf++;
  std::vector<double> Dc((size_t)Nc * 6), Dp((size_t)Np * 3);
  for (int c = 0; c < Nc; ++c) for (int i = 0; i < 6; ++i) Dc[6 * c + i] = clampd(B.U[(size_t)c * 36 + i * 7], 1e-6, 1e32) / radius;
  for (int p = 0; p < Np; ++p) for (int i = 0; i < 3; ++i) Dp[3 * p + i] = clampd(B.V[(size_t)p * 9 + i * 4], 1e-6, 1e32) / radius;
  std::vector<double> S, rhs, Vinv;
  if (!schur(P, A, B, fc, Nf, Dc.data(), Dp.data(), &S, &rhs, &Vinv)) return 1;
  std::memcpy(S_out, S.data(), S.size() * 8);
  std::memcpy(rhs_out, rhs.data(), rhs.size() * 8);
  return 0;
}

// Structural non-zero blocks of the reduced camera matrix over free cameras:
// (i,j) non-zero iff some point is observed by both free cameras i and j.  Sorted (i,j), full
// (both triangles + diagonal of every free camera that has >= 1 observation).
int oracle_schur_pattern(int Nc, int Np, int Nobs, const int32_t* obs_cam, const int32_t* obs_pt,
                         const uint8_t* cam_fixed, int32_t* rows, int32_t* cols, int capacity) {
  std::vector<int> fc(Nc); int Nf = 0;
  for (int c = 0; c < Nc; ++c) fc[c] = cam_fixed[c] ? -1 : Nf++;
  std::vector<std::vector<int>> per_pt(Np);
  for (int k = 0; k < Nobs; ++k) if (fc[obs_cam[k]] >= 0) per_pt[obs_pt[k]].push_back(fc[obs_cam[k]]);
  std::set<std::pair<int, int>> blocks;
  for (int p = 0; p < Np; ++p) for (int a : per_pt[p]) for (int b : per_pt[p]) blocks.insert({a, b});
  int n = 0;
  for (auto& b : blocks) { if (n < capacity) { rows[n] = b.first; cols[n] = b.second; } ++n; }
  return n;
}

// The full Ceres-rule LM solve.  poses [Nc*6] and points [Np*3] are updated in place.
int oracle_solve(int Nc, int Np, int Nobs, double* poses, double* points, const int32_t* obs_cam,
                 const int32_t* obs_pt, const float* obs_uv, const double* p_l, const double* p_r,
                 const uint8_t* cam_fixed, const oracle_options* opt, oracle_summary* sum) {
  using clk = std::chrono::steady_clock;
  auto t0 = clk::now();
  auto secs = [](clk::time_point a, clk::time_point b) { return std::chrono::duration<double>(b - a).count(); };
#ifdef _OPENMP
  const int saved_threads = omp_get_max_threads();
  omp_set_num_threads(opt->num_threads > 0 ? opt->num_threads : omp_get_num_procs());
#endif
  std::memset(sum, 0, sizeof(*sum));
  Problem P{Nc, Np, Nobs, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed, opt->huber_delta, opt->residual_dim, opt->analytic_jacobian != 0};
  Adjacency A; A.build(P);
  std::vector<int> fc(Nc); int Nf = 0;
  for (int c = 0; c < Nc; ++c) fc[c] = (cam_fixed && cam_fixed[c]) ? -1 : Nf++;
  const double lo = opt->point_lower_bound, hi = opt->point_upper_bound;
  const size_t nc6 = (size_t)Nc * 6, np3 = (size_t)Np * 3;

  // iteration zero: project onto the bounds, evaluate
  for (size_t i = 0; i < np3; ++i) points[i] = clampd(points[i], lo, hi);
  Linearization L; Blocks B;
  auto tl = clk::now();
  linearize(P, poses, points, &L); build_blocks(P, A, L, &B);
  sum->time_linearize_s += secs(tl, clk::now()); sum->linearizations = 1;
  double cost = L.cost; sum->initial_cost = cost;

  // Jacobi scaling, fixed for the whole solve: 1 / (1 + ||J_col||)
  std::vector<double> sc2c(nc6), sc2p(np3);  // squared scales
  for (size_t c = 0; c < (size_t)Nc; ++c) for (int i = 0; i < 6; ++i) { const double s = 1.0 / (1.0 + std::sqrt(B.U[c * 36 + i * 7])); sc2c[6 * c + i] = s * s; }
  for (size_t p = 0; p < (size_t)Np; ++p) for (int i = 0; i < 3; ++i) { const double s = 1.0 / (1.0 + std::sqrt(B.V[p * 9 + i * 4])); sc2p[3 * p + i] = s * s; }

  auto gradient_max_norm = [&]() {
    double m = 0.0;
    for (int c = 0; c < Nc; ++c) if (fc[c] >= 0) for (int i = 0; i < 6; ++i) m = std::max(m, std::fabs(B.gc[6 * c + i]));
    for (size_t i = 0; i < np3; ++i) { const double x = points[i]; m = std::max(m, std::fabs(x - clampd(x - B.gp[i], lo, hi))); }
    return m;
  };
  auto x_norm = [&]() {
    double s = 0.0;
    for (int c = 0; c < Nc; ++c) if (fc[c] >= 0) for (int i = 0; i < 6; ++i) s += poses[6 * c + i] * poses[6 * c + i];
    for (size_t i = 0; i < np3; ++i) s += points[i] * points[i];
    return std::sqrt(s);
  };

  double radius = opt->initial_radius, decrease_factor = 2.0;
  int term = 1, reason = 4, invalid = 0;
  double gmax = gradient_max_norm();
  if (gmax <= opt->gradient_tolerance) { term = 0; reason = 3; }
  std::vector<double> Dc(nc6), Dp(np3), S, rhs, Vinv, dc(nc6), dp(np3), cposes(nc6), cpoints(np3);
  const bool envelope = use_envelope(6 * Nf);
  Envelope E;
  if (envelope) E.build(P, A, fc, Nf);
  int iter = 0;
  while (term == 1) {
    if (iter >= opt->max_iterations) { reason = 4; break; }
    if (opt->max_time_s > 0 && secs(t0, clk::now()) >= opt->max_time_s) { reason = 5; break; }
    if (radius < 1e-32) { term = 0; reason = 6; break; }
    ++iter;
    // LM diagonal in the Jacobi-scaled space, mapped back: D = clamp(s^2 H_jj, 1e-6, 1e32) / (s^2 radius)
    for (size_t c = 0; c < (size_t)Nc; ++c) for (int i = 0; i < 6; ++i)
      Dc[6 * c + i] = clampd(sc2c[6 * c + i] * B.U[c * 36 + i * 7], 1e-6, 1e32) / (sc2c[6 * c + i] * radius);
    for (size_t p = 0; p < (size_t)Np; ++p) for (int i = 0; i < 3; ++i)
      Dp[3 * p + i] = clampd(sc2p[3 * p + i] * B.V[p * 9 + i * 4], 1e-6, 1e32) / (sc2p[3 * p + i] * radius);
    auto ts = clk::now();
    const int n = 6 * Nf;
    bool ok;
    if (envelope) {
      ok = schur_envelope(P, A, B, fc, Nf, Dc.data(), Dp.data(), &E, &rhs, &Vinv);
      sum->time_schur_s += secs(ts, clk::now());
      auto tv = clk::now();
      if (ok && n > 0) { ok = envelope_cholesky(E); if (ok) envelope_solve(E, rhs.data()); }
      sum->time_solve_s += secs(tv, clk::now());
    } else {
      ok = schur(P, A, B, fc, Nf, Dc.data(), Dp.data(), &S, &rhs, &Vinv);
      sum->time_schur_s += secs(ts, clk::now());
      auto tv = clk::now();
      if (ok && n > 0) { ok = cholesky(S, n); if (ok) cholesky_solve(S, n, rhs.data()); }
      sum->time_solve_s += secs(tv, clk::now());
    }
    double model_change = 0.0;
    if (ok) {
      std::fill(dc.begin(), dc.end(), 0.0);
      for (int c = 0; c < Nc; ++c) if (fc[c] >= 0) for (int i = 0; i < 6; ++i) dc[6 * c + i] = rhs[6 * fc[c] + i];
      // back-substitute: dp = -(V+Dp)^-1 (gp + W^T dc)
#pragma omp parallel for schedule(static)
      for (int p = 0; p < Np; ++p) {
        double t[3] = {B.gp[3 * p], B.gp[3 * p + 1], B.gp[3 * p + 2]};
        for (int e = A.pt_off[p]; e < A.pt_off[p + 1]; ++e) {
          const int k = A.pt_obs[e]; const int c = obs_cam[k];
          if (fc[c] < 0) continue;
          const double* W = &B.W[(size_t)k * 18];
          for (int i = 0; i < 6; ++i) for (int j = 0; j < 3; ++j) t[j] += W[i * 3 + j] * dc[6 * c + i];
        }
        const double* Vi = &Vinv[(size_t)p * 9];
        for (int i = 0; i < 3; ++i) dp[3 * p + i] = -(Vi[i * 3] * t[0] + Vi[i * 3 + 1] * t[1] + Vi[i * 3 + 2] * t[2]);
      }
      // model_cost_change = -(J d)^T (r + J d / 2)   (trust_region_minimizer.cc)
      double mc = 0.0;
#pragma omp parallel for reduction(+ : mc) schedule(static)
      for (int k = 0; k < Nobs; ++k) {
        const double* Jc = &L.Jc[(size_t)k * 24]; const double* Jp = &L.Jp[(size_t)k * 12]; const double* r = &L.r[(size_t)k * 4];
        const double* a = &dc[6 * (size_t)obs_cam[k]]; const double* b = &dp[3 * (size_t)obs_pt[k]];
        for (int m = 0; m < 4; ++m) {
          double jd = 0; for (int i = 0; i < 6; ++i) jd += Jc[m * 6 + i] * a[i]; for (int i = 0; i < 3; ++i) jd += Jp[m * 3 + i] * b[i];
          mc -= jd * (r[m] + 0.5 * jd);
        }
      }
      model_change = mc;
      for (size_t i = 0; i < nc6; ++i) if (!std::isfinite(dc[i])) ok = false;
    }
    if (!ok || !(model_change > 0.0)) {
      if (++invalid >= 5) { term = 2; reason = 7; break; }
      radius *= 0.5;
      if (opt->verbose) std::printf("[oracle] it %d invalid step, radius %.3e\n", iter, radius);
      continue;
    }
    invalid = 0;
    auto candidate = [&](double step) {
      for (size_t i = 0; i < nc6; ++i) cposes[i] = poses[i] + step * dc[i];
      for (size_t i = 0; i < np3; ++i) cpoints[i] = clampd(points[i] + step * dp[i], lo, hi);
      return evaluate_cost(P, cposes.data(), cpoints.data());
    };
    double step_size = 1.0;
    double ccost = candidate(1.0);
    if (opt->line_search) {
      // Projected Armijo line search along delta — Ceres runs it after every trust-region step when any
      // parameter is bounded (TrustRegionMinimizer::DoLineSearch; the reference bounds every point,
      // src/bundle_adjuster.cpp:104-108).  ArmijoLineSearch, CUBIC interpolation => value AND directional
      // derivative at every trial point; sufficient_decrease 1e-4, contraction in [1e-3, 0.6],
      // <= 20 iterations, min step 1e-9 (against the direction's max-norm).  On failure delta is kept whole.
      double gd = 0.0, dmax = 0.0;
      for (int c = 0; c < Nc; ++c) if (fc[c] >= 0) for (int i = 0; i < 6; ++i) { gd += B.gc[6 * c + i] * dc[6 * c + i]; dmax = std::max(dmax, std::fabs(dc[6 * c + i])); }
      for (size_t i = 0; i < np3; ++i) { gd += B.gp[i] * dp[i]; dmax = std::max(dmax, std::fabs(dp[i])); }
      // cost and directional derivative  d/ds cost(x + s delta)  at the (projected) trial point
      auto candidate_grad = [&](double step, double* g_out) {
        for (size_t i = 0; i < nc6; ++i) cposes[i] = poses[i] + step * dc[i];
        for (size_t i = 0; i < np3; ++i) cpoints[i] = clampd(points[i] + step * dp[i], lo, hi);
        double f = 0.0, g = 0.0;
#pragma omp parallel for reduction(+ : f, g) schedule(static)
        for (int k = 0; k < Nobs; ++k) {
          double r[4], Jc[24], Jp[12];
          f += eval_obs(P, cposes.data(), cpoints.data(), k, r, Jc, Jp);
          const double* a = &dc[6 * (size_t)obs_cam[k]]; const double* b = &dp[3 * (size_t)obs_pt[k]];
          for (int m = 0; m < 4; ++m) {
            double jd = 0; for (int i = 0; i < 6; ++i) jd += Jc[m * 6 + i] * a[i]; for (int i = 0; i < 3; ++i) jd += Jp[m * 3 + i] * b[i];
            g += r[m] * jd;
          }
        }
        *g_out = g;
        return f;
      };
      const Sample s0{0.0, cost, gd, true};
      Sample prev{0.0, 0.0, 0.0, false}; bool have_prev = false;
      Sample cur{1.0, ccost, 0.0, false};
      bool cur_valid = std::isfinite(ccost);
      bool success = true;
      int it_ls = 0;
      bool have_grad = false;
      while (!cur_valid || cur.f > cost + 1e-4 * gd * cur.x) {
        if (++it_ls >= 20) { success = false; break; }
        if (cur_valid && !have_grad) { double g; candidate_grad(cur.x, &g); cur.g = g; cur.has_g = std::isfinite(g); have_grad = true; }
        double nx;
        const double lo_x = cur.x * 1e-3, hi_x = cur.x * 0.6;
        if (!cur_valid) nx = std::min(std::max(cur.x * 0.5, lo_x), hi_x);
        else {
          std::vector<Sample> sm; sm.push_back(s0); sm.push_back(cur);
          if (have_prev) sm.push_back(prev);
          nx = minimize_interpolating(sm, lo_x, hi_x);
        }
        if (nx * dmax < 1e-9) { success = false; break; }
        prev = cur; have_prev = cur_valid;
        double g;
        cur.x = nx; cur.f = candidate_grad(nx, &g); cur.g = g;
        cur_valid = std::isfinite(cur.f) && std::isfinite(g); cur.has_g = cur_valid; have_grad = true;
      }
      if (success) { step_size = cur.x; ccost = cur.f; }
      else { step_size = 1.0; }
      ccost = candidate(step_size);     // leaves cposes / cpoints at the chosen step
    }
    double step_norm2 = 0.0;
    for (int c = 0; c < Nc; ++c) if (fc[c] >= 0) for (int i = 0; i < 6; ++i) { const double d = cposes[6 * c + i] - poses[6 * c + i]; step_norm2 += d * d; }
    for (size_t i = 0; i < np3; ++i) { const double d = cpoints[i] - points[i]; step_norm2 += d * d; }
    const double step_norm = std::sqrt(step_norm2);
    const double xn = x_norm();
    if (step_norm <= opt->parameter_tolerance * (xn + opt->parameter_tolerance)) { term = 0; reason = 1; break; }
    const double cost_change = cost - ccost;
    if (std::fabs(cost_change) <= opt->function_tolerance * cost) { term = 0; reason = 2; break; }
    const double rho = cost_change / model_change;
    if (opt->verbose) std::printf("[oracle] it %d cost %.12e -> %.12e rho %.4f radius %.3e step %.3e ls %.3g\n", iter, cost, ccost, rho, radius, step_norm, step_size);
    if (std::isfinite(ccost) && rho > 1e-3) {
      std::memcpy(poses, cposes.data(), nc6 * 8); std::memcpy(points, cpoints.data(), np3 * 8);
      auto tl2 = clk::now();
      linearize(P, poses, points, &L); build_blocks(P, A, L, &B);
      sum->time_linearize_s += secs(tl2, clk::now()); sum->linearizations++;
      cost = L.cost; sum->successful_steps++;
      gmax = gradient_max_norm();
      if (gmax <= opt->gradient_tolerance) { term = 0; reason = 3; break; }
      const double t = 2.0 * rho - 1.0;
      radius = radius / std::max(1.0 / 3.0, 1.0 - t * t * t);
      radius = std::min(1e16, radius);
      decrease_factor = 2.0;
    } else {
      radius = radius / decrease_factor;
      decrease_factor *= 2.0;
    }
  }
  sum->final_cost = cost; sum->iterations = iter; sum->termination = term; sum->reason = reason;
  sum->final_radius = radius; sum->final_gradient_max_norm = gmax;
  sum->time_total_s = secs(t0, clk::now());
#ifdef _OPENMP
  omp_set_num_threads(saved_threads);
#endif
  return term == 2 ? 2 : 0;
}

// Timed K1-K3 (linearise + blocks + Schur) for the CPU baseline: returns seconds per pass.
double oracle_time_linearize_schur(int Nc, int Np, int Nobs, const double* poses, const double* points,
                                   const int32_t* obs_cam, const int32_t* obs_pt, const float* obs_uv,
                                   const double* p_l, const double* p_r, const uint8_t* cam_fixed,
                                   double huber_delta, int rdim, int num_threads, int repeats) {
#ifdef _OPENMP
  const int saved_threads = omp_get_max_threads();
  omp_set_num_threads(num_threads > 0 ? num_threads : omp_get_num_procs());
#endif
  Problem P{Nc, Np, Nobs, obs_cam, obs_pt, obs_uv, p_l, p_r, cam_fixed, huber_delta, rdim, false};
  Adjacency A; A.build(P);
  std::vector<int> fc(Nc); int Nf = 0;
  for (int c = 0; c < Nc; ++c) fc[c] = cam_fixed[c] ? -1 : Nf++;
  auto t0 = std::chrono::steady_clock::now();
  for (int rep = 0; rep < repeats; ++rep) {
    Linearization L; linearize(P, poses, points, &L);
    Blocks B; build_blocks(P, A, L, &B);
    std::vector<double> Dc((size_t)Nc * 6), Dp((size_t)Np * 3), S, rhs, Vinv;
    for (int c = 0; c < Nc; ++c) for (int i = 0; i < 6; ++i) Dc[6 * c + i] = clampd(B.U[(size_t)c * 36 + i * 7], 1e-6, 1e32) / 1e4;
    for (int p = 0; p < Np; ++p) for (int i = 0; i < 3; ++i) Dp[3 * p + i] = clampd(B.V[(size_t)p * 9 + i * 4], 1e-6, 1e32) / 1e4;
    schur(P, A, B, fc, Nf, Dc.data(), Dp.data(), &S, &rhs, &Vinv);
  }
  const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
#ifdef _OPENMP
  omp_set_num_threads(saved_threads);
#endif
  return dt / repeats;
}

}  // extern "C"
