// Test-only: evaluates the per-observation algebra of csrc/sba_math.cuh on the CPU so the formulas
// the CUDA kernels are built from can be checked against the oracle without a GPU.  NOT a product
// path (libsba.so has no CPU fallback).
#include "../../stereo_orb_slam_b200/csrc/sba_math.cuh"
#include "../../stereo_orb_slam_b200/csrc/sba_linesearch.cuh"
extern "C" {
void emu_observation(const double* pose, const double* X, const float* uv, const double* pl, const double* pr,
                     int rdim, double delta, double* r_corrected, double* U21, double* gc, double* V6, double* gp,
                     double* W18, double* cost) {
  sba::CamPre c; sba::cam_precompute(pose, c);
  double q[3], r[4], D[12];
  const double s = sba::project(pl, pr, c, X, uv, rdim, q, r, D);
  double rho, w; sba::huber(delta, s, rho, w);
  *cost = 0.5 * rho;
  const double sq = sqrt(w);
  for (int i = 0; i < 4; ++i) r_corrected[i] = sq * r[i];
  double A[6], h[3], G[9];
  sba::form_A_h(D, r, w, A, h); sba::form_G(c, q, G);
  for (int i = 0; i < 21; ++i) U21[i] = 0; for (int i = 0; i < 6; ++i) { gc[i] = 0; V6[i] = 0; }
  for (int i = 0; i < 3; ++i) gp[i] = 0;
  sba::accum_camera(A, h, G, U21, gc); sba::accum_point(c, A, h, G, V6, gp, W18);
}
double emu_obs_cost(const double* pose, const double* X, const float* uv, const double* pl, const double* pr, int rdim, double delta) {
  sba::CamPre c; sba::cam_precompute(pose, c); return sba::obs_cost(pl, pr, c, X, uv, rdim, delta);
}
int emu_inv_spd6(double* M) { return sba::inv_spd6(M) ? 1 : 0; }
int emu_inv_sym3(const double* M, double* Mi) { return sba::inv_sym3(M, Mi) ? 1 : 0; }
double emu_ls_minimize(int n, const double* x, const double* f, const double* g, const int* has_g, double xmin, double xmax) {
  sba::LsSample s[3];
  for (int i = 0; i < n && i < 3; ++i) { s[i].x = x[i]; s[i].f = f[i]; s[i].g = g[i]; s[i].has_g = has_g[i]; }
  return sba::ls_minimize_interpolating(s, n, xmin, xmax);
}
}
extern "C" int emu_ls_roots(const double* c, int deg, double* out) { return sba::ls_roots_real_parts(c, deg, out); }
