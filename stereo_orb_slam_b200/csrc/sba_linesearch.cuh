// sba_linesearch.cuh — step-size selection of the projected Armijo line search that Ceres runs
// inside its trust-region loop whenever a parameter has bounds.  The reference sets bounds on every
// point (reference: src/bundle_adjuster.cpp:104-108), so every LM step of the reference goes through
// it.  Restated from Ceres's published algorithm (line_search.cc: ArmijoLineSearch with CUBIC
// interpolation, i.e. value AND directional derivative at every trial point; polynomial.cc:
// FindInterpolatingPolynomial / MinimizePolynomial / FindPolynomialRoots):
//   samples   : (0, f0, g0), the current trial (x, f, g) and, once there is one, the previous trial
//   polynomial: the lowest-degree interpolant of all values and derivatives (cubic, then quintic)
//   next step : argmin of the polynomial over {x_min, x_max, their midpoint, and the real parts of the
//               roots of its derivative that fall inside [x_min, x_max]}
// Host + device; one thread per window runs it (rare path: only when the full step fails Armijo).
#pragma once
#include <math.h>

#include "sba_math.cuh"

namespace sba {

struct LsSample {
  double x, f, g;
  int has_g;   // value is always valid for a sample that is passed in
};

SBA_HD double ls_poly_eval(const double* c, int deg, double x) {
  double v = 0.0;
  for (int j = 0; j <= deg; ++j) v = v * x + c[j];
  return v;
}

// Interpolating polynomial through the samples (values and, where present, derivatives).
// c[0..deg], highest power first.  Returns deg, or -1 if the system is singular.
SBA_HD int ls_fit_polynomial(const LsSample* s, int n, double* c /* [6] */) {
  int neq = 0;
  for (int i = 0; i < n; ++i) neq += s[i].has_g ? 2 : 1;
  if (neq > 6 || neq < 1) return -1;
  const int deg = neq - 1;
  double M[36], b[6];
  for (int i = 0; i < 36; ++i) M[i] = 0.0;
  int row = 0;
  for (int i = 0; i < n; ++i) {
    double pw[6];                      // x^0 .. x^5 by repeated multiplication (one thread runs this: no pow())
    pw[0] = 1.0;
    for (int j = 1; j < 6; ++j) pw[j] = pw[j - 1] * s[i].x;
    for (int j = 0; j <= deg; ++j) M[row * 6 + j] = pw[deg - j];
    b[row++] = s[i].f;
    if (s[i].has_g) {
      for (int j = 0; j < deg; ++j) M[row * 6 + j] = (double)(deg - j) * pw[deg - j - 1];
      b[row++] = s[i].g;
    }
  }
  // Gaussian elimination, full pivoting (Ceres: fullPivLu)
  int perm[6];
  for (int i = 0; i < neq; ++i) perm[i] = i;
  for (int k = 0; k < neq; ++k) {
    int pr = k, pc = k;
    double best = -1.0;
    for (int i = k; i < neq; ++i)
      for (int j = k; j < neq; ++j)
        if (fabs(M[i * 6 + j]) > best) { best = fabs(M[i * 6 + j]); pr = i; pc = j; }
    if (!(best > 0.0)) return -1;
    if (pr != k) { for (int j = 0; j < neq; ++j) { const double t = M[k * 6 + j]; M[k * 6 + j] = M[pr * 6 + j]; M[pr * 6 + j] = t; } const double t = b[k]; b[k] = b[pr]; b[pr] = t; }
    if (pc != k) { for (int i = 0; i < neq; ++i) { const double t = M[i * 6 + k]; M[i * 6 + k] = M[i * 6 + pc]; M[i * 6 + pc] = t; } const int t = perm[k]; perm[k] = perm[pc]; perm[pc] = t; }
    const double d = M[k * 6 + k];
    for (int i = k + 1; i < neq; ++i) {
      const double f = M[i * 6 + k] / d;
      if (f == 0.0) continue;
      for (int j = k; j < neq; ++j) M[i * 6 + j] -= f * M[k * 6 + j];
      b[i] -= f * b[k];
    }
  }
  double y[6];
  for (int i = neq - 1; i >= 0; --i) {
    double acc = b[i];
    for (int j = i + 1; j < neq; ++j) acc -= M[i * 6 + j] * y[j];
    y[i] = acc / M[i * 6 + i];
  }
  for (int i = 0; i < neq; ++i) c[perm[i]] = y[i];
  return deg;
}

// One root of the complex polynomial sum_k (ar[k] + i ai[k]) x^k of degree m by Laguerre's iteration, started
// from (and returned in) xr + i xi.  Textbook algorithm (p, p', p'' by Horner; step m / (G +- sqrt((m-1)(m H - G^2)))
// with the larger denominator; an occasional fractional step breaks limit cycles).
SBA_HD void sba_ls_laguerre(const double* ar, const double* ai, int m, double& xr, double& xi) {
  const double frac[8] = {0.5, 0.25, 0.75, 0.13, 0.38, 0.62, 0.88, 1.0};
  for (int it = 1; it <= 80; ++it) {
    double br = ar[m], bi = ai[m], dr = 0.0, di = 0.0, fr = 0.0, fi = 0.0;
    double err = sqrt(br * br + bi * bi);
    const double abx = sqrt(xr * xr + xi * xi);
    for (int j = m - 1; j >= 0; --j) {
      double tr = xr * fr - xi * fi + dr, ti = xr * fi + xi * fr + di; fr = tr; fi = ti;
      tr = xr * dr - xi * di + br; ti = xr * di + xi * dr + bi; dr = tr; di = ti;
      tr = xr * br - xi * bi + ar[j]; ti = xr * bi + xi * br + ai[j]; br = tr; bi = ti;
      err = sqrt(br * br + bi * bi) + abx * err;
    }
    const double babs2 = br * br + bi * bi;
    if (!(babs2 > 0.0) || sqrt(babs2) <= 1e-15 * err) return;          // at a root (to rounding)
    // G = p'/p, H = G^2 - 2 (p''/2)/p
    const double gr = (dr * br + di * bi) / babs2, gi = (di * br - dr * bi) / babs2;
    const double g2r = gr * gr - gi * gi, g2i = 2.0 * gr * gi;
    const double qr = (fr * br + fi * bi) / babs2, qi = (fi * br - fr * bi) / babs2;
    const double hr = g2r - 2.0 * qr, hi = g2i - 2.0 * qi;
    // sq = sqrt((m-1) (m H - G^2))
    const double sr0 = (double)(m - 1) * ((double)m * hr - g2r), si0 = (double)(m - 1) * ((double)m * hi - g2i);
    const double sabs = sqrt(sr0 * sr0 + si0 * si0);
    double sqr = sqrt(0.5 * fmax(sabs + sr0, 0.0)), sqi = sqrt(0.5 * fmax(sabs - sr0, 0.0));
    if (si0 < 0.0) sqi = -sqi;
    double pr = gr + sqr, pi = gi + sqi;
    const double mr = gr - sqr, mi = gi - sqi;
    if (pr * pr + pi * pi < mr * mr + mi * mi) { pr = mr; pi = mi; }
    const double pabs2 = pr * pr + pi * pi;
    double dxr, dxi;
    if (pabs2 > 0.0) { dxr = (double)m * pr / pabs2; dxi = -(double)m * pi / pabs2; }
    else { dxr = (1.0 + abx) * cos((double)it); dxi = (1.0 + abx) * sin((double)it); }
    const double nr = xr - dxr, ni = xi - dxi;
    if (nr == xr && ni == xi) return;                                   // converged
    if (!isfinite(nr) || !isfinite(ni)) return;
    if (it % 10) { xr = nr; xi = ni; }
    else { const double f = frac[(it / 10) & 7]; xr -= f * dxr; xi -= f * dxi; }
  }
}

// Real parts of ALL roots (complex ones included — Ceres's MinimizePolynomial looks at the real part of
// every eigenvalue of the companion matrix) of p(x) = c[0] x^deg + ... + c[deg].  deg <= 4.
// Returns the number of values written.  Degrees 1 and 2 are closed form (FindLinear/QuadraticPolynomialRoots);
// 3 and 4 by Laguerre's iteration with deflation and polishing, in complex arithmetic.
SBA_HD int ls_roots_real_parts(const double* cin, int deg, double* out /* [4] */) {
  int lead = 0;
  while (lead < deg && cin[lead] == 0.0) ++lead;     // RemoveLeadingZeros
  const double* c = cin + lead;
  deg -= lead;
  if (deg <= 0) return 0;
  if (deg == 1) { out[0] = -c[1] / c[0]; return 1; }
  if (deg == 2) {
    const double a = c[0], b = c[1], cc = c[2];
    const double D = b * b - 4.0 * a * cc;
    const double sD = sqrt(fabs(D));
    if (D >= 0.0) {
      if (b >= 0.0) { out[0] = (-b - sD) / (2.0 * a); out[1] = (2.0 * cc) / (-b - sD); }
      else { out[0] = (2.0 * cc) / (-b + sD); out[1] = (-b + sD) / (2.0 * a); }
    } else {
      out[0] = -b / (2.0 * a); out[1] = out[0];
    }
    return 2;
  }
  // Degrees 3 and 4: Laguerre's iteration in complex arithmetic with deflation, every root polished on the
  // undeflated polynomial.  (A simultaneous Durand-Kerner iteration started on the Cauchy circle needs hundreds
  // of sweeps when the leading coefficient is tiny — the usual case here, a quintic fitted through data that is
  // almost cubic — and its single thread then costs > 100 us per line-search round.)
  // a[k] = coefficient of x^k, complex, of the current (deflated) polynomial
  double ar[5], ai[5];
  for (int k = 0; k <= deg; ++k) { ar[k] = c[deg - k]; ai[k] = 0.0; }
  double zr[4], zi[4];
  int m = deg;
  for (int found = 0; found < deg; ++found, --m) {
    double xr = 0.0, xi = 0.0;
    sba_ls_laguerre(ar, ai, m, xr, xi);
    zr[found] = xr; zi[found] = xi;
    // deflate: synthetic division by (x - root)
    double br = ar[m], bi = ai[m];
    for (int k = m - 1; k >= 0; --k) {
      const double tr = ar[k], ti = ai[k];
      ar[k] = br; ai[k] = bi;
      const double nr = xr * br - xi * bi + tr, ni = xr * bi + xi * br + ti;
      br = nr; bi = ni;
    }
  }
  // polish on the original polynomial
  for (int k = 0; k <= deg; ++k) { ar[k] = c[deg - k]; ai[k] = 0.0; }
  for (int k = 0; k < deg; ++k) sba_ls_laguerre(ar, ai, deg, zr[k], zi[k]);
  for (int k = 0; k < deg; ++k) out[k] = zr[k];
  return deg;
}

// MinimizeInterpolatingPolynomial: argmin over [xmin, xmax] as described in the file header.
SBA_HD double ls_minimize_interpolating(const LsSample* s, int n, double xmin, double xmax) {
  double c[6];
  const int deg = ls_fit_polynomial(s, n, c);
  double best_x = 0.5 * (xmin + xmax);
  if (deg < 0) return best_x;
  double best_v = ls_poly_eval(c, deg, best_x);
  const double vmin = ls_poly_eval(c, deg, xmin), vmax = ls_poly_eval(c, deg, xmax);
  if (vmin < best_v) { best_v = vmin; best_x = xmin; }
  if (vmax < best_v) { best_v = vmax; best_x = xmax; }
  if (deg < 2) return best_x;
  // derivative in the scaled variable y = x / xmax: the search interval becomes [xmin / xmax, 1], the roots that
  // matter are O(1) and the simultaneous iteration converges in a few tens of steps instead of running to its
  // cap when the step sizes are ~1e-11 (coefficients spanning 50 orders of magnitude)
  const double sc = (xmax > 0.0 && isfinite(xmax)) ? xmax : 1.0;
  double dcoef[5];
  {
    double pw = 1.0;                                  // sc^(deg-1-j), built from the constant term upwards
    for (int j = deg - 1; j >= 0; --j) { dcoef[j] = (double)(deg - j) * c[j] * pw; pw *= sc; }
  }
  double roots[4];
  const int nr = ls_roots_real_parts(dcoef, deg - 1, roots);
  for (int i = 0; i < nr; ++i) {
    const double r = roots[i] * sc;
    if (!(r >= xmin) || !(r <= xmax)) continue;       // also drops NaN
    const double v = ls_poly_eval(c, deg, r);
    if (v < best_v) { best_v = v; best_x = r; }
  }
  return best_x;
}

// One Armijo contraction (ArmijoLineSearch::DoSearch loop body): the next trial step from the initial
// point (0, f0, g0), the previous trial (if any) and the current one.  CUBIC interpolation; bounds
// [max_step_contraction, min_step_contraction] * cur.x = [1e-3, 0.6] * cur.x.  A non-finite current
// value bisects (InterpolatingPolynomialMinimizingStepSize, first branch).
SBA_HD double ls_next_step(double f0, double g0, const LsSample& prev, int prev_valid, const LsSample& cur, int cur_valid) {
  const double lo = 1e-3 * cur.x, hi = 0.6 * cur.x;
  if (!cur_valid) return fmin(fmax(cur.x * 0.5, lo), hi);
  LsSample s[3];
  s[0].x = 0.0; s[0].f = f0; s[0].g = g0; s[0].has_g = 1;
  s[1] = cur;
  int n = 2;
  if (prev_valid) s[n++] = prev;
  return ls_minimize_interpolating(s, n, lo, hi);
}

}  // namespace sba
