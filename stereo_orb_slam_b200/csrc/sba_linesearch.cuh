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

// Real parts of ALL roots (complex ones included — Ceres's MinimizePolynomial looks at the real part of
// every eigenvalue of the companion matrix) of p(x) = c[0] x^deg + ... + c[deg].  deg <= 4.
// Returns the number of values written.  Degrees 1 and 2 are closed form (FindLinear/QuadraticPolynomialRoots);
// 3 and 4 by simultaneous (Durand-Kerner) iteration in complex arithmetic.
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
  // monic coefficients m[0..deg-1]: x^deg + m[0] x^(deg-1) + ... + m[deg-1]
  double m[4];
  double bound = 0.0;
  for (int i = 0; i < deg; ++i) { m[i] = c[i + 1] / c[0]; bound = fmax(bound, fabs(m[i])); }
  bound = 1.0 + bound;                                // Cauchy bound on |root|
  double zr[4], zi[4];
  for (int k = 0; k < deg; ++k) {                     // spread starting points on a circle, off the real axis
    const double ang = 0.4 + 6.283185307179586 * k / deg;
    zr[k] = 0.5 * bound * cos(ang); zi[k] = 0.5 * bound * sin(ang);
  }
  for (int it = 0; it < 500; ++it) {
    double change = 0.0;
    for (int k = 0; k < deg; ++k) {
      // p(z_k), Horner in complex arithmetic
      double pr = 1.0, pi = 0.0;
      for (int i = 0; i < deg; ++i) {
        const double tr = pr * zr[k] - pi * zi[k] + m[i];
        const double ti = pr * zi[k] + pi * zr[k];
        pr = tr; pi = ti;
      }
      double qr = 1.0, qi = 0.0;                      // prod_{j != k} (z_k - z_j)
      for (int j = 0; j < deg; ++j) {
        if (j == k) continue;
        const double dr = zr[k] - zr[j], di = zi[k] - zi[j];
        const double tr = qr * dr - qi * di, ti = qr * di + qi * dr;
        qr = tr; qi = ti;
      }
      const double den = qr * qr + qi * qi;
      if (!(den > 0.0)) continue;
      const double wr = (pr * qr + pi * qi) / den, wi = (pi * qr - pr * qi) / den;
      zr[k] -= wr; zi[k] -= wi;
      change = fmax(change, fabs(wr) + fabs(wi));
    }
    if (change <= 1e-9 * bound) break;
  }
  // polish: Newton on p(z) in complex arithmetic from the separated Durand-Kerner estimates (quadratic
  // convergence: three steps take 1e-9 to rounding level; the slow linear tail of the simultaneous iteration
  // is what made this routine cost > 100 us in its single thread)
  for (int k = 0; k < deg; ++k) {
    for (int it = 0; it < 4; ++it) {
      double pr = 1.0, pi = 0.0, dr = 0.0, di = 0.0;   // p and p' by Horner
      for (int i = 0; i < deg; ++i) {
        const double ndr = dr * zr[k] - di * zi[k] + pr, ndi = dr * zi[k] + di * zr[k] + pi;
        dr = ndr; di = ndi;
        const double tr = pr * zr[k] - pi * zi[k] + m[i], ti = pr * zi[k] + pi * zr[k];
        pr = tr; pi = ti;
      }
      const double den = dr * dr + di * di;
      if (!(den > 0.0) || !isfinite(den)) break;
      const double wr = (pr * dr + pi * di) / den, wi = (pi * dr - pr * di) / den;
      if (!isfinite(wr) || !isfinite(wi)) break;
      zr[k] -= wr; zi[k] -= wi;
      if (fabs(wr) + fabs(wi) <= 1e-16 * bound) break;
    }
  }
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
  double dcoef[5];
  for (int j = 0; j < deg; ++j) dcoef[j] = (double)(deg - j) * c[j];
  double roots[4];
  const int nr = ls_roots_real_parts(dcoef, deg - 1, roots);
  for (int i = 0; i < nr; ++i) {
    const double r = roots[i];
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
