// d4b_tables.cpp -- 1-D reference-element tables for the device kernels.
//
// Replaces the per-degree tables the reference builds lazily in
// d4est_operators_t (src/dGMath/d4est_operators.c): LGL / Gauss nodes and
// weights (table file GL_and_GLL_nodes_and_weights.h:6,4082), the nodal
// derivative matrix D = V_r V^-1 (:855-872), the LGL->Gauss interpolation
// (:411-438) and the p / hp prolongations (:944-1012).  Here every matrix is a
// Lagrange-basis evaluation computed with barycentric weights in long double;
// nodes come from Newton iteration on the Legendre recurrence.
#include <cmath>
#include <vector>
#include "d4b_internal.h"

namespace {

typedef long double ld;
const ld kPi = 3.14159265358979323846264338327950288L;

// P_n, P_n' at x via Bonnet's recurrence
void leg(int n, ld x, ld &P, ld &dP) {
  ld pm = 1.0L, p = x;
  if (n == 0) { P = 1.0L; dP = 0.0L; return; }
  for (int k = 1; k < n; ++k) {
    ld pn = ((2 * k + 1) * x * p - k * pm) / (k + 1);
    pm = p; p = pn;
  }
  P = p;
  // (1-x^2) P_n' = n (P_{n-1} - x P_n)
  ld one_m = 1.0L - x * x;
  if (fabsl(one_m) > 1e-30L) dP = n * (pm - x * p) / one_m;
  else dP = (x > 0 ? 1.0L : ((n % 2) ? 1.0L : -1.0L)) * 0.5L * n * (n + 1);
}

void bary_weights(int n, const double *x, std::vector<ld> &bw) {
  bw.assign(n, 1.0L);
  for (int j = 0; j < n; ++j)
    for (int m = 0; m < n; ++m)
      if (m != j) bw[j] /= ((ld)x[j] - (ld)x[m]);
}

}  // namespace

void d4b_tab_gauss(int n, double *x, double *w) {
  for (int i = 0; i < n; ++i) {
    ld z = -cosl(kPi * (4 * i + 3) / (4 * n + 2));
    ld P, dP;
    for (int it = 0; it < 60; ++it) {
      leg(n, z, P, dP);
      ld dz = P / dP;
      z -= dz;
      if (fabsl(dz) < 1e-20L) break;
    }
    leg(n, z, P, dP);
    x[i] = (double)z;
    w[i] = (double)(2.0L / ((1.0L - z * z) * dP * dP));
  }
}

void d4b_tab_lobatto(int n, double *x, double *w) {
  int p = n - 1;
  for (int i = 0; i < n; ++i) {
    ld z;
    if (i == 0) z = -1.0L;
    else if (i == p) z = 1.0L;
    else {
      // interior LGL nodes are the roots of P_p'
      z = -cosl(kPi * i / p);
      for (int it = 0; it < 60; ++it) {
        ld P, dP;
        leg(p, z, P, dP);
        ld d2P = (2.0L * z * dP - (ld)p * (p + 1) * P) / (1.0L - z * z);
        ld dz = dP / d2P;
        z -= dz;
        if (fabsl(dz) < 1e-20L) break;
      }
    }
    ld P, dP;
    leg(p, z, P, dP);
    x[i] = (double)z;
    w[i] = (double)(2.0L / ((ld)p * (p + 1) * P * P));
  }
}

void d4b_tab_deriv(int n, const double *x, double *D) {
  std::vector<ld> bw;
  bary_weights(n, x, bw);
  for (int i = 0; i < n; ++i) {
    ld diag = 0.0L;
    for (int j = 0; j < n; ++j) {
      if (j == i) continue;
      ld v = (bw[j] / bw[i]) / ((ld)x[i] - (ld)x[j]);
      D[i * n + j] = (double)v;
      diag -= v;
    }
    D[i * n + i] = (double)diag;
  }
}

void d4b_tab_eval(int n, const double *x, int nq, const double *y, double *E) {
  std::vector<ld> bw;
  bary_weights(n, x, bw);
  for (int k = 0; k < nq; ++k) {
    int hit = -1;
    for (int j = 0; j < n; ++j)
      if ((ld)y[k] == (ld)x[j]) hit = j;
    if (hit >= 0) {
      for (int j = 0; j < n; ++j) E[k * n + j] = (j == hit) ? 1.0 : 0.0;
      continue;
    }
    ld den = 0.0L;
    std::vector<ld> t(n);
    for (int j = 0; j < n; ++j) { t[j] = bw[j] / ((ld)y[k] - (ld)x[j]); den += t[j]; }
    for (int j = 0; j < n; ++j) E[k * n + j] = (double)(t[j] / den);
  }
}
