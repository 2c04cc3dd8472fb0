/* d4b_oracle.c -- CPU restatement ("port") of the reference's SIPG apply.
 *
 * TEST INFRASTRUCTURE - not part of the product.  Only tests/, smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library.
 * Parity status: PINNED - checked in tests/test_oracle.py against outputs of
 * the unmodified reference run here (oracle/_ref/ref_driver, fixtures under
 * tests/golden/) and against the reference's own known-answer tests
 * (plain == with_opt, A(x^2+y^2+z^2) = M(-6), symmetry).
 *
 * Plain C, single thread, side-centric like the reference: one pass over the
 * elements for the stiffness term, one pass over the flux "side" records for
 * the SIPG terms.  Each function cites the reference file:line it follows.
 * The one deliberate difference: side->mortar projection followed by
 * LGL->Gauss interpolation (d4est_laplacian_flux.c:640-700) is collapsed into
 * one evaluation matrix (side Lagrange basis evaluated at the mortar's Gauss
 * points); both are exact for polynomials of the side's degree, so results
 * agree to round-off.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include "../include/d4est_b200.h"

#define MAXN 24                 /* max nodes per direction the oracle handles */
#define D4BO_PI 3.14159265358979323846264338327950288L

/* ------------------------------------------------------------------ tables */

/* Legendre P_n(x) and derivative by the three-term recurrence
 * (what d4est_lgl_jacobi(x,0,0,n) evaluates up to normalisation,
 *  src/dGMath/d4est_lgl.c:14-48) */
static void
legendre (int n, long double x, long double *P, long double *dP)
{
  long double p0 = 1.0L, p1 = x, d0 = 0.0L, d1 = 1.0L;
  if (n == 0) { *P = p0; *dP = d0; return; }
  for (int k = 2; k <= n; k++) {
    long double p2 = ((2 * k - 1) * x * p1 - (k - 1) * p0) / k;
    long double d2 = d0 + (2 * k - 1) * p1;
    p0 = p1; p1 = p2; d0 = d1; d1 = d2;
  }
  *P = p1; *dP = d1;
}

/* Gauss-Legendre nodes/weights, n points (reference: table
 * GL_and_GLL_nodes_and_weights.h:4082; here by Newton iteration) */
void
d4bo_gauss (int n, double *x, double *w)
{
  for (int i = 0; i < n; i++) {
    long double z = -cosl (D4BO_PI * (i + 0.75L) / (n + 0.5L));
    long double P, dP;
    for (int it = 0; it < 100; it++) {
      legendre (n, z, &P, &dP);
      long double dz = P / dP;
      z -= dz;
      if (fabsl (dz) < 1e-19L) break;
    }
    legendre (n, z, &P, &dP);
    x[i] = (double) z;
    w[i] = (double) (2.0L / ((1.0L - z * z) * dP * dP));
  }
}

/* Gauss-Lobatto-Legendre nodes/weights, n = deg+1 points (table :6) */
void
d4bo_lobatto (int n, double *x, double *w)
{
  int p = n - 1;
  for (int i = 0; i < n; i++) {
    long double z;
    if (i == 0) z = -1.0L;
    else if (i == n - 1) z = 1.0L;
    else {
      /* roots of P'_p: Newton on q = P'_p with q' from the Legendre ODE */
      z = -cosl (D4BO_PI * i / p);
      for (int it = 0; it < 100; it++) {
        long double P, dP;
        legendre (p, z, &P, &dP);
        long double ddP = (2 * z * dP - p * (p + 1) * P) / (1 - z * z);
        long double dz = dP / ddP;
        z -= dz;
        if (fabsl (dz) < 1e-19L) break;
      }
    }
    long double P, dP;
    legendre (p, z, &P, &dP);
    x[i] = (double) z;
    w[i] = (double) (2.0L / (p * (p + 1) * P * P));
  }
}

/* Lagrange basis on nodes xn[0..n) evaluated at y: L[j], and derivative dL[j] */
static void
lagrange_eval (int n, const double *xn, long double y, long double *L, long double *dL)
{
  for (int j = 0; j < n; j++) {
    long double num = 1.0L, den = 1.0L;
    for (int m = 0; m < n; m++)
      if (m != j) { num *= (y - xn[m]); den *= ((long double) xn[j] - xn[m]); }
    L[j] = num / den;
    long double s = 0.0L;
    for (int k = 0; k < n; k++) {
      if (k == j) continue;
      long double t = 1.0L;
      for (int m = 0; m < n; m++)
        if (m != j && m != k) t *= (y - xn[m]);
      s += t;
    }
    dL[j] = s / den;
  }
}

/* evaluation matrix E[nq x n]: LGL(deg) basis at Gauss(deg_q) points mapped
 * to the whole interval (half = 0) or to the lower / upper half
 * (half = 1, 2: r -> (r -+ 1)/2, d4est_reference_child_to_parent_coords).
 * half=0 equals build_lobatto_to_gauss_interp_1d (d4est_operators.c:411-438);
 * half=1,2 equals interp * hp_prolong (d4est_operators.c:944-993). */
void
d4bo_eval_matrix (int deg, int deg_q, int half, double *E)
{
  int n = deg + 1, nq = deg_q + 1;
  double xl[MAXN], wl[MAXN], xg[MAXN], wg[MAXN];
  long double L[MAXN], dL[MAXN];
  d4bo_lobatto (n, xl, wl);
  d4bo_gauss (nq, xg, wg);
  for (int k = 0; k < nq; k++) {
    long double y = xg[k];
    if (half == 1) y = 0.5L * (y - 1.0L);
    if (half == 2) y = 0.5L * (y + 1.0L);
    lagrange_eval (n, xl, y, L, dL);
    for (int j = 0; j < n; j++) E[k * n + j] = (double) L[j];
  }
}

/* derivative matrix D[n x n] on LGL(deg): D_ij = L_j'(x_i)
 * (build_dij_1d, d4est_operators.c:855-872: D = V_r V^-1) */
void
d4bo_dij (int deg, double *D)
{
  int n = deg + 1;
  double xl[MAXN], wl[MAXN];
  long double L[MAXN], dL[MAXN];
  d4bo_lobatto (n, xl, wl);
  for (int i = 0; i < n; i++) {
    lagrange_eval (n, xl, xl[i], L, dL);
    for (int j = 0; j < n; j++) D[i * n + j] = (double) dL[j];
  }
}

/* ---------------------------------------------------------- tensor products */

/* y = (A_t (x) A_s (x) A_r) x, each A row-major [nr x nc], r fastest in x/y
 * (d4est_kron_A1A2A3x_nonsqr, src/Kron/d4est_kron.h:531-548, A3 <-> r).
 * A NULL matrix means identity in that direction (then nr == nc). */
static void
kron3 (double *y, const double *At, const double *As, const double *Ar, const double *x,
       int nr_t, int nc_t, int nr_s, int nc_s, int nr_r, int nc_r)
{
  double *t1 = (double *) malloc (sizeof (double) * (size_t) nc_t * nc_s * nr_r);
  double *t2 = (double *) malloc (sizeof (double) * (size_t) nc_t * nr_s * nr_r);
  for (int c = 0; c < nc_t; c++)
    for (int b = 0; b < nc_s; b++)
      for (int a = 0; a < nr_r; a++) {
        double s = 0.;
        if (Ar) { for (int l = 0; l < nc_r; l++) s += Ar[a * nc_r + l] * x[(c * nc_s + b) * nc_r + l]; }
        else s = x[(c * nc_s + b) * nc_r + a];
        t1[(c * nc_s + b) * nr_r + a] = s;
      }
  for (int c = 0; c < nc_t; c++)
    for (int b = 0; b < nr_s; b++)
      for (int a = 0; a < nr_r; a++) {
        double s = 0.;
        if (As) { for (int l = 0; l < nc_s; l++) s += As[b * nc_s + l] * t1[(c * nc_s + l) * nr_r + a]; }
        else s = t1[(c * nc_s + b) * nr_r + a];
        t2[(c * nr_s + b) * nr_r + a] = s;
      }
  for (int c = 0; c < nr_t; c++)
    for (int b = 0; b < nr_s; b++)
      for (int a = 0; a < nr_r; a++) {
        double s = 0.;
        if (At) { for (int l = 0; l < nc_t; l++) s += At[c * nc_t + l] * t2[(l * nr_s + b) * nr_r + a]; }
        else s = t2[(c * nr_s + b) * nr_r + a];
        y[(c * nr_s + b) * nr_r + a] = s;
      }
  free (t1);
  free (t2);
}

/* y = (A_y (x) A_x) x in 2-D (d4est_kron_A1A2x_nonsqr, d4est_kron.h:443-467) */
static void
kron2 (double *y, const double *Ay, const double *Ax, const double *x,
       int nr_y, int nc_y, int nr_x, int nc_x)
{
  double t[MAXN * MAXN];
  for (int b = 0; b < nc_y; b++)
    for (int a = 0; a < nr_x; a++) {
      double s = 0.;
      for (int l = 0; l < nc_x; l++) s += Ax[a * nc_x + l] * x[b * nc_x + l];
      t[b * nr_x + a] = s;
    }
  for (int b = 0; b < nr_y; b++)
    for (int a = 0; a < nr_x; a++) {
      double s = 0.;
      for (int l = 0; l < nc_y; l++) s += Ay[b * nc_y + l] * t[l * nr_x + a];
      y[b * nr_x + a] = s;
    }
}

static void
transpose (const double *A, int nr, int nc, double *AT)
{
  for (int i = 0; i < nr; i++)
    for (int j = 0; j < nc; j++) AT[j * nr + i] = A[i * nc + j];
}

/* ----------------------------------------------------------------- context */

typedef struct
{
  double D[MAXN * MAXN], DT[MAXN * MAXN];
} deg_tab_t;

typedef struct
{
  double E[3][MAXN * MAXN];     /* [half][nq x n] */
  double ET[3][MAXN * MAXN];
  int built;
} pair_tab_t;

typedef struct d4bo_ctx
{
  d4b_mesh_desc_t d;            /* shallow copy; arrays owned by the caller */
  deg_tab_t deg[MAXN];
  int deg_built[MAXN];
  pair_tab_t pair[MAXN][MAXN];  /* [deg][deg_q] */
  double gw[MAXN][MAXN];        /* Gauss weights per deg_q */
  int gw_built[MAXN];
} d4bo_ctx_t;

static deg_tab_t *
get_deg (d4bo_ctx_t * c, int deg)
{
  if (!c->deg_built[deg]) {
    d4bo_dij (deg, c->deg[deg].D);
    transpose (c->deg[deg].D, deg + 1, deg + 1, c->deg[deg].DT);
    c->deg_built[deg] = 1;
  }
  return &c->deg[deg];
}

static pair_tab_t *
get_pair (d4bo_ctx_t * c, int deg, int deg_q)
{
  pair_tab_t *p = &c->pair[deg][deg_q];
  if (!p->built) {
    for (int h = 0; h < 3; h++) {
      d4bo_eval_matrix (deg, deg_q, h, p->E[h]);
      transpose (p->E[h], deg_q + 1, deg + 1, p->ET[h]);
    }
    p->built = 1;
  }
  return p;
}

static const double *
get_gw (d4bo_ctx_t * c, int deg_q)
{
  if (!c->gw_built[deg_q]) {
    double x[MAXN];
    d4bo_gauss (deg_q + 1, x, c->gw[deg_q]);
    c->gw_built[deg_q] = 1;
  }
  return c->gw[deg_q];
}

d4bo_ctx_t *
d4bo_create (const d4b_mesh_desc_t * desc)
{
  d4bo_ctx_t *c = (d4bo_ctx_t *) calloc (1, sizeof (*c));
  c->d = *desc;
  return c;
}

void
d4bo_destroy (d4bo_ctx_t * c)
{
  free (c);
}
