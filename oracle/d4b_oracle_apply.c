/* d4b_oracle_apply.c -- stiffness + SIPG flux terms of the CPU restatement.
 * TEST INFRASTRUCTURE (see d4b_oracle.c).  Unity build: this file includes
 * d4b_oracle.c (tables, tensor products, context).
 */
#include "d4b_oracle.c"

/* element field pointer: local vector or ghost buffer
 * (d4est_mesh_get_field_on_element, src/Mesh/d4est_mesh.c) */
static const double *
elem_field (const d4bo_ctx_t * c, int32_t ref, const double *u, const double *ughost)
{
  if (ref >= 0) return u + c->d.nodal_stride[ref];
  return ughost + c->d.ghost_nodal_stride[-ref - 1];
}

static int
elem_deg (const d4bo_ctx_t * c, int32_t ref)
{
  return (ref >= 0) ? c->d.deg[ref] : c->d.ghost_deg[-ref - 1];
}

static int
elem_deg_quad (const d4bo_ctx_t * c, int32_t ref)
{
  return (ref >= 0) ? c->d.deg_quad[ref] : c->d.ghost_deg_quad[-ref - 1];
}

/* ---------------------------------------------------------------- stiffness */

/* K_e u_e = sum_{k,l',l} D_l'^T V^T diag(w J rst[l][k] rst[l'][k]) V D_l u
 * (d4est_quadrature_apply_stiffness_matrix, src/Quadrature/d4est_quadrature.c:263-382;
 *  loop over elements: d4est_laplacian_apply_stiffness_matrix,
 *  src/dGMath/d4est_laplacian.c:197-234).  Overwrites Au. */
void
d4bo_apply_stiffness (d4bo_ctx_t * c, const double *u, double *Au)
{
  const d4b_mesh_desc_t *d = &c->d;
  int64_t LQ = d->local_nodes_quad;
  for (int e = 0; e < d->n_local; e++) {
    int p = d->deg[e], q = d->deg_quad[e];
    int N = p + 1, Nq = q + 1, n3 = N * N * N, q3 = Nq * Nq * Nq;
    deg_tab_t *dt = get_deg (c, p);
    pair_tab_t *pt = get_pair (c, p, q);
    const double *w = get_gw (c, q);
    const double *ue = u + d->nodal_stride[e];
    double *oe = Au + d->nodal_stride[e];
    int64_t qs = d->quad_stride[e];
    double *du = (double *) malloc (sizeof (double) * 3 * n3);
    double *gq = (double *) malloc (sizeof (double) * 3 * q3);
    double *fq = (double *) malloc (sizeof (double) * 3 * q3);
    double *tmp = (double *) malloc (sizeof (double) * n3);
    double *tmp2 = (double *) malloc (sizeof (double) * n3);
    /* D_l u (apply_dij, d4est_operators.c:1385-1410) then V (LGL->Gauss) */
    kron3 (du, NULL, NULL, dt->D, ue, N, N, N, N, N, N);
    kron3 (du + n3, NULL, dt->D, NULL, ue, N, N, N, N, N, N);
    kron3 (du + 2 * n3, dt->D, NULL, NULL, ue, N, N, N, N, N, N);
    for (int l = 0; l < 3; l++)
      kron3 (gq + l * q3, pt->E[0], pt->E[0], pt->E[0], du + l * n3, Nq, N, Nq, N, Nq, N);
    for (int kk = 0; kk < q3; kk++) {
      int a = kk % Nq, b = (kk / Nq) % Nq, cc = kk / (Nq * Nq);
      double wj = w[a] * w[b] * w[cc] * d->J_quad[qs + kk];
      for (int lp = 0; lp < 3; lp++) {
        double s = 0.;
        for (int l = 0; l < 3; l++) {
          double g = 0.;
          for (int k = 0; k < 3; k++)
            g += d->rst_xyz_quad[(l * 3 + k) * LQ + qs + kk] * d->rst_xyz_quad[(lp * 3 + k) * LQ + qs + kk];
          s += g * gq[l * q3 + kk];
        }
        fq[lp * q3 + kk] = wj * s;
      }
    }
    for (int i = 0; i < n3; i++) oe[i] = 0.;
    for (int lp = 0; lp < 3; lp++) {
      kron3 (tmp, pt->ET[0], pt->ET[0], pt->ET[0], fq + lp * q3, N, Nq, N, Nq, N, Nq);
      /* apply_dij_transpose, d4est_operators.c:2259-2284 */
      if (lp == 0) kron3 (tmp2, NULL, NULL, dt->DT, tmp, N, N, N, N, N, N);
      if (lp == 1) kron3 (tmp2, NULL, dt->DT, NULL, tmp, N, N, N, N, N, N);
      if (lp == 2) kron3 (tmp2, dt->DT, NULL, NULL, tmp, N, N, N, N, N, N);
      for (int i = 0; i < n3; i++) oe[i] += tmp2[i];
    }
    free (du); free (gq); free (fq); free (tmp); free (tmp2);
  }
}

/* ------------------------------------------------------ volume quadrature ops */

/* u_quad = (V x V x V) u on every element, LGL(deg) -> Gauss(deg_quad)
 * (d4est_quadrature_interpolate, src/Quadrature/d4est_quadrature.c:966-1016) */
void
d4bo_interp_to_quad (d4bo_ctx_t * c, const double *u, double *u_quad)
{
  const d4b_mesh_desc_t *d = &c->d;
  for (int e = 0; e < d->n_local; e++) {
    int N = d->deg[e] + 1, Nq = d->deg_quad[e] + 1;
    pair_tab_t *pt = get_pair (c, d->deg[e], d->deg_quad[e]);
    kron3 (u_quad + d->quad_stride[e], pt->E[0], pt->E[0], pt->E[0], u + d->nodal_stride[e], Nq, N, Nq, N, Nq, N);
  }
}

/* out = V^T (w . jac . in_quad)
 * (d4est_quadrature_apply_galerkin_integral, d4est_quadrature.c:142-213) */
void
d4bo_apply_galerkin (d4bo_ctx_t * c, const double *in_quad, const double *jac_quad, double *out)
{
  const d4b_mesh_desc_t *d = &c->d;
  for (int e = 0; e < d->n_local; e++) {
    int N = d->deg[e] + 1, Nq = d->deg_quad[e] + 1, q3 = Nq * Nq * Nq;
    pair_tab_t *pt = get_pair (c, d->deg[e], d->deg_quad[e]);
    const double *w = get_gw (c, d->deg_quad[e]);
    int64_t qs = d->quad_stride[e];
    double *wji = (double *) malloc (sizeof (double) * q3);
    for (int kk = 0; kk < q3; kk++) {
      int a = kk % Nq, b = (kk / Nq) % Nq, cc = kk / (Nq * Nq);
      wji[kk] = w[cc] * w[b] * w[a] * (jac_quad ? jac_quad[qs + kk] : 1.) * in_quad[qs + kk];
    }
    kron3 (out + d->nodal_stride[e], pt->ET[0], pt->ET[0], pt->ET[0], wji, N, Nq, N, Nq, N, Nq);
    free (wji);
  }
}

/* out = V^T diag(w jac) V in: the mass matrix for jac = J
 * (d4est_quadrature_apply_mass_matrix, d4est_quadrature.c:385-477) and, for
 * jac = J f(x, V u0), the Newton-linearised mass-type term of
 * d4est_quadrature_apply_fofufofvlilj (QUAD_APPLY_MATRIX, :593-774) */
void
d4bo_apply_mass (d4bo_ctx_t * c, const double *in, const double *jac_quad, double *out)
{
  const d4b_mesh_desc_t *d = &c->d;
  double *in_quad = (double *) malloc (sizeof (double) * (d->local_nodes_quad + 1));
  d4bo_interp_to_quad (c, in, in_quad);
  d4bo_apply_galerkin (c, in_quad, jac_quad, out);
  free (in_quad);
}

/* --------------------------------------------------------------- face tools */

/* face f -> normal direction f/2, side f%2; face nodes ordered by the two
 * remaining directions, lower direction fastest
 * (d4est_operators_apply_slicer, d4est_operators.c:1521-1582) */
static inline int
vol_index (int N, int f, int a, int b, int n)
{
  int dir = f / 2;
  if (dir == 0) return n + N * (a + N * b);
  if (dir == 1) return a + N * (n + N * b);
  return a + N * (b + N * n);
}

static void
slice_face (const double *v, int N, int f, double *out)
{
  int n = (f % 2) ? N - 1 : 0;
  for (int b = 0; b < N; b++)
    for (int a = 0; a < N; a++) out[b * N + a] = v[vol_index (N, f, a, b, n)];
}

/* trace of dv/d(xi_dir) on face f, dir = 0,1,2 (compute_dudr + slicer,
 * d4est_laplacian.c:236-281, d4est_laplacian_flux.c:700-745) */
static void
slice_face_deriv (const double *v, int N, int f, int dir, const double *D, double *out)
{
  int n = (f % 2) ? N - 1 : 0;
  int stride = (dir == 0) ? 1 : (dir == 1) ? N : N * N;
  for (int b = 0; b < N; b++)
    for (int a = 0; a < N; a++) {
      int idx = vol_index (N, f, a, b, n);
      int i_dir = (dir == 0) ? idx % N : (dir == 1) ? (idx / N) % N : idx / (N * N);
      int base = idx - i_dir * stride;
      double s = 0.;
      for (int l = 0; l < N; l++) s += D[i_dir * N + l] * v[base + l * stride];
      out[b * N + a] = s;
    }
}

/* (+)-ordered N x N face array -> (-) ordering: flip dir0, flip dir1,
 * transpose (d4est_operators_reorient_face_data, d4est_operators.c:1993-2087) */
static void
reorient (const double *in, int N, int perm, double *out)
{
  double t0[MAXN * MAXN], t1[MAXN * MAXN];
  for (int b = 0; b < N; b++)
    for (int a = 0; a < N; a++)
      t0[b * N + a] = (perm & 1) ? in[b * N + (N - 1 - a)] : in[b * N + a];
  for (int b = 0; b < N; b++)
    for (int a = 0; a < N; a++)
      t1[b * N + a] = (perm & 2) ? t0[(N - 1 - b) * N + a] : t0[b * N + a];
  for (int b = 0; b < N; b++)
    for (int a = 0; a < N; a++)
      out[b * N + a] = (perm & 4) ? t1[a * N + b] : t1[b * N + a];
}

/* penalty functions, d4est_laplacian_flux_sipg.c:945-1005 */
static double
penalty (int fcn, int deg_m, double h_m, int deg_p, double h_p, double pre)
{
  switch (fcn) {
  case D4B_PENALTY_MAXP_SQR_OVER_MINH: {
    double md = (deg_m > deg_p) ? deg_m : deg_p, mh = (h_m < h_p) ? h_m : h_p;
    return (pre * md * md) / mh; }
  case D4B_PENALTY_MAXPP1_SQR_OVER_MINH: {
    double md = (deg_m > deg_p) ? deg_m : deg_p, mh = (h_m < h_p) ? h_m : h_p;
    return (pre * (md + 1) * (md + 1)) / mh; }
  case D4B_PENALTY_MEANP_SQR_OVER_MEANH: {
    double mp = .5 * (deg_m + deg_p), mh = .5 * (h_m + h_p);
    return (pre * mp * mp) / mh; }
  default: {
    double mpen = .5 * (deg_m * deg_m / h_m + deg_p * deg_p / h_p);
    return pre * mpen; }
  }
}
