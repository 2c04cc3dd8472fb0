/* d4b_oracle_flux.c -- SIPG interface / Dirichlet boundary terms and the full
 * apply of the CPU restatement.  TEST INFRASTRUCTURE (see d4b_oracle.c).
 * Unity build root: compile THIS file into liboracle.
 */
#include "d4b_oracle_apply.c"

#define FQ (MAXN * MAXN)

/* traces of one side element at the Gauss points of one mortar:
 * uq[k], dq[3][k] (reference-space gradient), in the element's own face order */
static void
side_traces_on_mortar (d4bo_ctx_t * c, const double *v, int deg, int f, int dq_deg,
                       int hx, int hy, double *uq, double *dq)
{
  int N = deg + 1, Nq = dq_deg + 1;
  deg_tab_t *dt = get_deg (c, deg);
  pair_tab_t *pt = get_pair (c, deg, dq_deg);
  double tr[FQ];
  slice_face (v, N, f, tr);
  kron2 (uq, pt->E[hy], pt->E[hx], tr, Nq, N, Nq, N);
  for (int d = 0; d < 3; d++) {
    slice_face_deriv (v, N, f, d, dt->D, tr);
    kron2 (dq + d * FQ, pt->E[hy], pt->E[hx], tr, Nq, N, Nq, N);
  }
}

/* V^T w (galerkin integral, d4est_quadrature.c:142-213), mortar->side
 * projection (d4est_mortars.c:510-547), lift (d4est_operators.c:1454-1519) and
 * D_l^T for term2 (d4est_laplacian_flux_sipg.c:745-790); accumulates into the
 * element's Au.  t13 = term1 + term3, t2[l] = term2_l, all at mortar Gauss
 * points in the element's face order. */
static void
test_and_lift (d4bo_ctx_t * c, double *Au_e, int deg, int f, int dq_deg, int hx, int hy,
               const double *t13, const double *t2, double scale2)
{
  int N = deg + 1, Nq = dq_deg + 1;
  deg_tab_t *dt = get_deg (c, deg);
  pair_tab_t *pt = get_pair (c, deg, dq_deg);
  const double *w = get_gw (c, dq_deg);
  double wt[FQ], r[FQ];
  int n = (f % 2) ? N - 1 : 0;
  for (int b = 0; b < Nq; b++)
    for (int a = 0; a < Nq; a++) wt[b * Nq + a] = w[a] * w[b] * t13[b * Nq + a];
  kron2 (r, pt->ET[hy], pt->ET[hx], wt, N, Nq, N, Nq);
  for (int b = 0; b < N; b++)
    for (int a = 0; a < N; a++) Au_e[vol_index (N, f, a, b, n)] += r[b * N + a];
  for (int l = 0; l < 3; l++) {
    for (int b = 0; b < Nq; b++)
      for (int a = 0; a < Nq; a++) wt[b * Nq + a] = w[a] * w[b] * t2[l * FQ + b * Nq + a];
    kron2 (r, pt->ET[hy], pt->ET[hx], wt, N, Nq, N, Nq);
    /* D_l^T applied to the lifted face vector: out[i] += D[j][i] * lifted[j]
     * along direction l */
    int stride = (l == 0) ? 1 : (l == 1) ? N : N * N;
    for (int b = 0; b < N; b++)
      for (int a = 0; a < N; a++) {
        int idx = vol_index (N, f, a, b, n);
        int j_l = (l == 0) ? idx % N : (l == 1) ? (idx / N) % N : idx / (N * N);
        int base = idx - j_l * stride;
        double val = scale2 * r[b * N + a];
        for (int i = 0; i < N; i++) Au_e[base + i * stride] += dt->D[j_l * N + i] * val;
      }
  }
}

/* Au += SIPG face terms, one flux "side" record at a time
 * (d4est_laplacian_flux_interface d4est_laplacian_flux.c:232-1014,
 *  d4est_laplacian_flux_boundary :23-230,
 *  d4est_laplacian_flux_sipg_interface_aux d4est_laplacian_flux_sipg.c:494-814,
 *  d4est_laplacian_flux_sipg_dirichlet_aux :15-237) */
void
d4bo_apply_flux (d4bo_ctx_t * c, const double *u, const double *ughost, double *Au)
{
  const d4b_mesh_desc_t *d = &c->d;
  for (int s = 0; s < d->n_sides; s++) {
    int faces_m = d->side_faces_m[s], faces_p = d->side_faces_p[s];
    int f_m = d->side_f_m[s], f_p = d->side_f_p[s], perm = d->side_perm[s];
    int M = (faces_m > faces_p) ? faces_m : faces_p;
    const int32_t *e_m = d->side_e_m + 4 * s, *e_p = d->side_e_p + 4 * s;
    const int32_t *sub = d->side_sub_order + 4 * s;
    int64_t S = d->side_mortar_stride[s];
    int dqd[4], off[4], offp[4], nqp[4], T = 0;
    int32_t em[4], ep[4];
    for (int i = 0; i < M; i++) {
      em[i] = e_m[(faces_m == 4) ? i : 0];
      ep[i] = (faces_p == 0) ? 0 : e_p[(faces_p == 4) ? sub[i] : 0];
      int qm = elem_deg_quad (c, em[i]);
      dqd[i] = qm;
      if (faces_p) { int qp = elem_deg_quad (c, ep[i]); if (qp > qm) dqd[i] = qp; }
      off[i] = T;
      T += (dqd[i] + 1) * (dqd[i] + 1);
    }
    for (int i = 0; i < M; i++) nqp[(M == 4) ? sub[i] : i] = (dqd[i] + 1) * (dqd[i] + 1);
    for (int j = 0, acc = 0; j < M; j++) { offp[j] = acc; acc += nqp[j]; }
    int m_big = (faces_m == 1 && M == 4), p_big = (faces_p == 1 && M == 4);

    for (int i = 0; i < M; i++) {
      if (em[i] < 0) continue;          /* ghost (-) element: owner rank updates it */
      int deg_m = elem_deg (c, em[i]);
      int Nq = dqd[i] + 1, nq2 = Nq * Nq;
      int hx = m_big ? 1 + (i & 1) : 0, hy = m_big ? 1 + (i >> 1) : 0;
      double um[FQ], dm[3 * FQ], dxm[3 * FQ];
      side_traces_on_mortar (c, elem_field (c, em[i], u, ughost), deg_m, f_m, dqd[i], hx, hy, um, dm);
      const double *sj = d->sj_m_mortar_quad + S + off[i];
      const double *hm = d->hm_mortar_quad + S + off[i];
      const double *hp = d->hp_mortar_quad + S + off[i];
      const double *nv[3], *drm[3][3];
      for (int a = 0; a < 3; a++) {
        nv[a] = d->n_m_mortar_quad + 3 * S + (int64_t) a * T + off[i];
        for (int b = 0; b < 3; b++)
          drm[a][b] = d->drst_dxyz_m_mortar_quad + 9 * S + (int64_t) (a + 3 * b) * T + off[i];
      }
      for (int j = 0; j < 3; j++)
        for (int k = 0; k < nq2; k++) {
          double g = 0.;
          for (int a = 0; a < 3; a++) g += drm[a][j][k] * dm[a * FQ + k];
          dxm[j * FQ + k] = m_big ? .5 * g : g;
        }
      double t13[FQ], t2[3 * FQ];
      if (faces_p == 0 && d->bc_type == D4B_BC_ROBIN) {
        /* Robin boundary, d4est_laplacian_flux_sipg.c:339-455: sj (coeff u_m - rhs), tested and lifted */
        const double *cf = d->robin_coeff_mortar_quad + S + off[i];
        const double *rh = d->robin_rhs_mortar_quad ? d->robin_rhs_mortar_quad + S + off[i] : NULL;
        for (int k = 0; k < nq2; k++) {
          t13[k] = sj[k] * (cf[k] * um[k] - (rh ? rh[k] : 0.));
          for (int l = 0; l < 3; l++) t2[l * FQ + k] = 0.;
        }
        test_and_lift (c, Au + d->nodal_stride[em[i]], deg_m, f_m, dqd[i], 0, 0, t13, t2, 1.);
        continue;
      }
      if (faces_p == 0) {
        /* Dirichlet boundary, d4est_laplacian_flux_sipg.c:86-137 */
        const double *g = d->dirichlet_g_mortar_quad ? d->dirichlet_g_mortar_quad + S + off[i] : NULL;
        for (int k = 0; k < nq2; k++) {
          double sigma = d->side_penalty_mult[s]
            * penalty (d->sipg_penalty_fcn, deg_m, hm[k], deg_m, hm[k], d->sipg_penalty_prefactor);
          double du = um[k] - (g ? g[k] : 0.);
          double t1 = 0.;
          for (int a = 0; a < 3; a++) t1 += -1. * nv[a][k] * sj[k] * dxm[a * FQ + k];
          for (int l = 0; l < 3; l++) {
            double t = 0.;
            for (int a = 0; a < 3; a++) t += -.5 * drm[l][a][k] * nv[a][k] * sj[k] * 2. * du;
            t2[l * FQ + k] = t;
          }
          t13[k] = t1 + sj[k] * sigma * du;
        }
        test_and_lift (c, Au + d->nodal_stride[em[i]], deg_m, f_m, dqd[i], 0, 0, t13, t2, 1.);
        continue;
      }
      /* (+) side in its own ordering, then reoriented at the Gauss points */
      int deg_p = elem_deg (c, ep[i]);
      int j = (M == 4) ? sub[i] : 0;
      int hxp = p_big ? 1 + (j & 1) : 0, hyp = p_big ? 1 + (j >> 1) : 0;
      double up_po[FQ], dp_po[3 * FQ], dxp_po[FQ], up[FQ], dxp[3 * FQ];
      side_traces_on_mortar (c, elem_field (c, ep[i], u, ughost), deg_p, f_p, dqd[i], hxp, hyp, up_po, dp_po);
      reorient (up_po, Nq, perm, up);
      for (int jj = 0; jj < 3; jj++) {
        for (int k = 0; k < nq2; k++) {
          double g = 0.;
          for (int a = 0; a < 3; a++)
            g += d->drst_dxyz_p_mortar_quad_porder[9 * S + (int64_t) (a + 3 * jj) * T + offp[j] + k] * dp_po[a * FQ + k];
          dxp_po[k] = p_big ? .5 * g : g;
        }
        reorient (dxp_po, Nq, perm, dxp + jj * FQ);
      }
      /* sigma uses the degrees of the elements facing this mortar
       * (d4est_laplacian_flux_sipg.c:568-582) */
      for (int k = 0; k < nq2; k++) {
        double sigma = d->side_penalty_mult[s]
          * penalty (d->sipg_penalty_fcn, deg_m, hm[k], deg_p, hp[k], d->sipg_penalty_prefactor);
        double du = um[k] - up[k];
        double t1 = 0.;
        for (int a = 0; a < 3; a++)
          t1 += -1. * nv[a][k] * sj[k] * .5 * (dxp[a * FQ + k] + dxm[a * FQ + k]);
        for (int l = 0; l < 3; l++) {
          double t = 0.;
          for (int a = 0; a < 3; a++) t += -.5 * drm[l][a][k] * sj[k] * nv[a][k] * du;
          t2[l * FQ + k] = t;
        }
        t13[k] = t1 + sj[k] * sigma * du;
      }
      test_and_lift (c, Au + d->nodal_stride[em[i]], deg_m, f_m, dqd[i], hx, hy, t13, t2, m_big ? .5 : 1.);
    }
  }
}

/* Au = A u (d4est_laplacian_apply_aij, src/dGMath/d4est_laplacian.c:317-399).
 * ughost: ghost-element nodal values in ghost-id order (what
 * d4est_ghost_data_exchange delivers); NULL on one rank. */
void
d4bo_apply (d4bo_ctx_t * c, const double *u, const double *ughost, double *Au)
{
  d4bo_apply_stiffness (c, u, Au);
  d4bo_apply_flux (c, u, ughost, Au);
}
