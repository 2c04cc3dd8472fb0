/* d4est_b200_adapter.c -- see d4est_b200_adapter.h.
 *
 * Reference-side binding (C99, compiled with -DDIM=3 against the reference's
 * headers).  Nothing here does arithmetic on the hot path: it records the
 * face-iteration topology once per mesh and hands reference-owned arrays to
 * libd4est_b200 by pointer.
 */
#include <stdlib.h>
#include <string.h>
#include <d4est_util.h>
#include <d4est_reference.h>
#include <d4est_lgl.h>
#include "d4est_b200_adapter.h"

#if (P4EST_DIM) != 3
#error "the B200 path is 3-D only (d8est); build with -DDIM=3"
#endif

/* ------------------------------------------------------------------ sides */

static void
side_reserve (d4est_b200_flat_t * fl, int n)
{
  if (n <= fl->side_capacity) return;
  int cap = fl->side_capacity ? 2 * fl->side_capacity : 1024;
  while (cap < n) cap *= 2;
#define GROW(p, T, k) fl->p = (T *) realloc (fl->p, sizeof (T) * (size_t) (k) * cap)
  GROW (side_faces_m, int32_t, 1);
  GROW (side_faces_p, int32_t, 1);
  GROW (side_f_m, int32_t, 1);
  GROW (side_f_p, int32_t, 1);
  GROW (side_perm, int32_t, 1);
  GROW (side_e_m, int32_t, 4);
  GROW (side_e_p, int32_t, 4);
  GROW (side_sub_order, int32_t, 4);
  GROW (side_mortar_stride, int64_t, 1);
  GROW (side_penalty_mult, double, 1);
#undef GROW
  fl->side_capacity = cap;
}

typedef struct
{
  d4est_b200_flat_t *fl;
  d4est_mesh_data_t *factors;
  d4est_laplacian_flux_sipg_params_t *sipg;
  int mpirank;
} iter_ctx_t;

/* same 3 bits d4est_operators_reorient_face_data derives
 * (src/dGMath/d4est_operators.c:2030-2044) */
static int
face_perm_code (int f_m, int f_p, int o)
{
  int ftransform[9];
  p4est_expand_face_transform (((f_m <= f_p) ? f_m : f_p),
                               (P4EST_FACES) * o + ((f_m <= f_p) ? f_p : f_m),
                               ftransform);
  int flip0 = ftransform[6];
  int flip1 = ftransform[7];
  int aligned = ((ftransform[1] - ftransform[0])
                 * (ftransform[4] - ftransform[3]) > 0);
  return (flip0 ? 1 : 0) | (flip1 ? 2 : 0) | (aligned ? 0 : 4);
}

static int32_t
elem_ref (d4est_element_data_t * ed, int is_ghost, d4est_element_data_t * ghost_base)
{
  if (is_ghost) return -(int32_t) (ed - ghost_base) - 1;
  return (int32_t) ed->id;
}

static void
push_side (iter_ctx_t * ic, d4est_element_data_t ** e_m, int *m_ghost, int faces_m,
           int f_m, d4est_element_data_t ** e_p, int *p_ghost, int faces_p, int f_p,
           int orientation, d4est_element_data_t * ghost_base)
{
  d4est_b200_flat_t *fl = ic->fl;
  int s = fl->desc.n_sides;
  side_reserve (fl, s + 1);
  fl->side_faces_m[s] = faces_m;
  fl->side_faces_p[s] = faces_p;
  fl->side_f_m[s] = f_m;
  fl->side_f_p[s] = f_p;
  int use_id = -1;
  for (int i = 0; i < 4; i++) {
    fl->side_e_m[4 * s + i] = (i < faces_m) ? elem_ref (e_m[i], m_ghost[i], ghost_base) : INT32_MIN;
    fl->side_e_p[4 * s + i] = (i < faces_p) ? elem_ref (e_p[i], p_ghost[i], ghost_base) : INT32_MIN;
    fl->side_sub_order[4 * s + i] = i;
    if (i < faces_m && use_id < 0 && !m_ghost[i]) use_id = e_m[i]->id;
  }
  fl->side_perm[s] = 0;
  if (faces_p > 0) {
    fl->side_perm[s] = face_perm_code (f_m, f_p, orientation);
    int faces_mortar = (faces_m > faces_p) ? faces_m : faces_p;
    if (faces_mortar == (P4EST_HALF)) {
      for (int i = 0; i < 4; i++)
        fl->side_sub_order[4 * s + i] =
          d4est_reference_reorient_face_order ((P4EST_DIM) - 1, f_m, f_p, orientation, i);
    }
  }
  fl->side_mortar_stride[s] = ic->factors->local_strides[use_id].mortar_quad_stride[f_m];
  int region_m = e_m[0]->region;
  int region_p = (faces_p > 0) ? e_p[0]->region : region_m;
  fl->side_penalty_mult[s] = d4est_laplacian_flux_sipg_get_multiplier (ic->sipg, region_m, region_p);
  fl->desc.n_sides = s + 1;
}

/* mirrors the side selection of d4est_mortars_compute_flux_on_local_elements_aux
 * (src/Mesh/d4est_mortars.c:601-802) */
static void
record_face (p4est_iter_face_info_t * info, void *user_data)
{
  iter_ctx_t *ic = (iter_ctx_t *) info->p4est->user_pointer;
  d4est_element_data_t *ghost_data = (d4est_element_data_t *) user_data;
  d4est_element_data_t *e_m[(P4EST_HALF)], *e_p[(P4EST_HALF)];
  int m_ghost[(P4EST_HALF)], p_ghost[(P4EST_HALF)];
  sc_array_t *sides = &(info->sides);

  if (sides->elem_count == 2) {
    p4est_iter_face_side_t *side[2];
    side[0] = p4est_iter_fside_array_index_int (sides, 0);
    side[1] = p4est_iter_fside_array_index_int (sides, 1);
    for (int s_m = 0; s_m < 2; s_m++) {
      int s_p = (s_m == 0) ? 1 : 0;
      for (int i = 0; i < (P4EST_HALF); i++) m_ghost[i] = p_ghost[i] = 0;
      if (side[s_m]->is_hanging) {
        int nghost = 0;
        for (int i = 0; i < (P4EST_HALF); i++) {
          if (!side[s_m]->is.hanging.is_ghost[i]) {
            e_m[i] = (d4est_element_data_t *) side[s_m]->is.hanging.quad[i]->p.user_data;
          }
          else {
            e_m[i] = &ghost_data[side[s_m]->is.hanging.quadid[i]];
            m_ghost[i] = 1;
            nghost++;
          }
        }
        if (side[s_p]->is.full.is_ghost) {
          e_p[0] = &ghost_data[side[s_p]->is.full.quadid];
          p_ghost[0] = 1;
        }
        else
          e_p[0] = (d4est_element_data_t *) side[s_p]->is.full.quad->p.user_data;
        if (nghost < (P4EST_HALF))
          push_side (ic, e_m, m_ghost, (P4EST_HALF), side[s_m]->face, e_p, p_ghost, 1,
                     side[s_p]->face, info->orientation, ghost_data);
      }
      else if (!side[s_m]->is.full.is_ghost) {
        e_m[0] = (d4est_element_data_t *) side[s_m]->is.full.quad->p.user_data;
        if (side[s_p]->is_hanging) {
          for (int i = 0; i < (P4EST_HALF); i++) {
            if (side[s_p]->is.hanging.is_ghost[i]) {
              e_p[i] = &ghost_data[side[s_p]->is.hanging.quadid[i]];
              p_ghost[i] = 1;
            }
            else
              e_p[i] = (d4est_element_data_t *) side[s_p]->is.hanging.quad[i]->p.user_data;
          }
          push_side (ic, e_m, m_ghost, 1, side[s_m]->face, e_p, p_ghost, (P4EST_HALF),
                     side[s_p]->face, info->orientation, ghost_data);
        }
        else {
          if (side[s_p]->is.full.is_ghost) {
            e_p[0] = &ghost_data[side[s_p]->is.full.quadid];
            p_ghost[0] = 1;
          }
          else
            e_p[0] = (d4est_element_data_t *) side[s_p]->is.full.quad->p.user_data;
          push_side (ic, e_m, m_ghost, 1, side[s_m]->face, e_p, p_ghost, 1,
                     side[s_p]->face, info->orientation, ghost_data);
        }
      }
    }
  }
  else {
    p4est_iter_face_side_t *side = p4est_iter_fside_array_index_int (sides, 0);
    e_m[0] = (d4est_element_data_t *) side->is.full.quad->p.user_data;
    m_ghost[0] = 0;
    push_side (ic, e_m, m_ghost, 1, side->face, e_p, p_ghost, 0, -1, 0, ghost_data);
  }
}

/* The penalty functions are file-static in the reference
 * (d4est_laplacian_flux_sipg.c:945-1005); identify the pointer by probing it
 * at one point where the four functions take four different values. */
static int
identify_penalty (penalty_calc_t fcn)
{
  double v = fcn (2, 1.0, 3, 2.0, 1.0);
  if (fabs (v - 9.0) < 1e-12) return D4B_PENALTY_MAXP_SQR_OVER_MINH;
  if (fabs (v - 16.0) < 1e-12) return D4B_PENALTY_MAXPP1_SQR_OVER_MINH;
  if (fabs (v - 6.25 / 1.5) < 1e-12) return D4B_PENALTY_MEANP_SQR_OVER_MEANH;
  if (fabs (v - 4.25) < 1e-12) return D4B_PENALTY_MEAN_P_SQR_OVER_H;
  D4EST_ABORT ("d4est_b200: unknown sipg_penalty_fcn");
  return -1;
}

d4est_b200_flat_t *
d4est_b200_flatten (p4est_t * p4est, d4est_ghost_t * d4est_ghost,
                    d4est_mesh_data_t * d4est_factors,
                    d4est_laplacian_flux_data_t * flux_data)
{
  if (flux_data->flux_type != FLUX_SIPG)
    D4EST_ABORT ("d4est_b200: only the SIPG flux is implemented");
  d4est_laplacian_flux_sipg_params_t *sipg =
    (d4est_laplacian_flux_sipg_params_t *) flux_data->flux_data;

  d4est_b200_flat_t *fl = (d4est_b200_flat_t *) calloc (1, sizeof (*fl));
  d4b_mesh_desc_t *D = &fl->desc;
  D->abi_version = D4B_ABI_VERSION;
  int n_local = p4est->local_num_quadrants;
  int n_ghost = (int) d4est_ghost->ghost->ghosts.elem_count;
  D->n_local = n_local;
  D->n_ghost = n_ghost;
  D->local_nodes = d4est_factors->local_sizes.local_nodes;
  D->local_nodes_quad = d4est_factors->local_sizes.local_nodes_quad;
  D->local_mortar_nodes_quad = d4est_factors->local_sizes.local_mortar_nodes_quad;

  fl->deg = (int32_t *) malloc (sizeof (int32_t) * (n_local + 1));
  fl->deg_quad = (int32_t *) malloc (sizeof (int32_t) * (n_local + 1));
  fl->nodal_stride = (int64_t *) malloc (sizeof (int64_t) * (n_local + 1));
  fl->quad_stride = (int64_t *) malloc (sizeof (int64_t) * (n_local + 1));
  for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
    p4est_tree_t *tree = p4est_tree_array_index (p4est->trees, tt);
    sc_array_t *tquadrants = &tree->quadrants;
    int Q = (p4est_locidx_t) tquadrants->elem_count;
    for (int q = 0; q < Q; ++q) {
      p4est_quadrant_t *quad = p4est_quadrant_array_index (tquadrants, q);
      d4est_element_data_t *ed = (d4est_element_data_t *) quad->p.user_data;
      fl->deg[ed->id] = ed->deg;
      fl->deg_quad[ed->id] = ed->deg_quad;
      fl->nodal_stride[ed->id] = ed->nodal_stride;
      fl->quad_stride[ed->id] = ed->quad_stride;
    }
  }
  fl->ghost_deg = (int32_t *) malloc (sizeof (int32_t) * (n_ghost + 1));
  fl->ghost_deg_quad = (int32_t *) malloc (sizeof (int32_t) * (n_ghost + 1));
  fl->ghost_nodal_stride = (int64_t *) malloc (sizeof (int64_t) * (n_ghost + 1));
  int64_t gstride = 0;
  for (int g = 0; g < n_ghost; g++) {
    d4est_element_data_t *ged = &d4est_ghost->ghost_elements[g];
    fl->ghost_deg[g] = ged->deg;
    fl->ghost_deg_quad[g] = ged->deg_quad;
    fl->ghost_nodal_stride[g] = gstride;
    gstride += d4est_lgl_get_nodes ((P4EST_DIM), ged->deg);
  }
  D->ghost_nodes = gstride;

  D->deg = fl->deg;
  D->deg_quad = fl->deg_quad;
  D->nodal_stride = fl->nodal_stride;
  D->quad_stride = fl->quad_stride;
  D->ghost_deg = fl->ghost_deg;
  D->ghost_deg_quad = fl->ghost_deg_quad;
  D->ghost_nodal_stride = fl->ghost_nodal_stride;

  /* aliases of reference arrays */
  D->J_quad = d4est_factors->J_quad;
  D->rst_xyz_quad = d4est_factors->rst_xyz_quad;
  D->sj_m_mortar_quad = d4est_factors->sj_m_mortar_quad;
  D->n_m_mortar_quad = d4est_factors->n_m_mortar_quad;
  D->drst_dxyz_m_mortar_quad = d4est_factors->drst_dxyz_m_mortar_quad;
  D->drst_dxyz_p_mortar_quad_porder = d4est_factors->drst_dxyz_p_mortar_quad_porder;
  D->hm_mortar_quad = d4est_factors->hm_mortar_quad;
  D->hp_mortar_quad = d4est_factors->hp_mortar_quad;
  D->dirichlet_g_mortar_quad = NULL;

  /* side records */
  iter_ctx_t ic;
  ic.fl = fl;
  ic.factors = d4est_factors;
  ic.sipg = sipg;
  ic.mpirank = p4est->mpirank;
  void *tmpptr = p4est->user_pointer;
  p4est->user_pointer = &ic;
  p4est_iterate (p4est, d4est_ghost->ghost, (void *) d4est_ghost->ghost_elements,
                 NULL, record_face, NULL, NULL);
  p4est->user_pointer = tmpptr;
  D->side_faces_m = fl->side_faces_m;
  D->side_faces_p = fl->side_faces_p;
  D->side_f_m = fl->side_f_m;
  D->side_f_p = fl->side_f_p;
  D->side_e_m = fl->side_e_m;
  D->side_e_p = fl->side_e_p;
  D->side_perm = fl->side_perm;
  D->side_sub_order = fl->side_sub_order;
  D->side_mortar_stride = fl->side_mortar_stride;
  D->side_penalty_mult = fl->side_penalty_mult;

  D->sipg_penalty_prefactor = sipg->sipg_penalty_prefactor;
  D->sipg_penalty_fcn = identify_penalty (sipg->sipg_penalty_fcn);
  D->bc_type = (flux_data->bc_type == BC_ROBIN) ? D4B_BC_ROBIN : D4B_BC_DIRICHLET;
  if (flux_data->bc_type == BC_ROBIN) {
    /* the reference evaluates robin_coeff / robin_rhs at the boundary quadrature points inside every
     * apply (d4est_laplacian_flux_sipg.c:385-410); they depend on the position only in every shipped
     * problem (multi_puncture_fcns.h:772-804), so they are evaluated once per mesh here.  The
     * callbacks also receive the boundary data of the face: position, normal and sj are provided. */
    d4est_laplacian_robin_bc_t *bc = (d4est_laplacian_robin_bc_t *) flux_data->bc_data;
    fl->robin_coeff = (double *) calloc (D->local_mortar_nodes_quad + 1, sizeof (double));
    fl->robin_rhs = (double *) calloc (D->local_mortar_nodes_quad + 1, sizeof (double));
    for (int s = 0; s < D->n_sides; s++) {
      if (fl->side_faces_p[s] != 0) continue;
      const int e = fl->side_e_m[4 * s], f_m = fl->side_f_m[s];
      const int fnq = d4est_lgl_get_nodes ((P4EST_DIM) - 1, fl->deg_quad[e]);
      const int64_t S = fl->side_mortar_stride[s];
      const int bstride = (P4EST_DIM) * d4est_factors->local_strides[e].boundary_quad_stride[f_m];
      d4est_laplacian_flux_boundary_data_t bd;
      memset (&bd, 0, sizeof (bd));
      bd.face_nodes_m_quad = fnq;
      bd.deg_mortar_quad = fl->deg_quad[e];
      bd.sj_on_f_m_quad = &d4est_factors->sj_m_mortar_quad[S];
      bd.h_quad = &d4est_factors->hm_mortar_quad[S];
      for (int d = 0; d < (P4EST_DIM); d++) {
        bd.xyz_on_f_m_quad[d] = &d4est_factors->xyz_m_mortar_quad[bstride + d * fnq];
        bd.xyz_on_f_m_lobatto[d] = &d4est_factors->xyz_m_mortar_lobatto[bstride + d * fnq];
        bd.n_on_f_m_quad[d] = &d4est_factors->n_m_mortar_quad[(P4EST_DIM) * S + d * fnq];
      }
      for (int i = 0; i < fnq; i++) {
        const double x = bd.xyz_on_f_m_quad[0][i], y = bd.xyz_on_f_m_quad[1][i], z = bd.xyz_on_f_m_quad[2][i];
        fl->robin_coeff[S + i] = bc->robin_coeff (x, y, z, bc->user, &bd, i);
        fl->robin_rhs[S + i] = bc->robin_rhs (x, y, z, bc->user, &bd, i);
      }
    }
    D->robin_coeff_mortar_quad = fl->robin_coeff;
    D->robin_rhs_mortar_quad = fl->robin_rhs;
  }

  /* ghost exchange plan, as d4est_ghost_data_exchange reads it
   * (src/Mesh/d4est_ghost_data.c:171-199) */
  p4est_ghost_t *gh = d4est_ghost->ghost;
  D->mpirank = p4est->mpirank;
  D->mpisize = p4est->mpisize;
  D->n_mirrors = (int) gh->mirrors.elem_count;
  fl->mirror_elem = (int32_t *) malloc (sizeof (int32_t) * (D->n_mirrors + 1));
  for (int m = 0; m < D->n_mirrors; m++)
    fl->mirror_elem[m] = d4est_ghost->mirror_elements[m].id;
  fl->mirror_proc_offsets = (int32_t *) malloc (sizeof (int32_t) * (D->mpisize + 1));
  fl->ghost_proc_offsets = (int32_t *) malloc (sizeof (int32_t) * (D->mpisize + 1));
  for (int r = 0; r <= D->mpisize; r++) {
    fl->mirror_proc_offsets[r] = gh->mirror_proc_offsets[r];
    fl->ghost_proc_offsets[r] = gh->proc_offsets[r];
  }
  int nmpm = gh->mirror_proc_offsets[D->mpisize];
  fl->mirror_proc_mirrors = (int32_t *) malloc (sizeof (int32_t) * (nmpm + 1));
  for (int i = 0; i < nmpm; i++)
    fl->mirror_proc_mirrors[i] = gh->mirror_proc_mirrors[i];
  D->mirror_elem = fl->mirror_elem;
  D->mirror_proc_offsets = fl->mirror_proc_offsets;
  D->mirror_proc_mirrors = fl->mirror_proc_mirrors;
  D->ghost_proc_offsets = fl->ghost_proc_offsets;
  return fl;
}

void
d4est_b200_flat_destroy (d4est_b200_flat_t * fl)
{
  if (fl == NULL) return;
  free (fl->deg); free (fl->deg_quad); free (fl->nodal_stride); free (fl->quad_stride);
  free (fl->ghost_deg); free (fl->ghost_deg_quad); free (fl->ghost_nodal_stride);
  free (fl->side_faces_m); free (fl->side_faces_p); free (fl->side_f_m); free (fl->side_f_p);
  free (fl->side_e_m); free (fl->side_e_p); free (fl->side_perm); free (fl->side_sub_order);
  free (fl->side_mortar_stride); free (fl->side_penalty_mult); free (fl->dirichlet_g);
  free (fl->robin_coeff); free (fl->robin_rhs);
  free (fl->mirror_elem); free (fl->mirror_proc_offsets); free (fl->mirror_proc_mirrors);
  free (fl->ghost_proc_offsets);
  free (fl);
}

/* ------------------------------------------------------------ apply_lhs */

#ifndef D4EST_B200_ADAPTER_NO_DEVICE
void
d4est_b200_comm_init_from_p4est (p4est_t * p4est, int device)
{
  char id[128];
  d4b_init (device);
  if (p4est->mpisize <= 1) return;
  memset (id, 0, sizeof (id));
  if (p4est->mpirank == 0) d4b_comm_get_unique_id (id);
  int mpiret = sc_MPI_Bcast (id, (int) sizeof (id), sc_MPI_BYTE, 0, p4est->mpicomm);
  SC_CHECK_MPI (mpiret);
  d4b_comm_init (id, p4est->mpirank, p4est->mpisize);
}

void
d4est_b200_comm_finalize (void)
{
  d4b_comm_destroy ();
}
#endif

#ifndef D4EST_B200_ADAPTER_NO_DEVICE
void
d4est_b200_pin_vectors (d4est_elliptic_data_t * prob_vecs)
{
  double *v[3] = { prob_vecs->u, prob_vecs->Au, prob_vecs->rhs };
  for (int i = 0; i < 3; i++)
    if (v[i] != NULL && (i == 0 || (v[i] != v[0] && (i == 1 || v[i] != v[1]))))
      (void) d4b_host_pin (v[i], (int64_t) prob_vecs->local_nodes * (int64_t) sizeof (double));
}

void
d4est_b200_unpin_vectors (d4est_elliptic_data_t * prob_vecs)
{
  double *v[3] = { prob_vecs->u, prob_vecs->Au, prob_vecs->rhs };
  for (int i = 0; i < 3; i++)
    if (v[i] != NULL && (i == 0 || (v[i] != v[0] && (i == 1 || v[i] != v[1]))))
      (void) d4b_host_unpin (v[i]);
}
#endif

d4est_b200_ctx_t *
d4est_b200_ctx_new (d4est_laplacian_flux_data_t * flux_data)
{
  d4est_b200_ctx_t *ctx = (d4est_b200_ctx_t *) calloc (1, sizeof (*ctx));
  ctx->flux_data = flux_data;
  return ctx;
}

void
d4est_b200_ctx_invalidate (d4est_b200_ctx_t * ctx)
{
  if (ctx == NULL) return;
#ifndef D4EST_B200_ADAPTER_NO_DEVICE
  for (int i = 0; i < D4EST_B200_CACHE; i++)
    if (ctx->cache[i].mesh) d4b_mesh_destroy (ctx->cache[i].mesh);
#endif
  memset (ctx->cache, 0, sizeof (ctx->cache));
}

void
d4est_b200_ctx_destroy (d4est_b200_ctx_t * ctx)
{
  if (ctx == NULL) return;
  d4est_b200_ctx_invalidate (ctx);
  free (ctx);
}

/* ---- identity of the reference's current mesh (see d4est_b200_cached_mesh_t) ----
 * One walk over the local quadrants: position in the forest and the two degrees of every element.
 * (d4est_mesh_data_t::local_strides is no substitute: its nodal / quadrature strides are never filled,
 * src/Mesh/d4est_mesh.c:502-503, :962-963 set the mortar strides only.) */
static uint64_t
mesh_signature (p4est_t * p4est)
{
  uint64_t h[2] = { 0x5851F42D4C957F2Dull, 0x9E3779B97F4A7C15ull };
  for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
    p4est_tree_t *tree = p4est_tree_array_index (p4est->trees, tt);
    sc_array_t *tq = &tree->quadrants;
    h[0] = (h[0] ^ (uint64_t) tt) * 0x100000001B3ull;
    for (size_t q = 0; q < tq->elem_count; ++q) {
      const p4est_quadrant_t *quad = p4est_quadrant_array_index (tq, q);
      const d4est_element_data_t *ed = (const d4est_element_data_t *) quad->p.user_data;
      h[0] = (h[0] ^ (((uint64_t) (uint32_t) quad->x << 32) | (uint32_t) quad->y)) * 0x100000001B3ull;
      h[0] ^= h[0] >> 29;
      h[1] = (h[1] ^ (((uint64_t) (uint32_t) quad->z << 32) | ((uint32_t) quad->level << 16)
                      | ((uint32_t) ed->deg << 8) | (uint32_t) ed->deg_quad)) * 0xC2B2AE3D27D4EB4Full;
      h[1] ^= h[1] >> 31;
    }
  }
  return h[0] * 3 ^ h[1] * 5;
}

static void eval_dirichlet (d4est_b200_flat_t * fl, d4est_operators_t * d4est_ops, d4est_geometry_t * d4est_geom,
                           d4est_quadrature_t * d4est_quad, d4est_mesh_data_t * d4est_factors,
                           d4est_laplacian_dirichlet_bc_t * bc);

void
d4est_b200_flat_eval_dirichlet (d4est_b200_flat_t * fl, d4est_operators_t * d4est_ops,
                                d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                                d4est_mesh_data_t * d4est_factors, d4est_laplacian_flux_data_t * flux_data)
{
  if (flux_data->bc_type != BC_DIRICHLET) return;
  eval_dirichlet (fl, d4est_ops, d4est_geom, d4est_quad, d4est_factors,
                  (d4est_laplacian_dirichlet_bc_t *) flux_data->bc_data);
}

static void
eval_dirichlet (d4est_b200_flat_t * fl, d4est_operators_t * d4est_ops, d4est_geometry_t * d4est_geom,
                d4est_quadrature_t * d4est_quad, d4est_mesh_data_t * d4est_factors,
                d4est_laplacian_dirichlet_bc_t * bc)
{
  free (fl->dirichlet_g);
  fl->dirichlet_g = (double *) calloc (fl->desc.local_mortar_nodes_quad + 1, sizeof (double));
  for (int s = 0; s < fl->desc.n_sides; s++) {
    if (fl->desc.side_faces_p[s] != 0) continue;
    const int e = fl->desc.side_e_m[4 * s], f_m = fl->desc.side_f_m[s];
    const int deg = fl->desc.deg[e], deg_quad = fl->desc.deg_quad[e];
    const int fn = d4est_lgl_get_nodes ((P4EST_DIM) - 1, deg), fnq = d4est_lgl_get_nodes ((P4EST_DIM) - 1, deg_quad);
    const int bstride = (P4EST_DIM) * d4est_factors->local_strides[e].boundary_quad_stride[f_m];
    double *g = &fl->dirichlet_g[fl->desc.side_mortar_stride[s]];
    if (bc->eval_method == EVAL_BNDRY_FCN_ON_QUAD) {
      for (int i = 0; i < fnq; i++)
        g[i] = bc->dirichlet_fcn (d4est_factors->xyz_m_mortar_quad[bstride + 0 * fnq + i],
                                  d4est_factors->xyz_m_mortar_quad[bstride + 1 * fnq + i],
                                  d4est_factors->xyz_m_mortar_quad[bstride + 2 * fnq + i], bc->user);
    }
    else {
      double *g_lob = (double *) malloc (sizeof (double) * fn);
      for (int i = 0; i < fn; i++)
        g_lob[i] = bc->dirichlet_fcn (d4est_factors->xyz_m_mortar_lobatto[bstride + 0 * fnq + i],
                                      d4est_factors->xyz_m_mortar_lobatto[bstride + 1 * fnq + i],
                                      d4est_factors->xyz_m_mortar_lobatto[bstride + 2 * fnq + i], bc->user);
      d4est_quadrature_mortar_t face_object;
      memset (&face_object, 0, sizeof (face_object));
      d4est_quadrature_interpolate (d4est_ops, d4est_quad, d4est_geom, &face_object, QUAD_OBJECT_MORTAR,
                                    QUAD_INTEGRAND_UNKNOWN, g_lob, deg, g, deg_quad);
      free (g_lob);
    }
  }
  fl->desc.dirichlet_g_mortar_quad = fl->dirichlet_g;
}

void
d4est_b200_geom_desc_fill (d4b_geom_desc_t * g, p4est_t * p4est, d4est_ghost_t * ghost,
                           d4est_geometry_t * d4est_geom, d4est_mesh_face_h_t face_h_type)
{
  memset (g, 0, sizeof (*g));
  if (d4est_geom->DX_compute_method != GEOM_COMPUTE_ANALYTIC || d4est_geom->JAC_compute_method != GEOM_COMPUTE_NUMERICAL)
    D4EST_ABORT ("d4est_b200: device geometry needs DX_compute_method = analytic, JAC_compute_method = numerical");
  if (d4est_geom->geom_type == GEOM_BRICK) {
    d4est_geometry_brick_attr_t *b = (d4est_geometry_brick_attr_t *) d4est_geom->user;
    g->geom_type = D4B_GEOM_BRICK;
    g->X0 = b->X0; g->X1 = b->X1; g->Y0 = b->Y0; g->Y1 = b->Y1; g->Z0 = b->Z0; g->Z1 = b->Z1;
  }
  else if (d4est_geom->geom_type == GEOM_CUBED_SPHERE_13TREE) {
    d4est_geometry_cubed_sphere_attr_t *sp = (d4est_geometry_cubed_sphere_attr_t *) d4est_geom->user;
    g->geom_type = D4B_GEOM_CUBED_SPHERE;
    g->R0 = sp->R0; g->R1 = sp->R1; g->R2 = sp->R2;
    g->compactify_inner_shell = sp->compactify_inner_shell;
    g->compactify_outer_shell = sp->compactify_outer_shell;
  }
  else
    D4EST_ABORT ("d4est_b200: device geometry supports the brick and the 13-tree cubed sphere");
  g->face_h_type = (int32_t) face_h_type;   /* d4b_face_h_t mirrors d4est_mesh_face_h_t */
  const int nl = p4est->local_num_quadrants, ng = (int) ghost->ghost->ghosts.elem_count;
  int32_t *et = (int32_t *) malloc (sizeof (int32_t) * (nl > 0 ? nl : 1));
  int32_t *eq = (int32_t *) malloc (sizeof (int32_t) * 3 * (nl > 0 ? nl : 1));
  int32_t *ed_ = (int32_t *) malloc (sizeof (int32_t) * (nl > 0 ? nl : 1));
  int32_t *gt = (int32_t *) malloc (sizeof (int32_t) * (ng > 0 ? ng : 1));
  int32_t *gq = (int32_t *) malloc (sizeof (int32_t) * 3 * (ng > 0 ? ng : 1));
  int32_t *gd = (int32_t *) malloc (sizeof (int32_t) * (ng > 0 ? ng : 1));
  for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
    p4est_tree_t *tree = p4est_tree_array_index (p4est->trees, tt);
    sc_array_t *tquadrants = &tree->quadrants;
    for (int qq = 0; qq < (int) tquadrants->elem_count; ++qq) {
      p4est_quadrant_t *quad = p4est_quadrant_array_index (tquadrants, qq);
      d4est_element_data_t *ed = (d4est_element_data_t *) (quad->p.user_data);
      et[ed->id] = ed->tree; ed_[ed->id] = ed->dq;
      for (int d = 0; d < 3; d++) eq[3 * ed->id + d] = ed->q[d];
    }
  }
  for (int i = 0; i < ng; i++) {
    d4est_element_data_t *ged = &ghost->ghost_elements[i];
    gt[i] = ged->tree; gd[i] = ged->dq;
    for (int d = 0; d < 3; d++) gq[3 * i + d] = ged->q[d];
  }
  g->elem_tree = et; g->elem_q = eq; g->elem_dq = ed_;
  g->ghost_tree = gt; g->ghost_q = gq; g->ghost_dq = gd;
}

void
d4est_b200_geom_desc_free (d4b_geom_desc_t * g)
{
  free ((void *) g->elem_tree); free ((void *) g->elem_q); free ((void *) g->elem_dq);
  free ((void *) g->ghost_tree); free ((void *) g->ghost_q); free ((void *) g->ghost_dq);
  memset (g, 0, sizeof (*g));
}

void
d4est_b200_ctx_use_device_geometry (d4est_b200_ctx_t * ctx, d4est_geometry_t * d4est_geom,
                                    d4est_mesh_face_h_t face_h_type)
{
  ctx->device_geom = d4est_geom;
  ctx->device_face_h_type = (int) face_h_type;
}

#ifndef D4EST_B200_ADAPTER_NO_DEVICE
d4b_mesh_t *
d4est_b200_get_mesh (d4est_b200_ctx_t * ctx, p4est_t * p4est, d4est_ghost_t * ghost,
                     d4est_mesh_data_t * d4est_factors)
{
  const int ne = p4est->local_num_quadrants, nn = d4est_factors->local_sizes.local_nodes;
  const int nq = d4est_factors->local_sizes.local_nodes_quad, nm = d4est_factors->local_sizes.local_mortar_nodes_quad;
  const int has_bc = ctx->bc_ops != NULL;
  int victim = 0;
  int have_sig = 0;
  uint64_t sig = 0;
  ctx->clock++;
  for (int i = 0; i < D4EST_B200_CACHE; i++) {
    d4est_b200_cached_mesh_t *c = &ctx->cache[i];
    if (!(c->mesh && c->factors == d4est_factors && c->local_elements == ne && c->local_nodes == nn
          && c->local_nodes_quad == nq && c->local_mortar_nodes_quad == nm)) continue;
    /* same object and counts: is it still the same mesh?  (the reference's multigrid coarsens and
     * refines one p4est in place and re-runs d4est_mesh_update on one d4est_mesh_data_t after AMR) */
    if (!have_sig) { sig = mesh_signature (p4est); have_sig = 1; }
    const int same = (c->mesh_sig == sig);
    if (same && c->has_dirichlet_data == has_bc && c->device_geom == (void *) ctx->device_geom) {
      c->stamp = ctx->clock;
      return c->mesh;
    }
    /* the reference reused this d4est_mesh_data_t for another mesh (AMR step with equal counts), or the
     * context gained boundary data / device geometry: the cached operator is stale */
    d4b_mesh_destroy (c->mesh);
    memset (c, 0, sizeof (*c));
  }
  /* first free slot, else the least recently used one */
  victim = -1;
  for (int i = 0; i < D4EST_B200_CACHE && victim < 0; i++)
    if (ctx->cache[i].mesh == NULL) victim = i;
  if (victim < 0) {
    victim = 0;
    for (int i = 1; i < D4EST_B200_CACHE; i++)
      if (ctx->cache[i].stamp < ctx->cache[victim].stamp) victim = i;
    d4b_mesh_destroy (ctx->cache[victim].mesh);
  }
  d4est_b200_flat_t *fl = d4est_b200_flatten (p4est, ghost, d4est_factors, ctx->flux_data);
  d4est_b200_cached_mesh_t *c = &ctx->cache[victim];
  if (ctx->bc_ops)
    d4est_b200_flat_eval_dirichlet (fl, ctx->bc_ops, ctx->bc_geom, ctx->bc_quad, d4est_factors, ctx->flux_data);
  if (ctx->device_geom) {
    d4b_geom_desc_t gd;
    d4est_b200_geom_desc_fill (&gd, p4est, ghost, ctx->device_geom, (d4est_mesh_face_h_t) ctx->device_face_h_type);
    c->mesh = d4b_mesh_create_with_geometry (&fl->desc, &gd);
    d4est_b200_geom_desc_free (&gd);
  }
  else
    c->mesh = d4b_mesh_create (&fl->desc);
  d4est_b200_flat_destroy (fl);
  if (ctx->has_nonlinear) {
    d4b_nonlinear_desc_t nl = ctx->nonlinear;
    nl.xyz_quad = d4est_factors->xyz_quad;
    nl.jac_quad = d4est_factors->J_quad;
    d4b_mesh_set_nonlinear_term (c->mesh, &nl);
  }
  c->factors = d4est_factors;
  c->local_elements = ne;
  c->local_nodes = nn;
  c->local_nodes_quad = nq;
  c->local_mortar_nodes_quad = nm;
  c->mesh_sig = have_sig ? sig : mesh_signature (p4est);
  c->has_dirichlet_data = has_bc;
  c->device_geom = (void *) ctx->device_geom;
  c->stamp = ctx->clock;
  c->serial = ++ctx->serial;
  ctx->uploads++;
  return c->mesh;
}

void
d4est_b200_set_nonlinear_term (d4est_b200_ctx_t * ctx, const d4b_nonlinear_desc_t * desc, d4est_b200_ctx_t * residual_ctx)
{
  d4est_b200_ctx_invalidate (ctx);
  ctx->has_nonlinear = (desc != NULL);
  if (desc) ctx->nonlinear = *desc;
  ctx->residual_ctx = residual_ctx;
  ctx->lin_u0 = NULL; ctx->lin_mesh = NULL;
  if (residual_ctx && residual_ctx != ctx) {
    d4est_b200_ctx_invalidate (residual_ctx);
    residual_ctx->has_nonlinear = ctx->has_nonlinear;
    residual_ctx->nonlinear = ctx->nonlinear;
  }
}

/* strided checksum of a linearisation point: <= 4096 samples, position-weighted */
static double
u0_signature (const double *u0, int n)
{
  const int stride = n > 4096 ? n / 4096 : 1;
  double s = 0.;
  for (int i = 0, k = 1; i < n; i += stride, k++) s += u0[i] * (1. + 1e-3 * (k % 97));
  return s + u0[n - 1];
}

/* Newton: make `mesh` the Jacobian about u0 unless it already is (same address, size and checksum) */
static void
d4est_b200_ensure_linearised (d4est_b200_ctx_t * ctx, d4b_mesh_t * mesh, const double *u0, int n)
{
  if (!ctx->has_nonlinear || u0 == NULL || n <= 0) return;
  const double sig = u0_signature (u0, n);
  if (ctx->lin_mesh == mesh && ctx->lin_u0 == u0 && ctx->lin_nodes == n && ctx->lin_sig == sig) return;
  d4b_mesh_linearize (mesh, u0);
  ctx->lin_mesh = mesh; ctx->lin_u0 = u0; ctx->lin_nodes = n; ctx->lin_sig = sig;
  ctx->linearisations++;
}

void
d4est_b200_apply_lhs (p4est_t * p4est, d4est_ghost_t * ghost,
                      d4est_ghost_data_t * ghost_data,
                      d4est_elliptic_data_t * prob_vecs,
                      d4est_operators_t * d4est_ops,
                      d4est_geometry_t * d4est_geom,
                      d4est_quadrature_t * d4est_quad,
                      d4est_mesh_data_t * d4est_factors, void *user)
{
  d4est_b200_ctx_t *ctx = (d4est_b200_ctx_t *) user;
  d4b_mesh_t *mesh = d4est_b200_get_mesh (ctx, p4est, ghost, d4est_factors);
  ctx->applies++;
  /* the Jacobian of a nonlinear problem: only the mesh the Newton iterate lives on carries u0 */
  /* (inside the reference's own multigrid the coarse levels see the fine level's u0 pointer: skipped) */
  if (ctx->has_nonlinear && prob_vecs->u0 != NULL
      && (ctx->newton_nodes == 0 || ctx->newton_nodes == prob_vecs->local_nodes))
    d4est_b200_ensure_linearised (ctx, mesh, prob_vecs->u0, prob_vecs->local_nodes);
  d4b_apply (mesh, prob_vecs->u, prob_vecs->Au);
}

void
d4est_b200_build_residual (p4est_t * p4est, d4est_ghost_t * ghost, d4est_ghost_data_t * ghost_data,
                           d4est_elliptic_data_t * prob_vecs, d4est_operators_t * d4est_ops,
                           d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                           d4est_mesh_data_t * d4est_factors, void *user)
{
  d4est_b200_ctx_t *ctx = (d4est_b200_ctx_t *) user;
  d4est_b200_ctx_t *rctx = ctx->residual_ctx ? ctx->residual_ctx : ctx;
  d4b_mesh_t *mesh = d4est_b200_get_mesh (rctx, p4est, ghost, d4est_factors);
  (void) ghost_data; (void) d4est_ops; (void) d4est_geom; (void) d4est_quad;
  ctx->newton_nodes = prob_vecs->local_nodes;
  d4b_build_residual (mesh, prob_vecs->u, ctx->has_nonlinear ? NULL : prob_vecs->rhs, prob_vecs->Au);
}

/* ------------------------------------------------------------ right-hand side */
void
d4est_b200_build_rhs_with_strong_bc (d4est_b200_ctx_t * ctx, p4est_t * p4est, d4est_ghost_t * ghost,
                                     d4est_ghost_data_t * ghost_data, d4est_operators_t * d4est_ops,
                                     d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                                     d4est_mesh_data_t * d4est_factors, d4est_elliptic_data_t * prob_vecs,
                                     d4est_laplacian_flux_data_t * flux_fcn_data_for_build_rhs,
                                     double *rhs, d4est_xyz_fcn_t problem_rhs_fcn,
                                     d4est_mesh_init_field_option_t init_option, void *fcn_ctx, int which_field)
{
  (void) ghost_data; (void) prob_vecs; (void) which_field;
  if (ctx->flux_data != flux_fcn_data_for_build_rhs)
    D4EST_ABORT ("d4est_b200_build_rhs_with_strong_bc: the context was created from another flux data");
  ctx->bc_ops = d4est_ops; ctx->bc_geom = d4est_geom; ctx->bc_quad = d4est_quad;
  d4b_mesh_t *mesh = d4est_b200_get_mesh (ctx, p4est, ghost, d4est_factors);
  const int on_quad = (init_option == INIT_FIELD_ON_QUAD);
  if (!on_quad && init_option != INIT_FIELD_ON_LOBATTO) D4EST_ABORT ("Not a support init option");
  const size_t nf = on_quad ? (size_t) d4est_factors->local_sizes.local_nodes_quad
                            : (size_t) d4est_factors->local_sizes.local_nodes;
  double *f = (double *) malloc (sizeof (double) * (nf + 1));
  d4est_mesh_init_field (p4est, f, problem_rhs_fcn, d4est_ops, d4est_geom, d4est_factors, init_option, fcn_ctx);
  d4b_build_rhs (mesh, f, on_quad, d4est_factors->J_quad, rhs);
  free (f);
}

/* ------------------------------------------------------------ estimator */
static int
identify_estimator_family (d4est_estimator_bi_penalty_data_t * pd)
{
  /* values of the (gradu, u, u_dirichlet) callbacks at one point where the families differ */
  const int dm = 2, dp = 3;
  const double hm = 0.3, hp = 0.7, pre = 1.7;
  const double g = pd->gradu_penalty_fcn (dm, hm, dp, hp, pre), u = pd->u_penalty_fcn (dm, hm, dp, hp, pre);
  const double ud = pd->u_dirichlet_penalty_fcn (dm, hm, dp, hp, pre);
  const double maxp = 3., minh = 0.3, max_hop = 0.7 / 3., max_p2h = 4. / 0.3;
  const double fam[4][3] = {
    {sqrt (.5 * minh / maxp), sqrt (.5 * pre * maxp * maxp / minh), sqrt (pre * maxp * maxp / minh)},
    {sqrt (.5 * max_hop), sqrt (.5 * pre * max_p2h), sqrt (pre * max_p2h)},
    {sqrt (minh / maxp), sqrt (pre * maxp * maxp / minh), sqrt (pre * maxp * maxp / minh)},
    {sqrt (max_hop), sqrt (pre * max_p2h), sqrt (pre * max_p2h)}};
  for (int f = 0; f < 4; f++)
    if (fabs (g - fam[f][0]) < 1e-12 && fabs (u - fam[f][1]) < 1e-12 && fabs (ud - fam[f][2]) < 1e-12) return f;
  D4EST_ABORT ("d4est_b200: unknown combination of estimator penalty callbacks");
  return -1;
}

#ifndef D4EST_B200_ADAPTER_NO_DEVICE
double *
d4est_b200_estimator_bi_compute (d4est_b200_ctx_t * ctx, p4est_t * p4est,
                                 d4est_elliptic_data_t * d4est_elliptic_data, d4est_elliptic_eqns_t * fcns,
                                 d4est_estimator_bi_penalty_data_t penalty_data, d4est_xyz_fcn_t u_bndry_fcn,
                                 void *bndry_ctx, d4est_ghost_t * d4est_ghost, d4est_ghost_data_t * d4est_ghost_data,
                                 d4est_operators_t * d4est_ops, d4est_geometry_t * d4est_geom_for_residual,
                                 d4est_mesh_data_t * d4est_factors_for_residual,
                                 d4est_geometry_t * d4est_geom_for_integrals,
                                 d4est_mesh_data_t * d4est_factors_for_integrals, d4est_quadrature_t * d4est_quad,
                                 int which_field, double *estimator_vtk, double *estimator_vtk_per_face)
{
  (void) which_field;
  const int nl = p4est->local_num_quadrants;
  d4est_elliptic_eqns_build_residual (p4est, d4est_ghost, d4est_ghost_data, fcns, d4est_elliptic_data, d4est_ops,
                                      d4est_geom_for_residual, d4est_quad, d4est_factors_for_residual);
  d4b_mesh_t *mesh = d4est_b200_get_mesh (ctx, p4est, d4est_ghost, d4est_factors_for_integrals);
  /* boundary data the way d4est_estimator_bi_dirichlet evaluates it: on the face LGL nodes, then interpolated */
  d4est_b200_flat_t *fl = d4est_b200_flatten (p4est, d4est_ghost, d4est_factors_for_integrals, ctx->flux_data);
  d4est_laplacian_dirichlet_bc_t bc;
  bc.dirichlet_fcn = u_bndry_fcn; bc.user = bndry_ctx; bc.eval_method = EVAL_BNDRY_FCN_ON_LOBATTO;
  eval_dirichlet (fl, d4est_ops, d4est_geom_for_integrals, d4est_quad, d4est_factors_for_integrals, &bc);
  d4b_estimator_desc_t ed;
  memset (&ed, 0, sizeof (ed));
  ed.u = d4est_elliptic_data->u; ed.residual = d4est_elliptic_data->Au;
  ed.jac_quad = d4est_factors_for_integrals->J_quad;
  ed.diam_volume = d4est_factors_for_integrals->diam_volume;
  ed.sj_m_mortar_quad = d4est_factors_for_integrals->sj_m_mortar_quad;
  ed.hm_mortar_quad = d4est_factors_for_integrals->hm_mortar_quad;
  ed.hp_mortar_quad = d4est_factors_for_integrals->hp_mortar_quad;
  ed.dirichlet_g_mortar_quad = fl->dirichlet_g;
  ed.penalty_family = identify_estimator_family (&penalty_data);
  ed.penalty_prefactor = penalty_data.penalty_prefactor;
  double *estimator = P4EST_ALLOC (double, nl);
  d4b_estimator_bi (mesh, &ed, estimator, estimator_vtk);
  if (estimator_vtk_per_face)
    for (int i = 0; i < 18 * nl; i++) estimator_vtk_per_face[i] = 0.;
  d4est_b200_flat_destroy (fl);
  return estimator;
}
#endif

/* ------------------------------------------------------------ mass term */
void
d4est_b200_set_mass_term (d4est_b200_ctx_t * ctx, p4est_t * p4est, d4est_ghost_t * ghost,
                          d4est_operators_t * d4est_ops, d4est_geometry_t * d4est_geom,
                          d4est_quadrature_t * d4est_quad, d4est_mesh_data_t * d4est_factors,
                          double *u0, d4est_xyzu_fcn_t f, void *f_ctx)
{
  d4b_mesh_t *mesh = d4est_b200_get_mesh (ctx, p4est, ghost, d4est_factors);
  if (f == NULL) { d4b_mesh_set_mass_diag (mesh, NULL); return; }
  double *jf = (double *) malloc (sizeof (double) * (d4est_factors->local_sizes.local_nodes_quad + 1));
  for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
    p4est_tree_t *tree = p4est_tree_array_index (p4est->trees, tt);
    sc_array_t *tquadrants = &tree->quadrants;
    for (size_t q = 0; q < tquadrants->elem_count; ++q) {
      p4est_quadrant_t *quad = p4est_quadrant_array_index (tquadrants, q);
      d4est_element_data_t *ed = (d4est_element_data_t *) quad->p.user_data;
      d4est_quadrature_volume_t mo;
      mo.dq = ed->dq; mo.tree = ed->tree; mo.element_id = ed->id;
      mo.q[0] = ed->q[0]; mo.q[1] = ed->q[1];
#if (P4EST_DIM)==3
      mo.q[2] = ed->q[2];
#endif
      double *J_quad = d4est_mesh_get_jacobian_on_quadrature_points (d4est_factors, ed);
      d4est_mesh_data_on_element_t md = d4est_mesh_data_on_element (d4est_factors, ed);
      const int nq = d4est_lgl_get_nodes ((P4EST_DIM), ed->deg_quad);
      double *out = jf + ed->quad_stride;
      /* u0 at the quadrature points, then J f as d4est_quadrature.c:660-680 forms it */
      d4est_quadrature_interpolate (d4est_ops, d4est_quad, d4est_geom, &mo, QUAD_OBJECT_VOLUME,
                                    QUAD_INTEGRAND_UNKNOWN, &u0[ed->nodal_stride], ed->deg, out, ed->deg_quad);
      for (int k = 0; k < nq; k++)
        out[k] = J_quad[k] * f (md.xyz_quad[0][k], md.xyz_quad[1][k],
#if (P4EST_DIM)==3
                                md.xyz_quad[2][k],
#endif
                                out[k], f_ctx);
    }
  }
  d4b_mesh_set_mass_diag (mesh, jf);
  free (jf);
}

/* ------------------------------------------------------------ smoother */
#include <d4est_solver_multigrid_smoother_cheby.h>

/* one context per process is enough for the reference's single-threaded solvers */
static d4est_b200_ctx_t *d4est_b200_smoother_ctx = NULL;

static void
d4est_b200_smoother_cheby_smooth (p4est_t * p4est, d4est_elliptic_data_t * vecs,
                                  d4est_elliptic_eqns_t * fcns, double *r, int level)
{
  d4est_solver_multigrid_t *mg_data = (d4est_solver_multigrid_t *) p4est->user_pointer;
  d4est_solver_multigrid_smoother_cheby_t *cheby = mg_data->smoother->user;
  d4est_solver_multigrid_element_data_updater_t *updater = mg_data->elem_data_updater;
  d4b_mesh_t *mesh = d4est_b200_get_mesh (d4est_b200_smoother_ctx, p4est, updater->current_d4est_ghost,
                                          updater->current_d4est_factors);
  /* same decisions as d4est_solver_multigrid_smoother_cheby (:264-376); the flag
   * cheby_eigs_compute is maintained by the reference's own update callback (:222-262) */
  if (cheby->cheby_eigs_compute) {
    cheby->eigs[level] = d4b_cg_eigs (mesh, vecs->u, vecs->rhs, cheby->cheby_eigs_cg_imax,
                                      cheby->cheby_use_new_cg_eigs, cheby->cheby_use_zero_guess_for_eigs);
    cheby->eigs[level] *= cheby->cheby_eigs_max_multiplier;
  }
  if (cheby->cheby_use_zero_guess_for_eigs == 1 && cheby->cheby_eigs_reuse_fromdownvcycle != 1) {
    D4EST_ABORT ("If you set cheby_use_zero_guess_for_eigs == 1, please set cheby_eigs_reuse_fromdownvcycle = 1");
  }
  const double lmax = cheby->eigs[level];
  const double lmin = lmax / cheby->cheby_eigs_lmax_lmin_ratio;
  d4b_cheby_smooth (mesh, vecs->u, vecs->rhs, r, cheby->cheby_imax, lmin, lmax);
}

d4est_solver_multigrid_smoother_t *
d4est_b200_smoother_cheby_init (p4est_t * p4est, int num_of_levels, const char *input_file,
                                d4est_b200_ctx_t * ctx)
{
  /* the reference's factory parses [mg_smoother_cheby] and installs its update callback */
  d4est_solver_multigrid_smoother_t *smoother =
    d4est_solver_multigrid_smoother_cheby_init (p4est, num_of_levels, input_file);
  smoother->smooth = d4est_b200_smoother_cheby_smooth;
  d4est_b200_smoother_ctx = ctx;
  return smoother;
}

/* ------------------------------------------------------- preconditioner */
#include <d4est_solver_multigrid_bottom_solver_cg.h>
#include <d4est_solver_multigrid_bottom_solver_cheby.h>

static d4est_b200_pc_multigrid_data_t *d4est_b200_pc_recording = NULL;

static void
d4est_b200_pc_record_level (p4est_t * p4est, int level, double *r, int local_nodes)
{
  d4est_b200_pc_multigrid_data_t *pcd = d4est_b200_pc_recording;
  d4est_solver_multigrid_t *mg_data = (d4est_solver_multigrid_t *) p4est->user_pointer;
  d4est_solver_multigrid_element_data_updater_t *updater = mg_data->elem_data_updater;
  if (pcd->level_mesh[level] == NULL) {     /* down leg; the up leg revisits the same meshes */
    pcd->level_mesh[level] = d4est_b200_get_mesh (pcd->ctx, p4est, updater->current_d4est_ghost,
                                                  updater->current_d4est_factors);
    pcd->level_elements[level] = p4est->local_num_quadrants;
    pcd->level_serial[level] = 0;
    for (int i = 0; i < D4EST_B200_CACHE; i++)
      if (pcd->ctx->cache[i].mesh == pcd->level_mesh[level]) pcd->level_serial[level] = pcd->ctx->cache[i].serial;
  }
  /* zero residual -> zero coarse right-hand sides, zero corrections: u is left as it was */
  memset (r, 0, sizeof (double) * (size_t) local_nodes);
}

static void
d4est_b200_pc_record_smooth (p4est_t * p4est, d4est_elliptic_data_t * vecs,
                             d4est_elliptic_eqns_t * fcns, double *r, int level)
{
  d4est_b200_pc_record_level (p4est, level, r, vecs->local_nodes);
}

static void
d4est_b200_pc_record_bottom (p4est_t * p4est, d4est_elliptic_data_t * vecs,
                             d4est_elliptic_eqns_t * fcns, double *r)
{
  d4est_b200_pc_record_level (p4est, 0, r, vecs->local_nodes);
}

static void
d4est_b200_pc_no_update (p4est_t * p4est, int level, d4est_elliptic_data_t * vecs)
{
}

static void
d4est_b200_pc_drop_hierarchy (d4est_b200_pc_multigrid_data_t * pcd)
{
  if (pcd->mg) d4b_mg_destroy (pcd->mg);
  pcd->mg = NULL;
  for (int l = 0; l < D4EST_B200_MAX_LEVELS; l++) pcd->level_mesh[l] = NULL;
}

/* one recording V-cycle of the reference, then d4b_mg_create from what it left behind */
static void
d4est_b200_pc_build_hierarchy (d4est_b200_pc_multigrid_data_t * pcd, p4est_t * p4est,
                               d4est_elliptic_data_t * vecs, d4est_elliptic_eqns_t * fcns)
{
  d4est_solver_multigrid_t *mg_data = pcd->mg_data;
  const int nl = mg_data->num_of_levels, top = nl - 1;
  if (nl < 2 || nl > D4EST_B200_MAX_LEVELS) D4EST_ABORT ("d4est_b200_pc_multigrid: unsupported number of levels");
  if (strcmp (mg_data->smoother_name, "mg_smoother_cheby")) D4EST_ABORT ("d4est_b200_pc_multigrid needs mg_smoother_cheby");
  d4est_b200_pc_drop_hierarchy (pcd);

  d4est_solver_multigrid_smoother_fcn_t smooth0 = mg_data->smoother->smooth;
  d4est_solver_multigrid_update_callback_fcn_t supdate0 = mg_data->smoother->update;
  d4est_solver_multigrid_bottom_solver_fcn_t solve0 = mg_data->bottom_solver->solve;
  d4est_solver_multigrid_update_callback_fcn_t bupdate0 = mg_data->bottom_solver->update;
  const int imax0 = mg_data->vcycle_imax;
  mg_data->smoother->smooth = d4est_b200_pc_record_smooth;
  mg_data->smoother->update = d4est_b200_pc_no_update;
  mg_data->bottom_solver->solve = d4est_b200_pc_record_bottom;
  if (bupdate0) mg_data->bottom_solver->update = d4est_b200_pc_no_update;
  mg_data->vcycle_imax = 1;
  d4est_b200_pc_recording = pcd;
  d4est_solver_multigrid_solve (p4est, vecs, fcns, mg_data);
  d4est_b200_pc_recording = NULL;
  mg_data->smoother->smooth = smooth0;
  mg_data->smoother->update = supdate0;
  mg_data->bottom_solver->solve = solve0;
  mg_data->bottom_solver->update = bupdate0;
  mg_data->vcycle_imax = imax0;

  /* families: the records of level l's coarsening follow those of level l+1's
   * (mg_data->stride bookkeeping, d4est_solver_multigrid.c:975-1046) */
  int32_t *first[D4EST_B200_MAX_LEVELS], *count[D4EST_B200_MAX_LEVELS];
  const d4est_solver_multigrid_refine_data_t *cgr = mg_data->coarse_grid_refinement;
  int rec = 0;
  for (int l = top - 1; l >= 0; l--) {
    if (pcd->level_mesh[l] == NULL || pcd->level_mesh[l + 1] == NULL)
      D4EST_ABORT ("d4est_b200_pc_multigrid: the recording V-cycle did not visit every level");
    const int n_fine = pcd->level_elements[l + 1], n_coarse = pcd->level_elements[l];
    first[l] = (int32_t *) malloc (sizeof (int32_t) * (n_coarse + 8));
    count[l] = (int32_t *) malloc (sizeof (int32_t) * (n_coarse + 8));
    int consumed = 0, ce = 0;
    while (consumed < n_fine && ce < n_coarse) {
      const int h = cgr[rec++].hrefine;
      if (h == 1) { first[l][ce] = consumed; count[l][ce] = (P4EST_CHILDREN); ce++; consumed += (P4EST_CHILDREN); }
      else if (h == 0) { first[l][ce] = consumed; count[l][ce] = 1; ce++; consumed++; }
      else                                  /* coarsened, then re-split by the 2:1 balance: eight 1:1 copies */
        for (int i = 0; i < (P4EST_CHILDREN); i++) { first[l][ce] = consumed; count[l][ce] = 1; ce++; consumed++; }
    }
    if (consumed != n_fine || ce != n_coarse) D4EST_ABORT ("d4est_b200_pc_multigrid: refinement records do not match the level sizes");
  }

  d4b_mg_desc_t D;
  memset (&D, 0, sizeof (D));
  D.n_levels = nl;
  D.meshes = pcd->level_mesh;
  D.child_first = (const int32_t * const *) first;
  D.child_count = (const int32_t * const *) count;
  D.vcycle_imax = mg_data->vcycle_imax;
  D.vcycle_rtol = mg_data->vcycle_rtol;
  D.vcycle_atol = mg_data->vcycle_atol;
  const d4est_solver_multigrid_smoother_cheby_t *ch = mg_data->smoother->user;
  D.smoother.cheby_imax = ch->cheby_imax;
  D.smoother.cheby_eigs_cg_imax = ch->cheby_eigs_cg_imax;
  D.smoother.cheby_eigs_lmax_lmin_ratio = ch->cheby_eigs_lmax_lmin_ratio;
  D.smoother.cheby_eigs_max_multiplier = ch->cheby_eigs_max_multiplier;
  D.smoother.cheby_use_new_cg_eigs = ch->cheby_use_new_cg_eigs;
  D.smoother.cheby_use_zero_guess_for_eigs = ch->cheby_use_zero_guess_for_eigs;
  D.smoother.cheby_eigs_reuse_fromdownvcycle = ch->cheby_eigs_reuse_fromdownvcycle;
  D.smoother.cheby_eigs_reuse_fromlastvcycle = ch->cheby_eigs_reuse_fromlastvcycle;
  if (!strcmp (mg_data->bottom_solver_name, "mg_bottom_solver_cg")) {
    const d4est_solver_multigrid_bottom_solver_cg_t *b = mg_data->bottom_solver->user;
    D.bottom_solver = 0;
    D.bottom_iter = b->bottom_imax;
    D.bottom_rtol = b->bottom_rtol;
    D.bottom_atol = b->bottom_atol;
  }
  else if (!strcmp (mg_data->bottom_solver_name, "mg_bottom_solver_cheby")) {
    const d4est_solver_multigrid_bottom_solver_cheby_t *b = mg_data->bottom_solver->user;
    D.bottom_solver = 1;
    D.bottom_cheby.cheby_imax = b->cheby_imax;
    D.bottom_cheby.cheby_eigs_cg_imax = b->cheby_eigs_cg_imax;
    D.bottom_cheby.cheby_eigs_lmax_lmin_ratio = b->cheby_eigs_lmax_lmin_ratio;
    D.bottom_cheby.cheby_eigs_max_multiplier = b->cheby_eigs_max_multiplier;
    D.bottom_cheby.cheby_use_new_cg_eigs = b->cheby_use_new_cg_eigs;
  }
  else
    D4EST_ABORT ("d4est_b200_pc_multigrid needs mg_bottom_solver_cg or mg_bottom_solver_cheby");
  pcd->mg = d4b_mg_create (&D);
  if (pcd->ctx->has_nonlinear) d4b_mg_set_coarse_mass (pcd->mg, D4B_COARSE_MASS_REFERENCE);
  for (int l = 0; l < top; l++) { free (first[l]); free (count[l]); }
  pcd->n_levels = nl;
  pcd->top_elements = p4est->local_num_quadrants;
  pcd->top_nodes = vecs->local_nodes;
  pcd->builds++;
}

/* is the recorded hierarchy still the one of the reference's current mesh?  Every level's device mesh must
 * still be the cached upload it was built from, and the top entry must still describe p4est (same test as
 * d4est_b200_get_mesh: an AMR step can keep all counts) */
static int
d4est_b200_pc_hierarchy_current (d4est_b200_pc_multigrid_data_t * pcd, p4est_t * p4est, int local_nodes)
{
  if (pcd->mg == NULL || pcd->top_elements != p4est->local_num_quadrants || pcd->top_nodes != local_nodes) return 0;
  d4est_b200_cached_mesh_t *top = NULL;
  for (int l = 0; l < pcd->n_levels; l++) {
    d4est_b200_cached_mesh_t *hit = NULL;
    for (int i = 0; i < D4EST_B200_CACHE && !hit; i++)
      if (pcd->ctx->cache[i].mesh == pcd->level_mesh[l] && pcd->ctx->cache[i].serial == pcd->level_serial[l])
        hit = &pcd->ctx->cache[i];
    if (!hit) return 0;
    if (l == pcd->n_levels - 1) top = hit;
  }
  return top->mesh_sig == mesh_signature (p4est);
}

static void
d4est_b200_pc_multigrid_apply (d4est_krylov_pc_t * kpc, double *xp, double *yp)
{
  d4est_b200_pc_multigrid_data_t *pcd = kpc->pc_data;
  krylov_ctx_t *kct = kpc->pc_ctx;
  const int local_nodes = kct->vecs->local_nodes;
  /* zero initial guess, d4est_krylov_pc_multigrid.c:52 */
  memset (yp, 0, sizeof (double) * (size_t) local_nodes);
  if (!d4est_b200_pc_hierarchy_current (pcd, kct->p4est, local_nodes)) {
    double *Au = P4EST_ALLOC (double, local_nodes);
    d4est_elliptic_data_t mg_prob_vecs;
    d4est_elliptic_data_copy_ptrs (kct->vecs, &mg_prob_vecs);
    mg_prob_vecs.u = yp;
    mg_prob_vecs.rhs = xp;
    mg_prob_vecs.Au = Au;
    d4est_b200_pc_build_hierarchy (pcd, kct->p4est, &mg_prob_vecs, kct->fcns);
    P4EST_FREE (Au);
  }
  if (pcd->ctx->has_nonlinear)
    d4est_b200_ensure_linearised (pcd->ctx, pcd->level_mesh[pcd->n_levels - 1], kct->vecs->u0, local_nodes);
  pcd->vcycles += d4b_mg_solve (pcd->mg, yp, xp, NULL);
  pcd->applies++;
}

static void
d4est_b200_pc_multigrid_setup (d4est_krylov_pc_t * kpc)
{
  /* like d4est_krylov_pc_multigrid_setup (d4est_krylov_pc_multigrid.c:8-26): only the user's
   * hook runs.  The reference's FCG calls this before EVERY pc_apply
   * (d4est_solver_fcg_improved.c:206-210); the device hierarchy depends on the meshes alone and
   * is rebuilt by pc_apply when the top mesh changed, so there is nothing to drop here. */
  d4est_b200_pc_multigrid_data_t *pcd = kpc->pc_data;
  if (pcd->user_setup_fcn != NULL) pcd->user_setup_fcn (kpc);
}

d4est_krylov_pc_t *
d4est_b200_pc_multigrid_create (d4est_solver_multigrid_t * mg_data, d4est_b200_ctx_t * ctx,
                                void (*user_setup_fcn) (d4est_krylov_pc_t *))
{
  d4est_krylov_pc_t *pc = P4EST_ALLOC (d4est_krylov_pc_t, 1);
  d4est_b200_pc_multigrid_data_t *pcd = P4EST_ALLOC_ZERO (d4est_b200_pc_multigrid_data_t, 1);
  pcd->mg_data = mg_data;
  pcd->ctx = ctx;
  pcd->user_setup_fcn = user_setup_fcn;
  pc->pc_apply = d4est_b200_pc_multigrid_apply;
  pc->pc_setup = d4est_b200_pc_multigrid_setup;
  pc->pc_data = pcd;
  pc->pc_ctx = NULL;
  return pc;
}

void
d4est_b200_pc_multigrid_destroy (d4est_krylov_pc_t * kpc)
{
  d4est_b200_pc_drop_hierarchy ((d4est_b200_pc_multigrid_data_t *) kpc->pc_data);
  P4EST_FREE (kpc->pc_data);
  P4EST_FREE (kpc);
}
#endif
