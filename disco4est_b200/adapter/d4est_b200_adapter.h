/* d4est_b200_adapter.h -- reference-side binding of libd4est_b200.
 *
 * This file is compiled AGAINST THE REFERENCE'S HEADERS (p4est, d4est_*.h);
 * it is what a maintainer of the reference adds to their tree.  It flattens
 * p4est_t + d4est_ghost_t + d4est_mesh_data_t + flux parameters into the
 * plain-C d4b_mesh_desc_t of include/d4est_b200.h and provides a function
 * with the exact d4est_apply_operator_fcn_t signature
 * (src/EllipticSystem/d4est_elliptic_eqns.h:12-24).
 */
#ifndef D4EST_B200_ADAPTER_H
#define D4EST_B200_ADAPTER_H

#include <pXest.h>
#include <d4est_ghost.h>
#include <d4est_ghost_data.h>
#include <d4est_mesh.h>
#include <d4est_elliptic_data.h>
#include <d4est_elliptic_eqns.h>
#include <d4est_laplacian_flux.h>
#include <d4est_laplacian_flux_sipg.h>
#include <d4est_solver_multigrid.h>
#include <d4est_krylov_pc.h>
#include <d4est_quadrature.h>
#include <d4est_xyz_functions.h>
#include <d4est_geometry.h>
#include <d4est_geometry_brick.h>
#include <d4est_geometry_cubed_sphere.h>
#include <d4est_estimator_bi.h>
#include "d4est_b200.h"

/* Owns the arrays a d4b_mesh_desc_t points to (those that are not aliases of
 * reference arrays). */
typedef struct
{
  d4b_mesh_desc_t desc;
  /* owned */
  int32_t *deg, *deg_quad;
  int64_t *nodal_stride, *quad_stride;
  int32_t *ghost_deg, *ghost_deg_quad;
  int64_t *ghost_nodal_stride;
  int32_t *side_faces_m, *side_faces_p, *side_f_m, *side_f_p;
  int32_t *side_e_m, *side_e_p, *side_perm, *side_sub_order;
  int64_t *side_mortar_stride;
  double  *side_penalty_mult;
  double  *dirichlet_g;
  double  *robin_coeff, *robin_rhs;   /* BC_ROBIN: the two callbacks at the boundary quadrature points */
  int32_t *mirror_elem, *mirror_proc_offsets, *mirror_proc_mirrors, *ghost_proc_offsets;
  int side_capacity;
} d4est_b200_flat_t;

/* One uploaded mesh.  The reference re-runs d4est_mesh_update after AMR and keeps
 * one d4est_mesh_data_t per multigrid level alive for the whole solve
 * (d4est_solver_multigrid_element_data_updater.c:89-205); a small cache makes every
 * level a one-time upload.  The reference also REUSES one d4est_mesh_data_t across
 * d4est_mesh_update calls, and an AMR step can keep every count (refine one element and
 * coarsen one family; a +1/-1 p-adapt), so a cached mesh is identified by
 *   (factors pointer, element / node / quadrature-node / mortar-node counts)   O(1)
 *   + a 64-bit hash of (tree, x, y, z, level, deg, deg_quad) of every local quadrant: one walk over
 *     the forest per look-up (about 1 ms per 250 k elements, against a PCIe-bound host-vector apply)
 *   + how the mesh was built: with / without evaluated Dirichlet data, copied / generated factors.
 * d4est_b200_ctx_invalidate drops every cached mesh (for callers that overwrite factor arrays
 * in place without changing the forest or the degrees).                        */
#define D4EST_B200_CACHE 32
typedef struct
{
  d4b_mesh_t *mesh;
  d4est_mesh_data_t *factors;
  int local_elements, local_nodes, local_nodes_quad, local_mortar_nodes_quad;
  uint64_t mesh_sig;
  int has_dirichlet_data;                  /* built after bc_ops was set     */
  void *device_geom;                       /* built with generated factors   */
  unsigned long stamp;                     /* last use, for eviction         */
  unsigned long serial;                    /* unique per upload: holders of a d4b_mesh_t* (the pc hierarchy)
                                              check it to notice that their mesh was replaced */
} d4est_b200_cached_mesh_t;

/* user context stored in d4est_elliptic_eqns_t.user */
typedef struct d4est_b200_ctx
{
  d4est_laplacian_flux_data_t *flux_data;  /* as the reference callbacks use */
  d4est_b200_cached_mesh_t cache[D4EST_B200_CACHE];
  unsigned long clock, serial;
  long uploads, applies;                   /* statistics                     */
  /* device-generated geometric factors (d4est_b200_ctx_use_device_geometry) */
  d4est_geometry_t *device_geom;
  int device_face_h_type;
  /* set by d4est_b200_build_rhs_with_strong_bc: evaluate the Dirichlet function when a mesh is built */
  d4est_operators_t *bc_ops;
  d4est_geometry_t *bc_geom;
  d4est_quadrature_t *bc_quad;
  /* nonlinear mass-type term (d4est_b200_set_nonlinear_term): attached to every mesh this context uploads */
  int has_nonlinear;
  d4b_nonlinear_desc_t nonlinear;
  /* the context whose operator build_residual applies (flux_data_for_res of the problem files: same faces,
   * boundary data NOT zeroed); NULL = this context */
  struct d4est_b200_ctx *residual_ctx;
  /* linearisation point last uploaded for the top mesh (d4est_b200_apply_lhs with prob_vecs->u0) */
  const double *lin_u0;
  double lin_sig;
  int lin_nodes;
  int newton_nodes;                        /* local_nodes of the mesh build_residual ran on last (0 = unknown) */
  d4b_mesh_t *lin_mesh;
  long linearisations;                     /* statistics: u0 uploads          */
} d4est_b200_ctx_t;

/* Dirichlet data of flux_data->bc_data at the boundary-face quadrature points, evaluated the way
 * d4est_laplacian_flux_sipg_dirichlet_aux does (src/dGMath/d4est_laplacian_flux_sipg.c:60-130):
 * EVAL_BNDRY_FCN_ON_LOBATTO -- on the face LGL nodes (xyz_m_mortar_lobatto), then interpolated to
 * the Gauss points; EVAL_BNDRY_FCN_ON_QUAD -- directly at xyz_m_mortar_quad.  Fills
 * fl->dirichlet_g and points fl->desc.dirichlet_g_mortar_quad at it.                          */
void d4est_b200_flat_eval_dirichlet (d4est_b200_flat_t * fl, d4est_operators_t * d4est_ops,
                                     d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                                     d4est_mesh_data_t * d4est_factors,
                                     d4est_laplacian_flux_data_t * flux_data);

/* drop-in for d4est_laplacian_build_rhs_with_strong_bc (src/dGMath/d4est_laplacian.c:15-140), same
 * arguments: rhs = M f - A(0) on the device.  The mesh is built (and cached in ctx_for_rhs, a context
 * created from flux_fcn_data_for_build_rhs) with that flux data's Dirichlet function evaluated at the
 * boundary quadrature points; f is evaluated on the host through problem_rhs_fcn.            */
void d4est_b200_build_rhs_with_strong_bc (d4est_b200_ctx_t * ctx_for_rhs, p4est_t * p4est, d4est_ghost_t * ghost,
                                          d4est_ghost_data_t * ghost_data, d4est_operators_t * d4est_ops,
                                          d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                                          d4est_mesh_data_t * d4est_factors, d4est_elliptic_data_t * prob_vecs,
                                          d4est_laplacian_flux_data_t * flux_fcn_data_for_build_rhs,
                                          double *rhs, d4est_xyz_fcn_t problem_rhs_fcn,
                                          d4est_mesh_init_field_option_t init_option, void *fcn_ctx,
                                          int which_field);

/* drop-in for d4est_estimator_bi_compute (src/Estimators/d4est_estimator_bi.c:349-602), same arguments
 * behind the context: runs the problem's build_residual callback (fcns->build_residual, which may
 * itself be device-backed), then eta^2 of every local element on the device (d4b_estimator_bi).
 * The three prefactor callbacks of penalty_data are identified by probing (they are static inline
 * in d4est_estimator_bi.h); an unknown combination aborts.  Returns a P4EST_ALLOC'd array like
 * the reference; estimator_vtk (may be NULL) receives the four parts, estimator_vtk_per_face is
 * not filled (zeros).                                                                          */
double *d4est_b200_estimator_bi_compute (d4est_b200_ctx_t * ctx, p4est_t * p4est,
                                         d4est_elliptic_data_t * d4est_elliptic_data, d4est_elliptic_eqns_t * fcns,
                                         d4est_estimator_bi_penalty_data_t penalty_data, d4est_xyz_fcn_t u_bndry_fcn,
                                         void *bndry_ctx, d4est_ghost_t * d4est_ghost,
                                         d4est_ghost_data_t * d4est_ghost_data, d4est_operators_t * d4est_ops,
                                         d4est_geometry_t * d4est_geom_for_residual,
                                         d4est_mesh_data_t * d4est_factors_for_residual,
                                         d4est_geometry_t * d4est_geom_for_integrals,
                                         d4est_mesh_data_t * d4est_factors_for_integrals,
                                         d4est_quadrature_t * d4est_quad, int which_field, double *estimator_vtk,
                                         double *estimator_vtk_per_face);

/* Opt in to geometric factors generated on the device: from now on a mesh that is uploaded
 * for this context takes J, rst_xyz and all mortar quantities from d4b_geom_compute (the map of
 * d4est_geom evaluated at every quadrature point in a CUDA thread) instead of copying the
 * arrays d4est_mesh_data_compute / d4est_mesh_compute_mortar_quadrature_quantities filled
 * (src/Mesh/d4est_mesh.c:2543-2755, :481-1140); only the forest positions (tree, q, dq) and
 * degrees of the local and ghost elements cross the bus.  Supported maps: brick, 13-tree cubed
 * sphere with DX_compute_method = analytic, JAC_compute_method = numerical; aborts otherwise.
 * face_h_type: the mesh's d4est_mesh_face_h_t (initial_extents->face_h_type).              */
void d4est_b200_ctx_use_device_geometry (d4est_b200_ctx_t * ctx, d4est_geometry_t * d4est_geom,
                                         d4est_mesh_face_h_t face_h_type);
/* d4b_geom_desc_t of the reference's current forest; free with d4est_b200_geom_desc_free */
void d4est_b200_geom_desc_fill (d4b_geom_desc_t * g, p4est_t * p4est, d4est_ghost_t * ghost,
                                d4est_geometry_t * d4est_geom, d4est_mesh_face_h_t face_h_type);
void d4est_b200_geom_desc_free (d4b_geom_desc_t * g);

/* the device mesh for the reference's current mesh (uploads it on first use) */
d4b_mesh_t *d4est_b200_get_mesh(d4est_b200_ctx_t *ctx, p4est_t *p4est, d4est_ghost_t *ghost,
                                d4est_mesh_data_t *d4est_factors);

d4est_b200_flat_t *d4est_b200_flatten(p4est_t *p4est, d4est_ghost_t *d4est_ghost,
                                      d4est_mesh_data_t *d4est_factors,
                                      d4est_laplacian_flux_data_t *flux_data);
void d4est_b200_flat_destroy(d4est_b200_flat_t *flat);

/* drop-in for d4est_elliptic_eqns_t.apply_lhs */
void d4est_b200_apply_lhs(p4est_t *p4est, d4est_ghost_t *ghost,
                          d4est_ghost_data_t *ghost_data,
                          d4est_elliptic_data_t *prob_vecs,
                          d4est_operators_t *d4est_ops,
                          d4est_geometry_t *d4est_geom,
                          d4est_quadrature_t *d4est_quad,
                          d4est_mesh_data_t *d4est_factors, void *user);

/* drop-in for d4est_elliptic_eqns_t.build_residual (src/EllipticSystem/d4est_elliptic_eqns.h:26-37), same nine
 * arguments, user = the d4est_b200_ctx_t of apply_lhs:  prob_vecs->Au = A(u) + N(u) - rhs  on the device, with
 * A the operator of ctx->residual_ctx (or ctx) INCLUDING its boundary data, N the nonlinear term of the context
 * if one is set, and rhs = prob_vecs->rhs unless ctx->has_nonlinear (the nonlinear problem files put no rhs
 * into their residual, multi_punctures_build_residual; the linear ones do, poisson_build_residual,
 * src/Problems/Poisson/poisson_sinx_fcns.h:133-149).                                                    */
void d4est_b200_build_residual(p4est_t *p4est, d4est_ghost_t *ghost, d4est_ghost_data_t *ghost_data,
                               d4est_elliptic_data_t *prob_vecs, d4est_operators_t *d4est_ops,
                               d4est_geometry_t *d4est_geom, d4est_quadrature_t *d4est_quad,
                               d4est_mesh_data_t *d4est_factors, void *user);

/* Nonlinear mass-type term f(x, u) = a(x) (b(x) + u)^k as a DEVICE functor (include/d4est_b200.h,
 * d4b_nonlinear_desc_t): `desc` carries the kind and its parameters (puncture list / star); xyz_quad and jac_quad
 * are filled in from the reference's d4est_mesh_data_t whenever a mesh is uploaded.  From then on
 *   d4est_b200_build_residual adds V^T w J f(x, V u)                  (d4est_quadrature_apply_fofufofvlj),
 *   d4est_b200_apply_lhs applies the Jacobian about prob_vecs->u0      (..._fofufofvlilj); u0 is uploaded
 *     once per linearisation point -- recognised by its address and a strided checksum -- and f, f_u are
 *     evaluated on the device,
 *   d4est_b200_pc_multigrid_create's hierarchy applies the reference's multigrid matrix operator on its coarse
 *     levels (d4b_mg_set_coarse_mass, as the reference executes it).
 * residual_ctx: the context of the problem's flux_data_for_res (may be NULL = same operator).                */
void d4est_b200_set_nonlinear_term(d4est_b200_ctx_t *ctx, const d4b_nonlinear_desc_t *desc,
                                   d4est_b200_ctx_t *residual_ctx);

/* Collective context from the reference side: one NCCL communicator over the ranks of
 * p4est->mpicomm.  Rank 0 draws the id (d4b_comm_get_unique_id), sc_MPI_Bcast carries its 128 bytes, every rank
 * selects `device` (e.g. local rank) and calls d4b_comm_init; the neighbour lists of the halo exchange come with
 * every flattened mesh (mirror / ghost offsets of d4est_ghost_t, src/Mesh/d4est_ghost_data.c:171-199).  Call once
 * after p4est_new_ext; a no-op on one rank.  d4est_b200_comm_finalize before sc_MPI_Finalize.                  */
void d4est_b200_comm_init_from_p4est(p4est_t *p4est, int device);
void d4est_b200_comm_finalize(void);

/* The solver vectors of the reference are P4EST_ALLOC'd (pageable): page-lock prob_vecs->u / Au / rhs (those that are
 * set, local_nodes doubles each) once after they are allocated, so that d4est_b200_apply_lhs moves them at link speed
 * and can overlap the two directions (d4b_host_pin).  d4est_b200_unpin_vectors before they are freed.  Optional:
 * without it everything works on pageable memory, more slowly.                                                   */
void d4est_b200_pin_vectors(d4est_elliptic_data_t *prob_vecs);
void d4est_b200_unpin_vectors(d4est_elliptic_data_t *prob_vecs);

d4est_b200_ctx_t *d4est_b200_ctx_new(d4est_laplacian_flux_data_t *flux_data);
void d4est_b200_ctx_destroy(d4est_b200_ctx_t *ctx);
/* forget (and free) every cached device mesh: the next call re-flattens and uploads */
void d4est_b200_ctx_invalidate(d4est_b200_ctx_t *ctx);

/* Second entry point: the multigrid smoother.  Same factory signature, INI section
 * ([mg_smoother_cheby]) and behaviour as d4est_solver_multigrid_smoother_cheby_init
 * (src/Solver/d4est_solver_multigrid_smoother_cheby.c:20-63, :264-376, :378-430) plus the
 * context: `smooth` runs cg_eigs and the Chebyshev recurrence on the device (vectors cross
 * PCIe once per smoothing instead of once per apply) and leaves vecs->u and r = rhs - A u
 * exactly like the reference.  The operator is the SIPG Laplacian of ctx->flux_data: only
 * valid where fcns->apply_lhs is d4est_laplacian[_with_opt]_apply_aij.  Install with
 *   mg_data->smoother = d4est_b200_smoother_cheby_init(p4est, mg_data->num_of_levels, ini, ctx);
 * (what d4est_solver_multigrid_set_smoother does for "mg_smoother_cheby",
 * d4est_solver_multigrid.c:429-474); destroy with the reference's
 * d4est_solver_multigrid_smoother_cheby_destroy.                                        */
d4est_solver_multigrid_smoother_t *d4est_b200_smoother_cheby_init(p4est_t *p4est, int num_of_levels,
                                                                  const char *input_file,
                                                                  d4est_b200_ctx_t *ctx);

/* Mass-type term of the nonlinear problems' Jacobians.  The reference adds,
 * element by element, d4est_quadrature_apply_fofufofvlilj (vec = u, u = u0, fofu_fcn = f,
 * QUAD_APPLY_MATRIX) to the Laplacian apply (e.g. constant_density_star_apply_jac,
 * src/Problems/ConstantDensityStar/constant_density_star_fcns.h:528-600).  This call evaluates
 * J f(x, V u0) at the quadrature points with the caller's own d4est_xyzu_fcn_t -- once per
 * Newton linearisation -- and fuses it into the device operator of the current mesh, so that
 * d4est_b200_apply_lhs, the smoother and the preconditioner then apply the whole Jacobian
 * in one pass.  f == NULL removes the term.                                             */
void d4est_b200_set_mass_term (d4est_b200_ctx_t * ctx, p4est_t * p4est, d4est_ghost_t * ghost,
                               d4est_operators_t * d4est_ops, d4est_geometry_t * d4est_geom,
                               d4est_quadrature_t * d4est_quad, d4est_mesh_data_t * d4est_factors,
                               double *u0, d4est_xyzu_fcn_t f, void *f_ctx);

/* Third entry point: the whole multigrid preconditioner, device-resident.  Same role and
 * call convention as d4est_krylov_pc_multigrid_create (src/Solver/d4est_krylov_pc_multigrid.c:82-100):
 * the returned d4est_krylov_pc_t is handed to the Krylov solver, which calls
 * pc_apply(kpc, x, y) with host arrays (d4est_krylov_pc.h:69-89).  pc_apply does what
 * d4est_krylov_pc_multigrid_apply does -- y = 0, then d4est_solver_multigrid_solve on
 * A y = x with mg_data's vcycle_imax / rtol / atol, [mg_smoother_cheby] and bottom-solver
 * settings -- but as H2D(x), V-cycles on the device (d4b_mg_solve), D2H(y).
 *
 * The level hierarchy is taken from the reference itself: on the first apply (and again
 * when the top mesh changed) ONE V-cycle of
 * the unmodified d4est_solver_multigrid_solve is run with a recording smoother / bottom
 * solver that do no arithmetic.  It walks the reference's own coarsen / balance / mesh-update
 * sequence, so every level's d4est_mesh_data_t is uploaded through d4est_b200_get_mesh and
 * mg_data->coarse_grid_refinement afterwards holds, per level, which coarse element came from
 * which fine elements (hrefine 0 / 1 / 2, d4est_solver_multigrid.h:12-18).
 * Valid where fcns->apply_lhs is the SIPG Laplacian of ctx->flux_data and the smoother is
 * mg_smoother_cheby with mg_bottom_solver_cg or mg_bottom_solver_cheby.                  */
#define D4EST_B200_MAX_LEVELS 32
typedef struct
{
  d4est_solver_multigrid_t *mg_data;
  d4est_b200_ctx_t *ctx;
  void (*user_setup_fcn) (d4est_krylov_pc_t *);
  d4b_mg_t *mg;                            /* NULL until the first apply     */
  int n_levels;
  d4b_mesh_t *level_mesh[D4EST_B200_MAX_LEVELS];
  unsigned long level_serial[D4EST_B200_MAX_LEVELS];
  int level_elements[D4EST_B200_MAX_LEVELS];
  int top_elements, top_nodes;
  long builds, applies, vcycles;           /* statistics                     */
} d4est_b200_pc_multigrid_data_t;

d4est_krylov_pc_t *d4est_b200_pc_multigrid_create (d4est_solver_multigrid_t * mg_data,
                                                   d4est_b200_ctx_t * ctx,
                                                   void (*user_setup_fcn) (d4est_krylov_pc_t *));
void d4est_b200_pc_multigrid_destroy (d4est_krylov_pc_t * kpc);

#endif
