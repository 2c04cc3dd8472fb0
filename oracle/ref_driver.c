/* ref_driver.c -- drives the UNMODIFIED reference (compiled serially into
 * oracle/_ref/) through its own public API and dumps inputs + outputs.
 *
 * TEST INFRASTRUCTURE - not part of the product.  Only tests/, smoke() and
 * bench.py's cpu_baseline / --impl reference legs may execute this binary.
 *
 * Modelled on src/Tests/Unit/d4est_test_laplacian_speedup.c:155-540: build the
 * forest from an INI file, optional random hp-AMR steps, u = x^2+y^2+z^2 +
 * U(0,1) with srand(102321), apply the plain and the with_opt operator.
 *
 *   ref_driver <ini> <out.d4b|-> [--amr N] [--time REPS] [--cheby IMAX EIGS_IMAX] [--mg VCYCLES]
 *              [--bc zero|poly] [--uniform-deg P]
 */
#include <pXest.h>
#include <d4est_quadrature.h>
#include <d4est_element_data.h>
#include <d4est_quadrature_legendre.h>
#include <d4est_geometry.h>
#include <d4est_geometry_brick.h>
#include <d4est_geometry_cubed_sphere.h>
#include <d4est_mesh.h>
#include <d4est_linalg.h>
#include <d4est_mortars.h>
#include <d4est_amr.h>
#include <d4est_laplacian_with_opt.h>
#include <d4est_laplacian.h>
#include <d4est_solver_cg_eigs.h>
#include <d4est_solver_multigrid_smoother_cheby.h>
#include <d4est_solver_multigrid.h>
#include <d4est_krylov_pc_multigrid.h>
#include <d4est_solver_fcg_improved.h>
#include <d4est_solver_cg.h>
#include <d4est_estimator_bi.h>
#include <d4est_solver_newton.h>
#include <d4est_solver_multigrid_matrix_operator.h>
#include <d4est_util.h>
/* the two-puncture problem callbacks of the reference, used as they are
 * (src/Problems/TwoPunctures/multi_punctures_fcns_with_opt.h; found through -I src/Problems/TwoPunctures) */
#include <multi_punctures_fcns_with_opt.h>
#include <zlog.h>
#include <time.h>
#include <sys/time.h>
#include <unistd.h>
#include "d4b_dump.h"
#include "d4est_b200_adapter.h"

static double
wall (void)
{
  struct timeval tv;
  gettimeofday (&tv, NULL);
  return tv.tv_sec + 1e-6 * tv.tv_usec;
}

static double
drv_zero_fcn (double x, double y, double z, void *user)
{
  return 0.;
}

static double
drv_poly_fcn (double x, double y, double z, void *user)
{
  return x * x + y * y + z * z;
}

/* Robin boundary of the fixtures: coeff = 1 / |x| like multi_puncture_robin_coeff_sphere_fcn
 * (src/Problems/TwoPunctures/multi_puncture_fcns.h:772-786, shifted so that it is finite on the
 * brick), and a non-zero right-hand side so that both callbacks are exercised */
static double
drv_robin_coeff (double x, double y, double z, void *user, d4est_laplacian_flux_boundary_data_t * bd, int node)
{
  (void) user; (void) bd; (void) node;
  return 1. / sqrt (x * x + y * y + z * z + .25);
}

static double
drv_robin_rhs (double x, double y, double z, void *user, d4est_laplacian_flux_boundary_data_t * bd, int node)
{
  (void) user; (void) bd; (void) node;
  return .3 * x - .2 * y * z;
}

/* residual callback for the estimator fixture: Au <- rhs - A u, like poisson_build_residual */
static void
drv_build_residual (p4est_t * p4est, d4est_ghost_t * ghost, d4est_ghost_data_t * ghost_data,
                    d4est_elliptic_data_t * prob_vecs, d4est_operators_t * d4est_ops,
                    d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                    d4est_mesh_data_t * d4est_factors, void *user)
{
  d4est_laplacian_apply_aij (p4est, ghost, ghost_data, prob_vecs, (d4est_laplacian_flux_data_t *) user, d4est_ops,
                             d4est_geom, d4est_quad, d4est_factors, 0);
  d4est_linalg_vec_xpby (prob_vecs->rhs, -1., prob_vecs->Au, prob_vecs->local_nodes);
}

/* right-hand-side density used for the build_rhs fixtures */
static double
drv_rhs_fcn (double x, double y, double z, void *user)
{
  (void) user;
  return -6. + sin (x) * cos (2. * y) + z * z * x;
}

static void
apply_plain (p4est_t * p4est, d4est_ghost_t * ghost, d4est_ghost_data_t * ghost_data,
             d4est_elliptic_data_t * prob_vecs, d4est_operators_t * d4est_ops,
             d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
             d4est_mesh_data_t * d4est_factors, void *user)
{
  d4est_laplacian_apply_aij (p4est, ghost, ghost_data, prob_vecs,
                             (d4est_laplacian_flux_data_t *) user, d4est_ops,
                             d4est_geom, d4est_quad, d4est_factors, 0);
}

static void
apply_with_opt (p4est_t * p4est, d4est_ghost_t * ghost, d4est_ghost_data_t * ghost_data,
                d4est_elliptic_data_t * prob_vecs, d4est_operators_t * d4est_ops,
                d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                d4est_mesh_data_t * d4est_factors, void *user)
{
  d4est_laplacian_with_opt_apply_aij (p4est, ghost, ghost_data, prob_vecs,
                                      (d4est_laplacian_with_opt_flux_data_t *) user,
                                      d4est_ops, d4est_geom, d4est_quad, d4est_factors, 0);
}

/* count-preserving p-adapt for the stale-cache check: elements g_swap_a / g_swap_b exchange degrees */
static int g_swap_a = -1, g_swap_b = -1, g_swap_deg_a, g_swap_deg_b;
static void
swap_deg_cb (d4est_element_data_t * elem_data, void *user_ctx)
{
  d4est_mesh_initial_extents_t *input = user_ctx;
  if (elem_data->id == g_swap_a) elem_data->deg = g_swap_deg_b;
  if (elem_data->id == g_swap_b) elem_data->deg = g_swap_deg_a;
  elem_data->deg_quad = input->deg_quad_inc[elem_data->region] + elem_data->deg;
}

static int g_uniform_deg = -1;
static void
set_uniform_deg (d4est_element_data_t * elem_data, void *user_ctx)
{
  d4est_mesh_initial_extents_t *input = user_ctx;
  elem_data->deg = g_uniform_deg;
  elem_data->deg_quad = input->deg_quad_inc[elem_data->region] + elem_data->deg;
}

static void
dump_desc (FILE * f, const d4b_mesh_desc_t * D)
{
  d4b_dump_scalar_i64 (f, "n_local", D->n_local);
  d4b_dump_scalar_i64 (f, "n_ghost", D->n_ghost);
  d4b_dump_scalar_i64 (f, "n_sides", D->n_sides);
  d4b_dump_scalar_i64 (f, "local_nodes", D->local_nodes);
  d4b_dump_scalar_i64 (f, "local_nodes_quad", D->local_nodes_quad);
  d4b_dump_scalar_i64 (f, "local_mortar_nodes_quad", D->local_mortar_nodes_quad);
  d4b_dump_scalar_i64 (f, "ghost_nodes", D->ghost_nodes);
  d4b_dump_put1 (f, "deg", D4B_I32, D->n_local, D->deg);
  d4b_dump_put1 (f, "deg_quad", D4B_I32, D->n_local, D->deg_quad);
  d4b_dump_put1 (f, "nodal_stride", D4B_I64, D->n_local, D->nodal_stride);
  d4b_dump_put1 (f, "quad_stride", D4B_I64, D->n_local, D->quad_stride);
  d4b_dump_put1 (f, "J_quad", D4B_F64, D->local_nodes_quad, D->J_quad);
  d4b_dump_put1 (f, "rst_xyz_quad", D4B_F64, 9 * D->local_nodes_quad, D->rst_xyz_quad);
  uint64_t m = D->local_mortar_nodes_quad;
  d4b_dump_put1 (f, "sj_m_mortar_quad", D4B_F64, m, D->sj_m_mortar_quad);
  d4b_dump_put1 (f, "n_m_mortar_quad", D4B_F64, 3 * m, D->n_m_mortar_quad);
  d4b_dump_put1 (f, "drst_dxyz_m_mortar_quad", D4B_F64, 9 * m, D->drst_dxyz_m_mortar_quad);
  d4b_dump_put1 (f, "drst_dxyz_p_mortar_quad_porder", D4B_F64, 9 * m, D->drst_dxyz_p_mortar_quad_porder);
  d4b_dump_put1 (f, "hm_mortar_quad", D4B_F64, m, D->hm_mortar_quad);
  d4b_dump_put1 (f, "hp_mortar_quad", D4B_F64, m, D->hp_mortar_quad);
  if (D->dirichlet_g_mortar_quad)
    d4b_dump_put1 (f, "dirichlet_g_mortar_quad", D4B_F64, m, D->dirichlet_g_mortar_quad);
  if (D->robin_coeff_mortar_quad) {
    d4b_dump_put1 (f, "robin_coeff_mortar_quad", D4B_F64, m, D->robin_coeff_mortar_quad);
    d4b_dump_put1 (f, "robin_rhs_mortar_quad", D4B_F64, m, D->robin_rhs_mortar_quad);
  }
  uint64_t s = D->n_sides;
  d4b_dump_put1 (f, "side_faces_m", D4B_I32, s, D->side_faces_m);
  d4b_dump_put1 (f, "side_faces_p", D4B_I32, s, D->side_faces_p);
  d4b_dump_put1 (f, "side_f_m", D4B_I32, s, D->side_f_m);
  d4b_dump_put1 (f, "side_f_p", D4B_I32, s, D->side_f_p);
  d4b_dump_put1 (f, "side_e_m", D4B_I32, 4 * s, D->side_e_m);
  d4b_dump_put1 (f, "side_e_p", D4B_I32, 4 * s, D->side_e_p);
  d4b_dump_put1 (f, "side_perm", D4B_I32, s, D->side_perm);
  d4b_dump_put1 (f, "side_sub_order", D4B_I32, 4 * s, D->side_sub_order);
  d4b_dump_put1 (f, "side_mortar_stride", D4B_I64, s, D->side_mortar_stride);
  d4b_dump_put1 (f, "side_penalty_mult", D4B_F64, s, D->side_penalty_mult);
  d4b_dump_scalar_f64 (f, "sipg_penalty_prefactor", D->sipg_penalty_prefactor);
  d4b_dump_scalar_i64 (f, "sipg_penalty_fcn", D->sipg_penalty_fcn);
  d4b_dump_scalar_i64 (f, "bc_type", D->bc_type);
}

/* reference 1-D tables, for pinning the product's own table builders */
static void
dump_tables (FILE * f, d4est_operators_t * ops, int max_deg)
{
  char name[64];
  for (int p = 1; p <= max_deg; p++) {
    int N = p + 1;
    snprintf (name, 64, "tab_lgl_nodes_%d", p);
    d4b_dump_put1 (f, name, D4B_F64, N, d4est_operators_fetch_lobatto_nodes_1d (ops, p));
    snprintf (name, 64, "tab_lgl_weights_%d", p);
    d4b_dump_put1 (f, name, D4B_F64, N, d4est_operators_fetch_lobatto_weights_1d (ops, p));
    snprintf (name, 64, "tab_gauss_nodes_%d", p);
    d4b_dump_put1 (f, name, D4B_F64, N, d4est_operators_fetch_gauss_nodes_1d (ops, p));
    snprintf (name, 64, "tab_gauss_weights_%d", p);
    d4b_dump_put1 (f, name, D4B_F64, N, d4est_operators_fetch_gauss_weights_1d (ops, p));
    snprintf (name, 64, "tab_dij_%d", p);
    {
      /* the 1-D derivative table is file-static in the reference; recover it
       * column by column through the public apply_dij in one dimension */
      double Dm[64], ej[8], col[8];
      for (int j = 0; j < N; j++) {
        for (int i = 0; i < N; i++) ej[i] = (i == j) ? 1. : 0.;
        d4est_operators_apply_dij (ops, ej, 1, p, 0, col);
        for (int i = 0; i < N; i++) Dm[i * N + j] = col[i];
      }
      d4b_dump_put1 (f, name, D4B_F64, N * N, Dm);
    }
    snprintf (name, 64, "tab_mij_%d", p);
    d4b_dump_put1 (f, name, D4B_F64, N * N, d4est_operators_fetch_mij_1d (ops, p));
    for (int q = p; q <= max_deg; q++) {
      snprintf (name, 64, "tab_lgl_to_gauss_%d_%d", p, q);
      d4b_dump_put1 (f, name, D4B_F64, (q + 1) * N,
                     d4est_operators_fetch_lobatto_to_gauss_interp_1d (ops, p, q));
      snprintf (name, 64, "tab_p_prolong_%d_%d", p, q);
      d4b_dump_put1 (f, name, D4B_F64, (q + 1) * N,
                     d4est_operators_fetch_p_prolong_1d (ops, p, q));
    }
  }
}

/* the reference's in-house Krylov solvers report progress through zlog only: count their
 * iterations and record the residual norms by wrapping the callbacks they are handed */
static d4est_krylov_pc_apply_fcn_t fcg_inner_pc_apply;
static int fcg_pc_calls, fcg_nodes;
static double fcg_r2_hist[512];

static double *fcg_first_pc_x, *fcg_first_pc_y;   /* when set: input / output of the first preconditioner call */

static void
fcg_counting_pc_apply (d4est_krylov_pc_t * kpc, double *x, double *y)
{
  if (fcg_pc_calls < 512) fcg_r2_hist[fcg_pc_calls] = d4est_linalg_vec_dot (x, x, fcg_nodes);
  if (fcg_pc_calls == 0 && fcg_first_pc_x) memcpy (fcg_first_pc_x, x, sizeof (double) * fcg_nodes);
  fcg_pc_calls++;
  fcg_inner_pc_apply (kpc, x, y);
  if (fcg_pc_calls == 1 && fcg_first_pc_y) memcpy (fcg_first_pc_y, y, sizeof (double) * fcg_nodes);
}

static d4est_apply_operator_fcn_t cg_inner_apply_lhs;
static int cg_apply_calls;

static void
cg_counting_apply_lhs (p4est_t * p4est, d4est_ghost_t * ghost, d4est_ghost_data_t * ghost_data,
                       d4est_elliptic_data_t * prob_vecs, d4est_operators_t * d4est_ops,
                       d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                       d4est_mesh_data_t * d4est_factors, void *user)
{
  cg_apply_calls++;
  cg_inner_apply_lhs (p4est, ghost, ghost_data, prob_vecs, d4est_ops, d4est_geom, d4est_quad, d4est_factors, user);
}

/* mass-type nonlinearities of the kind the shipped problems use (ConstantDensityStar:
 * residual -2 pi rho (1+u)^5, Jacobian -10 pi rho (1+u)^4 with rho a step in radius,
 * src/Problems/ConstantDensityStar/constant_density_star_fcns.h:64-139) */
static double
star_like_rho (double x, double y, double z)
{
  return (x * x + y * y + z * z < 1.44) ? 0.3 : 0.02;
}

static double
star_like_jac_fcn (double x, double y, double z, double u, void *user)
{
  return -10. * M_PI * star_like_rho (x, y, z) * (1. + u) * (1. + u) * (1. + u) * (1. + u);
}

static double
star_like_res_fcn (double x, double y, double z, double u, void *user)
{
  return -2. * M_PI * star_like_rho (x, y, z) * (1. + u) * (1. + u) * (1. + u) * (1. + u) * (1. + u);
}

/* ---- BASELINE configs[4]: the reference's in-house Newton-Krylov-multigrid stack on the two-puncture problem,
 * wired as src/Problems/TwoPunctures/multi_punctures_anares_inhouse.c:849-964 does, with recording wrappers
 * around the callbacks (the solvers report through zlog only). */
#define NEWTON_MAX_STEPS 64
static d4est_apply_operator_fcn_t nw_inner_build_residual;
static int nw_res_calls, nw_nodes;
static double nw_fnrm[NEWTON_MAX_STEPS + 1];
static int nw_pc_calls_at_res[NEWTON_MAX_STEPS + 1];

static void
nw_recording_build_residual (p4est_t * p4est, d4est_ghost_t * ghost, d4est_ghost_data_t * ghost_data,
                             d4est_elliptic_data_t * prob_vecs, d4est_operators_t * d4est_ops,
                             d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                             d4est_mesh_data_t * d4est_factors, void *user)
{
  nw_inner_build_residual (p4est, ghost, ghost_data, prob_vecs, d4est_ops, d4est_geom, d4est_quad, d4est_factors, user);
  if (nw_res_calls <= NEWTON_MAX_STEPS) {
    nw_fnrm[nw_res_calls] = sqrt (d4est_linalg_vec_dot (prob_vecs->Au, prob_vecs->Au, nw_nodes));
    nw_pc_calls_at_res[nw_res_calls] = fcg_pc_calls;
  }
  nw_res_calls++;
}

/* D4B_NW_TRACE=1: one line per operator application during the first preconditioner call (debugging aid) */
static d4est_apply_operator_fcn_t nw_inner_apply_lhs;
static void
nw_tracing_apply_lhs (p4est_t * p4est, d4est_ghost_t * ghost, d4est_ghost_data_t * ghost_data,
                      d4est_elliptic_data_t * prob_vecs, d4est_operators_t * d4est_ops,
                      d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
                      d4est_mesh_data_t * d4est_factors, void *user)
{
  nw_inner_apply_lhs (p4est, ghost, ghost_data, prob_vecs, d4est_ops, d4est_geom, d4est_quad, d4est_factors, user);
  if (getenv ("D4B_NW_TRACE") && fcg_pc_calls == atoi (getenv ("D4B_NW_TRACE")))
    fprintf (stderr, "APPLY pc_calls=%d n=%d |u|=%.15e |Au|=%.15e\n", fcg_pc_calls, prob_vecs->local_nodes,
             sqrt (d4est_linalg_vec_dot (prob_vecs->u, prob_vecs->u, prob_vecs->local_nodes)),
             sqrt (d4est_linalg_vec_dot (prob_vecs->Au, prob_vecs->Au, prob_vecs->local_nodes)));
}

static void
run_newton (p4est_t * p4est, d4est_ghost_t ** d4est_ghost, d4est_ghost_data_t ** d4est_ghost_data,
            d4est_operators_t * d4est_ops, d4est_geometry_t * d4est_geom, d4est_quadrature_t * d4est_quad,
            d4est_mesh_data_t * d4est_factors, d4est_mesh_initial_extents_t * initial_grid_input,
            const char *ini, const char *out, int local_nodes, int use_b200)
{
  d4est_field_type_t field_type = NODAL;
  multi_punctures_params_t *mp = multi_punctures_params_init (p4est, ini);
  problem_ctx_t ctx;
  memset (&ctx, 0, sizeof (ctx));
  ctx.multi_punctures_params = mp;

  /* Robin outer boundary exactly as the problem file sets it up (multi_punctures_anares_inhouse.c:640-668) */
  d4est_laplacian_with_opt_robin_bc_t bc_robin;
  bc_robin.robin_coeff = multi_punctures_robin_coeff_sphere_fcn;
  bc_robin.robin_rhs = multi_punctures_robin_bc_rhs_fcn;
  bc_robin.user = NULL;
  d4est_laplacian_with_opt_flux_data_t *flux_jac = d4est_laplacian_with_opt_flux_new (p4est, ini, BC_ROBIN, &bc_robin);
  d4est_laplacian_with_opt_flux_data_t *flux_res = d4est_laplacian_with_opt_flux_new (p4est, ini, BC_ROBIN, &bc_robin);
  ctx.flux_data_for_jac = flux_jac;
  ctx.flux_data_for_res = flux_res;
  /* the same boundary through the plain flux: only used to flatten the mesh for the fixture */
  d4est_laplacian_robin_bc_t bc_plain;
  bc_plain.robin_coeff = (d4est_laplacian_mortar_xyz_fcn_t) multi_punctures_robin_coeff_sphere_fcn;
  bc_plain.robin_rhs = (d4est_laplacian_mortar_xyz_fcn_t) multi_punctures_robin_bc_rhs_fcn;
  bc_plain.user = NULL;
  d4est_laplacian_flux_data_t *flux_plain = d4est_laplacian_flux_new (p4est, ini, BC_ROBIN, &bc_plain);

  d4est_elliptic_eqns_t prob_fcns;
  prob_fcns.build_residual = multi_punctures_build_residual;
  prob_fcns.apply_lhs = multi_punctures_apply_jac;
  prob_fcns.user = &ctx;

  double *u = P4EST_ALLOC_ZERO (double, local_nodes);
  double *Au = P4EST_ALLOC_ZERO (double, local_nodes);
  double *rhs = P4EST_ALLOC_ZERO (double, local_nodes);
  d4est_elliptic_data_t vecs;
  memset (&vecs, 0, sizeof (vecs));
  vecs.u = u; vecs.Au = Au; vecs.rhs = rhs; vecs.local_nodes = local_nodes;
  vecs.field_types = &field_type; vecs.num_of_fields = 1;

  /* operator-level samples before the solve: residual F(us) and Jacobian J(us) v for seeded us, v
   * (matrix operator off: the fine level evaluates the quadrature form, multi_punctures_apply_jac) */
  double *us = P4EST_ALLOC (double, local_nodes), *v = P4EST_ALLOC (double, local_nodes);
  double *Fus = P4EST_ALLOC_ZERO (double, local_nodes), *Jv = P4EST_ALLOC_ZERO (double, local_nodes);
  srand (102321);
  for (int i = 0; i < local_nodes; i++) us[i] = 0.2 * (double) rand () / (double) RAND_MAX - 0.05;
  for (int i = 0; i < local_nodes; i++) v[i] = (double) rand () / (double) RAND_MAX - 0.5;
  ctx.use_matrix_operator = 0;
  {
    d4est_elliptic_data_t sv = vecs;
    sv.u = us; sv.Au = Fus;
    d4est_elliptic_eqns_build_residual (p4est, *d4est_ghost, *d4est_ghost_data, &prob_fcns, &sv, d4est_ops, d4est_geom,
                                        d4est_quad, d4est_factors);
    sv.u = v; sv.u0 = us; sv.Au = Jv;
    d4est_elliptic_eqns_apply_lhs (p4est, *d4est_ghost, *d4est_ghost_data, &prob_fcns, &sv, d4est_ops, d4est_geom,
                                   d4est_quad, d4est_factors);
  }

  d4est_solver_multigrid_t *mg_data =
    d4est_solver_multigrid_data_init (p4est, d4est_ops, d4est_geom, d4est_quad, d4est_ghost, d4est_ghost_data,
                                      d4est_factors, initial_grid_input, ini);
  d4est_solver_multigrid_user_callbacks_t *user_callbacks =
    d4est_solver_multigrid_matrix_operator_init (p4est, mg_data->num_of_levels, 0);
  d4est_solver_multigrid_set_user_callbacks (mg_data, user_callbacks);
  d4est_krylov_pc_t *pc = d4est_krylov_pc_multigrid_create (mg_data, multi_punctures_d4est_krylov_pc_setup_fcn);
  ctx.use_matrix_operator = 1;
  ctx.mg_data = mg_data;

  d4est_solver_fcg_params_t fcg_params;
  d4est_solver_fcg_input (p4est, ini, "d4est_solver_fcg", &fcg_params);
  d4est_solver_newton_params_t np = d4est_solver_newton_input (p4est, ini);

  /* recording wrappers */
  fcg_inner_pc_apply = pc->pc_apply;
  pc->pc_apply = fcg_counting_pc_apply;
  fcg_pc_calls = 0; fcg_nodes = local_nodes;
  nw_inner_build_residual = prob_fcns.build_residual;
  prob_fcns.build_residual = nw_recording_build_residual;
  nw_inner_apply_lhs = prob_fcns.apply_lhs;
  prob_fcns.apply_lhs = nw_tracing_apply_lhs;
  nw_res_calls = 0; nw_nodes = local_nodes;
  fcg_first_pc_x = P4EST_ALLOC_ZERO (double, local_nodes);
  fcg_first_pc_y = P4EST_ALLOC_ZERO (double, local_nodes);

  double t0 = wall ();
  int ierr = d4est_solver_newton_solve (p4est, &vecs, &prob_fcns, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom,
                                        d4est_quad, d4est_factors, (char *) ini, d4est_solver_fcg_solve, &fcg_params,
                                        (mg_data->num_of_levels <= 1) ? NULL : pc, 0);
  double t_newton = wall () - t0;
  const int steps = nw_res_calls - 1;
  double fcg_its[NEWTON_MAX_STEPS];
  for (int k = 0; k < steps && k < NEWTON_MAX_STEPS; k++)
    fcg_its[k] = (double) (nw_pc_calls_at_res[k + 1] - nw_pc_calls_at_res[k]);
  printf ("{\"elements\": %d, \"local_nodes\": %d, \"mg_levels\": %d, \"newton_steps\": %d, \"newton_ierr\": %d, "
          "\"fcg_iterations_per_step\": [", (int) p4est->local_num_quadrants, local_nodes, mg_data->num_of_levels, steps, ierr);
  for (int k = 0; k < steps; k++) printf ("%s%d", k ? ", " : "", (int) fcg_its[k]);
  printf ("], \"residual_norms\": [");
  for (int k = 0; k <= steps; k++) printf ("%s%.17g", k ? ", " : "", nw_fnrm[k]);
  printf ("], \"t_newton_s\": %.3f}\n", t_newton);

#ifdef D4B_WITH_DEVICE
  if (use_b200) {
    /* DROP-IN CHECK: the same UNMODIFIED d4est_solver_newton_solve + d4est_solver_fcg_solve, with
     * build_residual = d4est_b200_build_residual, apply_lhs = d4est_b200_apply_lhs and
     * pc = d4est_b200_pc_multigrid_create; the puncture terms are device functors, f is never evaluated on the host */
    d4est_laplacian_flux_data_t *flux_plain_res = d4est_laplacian_flux_new (p4est, ini, BC_ROBIN, &bc_plain);
    d4est_b200_ctx_t *jctx = d4est_b200_ctx_new (flux_plain), *rctx = d4est_b200_ctx_new (flux_plain_res);
    d4b_nonlinear_desc_t nl;
    memset (&nl, 0, sizeof (nl));
    nl.kind = D4B_NL_PUNCTURES;
    nl.n_punctures = mp->number_of_punctures;
    for (int i = 0; i < mp->number_of_punctures; i++) {
      nl.punctures[i][0] = mp->bh[i].M;
      for (int d = 0; d < 3; d++) {
        nl.punctures[i][1 + d] = mp->bh[i].XYZ[d];
        nl.punctures[i][4 + d] = mp->bh[i].P[d];
        nl.punctures[i][7 + d] = mp->bh[i].S[d];
      }
    }
    d4est_b200_set_nonlinear_term (jctx, &nl, rctx);
    d4est_elliptic_eqns_t b_fcns;
    b_fcns.build_residual = d4est_b200_build_residual;
    b_fcns.apply_lhs = d4est_b200_apply_lhs;
    b_fcns.user = jctx;
    d4est_solver_multigrid_t *bmg =
      d4est_solver_multigrid_data_init (p4est, d4est_ops, d4est_geom, d4est_quad, d4est_ghost, d4est_ghost_data,
                                        d4est_factors, initial_grid_input, ini);
    d4est_krylov_pc_t *bpc = d4est_b200_pc_multigrid_create (bmg, jctx, NULL);
    const int ref_steps = steps;
    int ref_its[NEWTON_MAX_STEPS];
    double ref_fnrm[NEWTON_MAX_STEPS + 1];
    for (int k = 0; k < steps; k++) ref_its[k] = (int) fcg_its[k];
    for (int k = 0; k <= steps; k++) ref_fnrm[k] = nw_fnrm[k];
    fcg_inner_pc_apply = bpc->pc_apply;
    bpc->pc_apply = fcg_counting_pc_apply;
    fcg_pc_calls = 0;
    nw_inner_build_residual = b_fcns.build_residual;
    b_fcns.build_residual = nw_recording_build_residual;
    nw_res_calls = 0;
    double *bu = P4EST_ALLOC_ZERO (double, local_nodes), *bAu = P4EST_ALLOC_ZERO (double, local_nodes);
    d4est_elliptic_data_t bv = vecs;
    bv.u = bu; bv.Au = bAu;
    double tb = wall ();
    int bierr = d4est_solver_newton_solve (p4est, &bv, &b_fcns, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom,
                                           d4est_quad, d4est_factors, (char *) ini, d4est_solver_fcg_solve, &fcg_params,
                                           bpc, 0);
    tb = wall () - tb;
    const int bsteps = nw_res_calls - 1;
    int same_its = (bsteps == ref_steps), k;
    double worst_norm = 0., un = 0., ud = 0.;
    printf ("{\"b200_newton_steps\": %d, \"reference_newton_steps\": %d, \"b200_newton_ierr\": %d, "
            "\"b200_fcg_iterations_per_step\": [", bsteps, ref_steps, bierr);
    for (k = 0; k < bsteps; k++) {
      const int its_k = nw_pc_calls_at_res[k + 1] - nw_pc_calls_at_res[k];
      printf ("%s%d", k ? ", " : "", its_k);
      if (k >= ref_steps || its_k != ref_its[k]) same_its = 0;
    }
    for (k = 0; k <= bsteps && k <= ref_steps; k++) {
      /* relative, with a floor at 1e-8 of the initial residual (the last norms are round-off of the first) */
      const double d = fabs (nw_fnrm[k] - ref_fnrm[k]) / (ref_fnrm[k] + 1e-8 * ref_fnrm[0]);
      if (d > worst_norm) worst_norm = d;
    }
    for (k = 0; k < local_nodes; k++) { un += (bu[k] - vecs.u[k]) * (bu[k] - vecs.u[k]); ud += vecs.u[k] * vecs.u[k]; }
    printf ("], \"same_iteration_counts\": %d, \"residual_norm_worst_rel_diff\": %.3e, \"newton_u_rel_l2_diff\": %.3e, "
            "\"u0_uploads\": %ld, \"device_mesh_uploads\": %ld, \"t_newton_reference_s\": %.3f, \"t_newton_b200_s\": %.3f}\n",
            same_its, worst_norm, sqrt (un / ud), jctx->linearisations, jctx->uploads + rctx->uploads, t_newton, tb);
    bpc->pc_apply = fcg_inner_pc_apply;
    d4est_b200_pc_multigrid_destroy (bpc);
    d4est_solver_multigrid_data_destroy (bmg);
    d4est_b200_ctx_destroy (jctx); d4est_b200_ctx_destroy (rctx);
    d4est_laplacian_flux_destroy (flux_plain_res);
    P4EST_FREE (bu); P4EST_FREE (bAu);
  }
#endif

  if (strcmp (out, "-")) {
    d4est_b200_flat_t *flat = d4est_b200_flatten (p4est, *d4est_ghost, d4est_factors, flux_plain);
    FILE *f = d4b_dump_open (out);
    dump_desc (f, &flat->desc);
    const int64_t lnq = d4est_factors->local_sizes.local_nodes_quad;
    d4b_dump_put1 (f, "xyz", D4B_F64, 3 * (uint64_t) local_nodes, d4est_factors->xyz);
    d4b_dump_put1 (f, "xyz_quad", D4B_F64, 3 * (uint64_t) lnq, d4est_factors->xyz_quad);
    double pun[MAX_PUNCTURES * 10];
    for (int i = 0; i < mp->number_of_punctures; i++) {
      pun[10 * i] = mp->bh[i].M;
      for (int d = 0; d < 3; d++) {
        pun[10 * i + 1 + d] = mp->bh[i].XYZ[d];
        pun[10 * i + 4 + d] = mp->bh[i].P[d];
        pun[10 * i + 7 + d] = mp->bh[i].S[d];
      }
    }
    d4b_dump_put1 (f, "punctures", D4B_F64, 10 * (uint64_t) mp->number_of_punctures, pun);
    d4b_dump_put1 (f, "nw_us", D4B_F64, local_nodes, us);
    d4b_dump_put1 (f, "nw_v", D4B_F64, local_nodes, v);
    d4b_dump_put1 (f, "nw_F_us", D4B_F64, local_nodes, Fus);
    d4b_dump_put1 (f, "nw_J_us_v", D4B_F64, local_nodes, Jv);
    double prm[8] = { np.atol, np.rtol, (double) np.imin, (double) np.imax, fcg_params.atol, fcg_params.rtol,
                      (double) fcg_params.imax, (double) mg_data->num_of_levels };
    d4b_dump_put1 (f, "nw_params", D4B_F64, 8, prm);
    d4b_dump_put1 (f, "nw_residual_norms", D4B_F64, steps + 1, nw_fnrm);
    d4b_dump_put1 (f, "nw_fcg_iterations", D4B_F64, steps, fcg_its);
    d4b_dump_put1 (f, "nw_fcg_r2", D4B_F64, fcg_pc_calls < 512 ? fcg_pc_calls : 512, fcg_r2_hist);
    d4b_dump_put1 (f, "nw_u", D4B_F64, local_nodes, vecs.u);
    /* the first preconditioner application of the first Newton step (linearised about u = 0): x -> y */
    d4b_dump_put1 (f, "nw_pc_x0", D4B_F64, local_nodes, fcg_first_pc_x);
    d4b_dump_put1 (f, "nw_pc_y0", D4B_F64, local_nodes, fcg_first_pc_y);
    d4est_solver_multigrid_smoother_cheby_t *cheby = mg_data->smoother->user;
    double cprm[5] = { (double) cheby->cheby_imax, (double) cheby->cheby_eigs_cg_imax,
                       cheby->cheby_eigs_lmax_lmin_ratio, cheby->cheby_eigs_max_multiplier,
                       (double) cheby->cheby_use_new_cg_eigs };
    d4b_dump_put1 (f, "mg_cheby_params", D4B_F64, 5, cprm);
    fclose (f);
    d4est_b200_flat_destroy (flat);
  }
  pc->pc_apply = fcg_inner_pc_apply;
  P4EST_FREE (fcg_first_pc_x); P4EST_FREE (fcg_first_pc_y);
  fcg_first_pc_x = fcg_first_pc_y = NULL;
  d4est_krylov_pc_multigrid_destroy (pc);
  d4est_solver_multigrid_matrix_operator_destroy (user_callbacks);
  d4est_solver_multigrid_data_destroy (mg_data);
  P4EST_FREE (us); P4EST_FREE (v); P4EST_FREE (Fus); P4EST_FREE (Jv);
  P4EST_FREE (u); P4EST_FREE (Au); P4EST_FREE (rhs);
  d4est_laplacian_flux_destroy (flux_plain);
  d4est_laplacian_with_opt_flux_destroy (flux_jac);
  d4est_laplacian_with_opt_flux_destroy (flux_res);
  multi_punctures_params_destroy (mp);
}

int
main (int argc, char *argv[])
{
  if (argc < 3) {
    fprintf (stderr, "usage: %s <ini> <out.d4b|-> [--amr N] [--time REPS] [--bc zero|poly]"
             " [--uniform-deg P] [--cheby IMAX EIGS_IMAX] [--mg VCYCLES]\n", argv[0]);
    return 2;
  }
  const char *ini = argv[1];
  const char *out = argv[2];
  int amr_steps = 0, time_reps = 0, bc_poly = 0, cheby_imax = 0, cheby_eigs_imax = 0, mg_vcycles = 0, use_b200 = 0;
  int dump_quad = 0, fcg_imax = 0, bc_robin = 0, b200_apply_only = 0, no_opt = 0, newton = 0;
  const char *geom_out = NULL;
  for (int a = 3; a < argc; a++) {
    if (!strcmp (argv[a], "--amr")) amr_steps = atoi (argv[++a]);
    else if (!strcmp (argv[a], "--time")) time_reps = atoi (argv[++a]);
    else if (!strcmp (argv[a], "--bc")) { ++a; bc_poly = !strcmp (argv[a], "poly"); bc_robin = !strcmp (argv[a], "robin"); }
    else if (!strcmp (argv[a], "--uniform-deg")) g_uniform_deg = atoi (argv[++a]);
    else if (!strcmp (argv[a], "--cheby")) { cheby_imax = atoi (argv[++a]); cheby_eigs_imax = atoi (argv[++a]); }
    else if (!strcmp (argv[a], "--mg")) { mg_vcycles = atoi (argv[++a]); }
    else if (!strcmp (argv[a], "--quad")) { dump_quad = 1; }
    else if (!strcmp (argv[a], "--fcg")) { fcg_imax = atoi (argv[++a]); }
    else if (!strcmp (argv[a], "--b200")) { use_b200 = 1; }
    else if (!strcmp (argv[a], "--b200-apply-only")) { use_b200 = 1; b200_apply_only = 1; }
    else if (!strcmp (argv[a], "--geom")) { geom_out = argv[++a]; }
    /* the with_opt flux knows no maxpp1_sqr_over_minh (d4est_laplacian_with_opt_flux_sipg.c:1350-1366 aborts):
     * both outputs then come from the plain operator */
    else if (!strcmp (argv[a], "--no-opt")) { no_opt = 1; }
    else if (!strcmp (argv[a], "--newton")) { newton = 1; }
    else { fprintf (stderr, "unknown option %s\n", argv[a]); return 2; }
  }

  sc_MPI_Comm mpicomm = sc_MPI_COMM_WORLD;
  int mpiret = sc_MPI_Init (&argc, &argv);
  SC_CHECK_MPI (mpiret);
  sc_init (mpicomm, 1, 1, NULL, SC_LP_ERROR);
  p4est_init (NULL, SC_LP_ERROR);

  /* zlog must be initialised before any reference module logs */
  char conf[256];
  snprintf (conf, sizeof (conf), "/tmp/d4b_zlog_%d.conf", (int) getpid ());
  FILE *cf = fopen (conf, "w");
  fprintf (cf, "[formats]\nplain = \"%%m%%n\"\n[rules]\n*.ERROR >stderr; plain\n");
  fclose (cf);
  if (zlog_init (conf)) { fprintf (stderr, "zlog_init failed\n"); return 2; }
  unlink (conf);

  zlog_category_t *c_geom = zlog_get_category ("d4est_geometry");
  d4est_geometry_t *d4est_geom = d4est_geometry_new (0, ini, "geometry", c_geom);
  d4est_mesh_initial_extents_t *initial_grid_input =
    d4est_mesh_initial_extents_parse (ini, d4est_geom);

  p4est_t *p4est = p4est_new_ext (mpicomm, d4est_geom->p4est_conn,
                                  initial_grid_input->min_quadrants,
                                  initial_grid_input->min_level,
                                  initial_grid_input->fill_uniform,
                                  sizeof (d4est_element_data_t), NULL, NULL);
  p4est_partition (p4est, 1, NULL);
  p4est_balance (p4est, P4EST_CONNECT_FULL, NULL);
#ifdef D4B_WITH_DEVICE
  /* the reference-side bootstrap of the collective context: device selection + (on > 1 rank) the NCCL id over
   * p4est->mpicomm; this serial build has one rank */
  if (use_b200) d4est_b200_comm_init_from_p4est (p4est, 0);
#endif

  d4est_ghost_t *d4est_ghost = d4est_ghost_init (p4est);
  d4est_operators_t *d4est_ops = d4est_ops_init (20);
  d4est_mesh_data_t *d4est_factors = d4est_mesh_data_init (p4est);
  d4est_quadrature_t *d4est_quad =
    d4est_quadrature_new (p4est, d4est_ops, d4est_geom, ini, "quadrature");

  double t0 = wall ();
  d4est_mesh_local_sizes_t local_sizes =
    d4est_mesh_update (p4est, &d4est_ghost, d4est_ops, d4est_geom, d4est_quad,
                       d4est_factors, initial_grid_input, INITIALIZE_GHOST,
                       INITIALIZE_QUADRATURE_DATA, INITIALIZE_GEOMETRY_DATA,
                       INITIALIZE_GEOMETRY_ALIASES,
                       (g_uniform_deg > 0) ? set_uniform_deg : d4est_mesh_set_initial_extents,
                       (void *) initial_grid_input);
  initial_grid_input->initial_nodes = local_sizes.local_nodes;

  if (amr_steps > 0) {
    d4est_amr_t *d4est_amr_random = d4est_amr_init_random_hp (p4est, amr_steps);
    for (int i = 0; i < amr_steps; i++) {
      d4est_amr_step (p4est, d4est_ops, d4est_amr_random, NULL, NULL, NULL, NULL);
      local_sizes =
        d4est_mesh_update (p4est, &d4est_ghost, d4est_ops, d4est_geom, d4est_quad,
                           d4est_factors, initial_grid_input, INITIALIZE_GHOST,
                           INITIALIZE_QUADRATURE_DATA, INITIALIZE_GEOMETRY_DATA,
                           INITIALIZE_GEOMETRY_ALIASES,
                           d4est_mesh_set_quadratures_after_amr,
                           (void *) initial_grid_input);
    }
    d4est_amr_destroy (d4est_amr_random);
  }
  double t_mesh = wall () - t0;

  d4est_field_type_t field_type = NODAL;
  d4est_ghost_data_t *d4est_ghost_data =
    d4est_ghost_data_init (p4est, d4est_ghost, &field_type, 1);
  int local_nodes = local_sizes.local_nodes;

  if (newton) {
    run_newton (p4est, &d4est_ghost, &d4est_ghost_data, d4est_ops, d4est_geom, d4est_quad, d4est_factors,
                initial_grid_input, ini, out, local_nodes, use_b200);
    d4est_ghost_data_destroy (d4est_ghost_data);
    d4est_ghost_destroy (d4est_ghost);
    d4est_mesh_initial_extents_destroy (initial_grid_input);
    d4est_quadrature_destroy (p4est, d4est_ops, d4est_geom, d4est_quad);
    d4est_mesh_data_destroy (d4est_factors);
    d4est_ops_destroy (d4est_ops);
    p4est_destroy (p4est);
    d4est_geometry_destroy (d4est_geom);
    zlog_fini ();
    sc_finalize ();
    return 0;
  }
  d4est_laplacian_dirichlet_bc_t bc_data_for_lhs;
  bc_data_for_lhs.dirichlet_fcn = bc_poly ? drv_poly_fcn : drv_zero_fcn;
  bc_data_for_lhs.eval_method = EVAL_BNDRY_FCN_ON_LOBATTO;
  bc_data_for_lhs.user = NULL;

  d4est_laplacian_robin_bc_t bc_robin_data;
  bc_robin_data.robin_coeff = drv_robin_coeff;
  bc_robin_data.robin_rhs = drv_robin_rhs;
  bc_robin_data.user = NULL;
  d4est_laplacian_with_opt_robin_bc_t bc_robin_opt;
  bc_robin_opt.robin_coeff = (d4est_laplacian_with_opt_mortar_xyz_fcn_t) drv_robin_coeff;
  bc_robin_opt.robin_rhs = (d4est_laplacian_with_opt_mortar_xyz_fcn_t) drv_robin_rhs;
  bc_robin_opt.user = NULL;

  d4est_laplacian_flux_data_t *flux_plain = bc_robin
    ? d4est_laplacian_flux_new (p4est, ini, BC_ROBIN, &bc_robin_data)
    : d4est_laplacian_flux_new (p4est, ini, BC_DIRICHLET, &bc_data_for_lhs);
  d4est_laplacian_with_opt_flux_data_t *flux_opt = no_opt ? NULL : bc_robin
    ? d4est_laplacian_with_opt_flux_new (p4est, ini, BC_ROBIN, &bc_robin_opt)
    : d4est_laplacian_with_opt_flux_new (p4est, ini, BC_DIRICHLET, &bc_data_for_lhs);

  d4est_elliptic_eqns_t fcns_plain = {.build_residual = NULL,.apply_lhs = apply_plain,.user = flux_plain };
  d4est_elliptic_eqns_t fcns_opt = {.build_residual = NULL,.apply_lhs = apply_with_opt,.user = flux_opt };
  if (no_opt) fcns_opt = fcns_plain;

  double *u = P4EST_ALLOC (double, local_nodes);
  double *Au_plain = P4EST_ALLOC_ZERO (double, local_nodes);
  double *Au_opt = P4EST_ALLOC_ZERO (double, local_nodes);
  d4est_mesh_init_field (p4est, u, drv_poly_fcn, d4est_ops, d4est_geom, d4est_factors,
                         INIT_FIELD_ON_LOBATTO, NULL);
  srand (102321);
  for (int i = 0; i < local_nodes; i++) u[i] += (double) rand () / (double) RAND_MAX;

  d4est_elliptic_data_t vecs;
  memset (&vecs, 0, sizeof (vecs));
  vecs.u = u;
  vecs.Au = Au_plain;
  vecs.local_nodes = local_nodes;
  vecs.field_types = &field_type;
  vecs.num_of_fields = 1;

  t0 = wall ();
  d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_plain, &vecs,
                                 d4est_ops, d4est_geom, d4est_quad, d4est_factors);
  double t_plain = wall () - t0;
  vecs.Au = Au_opt;
  t0 = wall ();
  d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_opt, &vecs,
                                 d4est_ops, d4est_geom, d4est_quad, d4est_factors);
  double t_opt = wall () - t0;

  /* timing loop for the CPU baseline: best of REPS for each variant */
  double best_plain = t_plain, best_opt = t_opt;
  double *Au_tmp = P4EST_ALLOC_ZERO (double, local_nodes);
  vecs.Au = Au_tmp;
  for (int r = 0; r < time_reps; r++) {
    t0 = wall ();
    d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_opt, &vecs,
                                   d4est_ops, d4est_geom, d4est_quad, d4est_factors);
    double t = wall () - t0;
    if (t < best_opt) best_opt = t;
    t0 = wall ();
    d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_plain, &vecs,
                                   d4est_ops, d4est_geom, d4est_quad, d4est_factors);
    t = wall () - t0;
    if (t < best_plain) best_plain = t;
  }

  double maxdiff = 0.;
  for (int i = 0; i < local_nodes; i++) {
    double d = fabs (Au_plain[i] - Au_opt[i]);
    if (d > maxdiff) maxdiff = d;
  }
  printf ("{\"elements\": %d, \"local_nodes\": %d, \"t_mesh_s\": %.6f, \"t_apply_plain_s\": %.6f, "
          "\"t_apply_with_opt_s\": %.6f, \"plain_vs_opt_maxabs\": %.3e, \"u_sum\": %.17g, \"Au_sum\": %.17g}\n",
          (int) p4est->local_num_quadrants, local_nodes, t_mesh, best_plain, best_opt, maxdiff,
          d4est_util_sum_array_dbl (u, local_nodes), d4est_util_sum_array_dbl (Au_plain, local_nodes));

#ifdef D4B_WITH_DEVICE
  /* DROP-IN CHECK (ref_driver_b200 only, needs a GPU): the reference's own code calls
   * eqns->apply_lhs = d4est_b200_apply_lhs instead of d4est_laplacian_apply_aij --
   * once directly, then from inside its UNMODIFIED multigrid solver (smoother,
   * cg_eigs, bottom CG all call apply_lhs through the same function pointer). */
  if (use_b200 && !bc_poly) {
    d4est_b200_ctx_t *bctx = d4est_b200_ctx_new (flux_plain);
    d4est_elliptic_eqns_t fcns_b200 = {.build_residual = NULL,.apply_lhs = d4est_b200_apply_lhs,.user = bctx };
    double *Au_b = P4EST_ALLOC_ZERO (double, local_nodes);
    vecs.Au = Au_b;
    d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_b200, &vecs,
                                   d4est_ops, d4est_geom, d4est_quad, d4est_factors);
    double t1 = wall (), t_b200_best = 1e30;
    for (int r = 0; r < 5; r++) {
      double t1r = wall ();
      d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_b200, &vecs,
                                     d4est_ops, d4est_geom, d4est_quad, d4est_factors);
      t1r = wall () - t1r;
      if (t1r < t_b200_best) t_b200_best = t1r;
    }
    double t_b200 = (wall () - t1) / 5.;
    double num = 0., den = 0.;
    for (int i = 0; i < local_nodes; i++) {
      num += (Au_b[i] - Au_plain[i]) * (Au_b[i] - Au_plain[i]);
      den += Au_plain[i] * Au_plain[i];
    }
    /* the same call with the reference's vectors page-locked once (d4est_b200_pin_vectors) */
    d4est_b200_pin_vectors (&vecs);
    d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_b200, &vecs,
                                   d4est_ops, d4est_geom, d4est_quad, d4est_factors);
    double t2 = wall (), t_pin_best = 1e30;
    for (int r = 0; r < 5; r++) {
      double t2r = wall ();
      d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_b200, &vecs,
                                     d4est_ops, d4est_geom, d4est_quad, d4est_factors);
      t2r = wall () - t2r;
      if (t2r < t_pin_best) t_pin_best = t2r;
    }
    double t_pin = (wall () - t2) / 5.;
    double num2 = 0.;
    for (int i = 0; i < local_nodes; i++) num2 += (Au_b[i] - Au_plain[i]) * (Au_b[i] - Au_plain[i]);
    d4est_b200_unpin_vectors (&vecs);
    printf ("{\"b200_apply_rel_l2\": %.3e, \"t_apply_b200_host_vectors_s\": %.6f, \"t_apply_b200_host_vectors_best_s\": %.6f, "
            "\"t_apply_reference_s\": %.6f, \"t_apply_b200_pinned_vectors_s\": %.6f, "
            "\"t_apply_b200_pinned_vectors_best_s\": %.6f, \"b200_apply_pinned_rel_l2\": %.3e}\n",
            sqrt (num / den), t_b200, t_b200_best, best_plain < best_opt ? best_plain : best_opt, t_pin, t_pin_best,
            sqrt (num2 / den));
    if (!b200_apply_only) {
      /* the same callback on a context that generates the geometric factors on the device
       * (d4est_b200_ctx_use_device_geometry) instead of copying the reference's arrays */
      d4est_b200_ctx_t *gctx = d4est_b200_ctx_new (flux_plain);
      d4est_b200_ctx_use_device_geometry (gctx, d4est_geom, initial_grid_input->face_h_type);
      d4est_elliptic_eqns_t fcns_g = {.build_residual = NULL,.apply_lhs = d4est_b200_apply_lhs,.user = gctx };
      double *Au_g = P4EST_ALLOC_ZERO (double, local_nodes);
      vecs.Au = Au_g;
      d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_g, &vecs,
                                     d4est_ops, d4est_geom, d4est_quad, d4est_factors);
      double n2 = 0., d2 = 0.;
      for (int i = 0; i < local_nodes; i++) {
        n2 += (Au_g[i] - Au_plain[i]) * (Au_g[i] - Au_plain[i]);
        d2 += Au_plain[i] * Au_plain[i];
      }
      printf ("{\"b200_device_geometry_apply_rel_l2\": %.3e}\n", sqrt (n2 / d2));
      vecs.Au = Au_b;
      P4EST_FREE (Au_g);
      d4est_b200_ctx_destroy (gctx);
    }
    if (!b200_apply_only) {
      /* Jacobian apply of a nonlinear problem: Laplacian + mass-type term.  Reference: the
       * problem files' loop over d4est_quadrature_apply_fofufofvlilj added to apply_aij
       * (constant_density_star_fcns.h:528-600); ours: d4est_b200_set_mass_term, then the same
       * apply_lhs callback as before. */
      double *u0 = P4EST_ALLOC (double, local_nodes);
      double *term = P4EST_ALLOC_ZERO (double, local_nodes);
      for (int i = 0; i < local_nodes; i++) u0[i] = 0.1 * u[i] - 0.05;
      for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
        p4est_tree_t *tree = p4est_tree_array_index (p4est->trees, tt);
        sc_array_t *tquadrants = &tree->quadrants;
        for (int q = 0; q < (int) tquadrants->elem_count; ++q) {
          p4est_quadrant_t *quad = p4est_quadrant_array_index (tquadrants, q);
          d4est_element_data_t *ed = quad->p.user_data;
          d4est_quadrature_volume_t mo;
          mo.dq = ed->dq; mo.tree = ed->tree; mo.element_id = ed->id;
          mo.q[0] = ed->q[0]; mo.q[1] = ed->q[1]; mo.q[2] = ed->q[2];
          double *J_quad = d4est_mesh_get_jacobian_on_quadrature_points (d4est_factors, ed);
          d4est_mesh_data_on_element_t md = d4est_mesh_data_on_element (d4est_factors, ed);
          d4est_quadrature_apply_fofufofvlilj (d4est_ops, d4est_geom, d4est_quad, &mo, QUAD_OBJECT_VOLUME,
                                               QUAD_INTEGRAND_UNKNOWN, &u[ed->nodal_stride], &u0[ed->nodal_stride],
                                               NULL, ed->deg, md.xyz_quad, J_quad, ed->deg_quad,
                                               &term[ed->nodal_stride], star_like_jac_fcn, NULL, NULL, NULL,
                                               QUAD_APPLY_MATRIX, 0, NULL);
        }
      }
      d4est_b200_set_mass_term (bctx, p4est, d4est_ghost, d4est_ops, d4est_geom, d4est_quad, d4est_factors,
                                u0, star_like_jac_fcn, NULL);
      d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_b200, &vecs,
                                     d4est_ops, d4est_geom, d4est_quad, d4est_factors);
      double jn = 0., jd = 0.;
      for (int i = 0; i < local_nodes; i++) {
        const double want = Au_plain[i] + term[i];
        jn += (Au_b[i] - want) * (Au_b[i] - want);
        jd += want * want;
      }
      printf ("{\"b200_jacobian_apply_rel_l2\": %.3e}\n", sqrt (jn / jd));
      d4est_b200_set_mass_term (bctx, p4est, d4est_ghost, d4est_ops, d4est_geom, d4est_quad, d4est_factors,
                                NULL, NULL, NULL);
      P4EST_FREE (u0); P4EST_FREE (term);
    }
    if (mg_vcycles > 0) {
      double r2s[3][64], tsolve[3];
      double *usol[3];
      /* 0: stock apply_lhs, 1: B200 apply_lhs under the reference's smoother,
       * 2: B200 smoother (d4est_b200_smoother_cheby_init) + B200 apply_lhs for the rest */
      for (int which = 0; which < 3; which++) {
        d4est_solver_multigrid_t *mg_data =
          d4est_solver_multigrid_data_init (p4est, d4est_ops, d4est_geom, d4est_quad, &d4est_ghost,
                                            &d4est_ghost_data, d4est_factors, initial_grid_input, ini);
        if (which == 2) {
          d4est_solver_multigrid_smoother_cheby_destroy (mg_data->smoother);
          mg_data->smoother = d4est_b200_smoother_cheby_init (p4est, mg_data->num_of_levels, ini, bctx);
        }
        usol[which] = P4EST_ALLOC_ZERO (double, local_nodes);
        double *mAu = P4EST_ALLOC_ZERO (double, local_nodes);
        d4est_elliptic_data_t mv = vecs;
        mv.u = usol[which]; mv.Au = mAu; mv.rhs = Au_plain;
        double t2 = wall ();
        for (int k = 0; k < mg_vcycles && k < 63; k++) {
          d4est_solver_multigrid_solve (p4est, &mv, which ? &fcns_b200 : &fcns_plain, mg_data);
          r2s[which][k] = mg_data->vcycle_r2_global_current;
        }
        tsolve[which] = wall () - t2;
        P4EST_FREE (mAu);
        d4est_solver_multigrid_data_destroy (mg_data);
      }
      double worst = 0., unum = 0., uden = 0., worst2 = 0., unum2 = 0.;
      for (int k = 0; k < mg_vcycles && k < 63; k++) {
        double rr = fabs (r2s[0][k] - r2s[1][k]) / r2s[0][k];
        if (rr > worst && r2s[0][k] > 1e-20 * r2s[0][0]) worst = rr;
        rr = fabs (r2s[0][k] - r2s[2][k]) / r2s[0][k];
        if (rr > worst2 && r2s[0][k] > 1e-20 * r2s[0][0]) worst2 = rr;
      }
      for (int i = 0; i < local_nodes; i++) {
        unum += (usol[0][i] - usol[1][i]) * (usol[0][i] - usol[1][i]);
        unum2 += (usol[0][i] - usol[2][i]) * (usol[0][i] - usol[2][i]);
        uden += usol[0][i] * usol[0][i];
      }
      printf ("{\"mg_vcycles\": %d, \"r2_reference\": [", mg_vcycles);
      for (int k = 0; k < mg_vcycles && k < 63; k++) printf ("%s%.6e", k ? ", " : "", r2s[0][k]);
      printf ("], \"r2_b200_apply\": [");
      for (int k = 0; k < mg_vcycles && k < 63; k++) printf ("%s%.6e", k ? ", " : "", r2s[1][k]);
      printf ("], \"r2_worst_rel_diff\": %.3e, \"u_rel_l2_diff\": %.3e, \"t_mg_reference_s\": %.4f, "
              "\"t_mg_b200_apply_s\": %.4f, \"device_mesh_uploads\": %ld, \"b200_applies\": %ld, "
              "\"smoother_r2_worst_rel_diff\": %.3e, \"smoother_u_rel_l2_diff\": %.3e, \"t_mg_b200_smoother_s\": %.4f}\n",
              worst, sqrt (unum / uden), tsolve[0], tsolve[1], bctx->uploads, bctx->applies,
              worst2, sqrt (unum2 / uden), tsolve[2]);
      P4EST_FREE (usol[0]); P4EST_FREE (usol[1]); P4EST_FREE (usol[2]);

      /* third entry point: the whole preconditioner.  y = PC(x) through the reference's
       * d4est_krylov_pc_multigrid (stock apply_lhs) against d4est_b200_pc_multigrid_create,
       * both called the way the Krylov wrapper calls them (pc->pc_apply (pc, x, y),
       * src/Solver/d4est_krylov_pc.h:69-89) with vcycle_imax = --mg. */
      {
        double *ypc[2];
        double tpc[2] = { 0., 0. }, tpc_first = 0.;
        long pc_builds = 0, pc_vcycles = 0;
        int ref_vcycles = 0;
        for (int which = 0; which < 2; which++) {
          d4est_solver_multigrid_t *mg_data =
            d4est_solver_multigrid_data_init (p4est, d4est_ops, d4est_geom, d4est_quad, &d4est_ghost,
                                              &d4est_ghost_data, d4est_factors, initial_grid_input, ini);
          mg_data->vcycle_imax = mg_vcycles;
          ypc[which] = P4EST_ALLOC_ZERO (double, local_nodes);
          d4est_elliptic_data_t kv = vecs;
          krylov_ctx_t kct;
          memset (&kct, 0, sizeof (kct));
          kct.p4est = p4est;
          kct.vecs = &kv;
          kct.fcns = which ? &fcns_b200 : &fcns_plain;
          d4est_krylov_pc_t *pc = which ? d4est_b200_pc_multigrid_create (mg_data, bctx, NULL)
                                        : d4est_krylov_pc_multigrid_create (mg_data, NULL);
          pc->pc_ctx = &kct;
          double t3 = wall ();
          pc->pc_apply (pc, Au_plain, ypc[which]);
          tpc_first = wall () - t3;
          if (which) {                       /* second call: hierarchy already on the device */
            t3 = wall ();
            pc->pc_apply (pc, Au_plain, ypc[which]);
            tpc[1] = wall () - t3;
            d4est_b200_pc_multigrid_data_t *pcd = pc->pc_data;
            pc_builds = pcd->builds; pc_vcycles = pcd->vcycles / pcd->applies;
            d4est_b200_pc_multigrid_destroy (pc);
          }
          else {
            tpc[0] = tpc_first;
            ref_vcycles = mg_data->vcycle_num_finished;
            d4est_krylov_pc_multigrid_destroy (pc);
          }
          d4est_solver_multigrid_data_destroy (mg_data);
        }
        double pnum = 0., pden = 0.;
        for (int i = 0; i < local_nodes; i++) {
          pnum += (ypc[0][i] - ypc[1][i]) * (ypc[0][i] - ypc[1][i]);
          pden += ypc[0][i] * ypc[0][i];
        }
        printf ("{\"pc_vcycles_reference\": %d, \"pc_vcycles_b200\": %ld, \"pc_y_rel_l2_diff\": %.3e, "
                "\"t_pc_reference_s\": %.4f, \"t_pc_b200_s\": %.4f, \"pc_hierarchy_builds\": %ld}\n",
                ref_vcycles, pc_vcycles, sqrt (pnum / pden), tpc[0], tpc[1], pc_builds);
        P4EST_FREE (ypc[0]); P4EST_FREE (ypc[1]);
      }

      /* the reference's own Krylov solver (d4est_solver_fcg_solve, unmodified) driving the two
       * preconditioners: iteration counts and solutions must agree */
      if (fcg_imax > 0) {
        int its[2];
        double tf[2], *fus[2];
        for (int which = 0; which < 2; which++) {
          d4est_solver_multigrid_t *mg_data =
            d4est_solver_multigrid_data_init (p4est, d4est_ops, d4est_geom, d4est_quad, &d4est_ghost,
                                              &d4est_ghost_data, d4est_factors, initial_grid_input, ini);
          d4est_krylov_pc_t *pc = which ? d4est_b200_pc_multigrid_create (mg_data, bctx, NULL)
                                        : d4est_krylov_pc_multigrid_create (mg_data, NULL);
          fcg_inner_pc_apply = pc->pc_apply;
          pc->pc_apply = fcg_counting_pc_apply;
          fcg_pc_calls = 0; fcg_nodes = local_nodes;
          d4est_solver_fcg_params_t fp;
          memset (&fp, 0, sizeof (fp));
          fp.imax = fcg_imax; fp.rtol = 1e-10; fp.atol = 0.;
          fus[which] = P4EST_ALLOC_ZERO (double, local_nodes);
          double *fAu = P4EST_ALLOC_ZERO (double, local_nodes);
          d4est_elliptic_data_t fv = vecs;
          fv.u = fus[which]; fv.Au = fAu; fv.rhs = Au_plain;
          double t4 = wall ();
          d4est_solver_fcg_solve (p4est, &fv, which ? &fcns_b200 : &fcns_plain, &d4est_ghost, &d4est_ghost_data,
                                  d4est_ops, d4est_geom, d4est_quad, d4est_factors, &fp, pc, -1, -1);
          tf[which] = wall () - t4;
          its[which] = fcg_pc_calls;
          pc->pc_apply = fcg_inner_pc_apply;
          if (which) d4est_b200_pc_multigrid_destroy (pc); else d4est_krylov_pc_multigrid_destroy (pc);
          P4EST_FREE (fAu);
          d4est_solver_multigrid_data_destroy (mg_data);
        }
        double fn = 0., fd = 0.;
        for (int i = 0; i < local_nodes; i++) {
          fn += (fus[0][i] - fus[1][i]) * (fus[0][i] - fus[1][i]);
          fd += fus[0][i] * fus[0][i];
        }
        printf ("{\"fcg_iters_reference\": %d, \"fcg_iters_b200_pc\": %d, \"fcg_u_rel_l2_diff\": %.3e, "
                "\"t_fcg_reference_s\": %.4f, \"t_fcg_b200_pc_s\": %.4f}\n",
                its[0], its[1], sqrt (fn / fd), tf[0], tf[1]);
        P4EST_FREE (fus[0]); P4EST_FREE (fus[1]);
      }
    }
    if (amr_steps > 0) {
      /* STALE-CACHE CHECK: the reference reuses one d4est_mesh_data_t across d4est_mesh_update calls, and a
       * +1/-1 p-adapt keeps the element, node, quadrature-node and mortar-node counts.  Two elements of the
       * same region with different degrees exchange them; the context must notice and upload a new mesh. */
      for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree && g_swap_b < 0; ++tt) {
        p4est_tree_t *tree = p4est_tree_array_index (p4est->trees, tt);
        sc_array_t *tquadrants = &tree->quadrants;
        g_swap_a = -1;
        for (int q = 0; q < (int) tquadrants->elem_count && g_swap_b < 0; ++q) {
          p4est_quadrant_t *quad = p4est_quadrant_array_index (tquadrants, q);
          d4est_element_data_t *ed = quad->p.user_data;
          if (g_swap_a < 0) { g_swap_a = ed->id; g_swap_deg_a = ed->deg; }
          else if (ed->deg != g_swap_deg_a) { g_swap_b = ed->id; g_swap_deg_b = ed->deg; }
        }
      }
      if (g_swap_b >= 0) {
        const long uploads0 = bctx->uploads;
        d4est_mesh_local_sizes_t ls2 =
          d4est_mesh_update (p4est, &d4est_ghost, d4est_ops, d4est_geom, d4est_quad, d4est_factors,
                             initial_grid_input, INITIALIZE_GHOST, INITIALIZE_QUADRATURE_DATA,
                             INITIALIZE_GEOMETRY_DATA, INITIALIZE_GEOMETRY_ALIASES, swap_deg_cb,
                             (void *) initial_grid_input);
        double *Au_r2 = P4EST_ALLOC_ZERO (double, local_nodes), *Au_b2 = P4EST_ALLOC_ZERO (double, local_nodes);
        vecs.Au = Au_r2;
        d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_plain, &vecs,
                                       d4est_ops, d4est_geom, d4est_quad, d4est_factors);
        vecs.Au = Au_b2;
        d4est_elliptic_eqns_apply_lhs (p4est, d4est_ghost, d4est_ghost_data, &fcns_b200, &vecs,
                                       d4est_ops, d4est_geom, d4est_quad, d4est_factors);
        double sn = 0., sd = 0., moved = 0.;
        for (int i = 0; i < local_nodes; i++) {
          sn += (Au_b2[i] - Au_r2[i]) * (Au_b2[i] - Au_r2[i]);
          sd += Au_r2[i] * Au_r2[i];
          moved += (Au_r2[i] - Au_plain[i]) * (Au_r2[i] - Au_plain[i]);
        }
        printf ("{\"b200_apply_after_degree_swap_rel_l2\": %.3e, \"same_counts\": %d, \"uploads_for_swapped_mesh\": %ld, "
                "\"reference_change_rel_l2\": %.3e}\n", sqrt (sn / sd),
                (int) (ls2.local_nodes == local_nodes), bctx->uploads - uploads0, sqrt (moved / sd));
        P4EST_FREE (Au_r2); P4EST_FREE (Au_b2);
      }
    }
    vecs.Au = Au_tmp;
    P4EST_FREE (Au_b);
    d4est_b200_ctx_destroy (bctx);
  }
  if (use_b200 && bc_poly) {
    /* right-hand side with strong boundary data: the reference's builder against the drop-in */
    d4est_b200_ctx_t *rctx = d4est_b200_ctx_new (flux_plain);
    double worst = 0.;
    for (int opt = 0; opt < 2; opt++) {
      d4est_mesh_init_field_option_t io = opt ? INIT_FIELD_ON_QUAD : INIT_FIELD_ON_LOBATTO;
      double *rhs_ref = P4EST_ALLOC_ZERO (double, local_nodes), *rhs_b = P4EST_ALLOC_ZERO (double, local_nodes);
      d4est_laplacian_build_rhs_with_strong_bc (p4est, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom, d4est_quad,
                                                d4est_factors, &vecs, flux_plain, rhs_ref, drv_rhs_fcn, io, NULL, 0);
      d4est_b200_build_rhs_with_strong_bc (rctx, p4est, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom, d4est_quad,
                                           d4est_factors, &vecs, flux_plain, rhs_b, drv_rhs_fcn, io, NULL, 0);
      double num = 0., den = 0.;
      for (int i = 0; i < local_nodes; i++) {
        num += (rhs_b[i] - rhs_ref[i]) * (rhs_b[i] - rhs_ref[i]);
        den += rhs_ref[i] * rhs_ref[i];
      }
      if (sqrt (num / den) > worst) worst = sqrt (num / den);
      P4EST_FREE (rhs_ref); P4EST_FREE (rhs_b);
    }
    printf ("{\"b200_build_rhs_rel_l2\": %.3e}\n", worst);
    d4est_b200_ctx_destroy (rctx);
    {
      /* estimator: the reference's d4est_estimator_bi_compute against the drop-in with the same arguments */
      const int nl = p4est->local_num_quadrants;
      double *rhs_e = P4EST_ALLOC_ZERO (double, local_nodes), *res = P4EST_ALLOC_ZERO (double, local_nodes);
      d4est_laplacian_build_rhs_with_strong_bc (p4est, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom, d4est_quad,
                                                d4est_factors, &vecs, flux_plain, rhs_e, drv_rhs_fcn,
                                                INIT_FIELD_ON_LOBATTO, NULL, 0);
      d4est_elliptic_data_t ev = vecs;
      ev.u = u; ev.Au = res; ev.rhs = rhs_e;
      d4est_elliptic_eqns_t fcns_res = {.build_residual = drv_build_residual,.apply_lhs = apply_plain,.user = flux_plain };
      d4est_estimator_bi_penalty_data_t pd;
      memset (&pd, 0, sizeof (pd));
      pd.u_penalty_fcn = houston_u_prefactor_maxp_minh;
      pd.gradu_penalty_fcn = houston_gradu_prefactor_maxp_minh;
      pd.u_dirichlet_penalty_fcn = houston_u_dirichlet_prefactor_maxp_minh;
      pd.penalty_prefactor = ((d4est_laplacian_flux_sipg_params_t *) flux_plain->flux_data)->sipg_penalty_prefactor;
      double *vtk = P4EST_ALLOC_ZERO (double, 4 * nl), *vtkf = P4EST_ALLOC_ZERO (double, 18 * nl);
      double te = wall ();
      double *eta_ref = d4est_estimator_bi_compute (p4est, &ev, &fcns_res, pd, drv_poly_fcn, NULL, d4est_ghost,
                                                    d4est_ghost_data, d4est_ops, d4est_geom, d4est_factors, d4est_geom,
                                                    d4est_factors, d4est_quad, 0, vtk, vtkf);
      double t_ref = wall () - te;
      d4est_b200_ctx_t *ectx = d4est_b200_ctx_new (flux_plain);
      te = wall ();
      double *eta_b = d4est_b200_estimator_bi_compute (ectx, p4est, &ev, &fcns_res, pd, drv_poly_fcn, NULL, d4est_ghost,
                                                       d4est_ghost_data, d4est_ops, d4est_geom, d4est_factors, d4est_geom,
                                                       d4est_factors, d4est_quad, 0, vtk, vtkf);
      double t_b = wall () - te;
      double w = 0., big = 0.;
      for (int i = 0; i < nl; i++) {
        if (fabs (eta_b[i] - eta_ref[i]) > w) w = fabs (eta_b[i] - eta_ref[i]);
        if (fabs (eta_ref[i]) > big) big = fabs (eta_ref[i]);
      }
      printf ("{\"b200_estimator_rel_max\": %.3e, \"t_estimator_reference_s\": %.4f, \"t_estimator_b200_incl_upload_s\": %.4f}\n",
              w / big, t_ref, t_b);
      P4EST_FREE (eta_ref); P4EST_FREE (eta_b); P4EST_FREE (vtk); P4EST_FREE (vtkf); P4EST_FREE (rhs_e); P4EST_FREE (res);
      d4est_b200_ctx_destroy (ectx);
    }
  }
#else
  if (use_b200) { fprintf (stderr, "--b200 needs the device-linked build (ref_driver_b200)\n"); return 2; }
#endif

  if (strcmp (out, "-")) {
    d4est_b200_flat_t *flat = d4est_b200_flatten (p4est, d4est_ghost, d4est_factors, flux_plain);
    if (bc_poly)
      d4est_b200_flat_eval_dirichlet (flat, d4est_ops, d4est_geom, d4est_quad, d4est_factors, flux_plain);
    FILE *f = d4b_dump_open (out);
    dump_desc (f, &flat->desc);
    d4b_dump_put1 (f, "xyz", D4B_F64, 3 * (uint64_t) local_nodes, d4est_factors->xyz);
    d4b_dump_put1 (f, "u", D4B_F64, local_nodes, u);
    d4b_dump_put1 (f, "Au_plain", D4B_F64, local_nodes, Au_plain);
    d4b_dump_put1 (f, "Au_with_opt", D4B_F64, local_nodes, Au_opt);
    int max_deg = 1;
    for (int e = 0; e < flat->desc.n_local; e++)
      if (flat->desc.deg_quad[e] > max_deg) max_deg = flat->desc.deg_quad[e];
    dump_tables (f, d4est_ops, max_deg);
    if (bc_poly) {
      /* d4est_laplacian_build_rhs_with_strong_bc of the reference with the fixture's Dirichlet data,
       * f given on the LGL nodes and on the quadrature points (src/dGMath/d4est_laplacian.c:15-140) */
      const int64_t lnq = d4est_factors->local_sizes.local_nodes_quad;
      double *rhs_l = P4EST_ALLOC_ZERO (double, local_nodes), *rhs_q = P4EST_ALLOC_ZERO (double, local_nodes);
      double *f_l = P4EST_ALLOC_ZERO (double, local_nodes), *f_q = P4EST_ALLOC_ZERO (double, lnq);
      d4est_mesh_init_field (p4est, f_l, drv_rhs_fcn, d4est_ops, d4est_geom, d4est_factors, INIT_FIELD_ON_LOBATTO, NULL);
      d4est_mesh_init_field (p4est, f_q, drv_rhs_fcn, d4est_ops, d4est_geom, d4est_factors, INIT_FIELD_ON_QUAD, NULL);
      d4est_laplacian_build_rhs_with_strong_bc (p4est, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom, d4est_quad,
                                                d4est_factors, &vecs, flux_plain, rhs_l, drv_rhs_fcn,
                                                INIT_FIELD_ON_LOBATTO, NULL, 0);
      d4est_laplacian_build_rhs_with_strong_bc (p4est, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom, d4est_quad,
                                                d4est_factors, &vecs, flux_plain, rhs_q, drv_rhs_fcn,
                                                INIT_FIELD_ON_QUAD, NULL, 0);
      d4b_dump_put1 (f, "rhs_f_lobatto", D4B_F64, local_nodes, f_l);
      d4b_dump_put1 (f, "rhs_f_quad", D4B_F64, lnq, f_q);
      d4b_dump_put1 (f, "rhs_strong_bc_lobatto", D4B_F64, local_nodes, rhs_l);
      d4b_dump_put1 (f, "rhs_strong_bc_quad", D4B_F64, local_nodes, rhs_q);
      {
        /* d4est_estimator_bi_compute of the reference (src/Estimators/d4est_estimator_bi.c:349-602) for
         * the seeded u against rhs_l, with two prefactor families; estimator_vtk gives the four parts */
        const int nl = p4est->local_num_quadrants;
        double *res = P4EST_ALLOC_ZERO (double, local_nodes);
        d4est_elliptic_data_t ev = vecs;
        ev.u = u; ev.Au = res; ev.rhs = rhs_l;
        d4est_elliptic_eqns_t fcns_res = {.build_residual = drv_build_residual,.apply_lhs = apply_plain,.user = flux_plain };
        double pre = ((d4est_laplacian_flux_sipg_params_t *) flux_plain->flux_data)->sipg_penalty_prefactor;
        for (int fam = 0; fam < 2; fam++) {
          d4est_estimator_bi_penalty_data_t pd;
          memset (&pd, 0, sizeof (pd));
          pd.u_penalty_fcn = fam ? bi_u_prefactor_conforming_max_p2_over_h : houston_u_prefactor_maxp_minh;
          pd.gradu_penalty_fcn = fam ? bi_gradu_prefactor_max_h_over_p : houston_gradu_prefactor_maxp_minh;
          pd.u_dirichlet_penalty_fcn = fam ? bi_u_prefactor_conforming_max_p2_over_h : houston_u_dirichlet_prefactor_maxp_minh;
          pd.penalty_prefactor = pre;
          double *vtk = P4EST_ALLOC_ZERO (double, 4 * nl), *vtkf = P4EST_ALLOC_ZERO (double, 18 * nl);
          double *eta2 = d4est_estimator_bi_compute (p4est, &ev, &fcns_res, pd, drv_poly_fcn, NULL, d4est_ghost,
                                                     d4est_ghost_data, d4est_ops, d4est_geom, d4est_factors, d4est_geom,
                                                     d4est_factors, d4est_quad, 0, vtk, vtkf);
          d4b_dump_put1 (f, fam ? "est_eta2_bi_p2h" : "est_eta2_houston", D4B_F64, nl, eta2);
          d4b_dump_put1 (f, fam ? "est_terms_bi_p2h" : "est_terms_houston", D4B_F64, 4 * (uint64_t) nl, vtk);
          P4EST_FREE (eta2); P4EST_FREE (vtk); P4EST_FREE (vtkf);
        }
        d4b_dump_put1 (f, "est_residual", D4B_F64, local_nodes, res);
        d4b_dump_put1 (f, "est_diam_volume", D4B_F64, nl, d4est_factors->diam_volume);
        d4b_dump_scalar_f64 (f, "est_penalty_prefactor", pre);
        P4EST_FREE (res);
      }
      P4EST_FREE (rhs_l); P4EST_FREE (rhs_q); P4EST_FREE (f_l); P4EST_FREE (f_q);
    }

    if (dump_quad) {
      /* volume quadrature operators of the reference, element by element the way the
       * problem files call them (src/Problems/ConstantDensityStar/constant_density_star_fcns.h:352-440,
       * :528-600): interpolation to the quadrature points, Galerkin integral, mass matrix, and
       * the Newton-linearised mass-type term V^T diag(w J f(x, V u0)) V u of
       * d4est_quadrature_apply_fofufofvlilj with a ConstantDensityStar-like f. */
      const int64_t lnq = d4est_factors->local_sizes.local_nodes_quad;
      double *q_interp = P4EST_ALLOC_ZERO (double, lnq);
      double *q_jf = P4EST_ALLOC_ZERO (double, lnq);
      double *q_galerkin = P4EST_ALLOC_ZERO (double, local_nodes);
      double *q_mass = P4EST_ALLOC_ZERO (double, local_nodes);
      double *q_jacterm = P4EST_ALLOC_ZERO (double, local_nodes);
      double *q_resterm = P4EST_ALLOC_ZERO (double, local_nodes);
      double *u0 = P4EST_ALLOC (double, local_nodes);
      for (int i = 0; i < local_nodes; i++) u0[i] = 0.1 * u[i] - 0.05;
      for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
        p4est_tree_t *tree = p4est_tree_array_index (p4est->trees, tt);
        sc_array_t *tquadrants = &tree->quadrants;
        for (int q = 0; q < (int) tquadrants->elem_count; ++q) {
          p4est_quadrant_t *quad = p4est_quadrant_array_index (tquadrants, q);
          d4est_element_data_t *ed = quad->p.user_data;
          d4est_quadrature_volume_t mo;
          mo.dq = ed->dq; mo.tree = ed->tree; mo.element_id = ed->id;
          mo.q[0] = ed->q[0]; mo.q[1] = ed->q[1]; mo.q[2] = ed->q[2];
          double *J_quad = d4est_mesh_get_jacobian_on_quadrature_points (d4est_factors, ed);
          d4est_mesh_data_on_element_t md = d4est_mesh_data_on_element (d4est_factors, ed);
          const int nq3 = d4est_lgl_get_nodes ((P4EST_DIM), ed->deg_quad);
          d4est_quadrature_interpolate (d4est_ops, d4est_quad, d4est_geom, &mo, QUAD_OBJECT_VOLUME,
                                        QUAD_INTEGRAND_UNKNOWN, &u[ed->nodal_stride], ed->deg,
                                        &q_interp[ed->quad_stride], ed->deg_quad);
          d4est_quadrature_apply_galerkin_integral (d4est_ops, d4est_geom, d4est_quad, &mo, QUAD_OBJECT_VOLUME,
                                                    QUAD_INTEGRAND_UNKNOWN, &q_interp[ed->quad_stride], ed->deg,
                                                    J_quad, ed->deg_quad, &q_galerkin[ed->nodal_stride]);
          d4est_quadrature_apply_mass_matrix (d4est_ops, d4est_geom, d4est_quad, &mo, QUAD_OBJECT_VOLUME,
                                              QUAD_INTEGRAND_UNKNOWN, &u[ed->nodal_stride], ed->deg, J_quad,
                                              ed->deg_quad, &q_mass[ed->nodal_stride]);
          d4est_quadrature_apply_fofufofvlilj (d4est_ops, d4est_geom, d4est_quad, &mo, QUAD_OBJECT_VOLUME,
                                               QUAD_INTEGRAND_UNKNOWN, &u[ed->nodal_stride], &u0[ed->nodal_stride],
                                               NULL, ed->deg, md.xyz_quad, J_quad, ed->deg_quad,
                                               &q_jacterm[ed->nodal_stride], star_like_jac_fcn, NULL, NULL, NULL,
                                               QUAD_APPLY_MATRIX, 0, NULL);
          d4est_quadrature_apply_fofufofvlj (d4est_ops, d4est_geom, d4est_quad, &mo, QUAD_OBJECT_VOLUME,
                                             QUAD_INTEGRAND_UNKNOWN, &u0[ed->nodal_stride], NULL, ed->deg, J_quad,
                                             md.xyz_quad, ed->deg_quad, &q_resterm[ed->nodal_stride],
                                             star_like_res_fcn, NULL, NULL, NULL, 0, NULL);
          /* J f(x, V u0) at the quadrature points: what a caller hands to d4b_mesh_set_mass_diag */
          double *u0q = P4EST_ALLOC (double, nq3);
          d4est_quadrature_interpolate (d4est_ops, d4est_quad, d4est_geom, &mo, QUAD_OBJECT_VOLUME,
                                        QUAD_INTEGRAND_UNKNOWN, &u0[ed->nodal_stride], ed->deg, u0q, ed->deg_quad);
          for (int k = 0; k < nq3; k++)
            q_jf[ed->quad_stride + k] = J_quad[k] * star_like_jac_fcn (md.xyz_quad[0][k], md.xyz_quad[1][k],
                                                                       md.xyz_quad[2][k], u0q[k], NULL);
          P4EST_FREE (u0q);
        }
      }
      d4b_dump_put1 (f, "q_u0", D4B_F64, local_nodes, u0);
      d4b_dump_put1 (f, "q_interp", D4B_F64, lnq, q_interp);
      d4b_dump_put1 (f, "q_galerkin", D4B_F64, local_nodes, q_galerkin);
      d4b_dump_put1 (f, "q_mass", D4B_F64, local_nodes, q_mass);
      d4b_dump_put1 (f, "q_jf", D4B_F64, lnq, q_jf);
      d4b_dump_put1 (f, "q_jacterm", D4B_F64, local_nodes, q_jacterm);
      d4b_dump_put1 (f, "q_resterm", D4B_F64, local_nodes, q_resterm);
      P4EST_FREE (q_interp); P4EST_FREE (q_jf); P4EST_FREE (q_galerkin); P4EST_FREE (q_mass);
      P4EST_FREE (q_jacterm); P4EST_FREE (q_resterm); P4EST_FREE (u0);
    }

    if (cheby_imax > 0) {
      /* stiffness-only apply, for the volume kernel in isolation */
      double *Ku = P4EST_ALLOC_ZERO (double, local_nodes);
      d4est_laplacian_apply_stiffness_matrix (p4est, d4est_ops, d4est_geom, d4est_quad,
                                              d4est_factors, u, Ku, local_nodes, 0);
      d4b_dump_put1 (f, "Ku", D4B_F64, local_nodes, Ku);
      P4EST_FREE (Ku);

      /* the smoother exactly as d4est_solver_multigrid_smoother_cheby does it
       * (src/Solver/d4est_solver_multigrid_smoother_cheby.c:264-376): cg_eigs
       * from the current guess (which advances the guess in place), lmax = bound,
       * lmin = lmax / ratio, then cheby_imax iterations leaving u and r = rhs - A u.
       * Problem: rhs = A u_seeded (homogeneous operator), initial guess u_seeded / 2. */
      if (!bc_poly) {
        const double ratio = 30.;
        double *cu = P4EST_ALLOC (double, local_nodes);
        double *cAu = P4EST_ALLOC_ZERO (double, local_nodes);
        double *cr = P4EST_ALLOC_ZERO (double, local_nodes);
        d4est_elliptic_data_t cv = vecs;
        cv.u = cu; cv.Au = cAu; cv.rhs = Au_plain;
        double eig_new = -1., eig_old = -1.;
        for (int i = 0; i < local_nodes; i++) cu[i] = .5 * u[i];
        cg_eigs (p4est, &cv, &fcns_plain, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom,
                 d4est_quad, d4est_factors, cheby_eigs_imax, 0, 1, &eig_new);
        for (int i = 0; i < local_nodes; i++) cu[i] = .5 * u[i];
        cg_eigs (p4est, &cv, &fcns_plain, d4est_ghost, d4est_ghost_data, d4est_ops, d4est_geom,
                 d4est_quad, d4est_factors, cheby_eigs_imax, 0, 0, &eig_old);
        d4b_dump_put1 (f, "cheby_rhs", D4B_F64, local_nodes, Au_plain);
        d4b_dump_put1 (f, "cheby_eig_new", D4B_F64, 1, &eig_new);
        d4b_dump_put1 (f, "cheby_eig_old", D4B_F64, 1, &eig_old);
        d4b_dump_put1 (f, "cheby_u_after_eigs", D4B_F64, local_nodes, cu);
        double lmax = eig_old, lmin = lmax / ratio;
        d4est_solver_multigrid_smoother_cheby_iterate_aux (p4est, d4est_ops, d4est_geom, d4est_quad,
                                                           d4est_factors, d4est_ghost, d4est_ghost_data,
                                                           &cv, &fcns_plain, cr, cheby_imax, lmin, lmax,
                                                           0, 0, 1);
        double params[4] = { (double) cheby_imax, (double) cheby_eigs_imax, ratio, lmax };
        d4b_dump_put1 (f, "cheby_params", D4B_F64, 4, params);
        d4b_dump_put1 (f, "cheby_u", D4B_F64, local_nodes, cu);
        d4b_dump_put1 (f, "cheby_r", D4B_F64, local_nodes, cr);
        P4EST_FREE (cu); P4EST_FREE (cAu); P4EST_FREE (cr);
      }
    }

    /* the reference's multigrid solve (src/Solver/d4est_solver_multigrid.c:1420-1506), one
     * V-cycle per call so that the residual and the iterate after every V-cycle can be
     * recorded.  rhs = A u_seeded, initial guess 0, parameters from the [multigrid],
     * [mg_smoother_cheby], [mg_bottom_solver_cg] sections of the ini file. */
    if (mg_vcycles > 0 && !bc_poly) {
      d4est_solver_multigrid_t *mg_data =
        d4est_solver_multigrid_data_init (p4est, d4est_ops, d4est_geom, d4est_quad, &d4est_ghost,
                                          &d4est_ghost_data, d4est_factors, initial_grid_input, ini);
      double *mu = P4EST_ALLOC_ZERO (double, local_nodes);
      double *mAu = P4EST_ALLOC_ZERO (double, local_nodes);
      double *r2 = P4EST_ALLOC_ZERO (double, mg_vcycles + 1);
      double *u_hist = P4EST_ALLOC_ZERO (double, (size_t) mg_vcycles * local_nodes);
      d4est_elliptic_data_t mv = vecs;
      mv.u = mu; mv.Au = mAu; mv.rhs = Au_plain;
      double t_mg = 0.;
      for (int k = 0; k < mg_vcycles; k++) {
        double t1 = wall ();
        d4est_solver_multigrid_solve (p4est, &mv, &fcns_plain, mg_data);
        t_mg += wall () - t1;
        if (k == 0) r2[0] = mg_data->vcycle_r2_global_last;   /* initial residual (only while no break) */
        r2[k + 1] = mg_data->vcycle_r2_global_current;
        memcpy (u_hist + (size_t) k * local_nodes, mu, sizeof (double) * local_nodes);
      }
      /* initial residual with u = 0 is |rhs|^2 */
      r2[0] = d4est_linalg_vec_dot (Au_plain, Au_plain, local_nodes);
      d4est_solver_multigrid_smoother_cheby_t *cheby = mg_data->smoother->user;
      double levels = (double) mg_data->num_of_levels;
      d4b_dump_put1 (f, "mg_rhs", D4B_F64, local_nodes, Au_plain);
      d4b_dump_put1 (f, "mg_num_levels", D4B_F64, 1, &levels);
      d4b_dump_put1 (f, "mg_r2", D4B_F64, mg_vcycles + 1, r2);
      d4b_dump_put1 (f, "mg_u_hist", D4B_F64, (uint64_t) mg_vcycles * local_nodes, u_hist);
      d4b_dump_put1 (f, "mg_eigs_last", D4B_F64, mg_data->num_of_levels, cheby->eigs);
      double prm[5] = { (double) cheby->cheby_imax, (double) cheby->cheby_eigs_cg_imax,
                        cheby->cheby_eigs_lmax_lmin_ratio, cheby->cheby_eigs_max_multiplier,
                        (double) cheby->cheby_use_new_cg_eigs };
      d4b_dump_put1 (f, "mg_cheby_params", D4B_F64, 5, prm);
      printf ("{\"mg_levels\": %d, \"mg_vcycles\": %d, \"t_mg_s\": %.6f, \"r2_first\": %.6e, \"r2_last\": %.6e}\n",
              mg_data->num_of_levels, mg_vcycles, t_mg, r2[0], r2[mg_vcycles]);
      P4EST_FREE (mu); P4EST_FREE (mAu); P4EST_FREE (r2); P4EST_FREE (u_hist);
      d4est_solver_multigrid_data_destroy (mg_data);
    }
    /* the reference's in-house Krylov solvers on A u = rhs (rhs = A u_seeded, u0 = 0):
     * d4est_solver_fcg_solve (src/Solver/d4est_solver_fcg_improved.c:103-346) preconditioned
     * with its multigrid (d4est_krylov_pc_multigrid, one V-cycle per application as the ini
     * file says), and the unpreconditioned d4est_solver_cg_solve (src/Solver/d4est_solver_cg.c:75-197) */
    if (fcg_imax > 0 && !bc_poly) {
      d4est_solver_multigrid_t *mg_data =
        d4est_solver_multigrid_data_init (p4est, d4est_ops, d4est_geom, d4est_quad, &d4est_ghost,
                                          &d4est_ghost_data, d4est_factors, initial_grid_input, ini);
      d4est_krylov_pc_t *pc = d4est_krylov_pc_multigrid_create (mg_data, NULL);
      fcg_inner_pc_apply = pc->pc_apply;
      pc->pc_apply = fcg_counting_pc_apply;
      fcg_pc_calls = 0; fcg_nodes = local_nodes;
      d4est_solver_fcg_params_t fp;
      memset (&fp, 0, sizeof (fp));
      fp.imax = fcg_imax; fp.rtol = 1e-10; fp.atol = 0.; fp.do_not_use_preconditioner = 0;
      fp.checkpoint_every_n_krylov_its = 0;
      double *fu = P4EST_ALLOC_ZERO (double, local_nodes);
      double *fAu = P4EST_ALLOC_ZERO (double, local_nodes);
      d4est_elliptic_data_t fv = vecs;
      fv.u = fu; fv.Au = fAu; fv.rhs = Au_plain;
      double t1 = wall ();
      d4est_solver_fcg_solve (p4est, &fv, &fcns_plain, &d4est_ghost, &d4est_ghost_data, d4est_ops, d4est_geom,
                              d4est_quad, d4est_factors, &fp, pc, -1, -1);
      double t_fcg = wall () - t1;
      double fprm[4] = { (double) fp.imax, fp.rtol, fp.atol, (double) fcg_pc_calls };
      d4b_dump_put1 (f, "fcg_params", D4B_F64, 4, fprm);
      d4b_dump_put1 (f, "fcg_u", D4B_F64, local_nodes, fu);
      d4b_dump_put1 (f, "fcg_r2", D4B_F64, fcg_pc_calls < 512 ? fcg_pc_calls : 512, fcg_r2_hist);
      /* same, identity preconditioner */
      const int pc_its = fcg_pc_calls;
      fp.do_not_use_preconditioner = 1;
      cg_inner_apply_lhs = fcns_plain.apply_lhs; cg_apply_calls = 0;
      d4est_elliptic_eqns_t fcns_count = fcns_plain;
      fcns_count.apply_lhs = cg_counting_apply_lhs;
      memset (fu, 0, sizeof (double) * local_nodes);
      d4est_solver_fcg_solve (p4est, &fv, &fcns_count, &d4est_ghost, &d4est_ghost_data, d4est_ops, d4est_geom,
                              d4est_quad, d4est_factors, &fp, NULL, -1, -1);
      double nopc_its = (double) (cg_apply_calls - 1);
      d4b_dump_put1 (f, "fcg_nopc_iters", D4B_F64, 1, &nopc_its);
      d4b_dump_put1 (f, "fcg_nopc_u", D4B_F64, local_nodes, fu);
      /* plain CG */
      d4est_solver_cg_params_t cp;
      memset (&cp, 0, sizeof (cp));
      cp.iter = 10 * fcg_imax; cp.rtol = 1e-10; cp.atol = 0.; cp.monitor = 0;
      cg_apply_calls = 0;
      memset (fu, 0, sizeof (double) * local_nodes);
      d4est_solver_cg_solve (p4est, &fv, &fcns_count, &d4est_ghost, &d4est_ghost_data, d4est_ops, d4est_geom,
                             d4est_quad, d4est_factors, &cp, NULL);
      double cprm[4] = { (double) cp.iter, cp.rtol, cp.atol, (double) (cg_apply_calls - 1) };
      d4b_dump_put1 (f, "cg_params", D4B_F64, 4, cprm);
      d4b_dump_put1 (f, "cg_u", D4B_F64, local_nodes, fu);
      printf ("{\"fcg_mg_iters\": %d, \"fcg_nopc_iters\": %d, \"cg_iters\": %d, \"t_fcg_mg_s\": %.4f}\n",
              pc_its, (int) nopc_its, cg_apply_calls - 1, t_fcg);
      P4EST_FREE (fu); P4EST_FREE (fAu);
      d4est_krylov_pc_multigrid_destroy (pc);
      d4est_solver_multigrid_data_destroy (mg_data);
    }
    fclose (f);
    d4est_b200_flat_destroy (flat);
  }

  if (geom_out) {
    /* sidecar for the device geometry generator (d4b_geom_compute): where every element sits
     * in the forest, the parameters of the map, and -- recomputed by the reference's own
     * d4est_mesh_compute_mortar_quadrature_quantities at the very end of the run, when nothing
     * else reads them any more -- hm / hp under every face_h_type the generator supports. */
    FILE *gf = d4b_dump_open (geom_out);
    int nl = p4est->local_num_quadrants;
    int32_t *et = P4EST_ALLOC (int32_t, nl), *eq = P4EST_ALLOC (int32_t, 3 * nl), *edq = P4EST_ALLOC (int32_t, nl);
    for (p4est_topidx_t tt = p4est->first_local_tree; tt <= p4est->last_local_tree; ++tt) {
      p4est_tree_t *tree = p4est_tree_array_index (p4est->trees, tt);
      sc_array_t *tquadrants = &tree->quadrants;
      for (int qq = 0; qq < (int) tquadrants->elem_count; ++qq) {
        p4est_quadrant_t *quad = p4est_quadrant_array_index (tquadrants, qq);
        d4est_element_data_t *ed = (d4est_element_data_t *) (quad->p.user_data);
        et[ed->id] = ed->tree; edq[ed->id] = ed->dq;
        for (int d = 0; d < 3; d++) eq[3 * ed->id + d] = ed->q[d];
      }
    }
    d4b_dump_put1 (gf, "elem_tree", D4B_I32, nl, et);
    d4b_dump_put1 (gf, "elem_q", D4B_I32, 3 * (uint64_t) nl, eq);
    d4b_dump_put1 (gf, "elem_dq", D4B_I32, nl, edq);
    P4EST_FREE (et); P4EST_FREE (eq); P4EST_FREE (edq);
    double prm[9] = { 0 };
    int gtype = -1;
    if (d4est_geom->geom_type == GEOM_BRICK) {
      d4est_geometry_brick_attr_t *b = d4est_geom->user;
      gtype = 0;
      prm[0] = b->X0; prm[1] = b->X1; prm[2] = b->Y0; prm[3] = b->Y1; prm[4] = b->Z0; prm[5] = b->Z1;
    }
    else if (d4est_geom->geom_type == GEOM_CUBED_SPHERE_13TREE) {
      d4est_geometry_cubed_sphere_attr_t *sp = d4est_geom->user;
      gtype = 1;
      prm[0] = sp->R0; prm[1] = sp->R1; prm[2] = sp->R2;
      prm[3] = sp->compactify_inner_shell; prm[4] = sp->compactify_outer_shell;
    }
    d4b_dump_scalar_i64 (gf, "geom_type", gtype);
    d4b_dump_put1 (gf, "geom_params", D4B_F64, 9, prm);
    d4b_dump_scalar_i64 (gf, "face_h_type", initial_grid_input->face_h_type);
    d4b_dump_scalar_i64 (gf, "root_len", P4EST_ROOT_LEN);
    uint64_t m = d4est_factors->local_sizes.local_mortar_nodes_quad;
    static const d4est_mesh_face_h_t types[] = { FACE_H_EQ_J_DIV_SJ_QUAD, FACE_H_EQ_J_DIV_SJ_MIN_LOBATTO,
      FACE_H_EQ_J_DIV_SJ_MEAN_LOBATTO, FACE_H_EQ_J_DIV_SJ_MAX_LOBATTO, FACE_H_EQ_TREE_H, FACE_H_EQ_VOLUME_DIV_AREA };
    for (int t = 0; t < 6; t++) {
      char name[64];
      memset (d4est_factors->hm_mortar_quad, 0, sizeof (double) * m);
      memset (d4est_factors->hp_mortar_quad, 0, sizeof (double) * m);
      d4est_factors->aux_mortar_quad_stride = 0;
      d4est_factors->aux_boundary_quad_stride = 0;
      d4est_mesh_compute_mortar_quadrature_quantities (p4est, d4est_ghost, d4est_ops, d4est_geom, d4est_quad,
                                                       d4est_factors, types[t]);
      snprintf (name, sizeof (name), "hm_type%d", (int) types[t]);
      d4b_dump_put1 (gf, name, D4B_F64, m, d4est_factors->hm_mortar_quad);
      snprintf (name, sizeof (name), "hp_type%d", (int) types[t]);
      d4b_dump_put1 (gf, name, D4B_F64, m, d4est_factors->hp_mortar_quad);
    }
    fclose (gf);
  }

  P4EST_FREE (Au_tmp);
  P4EST_FREE (u);
  P4EST_FREE (Au_plain);
  P4EST_FREE (Au_opt);
  d4est_ghost_data_destroy (d4est_ghost_data);
  d4est_laplacian_flux_destroy (flux_plain);
  if (flux_opt) d4est_laplacian_with_opt_flux_destroy (flux_opt);
  d4est_ghost_destroy (d4est_ghost);
  d4est_mesh_initial_extents_destroy (initial_grid_input);
  d4est_quadrature_destroy (p4est, d4est_ops, d4est_geom, d4est_quad);
  d4est_mesh_data_destroy (d4est_factors);
  d4est_ops_destroy (d4est_ops);
  p4est_destroy (p4est);
  d4est_geometry_destroy (d4est_geom);
  zlog_fini ();
  sc_finalize ();
  sc_MPI_Finalize ();
  return 0;
}
