/* ref_stubs.c -- link-time stand-ins for reference entry points that live in
 * files needing PETSc, HDF5, raw MPI or VTK output (none of them on the
 * apply / Chebyshev / multigrid path).  Each aborts if it is ever reached.
 * TEST INFRASTRUCTURE for oracle/_ref only. */
#include <stdio.h>
#include <stdlib.h>
#include <stdarg.h>
#include <stddef.h>

#define STUB(name) void *name () { fprintf (stderr, "oracle/_ref: %s is not built (needs PETSc/HDF5/MPI)\n", #name); abort (); return NULL; }

STUB (d4est_vtk_save_aux)
STUB (d4est_solver_multigrid_smoother_schwarz_init)
STUB (d4est_solver_multigrid_smoother_schwarz_destroy)
STUB (d4est_solver_multigrid_smoother_krylov_petsc_init)
STUB (d4est_solver_multigrid_smoother_krylov_petsc_destroy)
STUB (d4est_solver_multigrid_bottom_solver_krylov_petsc_init)
STUB (d4est_solver_multigrid_bottom_solver_krylov_petsc_destroy)
STUB (d4est_geometry_cubed_sphere_vtk_new)
STUB (d4est_amr_smooth_pred_init)
STUB (mpsort_mpi)
STUB (d4est_checkpoint_save)
STUB (d4est_solver_krylov_petsc_solve)
STUB (d4est_hessian_compute_hessian_trace_of_field_on_quadrature_points)

double
d4est_power_method ()
{
  fprintf (stderr, "oracle/_ref: d4est_power_method is not built\n");
  abort ();
  return 0.;
}

/* serial: one rank */
int MPI_Comm_size (int comm, int *size) { (void) comm; *size = 1; return 0; }
int MPI_Comm_rank (int comm, int *rank) { (void) comm; *rank = 0; return 0; }

int
PPF_Print (int comm, const char *string, ...)
{
  va_list ap;
  (void) comm;
  va_start (ap, string);
  vprintf (string, ap);
  va_end (ap);
  return 0;
}
