/* Type-only stand-in for PETSc headers so that the reference's solver headers
 * parse.  No PETSc function is provided: reference files that call PETSc are
 * not compiled into oracle/_ref (PETSc 3.12.4 tarball is absent). */
#ifndef D4B_SHIM_PETSC_H
#define D4B_SHIM_PETSC_H
typedef struct _p_KSP *KSP;
typedef struct _p_PC *PC;
typedef struct _p_SNES *SNES;
typedef struct _p_Vec *Vec;
typedef struct _p_Mat *Mat;
typedef int PetscErrorCode;
typedef int PetscInt;
typedef double PetscReal;
typedef double PetscScalar;
typedef int PetscBool;
typedef int KSPConvergedReason;
typedef int SNESConvergedReason;
#endif
