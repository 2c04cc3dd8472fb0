/* hdf5.h shim for oracle/_ref (TEST INFRASTRUCTURE): the reference's in-house FCG
 * (src/Solver/d4est_solver_fcg_improved.c) names hid_t / H5T_NATIVE_DOUBLE in a checkpoint
 * call that the oracle never reaches (checkpoint_every_n_krylov_its = 0). */
#ifndef D4B_HDF5_SHIM_H
#define D4B_HDF5_SHIM_H
typedef long hid_t;
#define H5T_NATIVE_DOUBLE ((hid_t) 1)
#define H5T_NATIVE_INT ((hid_t) 2)
#endif
