/* Serial stand-in for <mpi.h>: maps the few raw MPI_* names the reference
 * uses onto libsc's built-in no-MPI wrappers (sc_mpi.h, !SC_ENABLE_MPI).
 * TEST INFRASTRUCTURE for oracle/_ref only. */
#ifndef D4B_SHIM_MPI_H
#define D4B_SHIM_MPI_H
#include <sc.h>
typedef sc_MPI_Comm MPI_Comm;
typedef sc_MPI_Datatype MPI_Datatype;
typedef sc_MPI_Op MPI_Op;
#define MPI_COMM_WORLD sc_MPI_COMM_WORLD
#define MPI_Comm_size sc_MPI_Comm_size
#define MPI_Comm_rank sc_MPI_Comm_rank
#define MPI_Barrier sc_MPI_Barrier
#define MPI_DOUBLE sc_MPI_DOUBLE
#define MPI_INT sc_MPI_INT
#define MPI_SUM sc_MPI_SUM
#define MPI_MAX sc_MPI_MAX
#define MPI_MIN sc_MPI_MIN
#define MPI_Allreduce sc_MPI_Allreduce
#define MPI_Reduce sc_MPI_Reduce
#define MPI_Bcast sc_MPI_Bcast
#define MPI_Gather sc_MPI_Gather
#define MPI_Allgather sc_MPI_Allgather
#define MPI_SUCCESS sc_MPI_SUCCESS
#endif
