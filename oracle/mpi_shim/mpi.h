/*
 * mpi.h -- TEST INFRASTRUCTURE. A minimal stand-in for the 15 MPI symbols that the reference's
 * mpi.c uses (mpi.c:142-144,160,171-182,227-236,189,239,264-266), so that the reference's own
 * source compiles and runs in an image that has no MPI (no mpicc/mpirun/mpi.h/libmpi).
 *
 * Ranks are forked processes (GS_MPI_NP, default 1) that exchange data through one shared
 * anonymous mapping; see mpi_shim.c.  This is a deviation from a real `mpirun` and is stated
 * beside every number measured with it.
 *
 * Run-time parameters for the build-time-patched reference (oracle/Makefile): the patched
 * source calls gs_ref_n()/gs_ref_steps()/gs_ref_dt() where the original hardcodes
 * 8 / 500 / 3600.0 (mpi.c:97,107,146-148).
 */
#ifndef GS_MPI_SHIM_H
#define GS_MPI_SHIM_H
#include <stddef.h>

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef ptrdiff_t MPI_Aint;

#define MPI_COMM_WORLD 0
#define MPI_INT 1
#define MPI_DOUBLE 2
#define MPI_SUCCESS 0

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Bcast(void *buf, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Get_address(const void *location, MPI_Aint *address);
MPI_Aint MPI_Aint_diff(MPI_Aint a, MPI_Aint b);
int MPI_Type_create_struct(int count, const int blocklengths[], const MPI_Aint displacements[],
                           const MPI_Datatype types[], MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *type);
int MPI_Type_free(MPI_Datatype *type);
int MPI_Allgatherv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf,
                   const int recvcounts[], const int displs[], MPI_Datatype recvtype, MPI_Comm comm);
int MPI_Barrier(MPI_Comm comm);
double MPI_Wtime(void);

int gs_ref_n(void);       /* env GS_REF_N      (default 8,      mpi.c:97,107) */
int gs_ref_steps(void);   /* env GS_REF_STEPS  (default 500,    mpi.c:147)    */
double gs_ref_dt(void);   /* env GS_REF_DT     (default 3600.0, mpi.c:148)    */

#endif
