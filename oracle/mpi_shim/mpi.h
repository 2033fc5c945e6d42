/* mpi.h -- TEST INFRASTRUCTURE ONLY: a single-host stand-in for the dozen MPI calls the reference's
 * nbody/nbody-par.c makes (nbody-par.c:126-201,775-780,874,892,896,913-917,925,962-977), so that the
 * UNMODIFIED reference source can be compiled here (no mpicc/mpirun on this image) and used as a
 * text-level oracle and as the real "nbody-par MPI+OpenMP on host cores" CPU baseline.
 * Ranks are forked inside MPI_Init (MPISHIM_NP of them) and exchange through an anonymous shared
 * mapping; see mpi_shim.c.  Not an MPI implementation: only what that program uses.
 */
#ifndef MPISHIM_MPI_H
#define MPISHIM_MPI_H
#include <stddef.h>

typedef int MPI_Datatype;
typedef int MPI_Comm;
typedef long MPI_Aint;

#define MPI_COMM_WORLD 0
#define MPI_DATATYPE_NULL 0
#define MPI_BYTE 1
#define MPI_INT 2
#define MPI_DOUBLE 3
#define MPI_IN_PLACE ((void *)1)
#define MPI_SUCCESS 0

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_rank(MPI_Comm comm, int *rank);
int MPI_Comm_size(MPI_Comm comm, int *size);
int MPI_Get_address(const void *location, MPI_Aint *address);
int MPI_Type_create_struct(int count, const int blocklengths[], const MPI_Aint displacements[],
                           const MPI_Datatype types[], MPI_Datatype *newtype);
int MPI_Type_commit(MPI_Datatype *type);
int MPI_Type_free(MPI_Datatype *type);
int MPI_Bcast(void *buffer, int count, MPI_Datatype type, int root, MPI_Comm comm);
int MPI_Barrier(MPI_Comm comm);
int MPI_Allgatherv(const void *sendbuf, int sendcount, MPI_Datatype sendtype, void *recvbuf,
                   const int recvcounts[], const int displs[], MPI_Datatype recvtype, MPI_Comm comm);
double MPI_Wtime(void);

#endif
