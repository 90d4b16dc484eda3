/*
 * Stub <mpi.h> used ONLY to compile the reference sources in place
 * (oracle/Makefile -> oracle/_ref/).  MPI is not installed in this image; the
 * reference includes <mpi.h> from src/fl/fl_common.cuh:5 and calls a handful
 * of entry points on its fl-mpi / fl-nccl paths, none of which the oracle
 * ever executes (only fl-cpu, and fl on a GPU box).  Every call is a no-op
 * reporting a single-rank world.
 */
#ifndef FLRL_ORACLE_MPI_STUB_H
#define FLRL_ORACLE_MPI_STUB_H

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS 0
#define MPI_UNSIGNED_LONG 1
#define MPI_UNSIGNED_CHAR 2
#define MPI_BYTE 3
#define MPI_STATUS_IGNORE ((MPI_Status *)0)

static inline int MPI_Init(int *, char ***) { return MPI_SUCCESS; }
static inline int MPI_Finalize(void) { return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm, int *r) { *r = 0; return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm, int *s) { *s = 1; return MPI_SUCCESS; }
static inline int MPI_Send(const void *, int, MPI_Datatype, int, int, MPI_Comm) { return MPI_SUCCESS; }
static inline int MPI_Recv(void *, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status *) { return MPI_SUCCESS; }
static inline int MPI_Bcast(void *, int, MPI_Datatype, int, MPI_Comm) { return MPI_SUCCESS; }
static inline int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
static inline int MPI_Allgather(const void *, int, MPI_Datatype, void *, int, MPI_Datatype, MPI_Comm) { return MPI_SUCCESS; }

#endif
