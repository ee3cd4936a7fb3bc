/*
 * mpi.h -- MPI stand-in, written for this repo (TEST INFRASTRUCTURE).
 *
 * The container and the GPU box have no MPI.  The reference's MPI program
 * (/root/reference/mpi/mpi_heat_improved_persistent_stat.c) is compiled
 * UNMODIFIED against this header by oracle/build_ref.py.  Only the 20 entry
 * points that file uses exist; their bodies live in ../ref_shim.c:
 *   - $HEAT_REF_RANKS unset or 1: one rank, a 1x1 cartesian grid, every
 *     neighbour MPI_PROC_NULL (Jacobi with exact halos is decomposition
 *     independent, so one rank produces the same bits as any rank count);
 *   - $HEAT_REF_RANKS = N: N forked processes exchanging messages through
 *     shared-memory rings -- the reference's own 2-D block decomposition, its
 *     row_type / col_type halos, persistent requests and LAND allreduce run
 *     as written ("shim, not real MPI": no network, one box).
 * The first MPI_Startall / first MPI_Request_free time-stamp the solver loop
 * (M:159-269) without touching the source.
 */
#ifndef HEAT_MPI_STUB_H
#define HEAT_MPI_STUB_H

typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;

#define MPI_SUCCESS       0
#define MPI_COMM_WORLD    ((MPI_Comm)1)
#define MPI_PROC_NULL     (-2)
#define MPI_FLOAT         ((MPI_Datatype)4)
#define MPI_INT           ((MPI_Datatype)5)
#define MPI_LAND          ((MPI_Op)1)
#define MPI_STATUS_IGNORE ((MPI_Status *)0)

int MPI_Init(int *argc, char ***argv);
int MPI_Finalize(void);
int MPI_Comm_size(MPI_Comm c, int *size);
int MPI_Comm_rank(MPI_Comm c, int *rank);
int MPI_Dims_create(int nnodes, int ndims, int *dims);
int MPI_Cart_create(MPI_Comm c, int ndims, const int *dims, const int *periods, int reorder, MPI_Comm *out);
int MPI_Cart_coords(MPI_Comm c, int rank, int maxdims, int *coords);
int MPI_Cart_shift(MPI_Comm c, int dir, int disp, int *src, int *dst);
int MPI_Type_contiguous(int count, MPI_Datatype old, MPI_Datatype *newt);
int MPI_Type_vector(int count, int blocklen, int stride, MPI_Datatype old, MPI_Datatype *newt);
int MPI_Type_commit(MPI_Datatype *t);
double MPI_Wtime(void);
int MPI_Send(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c);
int MPI_Recv(void *buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Status *st);
int MPI_Send_init(const void *buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm c, MPI_Request *r);
int MPI_Recv_init(void *buf, int count, MPI_Datatype t, int src, int tag, MPI_Comm c, MPI_Request *r);
int MPI_Startall(int count, MPI_Request *reqs);
int MPI_Waitall(int count, MPI_Request *reqs, MPI_Status *sts);
int MPI_Request_free(MPI_Request *r);
int MPI_Allreduce(const void *in, void *out, int count, MPI_Datatype t, MPI_Op op, MPI_Comm c);

#endif
