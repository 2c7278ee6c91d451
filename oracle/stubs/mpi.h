/* Single-rank stand-in for <mpi.h>, enough for /root/reference/main.cu to compile and run as one
 * process (no MPI implementation is installed in this image).  With size == 1 the reference never
 * reaches its MPI_Send / MPI_Recv calls (their loops run for ranks 1..size-1), so these only need
 * to exist.  TEST INFRASTRUCTURE ONLY. */
#ifndef SVDJ_STUB_MPI_H_
#define SVDJ_STUB_MPI_H_
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_DOUBLE 0
#define MPI_SUCCESS 0
static inline int MPI_Init(int *, char ***) { return MPI_SUCCESS; }
static inline int MPI_Finalize(void) { return MPI_SUCCESS; }
static inline int MPI_Comm_rank(MPI_Comm, int *r) { *r = 0; return MPI_SUCCESS; }
static inline int MPI_Comm_size(MPI_Comm, int *s) { *s = 1; return MPI_SUCCESS; }
static inline int MPI_Send(const void *, int, MPI_Datatype, int, int, MPI_Comm) { return MPI_SUCCESS; }
static inline int MPI_Recv(void *, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status *) { return MPI_SUCCESS; }
static inline int MPI_Barrier(MPI_Comm) { return MPI_SUCCESS; }
#endif
