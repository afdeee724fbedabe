/* TEST INFRASTRUCTURE ONLY: single-rank stand-in for <mpi.h> so that reference headers which merely mention MPI types
 * in their signatures can be compiled into oracle/_ref.  No reference code; nothing here is ever called. */
#pragma once
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_SUM 1
#define MPI_MIN 2
#define MPI_MAX 3
#define MPI_IN_PLACE ((void*)1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_CHAR 1
#define MPI_UNSIGNED_CHAR 2
#define MPI_INT 4
#define MPI_UNSIGNED 5
#define MPI_LONG 6
#define MPI_UNSIGNED_LONG 7
#define MPI_UNSIGNED_LONG_LONG 8
#define MPI_FLOAT 9
#define MPI_DOUBLE 10
#define MPI_SHORT 11
#define MPI_UNSIGNED_SHORT 12
#define MPI_LONG_LONG_INT 13
#define MPI_BYTE 14
#define MPI_SUCCESS 0
static inline int MPI_Reduce(const void*, void*, int, MPI_Datatype, MPI_Op, int, MPI_Comm) { return 1; }
static inline int MPI_Allreduce(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm) { return 1; }
static inline int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return 0; }
static inline int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return 0; }
