/* TEST INFRASTRUCTURE ONLY: a functional single-rank <mpi.h> for building the reference's own `sphexa` CPU front-end
 * (main/src/sphexa/sphexa.cpp) in a container without MPI (BASELINE.json configs[0]: `sphexa --init sedov -n 50 -s 10`).
 * One rank: reductions copy, point-to-point never happens (aborts if reached).  No reference code. */
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
typedef int MPI_Comm;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Request;
typedef int MPI_Info;
typedef int MPI_Group;
typedef long MPI_Aint;
typedef long long MPI_Offset;
typedef struct { int MPI_SOURCE, MPI_TAG, MPI_ERROR; } MPI_Status;
#define MPI_COMM_WORLD 0
#define MPI_COMM_SELF 1
#define MPI_COMM_NULL (-1)
#define MPI_INFO_NULL 0
#define MPI_SUM 1
#define MPI_MIN 2
#define MPI_MAX 3
#define MPI_PROD 4
#define MPI_LOR 5
#define MPI_LAND 6
#define MPI_IN_PLACE ((void*)1)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_ANY_SOURCE (-1)
#define MPI_ANY_TAG (-1)
#define MPI_REQUEST_NULL 0
#define MPI_SUCCESS 0
#define MPI_THREAD_SINGLE 0
#define MPI_THREAD_FUNNELED 1
#define MPI_THREAD_SERIALIZED 2
#define MPI_THREAD_MULTIPLE 3
#define MPI_MAX_PROCESSOR_NAME 256
/* datatype = (size in bytes << 8) | id */
#define SPHB_DT(size, id) (((size) << 8) | (id))
#define MPI_CHAR SPHB_DT(1, 1)
#define MPI_UNSIGNED_CHAR SPHB_DT(1, 2)
#define MPI_SIGNED_CHAR SPHB_DT(1, 3)
#define MPI_INT SPHB_DT(4, 4)
#define MPI_UNSIGNED SPHB_DT(4, 5)
#define MPI_LONG SPHB_DT(8, 6)
#define MPI_UNSIGNED_LONG SPHB_DT(8, 7)
#define MPI_UNSIGNED_LONG_LONG SPHB_DT(8, 8)
#define MPI_FLOAT SPHB_DT(4, 9)
#define MPI_DOUBLE SPHB_DT(8, 10)
#define MPI_SHORT SPHB_DT(2, 11)
#define MPI_UNSIGNED_SHORT SPHB_DT(2, 12)
#define MPI_LONG_LONG_INT SPHB_DT(8, 13)
#define MPI_LONG_LONG SPHB_DT(8, 13)
#define MPI_BYTE SPHB_DT(1, 14)
#define MPI_UINT64_T SPHB_DT(8, 15)
#define MPI_UINT32_T SPHB_DT(4, 16)
#define MPI_INT64_T SPHB_DT(8, 17)
#define MPI_C_BOOL SPHB_DT(1, 18)
#define MPI_CXX_BOOL SPHB_DT(1, 18)
#define MPI_UINT8_T SPHB_DT(1, 19)
static inline size_t sphb_dtsize(MPI_Datatype t) { return (size_t)(t >> 8); }
static inline int sphb_p2p(const char* what)
{
    fprintf(stderr, "single-rank MPI shim: %s reached\n", what);
    abort();
    return 1;
}
static inline int MPI_Init(int*, char***) { return 0; }
static inline int MPI_Init_thread(int*, char***, int, int* provided) { if (provided) *provided = MPI_THREAD_FUNNELED; return 0; }
static inline int MPI_Finalize(void) { return 0; }
static inline int MPI_Initialized(int* f) { *f = 1; return 0; }
static inline int MPI_Comm_rank(MPI_Comm, int* r) { *r = 0; return 0; }
static inline int MPI_Comm_size(MPI_Comm, int* s) { *s = 1; return 0; }
static inline int MPI_Barrier(MPI_Comm) { return 0; }
static inline int MPI_Abort(MPI_Comm, int code) { exit(code); return 0; }
static inline int MPI_Get_version(int* v, int* sv) { *v = 3; *sv = 1; return 0; }
static inline int MPI_Get_processor_name(char* name, int* len) { strcpy(name, "localhost"); *len = 9; return 0; }
static inline double MPI_Wtime(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return ts.tv_sec + 1e-9 * ts.tv_nsec;
}
static inline int MPI_Allreduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op, MPI_Comm)
{
    if (s != MPI_IN_PLACE) memcpy(r, s, (size_t)n * sphb_dtsize(t));
    return 0;
}
static inline int MPI_Reduce(const void* s, void* r, int n, MPI_Datatype t, MPI_Op, int, MPI_Comm)
{
    if (s != MPI_IN_PLACE) memcpy(r, s, (size_t)n * sphb_dtsize(t));
    return 0;
}
static inline int MPI_Bcast(void*, int, MPI_Datatype, int, MPI_Comm) { return 0; }
static inline int MPI_Allgather(const void* s, int n, MPI_Datatype t, void* r, int, MPI_Datatype, MPI_Comm)
{
    if (s != MPI_IN_PLACE) memcpy(r, s, (size_t)n * sphb_dtsize(t));
    return 0;
}
static inline int MPI_Allgatherv(const void* s, int n, MPI_Datatype t, void* r, const int*, const int* displs, MPI_Datatype,
                                 MPI_Comm)
{
    if (s != MPI_IN_PLACE) memcpy((char*)r + (size_t)(displs ? displs[0] : 0) * sphb_dtsize(t), s, (size_t)n * sphb_dtsize(t));
    return 0;
}
static inline int MPI_Gather(const void* s, int n, MPI_Datatype t, void* r, int, MPI_Datatype, int, MPI_Comm)
{
    if (s != MPI_IN_PLACE) memcpy(r, s, (size_t)n * sphb_dtsize(t));
    return 0;
}
static inline int MPI_Alltoall(const void* s, int n, MPI_Datatype t, void* r, int, MPI_Datatype, MPI_Comm)
{
    if (s != MPI_IN_PLACE) memcpy(r, s, (size_t)n * sphb_dtsize(t));
    return 0;
}
static inline int MPI_Exscan(const void*, void*, int, MPI_Datatype, MPI_Op, MPI_Comm) { return 0; }
static inline int MPI_Scan(const void* s, void* r, int n, MPI_Datatype t, MPI_Op, MPI_Comm)
{
    if (s != MPI_IN_PLACE) memcpy(r, s, (size_t)n * sphb_dtsize(t));
    return 0;
}
static inline int MPI_Isend(const void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*) { return sphb_p2p("MPI_Isend"); }
static inline int MPI_Send(const void*, int, MPI_Datatype, int, int, MPI_Comm) { return sphb_p2p("MPI_Send"); }
static inline int MPI_Irecv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Request*) { return sphb_p2p("MPI_Irecv"); }
static inline int MPI_Recv(void*, int, MPI_Datatype, int, int, MPI_Comm, MPI_Status*) { return sphb_p2p("MPI_Recv"); }
static inline int MPI_Probe(int, int, MPI_Comm, MPI_Status*) { return sphb_p2p("MPI_Probe"); }
static inline int MPI_Iprobe(int, int, MPI_Comm, int* flag, MPI_Status*) { *flag = 0; return 0; }
static inline int MPI_Get_count(const MPI_Status*, MPI_Datatype, int*) { return sphb_p2p("MPI_Get_count"); }
static inline int MPI_Waitall(int, MPI_Request*, MPI_Status*) { return 0; }
static inline int MPI_Wait(MPI_Request*, MPI_Status*) { return 0; }
static inline int MPI_Comm_split(MPI_Comm, int, int, MPI_Comm* out) { *out = 0; return 0; }
static inline int MPI_Comm_split_type(MPI_Comm, int, int, MPI_Info, MPI_Comm* out) { *out = 0; return 0; }
static inline int MPI_Comm_free(MPI_Comm*) { return 0; }
#define MPI_COMM_TYPE_SHARED 1
