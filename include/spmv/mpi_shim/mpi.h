/* mpi.h -- the message-passing subset LIBSPMV's host code needs, without MPI.
 *
 * The GPU box has no MPI.  The reference's public signatures
 * carry `MPI_Comm` (L2GMap.h:41, Matrix.h:106, cg.h:33, read_petsc.h:30), so this
 * header keeps those names and provides the ~25 calls the in-scope host code makes
 * (the list in SURVEY.md section 8c), over three transports:
 *   - self      : a size-1 world, the default;
 *   - threads   : N ranks as threads of one process (spmv_shim_run_ranks);
 *   - callbacks : N processes, one per GPU; the launcher (torchrun + the Python
 *                 harness, or any host program) supplies `barrier` and a byte-wise
 *                 `alltoallv`, everything else is built on those two.
 * Only SETUP traffic (halo plan construction, create_matrix) and host scalars go
 * through here.  The hot path never does: halo values and CG scalars move over
 * NCCL (include/b200spmv.h).
 *
 * Semantics pinned for reproducibility: MPI_Allreduce / MPI_Reduce fold the
 * contributions in rank order 0,1,...,size-1; neighbourhood collectives order their
 * blocks by the `sources` / `destinations` arrays given to
 * MPI_Dist_graph_create_adjacent (MPI-3, which L2GMap.cpp:415-447 relies on).
 * One-sided and shared-memory-window calls (out of scope, L2GMap.cpp:645-865) are
 * declared so such code compiles, and abort when called.
 */
#ifndef SPMV_MPI_SHIM_H
#define SPMV_MPI_SHIM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define SPMV_SHIM_API __attribute__((visibility("default")))
#else
#define SPMV_SHIM_API
#endif

#define SPMV_MPI_SHIM 1

typedef struct spmv_shim_comm* MPI_Comm;
typedef struct spmv_shim_request* MPI_Request;
typedef int MPI_Datatype;
typedef int MPI_Op;
typedef int MPI_Info;
typedef int MPI_Group;
typedef struct spmv_shim_win* MPI_Win;
typedef ptrdiff_t MPI_Aint;
typedef struct {
  int MPI_SOURCE, MPI_TAG, MPI_ERROR;
} MPI_Status;

#define MPI_SUCCESS 0
#define MPI_ERR_OTHER 15
#define MPI_MAX_ERROR_STRING 256
#define MPI_UNDEFINED (-32766)
#define MPI_COMM_NULL ((MPI_Comm)0)
#define MPI_REQUEST_NULL ((MPI_Request)0)
#define MPI_INFO_NULL 0
#define MPI_UNWEIGHTED ((int*)0)
#define MPI_STATUS_IGNORE ((MPI_Status*)0)
#define MPI_STATUSES_IGNORE ((MPI_Status*)0)
#define MPI_COMM_TYPE_SHARED 1
#define MPI_MODE_NOCHECK 1
#define MPI_MODE_NOPRECEDE 2
#define MPI_MODE_NOSUCCEED 4
#define MPI_LOCK_SHARED 1
#define MPI_THREAD_FUNNELED 1

/* datatypes: the value is an id, the size comes from spmv_shim_type_size */
#define MPI_CHAR 1
#define MPI_BYTE 2
#define MPI_INT 3
#define MPI_INT32_T 4
#define MPI_INT64_T 5
#define MPI_FLOAT 6
#define MPI_DOUBLE 7
#define MPI_C_FLOAT_COMPLEX 8
#define MPI_DOUBLE_COMPLEX 9
#define MPI_C_DOUBLE_COMPLEX 9

#define MPI_SUM 1
#define MPI_MIN 2
#define MPI_MAX 3

/* the calling thread's world: a thread-rank's world inside spmv_shim_run_ranks,
 * else the process world (callbacks transport if installed, else self) */
SPMV_SHIM_API MPI_Comm spmv_shim_comm_world(void);
#define MPI_COMM_WORLD (spmv_shim_comm_world())

/* ---- transports -------------------------------------------------------- */
/* Run fn(arg) on n thread-ranks; MPI_COMM_WORLD inside is the n-rank world.
 * Returns 0, or 1 if any rank threw / aborted. */
SPMV_SHIM_API int spmv_shim_run_ranks(int n, void (*fn)(void* arg), void* arg);

/* byte-wise alltoallv: counts/displacements in BYTES, arrays of `size` entries */
typedef int (*spmv_shim_alltoallv_fn)(void* ctx, const void* sendbuf, const int64_t* sendcounts,
                                      const int64_t* sdispls, void* recvbuf, const int64_t* recvcounts,
                                      const int64_t* rdispls);
typedef int (*spmv_shim_barrier_fn)(void* ctx);
/* Install a multi-process world as this process's MPI_COMM_WORLD (replaces any
 * previous one).  size == 1 with NULL callbacks re-installs the self world. */
SPMV_SHIM_API int spmv_shim_set_world_callbacks(int rank, int size, spmv_shim_barrier_fn barrier,
                                                spmv_shim_alltoallv_fn alltoallv, void* ctx);
SPMV_SHIM_API int spmv_shim_type_size(MPI_Datatype t);

/* ---- the MPI subset ---------------------------------------------------- */
SPMV_SHIM_API int MPI_Init(int* argc, char*** argv);
SPMV_SHIM_API int MPI_Init_thread(int* argc, char*** argv, int required, int* provided);
SPMV_SHIM_API int MPI_Finalize(void);
SPMV_SHIM_API int MPI_Comm_size(MPI_Comm comm, int* size);
SPMV_SHIM_API int MPI_Comm_rank(MPI_Comm comm, int* rank);
SPMV_SHIM_API int MPI_Comm_free(MPI_Comm* comm);
SPMV_SHIM_API int MPI_Barrier(MPI_Comm comm);
SPMV_SHIM_API double MPI_Wtime(void);
SPMV_SHIM_API int MPI_Pcontrol(const int level, ...);
SPMV_SHIM_API int MPI_Error_string(int errorcode, char* string, int* resultlen);

SPMV_SHIM_API int MPI_Bcast(void* buf, int count, MPI_Datatype t, int root, MPI_Comm comm);
SPMV_SHIM_API int MPI_Allgather(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf,
                                int recvcount, MPI_Datatype recvtype, MPI_Comm comm);
SPMV_SHIM_API int MPI_Alltoall(const void* sendbuf, int sendcount, MPI_Datatype sendtype, void* recvbuf,
                               int recvcount, MPI_Datatype recvtype, MPI_Comm comm);
SPMV_SHIM_API int MPI_Alltoallv(const void* sendbuf, const int* sendcounts, const int* sdispls,
                                MPI_Datatype sendtype, void* recvbuf, const int* recvcounts,
                                const int* rdispls, MPI_Datatype recvtype, MPI_Comm comm);
SPMV_SHIM_API int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype t, MPI_Op op,
                                MPI_Comm comm);
SPMV_SHIM_API int MPI_Reduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype t, MPI_Op op,
                             int root, MPI_Comm comm);

SPMV_SHIM_API int MPI_Dist_graph_create_adjacent(MPI_Comm comm_old, int indegree, const int sources[],
                                                 const int sourceweights[], int outdegree,
                                                 const int destinations[], const int destweights[],
                                                 MPI_Info info, int reorder, MPI_Comm* comm_dist_graph);
SPMV_SHIM_API int MPI_Neighbor_alltoall(const void* sendbuf, int sendcount, MPI_Datatype sendtype,
                                        void* recvbuf, int recvcount, MPI_Datatype recvtype, MPI_Comm comm);
SPMV_SHIM_API int MPI_Neighbor_alltoallv(const void* sendbuf, const int sendcounts[], const int sdispls[],
                                         MPI_Datatype sendtype, void* recvbuf, const int recvcounts[],
                                         const int rdispls[], MPI_Datatype recvtype, MPI_Comm comm);
SPMV_SHIM_API int MPI_Ineighbor_alltoallv(const void* sendbuf, const int sendcounts[], const int sdispls[],
                                          MPI_Datatype sendtype, void* recvbuf, const int recvcounts[],
                                          const int rdispls[], MPI_Datatype recvtype, MPI_Comm comm,
                                          MPI_Request* request);

/* point to point: threads transport only (the callbacks transport aborts) */
SPMV_SHIM_API int MPI_Send(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm comm);
SPMV_SHIM_API int MPI_Recv(void* buf, int count, MPI_Datatype t, int source, int tag, MPI_Comm comm,
                           MPI_Status* status);
SPMV_SHIM_API int MPI_Isend(const void* buf, int count, MPI_Datatype t, int dest, int tag, MPI_Comm comm,
                            MPI_Request* request);
SPMV_SHIM_API int MPI_Irecv(void* buf, int count, MPI_Datatype t, int source, int tag, MPI_Comm comm,
                            MPI_Request* request);
SPMV_SHIM_API int MPI_Wait(MPI_Request* request, MPI_Status* status);
SPMV_SHIM_API int MPI_Waitall(int count, MPI_Request requests[], MPI_Status statuses[]);

/* declared so out-of-scope code compiles; every one aborts when called */
SPMV_SHIM_API int MPI_Comm_split_type(MPI_Comm comm, int split_type, int key, MPI_Info info, MPI_Comm* newcomm);
SPMV_SHIM_API int MPI_Comm_group(MPI_Comm comm, MPI_Group* group);
SPMV_SHIM_API int MPI_Group_translate_ranks(MPI_Group g1, int n, const int r1[], MPI_Group g2, int r2[]);
SPMV_SHIM_API int MPI_Info_create(MPI_Info* info);
SPMV_SHIM_API int MPI_Info_set(MPI_Info info, const char* key, const char* value);
SPMV_SHIM_API int MPI_Win_create(void* base, MPI_Aint size, int disp_unit, MPI_Info info, MPI_Comm comm, MPI_Win* win);
SPMV_SHIM_API int MPI_Win_allocate_shared(MPI_Aint size, int disp_unit, MPI_Info info, MPI_Comm comm, void* baseptr, MPI_Win* win);
SPMV_SHIM_API int MPI_Win_shared_query(MPI_Win win, int rank, MPI_Aint* size, int* disp_unit, void* baseptr);
SPMV_SHIM_API int MPI_Win_free(MPI_Win* win);
SPMV_SHIM_API int MPI_Win_fence(int assert_, MPI_Win win);
SPMV_SHIM_API int MPI_Win_sync(MPI_Win win);
SPMV_SHIM_API int MPI_Win_lock(int lock_type, int rank, int assert_, MPI_Win win);
SPMV_SHIM_API int MPI_Win_unlock(int rank, MPI_Win win);
SPMV_SHIM_API int MPI_Win_lock_all(int assert_, MPI_Win win);
SPMV_SHIM_API int MPI_Win_unlock_all(MPI_Win win);
SPMV_SHIM_API int MPI_Put(const void* origin, int ocount, MPI_Datatype otype, int target_rank, MPI_Aint target_disp,
                          int tcount, MPI_Datatype ttype, MPI_Win win);

#ifdef __cplusplus
}
#endif
#endif /* SPMV_MPI_SHIM_H */
